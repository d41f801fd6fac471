timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1 || exit 1
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench11.json 2> gpurun_out/bench9.err
export GTN_NO_COOP=1
timeout 300 python tools/profile_step.py 64 800 > gpurun_out/prof_plain.log 2>&1 || exit 1
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1h_launches.csv python tools/profile_step.py 64 800 > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:decoder_ -f -o gpurun_out/r1h_decoder_full python tools/profile_step.py 64 800 > gpurun_out/ncu_full_dec.log 2>&1
unset GTN_NO_COOP
timeout 100 python tools/prof_train.py 64 32 > gpurun_out/prof_train.log 2>&1
GTN_PROF=1 timeout 200 python tools/time_infer.py 200 > gpurun_out/time_infer.log 2>&1
