GTN_PROF=1 python tools/time_infer.py 200 > gpurun_out/time_infer.log 2>&1
export GTN_NO_COOP=1
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1d_launches.csv python tools/profile_step.py 64 800 > gpurun_out/ncu_launch.log 2>&1
