export GTN_NO_COOP=1
python tools/profile_step.py 64 800 > gpurun_out/prof_plain.log 2>&1 || exit 1
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1c_launches.csv python tools/profile_step.py 64 800 > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:decoder_ -f -o gpurun_out/r1c_decoder_full python tools/profile_step.py 64 800 > gpurun_out/ncu_full_dec.log 2>&1
timeout 900 ncu --profile-from-start off --section SpeedOfLight --section MemoryWorkloadAnalysis --section LaunchStats --clock-control none -k regex:gemm_tc -f -o /tmp/r1c_gemm python tools/profile_step.py 64 800 > gpurun_out/ncu_gemm.log 2>&1
ncu -i /tmp/r1c_gemm.ncu-rep --page raw --csv > gpurun_out/r1c_gemm_raw.csv 2>/dev/null
ls -la gpurun_out
