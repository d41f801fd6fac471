export GTN_NO_COOP=1
timeout 300 python tools/profile_step.py 64 800 > gpurun_out/prof_plain.log 2>&1 || exit 1
timeout 900 ncu --profile-from-start off -k regex:gemm_tc_kernel --clock-control none --csv --log-file gpurun_out/r1h_gemm_metrics.csv \
  --metrics gpu__time_duration.sum,sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tc.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active,launch__grid_size,launch__registers_per_thread \
  python tools/profile_step.py 64 800 > gpurun_out/ncu_gemm.log 2>&1
unset GTN_NO_COOP
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
