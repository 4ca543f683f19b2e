#!/bin/bash
# ncu metrics of the pair kernel on multi-tile conv launches (16x16 level, K = 11520 .. 23040): what bounds the K loop
# (--set full ends in LaunchFailed on this kernel after two passes; the metric list below replays fine)
mkdir -p gpurun_out
O=gpurun_out
M=gpu__time_duration.sum,sm__cycles_elapsed.max,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed
M=$M,lts__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sector_hit_rate.pct
M=$M,lts__lts2xbar_cycles_active.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_bytes.sum,l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed
M=$M,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_sectors.sum.pct_of_peak_sustained_elapsed
timeout 240 ncu --metrics $M --clock-control none -k regex:lora_gemm_tc2 -s 8 -c 4 --csv --log-file $O/k4_conv_metrics.csv \
  python tools/bench_conv.py --hw 16 --no-torch --iters 2 > $O/k4_ncu.log 2>&1; echo "ncu rc=$?"
tail -3 $O/k4_ncu.log
