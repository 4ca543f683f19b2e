#!/bin/bash
# in-step capture of the GEMM on the final code (after the same command exited 0 without ncu)
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/e4_build.log 2>&1
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-stock > $O/e4_plain.json 2> $O/e4_plain.err && \
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:lora_gemm_tc2 --launch-skip 320 -c 160 --csv --log-file $O/c2_instep_gemm.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-stock > $O/e4_instep.json 2> $O/e4_instep.err
echo "instep rc=$?"
