# In-step ncu capture of our GEMM launches inside the bench command (warm L2, real operands): one full step
# = 160 launches of lora_gemm_tc2_kernel (80 forward, 80 backward-dX), after two eager steps.
# Metrics are the subset of `--set full` that profiles/traffic.json and DESIGN.md quote.
set -x
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:lora_gemm_tc2 --launch-skip 320 -c 160 --csv --log-file gpurun_out/instep_gemm_v11.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/instep_ncu.json 2> gpurun_out/instep_ncu.err
echo "rc=$?"
ls -la gpurun_out/instep_gemm_v11.csv
