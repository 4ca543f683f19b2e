set -x
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
echo "launchlist rc=$?"
python tools/run_one.py --what fwd --M 32768 --K 320 --N 320 --reps 3 > gpurun_out/plain_a.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lora_gemm_tc -s 1 -c 1 -o gpurun_out/prof_fwd_320 -f python tools/run_one.py --what fwd --M 32768 --K 320 --N 320 --reps 3 > gpurun_out/ncu_a.log 2>&1
echo "full a rc=$?"
python tools/run_one.py --what fwd --M 32768 --K 320 --N 2560 --reps 3 > gpurun_out/plain_b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lora_gemm_tc -s 1 -c 1 -o gpurun_out/prof_fwd_geglu -f python tools/run_one.py --what fwd --M 32768 --K 320 --N 2560 --reps 3 > gpurun_out/ncu_b.log 2>&1
echo "full b rc=$?"
python tools/run_one.py --what all --M 8192 --K 640 --N 640 --reps 3 > gpurun_out/plain_c.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none --csv --log-file gpurun_out/launches_layer640.csv python tools/run_one.py --what all --M 8192 --K 640 --N 640 --reps 3 > gpurun_out/ncu_c.log 2>&1
echo "c rc=$?"
ls -la gpurun_out
