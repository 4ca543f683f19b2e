# ncu launch list of the bench command (rule: the same command must first exit 0 without ncu).
# --launch-skip 3200 -c 6000: skip the first eager step (one-time cache builds), list the next ~2 steps
# (GraphedStep's eager warm-up: the very kernels that get captured); everything after runs unprofiled.
set -x
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_bench.json 2> gpurun_out/plain_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 3200 -c 6000 --csv \
    --log-file gpurun_out/launches_bench_v10.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline \
    > gpurun_out/ncu_bench.json 2> gpurun_out/ncu_bench.err
echo "launchlist rc=$?"
cut -c1-200 gpurun_out/plain_bench.json
ls -la gpurun_out | tail -5
