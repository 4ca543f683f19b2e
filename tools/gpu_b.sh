#!/bin/bash
# GPU batch B: tests, bench A/B of the pre-wait W loads, config 5 with merge kernel v2, ncu of the merge kernel.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/b_build.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=40 -p no:cacheprovider > gpurun_out/b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/b_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err; echo "bench rc=$?" > gpurun_out/b_rc.txt
B200LORA_WPREFETCH=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-stock --no-cpu-baseline > gpurun_out/b_bench_nowpre.json 2> gpurun_out/b_bench_nowpre.err; echo "nowpre rc=$?" >> gpurun_out/b_rc.txt
timeout 600 python tools/bench_layer.py --iters 20 --json gpurun_out/b_layer_wpre.json > gpurun_out/b_layer_wpre.log 2>&1
B200LORA_WPREFETCH=0 timeout 600 python tools/bench_layer.py --iters 20 --json gpurun_out/b_layer_nowpre.json > gpurun_out/b_layer_nowpre.log 2>&1
timeout 600 python bench.py --config 5 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b_bench5.json 2> gpurun_out/b_bench5.err; echo "bench5 rc=$?" >> gpurun_out/b_rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lora_merge_batch -c 3 -o gpurun_out/b_merge_ncu python bench.py --config 5 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b_ncu5.log 2>&1
echo "ncu rc=$?" >> gpurun_out/b_rc.txt
tail -4 gpurun_out/b_pytest.log; cat gpurun_out/b_rc.txt
