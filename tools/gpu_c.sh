#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/c_build.log 2>&1
timeout 1200 python -m pytest tests/test_gpu_linear.py tests/test_gpu_fused.py tests/test_gpu_conv.py tests/test_gpu_model.py tests/test_gpu_edges.py -m gpu -q --maxfail=20 -p no:cacheprovider > gpurun_out/c_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-stock --no-cpu-baseline > gpurun_out/c_bench_scope1.json 2> gpurun_out/c_bench_scope1.err; echo "scope1 rc=$?" > gpurun_out/c_rc.txt
B200LORA_CLUSTER_SCOPE=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-stock --no-cpu-baseline > gpurun_out/c_bench_scope0.json 2> gpurun_out/c_bench_scope0.err; echo "scope0 rc=$?" >> gpurun_out/c_rc.txt
timeout 600 python bench.py --config 5 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/c_bench5.json 2> gpurun_out/c_bench5.err; echo "bench5 rc=$?" >> gpurun_out/c_rc.txt
tail -4 gpurun_out/c_pytest.log; cat gpurun_out/c_rc.txt
