#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/d_build.log 2>&1
# relay mode first on the kernel tests alone, under a short timeout (a protocol bug traps through the watchdog)
timeout 600 python -m pytest tests/test_gpu_linear.py tests/test_gpu_fused.py tests/test_gpu_conv.py -m gpu -q --maxfail=5 -p no:cacheprovider > gpurun_out/d_pytest_kernels.log 2>&1
echo "kernels rc=$?" > gpurun_out/d_rc.txt
timeout 1500 python -m pytest tests -m gpu -q --maxfail=40 -p no:cacheprovider --deselect tests/test_gpu_linear.py --deselect tests/test_gpu_fused.py --deselect tests/test_gpu_conv.py > gpurun_out/d_pytest_rest.log 2>&1
echo "rest rc=$?" >> gpurun_out/d_rc.txt
for mode in 2 0; do
  B200LORA_CLUSTER_SCOPE=$mode timeout 600 python bench.py --steps 20 --warmup 5 --no-stock --no-cpu-baseline > gpurun_out/d_bench_scope$mode.json 2> gpurun_out/d_bench_scope$mode.err; echo "scope$mode rc=$?" >> gpurun_out/d_rc.txt
done
tail -3 gpurun_out/d_pytest_kernels.log; tail -3 gpurun_out/d_pytest_rest.log; cat gpurun_out/d_rc.txt
