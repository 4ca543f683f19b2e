#!/bin/bash
# GPU batch A: full GPU test suite, the default bench line (teardown check), config 5, a first sanitizer pass.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/a_build.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=40 -p no:cacheprovider > gpurun_out/a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/a_pytest.log
( time timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err ) 2> gpurun_out/a_bench.time
echo "bench rc=$?" >> gpurun_out/a_bench.time
timeout 600 python bench.py --config 5 --steps 10 --warmup 3 > gpurun_out/a_bench5.json 2> gpurun_out/a_bench5.err
echo "bench5 rc=$?" >> gpurun_out/a_bench.time
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_linear.py -q -x -k "tcgen05_matches_oracle and bfloat16" -p no:cacheprovider > gpurun_out/a_memcheck_linear.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/a_memcheck_linear.log
tail -5 gpurun_out/a_pytest.log; cat gpurun_out/a_bench.time; head -c 1500 gpurun_out/a_bench.json; tail -3 gpurun_out/a_memcheck_linear.log
