#!/bin/bash
# Round-2 GPU batch G: full GPU suite (+ the no-allocation entry point's test alone, in a fresh process), the bench
# lines of configs 2..5 and the reference arm, then the ncu passes (launch list, in-step GEMM, merge --set full).
# Every ncu pass runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/g_build.log 2>&1
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/g_smi.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/g_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
timeout 1700 python -m pytest tests -x -q -m gpu > $O/g_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/g_pytest.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > $O/g_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > $O/g_bench.json 2> $O/g_bench.err; echo "bench2 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/g_bench_ref.json 2> $O/g_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --config 3 --no-cpu-baseline > $O/g_bench3.json 2> $O/g_bench3.err; echo "bench3 rc=$?"
timeout 900 python bench.py --config 4 --no-cpu-baseline > $O/g_bench4.json 2> $O/g_bench4.err; echo "bench4 rc=$?"
timeout 900 python bench.py --config 5 > $O/g_bench5.json 2> $O/g_bench5.err; echo "bench5 rc=$?"
# launch list of the bench command
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-stock > $O/g_plain.json 2> $O/g_plain.err && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 3200 -c 6000 --csv \
    --log-file $O/g_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-stock > $O/g_ncu_bench.json 2> $O/g_ncu_bench.err
echo "launchlist rc=$?"
# in-step capture of the GEMM (warm L2, real operands)
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:lora_gemm_tc2 --launch-skip 320 -c 160 --csv --log-file $O/g_instep_gemm.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-stock > $O/g_instep.json 2> $O/g_instep.err
echo "instep rc=$?"
# merge kernel: one --set full capture per weight dtype
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lora_merge_batch -c 4 -o $O/g_merge_full -f \
    python bench.py --config 5 --steps 1 --warmup 1 --no-cpu-baseline > $O/g_merge_ncu.json 2> $O/g_merge_ncu.err
echo "merge ncu rc=$?"
ncu -i $O/g_merge_full.ncu-rep --page raw --csv > $O/g_merge_full_raw.csv 2>/dev/null
ls -la $O | tail -30
for f in g_bench g_bench3 g_bench4 g_bench5 g_bench_ref; do echo "== $f"; cut -c1-600 $O/$f.json; done
