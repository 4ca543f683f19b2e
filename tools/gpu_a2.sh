#!/bin/bash
# fp32 dX output from the pair kernel's epilogue: parity, then the headline step with and without it
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/a2_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_linear.py tests/test_gpu_fused.py tests/test_gpu_model.py tests/test_gpu_train.py tests/test_gpu_edges.py tests/test_gpu_stale_pack.py -x -q -m gpu > $O/a2_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/a2_pytest.log
timeout 900 python bench.py --no-cpu-baseline --no-stock > $O/a2_bench.json 2> $O/a2_bench.err; echo "bench rc=$?"
B200LORA_CASTRIDE=0 timeout 900 python bench.py --no-cpu-baseline --no-stock > $O/a2_bench_off.json 2> $O/a2_bench_off.err; echo "bench off rc=$?"
timeout 900 python bench.py --no-cpu-baseline --no-stock > $O/a2_bench_b.json 2> $O/a2_bench_b.err
B200LORA_CASTRIDE=0 timeout 900 python bench.py --no-cpu-baseline --no-stock > $O/a2_bench_off_b.json 2> $O/a2_bench_off_b.err
python - <<'PY'
import json
for f in ("a2_bench", "a2_bench_off", "a2_bench_b", "a2_bench_off_b"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["avg_launch_us"])
    except Exception as e:
        print(f, "ERR", e)
PY
