#!/bin/bash
# final verification batch: whole GPU suite, smoke(), headline line + reference arm + config 4
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/t_build.log 2>&1
timeout 1700 python -m pytest tests -x -q -m gpu > $O/t_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/t_pytest.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > $O/t_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 $O/t_smoke.log
timeout 900 python bench.py > $O/t_bench.json 2> $O/t_bench.err; echo "bench2 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/t_bench_ref.json 2> $O/t_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --config 3 --no-cpu-baseline > $O/t_bench3.json 2> $O/t_bench3.err; echo "bench3 rc=$?"
python - <<'PY'
import json
for f in ("t_bench", "t_bench3", "t_bench_ref"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d.get("e2e", {}).get("value"), (d.get("roofline") or {}).get("frac"), (d.get("stock_gpu") or {}).get("value"), d.get("gpu_launches"))
    except Exception as e:
        print(f, "ERR", e)
PY
