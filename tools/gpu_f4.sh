#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/f4_build.log 2>&1
timeout 1700 python -m pytest tests -x -q -m gpu > $O/f4_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/f4_pytest.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > $O/f4_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/f4_smoke.log
timeout 900 python bench.py > $O/f4_bench.json 2> $O/f4_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/f4_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["stock_gpu"]["value"], d["cpu_baseline"]["value"], d["gpu_launches"], d["clocks"])
PY
