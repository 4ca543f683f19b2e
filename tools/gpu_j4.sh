#!/bin/bash
# final library: smoke + the driver-style headline line
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/j4_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/j4_smoke.log
timeout 600 python bench.py > $O/j4_bench.json 2> $O/j4_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/j4_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["stock_gpu"]["value"], d["cpu_baseline"]["value"], d["gpu_launches"], d["clocks"])
PY
