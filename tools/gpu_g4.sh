#!/bin/bash
# conv launches bracketed by the launch profiler: conv tests + the config-3 line
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_conv.py -x -q -m gpu > $O/g4_pytest.log 2>&1; echo "pytest rc=$?"
tail -2 $O/g4_pytest.log
timeout 500 python bench.py --config 3 > $O/g4_bench3.json 2> $O/g4_bench3.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/g4_bench3.json").read().strip().splitlines()[-1])
r = d["roofline"]
print(d["value"], d["ms_per_step"], r["frac"], r.get("conv_launches"), r.get("linear_launches"))
PY
