#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/y_build.log 2>&1
timeout 300 python -m pytest tests/test_gpu_conv.py -x -q -m gpu > $O/y_pytest_conv.log 2>&1; echo "conv tests rc=$?"; tail -2 $O/y_pytest_conv.log
timeout 900 python bench.py --config 3 --no-cpu-baseline --no-stock > $O/y_bench3.json 2> $O/y_bench3.err; echo "bench3 rc=$?"
timeout 900 python bench.py --config 3 --no-cpu-baseline --no-stock --memory-format channels_last > $O/y_bench3_cl.json 2> $O/y_bench3_cl.err; echo "bench3 cl rc=$?"
python - <<'PY'
import json
for f in ("y_bench3", "y_bench3_cl"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"])
    except Exception as e:
        print(f, "ERR", e)
PY
