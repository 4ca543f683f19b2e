#!/bin/bash
# conv operand tiles as tiled-mode 4-D boxes instead of im2col-mode loads: parity, per-shape A/B, config-3 line
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python -m pytest tests/test_gpu_conv.py -x -q -m gpu > $O/i4_conv.log 2>&1; rc=$?; echo "conv pytest rc=$rc"; tail -2 $O/i4_conv.log
if [ $rc -ne 0 ]; then grep -n "Error\|assert" $O/i4_conv.log | head -20; exit 1; fi
timeout 200 python tools/bench_conv.py --no-torch --iters 10 --json $O/i4_conv_tiled.json > $O/i4_conv_tiled.log 2>&1
B200LORA_CONV_TILED=0 timeout 200 python tools/bench_conv.py --no-torch --iters 10 --json $O/i4_conv_im2col.json > $O/i4_conv_im2col.log 2>&1
python - <<'PY'
import json
a = json.load(open("gpurun_out/i4_conv_tiled.json")); b = json.load(open("gpurun_out/i4_conv_im2col.json"))
for x, y in zip(a["rows"], b["rows"]):
    print(x["Cin"], x["Cout"], x["H"], x["k"], "fwd", round(x["fwd_us"], 1), "vs", round(y["fwd_us"], 1), "fwdbwd", round(x["fwdbwd_us"], 1), "vs", round(y["fwdbwd_us"], 1))
print(a["summary"], b["summary"])
PY
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_conv.py > $O/i4_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/i4_pytest.log
timeout 500 python bench.py --config 3 > $O/i4_bench3.json 2> $O/i4_bench3.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/i4_bench3.json").read().strip().splitlines()[-1])
r = d["roofline"]
print(d["value"], d["ms_per_step"], r["frac"], r.get("conv_launches"), r.get("linear_launches"))
PY
