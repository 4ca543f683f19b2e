#!/bin/bash
# tile width on the 8x8- and 16x16-level convs: the heuristic's pick vs forced widths
mkdir -p gpurun_out
O=gpurun_out
for hw in 8 16; do
python tools/bench_conv.py --hw $hw --json $O/h4_conv_hw${hw}_auto.json > $O/h4_hw${hw}_auto.log 2>&1
for bn in 64 96 128 160 192; do
  B200LORA_BN=$bn python tools/bench_conv.py --hw $hw --no-torch --json $O/h4_conv_hw${hw}_bn$bn.json > $O/h4_hw${hw}_bn$bn.log 2>&1
done
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/h4_conv_*.json")):
    d = json.load(open(f))
    print(f.split("h4_conv_")[1], [(r["Cin"], r["k"], round(r["fwd_us"], 1), round(r["fwdbwd_us"], 1)) for r in d["rows"]])
PY
