#!/bin/bash
# second B box on a third producer thread (B200LORA_SPLIT_B=1): full GPU suite, conv A/B, headline line
mkdir -p gpurun_out
O=gpurun_out
B200LORA_SPLIT_B=1 timeout 150 python -m pytest tests -x -q -m gpu > $O/l4_pytest.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -2 $O/l4_pytest.log
if [ $rc -ne 0 ]; then grep -n "Error\|assert\|FAILED" $O/l4_pytest.log | head -20; exit 1; fi
B200LORA_SPLIT_B=1 timeout 60 python tools/bench_conv.py --hw 16 --no-torch --iters 10 > $O/l4_conv_split.log 2>&1
B200LORA_SPLIT_B=0 timeout 60 python tools/bench_conv.py --hw 16 --no-torch --iters 10 > $O/l4_conv_nosplit.log 2>&1
grep -h '"Cin"' $O/l4_conv_split.log | cut -c1-120; echo; grep -h '"Cin"' $O/l4_conv_nosplit.log | cut -c1-120
B200LORA_SPLIT_B=1 timeout 120 python bench.py --no-stock --no-cpu-baseline > $O/l4_bench.json 2> $O/l4_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/l4_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["avg_launch_us"])
PY
