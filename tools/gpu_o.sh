#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/o_build.log 2>&1
timeout 300 python tools/bench_geglu.py > $O/o_geglu_bench.json 2> $O/o_geglu_bench.err; echo "geglu bench rc=$?"; cat $O/o_geglu_bench.json | head -60
timeout 900 python bench.py --config 3 --no-cpu-baseline > $O/o_bench3.json 2> $O/o_bench3.err; echo "bench3 rc=$?"
B200LORA_BENCH_GEGLU=0 timeout 900 python bench.py --config 3 --no-cpu-baseline --no-stock > $O/o_bench3_nogeglu.json 2> $O/o_bench3_nogeglu.err; echo "bench3 nogeglu rc=$?"
python - <<'PY'
import json
for f in ("o_bench3", "o_bench3_nogeglu"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["roofline"]["frac"], d.get("stock_gpu", {}).get("value"))
    except Exception as e:
        print(f, "ERR", e)
PY
