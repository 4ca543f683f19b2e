#!/bin/bash
# merge stream kernel iteration: parity + bench
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/i_build.log 2>&1
timeout 600 python -m pytest tests/test_gpu_model.py -x -q -m gpu -k "merge" > $O/i_pytest_merge.log 2>&1; echo "pytest merge rc=$?"
tail -5 $O/i_pytest_merge.log
timeout 600 python bench.py --config 5 --no-cpu-baseline > $O/i_bench5_stream.json 2> $O/i_bench5_stream.err; echo "bench5 stream rc=$?"
python - <<'PY'
import json
for f in ("i_bench5_stream",):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, {k: (round(v["gbs"]), round(v["ms"], 4)) for k, v in d["merge"].items() if isinstance(v, dict) and "gbs" in v})
    except Exception as e:
        print(f, "ERR", e)
PY
