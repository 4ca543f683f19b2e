#!/bin/bash
# final verification batch: whole GPU suite, smoke(), headline line + reference arm + config 4
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/t_build.log 2>&1
for bn in 128 192; do
  B200LORA_GEGLU_BN=$bn timeout 300 python -m pytest tests/test_gpu_geglu.py -x -q -m gpu > $O/t_pytest_geglu_$bn.log 2>&1; echo "geglu tests BN=$bn rc=$?"
  B200LORA_GEGLU_BN=$bn timeout 300 python tools/bench_geglu.py > $O/t_geglu_bench_$bn.json 2> $O/t_geglu_bench_$bn.err
  python -c "
import json; d=json.load(open('gpurun_out/t_geglu_bench_$bn.json')); print('BN=$bn', [(r['M'], r['fused_fwd_us'], r['fused_fwdbwd_us']) for r in d['rows']])"
done
timeout 1700 python -m pytest tests -x -q -m gpu > $O/t_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/t_pytest.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > $O/t_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 $O/t_smoke.log
timeout 900 python bench.py > $O/t_bench.json 2> $O/t_bench.err; echo "bench2 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/t_bench_ref.json 2> $O/t_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --config 4 --no-cpu-baseline > $O/t_bench4.json 2> $O/t_bench4.err; echo "bench4 rc=$?"
python - <<'PY'
import json
for f in ("t_bench", "t_bench4", "t_bench_ref"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d.get("e2e", {}).get("value"), (d.get("roofline") or {}).get("frac"), (d.get("stock_gpu") or {}).get("value"), d.get("gpu_launches"))
    except Exception as e:
        print(f, "ERR", e)
PY
