#!/bin/bash
# merge stream kernel: compile-time variants A/B on the box (rebuilds libb200lora.so per variant)
mkdir -p gpurun_out
O=gpurun_out
run() {
  name=$1; shift
  B200LORA_NVCC_DEFS="$*" python -c "from diffusion_finetuning_b200 import build as b; b.build(force=True)" > $O/k_build_$name.log 2>&1 || { echo "$name build failed"; tail -3 $O/k_build_$name.log; return; }
  timeout 300 python bench.py --config 5 --no-cpu-baseline --steps 10 > $O/k_$name.json 2> $O/k_$name.err
  python - "$name" <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/k_{n}.json").read().strip().splitlines()[-1])
    print(n, {k: round(v["gbs"]) for k, v in d["merge"].items() if isinstance(v, dict) and "gbs" in v})
except Exception as e:
    print(n, "ERR", e)
PY
}
python -c "import __graft_entry__ as g; g.build()" > $O/k_build0.log 2>&1
timeout 600 python -m pytest tests/test_gpu_model.py -x -q -m gpu -k "merge" > $O/k_pytest_merge.log 2>&1; echo "pytest merge rc=$?"
tail -3 $O/k_pytest_merge.log
run default
B200LORA_MERGE_CONTIG=0 run default_rr
run g32_4 -DMS_G32=4 -DMS_RU32=2
B200LORA_MERGE_CONTIG=0 run g32_4_rr -DMS_G32=4 -DMS_RU32=2
run g32_6 -DMS_G32=6 -DMS_RU32=1
run g16_6_ru2 -DMS_RU16=2
python -c "from diffusion_finetuning_b200 import build as b; b.build(force=True)" > /dev/null 2>&1
