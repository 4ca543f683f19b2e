#!/bin/bash
# pair kernel with the epilogue mode as a template parameter (new) against run-time flags (old), same box
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/d2_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_linear.py tests/test_gpu_fused.py tests/test_gpu_geglu.py tests/test_gpu_conv.py -x -q -m gpu > $O/d2_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/d2_pytest.log
for i in 1 2; do timeout 900 python bench.py --no-cpu-baseline --no-stock > $O/d2_new_$i.json 2> $O/d2_new_$i.err; done
cp diffusion_finetuning_b200/build/ab_old/linear_tc2.cu diffusion_finetuning_b200/csrc/linear_tc2.cu
cp diffusion_finetuning_b200/build/ab_old/common.cuh diffusion_finetuning_b200/csrc/common.cuh
python -c "from diffusion_finetuning_b200 import build as b; b.build(force=True)" > $O/d2_build_old.log 2>&1; echo "old build rc=$?"
for i in 1 2; do timeout 900 python bench.py --no-cpu-baseline --no-stock > $O/d2_old_$i.json 2> $O/d2_old_$i.err; done
python - <<'PY'
import json
for f in ("d2_new_1", "d2_new_2", "d2_old_1", "d2_old_2"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["avg_launch_us"])
    except Exception as e:
        print(f, "ERR", e)
PY
