#!/bin/bash
# after the AccumulateGrad-stream fix: 2-GPU DP test (incl. extended injection) and config 3 at N = 2 with the replica check
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/r_build.log 2>&1
timeout 600 python -m pytest tests/test_gpu_dp2.py -x -q -m gpu -s > $O/r_pytest_dp2.log 2>&1; echo "pytest dp2 rc=$?"
tail -6 $O/r_pytest_dp2.log | cut -c1-1500
for i in 1 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$i \
      bench.py --gpus 2 --config 3 --steps 10 --warmup 3 --no-cpu-baseline --no-stock > $O/r_bench3_n2_$i.json 2> $O/r_bench3_n2_$i.err; echo "config 3 n2 run $i rc=$?"
done
python - <<'PY'
import json
for i in (1, 2):
    try:
        d = json.loads(open(f"gpurun_out/r_bench3_n2_{i}.json").read().strip().splitlines()[-1])
        c = d.get("dp_check", {})
        print(i, d["value"], d["ms_per_step"], c.get("params_bit_identical_across_ranks"), c.get("differing_elements"), c.get("differing_params"))
    except Exception as e:
        print(i, "ERR", e)
PY
