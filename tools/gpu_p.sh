#!/bin/bash
# N = 8: configs 3 and 4 (BASELINE north_star: "8xB200 data-parallel"), one rank per GPU over NCCL
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/p_build.log 2>&1
nvidia-smi -L | wc -l
for c in 3; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2950$c \
      bench.py --gpus 8 --config $c --steps 10 --warmup 3 --no-cpu-baseline > $O/p_bench${c}_n8.json 2> $O/p_bench${c}_n8.err
  echo "config $c n8 rc=$?"
  tail -c 600 $O/p_bench${c}_n8.json | head -c 300; echo
done
python - <<'PY'
import json
for c in (3,):
    try:
        d = json.loads(open(f"gpurun_out/p_bench{c}_n8.json").read().strip().splitlines()[-1])
        print(c, d["value"], d["ms_per_step"], d["n_gpus"], d.get("dp_check"), (d.get("stock_gpu") or {}).get("value"), d["grad_allreduce"])
    except Exception as e:
        print(c, "ERR", e)
PY
