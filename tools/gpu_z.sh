#!/bin/bash
# the driver's N = 2 launch of the headline bench, as it will run it at round end
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/z_build.log 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 3 > $O/z_bench_n2.json 2> $O/z_bench_n2.err; echo "bench n2 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/z_ref_n2.json 2> $O/z_ref_n2.err; echo "ref n2 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/z_bench_n2.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["dp_check"]["params_bit_identical_across_ranks"], d["stock_gpu"]["value"], d["grad_allreduce"]["exposed_us_per_step"], d["roofline"]["frac"])
d = json.loads(open("gpurun_out/z_ref_n2.json").read().strip().splitlines()[-1])
print("ref", d["value"], d.get("impl"))
PY
