#!/bin/bash
# the driver's N = 8 launch of the headline bench
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/p8_build.log 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --steps 10 --warmup 3 > $O/p8_bench_n8.json 2> $O/p8_bench_n8.err; echo "bench n8 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/p8_bench_n8.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["dp_check"]["params_bit_identical_across_ranks"], d["stock_gpu"]["value"], d["grad_allreduce"]["exposed_us_per_step"], d["roofline"]["frac"])
PY
