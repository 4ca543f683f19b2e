#!/bin/bash
# last full pass on 2 GPUs: the whole GPU suite (incl. the 2-rank NCCL test), smoke, the driver-style N=2 line
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/b2_build.log 2>&1
timeout 1700 python -m pytest tests -x -q -m gpu > $O/b2_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/b2_pytest.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > $O/b2_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/b2_smoke.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > $O/b2_bench_n2.json 2> $O/b2_bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/b2_bench_n2.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["dp_check"]["params_bit_identical_across_ranks"], d["stock_gpu"]["value"])
PY
