#!/bin/bash
# config 3 at N = 2: which parameters diverge between the replicas (with and without the GEGLU epilogue, graph and eager)
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/q_build.log 2>&1
run() { name=$1; shift; env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
      bench.py --gpus 2 --config 3 --steps 6 --warmup 3 --no-cpu-baseline --no-stock $EXTRA > $O/q_$name.json 2> $O/q_$name.err; echo "$name rc=$?"
  python - "$name" <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/q_{n}.json").read().strip().splitlines()[-1])
    c = d.get("dp_check", {})
    print(n, d["value"], c.get("params_bit_identical_across_ranks"), c.get("differing_elements"), c.get("differing_params"))
except Exception as e:
    print(n, "ERR", e)
PY
}
EXTRA="" run graph A=1
EXTRA="" run graph_nogeglu B200LORA_BENCH_GEGLU=0
EXTRA="--no-graph" run eager A=1
EXTRA="--no-fuse" run graph_nofuse A=1
