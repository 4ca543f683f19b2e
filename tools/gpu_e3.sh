#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
run() { name=$1; shift; echo "== $name" >> gpurun_out/e3.log; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/dp_debug.py $EXTRA >> gpurun_out/e3.log 2>gpurun_out/e3_$name.err; echo "rc=$?" >> gpurun_out/e3.log; }
rm -f gpurun_out/e3.log
EXTRA="" run graph A=1
EXTRA="--no-graph" run nograph A=1
EXTRA="" run graph_ring NCCL_ALGO=Ring NCCL_PROTO=Simple
EXTRA="" run graph_nopdl B200LORA_PDL=0
cat gpurun_out/e3.log
