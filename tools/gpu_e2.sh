#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/e2_build.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29671 tests/dp2_worker.py gpurun_out/e2_dp2.json > gpurun_out/e2_dp2.log 2>&1; echo "dp2 rc=$?" > gpurun_out/e2_rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-stock --no-graph > gpurun_out/e2_bench_nograph.json 2> gpurun_out/e2_bench_nograph.err; echo "nograph rc=$?" >> gpurun_out/e2_rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --no-stock > gpurun_out/e2_bench_graph.json 2> gpurun_out/e2_bench_graph.err; echo "graph rc=$?" >> gpurun_out/e2_rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-stock --no-fuse --eager-profile > gpurun_out/e2_bench_graph_nofuse.json 2> gpurun_out/e2_bench_graph_nofuse.err; echo "graph nofuse rc=$?" >> gpurun_out/e2_rc.txt
cat gpurun_out/e2_rc.txt; cat gpurun_out/e2_dp2.json
