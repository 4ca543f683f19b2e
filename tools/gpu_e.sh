#!/bin/bash
# 2-GPU batch: NCCL DP correctness test, bench at N=2 (teardown, replica check, exposed all-reduce, stock DDP arm)
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/e_build.log 2>&1
timeout 600 python -m pytest tests/test_gpu_dp2.py -m gpu -q -p no:cacheprovider > gpurun_out/e_dp2.log 2>&1; echo "dp2 rc=$?" > gpurun_out/e_rc.txt
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/e_bench_n2.json 2> gpurun_out/e_bench_n2.err ) 2> gpurun_out/e_time.txt; echo "bench n2 rc=$?" >> gpurun_out/e_rc.txt
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 1 --impl reference > gpurun_out/e_ref_n2.json 2> gpurun_out/e_ref_n2.err ) 2>> gpurun_out/e_time.txt; echo "ref n2 rc=$?" >> gpurun_out/e_rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 -m diffusion_finetuning_b200.train --synthetic --max-train-steps 30 --mixed-precision bf16 --use-cuda-graph --fuse-siblings --lr-warmup-steps 0 --save-steps 0 --output-dir gpurun_out/e_train_out > gpurun_out/e_train.log 2>&1; echo "train n2 rc=$?" >> gpurun_out/e_rc.txt
tail -3 gpurun_out/e_dp2.log; cat gpurun_out/e_rc.txt; cat gpurun_out/e_time.txt; tail -3 gpurun_out/e_train.log; tail -5 gpurun_out/e_bench_n2.err
