#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/v_build.log 2>&1
timeout 600 python tools/bench_vs_cublas.py > gpurun_out/v_gemm_vs_cublas.json 2> gpurun_out/v_gemm_vs_cublas.err; echo "rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/v_gemm_vs_cublas.json'))
for r in d['rows']: print(r)"
