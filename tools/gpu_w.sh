#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/w_build.log 2>&1
run() { name=$1; shift; env "$@" timeout 300 python tools/bench_vs_cublas.py > gpurun_out/w_$name.json 2> gpurun_out/w_$name.err
python -c "
import json; d=json.load(open('gpurun_out/w_$name.json'))
print('$name', [(r['M'],r['N'],r['ours_fused_us']) for r in d['rows']])"; }
run default A=1
run scope0 B200LORA_CLUSTER_SCOPE=0
run single_cta B200LORA_PAIR=0
run nowprefetch B200LORA_WPREFETCH=0
run nopdl B200LORA_PDL=0
