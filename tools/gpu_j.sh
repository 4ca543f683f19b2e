#!/bin/bash
# ncu --set full (+ source counters) of the streamed merge kernel, one fp32 and one fp16 launch
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/j_build.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"lora_merge_stream_kernel<float" -c 1 -o $O/j_merge_f32 -f \
    python bench.py --config 5 --steps 1 --warmup 1 --no-cpu-baseline > $O/j_ncu1.json 2> $O/j_ncu1.err
echo "ncu f32 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"lora_merge_stream_kernel<__half" -c 1 -o $O/j_merge_f16 -f \
    python bench.py --config 5 --steps 1 --warmup 1 --no-cpu-baseline > $O/j_ncu2.json 2> $O/j_ncu2.err
echo "ncu f16 rc=$?"
ls -la $O/j_*
