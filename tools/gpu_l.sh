#!/bin/bash
# Round-2 GPU batch L: full GPU suite on the final merge kernels, config-5 line, ncu --set full of the streamed merge (fp32 + fp16)
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/l_build.log 2>&1
timeout 1700 python -m pytest tests -x -q -m gpu > $O/l_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 $O/l_pytest.log
timeout 900 python bench.py --config 5 > $O/l_bench5.json 2> $O/l_bench5.err; echo "bench5 rc=$?"
B200LORA_MERGE_STREAM=0 timeout 600 python bench.py --config 5 --no-cpu-baseline > $O/l_bench5_regs.json 2> $O/l_bench5_regs.err; echo "bench5 regs rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"lora_merge_stream_kernel<float" -c 1 -o $O/l_merge_f32 -f \
    python bench.py --config 5 --steps 1 --warmup 1 --no-cpu-baseline > $O/l_ncu1.json 2> $O/l_ncu1.err
echo "ncu f32 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"lora_merge_stream_kernel<__half" -c 1 -o $O/l_merge_f16 -f \
    python bench.py --config 5 --steps 1 --warmup 1 --no-cpu-baseline > $O/l_ncu2.json 2> $O/l_ncu2.err
echo "ncu f16 rc=$?"
for f in l_merge_f32 l_merge_f16; do ncu -i $O/$f.ncu-rep --page raw --csv > $O/${f}_raw.csv 2>/dev/null; done
python - <<'PY'
import json
for f in ("l_bench5", "l_bench5_regs"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, {k: (round(v["gbs"]), round(v["ms"], 4)) for k, v in d["merge"].items() if isinstance(v, dict) and "gbs" in v}, d["e2e"]["value"])
    except Exception as e:
        print(f, "ERR", e)
PY
