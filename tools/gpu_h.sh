#!/bin/bash
# Round-2 GPU batch H: the streamed (bulk-copy ring) merge kernel — parity, A/B against the register-staged kernel,
# ncu --set full — and config 3 with the frozen UNet in channels_last.
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/h_build.log 2>&1
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_artefacts.py -x -q -m gpu -k "merge or apply or collapse or add" > $O/h_pytest_merge.log 2>&1; echo "pytest merge rc=$?"
tail -15 $O/h_pytest_merge.log
timeout 600 python bench.py --config 5 --no-cpu-baseline > $O/h_bench5_stream.json 2> $O/h_bench5_stream.err; echo "bench5 stream rc=$?"
B200LORA_MERGE_STREAM=0 timeout 600 python bench.py --config 5 --no-cpu-baseline > $O/h_bench5_regs.json 2> $O/h_bench5_regs.err; echo "bench5 regs rc=$?"
python - <<'EOF'
import json
for f in ("h_bench5_stream", "h_bench5_regs"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, {k: (round(v["gbs"]), round(v["ms"], 4)) for k, v in d["merge"].items() if isinstance(v, dict) and "gbs" in v})
    except Exception as e:
        print(f, "ERR", e)
EOF
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lora_merge_stream -c 4 -o $O/h_merge_stream_full -f \
    python bench.py --config 5 --steps 1 --warmup 1 --no-cpu-baseline > $O/h_merge_ncu.json 2> $O/h_merge_ncu.err
echo "merge ncu rc=$?"
ncu -i $O/h_merge_stream_full.ncu-rep --page raw --csv > $O/h_merge_stream_full_raw.csv 2>/dev/null
timeout 900 python bench.py --config 3 --no-cpu-baseline --memory-format channels_last > $O/h_bench3_cl.json 2> $O/h_bench3_cl.err; echo "bench3 cl rc=$?"
cut -c1-300 $O/h_bench3_cl.json
