#!/bin/bash
# launch list of the config-3 bench command (extended injection r = 16: resnet convs + GEGLU)
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.build()" > $O/s_build.log 2>&1
timeout 900 python bench.py --config 3 --steps 2 --warmup 1 --no-cpu-baseline --no-stock > $O/s_plain3.json 2> $O/s_plain3.err && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 4000 -c 8000 --csv \
    --log-file $O/s_launches3.csv python bench.py --config 3 --steps 2 --warmup 1 --no-cpu-baseline --no-stock > $O/s_ncu3.json 2> $O/s_ncu3.err
echo "launchlist rc=$?"
ls -la $O/s_launches3.csv
