# one ncu pass over every SURVEY section-8 shape (fwd, bwd-dX, gradient reductions, merge)
set -x
python tools/run_one.py --census --reps 1 > gpurun_out/census_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:'lora_g|lora_merge' -o /tmp/census -f \
    python tools/run_one.py --census --reps 1 > gpurun_out/census_ncu.log 2>&1
echo "rc=$?"
ncu -i /tmp/census.ncu-rep --page raw --csv > gpurun_out/census_raw.csv 2>/dev/null
ls -la /tmp/census.ncu-rep gpurun_out/
sz=$(stat -c %s /tmp/census.ncu-rep); if [ "$sz" -lt 40000000 ]; then cp /tmp/census.ncu-rep gpurun_out/; fi
