set -x
python tools/run_fused.py > gpurun_out/fused_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:'lora_g' -o /tmp/fusedrep -f python tools/run_fused.py > gpurun_out/fused_ncu.log 2>&1
echo "rc=$?"
ncu -i /tmp/fusedrep.ncu-rep --page raw --csv > gpurun_out/fused_raw.csv 2>/dev/null
ls -la /tmp/fusedrep.ncu-rep
