set -x
python tools/run_one.py --what fwd --M 32768 --K 320 --N 2560 --reps 3 > gpurun_out/plain_p.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lora_gemm_tc -s 1 -c 1 -o gpurun_out/prof_pair_geglu -f python tools/run_one.py --what fwd --M 32768 --K 320 --N 2560 --reps 3 > gpurun_out/ncu_p.log 2>&1
echo "rc=$?"
B200LORA_PAIR=0 python tools/run_one.py --what fwd --M 32768 --K 320 --N 2560 --reps 3 > gpurun_out/plain_s.log 2>&1 && \
B200LORA_PAIR=0 ncu --set full --clock-control none --import-source on -k regex:lora_gemm_tc -s 1 -c 1 -o gpurun_out/prof_single_geglu -f python tools/run_one.py --what fwd --M 32768 --K 320 --N 2560 --reps 3 > gpurun_out/ncu_s.log 2>&1
echo "rc=$?"
