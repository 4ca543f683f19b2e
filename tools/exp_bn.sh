for bn in 128 160 192 224; do echo "BN=$bn"; B200LORA_BN=$bn python tools/bench_layer.py --iters 10 --no-torch --shapes 0,1,2,4,5 --back2back 20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except: print(l.strip()[:200]); continue
    print(r['M'],r['K'],r['N'],'fwd',round(r['fwd_us'],1),'b2b',round(r['fwd_b2b_us'],1),'bdx',round(r['bwd_dx_us'],1),'b2b',round(r['bwd_dx_b2b_us'],1),'grads',round(r['grads_us'],1),'b2b',round(r['grads_b2b_us'],1))
"; done
