timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for pr in 1 0; do echo "PAIR=$pr"; B200LORA_PAIR=$pr python tools/bench_layer.py --iters 10 --no-torch --back2back 20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except: print(l.strip()[:200]); continue
    print(r['M'],r['K'],r['N'],'fwd',round(r['fwd_us'],1),'b2b',round(r['fwd_b2b_us'],1),'bdx',round(r['bwd_dx_us'],1),'b2b',round(r['bwd_dx_b2b_us'],1),'grads',round(r['grads_us'],1),'b2b',round(r['grads_b2b_us'],1), 'fwdTF(b2b)', round((2.0*r['M']*r['K']*r['N'])/r['fwd_b2b_us']/1e6,0))
"; done
