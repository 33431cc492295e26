# gather kernel tuning sweep (GPU box): piece size x CTAs per SM x stages -> gather us (1 and 2 streams), cycle us
for piece in 4096 4704 5648 7056 9408 14112; do for ctas in 1 2; do
echo -n "piece=$piece ctas=$ctas : "
RR_GATHER_PIECE=$piece RR_GATHER_CTAS_PER_SM=$ctas timeout 200 python tools/cycle_probe.py $((1 << 20)) 512 1500 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k: round(v,2) for k,v in d.items() if k.endswith('_us')})"
done; done
