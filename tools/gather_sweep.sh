for piece in 4096 4704 7056 9408 14112 28224; do for ctas in 1 2; do for st in 32; do
echo -n "piece=$piece ctas=$ctas stages=$st : "
RR_GATHER_PIECE=$piece RR_GATHER_CTAS_PER_SM=$ctas RR_GATHER_STAGES=$st timeout 200 python bench.py --no-cpu-baseline --steps 3000 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['roofline']['kernel_us'],2), round(d['roofline']['frac'],3), round(d['ms_per_step']*1e3,1))"
done; done; done
