mkdir -p gpurun_out
O=gpurun_out
{
for n in 148 128 111 96 74; do
echo "== beside CTAs $n"; RR_GATHER_BESIDE_CTAS=$n timeout 300 python tools/cycle_probe.py $((1<<21)) 512 2000 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print({k: round(v,2) for k,v in d.items() if k.startswith('c_loop') or k.startswith('gather')})"
done
} > $O/r2q_log.txt 2>&1
tail -40 $O/r2q_log.txt
