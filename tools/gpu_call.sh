mkdir -p gpurun_out
O=gpurun_out
{
echo "== gpu tests"; timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | grep -v "DeprecationWarning\|warnings.warn" | tail -30
echo "== cycle probe 2^21"; timeout 300 python tools/cycle_probe.py 2>&1 | tail -2
echo "== e2e breakdown prefetch 0 / 2 (capacity 2^20)"
timeout 200 python tools/e2e_breakdown.py $((1<<20)) 0 2>&1 | tail -2
timeout 200 python tools/e2e_breakdown.py $((1<<20)) 2 2>&1 | tail -2
} > $O/r2d_log.txt 2>&1
tail -60 $O/r2d_log.txt
