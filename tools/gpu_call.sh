mkdir -p gpurun_out
{
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "== gpu tests"; timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -25
echo "== cycle probe 2^21"; timeout 300 python tools/cycle_probe.py 2>&1 | tail -3
} > gpurun_out/r2b_log.txt 2>&1
tail -50 gpurun_out/r2b_log.txt
