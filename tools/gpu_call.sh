mkdir -p gpurun_out
O=gpurun_out
{
echo "== gpu tests"
timeout 1100 python -m pytest tests -m gpu -q --durations=12 2>&1 | grep -v "DeprecationWarning\|warnings.warn\|_script.py" | tail -45
echo "== smoke"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== bench default"
date +%s
timeout 800 python bench.py > $O/r2x_bench_1gpu.json 2> $O/r2x_bench_1gpu.err
echo "rc=$?"; date +%s; tail -c 800 $O/r2x_bench_1gpu.err
} > $O/r2x_log.txt 2>&1
tail -60 $O/r2x_log.txt
