mkdir -p gpurun_out
O=gpurun_out
{
echo "== host layer tests"
timeout 900 python -m pytest tests/test_gpu_host_layer.py -m gpu -q -x 2>&1 | grep -v "DeprecationWarning\|warnings.warn\|_script.py" | tail -60
echo "== remaining gpu tests"
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gpu_host_layer.py 2>&1 | grep -v "DeprecationWarning\|warnings.warn\|_script.py" | tail -15
echo "== timeline probe 2^21"
timeout 600 python tools/timeline_probe.py 2097152 2 12 2>&1 | cut -c1-2500 | tail -8
echo "== bench default"
timeout 900 python bench.py > $O/r2v_bench_1gpu.json 2> $O/r2v_bench_1gpu.err
echo "rc=$?"; tail -c 800 $O/r2v_bench_1gpu.err
} > $O/r2v_log.txt 2>&1
tail -40 $O/r2v_log.txt
