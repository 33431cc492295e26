mkdir -p gpurun_out
O=gpurun_out
{
echo "== host layer tests"
timeout 900 python -m pytest tests/test_gpu_host_layer.py -m gpu -q 2>&1 | grep -v "DeprecationWarning\|warnings.warn\|_script.py" | tail -60
echo "== remaining gpu tests"
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gpu_host_layer.py 2>&1 | grep -v "DeprecationWarning\|warnings.warn\|_script.py" | tail -15
echo "== bench default"
timeout 900 python bench.py > $O/r2w_bench_1gpu.json 2> $O/r2w_bench_1gpu.err
echo "rc=$?"; tail -c 800 $O/r2w_bench_1gpu.err
} > $O/r2w_log.txt 2>&1
tail -40 $O/r2w_log.txt
