mkdir -p gpurun_out
O=gpurun_out
{
nvidia-smi -L | wc -l
echo "== striped tests (world 2)"; timeout 900 python -m pytest tests/test_sharded.py -m gpu -q 2>&1 | grep -v "Warning\|warnings.warn\|_script.py" | tail -60
} > $O/r2z_log.txt 2>&1
tail -70 $O/r2z_log.txt
