mkdir -p gpurun_out
O=gpurun_out
{
nvidia-smi -L | wc -l
echo "== striped tests (2 GPUs: world-2 cases)"; timeout 900 python -m pytest tests/test_sharded.py -m gpu -q 2>&1 | grep -v "DeprecationWarning\|warnings.warn" | tail -8
echo "== bench --gpus 2"
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 300 --warmup 20 > $O/r2m_bench_2gpu.json 2> $O/r2m_bench_2gpu.err
echo "rc=$?"; tail -c 1500 $O/r2m_bench_2gpu.err
} > $O/r2m_log.txt 2>&1
tail -30 $O/r2m_log.txt
