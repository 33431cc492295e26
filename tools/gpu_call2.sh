mkdir -p gpurun_out
O=gpurun_out
{
nvidia-smi -L | wc -l
echo "== parity subset"; timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "large or groups" 2>&1 | tail -3
for st in 512 4096; do
echo "== bench --gpus 2, RR_UPD_STAGED=$st"
RR_UPD_STAGED=$st timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 300 --warmup 20 > $O/r2t_bench_2gpu_staged$st.json 2> $O/r2t_bench_2gpu.err
echo "rc=$?"; tail -c 600 $O/r2t_bench_2gpu.err
done
} > $O/r2t_log.txt 2>&1
tail -30 $O/r2t_log.txt
