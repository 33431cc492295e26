mkdir -p gpurun_out
O=gpurun_out
{
nvidia-smi -L | wc -l
N=$(nvidia-smi -L | wc -l)
echo "== bench --gpus $N (chunks by locality)"
date +%s
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 300 --warmup 20 > $O/r2b_bench_${N}gpu.json 2> $O/r2b_bench_${N}gpu.err
echo "rc=$?"; date +%s; tail -c 1500 $O/r2b_bench_${N}gpu.err
} > $O/r2b_log${N}.txt 2>&1
tail -30 $O/r2b_log*.txt
