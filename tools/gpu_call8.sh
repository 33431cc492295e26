mkdir -p gpurun_out
O=gpurun_out
{
nvidia-smi -L | wc -l
echo "== bench --gpus 8"
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 300 --warmup 20 > $O/r2o_bench_8gpu.json 2> $O/r2o_bench_8gpu.err
echo "rc=$?"; tail -c 2500 $O/r2o_bench_8gpu.err; echo; head -c 7000 $O/r2o_bench_8gpu.json
} > $O/r2o_log.txt 2>&1
tail -100 $O/r2o_log.txt
