mkdir -p gpurun_out
O=gpurun_out
{
echo "== 1: carve-out max for G only"; RR_DEV_CARVEOUT_G=100 timeout 300 python tools/timeline_probe.py $((1<<20)) 2 6 2>&1 | tail -2 | cut -c1-1000
echo "== 2: carve-out max for all"; RR_DEV_CARVEOUT_G=100 RR_DEV_CARVEOUT_CTL=100 timeout 300 python tools/timeline_probe.py $((1<<20)) 2 6 2>&1 | tail -2 | cut -c1-1000
echo "== 3: G 19 stages"; RR_GATHER_STAGES=19 timeout 300 python tools/timeline_probe.py $((1<<20)) 2 6 2>&1 | tail -2 | cut -c1-1000
echo "== 4: G 19 stages + carve-out max for G"; RR_GATHER_STAGES=19 RR_DEV_CARVEOUT_G=100 timeout 300 python tools/timeline_probe.py $((1<<20)) 2 6 2>&1 | tail -2 | cut -c1-1000
echo "== 5: G 14 stages"; RR_GATHER_STAGES=14 timeout 300 python tools/timeline_probe.py $((1<<20)) 2 6 2>&1 | tail -2 | cut -c1-1000
} > $O/r2l_log.txt 2>&1
tail -40 $O/r2l_log.txt
