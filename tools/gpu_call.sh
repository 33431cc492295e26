mkdir -p gpurun_out
O=gpurun_out
{
echo "== launch list at the full configuration"
CMD="python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-extras --no-parity"
timeout 600 $CMD > $O/r2_plain_full.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:select_kernel|update_kernel|row_copy_kernel" -s 300 -c 3000 --csv --log-file $O/r2_launches_full_capacity.csv $CMD > $O/r2_ncu1.log 2>&1
echo "rc=$?"
echo "== ncu --set full, capacity 131072"
CMD2="python bench.py --capacity 131072 --steps 40 --warmup 5 --no-cpu-baseline --no-extras --no-parity"
timeout 600 $CMD2 > $O/r2_plain_small.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:select_kernel|update_kernel|row_copy_kernel" -s 120 -c 9 -o $O/r2_full $CMD2 > $O/r2_ncu2.log 2>&1
echo "rc=$?"; tail -3 $O/r2_ncu2.log | cut -c1-200
} > $O/r2r_log.txt 2>&1
tail -30 $O/r2r_log.txt
