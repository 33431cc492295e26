#!/bin/bash
# First gpurun call of the next round: everything that was written after round 1's GPU budget ran
# out gets its parity check and its numbers in ONE box session (about 12 minutes).
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/round2_first_call.sh'
# Outputs under gpurun_out/r2a_*.
mkdir -p gpurun_out
O=gpurun_out
{
  echo "== 1. default GPU suite"
  timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
  echo "== 2. gather-stream experiment: its own parity tests, then the C ABI / host-layer suites under the switch"
  RELA_B200_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_gather_stream.py -x -q 2>&1 | tail -5
  RR_GATHER_STREAM=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_host_layer.py \
      tests/test_gpu_checkpoint.py tests/test_gpu_concurrent_add.py tests/test_gpu_gather_convert.py -x -q 2>&1 | tail -5
  echo "== 3. pipelined C ABI loop, one stream vs two (capacity 2^20 and the full 2^21)"
  timeout 300 python tools/bench_gather_stream.py $((1 << 20)) 512 2 5000
  timeout 300 python tools/bench_gather_stream.py $((1 << 21)) 512 2 5000
  timeout 300 python tools/bench_gather_stream.py $((1 << 21)) 512 3 5000
  echo "== 3b. config 4: actor-side math, reference leg on this host + device leg"
  timeout 300 python tools/bench_actor_math.py --iters 3000
  echo "== 4. bench line, default and under the switch (e2e with prefetch 2 is the number to compare)"
  timeout 600 python bench.py > $O/r2a_bench_default.json
  tail -c 400 $O/r2a_bench_default.json; echo
  RR_GATHER_STREAM=1 timeout 600 python bench.py --no-cpu-baseline > $O/r2a_bench_gather_stream.json
  python - <<'P'
import json
for n in ("default", "gather_stream"):
    try:
        d = json.loads(open("gpurun_out/r2a_bench_%s.json" % n).read().strip().splitlines()[-1])
        print(n, "value us", round(d["ms_per_step"] * 1e3, 2), "e2e", [(r["prefetch"], round(r["ms_per_step"] * 1e3, 2)) for r in d["e2e"]["runs"]],
              "gather us", round(d["roofline"]["kernel_us"], 2), "frac", round(d["roofline"]["frac"], 3))
    except Exception as e:  # noqa: BLE001
        print(n, "unreadable:", e)
P
} > $O/r2a_log.txt 2>&1
tail -60 $O/r2a_log.txt
