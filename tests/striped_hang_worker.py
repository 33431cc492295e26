"""Worker of tests/test_bench_striped.py::test_a_hung_rank_ends_the_run_with_a_partial_line: two
processes with a gloo process group stand in for the ranks of `bench.py --gpus 2`. Both have
published one measured leg; then rank 1 "hangs" (sleeps) and rank 0 waits for it in a barrier.
The watchdogs must end both processes with exit code 0 and rank 0 must print the line of what
was measured, marked incomplete."""
import os
import sys
import time

import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import bench  # noqa: E402


class Args:
    alpha, beta = 0.6, 0.4


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    state = {"args": Args(), "world": world, "B": 512, "Bg": 512 * world, "K": 10, "W": 3, "total_cap": world << 21,
             "parity": {"ok_on_every_rank": True, "rows_local_frac": 0.875}, "runs": [], "peak": (6538.3, "measured"),
             "t_start": time.time(), "fill_s": 1.0}
    wd = bench.PhaseWatchdog(lambda why: bench.striped_end_run(state, rank, world, why), poll_s=0.05)
    state["wd"] = wd
    wd.enter("a leg that works", 30)
    dist.barrier()
    state["runs"].append({"prefetch": 0, "learners": world, "value_ms": 0.05, "value_ms_python_loop": 0.05,
                          "value_loop": "interpreter", "e2e_ms": None, "gpu_launches": None, "slab_gb": 148.4})
    wd.enter("a leg that hangs", 1.5)
    if rank == 1:
        time.sleep(3600)  # a kernel waiting for a flag that never comes
    dist.barrier()        # rank 0 waits for the hung rank
    print("NOT REACHED", flush=True)


if __name__ == "__main__":
    main()
