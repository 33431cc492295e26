"""Why the priority exchange of a striped replay rotates through prefetch + 2 slot arrays
(rela_b200/cpp/sharded_core.cc, updateArrays): a model of the ordering constraints, checked over
random schedules. No GPU, no library: this pins the REASONING the host layer's constant rests on.

Model of one learner L and one shard C that only receives (a replay with ONE learner), P batches
drawn ahead. Per step j each rank's control stream runs, in order, S(j + P) (selection) and U(j)
(update). Constraints taken from the code:
  * C's gather G(y) starts after C's S(y) and finishes at some later time;
  * L's S(y + P) carries the wait for batch y, i.e. finishes only after C's G(y) has;
  * L's U(y) WRITES the words of exchange y into array y % n at its start and needs nobody;
  * C's U(y) POLLS array y % n from its start until L has written exchange y.
The hazard: L writes exchange y into an array C is still polling for an older exchange x (same
array, x < y): C then never sees x's tag. With two arrays that happens as soon as P >= 2; with
P + 2 it cannot."""
import random


def hazard(P, n_arrays, steps, rng):
    # events: ("L", kind, idx) / ("C", kind, idx); each rank's stream is a fixed sequence
    def stream(P):
        seq = [("S", k) for k in range(P)]          # the queue is filled first
        for j in range(steps):
            seq += [("S", j + P), ("U", j)]
        return seq

    Ls, Cs = stream(P), stream(P)
    li = ci = 0
    c_sel_done = set()          # C's selections finished
    c_gather_done = set()       # C's gathers finished (any time after the selection)
    l_written = set()           # exchanges L has written
    c_polling = None            # exchange C's running update is polling for
    while li < len(Ls):
        moves = []
        kind, idx = Ls[li]
        if kind == "U" or idx - P < 0 or (idx - P) in c_gather_done:   # S(y + P) waits for batch y
            moves.append("L")
        if ci < len(Cs) or c_polling is not None:
            moves.append("C")
        pend = sorted(c_sel_done - c_gather_done)
        if pend:
            moves.append("G")
        m = rng.choice(moves)
        if m == "G":
            c_gather_done.add(pend[0])
        elif m == "L":
            if kind == "U":
                if c_polling is not None and c_polling < idx and c_polling % n_arrays == idx % n_arrays:
                    return True   # overwrote the array C is polling for an older exchange
                l_written.add(idx)
            li += 1
        else:
            if c_polling is not None:
                if c_polling in l_written:
                    c_polling = None          # the update saw its words and finished
                continue
            ckind, cidx = Cs[ci]
            if ckind == "S":
                c_sel_done.add(cidx)
            else:
                c_polling = cidx
            ci += 1
    return False


def test_two_arrays_are_not_enough_for_a_single_learner_with_two_batches_drawn_ahead():
    rng = random.Random(1)
    assert any(hazard(2, 2, 12, rng) for _ in range(2000))
    assert any(hazard(4, 2, 12, rng) for _ in range(2000))


def test_prefetch_plus_two_arrays_are():
    rng = random.Random(2)
    for P in (0, 1, 2, 3, 4, 6):
        n = max(2, P + 2)
        assert not any(hazard(P, n, 14, rng) for _ in range(3000)), P
    # ... and one less than prefetch + 1 is not, so the bound is about tight
    assert any(hazard(4, 4, 14, rng) for _ in range(6000))
