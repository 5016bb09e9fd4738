"""Replays, on ONE GPU, the particle set every rank of an N-rank `bench.py --gpus N` run hands to its session
(owned slab + the ghosts of its neighbours, both halo widths), through the phases of start-up and one step.
A development aid for multi-rank problems when only one GPU is at hand: any kernel that misbehaves on a
particular rank's data shows up here.   python tools/emulate_ranks.py [world] [n_per_rank]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch


def slab(rank, world, n, seed=1):
    from sphb200 import ics
    st = ics.uniform_box(n, seed + rank)
    st["points"][:, 0] += st["side"] * (rank - 0.5 * (world - 1))
    return st


def local_state(rank, world, n, w, cache):
    from sphb200.session import STATE_FIELDS
    for r in (rank - 1, rank, rank + 1):
        if 0 <= r < world and r not in cache:
            cache[r] = slab(r, world, n)
    for r in list(cache):
        if abs(r - rank) > 1:
            del cache[r]
    me = cache[rank]
    side = me["side"]
    lo = side * (rank - 0.5 * world) if rank > 0 else -np.inf
    hi = side * (rank + 1 - 0.5 * world) if rank < world - 1 else np.inf
    parts = [me]
    n_ghost = 0
    for r in (rank - 1, rank + 1):
        if 0 <= r < world:
            x = cache[r]["points"][:, 0]
            m = (x >= lo - w) & (x < hi + w)
            parts.append({k: (np.asarray(cache[r][k])[m] if k in STATE_FIELDS and cache[r].get(k) is not None else None)
                          for k in STATE_FIELDS})
            n_ghost += int(m.sum())
    out = {}
    for k in STATE_FIELDS:
        if me.get(k) is None:
            continue
        out[k] = np.concatenate([np.asarray(p[k]) for p in parts])
    return out, len(me["points"]), n_ghost, me


def run(rank, world, n, w, cache, label):
    from sphb200.session import Session
    from sphb200 import constants as C
    st, n_own, n_ghost, me = local_state(rank, world, n, w, cache)
    s = Session(n_own + n_ghost, me["N_NEIGH"], me["dist"], dt_0=C.dt_0 * 1e-8)
    s.set_scalar("n_owned", n_own)
    s.upload(st)
    t0 = time.time()
    for ph in ("sort", "knn", "pack", "density", "pi", "force", "rho0",
               "crossing", "dt_clamp", "sort", "knn", "pack", "density", "pi", "force", "integrate"):
        s.phase(ph)
        torch.cuda.synchronize()
    s.check()
    h = s.download("sizes")[:n_own]
    print("rank %d/%d %s: n_own=%d ghosts=%d  hmax=%.4e  d=%.4e  %.2fs  worklist=%d" % (
        rank, world, label, n_own, n_ghost, h.max(), s.scalar("d"), time.time() - t0, s.ctx.counter("knn_widen")), flush=True)
    return float(h.max())


def main():
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    n = int(float(sys.argv[2])) if len(sys.argv) > 2 else 1000000
    import __graft_entry__ as ge
    ge.build()
    from sphb200 import ics
    side = ics.uniform_box(10, 1)["side"] * (n / 10.0) ** (1.0 / 3.0)
    w0 = 1.6 * (40 * (world * side ** 3) / (4.1887902047863905 * world * n)) ** (1.0 / 3.0)
    cache = {}
    for rank in range(world):
        hmax = run(rank, world, n, w0, cache, "start-up halo")
        run(rank, world, n, 1.25 * hmax, cache, "step halo")
    print("all ranks done", flush=True)


if __name__ == "__main__":
    main()
