"""Per-phase CUDA-event timing of the device-resident step (development aid; bench.py is the contract)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1000000
    kind = sys.argv[2] if len(sys.argv) > 2 else "uniform"
    opts = dict(a.split("=") for a in sys.argv[3:])
    ge.build()
    from sphb200 import ics, _lib
    from sphb200.session import Session, PHASES
    t0 = time.time()
    st = ics.uniform_box(n, 1) if kind == "uniform" else ics.clustered_cloud(n, 2)
    print("ics %.1fs" % (time.time() - t0), flush=True)
    from sphb200 import constants as C
    # dt_0 (the caller-set nsc.dt_0) scaled down so that the repaired viscosity block does not drive
    # temperatures negative within a few steps (DESIGN.md "stability of the reference step")
    s = Session(n, st["N_NEIGH"], st["dist"], dt_0=C.dt_0 * float(opts.pop("dt_scale", 1e-6)))
    for kopt, v in opts.items():
        s.ctx.set_option(kopt, float(v))
    s.upload(st)
    s.startup()
    s.check()
    torch.cuda.synchronize()
    for it in range(4):
        times = {}
        for ph in PHASES:
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record(); s.phase(ph); b.record(); torch.cuda.synchronize()
            times[ph] = a.elapsed_time(b)
        s.check()
        tot = sum(times.values())
        print("iter %d total %.3f ms  %.3e particle-steps/s | " % (it, tot, n / tot * 1e3)
              + " ".join("%s=%.3f" % kv for kv in times.items()), flush=True)
    print("rescan", s.ctx.counter("knn_rescan"), "widen", s.ctx.counter("knn_widen"))


if __name__ == "__main__":
    main()
