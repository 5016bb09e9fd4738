import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as ge
ge.build()
from sphb200 import ics, _lib
from sphb200.session import Session, PHASES
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
st = ics.uniform_box(n, 1)
ctx = _lib.context()
s = Session(n, st["N_NEIGH"], st["dist"]); s.upload(st)
s.startup()
print("startup overflow", ctx.counter("knn_overflow"))
for it in range(2):
    for ph in PHASES:
        t0 = time.time(); s.phase(ph); torch.cuda.synchronize()
        print(it, ph, "%.4fs" % (time.time() - t0), "overflow", ctx.counter("knn_overflow"), "rescan", ctx.counter("knn_rescan"),
              "grid", [ctx.counter(c) for c in ("grid_nx", "grid_ny", "grid_nz")], flush=True)
    for nm in ("dt", "ct", "d", "age", "max_dist", "n_slow"):
        print("  ", nm, s.scalar(nm))
    p = s.download("points"); print("  points finite", np.isfinite(p).all(), np.abs(p).max())
