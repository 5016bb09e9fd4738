"""Per-step k-NN statistics of a moving session (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as ge
ge.build()
from sphb200 import ics, constants as C
from sphb200.session import Session, PHASES
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 200000
move = float(sys.argv[2]) if len(sys.argv) > 2 else 0.2
opts = dict(a.split("=") for a in sys.argv[3:])
st = ics.uniform_box(n, 1)
s = Session(n, 40, st["dist"], dt_0=C.dt_0 * 1e-8)
for k, v in opts.items():
    s.ctx.set_option(k, float(v))
s.upload(st); s.startup(); s.check()
c = s.ctx.counter
print("startup: rescan %d widen %d crowded %d worklist %d cells %d" % (c("knn_rescan"), c("knn_widen"), c("knn_crowded"), c("knn_worklist"), c("grid_ncells")))
amp = move * s.median_h()
for it in range(8):
    if amp > 0: s.perturb(amp, 100 + it)
    t = {}
    for ph in PHASES:
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); s.phase(ph); b.record(); torch.cuda.synchronize(); t[ph] = a.elapsed_time(b)
    h = s.download_tensor("sizes")
    print("step %d: rescan %d widen %d crowded %d worklist %d | hmed %.4e | %s" % (it, c("knn_rescan"), c("knn_widen"), c("knn_crowded"), c("knn_worklist"), float(h.median()),
          " ".join("%s=%.3f" % kv for kv in t.items())), flush=True)
s.check()
