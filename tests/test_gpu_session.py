"""GPU parity of the device-resident time step against the oracle's
restatement of the driver loop (code_running.py:217-540)."""
import numpy as np
import pytest

from oracle import sph_oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-10


def rel(a, b):
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    assert a.shape == b.shape
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def _run(st, nsteps):
    from sphb200.session import Session
    s = Session(len(st["points"]), st["N_NEIGH"], st["dist"])
    s.upload(st)
    s.startup()
    ref = O.startup(st)
    out = []
    for _ in range(nsteps):
        s.step()
        s.check()
        ref = O.step(ref)
        out.append((s, ref))
        yield s, ref


@pytest.mark.parametrize("n,seed,n_dust", [(3000, 0, 0), (10000, 0, 1)])
def test_startup_matches_oracle(built_lib, n, seed, n_dust):
    from sphb200 import ics
    from sphb200.session import Session
    st = ics.uniform_box(n, seed, n_dust=n_dust)
    ref = O.startup(st)
    s = Session(len(st["points"]), st["N_NEIGH"], st["dist"])
    s.upload(st)
    s.startup()
    s.check()
    assert np.array_equal(s.download("neighbor"), ref["neighbor"])
    assert np.array_equal(s.download("sizes"), ref["sizes"])
    assert rel(s.download("densities"), ref["densities"]) < RTOL
    assert rel(s.download("densities_0"), ref["densities_0"]) < RTOL
    assert np.array_equal(s.download("num_nontrivial"), ref["num_nontrivial"])


def test_ten_steps_default_config(built_lib):
    """BASELINE config 1: default config at reduced N = 10^4 gas particles, 10 steps."""
    from sphb200 import ics
    st = ics.uniform_box(10000, 0)
    nstep = 0
    for s, ref in _run(st, 10):
        nstep += 1
        assert np.array_equal(s.download("neighbor"), ref["neighbor"]), "step %d" % nstep
        assert s.scalar("dt") == pytest.approx(ref["dt"], rel=1e-13)
        assert s.scalar("ct") == pytest.approx(ref["ct"], rel=1e-13)
        assert s.scalar("d") == pytest.approx(ref["d"], rel=1e-12)
        assert s.scalar("age") == pytest.approx(ref["age"], rel=1e-13)
        for name in ("points", "velocities", "total_accel", "E_internal", "T", "sizes", "densities", "densities_0",
                     "num_densities", "pressure_accel", "visc_accel", "visc_heat"):
            assert rel(s.download(name), ref[name]) < RTOL, (nstep, name, rel(s.download(name), ref[name]))
    assert nstep == 10


def test_steps_with_dust(built_lib):
    """dust particles present (types 0 and 2 mixed in the neighbour rows)."""
    from sphb200 import ics
    st = ics.uniform_box(6000, 3, n_dust=30)
    for s, ref in _run(st, 3):
        assert np.array_equal(s.download("neighbor"), ref["neighbor"])
        assert np.array_equal(s.download("num_nontrivial"), ref["num_nontrivial"])
        for name in ("points", "velocities", "sizes", "T", "densities", "dust_densities"):
            assert rel(s.download(name), ref[name]) < RTOL, name
