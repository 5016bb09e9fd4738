"""GPU: SURVEY row F2, self-gravity (nsc:252-415; code_running.py:188,448,477) against the oracle's direct sum of the
softened monopole law the reference applies (PARITY UNPINNED: the reference routine cannot run, oracle/sph_oracle.py)."""
import os

import numpy as np
import pytest

from oracle import sph_oracle as O
from _util import rel, rel_elem

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_gravity_function_against_oracle_and_golden(built_lib):
    from sphb200 import ics
    from sphb200 import navier_stokes_cleaned as nsc
    g = np.load(os.path.join(GOLD, "gravity_n2000.npz"))
    st = ics.uniform_box(2000, 7)
    got = nsc.grav_force_calculation_new(st["mass"], st["points"], g["sizes"])
    assert got.shape == (2000, 3)
    assert rel(got, g["grav"]) < 1e-10 and rel_elem(got, g["grav"]) < 1e-10
    # a clustered cloud, larger N (several tiles per target), against the oracle computed here
    st = ics.clustered_cloud(6000, 3)
    h = O.neighbors(st["points"], st["dist"], 40)[4]
    want = O.gravity_direct(st["mass"], st["points"], h)
    got = nsc.grav_force_calculation_new(st["mass"], st["points"], h)
    assert rel(got, want) < 1e-10 and rel_elem(got, want) < 1e-10
    # momentum conservation of the pair law (size-independent property)
    assert np.max(np.abs(np.sum(got * st["mass"][:, None], axis=0))) < 1e-10 * np.sum(np.abs(got * st["mass"][:, None]))


def test_steps_with_gravity(built_lib):
    """Session.set_gravity: grav_accel enters total_accel (code_running.py:477) with the reference's softening, the
    median of the sizes the step's own search returned (nsc:359 after code_running.py:437)."""
    from sphb200 import ics
    from sphb200.session import Session
    st = ics.clustered_cloud(5000, 9)
    st["dt_0"] = O.dt_0 * 1e-6
    s = Session(5000, st["N_NEIGH"], st["dist"], dt_0=st["dt_0"])
    s.upload(st)
    s.set_gravity(True)
    s.startup()
    ref = O.startup(st)
    plain = O.startup(st)
    for it in range(3):
        s.step(); s.check()
        ref = O.step(ref, grav_fn=O.gravity_direct)
        plain = O.step(plain)
        assert np.array_equal(s.download("neighbor"), ref["neighbor"])
        assert rel(s.download("grav_accel"), ref["grav_accel"]) < 1e-10
        for name in ("points", "velocities", "total_accel", "E_internal", "T", "sizes"):
            assert rel(s.download(name), ref[name]) < 1e-10, (it, name)
    assert rel(ref["total_accel"], plain["total_accel"]) > 1e-6        # gravity really changes the step
    s.set_gravity(False)
    s.step(); s.check()
    ref = O.step(ref)
    assert rel(s.download("total_accel"), ref["total_accel"]) < 1e-10
