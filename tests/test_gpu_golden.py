"""GPU: the CUDA path (through the drop-in module / the C ABI) against the committed outputs of the
reference itself (tests/golden/*.npz; generated from the reference's own text by make_golden.py)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL = 1e-10     # north_star tolerance for fp64 sums


def rel(a, b):
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def test_functions_against_reference_outputs(built_lib):
    from sphb200 import ics
    from sphb200 import navier_stokes_cleaned as nsc
    g = np.load(os.path.join(GOLD, "functions_n1000.npz"))
    st = ics.uniform_box(1000, 0, n_dust=8)
    idx, tree, dd, nontriv, sizes = nsc.neighbors(st["points"], st["dist"], 40)
    assert np.array_equal(idx, g["idx"]) and np.array_equal(dd, g["dists"])          # bit-exact
    assert np.array_equal(nontriv, g["nontriv"]) and np.array_equal(sizes, g["new_sizes"])
    out = nsc.hydro_update(idx, st["points"], st["mass"], sizes, st["f_un"], st["particle_type"], st["T"],
                           st["mu_array"], st["gamma_array"], st["velocities"])
    for name, a in zip(("hydro_accel", "visc_accel", "visc_heat", "density", "num_density", "f_un_neighbor",
                        "dust_density"), out):
        assert rel(a, g["hu_" + name]) < RTOL, name
    nsc.d = float(g["loop_d"])
    p, m, pt = st["points"], st["mass"], st["particle_type"]
    assert rel(nsc.density(p, m, pt, idx), g["lg_density"]) < RTOL
    assert rel(nsc.dust_density(p, m, idx, pt, sizes), g["lg_dust_density"]) < RTOL
    assert rel(nsc.num_dens(m, p, st["mu_array"], idx), g["lg_num_dens"]) < RTOL
    assert rel(nsc.del_pressure(p, m, pt, idx, st["E_internal"], st["gamma_array"]), g["lg_del_pressure"]) < RTOL
    va, vh = nsc.artificial_viscosity(idx, p, pt, sizes, m, g["lg_density"], st["velocities"], st["T"],
                                      st["gamma_array"], st["mu_array"])
    assert rel(va, g["lg_visc_accel"]) < RTOL and rel(vh, g["lg_visc_heat"]) < RTOL
    from sphb200 import constants as C
    nsc.dt_0 = C.dt_0
    assert nsc.crossing_time(idx, st["velocities"], sizes, pt) == pytest.approx(float(g["crossing_time"]), rel=1e-13)


def test_ten_steps_against_reference_outputs(built_lib):
    """BASELINE config 1: N = 10^4, 10 steps, device-resident session vs the reference-driven harness."""
    from sphb200 import ics, constants as C
    from sphb200.session import Session
    g = np.load(os.path.join(GOLD, "steps_n10000.npz"))
    st = ics.uniform_box(10000, 0)
    s = Session(10000, 40, st["dist"], dt_0=C.dt_0 * 1e-6)
    s.upload(st)
    s.startup()
    assert np.array_equal(s.download("neighbor")[::50], g["startup_neighbor"])
    assert rel(s.download("densities")[::50], g["startup_densities"]) < RTOL
    w = np.arange(40)[None, :] + 1
    for it in range(10):
        s.step()
        s.check()
        nb = s.download("neighbor")
        assert int(np.sum(nb * w)) == int(g["neighbor_checksum"][it]), it
        for k in ("dt", "ct", "d", "age"):
            assert s.scalar(k) == pytest.approx(float(g["scal_" + k][it]), rel=1e-12), (it, k)
        if it in (0, 9):
            assert np.array_equal(nb[::50], g["step%d_neighbor" % (it + 1)])
            for f in ("points", "velocities", "total_accel", "E_internal", "T", "sizes", "densities", "num_densities",
                      "pressure_accel", "visc_accel", "visc_heat"):
                assert rel(s.download(f)[::50], g["step%d_%s" % (it + 1, f)]) < RTOL, (it, f)


@pytest.mark.parametrize("config,n,kind,seed", [(2, 1000000, "uniform", 1), (3, 4000000, "clustered", 2),
                                                (4, 10000000, "uniform", 3)])
def test_full_size_properties(built_lib, config, n, kind, seed):
    """BASELINE configs 2-4 at their full sizes (uniform 10^6, clustered 4x10^6, uniform 10^7): size-independent
    properties of the neighbour search and the sums.  Rows sorted by distance with self first, h = the row's
    largest distance, rows identical to a brute-force search for a random sample of targets (including the
    distance cut), density positive and equal to a host sum over the returned rows."""
    from sphb200 import ics
    from sphb200 import navier_stokes_cleaned as nsc
    st = ics.uniform_box(n, seed) if kind == "uniform" else ics.clustered_cloud(n, seed)
    p, dist = st["points"], st["dist"]
    idx, _, dd, nontriv, sizes = nsc.neighbors(p, dist, 40)
    assert idx.shape == (n, 40) and np.array_equal(idx[:, 0], np.arange(n))
    full = nontriv == 40
    assert full.mean() > 0.9 and nontriv.min() >= 1
    cols = np.arange(40)[None, :]
    valid = cols < nontriv[:, None]
    assert np.all(idx[~valid] == n) and np.all(dd[~valid] == 0.0)           # missing neighbours: index N, distance 0
    step = np.diff(dd, axis=1)
    assert np.all(step[valid[:, 1:]] >= 0)                                   # ascending distance inside each row
    del step
    assert np.array_equal(sizes, dd.max(axis=1)) and np.all(dd[valid] < dist)
    rs = np.random.RandomState(0)
    for i in rs.randint(0, n, 40):
        dx = p - p[i]
        d2 = dx[:, 0] * dx[:, 0] + dx[:, 1] * dx[:, 1] + dx[:, 2] * dx[:, 2]
        part = np.argpartition(d2, 41)[:41]
        order = part[np.argsort(d2[part], kind="stable")]
        m = int(nontriv[i])
        assert np.array_equal(order[:m], idx[i, :m]), i
        assert np.array_equal(np.sqrt(d2[order[:m]]), dd[i, :m])
        assert m == 40 or np.sqrt(d2[order[m]]) >= dist                      # the cut, not the search, ended the row
    out = nsc.hydro_update(idx, p, st["mass"], sizes, None, st["particle_type"], st["T"], st["mu_array"],
                           st["gamma_array"], st["velocities"])
    rho = out[3]
    sub = rs.randint(0, n, 2000)
    jj = np.where(valid[sub], idx[sub], 0)
    hj = sizes[jj]
    W = 315. / (64 * np.pi * hj ** 9) * (hj ** 2 - dd[sub] ** 2) ** 3
    W[(W < 0) | ~valid[sub]] = 0
    assert rel(rho[sub], np.sum(st["mass"][jj] * W, axis=1)) < RTOL
    assert np.all(rho > 0) and np.all(np.isfinite(out[0])) and np.all(np.isfinite(out[1]))


def test_net_impulse_against_reference_outputs(built_lib):
    """F1, dust-gas drag (nsc:719-742).  `accel_onto` is a gather (same tolerance as the other sums);
    `accel_reaction` is a scatter accumulated with fp64 atomics, so the order of the adds differs."""
    from sphb200 import ics
    from sphb200 import navier_stokes_cleaned as nsc
    g = np.load(os.path.join(GOLD, "drag_n2000.npz"))
    st = ics.uniform_box(2000, 3, n_dust=150)
    assert np.array_equal(nsc.grain_mass(nsc.mineral_densities, nsc.mrn_constants), g["grain_mass"])
    assert np.array_equal(nsc.sigma_effective(nsc.mineral_densities, nsc.mrn_constants, nsc.mu_specie), g["sigma_effective"])
    idx = g["idx"].astype(np.int64)
    onto, reaction = nsc.net_impulse(st["points"], st["mass"], g["sizes"], st["velocities"], st["particle_type"], idx,
                                     st["f_un"])
    assert rel(onto, g["accel_onto"]) < RTOL and rel(reaction, g["accel_reaction"]) < RTOL
    assert np.array_equal(onto == 0, g["accel_onto"] == 0) and np.array_equal(reaction == 0, g["accel_reaction"] == 0)
    # ragged rows (list of lists), as the loop generation allows
    rs = np.random.RandomState(4)
    ragged = [list(idx[i, : rs.randint(1, 41)]) for i in range(len(idx))]
    from oracle import sph_oracle as O
    got_o, got_r = nsc.net_impulse(st["points"], st["mass"], g["sizes"], st["velocities"], st["particle_type"], ragged,
                                   st["f_un"])
    pad = np.full(idx.shape, -1, dtype=np.int64)
    for i, row in enumerate(ragged):
        pad[i, :len(row)] = row
    keep = pad >= 0
    full_o, full_r = _drag_rows(O, st, g["sizes"], idx, keep)
    assert rel(got_o, full_o) < RTOL and rel(got_r, full_r) < RTOL


def _drag_rows(O, st, sizes, idx, keep):
    """nsc:719-742 restated with a per-entry mask (ragged rows)."""
    meff, seff = O.grain_mass(), O.sigma_effective()
    p, v, m, pt, f = st["points"], st["velocities"], st["mass"], st["particle_type"], st["f_un"]
    vec = p[idx] - p[:, None, :]
    n2 = np.sum(vec ** 2, axis=2)
    nv = v[idx] - v[:, None, :]
    ds = sizes[idx]
    with np.errstate(all="ignore"):
        wf = m[idx] * 315 * (ds ** 2 - n2) ** 3 / (64 * np.pi * ds ** 9)
        net = (wf / np.sum(meff * f[idx], axis=2) * np.sum(seff * f[idx], axis=2) * np.sum(nv ** 2, axis=2) ** 0.5)[..., None] \
            * nv * (pt[idx] == 2)[..., None] * (wf > 0)[..., None] * keep[..., None]
    onto = np.sum(net, axis=1)
    back = -(net * (idx != np.arange(len(p))[:, None])[..., None])
    reaction = np.zeros_like(onto)
    for c in range(3):
        np.add.at(reaction[:, c], idx.ravel(), back[..., c].ravel())
    return onto, reaction
