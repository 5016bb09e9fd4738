"""GPU parity tests: the CUDA path (through the drop-in module, i.e. through the C ABI) against the
CPU oracle on the same seeded inputs.  Bit-exact for neighbour indices; 1e-10 relative (north_star)
for floating-point sums, measured against the largest magnitude of the quantity so that components
which cancel to ~0 are not compared against themselves."""
import numpy as np
import pytest

from oracle import sph_oracle as O

pytestmark = pytest.mark.gpu

RTOL = 1e-10     # north_star: density/acceleration within 1e-10 relative in fp64


@pytest.fixture(scope="module")
def nsc(built_lib):
    from sphb200 import navier_stokes_cleaned as m
    return m


def rel(a, b):
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def rel_rows(a, b):
    """per-particle vector error relative to that particle's vector norm (worst particle)."""
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    nb = np.sqrt(np.sum(b * b, axis=-1)) if b.ndim > 1 else np.abs(b)
    na = np.sqrt(np.sum((a - b) ** 2, axis=-1)) if b.ndim > 1 else np.abs(a - b)
    return float(np.max(na / np.maximum(nb, 1e-300 + 1e-6 * np.max(nb))))


def _state(n, seed, n_dust=0):
    from sphb200 import ics
    return ics.uniform_box(n, seed, n_dust=n_dust)


@pytest.mark.parametrize("n,seed", [(64, 5), (1000, 0), (10000, 0), (200000, 1)])
def test_neighbors_bit_exact(nsc, n, seed):
    st = _state(n, seed)
    k = min(40, n)
    gi, gt, gd, gn, gs = nsc.neighbors(st["points"], st["dist"], k)
    oi, _, od, on, os_ = O.neighbors(st["points"], st["dist"], k)
    assert gt is None
    assert gi.dtype == np.int64 and gi.shape == (n, k)
    assert np.array_equal(gi, oi)                      # index sets AND ascending-distance order
    assert np.array_equal(gn, on)
    assert np.array_equal(gd, od)                      # same d^2 arithmetic + IEEE sqrt => identical bits
    assert np.array_equal(gs, os_)


def test_neighbors_vs_bruteforce(nsc):
    st = _state(3000, 7)
    gi, _, gd, _, _ = nsc.neighbors(st["points"], st["dist"], 40)
    bi, _, bd, _, _ = O.neighbors_bruteforce(st["points"], st["dist"], 40)
    assert np.array_equal(gi, bi)
    assert np.array_equal(gd, bd)


def test_neighbors_radius_cut(nsc):
    """distance_upper_bound: missing neighbours are index N / distance 0 (nsc:545-548)."""
    st = _state(5000, 3)
    _, _, od_full, _, hs = O.neighbors(st["points"], st["dist"], 40)
    cut = float(np.median(hs))                         # about half the rows are truncated
    gi, _, gd, gn, gs = nsc.neighbors(st["points"], cut, 40)
    oi, _, od, on, os_ = O.neighbors(st["points"], cut, 40)
    assert (on < 40).any() and (on == 40).any()
    assert np.array_equal(gi, oi) and np.array_equal(gd, od) and np.array_equal(gn, on) and np.array_equal(gs, os_)


def test_neighbors_k_larger_than_n(nsc):
    st = _state(25, 11)
    gi, _, gd, gn, gs = nsc.neighbors(st["points"], st["dist"], 40)
    oi, _, od, on, os_ = O.neighbors(st["points"], st["dist"], 40)
    assert np.array_equal(gi, oi) and np.array_equal(gd, od) and np.array_equal(gn, on) and np.array_equal(gs, os_)


def test_neighbors_clustered(nsc):
    from sphb200 import ics
    st = ics.clustered_cloud(60000, 2)
    gi, _, gd, gn, gs = nsc.neighbors(st["points"], st["dist"], 40)
    oi, _, od, on, os_ = O.neighbors(st["points"], st["dist"], 40)
    assert np.array_equal(gi, oi) and np.array_equal(gd, od) and np.array_equal(gs, os_)


def test_neighbors_group_sizes(nsc):
    """the k-NN kernel gives identical rows for every lanes-per-target setting."""
    from sphb200 import _lib
    st = _state(20000, 4)
    oi = O.neighbors(st["points"], st["dist"], 40)[0]
    ctx = _lib.context()
    try:
        for g in (8, 16, 32):
            ctx.set_option("knn_group", g)
            assert np.array_equal(nsc.neighbors(st["points"], st["dist"], 40)[0], oi)
    finally:
        ctx.set_option("knn_group", 32)


@pytest.mark.parametrize("n,seed,n_dust", [(2000, 0, 0), (10000, 0, 0), (10000, 1, 50)])
def test_hydro_update(nsc, n, seed, n_dust):
    st = _state(n, seed, n_dust)
    nb, _, _, _, sizes = O.neighbors(st["points"], st["dist"], 40)
    args = (nb, st["points"], st["mass"], sizes, st["f_un"], st["particle_type"], st["T"], st["mu_array"],
            st["gamma_array"], st["velocities"])
    g = nsc.hydro_update(*args)
    o = O.hydro_update(*args)
    names = ["hydro_accel", "visc_accel", "visc_heat", "density", "num_density", "f_un_neighbor", "dust_density"]
    for name, a, b in zip(names, g, o):
        assert a.shape == b.shape, name
        if name in ("density", "num_density"):
            assert np.max(np.abs(a - b) / np.abs(b)) < RTOL, name        # elementwise relative
        assert rel(a, b) < RTOL, (name, rel(a, b))
    assert rel_rows(g[0], o[0]) < 1e-9
    assert rel_rows(g[1], o[1]) < 1e-9


def test_hydro_update_arbitrary_rows(nsc):
    """rows need not come from a k-NN search: column 0 is 'self' whatever it holds (nsc:580-581)."""
    st = _state(3000, 9)
    rs = np.random.RandomState(0)
    nb = rs.randint(0, 3000, size=(3000, 17))
    sizes = np.full(3000, 3e15) * (1 + 0.2 * rs.rand(3000))
    args = (nb, st["points"], st["mass"], sizes, st["f_un"], st["particle_type"], st["T"], st["mu_array"],
            st["gamma_array"], st["velocities"])
    g = nsc.hydro_update(*args)
    o = O.hydro_update(*args)
    for a, b in zip(g, o):
        assert rel(a, b) < RTOL


def test_hydro_update_bad_index(nsc):
    st = _state(500, 2)
    nb, _, _, _, sizes = O.neighbors(st["points"], st["dist"], 40)
    nb = nb.copy(); nb[3, 5] = 500                      # the reference's "missing neighbour" index
    with pytest.raises(IndexError):
        nsc.hydro_update(nb, st["points"], st["mass"], sizes, st["f_un"], st["particle_type"], st["T"],
                         st["mu_array"], st["gamma_array"], st["velocities"])


def test_loop_generation(nsc):
    st = _state(4000, 0, n_dust=40)
    nb, _, _, _, sizes = O.neighbors(st["points"], st["dist"], 40)
    d = float(np.median(sizes) / (st["mass"][0] / O.m_0) ** (1 / 3.) * 0.9)
    nsc.d = d
    p, m, pt = st["points"], st["mass"], st["particle_type"]
    E, g, mu, T, v = st["E_internal"], st["gamma_array"], st["mu_array"], st["T"], st["velocities"]
    rho = O.density(p, m, pt, nb, d)
    assert rel(nsc.density(p, m, pt, nb), rho) < RTOL
    assert rel(nsc.dust_density(p, m, nb, pt, sizes), O.dust_density(p, m, nb, pt, sizes, d)) < RTOL
    assert rel(nsc.num_dens(m, p, mu, nb), O.num_dens(m, p, mu, nb, d)) < RTOL
    assert rel(nsc.del_pressure(p, m, pt, nb, E, g), O.del_pressure(p, m, pt, nb, E, g, d)) < RTOL
    ga, gh = nsc.artificial_viscosity(nb, p, pt, sizes, m, rho, v, T, g, mu)
    oa, oh = O.artificial_viscosity(nb, p, pt, sizes, m, rho, v, T, g, mu, d)
    assert rel(ga, oa) < RTOL and rel(gh, oh) < RTOL
    nsc.dt_0 = O.dt_0
    assert nsc.crossing_time(nb, v, sizes, pt) == pytest.approx(O.crossing_time(nb, v, sizes, pt), rel=1e-14)


def test_loop_generation_ragged(nsc):
    """the loop generation indexes neighbor[j] row by row, so ragged lists of lists are legal."""
    st = _state(600, 1)
    nb = O.neighbors(st["points"], st["dist"], 40)[0]
    rs = np.random.RandomState(1)
    ragged = [list(nb[i, : rs.randint(1, 41)]) for i in range(600)]
    d = 2e17
    nsc.d = d
    got = nsc.density(st["points"], st["mass"], st["particle_type"], ragged)
    want = np.zeros(600)
    for i in range(600):
        row = np.array(ragged[i])
        x = st["points"][row] - st["points"][i]
        w = O._weigh2(np.sum(x ** 2, axis=1), st["mass"][row], d) * (st["particle_type"][row] == 0)
        want[i] = np.sum(w[w > 0])
    assert rel(got, want) < RTOL


@pytest.mark.parametrize("k", [1, 2, 50, 56, 57, 100])
def test_neighbors_other_k(nsc, k):
    """BASELINE config 2 mentions '~50 neighbours': k = 50 is the labelled variant; 56 is the widest row
    the cell-tiled kernel takes, 57 and 100 go to the general kernel; 1 and 2 are the degenerate ends."""
    st = _state(30000, 21)
    gi, _, gd, gn, gs = nsc.neighbors(st["points"], st["dist"] * 1.5, k)
    if k == 1:      # cKDTree squeezes the k axis away and nsc:546 raises AxisError; the brute force keeps (N,1)
        oi, _, od, on, os_ = O.neighbors_bruteforce(st["points"], st["dist"] * 1.5, k)
    else:
        oi, _, od, on, os_ = O.neighbors(st["points"], st["dist"] * 1.5, k)
    assert gi.shape == (30000, k)
    assert np.array_equal(gi, oi) and np.array_equal(gd, od)
    assert np.array_equal(gn, on) and np.array_equal(gs, os_)


def test_neighbors_degenerate_inputs(nsc):
    """one particle; particles on a line (a grid that is one cell thick in two directions); all particles
    on one point."""
    p1 = np.array([[1.0, 2.0, 3.0]])
    gi, _, gd, gn, gs = nsc.neighbors(p1, 10.0, 4)
    assert gi.tolist() == [[0, 1, 1, 1]] and gd.tolist() == [[0.0, 0.0, 0.0, 0.0]] and gn.tolist() == [1] and gs.tolist() == [0.0]
    rs = np.random.RandomState(3)
    line = np.zeros((5000, 3)); line[:, 0] = rs.rand(5000) * 1e17
    gi, _, gd, gn, gs = nsc.neighbors(line, np.inf, 40)
    oi, _, od, on, os_ = O.neighbors(line, np.inf, 40)
    assert np.array_equal(gi, oi) and np.array_equal(gd, od) and np.array_equal(gs, os_)
    same = np.ones((300, 3)) * 7.0                 # every distance ties: any 40 distinct indices are a right answer
    gi, _, gd, gn, gs = nsc.neighbors(same, 1.0, 40)
    assert np.all(gd == 0.0) and np.all(gn == 40) and np.all(gs == 0.0)
    assert gi.min() >= 0 and gi.max() < 300 and all(len(set(r)) == 40 for r in gi.tolist())


def test_empty_input_is_an_error(nsc):
    """cKDTree.query on zero points returns empty arrays; the C ABI refuses N = 0 instead (documented)."""
    from sphb200._lib import Sphb200Error
    with pytest.raises((Sphb200Error, ValueError, IndexError)):
        nsc.neighbors(np.zeros((0, 3)), 1.0, 4)
