"""Drop-in replacement for the hot path of the reference's physics library.

Same module name, function names, positional signatures and return tuples as
/root/reference/sph/navier_stokes_cleaned.py (imported as `nsc` by sph/code_running.py:12):
`neighbors` (:541), `hydro_update` (:556), `density` (:693), `dust_density` (:704), `num_dens`
(:744), `del_pressure` (:755), `crossing_time` (:776), `artificial_viscosity` (:788), and the
module globals `d`, `d_0`, `dt`, `dt_0` that the driver assigns from outside
(code_running.py:68-75,231,524) and that are read here at call time.

Arguments are numpy arrays owned by the caller and never mutated; results are fresh numpy arrays.
Every computation runs in libsphb200's sm_100a kernels through its C ABI (include/sphb200.h); torch
tensors only carry the bytes to and from the GPU.  There is no CPU fallback.

Deviations from the reference as written: the k-NN is exact (eps = 0, not the
traversal-order-dependent eps = 0.1 of :544), and the viscosity reduction of :649 sums over
neighbours (axis repaired) so that the block returns instead of raising.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .constants import (G, k, sb, AU, amu, h, c, wien, m_h, solar_mass, solar_luminosity, solar_lifespan,  # noqa: F401
                        t_cmb, t_solar, supernova_energy, m_0, year, species_labels, mu_specie, gamma, f_u,
                        mineral_densities, mrn_constants)
from . import constants as _C

# module globals assigned by the caller, exactly as with the reference module
d = _C.d_default
d_0 = _C.d_0_default
dt_0 = _C.dt_0
dt = _C.dt_0

_PAD = np.iinfo(np.int64).min      # padding sentinel for ragged neighbour rows (convert_list_kernel)


def _dev(a, dtype=np.float64):
    """host array -> contiguous device tensor (pinned staging for large arrays)."""
    a = np.ascontiguousarray(np.asarray(a), dtype=dtype)
    t = torch.from_numpy(a)
    return t.cuda(non_blocking=False)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _neighbor_array(neighbor, n):
    """(N,k) int64 array from the reference's neighbour structure (array or list of lists)."""
    try:
        nb = np.asarray(neighbor)
        if nb.dtype != object and nb.ndim == 2:
            return np.ascontiguousarray(nb, dtype=np.int64)
    except ValueError:
        pass
    rows = [np.asarray(r, dtype=np.int64).ravel() for r in neighbor]
    kmax = max([len(r) for r in rows] + [1])
    nb = np.full((len(rows), kmax), _PAD, dtype=np.int64)
    for i, r in enumerate(rows):
        nb[i, :len(r)] = r
    return nb


# The driver passes the SAME (N,k) neighbour array to six calls per step (code_running.py:451-458); at 320 B per
# particle its upload dominates each call.  The last uploaded array is kept on the device and reused when the
# caller passes the same buffer again with the same content fingerprint (address, shape and a strided checksum:
# an in-place edit of a sampled entry, or any new array, triggers a fresh upload).
_nb_cache = {"key": None, "dev": None}


def _neighbor_dev(neighbor, n):
    nb = _neighbor_array(neighbor, n)
    step = max(1, nb.size // 4096)
    flat = nb.reshape(-1)
    key = (nb.__array_interface__["data"][0], nb.shape, int(flat[::step].sum()), int(flat[-1]), int(flat[0]))
    if _nb_cache["key"] == key and _nb_cache["dev"] is not None and isinstance(neighbor, np.ndarray):
        return _nb_cache["dev"]
    dev = _dev(nb, np.int64)
    _nb_cache["key"], _nb_cache["dev"] = (key if isinstance(neighbor, np.ndarray) else None), dev
    return dev


def _check_bad_index(ctx, n):
    bad = ctx.counter("bad_index", _stream())
    if bad:
        raise IndexError("neighbour index out of bounds for axis 0 with size %d (%d entries)" % (n, bad))


def neighbors(points, dist, N_NEIGH):
    """nsc:541-552.  Returns (idx (N,k) int64, tree_handle, dists (N,k), nontriv (N,), new_sizes (N,)).
    Slot 1 (the cKDTree object in the reference, never used by live code) is None."""
    ctx = _lib.context()
    pts = _dev(points)
    if pts.dim() != 2 or pts.shape[1] != 3:
        raise ValueError("points must have shape (N, 3)")
    n = pts.shape[0]
    kk = int(N_NEIGH)
    idx = torch.empty((n, kk), dtype=torch.int64, device=pts.device)
    dd = torch.empty((n, kk), dtype=torch.float64, device=pts.device)
    nontriv = torch.empty((n,), dtype=torch.int64, device=pts.device)
    sizes = torch.empty((n,), dtype=torch.float64, device=pts.device)
    _lib.check(_lib.load().sphb200_neighbors(ctx.handle, _ptr(pts), n, kk, float(dist), _ptr(idx), _ptr(dd),
                                             _ptr(nontriv), _ptr(sizes), _stream()))
    over = ctx.counter("knn_overflow", _stream())
    if over:
        raise _lib.Sphb200Error("k-NN search did not converge for %d particles (coincident points?)" % over)
    return idx.cpu().numpy(), None, dd.cpu().numpy(), nontriv.cpu().numpy(), sizes.cpu().numpy()


def hydro_update(neighbor, points, mass, sizes, f_un, particle_type, T, mu_array, gamma_array, velocities):
    """nsc:556-671.  Returns (hydro_accel (N,3), visc_accel (N,3), visc_heat (N,), density (N,),
    num_density (N,), f_un_neighbor (S,N), dust_density (N,))."""
    ctx = _lib.context()
    nb = _neighbor_dev(neighbor, len(points))
    n, kk = nb.shape
    pts = _dev(points); vel = _dev(velocities)
    m = _dev(mass); hs = _dev(sizes); pt = _dev(particle_type); Tt = _dev(T); mu = _dev(mu_array); ga = _dev(gamma_array)
    fu = _dev(f_un) if f_un is not None else None
    S = int(fu.shape[1]) if fu is not None else 0
    dev = pts.device
    f64 = dict(dtype=torch.float64, device=dev)
    hydro_accel = torch.empty((n, 3), **f64); visc_accel = torch.empty((n, 3), **f64)
    visc_heat = torch.empty((n,), **f64); dens = torch.empty((n,), **f64); num = torch.empty((n,), **f64)
    dust = torch.empty((n,), **f64)
    fun = torch.empty((S, n), **f64) if S else None
    _lib.check(_lib.load().sphb200_hydro_update(
        ctx.handle, n, kk, _ptr(nb), _ptr(pts), _ptr(m), _ptr(hs), _ptr(fu), S, _ptr(pt), _ptr(Tt), _ptr(mu), _ptr(ga),
        _ptr(vel), _ptr(hydro_accel), _ptr(visc_accel), _ptr(visc_heat), _ptr(dens), _ptr(num), _ptr(fun), _ptr(dust),
        _stream()))
    _check_bad_index(ctx, n)
    fun_np = fun.cpu().numpy() if S else np.zeros((0, n))
    return (hydro_accel.cpu().numpy(), visc_accel.cpu().numpy(), visc_heat.cpu().numpy(), dens.cpu().numpy(),
            num.cpu().numpy(), fun_np, dust.cpu().numpy())


def _loop_density(mode, points, mass, neighbor, particle_type=None, sizes=None, mu_array=None):
    ctx = _lib.context()
    nb = _neighbor_dev(neighbor, len(points))
    n, kk = nb.shape
    pts = _dev(points); m = _dev(mass)
    pt = _dev(particle_type) if particle_type is not None else None
    hs = _dev(sizes) if sizes is not None else None
    mu = _dev(mu_array) if mu_array is not None else None
    out = torch.empty((n,), dtype=torch.float64, device=pts.device)
    _lib.check(_lib.load().sphb200_loop_density(ctx.handle, mode, n, kk, _ptr(nb), _ptr(pts), _ptr(m), _ptr(pt), _ptr(hs),
                                                _ptr(mu), float(d), _ptr(out), _stream()))
    _check_bad_index(ctx, n)
    return out.cpu().numpy()


def density(points, mass, particle_type, neighbor):
    """nsc:693-702 (reads module global `d`)."""
    return _loop_density(0, points, mass, neighbor, particle_type=particle_type)


def dust_density(points, mass, neighbor, particle_type, sizes):
    """nsc:704-717."""
    return _loop_density(1, points, mass, neighbor, particle_type=particle_type, sizes=sizes)


def num_dens(mass, points, mu_array, neighbor):
    """nsc:744-753 (reads module global `d`)."""
    return _loop_density(2, points, mass, neighbor, mu_array=mu_array)


def del_pressure(points, mass, particle_type, neighbor, E_internal, gamma_array):
    """nsc:755-774 (reads module global `d`)."""
    ctx = _lib.context()
    nb = _neighbor_dev(neighbor, len(points))
    n, kk = nb.shape
    pts = _dev(points); m = _dev(mass); pt = _dev(particle_type); E = _dev(E_internal); ga = _dev(gamma_array)
    out = torch.empty((n, 3), dtype=torch.float64, device=pts.device)
    _lib.check(_lib.load().sphb200_del_pressure(ctx.handle, n, kk, _ptr(nb), _ptr(pts), _ptr(m), _ptr(pt), _ptr(E), _ptr(ga),
                                                float(d), _ptr(out), _stream()))
    _check_bad_index(ctx, n)
    return out.cpu().numpy()


def crossing_time(neighbor, velocities, sizes, particle_type):
    """nsc:776-786 (reads module global `dt_0`)."""
    ctx = _lib.context()
    nb = _neighbor_dev(neighbor, len(velocities))
    n, kk = nb.shape
    vel = _dev(velocities); hs = _dev(sizes); pt = _dev(particle_type)
    out = ctypes.c_double(0.0)
    _lib.check(_lib.load().sphb200_crossing_time(ctx.handle, n, kk, _ptr(nb), _ptr(vel), _ptr(hs), _ptr(pt), float(dt_0),
                                                 ctypes.byref(out), _stream()))
    _check_bad_index(ctx, n)
    return float(out.value)


def artificial_viscosity(neighbor, points, particle_type, sizes, mass, densities, velocities, T, gamma_array, mu_array):
    """nsc:788-816 (reads module global `d`).  Returns (visc_accel (N,3), visc_heat (N,))."""
    ctx = _lib.context()
    nb = _neighbor_dev(neighbor, len(points))
    n, kk = nb.shape
    pts = _dev(points); pt = _dev(particle_type); hs = _dev(sizes); m = _dev(mass); de = _dev(densities)
    vel = _dev(velocities); Tt = _dev(T); ga = _dev(gamma_array); mu = _dev(mu_array)
    va = torch.empty((n, 3), dtype=torch.float64, device=pts.device)
    vh = torch.empty((n,), dtype=torch.float64, device=pts.device)
    _lib.check(_lib.load().sphb200_artificial_viscosity(
        ctx.handle, n, kk, _ptr(nb), _ptr(pts), _ptr(pt), _ptr(hs), _ptr(m), _ptr(de), _ptr(vel), _ptr(Tt), _ptr(ga),
        _ptr(mu), float(d), _ptr(va), _ptr(vh), _stream()))
    _check_bad_index(ctx, n)
    return va.cpu().numpy(), vh.cpu().numpy()


def grav_force_calculation_new(mass, points, sizes):
    """nsc:252-415.  The reference bins the cloud with a KD bisection and applies the Plummer-softened
    monopole law bin by bin with softening = median(sizes) (:359); its tree code cannot run on a current scipy.
    Here the same law is summed directly over all particles on the GPU (the exact limit of that approximation):
    a_i = G sum_j m_j (x_j - x_i) / (|x_j - x_i|^2 + eps^2)^(3/2).  Returns (N,3)."""
    ctx = _lib.context()
    pts = _dev(points)
    n = int(pts.shape[0])
    m = _dev(np.asarray(mass, dtype=np.float64))
    eps = float(np.median(np.asarray(sizes, dtype=np.float64)))
    out = torch.empty((n, 3), dtype=torch.float64, device=pts.device)
    _lib.check(_lib.load().sphb200_gravity(ctx.handle, n, _ptr(m), _ptr(pts), float(G), eps, _ptr(out), _stream()))
    return out.cpu().numpy()


def sigma_effective(mineral_densities, mrn_constants, mu_specie):
    """nsc:69-74: effective cross-section per dust particle for an MRN size distribution (host table)."""
    return mu_specie * amu / mineral_densities * (3. / 4.) * -np.diff(mrn_constants ** -0.5) / np.diff(mrn_constants ** 0.5)


def grain_mass(mineral_densities, mrn_constants):
    """nsc:76-79: mean grain mass per species for an MRN size distribution (host table)."""
    return mineral_densities * -np.diff(mrn_constants ** 0.5) / np.diff(mrn_constants ** -2.5) * (4. / 5.) * 4 * np.pi / 3.


def net_impulse(points, mass, sizes, velocities, particle_type, neighbor, f_un):
    """nsc:719-742, dust-gas drag.  Returns (accel_onto (N,3), accel_reaction (N,3))."""
    ctx = _lib.context()
    nb = _neighbor_dev(neighbor, len(points))
    n, kk = nb.shape
    pts = _dev(points); m = _dev(mass); hs = _dev(sizes); vel = _dev(velocities); pt = _dev(particle_type)
    f = _dev(np.ascontiguousarray(np.asarray(f_un, dtype=np.float64)))
    if f.dim() != 2 or f.shape[0] != n:
        raise ValueError("f_un must be (N, n_species)")
    ns = int(f.shape[1])
    meff = np.ascontiguousarray(np.broadcast_to(grain_mass(mineral_densities, mrn_constants), (ns,)), dtype=np.float64)
    seff = np.ascontiguousarray(np.broadcast_to(sigma_effective(mineral_densities, mrn_constants, mu_specie), (ns,)),
                                dtype=np.float64)
    onto = torch.empty((n, 3), dtype=torch.float64, device=pts.device)
    reaction = torch.empty((n, 3), dtype=torch.float64, device=pts.device)
    _lib.check(_lib.load().sphb200_net_impulse(
        ctx.handle, n, kk, _ptr(nb), _ptr(pts), _ptr(m), _ptr(hs), _ptr(vel), _ptr(pt), _ptr(f), ns,
        meff.ctypes.data_as(ctypes.c_void_p), seff.ctypes.data_as(ctypes.c_void_p), _ptr(onto), _ptr(reaction), _stream()))
    _check_bad_index(ctx, n)
    return onto.cpu().numpy(), reaction.cpu().numpy()
