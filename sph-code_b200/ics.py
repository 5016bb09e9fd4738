"""Seeded initial conditions for the measured configurations.

The reference driver builds its particle state inline and unseeded
(/root/reference/sph/code_running.py:114-177).  These generators reproduce that recipe with a
legacy `RandomState(seed)` and the driver's draw order (positions `rand(N,3)`, temperature
`rand(N)`, velocities `normal(N,3)`), at the driver's default number density, so the reference
functions and the CUDA path can be fed identical arrays.
"""
import numpy as np

from . import constants as C


def _composition(n, particle_type):
    """f_un, mu_array, gamma_array as code_running.py:149-164 builds them (gas rows = f_u)."""
    fgas = np.array([C.f_u] * n).T * (particle_type == 0)
    dust_base = np.array([.0, .0, 0., 0., 0., 0., 0., 0.025, 0.025, 0.025, 0.025, 0.025, 0.025, 0.025, 0.025])
    dust_base = dust_base / np.sum(dust_base)
    fdust = np.array([dust_base] * n).T * (particle_type == 2)
    f_un = (fgas + fdust).T
    mu_array = np.sum(f_un * C.mu_specie, axis=1) / np.sum(f_un, axis=1)
    gamma_array = np.sum(f_un * C.gamma, axis=1) / np.sum(f_un, axis=1)
    return f_un, mu_array, gamma_array


def _finish(points, rs, mass, particle_type, d, dist, extra):
    n = len(points)
    T = 10 * (np.ones([n]) + rs.rand(n))                       # code_running.py:145
    velocities = rs.normal(size=(n, 3)) * 1000.                  # code_running.py:177
    velocities[particle_type == 2] *= 0.
    f_un, mu_array, gamma_array = _composition(n, particle_type)
    sizes = (mass / C.m_0) ** (1. / 3.) * d                      # code_running.py:139
    E_internal = gamma_array * mass * C.k * T / (mu_array * C.m_h)  # code_running.py:166
    state = dict(points=points, velocities=velocities, total_accel=np.zeros((n, 3)), mass=mass, sizes=sizes,
                 T=T, E_internal=E_internal, mu_array=mu_array, gamma_array=gamma_array,
                 particle_type=particle_type, f_un=f_un, d=float(d), dist=float(dist), N_NEIGH=C.N_NEIGH)
    state.update(extra)
    return state


def uniform_box(n, seed, n_dust=0):
    """Uniform random cube at the default number density (configs 1, 2 and 4).

    Cube side L = DIAMETER * (n / 2 145 000)^(1/3); equal gas masses 0.4 Msun / 715
    (code_running.py:116-124); optional dust particles appended as code_running.py:126-130 does.
    """
    rs = np.random.RandomState(seed)
    ntot = n + n_dust
    side = C.DIAMETER * (float(n) / C.N_PARTICLES_DEFAULT) ** (1. / 3.)
    points = (rs.rand(ntot, 3) - 0.5) * side                     # code_running.py:132
    mass = np.full(ntot, 0.4 * C.solar_mass / C.raise_factor)
    particle_type = np.zeros(ntot)
    if n_dust:
        DUST_MASS = 0.05000000001 / C.raise_factor
        mass[n:] = DUST_MASS * C.solar_mass
        particle_type[n:] = 2
    d = C.d_default
    dist = (0.4 * C.solar_mass / C.m_0) ** (1. / 3.) * d         # code_running.py:171
    return _finish(points, rs, mass, particle_type, d, dist, dict(side=float(side), kind="uniform"))


def clustered_cloud(n, seed):
    """Clustered cloud (configs 3 and 5): radial profile of imf/lane_emden_modified.py:317-321.

    Shell radii r = sqrt(linspace(0, 25e14, 2001)) AU, rho(r) ~ 1/(1 + (r / 2e7 AU)^4), n0 = 1e6 m^-3,
    composition 0.86/0.14.  Radii are drawn by inverse CDF of M(<r) = sum 4 pi r^2 rho dr, directions
    isotropic, equal masses M_tot / n.  Centre-to-edge density contrast 1 + 2.5^4 ~ 40.
    """
    rs = np.random.RandomState(seed)
    r = np.linspace(0, 25e14, 2001) ** 0.5 * C.AU
    dr = np.diff(r)
    dr = np.append(dr[0], dr)
    rho_shape = (r ** 4 / (C.AU * 2e7) ** 4 + 1) ** -1
    mu = 0.86 * C.mu_specie[0] + 0.14 * C.mu_specie[1]
    rho = 1e6 * mu * C.amu * rho_shape
    m_cum = np.cumsum(4 * np.pi * r ** 2 * rho * dr)
    m_tot = m_cum[-1]
    u = rs.rand(n)
    radii = np.interp(u * m_tot, m_cum, r)
    dirs = rs.normal(size=(n, 3))
    dirs /= np.sqrt(np.sum(dirs ** 2, axis=1))[:, None]
    points = dirs * radii[:, None]
    mass = np.full(n, m_tot / n)
    particle_type = np.zeros(n)
    # global length scale as the driver's adaptive formula would give it (code_running.py:515-521)
    dist_sq = np.sum(points ** 2, axis=1)
    d = (np.percentile(dist_sq, 90) ** 1.5 / n / 0.9 * C.N_INT_PER_PARTICLE) ** (1. / 3.)
    dist = 2. * r[-1]          # no radius cut: every row holds k real neighbours
    return _finish(points, rs, mass, particle_type, d, dist, dict(side=float(2 * r[-1]), kind="clustered"))


def clustered_cloud_chunk(n_total, seed, chunk_index, chunk_size):
    """Particles [chunk_index * chunk_size, ...) of a clustered cloud of n_total equal-mass particles, drawn with
    their own RandomState(seed, chunk_index): any subset of chunks can be generated independently (BASELINE
    configs[4]: 10^8 particles over 8 GPUs, no rank ever holds them all).  Same radial profile as
    clustered_cloud; `gid` = global particle index; the length scale d is the analytic 90th-percentile radius
    of the profile (the full-sample percentile clustered_cloud uses is not available to a single chunk)."""
    first = chunk_index * chunk_size
    n = int(min(chunk_size, n_total - first))
    rs = np.random.RandomState([seed, chunk_index])
    r = np.linspace(0, 25e14, 2001) ** 0.5 * C.AU
    dr = np.diff(r)
    dr = np.append(dr[0], dr)
    rho_shape = (r ** 4 / (C.AU * 2e7) ** 4 + 1) ** -1
    mu = 0.86 * C.mu_specie[0] + 0.14 * C.mu_specie[1]
    rho = 1e6 * mu * C.amu * rho_shape
    m_cum = np.cumsum(4 * np.pi * r ** 2 * rho * dr)
    m_tot = m_cum[-1]
    radii = np.interp(rs.rand(n) * m_tot, m_cum, r)
    dirs = rs.normal(size=(n, 3))
    dirs /= np.sqrt(np.sum(dirs ** 2, axis=1))[:, None]
    points = dirs * radii[:, None]
    mass = np.full(n, m_tot / n_total)
    r90 = float(np.interp(0.9 * m_tot, m_cum, r))
    d = ((r90 ** 2) ** 1.5 / n_total / 0.9 * C.N_INT_PER_PARTICLE) ** (1. / 3.)
    st = _finish(points, rs, mass, np.zeros(n), d, 2. * r[-1], dict(side=float(2 * r[-1]), kind="clustered"))
    st["gid"] = np.arange(first, first + n)
    return st


def perturbation(ids, amplitude, seed):
    """Host mirror of sphb200_session_perturb: the (len(ids), 3) displacement amplitude * u(id, seed), u uniform
    in [-1,1)^3 from a splitmix64 hash of (3 id + component, seed).  Bit-identical to the device kernel."""
    ids = np.asarray(ids, dtype=np.uint64)
    out = np.empty((len(ids), 3))
    with np.errstate(over="ignore"):
        for c in range(3):
            z = (ids * np.uint64(3) + np.uint64(c + 1)) * np.uint64(0x9E3779B97F4A7C15) \
                + np.uint64(seed) * np.uint64(0xBF58476D1CE4E5B9)
            z ^= z >> np.uint64(30); z *= np.uint64(0xBF58476D1CE4E5B9)
            z ^= z >> np.uint64(27); z *= np.uint64(0x94D049BB133111EB)
            z ^= z >> np.uint64(31)
            out[:, c] = (z >> np.uint64(11)).astype(np.float64) * (1.0 / 4503599627370496.0) - 1.0
    return amplitude * out
