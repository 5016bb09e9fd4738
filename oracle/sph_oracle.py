"""TEST INFRASTRUCTURE — CPU oracle for the SPH particle time-step.  Not product code.

A numpy/scipy restatement of the hot path of /root/reference/sph/navier_stokes_cleaned.py
(Python 2, cannot be imported here) and of the inline integrator of
/root/reference/sph/code_running.py.  Only tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs may import this module; the product package never does.

Pinning: the reference holds no tests, fixtures or golden vectors, and the only
third-party arithmetic on the path is scipy.spatial.cKDTree (un-vendored, unpinned; scipy 1.18.1 in
this image).  This restatement is therefore pinned against OUTPUTS OF THE REFERENCE ITSELF: the
reference's text is executed by oracle/ref_extract.py in the build container and
tests/test_oracle_vs_reference.py asserts equality function by function;
tests/golden/make_golden.py commits the reference's outputs as fixtures under tests/golden/ which
tests/test_oracle_golden.py checks on any box.  The k-NN is additionally cross-checked against an
O(N^2) brute force (`neighbors_bruteforce`) so it does not rest on scipy alone.

Deviations from the reference as written: exact k-NN (eps=0 instead of the
traversal-order-dependent eps=0.1 of nsc:544), and the `axis=1`->`axis=0` repair of nsc:649.
"""
import numpy as np
from scipy import constants as _c
from scipy import spatial

# navier_stokes_cleaned.py:21-52
k = _c.Boltzmann
AU = _c.au
amu = _c.physical_constants['atomic mass constant'][0]
m_h = 1.0008 * amu
solar_mass = 1.989e30
m_0 = 10 ** 1.5 * solar_mass
year = 60. * 60. * 24. * 365.
dt_0 = year * 250000.


# --------------------------------------------------------------------------------------------
# A1  neighbour query                                            navier_stokes_cleaned.py:541-552
# --------------------------------------------------------------------------------------------
def neighbors(points, dist, N_NEIGH, eps=0.0, workers=1):
    kdt = spatial.cKDTree(points)
    dd, ii = kdt.query(points, N_NEIGH, eps, 2, dist, workers=workers)
    nontriv = np.sum(dd < np.inf, axis=1)
    dd[dd == np.inf] = 0
    new_sizes = np.max(dd, axis=1)
    return ii, kdt, dd, nontriv, new_sizes


def neighbors_bruteforce(points, dist, N_NEIGH, block=512):
    """O(N^2) exact k-NN with the same conventions (ascending distance, self first, missing
    neighbour -> index N / distance 0).  d^2 is the plain sum dx*dx + dy*dy + dz*dz."""
    n = len(points)
    idx = np.full((n, N_NEIGH), n, dtype=np.int64)
    dd = np.zeros((n, N_NEIGH))
    for a in range(0, n, block):
        b = min(n, a + block)
        dx = points[a:b, None, 0] - points[None, :, 0]
        dy = points[a:b, None, 1] - points[None, :, 1]
        dz = points[a:b, None, 2] - points[None, :, 2]
        d2 = dx * dx + dy * dy + dz * dz
        kk = min(N_NEIGH, n)
        part = np.argpartition(d2, kk - 1, axis=1)[:, :kk]
        pd2 = np.take_along_axis(d2, part, axis=1)
        order = np.argsort(pd2, axis=1, kind="stable")
        part = np.take_along_axis(part, order, axis=1)
        pd = np.sqrt(np.take_along_axis(pd2, order, axis=1))
        ok = pd < dist
        idx[a:b, :kk] = np.where(ok, part, n)
        dd[a:b, :kk] = np.where(ok, pd, 0.)
    nontriv = np.sum(idx < n, axis=1)
    return idx, None, dd, nontriv, np.max(dd, axis=1)


# --------------------------------------------------------------------------------------------
# A2-A6  vectorised hydro step                                   navier_stokes_cleaned.py:556-671
# --------------------------------------------------------------------------------------------
def _hydro_block(rows, neighbor, pts_2, mas_2, vels_2, sizes_2, temp_2, f_un_2, sizes, mass, T, mu_array,
                 gamma_array, particle_type, dens_2, Pi, want):
    """One row block of hydro_update.  `want` selects the pass:
       'dens'  -> density, dust density, number density, species densities, pressure sum  (:575-628)
       'pi'    -> per-particle viscosity Pi_i (needs neighbours' density)                  (:639-649)
       'visc'  -> visc_accel, visc_heat (needs neighbours' Pi)                             (:651-654)
    """
    nb = neighbor[rows]                                    # (n,k)
    self_ = nb[:, 0]
    delt_x = np.stack([pts_2[j][nb] - pts_2[j][self_][:, None] for j in range(3)])      # :580
    dist2 = np.sum(delt_x ** 2, axis=0)
    distances = dist2 ** (1. / 2.)                                                        # :586
    neigh_sizes = sizes_2[nb]                                                             # :587
    gas_n = (particle_type[nb] == 0)
    mu_neigh = mu_array[nb]
    t2_neigh = temp_2[nb]
    W6_grad_b = (-1. * 6. * 315. / (64. * np.pi * neigh_sizes ** 9) * (neigh_sizes ** 2 - distances ** 2) ** 2) * delt_x
    hs = sizes[rows][:, None]
    W6_grad_a = (-1. * 6. * 315. / (64. * np.pi * hs ** 9) * (hs ** 2 - distances ** 2) ** 2) * delt_x
    out = {}
    if want == 'dens':
        W6 = 315. / (64 * np.pi * neigh_sizes ** 9) * (neigh_sizes ** 2 - distances ** 2) ** 3      # :588
        W6[W6 < 0] = 0.                                                                               # :589
        out['density'] = np.sum(mas_2[nb] * gas_n * W6, axis=1)                                       # :605
        out['dust'] = np.sum(mas_2[nb] * (particle_type[nb] == 2) * W6, axis=1)                       # :606
        out['num'] = np.sum(mas_2[nb] * gas_n * W6 / amu / mu_neigh, axis=1)                          # :607
        a_self = (mass / mu_array / amu * k * T * (particle_type == 0))[rows][:, None]
        out['pgrad'] = -np.sum((mas_2[nb] / mu_neigh / amu * k * t2_neigh * gas_n * W6_grad_b
                                + a_self * W6_grad_a) / 2., axis=2)                                   # :615
        if f_un_2 is not None:
            base = mas_2[nb] / mu_neigh / amu * gas_n
            out['fun'] = np.stack([np.sum(base * f_un_2[s][nb] * W6, axis=1) for s in range(len(f_un_2))])  # :624-626
        return out
    delt_vels = np.stack([vels_2[j][nb] - vels_2[j][self_][:, None] for j in range(3)])  # :581
    if want == 'pi':
        w_ab = np.sum(delt_vels * delt_x, axis=0) / (dist2 + 0.01 * neigh_sizes ** 2) ** 0.5          # :643
        w_ab[w_ab > 0.] = 0.                                                                          # :644
        dens = dens_2[:-1]
        rho_avg = (dens_2[nb] + dens[rows][:, None]) / 2.                                             # :646
        c_self = ((gamma_array * k * T / (mu_array * amu) * (particle_type == 0)) ** (1. / 2.))[rows][:, None]
        c_sound = (1. / 2.) * ((gamma_array[nb] * k * t2_neigh / mu_neigh / amu * gas_n) ** (1. / 2.) + c_self)  # :647
        out['Pi'] = np.sum(-1. / 2. * (c_sound * 2 - 3 * w_ab) * w_ab / rho_avg, axis=1)              # :649 (repaired)
        return out
    bracket = (mas_2[nb] * Pi[nb] * gas_n * W6_grad_b
               + (mass * Pi * (particle_type == 0))[rows][:, None] * W6_grad_a) / 2.                   # :651
    out['visc_accel'] = -np.sum(bracket, axis=2)
    out['visc_heat'] = np.sum(bracket * delt_vels, axis=(0, 2)) * (mass[rows] / 2.)                   # :653-654
    return out


def hydro_update(neighbor, points, mass, sizes, f_un, particle_type, T, mu_array, gamma_array, velocities,
                 chunk=20000, species=True, threads=1):
    """Same signature and return tuple as navier_stokes_cleaned.py:556-671, evaluated on row blocks so
    the (3,N,k)/(15,N,k) temporaries stay small.  Every output row depends only on its own neighbour
    row plus gathered per-particle arrays, so chunking does not change results."""
    points = np.asarray(points).astype('float64')
    sizes = np.asarray(sizes).astype('float64')
    T = np.asarray(T).astype('float64')
    mu_array = np.asarray(mu_array).astype('float64')
    gamma_array = np.asarray(gamma_array).astype('float64')
    velocities = np.asarray(velocities).astype('float64')
    mass = np.asarray(mass).astype('float64')
    particle_type = np.asarray(particle_type)
    neighbor = np.asarray(neighbor)
    n = len(points)
    if neighbor.size and neighbor.max() >= n:
        # the reference indexes particle_type / mu_array (length N) with the neighbour array (:598-605)
        raise IndexError("index %d is out of bounds for axis 0 with size %d" % (int(neighbor.max()), n))
    pts_2 = np.append(points, [points[-1] * 1e10], axis=0).T                             # :568
    mas_2 = np.append(mass, 0.)
    vels_2 = np.append(velocities, [velocities[-1] * 0.], axis=0).T
    sizes_2 = np.append(sizes, 0.)
    temp_2 = np.append(T, 0)
    f_un_2 = None
    if species and f_un is not None:
        f_un = np.asarray(f_un).astype('float64')
        f_un_2 = np.append(f_un, np.array([f_un[-1]]), axis=0).T
    S = 0 if f_un_2 is None else len(f_un_2)
    density = np.zeros(n); dust = np.zeros(n); num = np.zeros(n)
    pgrad = np.zeros((3, n)); fun = np.zeros((S, n))
    args = (neighbor, pts_2, mas_2, vels_2, sizes_2, temp_2)
    rest = (sizes, mass, T, mu_array, gamma_array, particle_type)
    blocks = [slice(a, min(n, a + chunk)) for a in range(0, n, chunk)]

    def run(fn):
        """evaluate fn(rows) over the row blocks; `threads` > 1 only changes who runs a block
        (numpy releases the GIL in its large array ops), never a result."""
        if threads > 1 and len(blocks) > 1:
            from concurrent.futures import ThreadPoolExecutor
            with ThreadPoolExecutor(max_workers=threads) as ex:
                list(ex.map(fn, blocks))
        else:
            for rows in blocks:
                fn(rows)

    def p_dens(rows):
        o = _hydro_block(rows, *args, f_un_2, *rest, None, None, 'dens')
        density[rows] = o['density']; dust[rows] = o['dust']; num[rows] = o['num']; pgrad[:, rows] = o['pgrad']
        if S:
            fun[:, rows] = o['fun']
    run(p_dens)
    with np.errstate(divide='ignore', invalid='ignore'):
        hydro_accel = (pgrad / density).T                                                # :619,631
    dens_2 = np.append(density, 0)                                                       # :639
    Pi = np.zeros(n)
    visc_accel = np.zeros((n, 3)); visc_heat = np.zeros(n)

    def p_pi(rows):
        with np.errstate(divide='ignore', invalid='ignore'):
            Pi[rows] = _hydro_block(rows, *args, None, *rest, dens_2, None, 'pi')['Pi']

    def p_visc(rows):
        with np.errstate(divide='ignore', invalid='ignore'):
            o = _hydro_block(rows, *args, None, *rest, dens_2, Pi, 'visc')
        visc_accel[rows] = o['visc_accel'].T; visc_heat[rows] = o['visc_heat']
    run(p_pi)
    run(p_visc)
    return hydro_accel, visc_accel, visc_heat, density, num, fun, dust


# --------------------------------------------------------------------------------------------
# A7-A8  loop-generation kernels and sums                        navier_stokes_cleaned.py:673-816
# Restated row-vectorised over an (N,k) neighbour array; `d` is the module-global length scale the
# reference reads (nsc:700,715,765,804), passed explicitly here.
# --------------------------------------------------------------------------------------------
def _weigh2(norms_sq, m, d):                                                         # nsc:673-676
    return m * 315 * (m_0 / m) ** 3 * ((m / m_0) ** (2. / 3.) * d ** 2 - norms_sq) ** 3 / (64 * np.pi * d ** 9)


def _grad_weight(vec, norms_sq, m, d, gas):                                          # nsc:683-690
    W = -315 * 6 * (m_0 / m) ** 3 / (64 * np.pi * d ** 9) * ((m / m_0) ** (2. / 3.) * d ** 2 - norms_sq) ** 2 * gas
    W = W[..., None] * vec
    return np.nan_to_num(W.astype('float') * (((m / m_0) ** (2. / 3.) * d ** 2 - norms_sq) > 0)[..., None])


def _rows(points, neighbor):
    neighbor = np.asarray(neighbor)
    vec = points[neighbor] - points[:, None, :]
    return neighbor, vec, np.sum(vec ** 2, axis=2)


def density(points, mass, particle_type, neighbor, d):                              # nsc:693-702
    nb, _, n2 = _rows(points, neighbor)
    rho = _weigh2(n2, mass[nb], d) * (particle_type[nb] == 0)
    return np.sum(np.where(rho > 0, rho, 0.), axis=1)


def dust_density(points, mass, neighbor, particle_type, sizes, d):                  # nsc:704-717
    nb, _, n2 = _rows(points, neighbor)
    ds = sizes[nb]
    with np.errstate(divide='ignore', invalid='ignore'):
        rho = mass[nb] * 315 * (ds ** 2 - n2) ** 3 / (64 * np.pi * ds ** 9) * (particle_type[nb] == 2)  # nsc:678-681
    return np.sum(np.where(rho > 0, rho, 0.), axis=1)


def num_dens(mass, points, mu_array, neighbor, d):                                  # nsc:744-753
    nb, _, n2 = _rows(points, neighbor)
    nd = _weigh2(n2, mass[nb], d) / (mu_array[nb] * m_h)
    return np.sum(np.where(nd > 0, nd, 0.), axis=1)


def del_pressure(points, mass, particle_type, neighbor, E_internal, gamma_array, d):  # nsc:755-774
    nb, vec, n2 = _rows(points, neighbor)
    gw = _grad_weight(vec, n2, mass[nb], d, (particle_type[nb] == 0))
    base = 0.5 * gw * ((E_internal[nb] + E_internal[:, None]) / gamma_array[nb])[..., None]
    return np.sum(base, axis=1) * (particle_type == 0)[:, None]


def crossing_time(neighbor, velocities, sizes, particle_type, dt_0=dt_0):           # nsc:776-786
    neighbor = np.asarray(neighbor)
    max_relvel = np.max(np.sum((velocities[neighbor] - velocities[:, None, :]) ** 2, axis=2), axis=1) ** 0.5
    with np.errstate(divide='ignore', invalid='ignore'):
        ct = np.nan_to_num(sizes / max_relvel) * (particle_type == 0)
    if len(ct[ct != 0]) == 0:
        return dt_0 / 10.
    return min(ct[ct != 0]) + 0.0001


def artificial_viscosity(neighbor, points, particle_type, sizes, mass, densities, velocities, T, gamma_array,
                         mu_array, d):                                               # nsc:788-816
    nb, vec, n2 = _rows(points, neighbor)
    with np.errstate(divide='ignore', invalid='ignore'):
        css = np.nan_to_num((gamma_array * k * T / (mu_array * amu)) ** 0.5)
        dv = velocities[nb] - velocities[:, None, :]
        w = np.sum(dv * vec, axis=2) / n2 ** 0.5
        w[w > 0] = 0
        w = np.nan_to_num(w)
        vsig = css[nb] + css[:, None] - 3 * w
        rho = (densities[nb] + densities[:, None]) / 2.
        PI = -1. / 2. * vsig * w / rho
        gas_n = (particle_type[nb] == 0)
        gw = _grad_weight(vec, n2, mass[:, None] * np.ones_like(n2), d, gas_n)
        mm = (mass[nb] + mass[:, None]) / 2.
        accel = mm[..., None] * PI[..., None] * gw
        heat = 0.5 * mm * PI * np.sum(dv * gw, axis=2)
        act = (particle_type == 0) & (np.sum(gas_n, axis=1) > 0)
        va = np.sum(np.where(gas_n[..., None], accel, 0.), axis=1) * act[:, None]
        vh = np.sum(np.where(gas_n, heat, 0.), axis=1) * act
    return va, vh


# --------------------------------------------------------------------------------------------
# F1  dust-gas drag                                               navier_stokes_cleaned.py:44,52,69-79,719-742
mu_specie = np.array([2.0158, 4.0026, 1.0079, 1.0074, 4.0021, 4.0016, 0.0005, 140.69, 60.08, 12.0107, 28.0855,
                      55.834, 100.39, 131.93, 40.096])                               # nsc:41
mineral_densities = np.array([1.e19, 1e19, 1e19, 1e19, 1e19, 1e19, 1e19, 3320, 2260, 2266, 2329, 7870, 3250,
                              3250., 3166.])                                          # nsc:44
mrn_constants = np.array([50e-10, 5000e-10])                                          # nsc:52


def sigma_effective(mineral_densities=mineral_densities, mrn_constants=mrn_constants, mu_specie=mu_specie):  # nsc:69-74
    return mu_specie * amu / mineral_densities * (3. / 4.) * -np.diff(mrn_constants ** -0.5) / np.diff(mrn_constants ** 0.5)


def grain_mass(mineral_densities=mineral_densities, mrn_constants=mrn_constants):    # nsc:76-79
    return mineral_densities * -np.diff(mrn_constants ** 0.5) / np.diff(mrn_constants ** -2.5) * (4. / 5.) * 4 * np.pi / 3.


def net_impulse(points, mass, sizes, velocities, particle_type, neighbor, f_un, d=None):   # nsc:719-742
    """Drag of the dust neighbours on each particle (`accel_onto`) and the opposite push on those
    dust particles (`accel_reaction`, a scatter over the rows).  Rows are visited in order like the
    reference's loop, so the scatter adds in the same sequence."""
    meff, seff = grain_mass(), sigma_effective()
    nb, vec, n2 = _rows(points, neighbor)
    f = np.asarray(f_un, dtype=np.float64)
    onto = np.zeros(points.shape)
    reaction = np.zeros(points.shape)
    n = len(points)
    for a in range(0, n, 4096):
        b = min(n, a + 4096)
        r = nb[a:b]
        mg = np.sum(meff * f[r], axis=2)                                              # :727
        mc = np.sum(seff * f[r], axis=2)                                              # :728
        nv = velocities[r] - velocities[a:b, None, :]                                 # :729
        ds = sizes[r]
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            wf = mass[r] * 315 * (ds ** 2 - n2[a:b]) ** 3 / (64 * np.pi * ds ** 9)    # :736 (Weigh2_dust, :678-681)
            net = (wf / mg * mc * np.sum(nv ** 2, axis=2) ** 0.5)[..., None] * nv \
                * (particle_type[r] == 2)[..., None] * (wf > 0)[..., None]            # :738
        onto[a:b] = np.sum(net, axis=1)                                               # :740
        back = -(net * (r != np.arange(a, b)[:, None])[..., None])                    # :741
        for c in range(3):                                                            # row after row, as the loop does
            np.add.at(reaction[:, c], r.ravel(), back[..., c].ravel())
    return onto, reaction


# --------------------------------------------------------------------------------------------
# F2  self-gravity                                                navier_stokes_cleaned.py:252-415
# PARITY UNPINNED for this row: grav_force_calculation_new cannot run (it walks .less/.greater/.idx of scipy's
# removed pure-Python KDTree, :257-270) and the reference holds no output of it.  What the routine applies, bin
# by bin, is the Plummer-softened monopole law a = G M (x_c - x) / (|x_c - x|^2 + eps^2)^(3/2) with
# eps = median(sizes) (:359, :392-397, :411); this oracle sums that law directly over all particles, the limit
# the reference's KD-bisection bins approximate.
G_newton = _c.G


def gravity_direct(mass, points, sizes, block=1024):
    eps2 = float(np.median(sizes)) ** 2
    out = np.zeros(points.shape)
    for a in range(0, len(points), block):
        b = min(len(points), a + block)
        dx = points[None, :, :] - points[a:b, None, :]
        r2 = np.sum(dx * dx, axis=2) + eps2
        out[a:b] = G_newton * np.sum((mass[None, :] / r2 ** 1.5)[..., None] * dx, axis=1)
    return out


# --------------------------------------------------------------------------------------------
# A9-A12  one time step in the driver's order                    code_running.py:217-540
# Harness conventions: gas/dust only (no stars, so the radiative / supernova block
# :243-397 is skipped by its own guard), gravity, drag and chemistry zero, the start-up search radius
# of :171 instead of the bad in-loop radius of :386,437, `densities_0` seeded by the start-up
# hydro_update as the commented :183 would, and the vectorised generation (`hydro_update`) supplying
# pressure and viscosity.
# --------------------------------------------------------------------------------------------
RAISE_FACTOR = 715
N_INT_PER_PARTICLE = 3000.
MAX_AGE = 3e7 * year


def startup(state, neighbors_fn=neighbors, hydro_fn=hydro_update):
    """code_running.py:171-179: first neighbour query + hydro_update; seeds densities_0."""
    s = dict(state)
    nb, _, dd, nontriv, sizes = neighbors_fn(s['points'], s['dist'], s['N_NEIGH'])
    out = hydro_fn(nb, s['points'], s['mass'], sizes, s.get('f_un'), s['particle_type'], s['T'], s['mu_array'],
                   s['gamma_array'], s['velocities'])
    s.update(neighbor=nb, num_nontrivial=nontriv, sizes=sizes, densities=out[3], densities_0=out[3].copy(),
             age=0., dt=dt_0)
    return s


def append(s, new, neighbors_fn=neighbors):
    """Variable-N state: the concatenations of code_running.py:363-381 for a dict `new`
    of per-particle arrays, then the neighbour query of :386 (start-up radius, as in step()).
    `total_accel` and `densities_0` keep their OLD length, which is what makes the reference take
    dv = a dt (:484-485) and new_smoothing = 1 (:534-535) in the following step."""
    s = dict(s)
    for name in ('points', 'velocities', 'mass', 'sizes', 'T', 'E_internal', 'mu_array', 'gamma_array', 'particle_type'):
        s[name] = np.concatenate((s[name], np.asarray(new[name], dtype=np.float64)))
    if s.get('f_un') is not None and new.get('f_un') is not None:
        s['f_un'] = np.vstack([s['f_un'], new['f_un']])
    nb, _, dd, nontriv, sizes = neighbors_fn(s['points'], s['dist'], s['N_NEIGH'])                 # :386
    s.update(neighbor=nb, num_nontrivial=nontriv, sizes=sizes)
    return s


def step(s, neighbors_fn=neighbors, hydro_fn=hydro_update, crossing_fn=None, drag_fn=None, grav_fn=None):
    """One pass of the `while` body; returns a new state dict (inputs are not mutated).
    The hooks let tests/golden/make_golden.py drive the same harness with the reference's own
    (extracted) functions.  `drag_fn` (net_impulse's signature) switches the dust-gas drag of
    code_running.py:455,462-463,473 on; the state then needs `f_un`."""
    s = dict(s)
    points = s['points'].copy(); velocities = s['velocities'].copy()
    ptype = s['particle_type']; mass = s['mass']; mu = s['mu_array']; gam = s['gamma_array']
    dt0 = s.get('dt_0', dt_0)       # nsc.dt_0, a module global the caller sets (code_running.py:75)
    if crossing_fn is None:
        ct = crossing_time(s['neighbor'], velocities, s['sizes'], ptype, dt0)             # :222
    else:
        ct = crossing_fn(s['neighbor'], velocities, s['sizes'], ptype)
    dt = dt0 / 10 if s['age'] == 0 else max(dt0 / 5., min(dt0 * 2., ct))                  # :223-226
    if ct > MAX_AGE:
        dt = MAX_AGE / 100.                                                               # :228-229
    points[points > 1e11 * AU] = 1e11 * AU                                                # :233-236
    points[points < -1e11 * AU] = -1e11 * AU
    velocities[points > 1e11 * AU] = 0.1
    velocities[points < -1e11 * AU] = -0.1
    points = np.nan_to_num(points); velocities = np.nan_to_num(velocities)               # :237-238
    nb, _, dd, nontriv, sizes = neighbors_fn(points, s['dist'], s['N_NEIGH'])             # :437 (radius of :171)
    age = s['age'] + dt                                                                   # :447
    hydro_accel, av0, av1, densities, num_d, _, dust_d = hydro_fn(
        nb, points, mass, sizes, s.get('f_un'), ptype, s['T'], mu, gam, velocities)
    gas = (ptype == 0)
    pressure_accel = np.nan_to_num(hydro_accel * gas.astype('float')[:, None])            # :460
    old_accel = s['total_accel']                                                          # :471
    visc_accel = av0.copy()                                                               # :473 (drag = 0)
    if drag_fn is not None:
        onto, reaction = drag_fn(points, mass, sizes, velocities, ptype, nb, s['f_un'])  # :455
        with np.errstate(divide='ignore', invalid='ignore'):
            drag_gas = np.nan_to_num(((onto.T) * dust_d / densities * gas.astype('float')).T)   # :462
        visc_accel = drag_gas + np.nan_to_num(reaction) + av0                             # :463, :473
    vmag = np.sum(velocities ** 2, axis=1) ** 0.5
    with np.errstate(invalid='ignore'):
        clamp = vmag - np.sum(visc_accel ** 2, axis=1) ** 0.5 * dt < 0                    # :475
    visc_accel[clamp] = (-velocities / dt)[clamp]
    if grav_fn is not None:
        grav_accel = grav_fn(mass, points, sizes)                                         # :448 (sizes of the search at :437)
        total_accel = grav_accel + pressure_accel + visc_accel                            # :477
    else:
        grav_accel = None
        total_accel = pressure_accel + visc_accel                                         # :477 (grav = 0)
    points = points + ((total_accel * dt ** 2) / 2. + velocities * dt)                    # :481
    if np.shape(total_accel) == np.shape(old_accel):                                      # :482-486
        velocities = velocities + (total_accel + old_accel) / 2. * dt
    else:
        velocities = velocities + total_accel * dt
    E_internal = np.nan_to_num(s['E_internal']) + np.nan_to_num(av1 * dt)                 # :490
    with np.errstate(divide='ignore', invalid='ignore'):
        T = np.nan_to_num(E_internal * (mu * m_h) / (gam * mass * k))                     # :491
    vel_condition = np.sum(velocities ** 2, axis=1)                                       # :493
    slow = vel_condition < 80000 ** 2
    dist_sq = np.sum(points ** 2, axis=1)                                                 # :515
    max_dist = np.percentile(dist_sq[slow], 90)                                           # :518
    V_new = max_dist ** (3. / 2.)
    d = (V_new / len(points[slow]) / 0.9 * N_INT_PER_PARTICLE) ** (1. / 3.)               # :521
    count = np.bincount(np.asarray(nb).ravel(), minlength=len(points))[:len(points)]      # :527-530
    if len(s['densities_0']) == len(densities):                                           # :532
        with np.errstate(divide='ignore', invalid='ignore'):
            new_smoothing = (np.exp(np.nan_to_num(-(densities - s['densities_0']) / (3 * densities))) * 2) / 2. \
                + (count + nontriv - 1 < 2) * 0.1                                         # :533
    else:
        new_smoothing = 1.                                                                # :535
    sizes = sizes.copy()
    sizes[gas] = (new_smoothing * sizes)[gas]                                             # :536
    sizes[gas & (sizes > d)] = d                                                          # :537
    sizes[ptype == 2] = d / RAISE_FACTOR ** (1. / 3.)                                     # :538
    s.update(points=points, velocities=velocities, total_accel=total_accel, E_internal=E_internal, T=T,
             sizes=sizes, neighbor=nb, num_nontrivial=nontriv, densities=densities, densities_0=densities.copy(),
             d=float(d), age=age, dt=dt, ct=float(ct), num_densities=num_d, dust_densities=dust_d,
             pressure_accel=pressure_accel, visc_accel=visc_accel, visc_heat=av1, grav_accel=grav_accel)
    return s
