"""The reference's on-disk key/value files (SURVEY.md §8 rows "config" and F3), host side only.

The galaxy-cycle loop of the reference passes state between runs through `.npz` files:
  ../savefiles/config/config_<n>.npz    written by sph/config_bootstrap.py:33 and sph/config_helper.py:284,
                                        read by config_helper.py:58-76 and by the (commented) loader at
                                        sph/code_running.py:87-108;
  ../savefiles/outputs/<n>.npz          written by sph/code_running.py:656-670, merged by config_helper.py:99-131.
This module reads and writes the same keys with the same shapes, so those scripts can consume runs of
the B200 stepper unchanged.  The AGB yield arithmetic of config_helper.py is out of scope; only the
schema and the file-numbering convention (largest <n> wins, next file is <n>+1) are mirrored.
"""
import os
import re

import numpy as np

from . import constants as C

CONFIG_KEYS = ("specie_fraction_array", "dust_base_frac", "DUST_FRAC", "OVERALL_AGE", "overall_AGB_list",
               "overall_AGB_time_until", "overall_AGB_metallicity", "overall_AGB_composition", "timestep_time_coord",
               "timestep_dust_temps", "timestep_star_frac", "timestep_imf_measure", "timestep_chems_error",
               "timestep_sup_error")                                   # config_bootstrap.py:33 / config_helper.py:284-300
OUTPUT_KEYS = ("gas_mass_by_species", "star_mass_by_species", "dust_mass_by_species", "AGB_condition", "AGB_list",
               "AGB_time_until", "AGB_metallicity", "AGB_composition", "time_coord", "dust_temps", "star_frac",
               "imf_measure", "chems_error", "sup_error")              # code_running.py:670

_CONF_RE = re.compile(r"^config_(-?\d+)\.npz$")
_OUT_RE = re.compile(r"^(-?\d+)\.npz$")


def bootstrap_config():
    """The values config_bootstrap.py:16-31 writes into config_0.npz (14 species, as that script has it)."""
    gas = np.array([0.86, 0.14, 0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 0.])
    dust = np.array([.0, .0, 0., 0., 0., 0., 0.025, 0.025, 0.025, 0.025, 0.025, 0.025, 0.025, 0.025])
    dust = dust / np.sum(dust)
    far = 1e20 * 60 * 60 * 24 * 365
    return dict(specie_fraction_array=gas, dust_base_frac=dust, DUST_FRAC=0.000000001, OVERALL_AGE=0,
                overall_AGB_list=np.array([1.1, 1.1]), overall_AGB_time_until=np.array([far, far]),
                overall_AGB_metallicity=np.array([0.001, 0.001]), overall_AGB_composition=np.vstack([gas] * 2),
                timestep_time_coord=np.array([0.]), timestep_dust_temps=np.array([0.]),
                timestep_star_frac=np.array([0.]), timestep_imf_measure=np.array([1.]),
                timestep_chems_error=np.array([0.]), timestep_sup_error=np.array([0.]))


def _numbers(directory, regex):
    out = []
    if os.path.isdir(directory):
        for name in os.listdir(directory):
            m = regex.match(name)
            if m:
                out.append(int(m.group(1)))
    return out


def latest_config_number(config_dir):
    """Largest <n> among config_<n>.npz, or -1 when there is none (config_helper.py:58-63)."""
    return max(_numbers(config_dir, _CONF_RE) + [-1])


def save_config(config_dir, number=None, **values):
    """Write config_<number>.npz (default: latest + 1).  All CONFIG_KEYS are required, as np.savez is
    called with all of them in the reference."""
    missing = [k for k in CONFIG_KEYS if k not in values]
    if missing:
        raise KeyError("config file needs the keys %s" % ", ".join(missing))
    if number is None:
        number = latest_config_number(config_dir) + 1
    os.makedirs(config_dir, exist_ok=True)
    path = os.path.join(config_dir, "config_%d" % int(number))
    np.savez(path, **{k: values[k] for k in CONFIG_KEYS})
    return path + ".npz"


def load_config(config_dir, number=None):
    """Read config_<number>.npz (default: the latest) into a dict of arrays (config_helper.py:65-76)."""
    if number is None:
        number = latest_config_number(config_dir)
    if number < 0:
        raise IOError("no config_<n>.npz under %s (run the bootstrap first)" % config_dir)
    with np.load(os.path.join(config_dir, "config_%d.npz" % int(number))) as fh:
        missing = [k for k in CONFIG_KEYS if k not in fh.files]
        if missing:
            raise KeyError("config_%d.npz lacks the keys %s" % (number, ", ".join(missing)))
        return {k: fh[k] for k in CONFIG_KEYS}


def composition_from_config(cfg, n_species=C.N_SPECIES):
    """What the driver's (commented) loader would take from a config file (code_running.py:104-108):
    gas composition, dust composition, dust fraction and galaxy age.  The config scripts carry 14
    species, the drivers 15; shorter arrays are zero-padded to `n_species`."""
    def pad(a):
        a = np.asarray(a, dtype=np.float64)
        return np.concatenate([a, np.zeros(max(0, n_species - len(a)))])[:n_species]
    return dict(specie_fraction_array=pad(cfg["specie_fraction_array"]), dust_base_frac=pad(cfg["dust_base_frac"]),
                DUST_FRAC=float(cfg["DUST_FRAC"]), OVERALL_AGE=float(cfg["OVERALL_AGE"]))


def latest_output_number(outputs_dir):
    return max(_numbers(outputs_dir, _OUT_RE) + [-1])


def species_masses(f_un, mass, particle_type, which, mu_specie=C.mu_specie):
    """Mass per species of one particle type (code_running.py:656-658)."""
    sel = np.asarray(particle_type) == which
    w = (np.asarray(f_un, dtype=np.float64) * mu_specie)[sel]
    if len(w) == 0:
        return np.zeros(np.shape(f_un)[1])
    return np.sum(((w.T / np.sum(w, axis=1)) * np.asarray(mass, dtype=np.float64)[sel]).T, axis=0)


def save_output(outputs_dir, f_un, mass, particle_type, time_coord=(), dust_temps=(), star_frac=(), imf_measure=(),
                chems_error=(), sup_error=(), AGB=None, number=None):
    """Write outputs/<number>.npz (default: latest + 1) with the keys of code_running.py:670.
    `AGB` may carry the star-derived arrays (AGB_condition, AGB_list, AGB_time_until, AGB_metallicity,
    AGB_composition); with no stars (the hot path's gas/dust runs) they are empty, as they are in the
    reference when `particle_type == 1` selects nothing."""
    n = len(mass)
    agb = dict(AGB_condition=np.zeros(n, dtype=bool), AGB_list=np.zeros(0), AGB_time_until=np.zeros(0),
               AGB_metallicity=np.zeros(0), AGB_composition=np.zeros((0, np.shape(f_un)[1])))
    if AGB:
        agb.update(AGB)
    if number is None:
        number = latest_output_number(outputs_dir) + 1
    os.makedirs(outputs_dir, exist_ok=True)
    path = os.path.join(outputs_dir, "%d" % int(number))
    np.savez(path, gas_mass_by_species=species_masses(f_un, mass, particle_type, 0),
             star_mass_by_species=species_masses(f_un, mass, particle_type, 1),
             dust_mass_by_species=species_masses(f_un, mass, particle_type, 2),
             time_coord=np.asarray(time_coord, dtype=np.float64), dust_temps=np.asarray(dust_temps, dtype=np.float64),
             star_frac=np.asarray(star_frac, dtype=np.float64), imf_measure=np.asarray(imf_measure, dtype=np.float64),
             chems_error=np.asarray(chems_error, dtype=np.float64), sup_error=np.asarray(sup_error, dtype=np.float64), **agb)
    return path + ".npz"


def load_output(outputs_dir, number):
    with np.load(os.path.join(outputs_dir, "%d.npz" % int(number))) as fh:
        return {k: fh[k] for k in fh.files}


# -- particle checkpoint ---------------------------------------------------------------------------
# The reference has none: code_running_2.py:162-166 only sketches reloading positions, and a run that
# stops loses its particles.  One .npz holds everything a Session needs to continue bit-for-bit: the
# eleven state arrays under the driver's variable names, the k-NN trial radii, and the scalars.
CHECKPOINT_SCALARS = ("age", "dt", "d", "ct")
CHECKPOINT_PARAMS = ("dt_0", "MAX_AGE", "N_INT_PER_PARTICLE", "raise_factor", "AU")


def save_checkpoint(path, session, params=None, **extra):
    """Write the device-resident state of `session` (sphb200.session.Session) to `path` (.npz).
    `params` overrides the run parameters recorded (default: the module constants the session was
    created with, plus its dt_0); `extra` arrays (e.g. f_un) are stored alongside."""
    from .session import STATE_FIELDS
    out = {name: session.download(name) for name in STATE_FIELDS}
    out["knn_h"] = session.download_tensor("knn_h").cpu().numpy()
    for name in CHECKPOINT_SCALARS:
        out["scalar_" + name] = np.float64(session.scalar(name))
    p = dict(dt_0=session.dt_0, MAX_AGE=C.MAX_AGE, N_INT_PER_PARTICLE=C.N_INT_PER_PARTICLE, raise_factor=C.raise_factor,
             AU=C.AU)
    p.update(params or {})
    for name in CHECKPOINT_PARAMS:
        out["param_" + name] = np.float64(p[name])
    out["N_NEIGH"] = np.int64(session.k)
    out["dist"] = np.float64(session.dist)
    out["neighbor"] = session.download("neighbor").astype(np.int32)      # the list the next crossing time reads
    for name, a in extra.items():
        out["extra_" + name] = np.asarray(a)
    np.savez(path, **out)
    return path if str(path).endswith(".npz") else str(path) + ".npz"


def load_checkpoint(path, device=None):
    """Rebuild a Session from save_checkpoint's file: state uploaded, scalars restored, neighbour
    list restored, ready for step().  Returns (session, extra)."""
    from .session import Session, STATE_FIELDS
    with np.load(path) as fh:
        data = {k: fh[k] for k in fh.files}
    n = len(data["points"])
    s = Session(n, int(data["N_NEIGH"]), float(data["dist"]), device=device, dt_0=float(data["param_dt_0"]))
    for name in CHECKPOINT_PARAMS:
        s.set_scalar(name, float(data["param_" + name]))
    s.upload({name: data[name] for name in STATE_FIELDS})
    import torch
    s.upload_tensor("knn_h", torch.from_numpy(np.ascontiguousarray(data["knn_h"], dtype=np.float64)).to(s.device))
    for name in CHECKPOINT_SCALARS:
        s.set_scalar(name, float(data["scalar_" + name]))
    s.upload_neighbor(data["neighbor"])
    return s, {k[len("extra_"):]: v for k, v in data.items() if k.startswith("extra_")}
