"""CPU: the reference's .npz config / output schema (config_bootstrap.py:33, config_helper.py:58-76,284,
code_running.py:656-670) as mirrored by sphb200.config_io."""
import numpy as np
import pytest

from sphb200 import config_io as cio


def test_bootstrap_roundtrip_and_numbering(tmp_path):
    d = str(tmp_path / "config")
    assert cio.latest_config_number(d) == -1
    with pytest.raises(IOError):
        cio.load_config(d)
    p0 = cio.save_config(d, **cio.bootstrap_config())
    assert p0.endswith("config_0.npz")
    cfg = cio.load_config(d)
    assert sorted(cfg) == sorted(cio.CONFIG_KEYS)
    assert cfg["specie_fraction_array"].shape == (14,) and cfg["overall_AGB_composition"].shape == (2, 14)
    assert np.isclose(np.sum(cfg["dust_base_frac"]), 1.0)
    # the next file is <latest> + 1 and wins (config_helper.py:58-63,284)
    nxt = dict(cfg); nxt["OVERALL_AGE"] = 3.0e15
    assert cio.save_config(d, **nxt).endswith("config_1.npz")
    assert cio.latest_config_number(d) == 1 and float(cio.load_config(d)["OVERALL_AGE"]) == 3.0e15
    with np.load(p0) as raw:                      # plain np.load sees exactly the reference's keys
        assert sorted(raw.files) == sorted(cio.CONFIG_KEYS)


def test_missing_key_is_an_error(tmp_path):
    cfg = cio.bootstrap_config()
    del cfg["DUST_FRAC"]
    with pytest.raises(KeyError):
        cio.save_config(str(tmp_path), **cfg)


def test_composition_padding():
    c = cio.composition_from_config(cio.bootstrap_config())
    assert c["specie_fraction_array"].shape == (15,) and c["specie_fraction_array"][0] == 0.86
    assert c["DUST_FRAC"] == 0.000000001 and c["OVERALL_AGE"] == 0.0


def test_output_file(tmp_path):
    from sphb200 import ics, constants as C
    st = ics.uniform_box(500, 0, n_dust=5)
    d = str(tmp_path / "outputs")
    p = cio.save_output(d, st["f_un"], st["mass"], st["particle_type"], time_coord=[0., 1.], dust_temps=[10., 11.])
    assert p.endswith("0.npz") and cio.latest_output_number(d) == 0
    out = cio.load_output(d, 0)
    assert sorted(out) == sorted(cio.OUTPUT_KEYS)
    gas = st["particle_type"] == 0
    assert np.isclose(np.sum(out["gas_mass_by_species"]), np.sum(st["mass"][gas]), rtol=1e-12)
    assert np.isclose(np.sum(out["dust_mass_by_species"]), np.sum(st["mass"][~gas]), rtol=1e-12)
    assert np.all(out["star_mass_by_species"] == 0) and out["AGB_list"].shape == (0,)
    # mass fractions of the gas are f_un * mu normalised (code_running.py:656)
    w = C.f_u * C.mu_specie
    assert np.allclose(out["gas_mass_by_species"] / np.sum(out["gas_mass_by_species"]), w / np.sum(w))
