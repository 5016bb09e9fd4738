"""CPU: libsphb200.so builds for sm_100a, loads, and exports every symbol include/sphb200.h declares
(no compute calls: there is no GPU here).  Also: the product path fails loudly without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "sphb200.h")) as fh:
        text = fh.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sphb200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_path():
    names = declared_symbols()
    for must in ("sphb200_neighbors", "sphb200_hydro_update", "sphb200_crossing_time", "sphb200_loop_density",
                 "sphb200_del_pressure", "sphb200_artificial_viscosity", "sphb200_session_step"):
        assert must in names


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    for name in declared_symbols():
        assert hasattr(lib, name), "libsphb200.so does not export %s" % name


def test_binding_covers_header(built_lib):
    from sphb200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    _lib.load()


def test_library_is_sm100a_only(built_lib):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", built_lib], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import numpy as np
    from sphb200 import _lib
    from sphb200 import navier_stokes_cleaned as nsc
    with pytest.raises(_lib.Sphb200Error):
        nsc.neighbors(np.zeros((10, 3)), 1.0, 4)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "sph-code_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, f)) as fh:
                    src = fh.read()
                assert "import oracle" not in src and "from oracle" not in src, f
