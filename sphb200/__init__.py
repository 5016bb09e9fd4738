"""Importable alias for the product package, which lives in `sph-code_b200/` (a directory name
Python cannot import directly).  `import sphb200` executes `sph-code_b200/__init__.py` with this
package's `__path__` pointing there, so `sphb200.navier_stokes_cleaned`, `sphb200.session` … resolve
to the files under `sph-code_b200/`."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "sph-code_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _fh:
    exec(compile(_fh.read(), _os.path.join(_real, "__init__.py"), "exec"))
