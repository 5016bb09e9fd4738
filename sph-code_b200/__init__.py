"""sphb200 — B200-native SPH particle time-step behind the reference's `navier_stokes_cleaned` API.

Package contents (only what the hot path needs):
  csrc/                      CUDA kernels + the C ABI (include/sphb200.h) -> libsphb200.so
  navier_stokes_cleaned.py   drop-in module: neighbors, hydro_update, density, ... (ctypes binding)
  session.py                 device-resident state + whole time step
  ics.py, constants.py       seeded initial conditions of the measured configs, physical constants
  config_io.py               the reference's .npz config / output schema
  domain.py                  slab decomposition + ghost exchange for 1..8 GPUs
Importing the package does not load the CUDA library; the first call does, and fails loudly if it is
missing (there is no CPU fallback).
"""
__version__ = "0.1.0"
