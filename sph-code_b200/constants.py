"""Physical constants and species tables of the reference's physics library.

Values mirror /root/reference/sph/navier_stokes_cleaned.py:21-52 (they come from
scipy.constants there as well, so the numbers are identical on any box with the
same scipy).  Names are the reference's: `k` is Boltzmann, `h` is Planck.
"""
import numpy as np
from scipy import constants as _c

G = _c.G
k = _c.Boltzmann
sb = _c.Stefan_Boltzmann
AU = _c.au
amu = _c.physical_constants['atomic mass constant'][0]
h = _c.h
c = _c.c
wien = _c.Wien
m_h = 1.0008 * amu
solar_mass = 1.989e30
solar_luminosity = 3.846e26
solar_lifespan = 1e10
t_cmb = 2.732
t_solar = 5776
supernova_energy = 1e44
m_0 = 10 ** 1.5 * solar_mass
year = 60. * 60. * 24. * 365.
dt_0 = year * 250000.

species_labels = np.array(['H2', 'He', 'H', 'H+', 'He+', 'He++', 'e-', 'Mg2SiO4', 'SiO2', 'C', 'Si', 'Fe',
                           'MgSiO3', 'FeSiO3', 'SiC'])
mu_specie = np.array([2.0158, 4.0026, 1.0079, 1.0074, 4.0021, 4.0016, 0.0005, 140.69, 60.08, 12.0107, 28.0855,
                      55.834, 100.39, 131.93, 40.096])
gamma = np.array([7. / 5, 5. / 3, 5. / 3, 5. / 3, 5. / 3, 5. / 3, 5. / 3, 15.6354113, 4.913, 1.0125, 2.364, 3.02,
                  10., 10., 10.])
mineral_densities = np.array([1.e19, 1e19, 1e19, 1e19, 1e19, 1e19, 1e19, 3320, 2260, 2266, 2329, 7870, 3250, 3250.,
                              3166.])                                   # navier_stokes_cleaned.py:44
mrn_constants = np.array([50e-10, 5000e-10])                            # :52, MRN grain radius range
f_u = np.array([.86, .14, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0])
N_SPECIES = 15

# driver-level defaults (/root/reference/sph/code_running.py:59-81)
raise_factor = 715
DIAMETER = 1.25e6 * AU
N_PARTICLES_DEFAULT = 3000 * raise_factor
N_INT_PER_PARTICLE = N_PARTICLES_DEFAULT / raise_factor
N_NEIGH = 40
MAX_AGE = 3e7 * year
d_default = ((DIAMETER ** 3) / N_PARTICLES_DEFAULT * N_INT_PER_PARTICLE) ** (1. / 3.) * raise_factor ** (1. / 3.)
d_0_default = 1e5 * AU
