"""Minimal stand-in for astropy 3.1 (absent here) so the UNMODIFIED reference modules import.
TEST INFRASTRUCTURE (oracle/refload.py only).  Covers exactly what the hot path touches:
cosmo.py:6, utils.py:33, Cgg.py:60, cmb/lensing.py:19,36, modi/cosmology.py:32."""
from . import cosmology, constants, units  # noqa: F401
