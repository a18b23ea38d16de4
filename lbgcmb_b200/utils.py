"""utils.comoving_distance; mirrors reference utils.py:33-34 (chi in h^-1 Mpc)."""
from lbgcmb_b200.cosmo import cosmo
from lbgcmb_b200.params import get_params

params = get_params()


def comoving_distance(z):
    return params['h_100'] * cosmo.comoving_distance(z).value     # [h^-1 Mpc]
