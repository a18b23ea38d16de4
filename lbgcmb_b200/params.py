"""Cosmology / experiment constants; mirrors reference params.py:1-15 (RunPB set is the active one there)."""


def get_params():
    params = {'om_m': 0.292, 'om_b': 0.0468, 'om_L': 0.708, 'h_100': 0.69, 'sig_8': 0.819, 'Tcmb': 2.786,
              'zscatter': 1100, 'ns': 0.9688, 'As': 2.187e-9}
    params['om_cm'] = params['om_m'] - params['om_b']
    params['h_70'] = (10. / 7.) * params['h_100']
    return params
