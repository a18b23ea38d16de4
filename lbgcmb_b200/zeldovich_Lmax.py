"""Perturbation-theory L_max per redshift (k sigma_zel = 0.75, RunPB); the constants of reference zeldovich_Lmax.py:2-4."""
Lcutmax = {0.0: [474., 'c'], 1.0: [474., 'b'], 2.0: [1079.9, 'r'],
           3.0: [1752.6, 'k'], 4.0: [2464.8, 'g'], 5.0: [3203.6, 'y'],
           6.0: [3962.1, 'c']}
