"""Mirrors of the reference's `modi/` flavour (Modi et al. lensing code): nkk, halofit, cosmology, lensing."""
