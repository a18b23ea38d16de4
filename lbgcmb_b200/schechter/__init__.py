"""Dropout-sample input providers; mirrors the pieces of reference dropouts/schechter/ the forecast drivers import
(snr.py:47-49, interloper.py:93-95): get_pz.get_pz, get_shot.get_shot, gen_pz.peakz.  Host-side file readers, as in the
reference; the files live under $LBGCMB/dropouts/schechter/dat (the reference's own data tree, never copied here)."""
