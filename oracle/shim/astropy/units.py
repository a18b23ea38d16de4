"""empty: the hot path only imports the name (Cgg.py:3, snr.py:5)."""
