"""CMB experiment table; mirrors reference cmb/bolometers.py:4-12."""
bolometers = {'Planck': {'fsky': 1.0, 'thetab': 7.0, 'DeltaT': 30.0, 'iterative': False},
              'AdvACT': {'fsky': 0.2, 'thetab': 1.5, 'DeltaT': 12.0, 'iterative': False},
              'SO': {'fsky': 0.4, 'thetab': 1.4, 'DeltaT': 7.0, 'iterative': False},
              'CMBS4': {'fsky': 0.4, 'thetab': 1.4, 'DeltaT': 1.0, 'iterative': True}}
