class _C:
    def to(self, unit):
        assert unit == 'km/s'

        class _V:
            value = 299792.458
        return _V()


c = _C()


class _PC:
    value = 3.0856775814913673e16


pc = _PC()
