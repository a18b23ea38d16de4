"""Stand-in so `import prep_camb` succeeds (prep_camb.py:1,5).  CAMB itself is absent and
unpinned; nothing here computes anything."""


class CAMBparams:
    def __getattr__(self, name):
        raise RuntimeError("camb is not available in this container")


class _M:
    NonLinear_none = 0
    NonLinear_both = 3


model = _M()
initialpower = _M()
