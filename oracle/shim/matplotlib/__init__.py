"""No-op matplotlib so reference modules that style plots at import time load (cmb/Nkk.py:11)."""
import types, sys


class _Any:
    def __getattr__(self, name):
        return _Any()

    def __call__(self, *a, **k):
        return _Any()

    def __getitem__(self, k):
        return _Any()

    def __setitem__(self, k, v):
        pass

    def __iter__(self):
        return iter(())

    # utils.py:80-90 (set_size) does arithmetic on figure geometry
    def _num(self, *a):
        return 1.0
    __add__ = __radd__ = __sub__ = __rsub__ = __mul__ = __rmul__ = __truediv__ = __rtruediv__ = _num

    def __float__(self):
        return 1.0


rcParams = _Any()
style = _Any()


def use(*a, **k):
    pass


pyplot = types.ModuleType('matplotlib.pyplot')
pyplot.__getattr__ = lambda name: _Any()
sys.modules['matplotlib.pyplot'] = pyplot
pylab = types.ModuleType('pylab')
pylab.__getattr__ = lambda name: _Any()
sys.modules['pylab'] = pylab


def __getattr__(name):          # mpl.rc(...), mpl.rcdefaults(), ... (utils.py:120 latexify)
    if name.startswith('__'):
        raise AttributeError(name)
    return _Any()
