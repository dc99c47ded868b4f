"""Marker classes naming the integration methods, one per ``ozone.methods`` entry
(``ozone/methods.py:1-68``).  Each carries only ``.name``; the tableau is looked up in
``glms.get_integration_method``."""
from .glms import method_names as _names


class _Method(object):
    name = None

    def __repr__(self):
        return f'ozone_b200.methods.{type(self).__name__}()'


def _make(n):
    return type(n, (_Method,), {'name': n, '__doc__': f'Runge-Kutta method {n!r} (ozone/glms/GLMs.py:45-68).'})


for _n in _names:
    globals()[_n] = _make(_n)
__all__ = ['_Method'] + list(_names)
del _n
