"""Import stub (test infrastructure): plot.py:2-3 imports pyqtgraph for the OpenGL debug
viewer, which is out of scope (no graphics stack on this box)."""


def mkQApp(*a, **k):
    raise RuntimeError("pyqtgraph stub: the OpenGL viewer is unavailable on this box")
