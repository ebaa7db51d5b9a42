"""Import stub (test infrastructure): the reference imports ipdb at module load
(utils.py:3,5; plot.py:5) but never calls it on the hot path."""


def set_trace(*a, **k):
    raise RuntimeError("ipdb stub: debugger is not available")
