"""Import stub (test infrastructure): utils.py:6-7,11 imports colorama only for the
`f()` console formatter, which is not on the hot path."""


def init(*a, **k):
    return None


class _Codes:
    def __getattr__(self, name):
        return ""


Fore = _Codes()
Back = _Codes()
Style = _Codes()
