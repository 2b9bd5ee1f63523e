"""Stand-in for seaborn (see ../matplotlib/__init__.py)."""
from matplotlib import _Anything


def __getattr__(name):
    return _Anything()
