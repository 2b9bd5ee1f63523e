"""Stand-in for matplotlib on images that do not ship it (this one): lets the UNMODIFIED reference scripts import and
call their plotting code (client1.py:15-19,153-225) without drawing anything.  Put baseline/plot_stubs on PYTHONPATH
for the reference processes only; the product package has its own Pillow renderer (report/plots.py)."""


class _Anything:
    def __getattr__(self, name):
        return _Anything()

    def __call__(self, *a, **k):
        return _Anything()

    def __iter__(self):
        return iter((_Anything(), _Anything()))

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def use(*a, **k):
    return None


def __getattr__(name):
    return _Anything()
