"""Drop-in for the reference's `weighted_distance` package (submodules/weighted_distance/setup.py:10-20).

`from weighted_distance._C import weightedDistance` (compression/vq.py:12) resolves here once
`c3dgs_b200.install()` has registered this package under that name, or when this directory is
put on sys.path ahead of the reference's build.
"""
from . import _C  # noqa: F401
