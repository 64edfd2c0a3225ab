"""Import alias for the package directory ``game-of-life---mpi-cuda_b200/``.

The product directory carries the upstream project's name, which is not a legal
Python identifier; ``import gol_b200`` executes that directory's ``__init__``
under this name so that sub-packages (``gol_b200.ops`` ...) resolve there.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "game-of-life---mpi-cuda_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f, _os
