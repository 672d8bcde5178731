"""Importable alias for the framework package.

The package lives in ``cuda-mpi-gpu-cluster-programming_b200/`` (the name the project
brief fixes); a hyphen cannot appear in a Python ``import`` statement, so this thin
package points its ``__path__`` at that directory and executes its ``__init__``.
Every submodule is therefore importable as ``b200cnn.<sub>`` and exists exactly once
in ``sys.modules``.
"""
import os as _os

_REAL = _os.path.join(
    _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
    "cuda-mpi-gpu-cluster-programming_b200",
)
__path__ = [_REAL]
__file__ = _os.path.join(_REAL, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
