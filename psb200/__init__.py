"""Import alias for the framework package.

The source tree lives in ``distributed-tensorflow-examples_b200/`` (a directory name
Python cannot import because of the hyphens).  This shim makes it importable as
``psb200``: it points ``__path__`` at that directory and runs its ``__init__``.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
_SRC = _os.path.join(_ROOT, "distributed-tensorflow-examples_b200")
__path__ = [_SRC]
__file__ = _os.path.join(_SRC, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
