"""Import alias for the framework.

The source tree lives in ``distributed-tensorflow-guide_b200/`` (the project's
name, which is not a valid Python identifier).  ``import dtg_b200`` points the
package search path at that directory and runs its ``__init__``; every
submodule is then reachable as ``dtg_b200.<name>``.
"""
import os as _os

_SRC = _os.path.join(
    _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
    "distributed-tensorflow-guide_b200",
)
__path__ = [_SRC]
_init = _os.path.join(_SRC, "__init__.py")
with open(_init) as _f:
    exec(compile(_f.read(), _init, "exec"))
del _f, _init
