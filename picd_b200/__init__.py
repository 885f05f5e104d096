"""Import alias for the framework package.

The source tree lives in ``pytorch-imageclassification-distributed_b200/`` (the
directory name the project brief asks for).  A hyphen cannot appear in a Python
module name, so this tiny package points its ``__path__`` at that directory and
executes its ``__init__``; ``import picd_b200.models`` therefore resolves to
``pytorch-imageclassification-distributed_b200/models``.
"""
import os as _os

_SRC = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                     "pytorch-imageclassification-distributed_b200")
__path__ = [_SRC]
with open(_os.path.join(_SRC, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_SRC, "__init__.py"), "exec"))
del _f
