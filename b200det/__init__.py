"""Import alias: ``import b200det`` loads the package whose sources live in
``objectdetection.pytorch_b200/`` (a directory name Python cannot import
directly because of the dot)."""
import os as _os

_src = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                     "objectdetection.pytorch_b200")
__path__.insert(0, _src)
with open(_os.path.join(_src, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_src, "__init__.py"), "exec"))
del _os, _f
