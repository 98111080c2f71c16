"""Import shim: the package sources live in the directory `image-segmentation_b200/` (the name the
project layout prescribes, which is not a valid Python identifier); this module makes them importable
as `image_segmentation_b200`."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "image-segmentation_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _os, _f
