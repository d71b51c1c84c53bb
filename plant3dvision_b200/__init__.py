"""Importable name of the package whose sources live in ``plant-3d-vision_b200/``.

The directory name the project layout asks for contains hyphens and therefore cannot be
imported; this module forwards ``import plant3dvision_b200[.x]`` to it.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "plant-3d-vision_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
