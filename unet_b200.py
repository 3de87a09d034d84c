"""Import shim: the package directory is named ``unet-_b200`` (not a valid Python
identifier), so this module makes it importable as ``unet_b200``.

It behaves as a package: ``__path__`` points at ``unet-_b200/`` and the directory's
``__init__.py`` is executed in this module's namespace.
"""
import os as _os

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "unet-_b200")
__path__ = [_dir]
__package__ = __name__          # so `from .model import ...` inside __init__.py resolves
if __spec__ is not None:
    __spec__.submodule_search_locations = __path__
with open(_os.path.join(_dir, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_dir, "__init__.py"), "exec"))
del _f
