"""Import shim: the package directory is named `voxelnet-pytorch_b200/` (the layout the project
brief fixes), which is not a valid Python identifier.  This module makes it importable as
`voxelnet_pytorch_b200` by pointing `__path__` at that directory and executing its `__init__.py`
in this module's namespace.  Nothing else lives here."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "voxelnet-pytorch_b200")]
__package__ = __name__
if __spec__ is not None:
    __spec__.submodule_search_locations = __path__
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
del _f
