"""Import alias: the package directory is named ``cloud-server_b200`` (not a valid Python
identifier); ``import cloud_server_b200`` resolves to it."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "cloud-server_b200")
__path__[:] = [_real]
_init = _os.path.join(_real, "__init__.py")
with open(_init) as _f:
    exec(compile(_f.read(), _init, "exec"))
