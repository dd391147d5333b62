"""Import shim: the package directory is ``kubernetes-cloud_b200`` (not a Python identifier), so this importable
name points its ``__path__`` there and runs that directory's ``__init__.py`` in its own namespace."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "kubernetes-cloud_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
