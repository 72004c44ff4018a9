"""Importable alias of the package directory `twitter-stream-ml_b200/` (hyphens are not valid in a
Python module name): this package's search path IS that directory."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "twitter-stream-ml_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
del _f
