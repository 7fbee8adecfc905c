"""Importable alias for the package directory `multi-modal-trakcing-bechmark_b200/` (its name has
hyphens, so Python cannot import it directly)."""
import os as _os

_pkg = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                     "multi-modal-trakcing-bechmark_b200")
__path__ = [_pkg]
with open(_os.path.join(_pkg, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_pkg, "__init__.py"), "exec"))
