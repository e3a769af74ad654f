"""Importable alias of the product package, which lives in the directory `improved-body-parts_b200/`
(a hyphen cannot appear in a Python module name).  `import improved_body_parts_b200 as ibp` executes that
directory's __init__.py with this module's __path__ pointing at it, so sub-modules resolve there."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "improved-body-parts_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
