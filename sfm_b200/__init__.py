"""Importable alias of the `structure-from-motion_b200/` package directory (a hyphen cannot be
imported).  `import sfm_b200` executes structure-from-motion_b200/__init__.py under this name, so
`sfm_b200.engine`, `sfm_b200.ba_processor`, ... resolve to the files in that directory."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), 'structure-from-motion_b200')
__path__ = [_real]
with open(_os.path.join(_real, '__init__.py')) as _f:
    exec(compile(_f.read(), _os.path.join(_real, '__init__.py'), 'exec'))
del _f
