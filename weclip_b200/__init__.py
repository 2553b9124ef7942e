"""Importable alias for the ``weclip-vit-comer_b200/`` package directory.

The product package lives in ``weclip-vit-comer_b200/`` (a name Python cannot import
because of the hyphens); this shim points ``weclip_b200``'s package path at it so that
``import weclip_b200.par`` resolves to ``weclip-vit-comer_b200/par.py``.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "weclip-vit-comer_b200")
__path__ = [_real]
_init = _os.path.join(_real, "__init__.py")
with open(_init) as _f:
    exec(compile(_f.read(), _init, "exec"))
del _os, _f, _init
