"""Drop-in for the package gmat/remma/remma_epiDD/ (same function names and signatures)."""
from . import _core, _frontends

__all__ = _frontends.build(_core.DD, globals())
