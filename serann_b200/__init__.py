"""Import shim: the product package lives in the directory the build contract
names, ``self-replicating-artificial-neural-networks_b200/``.  That name has
hyphens and cannot be imported, so ``serann_b200`` points its ``__path__``
there; ``import serann_b200.frontend`` loads
``self-replicating-artificial-neural-networks_b200/frontend.py``.
"""
import pathlib as _pathlib

_real = _pathlib.Path(__file__).resolve().parent.parent / "self-replicating-artificial-neural-networks_b200"
__path__ = [str(_real)]
__version__ = "0.1.0"
