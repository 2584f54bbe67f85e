# Mirrors bnn/__init__.py:1-2 of BNN-PYNQ: the public names of the package.
from .bnn import *  # noqa: F401,F403
from .bnn import available_params, CnvClassifier, LfcClassifier, PynqBNN  # noqa: F401
