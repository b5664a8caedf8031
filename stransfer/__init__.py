"""Drop-in alias package: ``import stransfer.network`` / ``python -m stransfer ...`` resolve to
the B200 implementation in ``styletransfer_b200`` (same module names as the reference's
``stransfer/`` package: network, constants, img_utils, dataset, c_logging)."""
import importlib
import sys

for _name in ('constants', 'c_logging', 'img_utils', 'dataset', 'network'):
    _mod = importlib.import_module('styletransfer_b200.' + _name)
    sys.modules[__name__ + '.' + _name] = _mod
    globals()[_name] = _mod
