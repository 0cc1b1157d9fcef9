"""Drop-in alias: ``import kungfu.torch as kf`` / ``import kungfu.python`` keep
working on top of kungfu_b200 (reference package layout:
/root/reference/srcs/python/kungfu/)."""
import importlib
import sys

import kungfu_b200

__version__ = kungfu_b200.__version__

for _name in ('python', 'torch', 'torch.ops', 'torch.optimizers', 'cmd', 'info', 'optim', 'models', 'datasets',
              'benchmarks', 'variables', 'utils'):
    _mod = importlib.import_module('kungfu_b200.' + _name)
    sys.modules['kungfu.' + _name] = _mod
    if '.' not in _name:
        globals()[_name] = _mod
