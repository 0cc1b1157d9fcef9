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


_utils = importlib.import_module('kungfu_b200._utils')  # measure, map_maybe, show_duration, _log_event, ...
sys.modules['kungfu._utils'] = _utils


# ``kungfu.tensorflow.*``: the reference's main front end was TensorFlow 1.x. Its module layout and names are
# kept so that ported code finds everything in the same place; the objects are the PyTorch ones (there is no
# TensorFlow in a B200 stack built on torch): optimizers wrap torch optimizers, ops take torch tensors.
import types as _types

_tf = _types.ModuleType('kungfu.tensorflow', 'Reference layout of kungfu.tensorflow, backed by kungfu_b200.torch')
_v1 = _types.ModuleType('kungfu.tensorflow.v1', 'Reference layout of kungfu.tensorflow.v1')
_exp = _types.ModuleType('kungfu.tensorflow.experimental', 'Reference layout of kungfu.tensorflow.experimental')
_exp_hook = _types.ModuleType('kungfu.tensorflow.experimental.hook')
for _alias, _target in (('ops', 'torch.ops'), ('optimizers', 'torch.optimizers'), ('initializer', 'torch.initializer'),
                        ('policy', 'torch.policy'), ('hooks', 'torch.elastic'), ('variables', 'variables')):
    _mod = importlib.import_module('kungfu_b200.' + _target)
    sys.modules['kungfu.tensorflow.' + _alias] = _mod
    setattr(_tf, _alias, _mod)
for _alias, _target in (('benchmarks', 'benchmarks'), ('helpers', 'datasets'), ('datasets', 'datasets'), ('utils', 'utils'),
                        ('examples', 'torch.examples')):
    _mod = importlib.import_module('kungfu_b200.' + _target)
    sys.modules['kungfu.tensorflow.v1.' + _alias] = _mod
    setattr(_v1, _alias, _mod)
_exp_hook.ElasticHook = sys.modules['kungfu.tensorflow.hooks'].ElasticHook
_exp_hook.ResizeProfiler = sys.modules['kungfu.tensorflow.hooks'].ResizeProfiler
_exp.hook = _exp_hook
_tf.v1, _tf.experimental = _v1, _exp
sys.modules.update({'kungfu.tensorflow': _tf, 'kungfu.tensorflow.v1': _v1, 'kungfu.tensorflow.experimental': _exp,
                    'kungfu.tensorflow.experimental.hook': _exp_hook})
tensorflow = _tf
