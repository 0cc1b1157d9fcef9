"""Process-wide training variables the policies and monitors share.
Reference: /root/reference/srcs/python/kungfu/tensorflow/variables.py:84-122
(GraphKeys.{BATCH_SIZE,GRADIENT_NOISE_SCALE,TOTAL_SAMPLES,TRAINED_SAMPLES})."""


class GraphKeys(object):
    BATCH_SIZE = 'kungfu_batch_size'
    GRADIENT_NOISE_SCALE = 'kungfu_gradient_noise_scale'
    TOTAL_SAMPLES = 'kungfu_total_samples'
    TRAINED_SAMPLES = 'kungfu_trained_samples'


_vars = {}


def get_or_create(key, init=0):
    """Value stored under ``key``, created with ``init`` if absent."""
    return _vars.setdefault(key, init)


def set_value(key, value):
    """Store ``value`` under ``key``."""
    _vars[key] = value


def get_value(key, default=None):
    """Value stored under ``key`` or ``default``."""
    return _vars.get(key, default)


def create_setter(key):
    """A one-argument callable that stores its argument under ``key``."""
    return lambda v: set_value(key, v)


def get_or_create_batch_size(init=None):
    """The per-worker batch size variable (set once with ``init``)."""
    if init is not None and GraphKeys.BATCH_SIZE not in _vars:
        _vars[GraphKeys.BATCH_SIZE] = init
    return _vars.get(GraphKeys.BATCH_SIZE)


def eval_batch_size():
    """Current per-worker batch size or None."""
    return _vars.get(GraphKeys.BATCH_SIZE)


def eval_gradient_noise_scale():
    """Last gradient noise scale published by the GNS monitor, as a float, or None."""
    v = _vars.get(GraphKeys.GRADIENT_NOISE_SCALE)
    return float(v) if v is not None else None


# ---- generic accessors under the reference's names (tensorflow/variables.py:17-82) ---------------
def create_global_variable(key, init=0):
    """Create (or overwrite) ``key`` with ``init``."""
    _vars[key] = init
    return init


def get_global_variable(key):
    """Value of an existing variable (KeyError if absent)."""
    return _vars[key]


def get_or_create_global_variable(key, init=0):
    """Same as get_or_create (reference name)."""
    return get_or_create(key, init)


def eval_global_variable(key):
    """Value of ``key`` or None."""
    return _vars.get(key)


def create_assign_op_for(key):
    """A callable that stores its argument (the eager counterpart of an assign op)."""
    return create_setter(key)


def create_placeholder_for(key):
    """Eager code has no placeholders: the current value stands in for it."""
    return _vars.get(key)


def get_batch_size():
    """Same as eval_batch_size (reference name)."""
    return eval_batch_size()
