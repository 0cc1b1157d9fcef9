"""Initial state synchronisation. Reference:
/root/reference/srcs/python/kungfu/tensorflow/initializer/__init__.py:13-99
(broadcast_variables, BroadcastGlobalVariablesOp/Hook/Callback)."""
from .ops.collective import broadcast_parameters


def broadcast_variables(tensors):
    """Every tensor := rank 0's."""
    broadcast_parameters({'v%d' % i: t for i, t in enumerate(tensors)})


def broadcast_optimizer_state(optimizer):
    """Momentum buffers etc. := rank 0's (tensors in optimizer.state)."""
    sd = {}
    for gi, g in enumerate(optimizer.param_groups):
        for pi, p in enumerate(g['params']):
            for k, v in optimizer.state.get(p, {}).items():
                if hasattr(v, 'is_contiguous'):
                    sd['g%d.p%d.%s' % (gi, pi, k)] = v
    broadcast_parameters(sd)


class BroadcastGlobalVariablesCallback(object):
    """Call ``on_batch_end()`` after every batch: broadcasts model (and
    optimizer slots, which only exist after the first step) exactly once, after
    the first batch -- the reference's Keras callback semantics (:80-99)."""

    def __init__(self, model, optimizer=None):
        self.model, self.optimizer, self.done = model, optimizer, False

    def on_batch_end(self):
        if self.done:
            return
        broadcast_parameters(self.model.state_dict())
        if self.optimizer is not None:
            broadcast_optimizer_state(self.optimizer)
        self.done = True


def BroadcastGlobalVariablesOp(model, optimizer=None):
    """Reference name (initializer/__init__.py:30): broadcast everything now. Returns None (eager)."""
    broadcast_parameters(model.state_dict())
    if optimizer is not None:
        broadcast_optimizer_state(optimizer)


class BroadcastGlobalVariablesHook(BroadcastGlobalVariablesCallback):
    """Reference name for the session hook flavour: ``after_create_session()`` broadcasts once."""

    def after_create_session(self, *_):
        self.on_batch_end()
