"""Adaptation policies: user code that observes training and may resize the
cluster or change hyper-parameters at step / epoch boundaries.
Reference: /root/reference/srcs/python/kungfu/tensorflow/policy/{base_policy.py:5-31,
policy_hook.py:8-77}, /root/reference/examples/mnist_elastic_policy/elastic_policy.py:9-45."""
from ... import python as kfpy
from ... import variables as V
from ...utils import step_based_schedule
from ..elastic import ElasticState


class BasePolicy(object):
    def before_train(self):
        pass

    def before_epoch(self):
        pass

    def before_step(self):
        pass

    def after_step(self, state):
        pass

    def after_epoch(self, state):
        pass

    def after_train(self):
        pass


class PolicyRunner(object):
    """Drives policies from a plain training loop (the reference's PolicyHook
    for Estimators): tracks trained samples / epochs, stops detached workers."""

    def __init__(self, policies, epoch_size, epoch_num, model, optimizer=None, batch_size=None):
        self.policies = list(policies)
        self.epoch_size, self.epoch_num = epoch_size, epoch_num
        self.state = ElasticState(model, optimizer)
        if batch_size is not None:
            V.get_or_create_batch_size(batch_size)
        self._epoch = 0
        for p in self.policies:
            p.before_train()
        for p in self.policies:
            p.before_epoch()

    @property
    def total_samples(self):
        return self.epoch_size * self.epoch_num

    def before_step(self):
        for p in self.policies:
            p.before_step()

    def after_step(self):
        """Returns False when this worker should stop (detached or finished)."""
        bs = V.eval_batch_size() or 0
        self.state.end_step(bs)
        V.set_value(V.GraphKeys.TRAINED_SAMPLES, self.state.trained_samples)
        for p in self.policies:
            p.after_step(self.state)
        if self.state.trained_samples >= (self._epoch + 1) * self.epoch_size:
            for p in self.policies:
                p.after_epoch(self.state)
            self._epoch += 1
            if self._epoch < self.epoch_num:
                for p in self.policies:
                    p.before_epoch()
        if kfpy.detached() or self.state.stopped:
            return False
        if self.state.trained_samples >= self.total_samples:
            for p in self.policies:
                p.after_train()
            return False
        return True


class ScheduledElasticPolicy(BasePolicy):
    """Resize at fixed steps: {step: new_size} or a "size:steps,..." string."""

    def __init__(self, schedule):
        self.schedule = schedule

    def after_step(self, state):
        if isinstance(self.schedule, dict):
            want = self.schedule.get(state.step)
        else:
            want = step_based_schedule(self.schedule, state.step, default=kfpy.current_cluster_size(), strict=False)
        if want is not None and want != kfpy.current_cluster_size():
            state.resize(want)


class NoiseScaleElasticPolicy(BasePolicy):
    """Grow the cluster while the gradient noise scale says larger batches still
    pay off (BASELINE.json config 5: 1 -> 2 -> 4 -> 8 driven by the GNS
    monitor). `optimizer` is a MonitorGradientNoiseScaleOptimizer."""

    def __init__(self, optimizer, device_batch_size, max_size=8, check_every=10, min_steps_between=20):
        self.opt, self.bs = optimizer, device_batch_size
        self.max_size, self.check_every, self.cooldown = max_size, check_every, min_steps_between
        self._last_resize = 0

    def after_step(self, state):
        np_ = kfpy.current_cluster_size()
        # decisions use only state that ElasticState keeps identical on every worker
        # (a joiner's fresh policy object must not disagree with the survivors')
        if state.step % self.check_every or state.step - state.last_resize_step < self.cooldown or np_ >= self.max_size:
            return
        ns = getattr(self.opt, 'noise_scale', None)
        # single worker: no cross-worker norm yet, grow once to start measuring
        # B_noise = S/G. A non-positive estimate means the gradient signal is still
        # below the noise floor at this batch size: larger batches help as well.
        grow = np_ == 1 or (ns is not None and (float(ns) <= 0 or float(ns) > self.bs * np_))
        # every rank must take the same decision
        from ..ops import collective as tops
        import torch
        votes = int(tops.all_reduce_fn(torch.tensor([1 if grow else 0]), 'sum', name='kungfu::gns-policy')[0])
        if votes * 2 > np_:
            V.set_value(V.GraphKeys.GRADIENT_NOISE_SCALE, float(ns) if ns is not None else None)
            self._last_resize = state.step
            state.resize(min(self.max_size, np_ * 2))


# reference name (tensorflow/policy/policy_hook.py:8): the Estimator hook that drives the policies
PolicyHook = PolicyRunner
