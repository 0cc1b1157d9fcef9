"""Adaptation and elastic control. Reference:
/root/reference/srcs/python/kungfu/tensorflow/ops/adapt.py:5-62,
/root/reference/srcs/cpp/src/tensorflow/ops/cpu/{control,elastic,adaptation}.cpp."""
from ... import python as kfpy
from ...utils import step_based_schedule as _sbs


def resize(n):
    """Resize the cluster to n workers. Returns (changed, detached)."""
    return kfpy.resize_cluster(int(n))


resize_cluster = resize


def resize_cluster_from_url():
    return kfpy.resize_cluster_from_url()


def step_based_schedule(config, step, default=1, strict=True):
    return _sbs(config, step, default, strict)


def set_tree(tree):
    kfpy.set_tree([int(t) for t in tree])
    return True


def calc_stats():
    kfpy.calc_stats()


def log_stats():
    kfpy.log_stats()


def print_strategy_stats():
    kfpy.print_strategy_stats()


def check_interference():
    return kfpy.check_interference()
