"""Single-process checks of the Python API surface a KungFu user expects
(/root/reference/tests/python/unit/*: counter, EMA, step schedule, detached,
rank/size, variables, policy hook) plus the tools."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT, run_workers


def test_alias_package_and_surface():
    import kungfu.torch as kf
    from kungfu.python import current_cluster_size, current_rank
    assert current_rank() == 0 and current_cluster_size() == 1
    for name in ('SynchronousSGDOptimizer', 'SynchronousAveragingOptimizer', 'PairAveragingOptimizer',
                 'MonitorGradientNoiseScaleOptimizer', 'MonitorGradientVarianceOptimizer', 'AdaptiveSGDOptimizer'):
        assert hasattr(kf.optimizers, name)
    for name in ('all_reduce_fn', 'inplace_all_reduce_op', 'inplace_all_reduce_async_op', 'inplace_broadcast_async_op',
                 'wait_handle', 'wait_all_handles', 'broadcast_parameters', 'all_gather', 'barrier', 'consensus',
                 'monitored_all_reduce', 'save_variable', 'request_variable', 'resize', 'set_tree', 'calc_stats',
                 'egress_rates', 'global_noise_scale', 'counter', 'exponential_moving_average', 'peer_info',
                 'get_peer_latencies', 'global_minimum_spanning_tree', 'get_neighbour_mask', 'round_robin',
                 'group_all_reduce', 'fuse', 'defuse', 'step_based_schedule'):
        assert hasattr(kf.ops, name), name
    assert kf.nccl_built() is False
    assert kf.get_cuda_index() >= 0


def test_state_ops_and_schedule():
    from kungfu_b200.torch import ops
    c = ops.counter(3, 2)
    assert [c(), c(), c()] == [3, 5, 7]
    e = ops.exponential_moving_average(0.9)
    assert [round(e(x), 4) for x in (2, 3, 4, 5, 6)] == [2.0, 2.1, 2.29, 2.561, 2.9049]
    sched = '1:2,2:2,4:1'
    assert [ops.step_based_schedule(sched, s) for s in range(5)] == [1, 1, 2, 2, 4]
    assert ops.step_based_schedule(sched, 9, default=7, strict=False) == 7
    with pytest.raises(ValueError):
        ops.step_based_schedule(sched, 9)
    assert ops.group_all_reduce([torch.ones(2), None])[1] is None
    a, b = torch.arange(6.).view(2, 3), torch.arange(4.)
    flat = ops.fuse([a, b])
    x, y = ops.defuse(flat, [a.shape, b.shape])
    assert torch.equal(x, a) and torch.equal(y, b)


def test_noise_scale_formula_matches_reference_definition():
    from kungfu_b200.torch.ops.monitor import NoiseScale
    ns = NoiseScale(alpha=0.6)
    g_small, g_big = torch.tensor([3.0, 4.0]), torch.tensor([1.0, 2.0])
    out = ns(32, 128, g_small, g_big)
    gs, gb = 25.0, 5.0
    G = (128 * gb - 32 * gs) / 96
    S = (gs - gb) / (1 / 32 - 1 / 128)
    assert abs(out - S / G) < 1e-9
    out2 = ns(32, 128, g_small * 2, g_big)
    gs2 = 100.0
    G2, S2 = (128 * gb - 32 * gs2) / 96, (gs2 - gb) / (1 / 32 - 1 / 128)
    assert abs(out2 - (0.6 * S + 0.4 * S2) / (0.6 * G + 0.4 * G2)) < 1e-9


def test_topology_helpers():
    from kungfu_b200.torch import ops
    edges = ops.minimum_spanning_tree([[0, 1, 9], [1, 0, 2], [9, 2, 0]])
    assert sorted(map(tuple, edges.tolist())) == [(0, 1), (1, 2)]
    assert ops.mst_to_tree(edges, 3) == [0, 0, 1]
    rr = ops.round_robin([False, True, True])
    assert [rr(), rr(), rr()] == [1, 2, 1]
    assert ops.peer_info() == (0, 1)


def test_policy_runner_and_variables():
    from kungfu_b200 import variables as V
    from kungfu_b200.torch.policy import BasePolicy, PolicyRunner

    class Count(BasePolicy):
        def __init__(self):
            self.log = []

        def before_train(self):
            self.log.append('bt')

        def before_epoch(self):
            self.log.append('be')

        def after_step(self, state):
            self.log.append('s%d' % state.step)

        def after_epoch(self, state):
            self.log.append('ae')

        def after_train(self):
            self.log.append('at')

    p = Count()
    runner = PolicyRunner([p], epoch_size=4, epoch_num=2, model=torch.nn.Linear(1, 1), batch_size=2)
    steps = 0
    while True:
        runner.before_step()
        steps += 1
        if not runner.after_step():
            break
    assert steps == 4
    assert p.log == ['bt', 'be', 's1', 's2', 'ae', 'be', 's3', 's4', 'ae', 'at']
    assert V.get_value(V.GraphKeys.TRAINED_SAMPLES) == 8
    assert V.eval_batch_size() == 2


def test_datasets():
    from kungfu_b200.datasets import ElasticShardedDataset, load_mnist, read_idx, write_idx
    x, y, xe, ye = load_mnist(n_train=64, n_test=16)
    assert x.shape == (64, 1, 28, 28) and ye.shape == (16, )
    # the same global stream whatever the sharding
    seen = {}
    for size in (1, 2, 4):
        got = []
        for r in range(size):
            d = ElasticShardedDataset(torch.arange(64), y, 8, r, size)
            got.append([d.next_batch()[0] for _ in range(3)])
        for s in range(3):
            seen.setdefault(s, []).append(sorted(torch.cat([g[s] for g in got]).tolist()))
    for s in range(3):
        assert all(v == seen[s][0] for v in seen[s])


def test_idx_roundtrip(tmp_path):
    from kungfu_b200.datasets import read_idx, write_idx
    a = np.arange(24, dtype=np.uint8).reshape(2, 3, 4)
    write_idx(str(tmp_path / 'a.idx'), a)
    assert np.array_equal(read_idx(str(tmp_path / 'a.idx')), a)


def test_cli_tools():
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin')
    r = subprocess.run([os.path.join(exe, 'kungfu-run'), '-help'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    for flag in ('-np', '-H', '-hostfile', '-strategy', '-w', '-config-server', '-builtin-config-port', '-allow-nvlink', '-port-range'):
        assert flag in r.stdout
    r = subprocess.run([os.path.join(exe, 'kungfu-distribute'), '-H', '10.0.0.1:1,10.0.0.2:1', '-dry-run', 'hostname'],
                       stdout=subprocess.PIPE, text=True)
    assert r.stdout.count('ssh ') == 2
    r = subprocess.run([sys.executable, '-m', 'kungfu_b200.info'], stdout=subprocess.PIPE, text=True, cwd=ROOT)
    assert 'host runtime' in r.stdout and '[built]' in r.stdout
    # python console-script shim -> C++ kungfu-run main
    r = subprocess.run([sys.executable, '-m', 'kungfu_b200.cmd', '-q', '-np', '2', sys.executable, '-c',
                        'import kungfu_b200.torch as kf; print("hello", kf.current_rank())'],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT,
                       env=dict(os.environ, PYTHONPATH=ROOT))
    assert r.returncode == 0 and r.stdout.count('hello') == 2, r.stdout


def test_examples_run():
    ex = os.path.join(ROOT, 'examples')
    res = run_workers(3, os.path.join(ex, 'torch_simple_example.py'))
    assert res.returncode == 0 and res.stdout.count('tensor([[3., 3., 3.]') == 3, res.stdout[-2000:]
    res = run_workers(2, os.path.join(ex, 'torch_mnist_example.py'), args=['--no-cuda', '--n-train', '512', '--epochs', '1'])
    assert res.returncode == 0 and 'test accuracy' in res.stdout, res.stdout[-2000:]
    res = run_workers(4, os.path.join(ex, 'adapt_strategy.py'), strategy='BINARY_TREE')
    assert res.returncode == 0 and 'throughput=' in res.stdout, res.stdout[-2000:]


def test_elastic_policy_example():
    import socket
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    res = run_workers(1, os.path.join(ROOT, 'examples', 'mnist_elastic_policy.py'), timeout=200,
                      extra_flags=['-w', '-builtin-config-port', str(port)])
    assert res.returncode == 0, res.stdout[-4000:]
    assert 'done at step 50' in res.stdout, res.stdout[-4000:]
    for n in (2, 3, 4):
        assert 'np %d ' % n in res.stdout


def test_monitoring_endpoint():
    """KUNGFU_CONFIG_ENABLE_MONITORING: Prometheus-style /metrics on port+10000
    (reference: kungfu-test-monitor, /root/reference/srcs/go/monitor/server.go:15-26)."""
    script = os.path.join(ROOT, 'tests', 'workers', 'monitor_worker.py')
    res = run_workers(2, script, env={'KUNGFU_CONFIG_ENABLE_MONITORING': 'true', 'KUNGFU_CONFIG_MONITORING_PERIOD': '30s'})
    assert res.returncode == 0, res.stdout[-3000:]
    assert res.stdout.count('OK monitor') == 2


def test_reference_tensorflow_layout_resolves_to_the_torch_front_end():
    """`kungfu.tensorflow.*` keeps the reference's module layout and names (backed by the torch objects)."""
    import torch

    import kungfu
    from kungfu.tensorflow.experimental.hook import ElasticHook, ResizeProfiler  # noqa: F401
    from kungfu.tensorflow.hooks import KungFuElasticTrainHook  # noqa: F401
    from kungfu.tensorflow.initializer import (BroadcastGlobalVariablesCallback, BroadcastGlobalVariablesHook,  # noqa: F401
                                               BroadcastGlobalVariablesOp, broadcast_variables)
    from kungfu.tensorflow.ops import (all_reduce, barrier, broadcast, group_all_reduce,  # noqa: F401
                                       group_hierarchical_nccl_all_reduce, group_nccl_all_reduce, resize)
    from kungfu.tensorflow.optimizers import (AdaptiveSGDOptimizer, MonitorGradientNoiseScaleOptimizer,  # noqa: F401
                                              MonitorGradientVarianceOptimizer, PairAveragingOptimizer,
                                              SynchronousAveragingOptimizer, SynchronousSGDOptimizer)
    from kungfu.tensorflow.policy import BasePolicy, PolicyHook  # noqa: F401
    from kungfu.tensorflow.v1.helpers import load_cifar10, load_mnist  # noqa: F401
    from kungfu.tensorflow.variables import GraphKeys, get_or_create_batch_size  # noqa: F401
    assert kungfu.tensorflow.optimizers is kungfu.torch.optimizers
    assert torch.equal(all_reduce(torch.ones(3)), torch.ones(3))
    got = group_nccl_all_reduce([torch.ones(2), None, torch.full((1, ), 3.0)])
    assert got[1] is None and float(got[2]) == 3.0
    m = torch.nn.Linear(2, 2)
    BroadcastGlobalVariablesOp(m)
    hook = ElasticHook(8, 1, 32, m)
    steps = 1
    while hook.after_step():
        steps += 1
    assert steps == 4 and hook.state.trained_samples == 32


def test_every_public_python_name_of_the_reference_exists_here():
    """Top-level public functions / classes of the reference's Python package, module by module, against the
    union of our front-end modules. What may be missing is listed explicitly: TensorFlow/Keras/Horovod-only
    plumbing that has no meaning outside a TF graph."""
    import ast
    import importlib
    ref = '/root/reference/srcs/python/kungfu'
    if not os.path.isdir(ref):
        pytest.skip('reference tree not present')
    mods = ['python', 'torch', 'torch.ops', 'torch.optimizers', 'torch.initializer', 'torch.policy', 'torch.elastic',
            'variables', 'datasets', 'utils', 'benchmarks', 'benchmarks.model_sizes', 'cmd', 'info', 'models',
            'torch.ops.collective', 'torch.ops.adapt', 'torch.ops.monitor', 'torch.ops.local', 'torch.ops.state',
            'torch.ops.topology', 'torch.examples.__main__']
    ours = set()
    for m in mods:
        ours |= set(dir(importlib.import_module('kungfu_b200.' + m)))
    tf_only = {
        'AdaSGDHook', 'KungFuTFOptimizer', 'KungFuTFKerasOptimizer', 'KungFuKerasOptimizer',  # TF optimizer base classes
        'hvd_init', 'hvd_group_all_reduce', 'get_cluster_size', 'get_rank', 'parse_args', 'log_detailed_result',
        'log_final_result', 'all_reduce_benchmark', 'main',  # benchmark script internals (Horovod arm)
        'new_conv_kernel', 'new_bias', 'new_dense_weight', 'conv2d', 'max_pool', 'Layer', 'seq_apply', 'Pool', 'Dense',
        'Conv',  # TF layer helpers of the benchmark models
        'write_idx_header', 'write_idx_to', 'read_idx_header', 'read_idx_from',  # file-object halves of read/write_idx_file
        'record_to_labeled_image', 'create_dataset_from_files',  # TFRecord parsing
        'training_step', 'test_gradient_tape',  # TF2 GradientTape demo
        'import_op_lib',  # loader of the TF/torch op .so (ours is kungfu_b200.ops._cuda.lib)
    }
    missing = {}
    for base, _, files in os.walk(ref):
        for f in files:
            if not f.endswith('.py'):
                continue
            path = os.path.join(base, f)
            for node in ast.parse(open(path).read()).body:
                if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and not node.name.startswith('_'):
                    if node.name not in ours and node.name not in tf_only:
                        missing.setdefault(os.path.relpath(path, ref), []).append(node.name)
    assert not missing, missing
