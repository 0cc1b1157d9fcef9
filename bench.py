#!/usr/bin/env python
"""Headline benchmark: ResNet-50 S-SGD training throughput (img/s), bf16,
synthetic ImageNet-shaped data, 1..8 B200 (one process per GPU).

    python bench.py --gpus 1 --steps 30 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 8 --steps 30 --warmup 5

Protocol mirrors the reference's synthetic benchmark
(/root/reference/benchmarks/system/benchmark_kungfu_torch.py:61-135: fixed fake
batch, SGD lr 0.01 x np, optimizer wrapped by SynchronousSGDOptimizer) with the
timing rules of this project: device-timed with CUDA events, barrier +
synchronize on both sides, max over ranks, >= 3 warm-up steps, clocks sampled
during the timed region. The per-GPU batch (256) is the reference's largest
published point (global batch 4096 over 16 GPUs).

`value`  : device-timed steps on resident synthetic data (the reference's protocol)
`e2e`    : same steps, but every step copies its batch from pinned host memory
           and reads the loss back to the host.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASELINE_IMG_S = 5507.0  # KungFu S-SGD, ResNet-50, global batch 4096, 16 x V100
BASELINE_NOTE = ('vs_baseline divides by the only published figure (KungFu S-SGD over TCP, TensorFlow fp32, global batch '
                 '4096 on 16 x V100 = 5507 img/s): other hardware, precision and framework, context only; the same-box '
                 'comparison is the `same_box_baseline` block (stock PyTorch DDP over NCCL on these GPUs)')


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=30)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--model', default='resnet50')
    ap.add_argument('--batch-size', type=int, default=256, help='per-GPU batch')
    ap.add_argument('--image-size', type=int, default=224)
    ap.add_argument('--no-graph', action='store_true', help='eager launches instead of a captured CUDA graph')
    ap.add_argument('--fused-bn', type=int, default=int(os.environ.get('KUNGFU_FUSED_BN', '1')))
    ap.add_argument('--s2d-stem', type=int, default=int(os.environ.get('KUNGFU_S2D_STEM', '1')),
                    help='evaluate the 7x7/2 stem as a 4x4 conv over a space-to-depth input (same maths)')
    ap.add_argument('--momentum', type=float, default=0.0)
    ap.add_argument('--skip-e2e', action='store_true')
    ap.add_argument('--overlap', type=int, default=int(os.environ.get('KUNGFU_BENCH_OVERLAP', '0')),
                    help='launch the fused exchange+update bucket by bucket from autograd hooks, under backward '
                         '(measured on 2 and 8 B200: within +-0.3 %% of the single launch, see profiles/overlap_ab_b200.txt)')
    ap.add_argument('--same-box-baseline', default=os.environ.get('KUNGFU_BENCH_BASELINE', 'ddp'),
                    choices=['ddp', 'none'],
                    help='after the timed region, also time stock PyTorch (torchvision-style ResNet-50, torch BN, '
                         'torch.optim.SGD, DistributedDataParallel over NCCL) with the same protocol')
    ap.add_argument('--baseline-steps', type=int, default=10)
    ap.add_argument('--clock-interval-ms', type=int, default=100)
    ap.add_argument('--trace', action='store_true',
                    help='after the measurement, replay 4 steps under torch.profiler and print where the communication '
                         'kernels sit on the timeline of rank 0 (diagnostic; never part of a reported number)')
    return ap.parse_args()


class ClockSampler(object):
    """SM clock and throttle reasons of ONE device during the timed region, read
    through NVML in this process (a thread, one device handle). The earlier
    `nvidia-smi -lms 100` child enumerated every GPU of the box ten times a
    second; with several ranks that is N x 10 whole-box queries per second next
    to a measurement that waits for the slowest GPU every step."""
    REASONS = (('hw_slowdown', 0x8), ('hw_thermal_slowdown', 0x40), ('sw_thermal_slowdown', 0x20),
               ('sw_power_cap', 0x4))

    def __init__(self, index, interval_s=0.1):
        self.index = index
        self.interval = interval_s
        self.rows = []
        self.th = None
        self._stop = threading.Event()
        self.err = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get('CUDA_VISIBLE_DEVICES')
            idx = self.index
            if vis:
                try:
                    idx = int(vis.split(',')[self.index])
                except ValueError:
                    idx = self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:
            self.err = '%s: %s' % (type(e).__name__, e)
            return
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((mhz, mask))
            except Exception as e:
                self.err = '%s: %s' % (type(e).__name__, e)
                return
            self._stop.wait(self.interval)

    def stop(self):
        if self.th is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['NVML unavailable: %s' % self.err]}
        self._stop.set()
        self.th.join(timeout=5)
        sm = [r[0] for r in self.rows if r[0] > 0]
        reasons = sorted({name for _, mask in self.rows for name, bit in self.REASONS if mask & bit})
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': self.max_mhz, 'reasons': reasons,
                'samples': len(self.rows), 'how': 'NVML in-process, every %d ms' % int(self.interval * 1e3)}


def reference_arm(args):
    # The one allowed offline install was attempted ("Reference
    # arm"): the reference's build needs Go >= 1.11 for its runtime and
    # TensorFlow headers for its op library; neither exists in this image.
    if os.path.isdir(os.path.join(ROOT, 'baseline', '_ref', 'kungfu')):
        pass  # would run the unmodified reference here; it never installed
    print(json.dumps({'impl': 'reference',
                      'unavailable': 'reference does not build offline: setup.py needs TensorFlow '
                                     '(import tensorflow) and a Go toolchain for libkungfu-comm; neither is in the image'}))
    return 0


def correctness_check(kf, K, opt, dev, rank, np_):
    """Untimed, after the measurement: (1) every rank holds bit-identical
    parameters (device-side byte consensus over the whole parameter arena);
    (2) one fused exchange+update step on a 1 Mi-element bf16 buffer against a
    host-computed fp32 result. The reference's NCCL trainer swept int32 sums
    before timing (/root/reference/tests/cpp/integration/fake_nccl_trainer.cpp:34-63)."""
    import torch

    from kungfu_b200.ops import collective as coll
    from kungfu_b200.optim import FusedSGD
    comm = opt.comm
    out = {'np': np_, 'algo': opt.last_algo, 'multicast': bool(getattr(comm, 'multicast_ok', False))}
    identical = True
    if np_ > 1:
        agree = torch.zeros(1, dtype=torch.int32, device=dev)
        for _, arenas in opt.arenas():
            for a in arenas:
                coll.consensus(comm, a.params_buf, a.nbytes, agree)
                torch.cuda.synchronize()
                identical = identical and bool(int(agree))
    out['params_identical'] = identical
    n = 1 << 20
    w = torch.nn.Parameter(torch.zeros(n, device=dev, dtype=torch.bfloat16))
    chk = kf.optimizers.SynchronousSGDOptimizer(FusedSGD([w], lr=1.0), named_parameters=[('check', w)])
    idx = torch.arange(n, device=dev)
    w.grad = ((idx % 7) + rank).to(torch.bfloat16)
    chk.step()
    torch.cuda.synchronize()
    # mean over ranks of (i % 7 + r) = i % 7 + (np - 1) / 2: exact in bf16, so the only acceptable error is 0
    want = -((idx % 7).float() + (np_ - 1) / 2.0)
    out['fused_step_max_err'] = float((w.detach().float() - want).abs().max())
    out['fused_step_algo'] = chk.last_algo
    return out


def trace_steps(run_step, rank, kf):
    import tempfile

    import torch
    kf.run_barrier()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        for _ in range(4):
            run_step()
        torch.cuda.synchronize()
    kf.run_barrier()
    path = os.path.join(tempfile.gettempdir(), 'kf_trace_%d.json' % rank)
    prof.export_chrome_trace(path)
    ev = [e for e in json.load(open(path))['traceEvents'] if e.get('cat') == 'kernel']
    ev.sort(key=lambda e: e['ts'])
    if rank != 0 or not ev:
        return
    # steps are separated by the stem convolution's first kernel; use the fused-step kernels as anchors instead:
    # report, per occurrence of each of our communication / optimizer kernels, start offset and duration
    ours = [e for e in ev if 'ssgd_fused' in e['name'] or 'k_fuse' in e['name'] or 'k_sum_slots' in e['name']]
    t0 = ev[0]['ts']
    span = ev[-1]['ts'] + ev[-1]['dur'] - t0
    busy = sum(e['dur'] for e in ev)
    print('[trace] 4 steps: span %.3f ms/step, sum of kernel durations %.3f ms/step, %d kernels/step' %
          (span / 4e3, busy / 4e3, len(ev) // 4), file=sys.stderr)
    agg = {}
    for e in ev:
        k = e['name'][:120]
        t = agg.setdefault(k, [0.0, 0])
        t[0] += e['dur']
        t[1] += 1
    for name, (dur, cnt) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(os.environ.get('KUNGFU_TRACE_TOP', '0'))]:
        print('[trace-top] %8.3f ms/step %5d x  %s' % (dur / 4e3, cnt // 4, name), file=sys.stderr)
    last_model = None
    for e in ev:
        if e in ours:
            name = e['name'].split('<')[0].split('(')[0][-28:]
            gap = (e['ts'] - (last_model['ts'] + last_model['dur'])) if last_model else 0.0
            print('[trace] %-28s start %9.3f ms  dur %8.1f us  (%.1f us after the previous model kernel ended)' %
                  (name, (e['ts'] - t0) / 1e3, e['dur'], gap), file=sys.stderr)
        else:
            last_model = e


def ddp_baseline(args, kf, dev, rank, np_, local_rank):
    """Stock PyTorch on the same GPUs, same protocol: plain ResNet-50 (torch
    BatchNorm / ReLU / pooling, cuDNN convolutions), bf16 channels-last,
    torch.optim.SGD, DistributedDataParallel over NCCL when N > 1. None of this
    repo's kernels or optimizers are on this path."""
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F

    from kungfu_b200 import models
    torch.manual_seed(1234)
    B, S = args.batch_size, args.image_size
    model = models.resnet50(fused_bn=False, s2d_stem=False).to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
    model.train()
    net = model
    if np_ > 1:
        if not dist.is_initialized():
            if 'MASTER_ADDR' in os.environ and 'RANK' in os.environ:
                dist.init_process_group('nccl', device_id=dev)
            else:
                dist.init_process_group('nccl', init_method='tcp://127.0.0.1:%d' % (29200 + np_), rank=rank,
                                        world_size=np_, device_id=dev)
        net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], gradient_as_bucket_view=True)
    opt = torch.optim.SGD(model.parameters(), lr=0.01 * np_)
    x = torch.randn(B, 3, S, S, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (B, ), device=dev)

    def step():
        opt.zero_grad(set_to_none=True)
        F.cross_entropy(net(x).float(), y).backward()
        opt.step()

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    kf.run_barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.baseline_steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    kf.run_barrier()
    ms = float(kf.ops.all_reduce_fn(torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64), 'max')[0])
    ms /= args.baseline_steps
    if dist.is_initialized():
        dist.destroy_process_group()
    return {'impl': 'torch DDP (NCCL) + torch.optim.SGD, stock torch BN/ReLU, eager' if np_ > 1 else
            'torch.optim.SGD, stock torch BN/ReLU, eager (single GPU)',
            'value': B * np_ / (ms / 1e3), 'unit': 'img/s', 'ms_per_step': ms, 'steps': args.baseline_steps, 'warmup': 5}


def main():
    args = parse_args()
    if args.impl == 'reference':
        return reference_arm(args)

    import torch
    import torch.nn.functional as F

    import kungfu_b200.torch as kf
    from kungfu_b200 import models
    from kungfu_b200.ops import _cuda as K
    from kungfu_b200.optim import FusedSGD

    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    rank, np_ = kf.current_rank(), kf.current_cluster_size()
    if np_ != args.gpus and rank == 0:
        print('[bench] warning: --gpus %d but cluster size is %d' % (args.gpus, np_), file=sys.stderr)
    torch.backends.cudnn.benchmark = True

    torch.manual_seed(1234)
    model = getattr(models, args.model)(fused_bn=bool(args.fused_bn), s2d_stem=bool(args.s2d_stem))
    model = models.to_b200_layout(model.to(dev), torch.bfloat16)
    model.train()
    opt = FusedSGD(model.parameters(), lr=0.01 * np_, momentum=args.momentum)
    opt = kf.optimizers.SynchronousSGDOptimizer(opt, named_parameters=model.named_parameters(),
                                                overlap=bool(args.overlap))
    opt.arenas()  # move the parameters into symmetric memory before anything captures their addresses
    kf.broadcast_parameters(model.state_dict())

    B, S = args.batch_size, args.image_size
    # resident synthetic batch (the reference's "fixed fake data")
    x_dev = torch.randn(B, 3, S, S, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y_dev = torch.randint(0, 1000, (B, ), device=dev)
    loss_dev = torch.zeros((), device=dev)

    def train_step():
        opt.zero_grad(set_to_none=True)
        out = model(x_dev)
        loss = F.cross_entropy(out.float(), y_dev)
        loss.backward()
        opt.step()
        loss_dev.copy_(loss.detach())

    # ---- warm-up (eager; also cuDNN autotune and first-step momentum init) ---------------
    eager_warm = max(3, args.warmup) if args.no_graph else 3
    for _ in range(eager_warm):
        train_step()
    torch.cuda.synchronize()

    launches_per_step = None
    graph_holder = []
    if args.no_graph:
        run_step = train_step
    else:
        kf.run_barrier()
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            train_step()  # one more eager step on the capture stream's allocator state
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        kf.run_barrier()
        before = K.launch_count
        with torch.cuda.graph(graph):
            train_step()
        launches_per_step = K.launch_count - before
        graph_holder.append(graph)
        run_step = graph.replay
        for _ in range(max(0, args.warmup - 3) + 3):
            run_step()
        torch.cuda.synchronize()

    def timed(fn, steps):
        kf.run_barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        before = K.launch_count
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        kf.run_barrier()
        ms = e0.elapsed_time(e1)
        mx = kf.ops.all_reduce_fn(torch.tensor([ms], dtype=torch.float64), 'max')
        return float(mx[0]), K.launch_count - before

    sampler = ClockSampler(local_rank, args.clock_interval_ms / 1e3)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    total_ms, eager_launches = timed(run_step, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = total_ms / args.steps
    value = B * np_ / (ms_per_step / 1e3)
    gpu_launches = eager_launches if args.no_graph else launches_per_step * args.steps

    # ---- end to end through the public API: pinned host batch in, loss out, every step ----
    e2e = None
    if not args.skip_e2e:
        x_host = torch.randn(B, 3, S, S).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).pin_memory()
        y_host = torch.randint(0, 1000, (B, )).pin_memory()
        loss_host = torch.zeros(()).pin_memory()
        copy_stream = torch.cuda.Stream()
        x_next = torch.empty_like(x_dev)
        y_next = torch.empty_like(y_dev)
        ready = torch.cuda.Event()
        consumed = torch.cuda.Event()
        consumed.record()

        def prefetch():
            # next batch travels over PCIe on its own stream while this step computes
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed)
                x_next.copy_(x_host, non_blocking=True)
                y_next.copy_(y_host, non_blocking=True)
                ready.record(copy_stream)

        prefetch()

        def e2e_step():
            cur = torch.cuda.current_stream()
            cur.wait_event(ready)
            x_dev.copy_(x_next, non_blocking=True)  # device-side hand-over into the graph's static input
            y_dev.copy_(y_next, non_blocking=True)
            consumed.record(cur)
            prefetch()
            run_step()
            loss_host.copy_(loss_dev, non_blocking=True)
            cur.synchronize()  # the loss is on the host when the step returns

        for _ in range(3):
            e2e_step()
        e2e_ms, _ = timed(e2e_step, args.steps)
        e2e = {'value': B * np_ / (e2e_ms / args.steps / 1e3), 'unit': 'img/s',
               'h2d_bytes_per_step': x_host.numel() * x_host.element_size() + y_host.numel() * y_host.element_size(),
               'd2h_bytes_per_step': loss_host.element_size(), 'ms_per_step': e2e_ms / args.steps,
               'loss': float(loss_host)}

    if args.trace:
        trace_steps(run_step, rank, kf)
    check = correctness_check(kf, K, opt, dev, rank, np_)
    baseline = None
    if args.same_box_baseline == 'ddp':
        del graph_holder[:]
        try:
            baseline = ddp_baseline(args, kf, dev, rank, np_, local_rank)
        except Exception as e:  # never lose the bench line to the comparison arm
            baseline = {'impl': 'torch DDP (NCCL)', 'error': '%s: %s' % (type(e).__name__, str(e)[:200])}

    if rank == 0:
        out = {
            'metric': 'ResNet-50 S-SGD training throughput',
            'value': value, 'unit': 'img/s', 'n_gpus': np_, 'steps': args.steps, 'warmup': max(args.warmup, 3),
            'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': value / BASELINE_IMG_S, 'dtype': 'bf16', 'data': 'synthetic', 'impl': 'ours',
            'config': {'model': args.model, 'global_batch': B * np_, 'per_gpu_batch': B, 'seq_len': None,
                       'image_size': S, 'parallelism': 'dp%d' % np_, 'optimizer': 'SynchronousSGDOptimizer(FusedSGD)',
                       'cuda_graph': not args.no_graph, 'fused_bn': bool(args.fused_bn), 's2d_stem': bool(args.s2d_stem),
                       'overlap': bool(args.overlap), 'buckets': len(opt._ov['buckets']) if opt._ov else 1,
                       'fused_step': opt.last_algo,
                       'l2': 'working set per step (activations, bs%d) exceeds the 126 MB L2' % B,
                       'baseline': 'KungFu S-SGD ResNet-50 global batch 4096 on 16xV100 = 5507 img/s'},
            'clocks': clocks, 'e2e': e2e, 'gpu_launches': gpu_launches, 'loss': float(loss_dev),
            'check': check, 'same_box_baseline': baseline, 'vs_baseline_note': BASELINE_NOTE,
        }
        print(json.dumps(out))
    kf.run_barrier()
    return 0


if __name__ == '__main__':
    sys.exit(main())
