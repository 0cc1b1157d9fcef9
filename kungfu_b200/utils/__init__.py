"""Small helpers shared by examples, benchmarks and policies."""


class ExponentialMovingAverage(object):
    """v <- alpha v + (1 - alpha) x, first sample passes through
    (/root/reference/srcs/cpp/include/kungfu/utils/ema.hpp:19-28,
    /root/reference/srcs/python/kungfu/tensorflow/v1/utils/ema.py)."""

    def __init__(self, alpha):
        self.alpha = alpha
        self.value = None

    def update(self, x):
        self.value = x if self.value is None else self.alpha * self.value + (1 - self.alpha) * x
        return self.value


def step_based_schedule(config, step, default=1, strict=True):
    """Cluster size for `step` from a schedule string "size:steps,size:steps,...".
    Past the end: `default` (strict=False) or the last size is kept... the
    reference's op (/root/reference/srcs/cpp/src/tensorflow/ops/cpu/elastic.cpp:16-81)
    returns the default when the schedule is exhausted and errors when strict."""
    off = 0
    for part in config.split(','):
        part = part.strip()
        if not part:
            continue
        size, n = part.split(':')
        off += int(n)
        if step < off:
            return int(size)
    if strict:
        raise ValueError('step %d is beyond the schedule %r' % (step, config))
    return default


def list_nvidia_gpu_names():
    """Names of the visible NVIDIA GPUs, [] when there is no driver
    (/root/reference/srcs/go/utils/nvidia.go: `nvidia-smi -L` parser)."""
    import subprocess
    try:
        out = subprocess.run(['nvidia-smi', '-L'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
                             timeout=20).stdout
    except Exception:
        return []
    names = []
    for line in out.splitlines():
        # "GPU 0: NVIDIA B200 (UUID: GPU-...)"
        if line.startswith('GPU ') and ':' in line:
            names.append(line.split(':', 1)[1].split('(UUID')[0].strip())
    return names


def measure(f, *args, **kw):
    """(result, seconds) of one call (/root/reference/srcs/go/utils/utils.go Measure)."""
    import time
    t0 = time.perf_counter()
    r = f(*args, **kw)
    return r, time.perf_counter() - t0


def show_rate(nbytes, seconds):
    """Human-readable data rate (/root/reference/srcs/go/utils/utils.go ShowRate)."""
    rate = nbytes / max(seconds, 1e-12)
    for unit in ('B/s', 'KiB/s', 'MiB/s', 'GiB/s', 'TiB/s'):
        if rate < 1024 or unit == 'TiB/s':
            return '%.2f %s' % (rate, unit)
        rate /= 1024


def show_size(nbytes):
    for unit in ('B', 'KiB', 'MiB', 'GiB', 'TiB'):
        if nbytes < 1024 or unit == 'TiB':
            return '%.1f %s' % (nbytes, unit) if unit != 'B' else '%d B' % nbytes
        nbytes /= 1024.0


def poll(cond, timeout=10.0, interval=0.05):
    """Wait until cond() is truthy; returns its value or None on timeout
    (/root/reference/srcs/go/utils/utils.go Poll)."""
    import time
    end = time.monotonic() + timeout
    while True:
        v = cond()
        if v:
            return v
        if time.monotonic() >= end:
            return None
        time.sleep(interval)


class StallDetector(object):
    """Context manager that prints "<name> stalled for N s" every `period`
    seconds while the body has not finished (the runtime has the same watchdog
    around its blocking calls; /root/reference/srcs/go/utils/stalldetector.go:15-46)."""

    def __init__(self, name, period=3.0, out=None):
        import sys
        self.name, self.period, self.out = name, period, out or sys.stderr
        self.reports = 0

    def __enter__(self):
        import threading
        import time
        self._done = threading.Event()
        t0 = time.monotonic()

        def watch():
            while not self._done.wait(self.period):
                self.reports += 1
                print('%s stalled for %.0f s' % (self.name, time.monotonic() - t0), file=self.out, flush=True)

        self._th = threading.Thread(target=watch, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *exc):
        self._done.set()
        self._th.join()
        return False


def map_maybe(f, lst):
    """f over the non-None entries, None stays None (reference: kungfu/_utils.py:5-6)."""
    return [f(x) if x is not None else None for x in lst]


def one_based_range(n):
    return range(1, 1 + n)


def show_duration(seconds):
    """'12.00ms' / '3.20s' / '2m5s' / '1h2m3s' (reference: kungfu/_utils.py:20-30)."""
    if seconds < 1:
        return '%.2fms' % (seconds * 1e3)
    if seconds < 60:
        return '%.2fs' % seconds
    sec = int(seconds)
    mm, ss = sec // 60, sec % 60
    if seconds < 3600:
        return '%dm%ds' % (mm, ss)
    return '%dh%dm%ds' % (mm // 60, mm % 60, ss)


def since_job_start():
    """Seconds since kungfu-run started the job (KUNGFU_JOB_START_TIMESTAMP; reference: kungfu/_utils.py:33-36)."""
    import os
    import time
    return time.time() - int(os.getenv('KUNGFU_JOB_START_TIMESTAMP') or '0')


def since_proc_start():
    """Seconds since the runner spawned this worker (KUNGFU_PROC_START_TIMESTAMP; reference: kungfu/_utils.py:39-42)."""
    import os
    import time
    return time.time() - int(os.getenv('KUNGFU_PROC_START_TIMESTAMP') or '0')


def log_event(name):
    """One 'TS=<unix time> <name> :: <duration> since proc started' line, the format the scaling benchmark's
    log scraper expects (reference: kungfu/_utils.py:44-50)."""
    import time
    line = 'TS=%f %s :: %s since proc started' % (time.time(), name, show_duration(since_proc_start()))
    print(line, flush=True)
    return line


_since_job_start, _since_proc_start, _log_event = since_job_start, since_proc_start, log_event


class EMA(ExponentialMovingAverage):
    """Moving average whose input is clamped to [value / scale_cap, value * scale_cap] first: one wild
    reading (a noise-scale spike) cannot move the average by more than that factor
    (reference: kungfu/tensorflow/v1/utils/ema.py)."""

    def __init__(self, alpha, scale_cap=None):
        super().__init__(alpha)
        self.scale_cap = scale_cap

    def update(self, x):
        if self.value is not None and self.scale_cap is not None:
            lo, hi = sorted((self.value / self.scale_cap, self.value * self.scale_cap))
            x = min(max(x, lo), hi)
        return super().update(x)

    def get(self):
        return self.value

    def reset(self):
        self.value = None
