"""python -m kungfu_b200.info: versions of what this build runs on.
Reference: /root/reference/srcs/python/kungfu/info/__main__.py."""
import os

import torch

import kungfu_b200
from kungfu_b200 import _host
from kungfu_b200.ops import _cuda


def main():
    print('kungfu_b200 %s' % kungfu_b200.__version__)
    print('torch %s (CUDA %s)' % (torch.__version__, torch.version.cuda))
    print('host runtime: %s [%s]' % (_host.LIB_PATH, 'built' if _host.available() else 'MISSING'))
    print('device library: %s [%s]' % (_cuda.LIB_PATH, 'built' if _cuda.available() else 'MISSING'))
    if torch.cuda.is_available():
        for i in range(torch.cuda.device_count()):
            p = torch.cuda.get_device_properties(i)
            mc = _cuda.lib().kfb_multicast_supported(i) if _cuda.available() else '?'
            print('GPU %d: %s, %d SMs, cc %d.%d, %.0f GiB, NVLS multicast: %s' %
                  (i, p.name, p.multi_processor_count, p.major, p.minor, p.total_memory / 2**30, mc))
    else:
        print('no CUDA device visible')
    try:
        print('NCCL (baseline only): %s' % '.'.join(map(str, torch.cuda.nccl.version())))
    except Exception:
        pass
    for k, v in sorted(os.environ.items()):
        if k.startswith('KUNGFU_'):
            print('%s=%s' % (k, v))


if __name__ == '__main__':
    main()
