import os
import time
import urllib.request

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf

rank, np_ = kf.current_rank(), kf.current_cluster_size()
for i in range(20):
    kf.ops.all_reduce_fn(torch.ones(200000), name='x%d' % i)
    time.sleep(0.01)
port = int(os.environ['KUNGFU_SELF_SPEC'].split(':')[1]) + 10000
text = urllib.request.urlopen('http://127.0.0.1:%d/metrics' % port).read().decode()
assert 'egress_total_bytes' in text and 'ingress_total_bytes' in text, text
rates = kfpy.egress_rates()
assert len(rates) == np_ and max(rates) > 0, rates
kf.run_barrier()
print('OK monitor', rank)
