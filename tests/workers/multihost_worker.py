"""Several 'hosts' emulated on loopback aliases (127.0.0.x), one kungfu-run per
host: hierarchical strategies, local / cross collectives, host-aware ranks.
(The reference emulates this with docker-compose networks)"""
import ctypes as C

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf
from kungfu_b200 import _host as H

r, n = kf.current_rank(), kf.current_cluster_size()
p = kfpy.peer()
x = torch.full((300000, ), float(r + 1))
assert torch.all(kf.ops.all_reduce_fn(x) == n * (n + 1) / 2)
ones, out = torch.ones(10), torch.zeros(10)
L = H.lib()
# local reduce -> cross all-reduce among host masters -> local broadcast == global sum
assert L.kf_local_reduce(p._h, ones.data_ptr(), out.data_ptr(), 10, H.F32, 0, b'lr') == 0
if kfpy.current_local_rank() == 0:
    assert float(out[0]) == kfpy.current_local_size()
part = out.clone() if kfpy.current_local_rank() == 0 else torch.zeros(10)
tot = torch.zeros(10)
assert L.kf_cross_all_reduce(p._h, part.data_ptr(), tot.data_ptr(), 10, H.F32, 0, b'cross') == 0
res = torch.zeros(10)
assert L.kf_local_broadcast(p._h, tot.data_ptr(), res.data_ptr(), 10, H.F32, b'lb') == 0
assert float(res[0]) == n, (float(res[0]), n)
g = kf.ops.all_gather(torch.tensor([kfpy.current_local_rank(), kfpy.current_local_size(), kfpy.host_count()]))
import os
hosts = len({peer.split(':')[0] for peer in os.environ['KUNGFU_INIT_PEERS'].split(',')})
assert hosts > 1 and g[:, 2].tolist() == [hosts] * n
kf.run_barrier()
print('OK multihost rank %d/%d local %d/%d hosts %d' % (r, n, kfpy.current_local_rank(), kfpy.current_local_size(), kfpy.host_count()))
