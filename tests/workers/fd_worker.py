"""Descriptor passing between colocated workers (what carries the CUDA VMM handles of every symmetric
buffer): every rank creates a file, sends its descriptor to every other rank over the unix-socket control
plane (SCM_RIGHTS), and reads everybody else's file through the descriptors it received. Twice, with
different tags, and once as a broadcast from rank 1."""
import os
import tempfile

import kungfu_b200.python as kfpy
from kungfu_b200 import _host

rank, np_ = kfpy.current_rank(), kfpy.current_cluster_size()
x = _host.Exchange(kfpy.peer())
for rnd in range(2):
    f = tempfile.TemporaryFile()
    f.write(('round %d from rank %d' % (rnd, rank)).encode())
    f.flush()
    fds = x.exchange_fds(f.fileno(), 'test-fd-%d' % rnd)
    for r in range(np_):
        if r == rank:
            assert fds[r] is None
            continue
        assert fds[r] is not None and fds[r] >= 0 and fds[r] != f.fileno()
        got = os.pread(fds[r], 100, 0).decode()
        assert got == 'round %d from rank %d' % (rnd, r), (r, got)
        os.close(fds[r])
    x.barrier()
    f.close()
if np_ > 1:
    g = tempfile.TemporaryFile()
    g.write(b'broadcast payload')
    g.flush()
    fd = x.broadcast_fd(g.fileno(), 'test-bcast-fd', root=1)
    assert os.pread(fd, 100, 0) == (b'broadcast payload' if True else b'')
    x.barrier()
print('OK fd passing rank %d of %d' % (rank, np_))
