"""``kungfu-run -np 4 python -m kungfu_b200.torch.examples``

Three self-checking examples, the counterparts of ``python -m kungfu.tensorflow.v1.examples``
(/root/reference/srcs/python/kungfu/tensorflow/v1/examples/__main__.py:9-54): print who we are,
all-reduce a vector of ones, and run S-SGD on f(x) = x^2 where every step must give
x_k = (1 - 2 lr)^k to 1e-6 on every worker."""
import torch

import kungfu_b200.torch as kf


def show_info_example():
    print('rank %d of %d (local %d of %d), uid %x' % (kf.current_rank(), kf.current_cluster_size(),
                                                      kf.current_local_rank(), kf.current_local_size(), kf.uid()))


def all_reduce_example():
    np_ = kf.current_cluster_size()
    y = kf.ops.all_reduce_fn(torch.ones(10), name='ones')
    assert torch.equal(y, torch.full((10, ), float(np_))), y
    print('all_reduce of ones over %d workers = %s' % (np_, y[:3].tolist()))


def sgd_example(lr=0.1, steps=10):
    x = torch.nn.Parameter(torch.ones(()))
    opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD([x], lr=lr), named_parameters=[('x', x)])
    kf.broadcast_parameters({'x': x.data})
    for k in range(1, steps + 1):
        opt.zero_grad()
        (x * x).backward()
        opt.step()
        assert abs(float(x.detach()) - (1 - 2 * lr) ** k) < 1e-6, (k, float(x.detach()))
    print('S-SGD on x^2: x_%d = %.6f = (1 - 2 lr)^%d' % (steps, float(x.detach()), steps))


def main():
    show_info_example()
    all_reduce_example()
    sgd_example()


if __name__ == '__main__':
    main()
