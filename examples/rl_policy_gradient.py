#!/usr/bin/env python
"""Data-parallel policy-gradient (REINFORCE) on a small built-in paddle game.

Every worker plays its own episodes (different seeds); the policy gradients are
averaged by SynchronousSGDOptimizer, so N workers see N times the experience
per update. Weights can be initialised from / saved to a numpy checkpoint,
rank 0's copy is broadcast at start.

    python examples/rl_policy_gradient.py --episodes 200
    kungfu-run -np 4 python examples/rl_policy_gradient.py --episodes 200

Reference: /root/reference/experimental/atari_pingpong/ (gym Pong, a policy
gradient agent with numpy-checkpoint init, `kungfu-run -np 4`). gym and the
Atari ROMs are not available offline, so the environment here is a
self-contained "catch" game with the same structure: an observation vector, 3
actions (left / stay / right), a reward at the end of each rally.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import numpy as np
import torch
import torch.nn as nn

import kungfu_b200.torch as kf
from kungfu_b200.utils import ExponentialMovingAverage


class Catch(object):
    """A ball falls through a W x H grid with a random horizontal drift; the
    paddle (bottom row) moves one cell per tick. +1 for a catch, -1 for a miss."""

    def __init__(self, width=9, height=10, seed=0):
        self.W, self.H = width, height
        self.rng = np.random.RandomState(seed)

    def reset(self):
        self.bx, self.by = int(self.rng.randint(self.W)), 0
        self.dx = int(self.rng.randint(-1, 2))
        self.px = self.W // 2
        return self._obs()

    def _obs(self):
        return np.array([self.bx / (self.W - 1), self.by / (self.H - 1), self.px / (self.W - 1), self.dx,
                         (self.bx - self.px) / (self.W - 1)], dtype=np.float32)

    def step(self, action):  # 0 left, 1 stay, 2 right
        self.px = min(self.W - 1, max(0, self.px + action - 1))
        self.by += 1
        nx = self.bx + self.dx
        if nx < 0 or nx >= self.W:
            self.dx = -self.dx
            nx = self.bx + self.dx
        self.bx = nx
        done = self.by == self.H - 1
        reward = 0.0
        if done:
            reward = 1.0 if self.bx == self.px else -1.0
        return self._obs(), reward, done


class Policy(nn.Module):
    def __init__(self, n_obs=5, hidden=32, n_act=3):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(n_obs, hidden), nn.Tanh(), nn.Linear(hidden, n_act))

    def forward(self, x):
        return torch.distributions.Categorical(logits=self.net(x))


def load_numpy_checkpoint(model, path):
    data = np.load(path)
    with torch.no_grad():
        for k, v in model.state_dict().items():
            v.copy_(torch.from_numpy(data[k]))


def save_numpy_checkpoint(model, path):
    np.savez(path, **{k: v.numpy() for k, v in model.state_dict().items()})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--episodes', type=int, default=300, help='updates (each over --batch rallies per worker)')
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--lr', type=float, default=0.01)
    ap.add_argument('--init', default='', help='numpy checkpoint (.npz) to start from')
    ap.add_argument('--save', default='', help='write a numpy checkpoint here (rank 0)')
    args = ap.parse_args()

    rank, np_ = kf.current_rank(), kf.current_cluster_size()
    torch.manual_seed(0)
    model = Policy()
    if args.init and rank == 0:
        load_numpy_checkpoint(model, args.init)
    opt = torch.optim.Adam(model.parameters(), lr=args.lr)
    opt = kf.optimizers.SynchronousSGDOptimizer(opt, model.named_parameters())
    kf.broadcast_parameters(model.state_dict())
    env = Catch(seed=1000 + rank)  # every worker sees different rallies
    torch.manual_seed(1234 + rank)  # ... and samples its own actions
    running = ExponentialMovingAverage(0.95)

    for it in range(args.episodes):
        logps, returns = [], []
        for _ in range(args.batch):
            obs, done, lp = env.reset(), False, []
            while not done:
                dist = model(torch.from_numpy(obs))
                a = dist.sample()
                lp.append(dist.log_prob(a))
                obs, reward, done = env.step(int(a))
            logps.append(torch.stack(lp).sum())
            returns.append(reward)
        ret = torch.tensor(returns)
        # baseline = the mean return over ALL workers' rallies of this update
        mean_ret = kf.ops.all_reduce_fn(ret.mean().reshape(1), name='mean-return')[0] / np_
        loss = -(torch.stack(logps) * (ret - mean_ret)).mean()
        opt.zero_grad()
        loss.backward()
        opt.step()  # gradients averaged over workers
        avg = running.update(float(mean_ret))
        if rank == 0 and (it % 25 == 0 or it == args.episodes - 1):
            print('update %4d  np %d  mean return %+.3f  running %+.3f' % (it, np_, float(mean_ret), avg), flush=True)

    digest = kf.ops.all_reduce_fn(torch.cat([p.detach().flatten() for p in model.parameters()]), 'max', name='digest')
    same = torch.equal(digest, torch.cat([p.detach().flatten() for p in model.parameters()]))
    if rank == 0:
        print('final running return %+.3f, replicas identical: %s' % (running.value, same), flush=True)
        if args.save:
            save_numpy_checkpoint(model, args.save)
    return 0 if same else 1


if __name__ == '__main__':
    sys.exit(main())
