#!/usr/bin/env python
"""A few bs_actor_forward launches at config 5's size (3v3 discrete, 32,768 envs = 98,304 rows): the short command
the ncu capture of actor_kernel runs on; prints the live CUDA-event time per launch."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "deep-rl-battlespace_b200"))
from battlespace import TeamActors, parallel_env  # noqa: E402
N, E, A, L = 3, 32768, 4, 11
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
u = lambda *s, a=1.0: ((torch.rand(*s, generator=g) * 2 - 1) * a)
params = [{"fc1.weight": u(64, L, a=.125), "fc1.bias": u(64, a=.125), "bn1.weight": torch.ones(64), "bn1.bias": torch.zeros(64),
           "fc2.weight": u(64, 64, a=.125), "fc2.bias": u(64, a=.125), "bn2.weight": torch.ones(64), "bn2.bias": torch.zeros(64),
           "pi.weight": u(A, 64, a=.003), "pi.bias": u(A, a=.003)} for _ in range(N)]
actors = TeamActors(params, dev)
env = parallel_env(n_agents=N, n_envs=E, seed=0, device=dev)
env.reset_batched()
out = torch.empty((E, N, A), device=dev)
for _ in range(6):
    env.actor_actions(actors, team=0, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    env.actor_actions(actors, team=0, out=out)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 200 * 1e3
flops = 2.0 * E * N * (L * 64 + 64 * 64 + 64 * A)
print(f"actor_kernel {us:.2f} us per launch, {flops / us / 1e6:.2f} TFLOP/s fp32 = {flops / us / 1e6 / (148 * 128 * 2 * 1.965e-3):.3f} of the FFMA peak")
