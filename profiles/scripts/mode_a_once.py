#!/usr/bin/env python
"""A few single-tick bs_step launches at E envs (3v3 continuous, both teams' actions from HBM): the short command the
ncu capture of step_kernel runs on (Mode A; at 1,048,576 envs the 1.1 GB working set cannot sit in L2).
usage: mode_a_once.py [E]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "deep-rl-battlespace_b200"))
from battlespace import parallel_env  # noqa: E402
E = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
CF = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03, lose_punishment=-0.05)
env = parallel_env(n_agents=3, continuous_actions=True, n_envs=E, seed=0, device="cuda:0", **CF)
acts = [torch.empty((E, 6, 3), device="cuda:0").uniform_(-1, 1) for _ in range(2)]
env.reset_batched()
for k in range(12):
    env.step_batched(acts[k % 2])
torch.cuda.synchronize()
print("ok", int(env.stats()[6]))
