"""Is a T=1 fused rollout launch (dense bullet pool + ring<->pool conversion) cheaper than step_kernel?"""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "deep-rl-battlespace_b200"))
from battlespace import DeviceReplay, parallel_env
dev = torch.device("cuda:0")
CF = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03, lose_punishment=-0.05)
def timed(fn, iters=300, warm=10):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
for cont in (False, True):
    N, E, A = 3, 32768, 3 if cont else 4
    env = parallel_env(n_agents=N, continuous_actions=cont, n_envs=E, seed=0, device=dev, **CF)
    rp = DeviceReplay(1 << 20, 512, N, 11, A, dev)
    env.rollout(64, "random", "instinct")
    act = torch.empty((1, E, N, A), device=dev).uniform_(-1, 1)
    out = {"cont": cont}
    for T in (1, 2, 4, 32):
        acts = act.expand(T, E, N, A).contiguous()
        out[f"rollout_T{T}_ms_per_tick"] = timed(lambda: env.rollout(T, "external", "instinct", 0, acts, rp)) / T
    out["step_team_replay_autoreset_ms"] = timed(lambda: env.step_team_batched(act[0], 0, "instinct", replay=rp, auto_reset=True))
    print(json.dumps(out))
