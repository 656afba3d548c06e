"""Who is closer to float64: bs_actor_forward or torch's fp32 forward?  (completed_model actors)"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "deep-rl-battlespace_b200"))
from battlespace import TeamActors, parallel_env
dev = torch.device("cuda:0")
w = np.load(os.path.join(ROOT, "tests", "golden", "completed_model_actors.npz"))
KEYS = ("fc1.weight", "fc1.bias", "bn1.weight", "bn1.bias", "fc2.weight", "fc2.bias", "bn2.weight", "bn2.bias", "pi.weight", "pi.bias")
F = torch.nn.functional
def fwd(p, x, pre=False):
    x = torch.relu(F.layer_norm(F.linear(x, p["fc1.weight"], p["fc1.bias"]), (64,), p["bn1.weight"], p["bn1.bias"]))
    x = torch.relu(F.layer_norm(F.linear(x, p["fc2.weight"], p["fc2.bias"]), (64,), p["bn2.weight"], p["bn2.bias"]))
    z = F.linear(x, p["pi.weight"], p["pi.bias"])
    return z if pre else torch.tanh(z)
actors = TeamActors.from_npz(w, ["plane0", "plane1"], dev)
env = parallel_env(n_agents=2, n_envs=4096, seed=1)
obs = env.reset_batched()
for t in range(30):
    got = env.actor_actions(actors, team=0)
    obs, *_ = env.step_team_batched(got, 0, "instinct", auto_reset=True)
got = env.actor_actions(actors, team=0)
for j in range(2):
    p32 = {k: torch.as_tensor(w[f"plane{j}.{k}"], device=dev) for k in KEYS}
    p64 = {k: v.double() for k, v in p32.items()}
    t32, t64 = fwd(p32, obs[:, j]), fwd(p64, obs[:, j].double())
    print(j, "ours-f64", (got[:, j].double() - t64).abs().max().item(), "torch32-f64", (t32.double() - t64).abs().max().item(),
          "ours-torch32", (got[:, j] - t32).abs().max().item(), "max|logit|", fwd(p64, obs[:, j].double(), True).abs().max().item(),
          "max|w|", max(v.abs().max().item() for v in p32.values()))
