"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz from the unmodified reference.

Run in the build container (needs /root/reference):  python oracle/gen_golden.py
Each fixture is one ``run_reference`` trace (oracle/ref_harness.py): spawn draws,
per-step actions and unit draws, and everything the reference produced (float32
observations, rewards, dones, winner, plane / base / bullet state).  The shipped
2v2 actor weights (models/completed_model/actor_plane{0,1}, data not code) are
exported as plain arrays for the config-2 statistical test.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
KEYS = ("hit_base_reward", "hit_plane_reward", "miss_punishment", "die_punishment", "lose_punishment")
TRAINED = dict(zip(KEYS, (1.0, 0.9, -0.02, -0.03, -0.05)))   # main.py:33-37 / cf.json
DEFAULT = {}                                                  # battle_env.py:73 defaults

# name -> (N, continuous, rewards, red, blue, seed, episodes, extra_done_steps)
SCENARIOS = {
    "1v1_disc_random":        (1, False, DEFAULT, "random", "random", 1, 8, 2),
    "1v1_cont_random":        (1, True, TRAINED, "random", "random", 2, 8, 2),
    "1v1_cont_instinct":      (1, True, DEFAULT, "instinct", "instinct", 3, 6, 0),
    "2v2_disc_instinct":      (2, False, TRAINED, "instinct", "instinct", 4, 6, 1),
    "2v2_disc_logits":        (2, False, TRAINED, "random_logits", "random_logits", 5, 4, 0),
    "2v2_disc_model_instinct": (2, False, TRAINED, "model", "instinct", 6, 8, 0),
    "2v2_cont_rand_instinct": (2, True, TRAINED, "random", "instinct", 7, 6, 1),
    "3v3_disc_rand_instinct": (3, False, DEFAULT, "random", "instinct", 8, 4, 0),
    "3v3_cont_rand_instinct": (3, True, TRAINED, "random", "instinct", 9, 6, 2),
    "3v3_cont_instinct":      (3, True, TRAINED, "instinct", "instinct", 10, 6, 0),
    "3v3_disc_shoot_instinct": (3, False, TRAINED, "noop_shoot", "instinct", 11, 3, 0),
    "5v5_cont_instinct_ff":   (5, True, TRAINED, "instinct_ff", "instinct_ff", 12, 3, 0),
    "5v5_disc_instinct":      (5, False, DEFAULT, "instinct", "instinct", 13, 3, 1),
    "4v4_cont_instinct":      (4, True, TRAINED, "instinct", "instinct", 14, 3, 0),
    "8v8_cont_rand_instinct": (8, True, TRAINED, "random", "instinct", 15, 2, 0),
}


def main():
    os.makedirs(OUT, exist_ok=True)
    total = 0
    for name, (N, cont, rw, red, blue, seed, eps, extra) in SCENARIOS.items():
        tr = rh.run_reference(N, cont, rw, red, blue, seed=seed, n_episodes=eps, extra_done_steps=extra)
        path = os.path.join(OUT, f"trace_{name}.npz")
        np.savez_compressed(path, **tr)
        sz = os.path.getsize(path); total += sz
        print(f"{name:28s} steps={len(tr['act']):5d} winners={tr['counters'].tolist()} {sz/1024:.0f} KiB")
    # shipped actors as arrays
    import torch
    w = {}
    for a in ("plane0", "plane1"):
        sd = torch.load(os.path.join(rh.REF_ROOT, "models", "completed_model", f"actor_{a}"), map_location="cpu")
        for k, v in sd.items():
            w[f"{a}.{k}"] = v.numpy()
    path = os.path.join(OUT, "completed_model_actors.npz")
    np.savez_compressed(path, **w)
    total += os.path.getsize(path)
    print(f"total {total/1024:.0f} KiB")


if __name__ == "__main__":
    main()
