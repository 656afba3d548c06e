"""TEST INFRASTRUCTURE ONLY -- SURVEY 8(d) config 1 at its full size, as a digest.

Runs the UNMODIFIED reference (oracle/ref_harness.py) for 1v1, discrete (default rewards) and continuous
(cf.json rewards, actions uniform(-1.3, 1.3) to exercise the clip, battle_env.py:295-297), seeds 1..8,
300 full episodes each (~36 k ticks per run, 16 runs), and writes tests/golden/config1_digest.npz.

A full trace of that size would be ~40 MB.  What is committed instead is, per episode: its length, its
winner, the float64 reward sums, CRC-32s of the float32 observations / float32-rounded rewards / dones of all
its steps, and a CRC-32 of the final plane/base/bullet state.  The INPUTS are not stored at all: the test
(tests/test_config1_full.py) regenerates them the way the harness does -- actions from
numpy RandomState(seed * 31337 + 101), spawn and bullet-spread draws from Python's random.Random(seed * 7919 + 13)
consumed in the reference's own order (sprites.py:82-91,246-251,314) -- and injects them into the CUDA env.

Run in the build container (needs /root/reference):  python oracle/gen_config1_digest.py
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden", "config1_digest.npz")
KEYS = ("hit_base_reward", "hit_plane_reward", "miss_punishment", "die_punishment", "lose_punishment")
TRAINED = dict(zip(KEYS, (1.0, 0.9, -0.02, -0.03, -0.05)))
EPISODES, SEEDS = 300, range(1, 9)


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c


def digest(tr):
    """Per-episode digest of a run_reference trace; the test computes the same from the CUDA env's outputs."""
    ep = tr["ep_start"]
    n = len(ep) - 1
    d = dict(steps=np.zeros(n, np.int32), winner=np.zeros(n, np.int8), rew_sum=np.zeros((n, 2), np.float64),
             crc_obs=np.zeros(n, np.uint32), crc_rew=np.zeros(n, np.uint32), crc_done=np.zeros(n, np.uint32),
             crc_state=np.zeros(n, np.uint32))
    for e in range(n):
        a, b = int(ep[e]), int(ep[e + 1])
        d["steps"][e] = b - a
        d["winner"][e] = tr["winner"][b - 1]
        d["rew_sum"][e] = tr["rew"][a:b].sum(axis=0)
        d["crc_obs"][e] = crc(tr["obs0"][e].astype(np.float32), tr["obs"][a:b].astype(np.float32))
        d["crc_rew"][e] = crc(tr["rew"][a:b].astype(np.float32))
        d["crc_done"][e] = crc(tr["done"][a:b].astype(np.uint8), tr["env_done"][a:b].astype(np.uint8))
        alive = tr["php"][b - 1] > 0   # a dead plane's sprite is gone in the reference: its last position is not state
        d["crc_state"][e] = crc(np.where(alive, tr["px"][b - 1], 0).astype(np.int16), np.where(alive, tr["py"][b - 1], 0).astype(np.int16),
                                np.maximum(tr["php"][b - 1], 0).astype(np.int8), tr["bhp"][b - 1].astype(np.int16),
                                tr["bullets"][b - 1].astype(np.int16))
    return d


def main():
    out = {}
    for cont in (False, True):
        for seed in SEEDS:
            tr = rh.run_reference(1, cont, TRAINED if cont else {}, "random", "random", seed=seed, n_episodes=EPISODES)
            tag = f"{'cont' if cont else 'disc'}_{seed}"
            for k, v in digest(tr).items():
                out[f"{tag}_{k}"] = v
            out[f"{tag}_reset_draws"] = tr["reset_draws"].astype(np.int16)
            out[f"{tag}_counters"] = tr["counters"]
            out[f"{tag}_n_spread"] = np.int64(np.isfinite(tr["draws"][:, :, 4]).sum())
            print(tag, "ticks", len(tr["act"]), "counters", tr["counters"].tolist(), flush=True)
    np.savez_compressed(OUT, **out)
    print(OUT, os.path.getsize(OUT) // 1024, "KiB")


if __name__ == "__main__":
    main()
