"""TEST INFRASTRUCTURE ONLY -- live cross-check of the C port against the reference source.

Runs the unmodified reference (oracle/ref_harness.py) over many seeded games in every
mode and replays each trace into oracle/liboracle.so, demanding BIT-IDENTICAL
observations, rewards, dones, winners, plane/base/bullet state and instinct actions.
Needs /root/reference, so it runs in the build container only:

    python oracle/validate_port.py [--episodes 40] [--seeds 6]
"""
import argparse
import itertools
import sys
import time
import os

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_harness as rh  # noqa: E402
from port import OracleEnv, TRAINED_REWARDS, DEFAULT_REWARDS, tie_tick  # noqa: E402
from replay_trace import replay  # noqa: E402

RW = lambda r: dict(zip(("hit_base_reward", "hit_plane_reward", "miss_punishment", "die_punishment", "lose_punishment"), r))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--episodes", type=int, default=24)
    ap.add_argument("--seeds", type=int, default=3)
    ap.add_argument("--sizes", type=int, nargs="*", default=[1, 2, 3, 5])
    a = ap.parse_args()
    for n in range(1, 9):
        assert tie_tick(n) == rh.tie_tick(n)
    total, t0 = {}, time.time()
    modes = [("random", "random"), ("instinct", "instinct"), ("random", "instinct"), ("instinct_ff", "instinct_ff"),
             ("noop_shoot", "instinct")]
    for N, cont, (red, blue), seed in itertools.product(a.sizes, (False, True), modes, range(a.seeds)):
        rw = TRAINED_REWARDS if seed % 2 == 0 else DEFAULT_REWARDS
        tr = rh.run_reference(N, cont, RW(rw), red, blue, seed=seed, n_episodes=a.episodes, extra_done_steps=2)
        env = OracleEnv(N, cont, rw, n_envs=a.episodes)
        replay(env, tr, exact=True, counts=total)
        assert (env.stats()[:4] == tr["counters"]).all(), (env.stats(), tr["counters"])
    if not any(False for _ in ()):  # discrete argmax path
        for N, seed in itertools.product(a.sizes, range(a.seeds)):
            tr = rh.run_reference(N, False, RW(TRAINED_REWARDS), "random_logits", "random_logits", seed=seed,
                                  n_episodes=a.episodes)
            replay(OracleEnv(N, False, TRAINED_REWARDS, n_envs=a.episodes), tr, exact=True, counts=total)
    print(f"OK: {total} in {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main()
