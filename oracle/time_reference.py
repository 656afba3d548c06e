"""TEST / BENCH INFRASTRUCTURE ONLY -- times the UNMODIFIED pure-Python reference on the bench workload.

Runs /root/reference (envs/battle_env.py, envs/sprites.py, instinct/*.py) under oracle/stubs exactly as
oracle/ref_harness.py does, without any tracing: BASELINE configs[2] as the game loop of main.py:148-213
would play it -- 3v3, continuous actions, cf.json rewards, red = uniform[-1,1]^3 actions, blue = the
reference's instinct Team -- one env, one Python thread.  The reference tree is not on the GPU box, so
the rate is measured HERE (build container) and committed as profiles/python_reference_rate.json, which
bench.py quotes as cpu_baseline.python_reference (context only; the timed CPU arm is the C port).

The stub's pygame.transform.rotate does no pixel work, so this is an UPPER bound on the reference's
real speed.
usage: python oracle/time_reference.py [games]
"""
import json
import os
import platform
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_harness as rh  # noqa: E402

REWARDS = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03,
               lose_punishment=-0.05)


def main():
    games = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    be, sprites, iagent, iteam = rh._load()
    N = 3
    rng = np.random.RandomState(0)
    with rh._cwd(rh.REF_ROOT):
        env = be.parallel_env(n_agents=N, continuous_actions=True, **REWARDS)
        blue = iteam.Team(env.possible_blue, env.possible_red, env)
        ticks, t0 = 0, time.perf_counter()
        for _ in range(games):
            obs = env.reset()
            while not env.env_done:
                actions = {a: rng.uniform(-1, 1, 3) for a in env.possible_red}
                # numpy-1.23.1 promotion (requirements.txt:3): the instinct policy computes in float64
                actions.update(blue.choose_actions({a: obs[a].astype(np.float64) for a in env.possible_blue}))
                obs, rew, dones, infos = env.step(actions)
                ticks += 1
        dt = time.perf_counter() - t0
    out = {"agent_steps_per_s": ticks * 2 * N / dt, "env_steps_per_s": ticks / dt, "games": games, "ticks": ticks,
           "seconds": dt, "cores": 1, "workload": "3v3 continuous, red uniform random, blue reference instinct Team, 1 env",
           "where": "build container (not the GPU box): " + platform.processor() + f" python {platform.python_version()} numpy {np.__version__}",
           "note": "unmodified reference source under oracle/stubs; the stubbed pygame.transform.rotate does no pixel work: upper bound"}
    print(json.dumps(out))
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "python_reference_rate.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
