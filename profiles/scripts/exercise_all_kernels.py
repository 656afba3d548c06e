"""Small end-to-end exercise of every kernel and entry point at ragged sizes (E = 203; N = 3, 2, 8).
Run plain, or under compute-sanitizer (profiles/scripts/job_sanitizer.sh) where the pool allows it."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "deep-rl-battlespace_b200"))
from battlespace import DeviceOUNoise, DeviceReplay, EpisodeLog, TeamActors, parallel_env
dev = torch.device("cuda:0")
for N, cont in ((3, True), (2, False), (8, True)):
    E, L, A = 203, 3 * N + 2, 3 if cont else 4
    env = parallel_env(n_agents=N, continuous_actions=cont, n_envs=E, seed=1)
    log = EpisodeLog(env, 1024)
    rp = DeviceReplay(8 * E, 64, N, L, A, dev)
    env.reset_batched()
    env.rollout(40, "random", "instinct_ff" if cont else "instinct", 0, None, rp)
    act = torch.zeros((E, 2 * N, A), device=dev).uniform_(-1, 1) if cont else torch.randint(0, 4, (E, 2 * N), device=dev, dtype=torch.int32)
    for _ in range(5):
        env.step_batched(act); env.reset_done()
    g = torch.Generator().manual_seed(0)
    u = lambda *s, a=1.0: ((torch.rand(*s, generator=g) * 2 - 1) * a)
    params = [{"fc1.weight": u(64, L, a=.2), "fc1.bias": u(64, a=.1), "bn1.weight": torch.ones(64), "bn1.bias": torch.zeros(64),
               "fc2.weight": u(64, 64, a=.2), "fc2.bias": u(64, a=.1), "bn2.weight": torch.ones(64), "bn2.bias": torch.zeros(64),
               "pi.weight": u(A, 64, a=.3), "pi.bias": u(A, a=.1)} for _ in range(N)]
    actors = TeamActors(params, dev)
    noise = DeviceOUNoise(E, N, A, dev, seed=2)
    env.rollout_actor(6, actors, 0, "instinct", noise=noise, replay=rp)
    rp.sample()
    env.render_frame(E - 1)
    env.instinct_actions(1)
    st = env.get_state(); env.set_state(st)
    if N <= 5:   # the team-per-thread rollout mapping and both bullet-layout converters
        os.environ["BS_ROLLOUT_V2"] = "1"
        env2 = parallel_env(n_agents=N, continuous_actions=cont, n_envs=E, seed=1)
        os.environ.pop("BS_ROLLOUT_V2")
        env2.rollout(24, "random", "instinct", 0, None, rp)
        env2.step_batched(act)
        env2.rollout(8, "instinct", "random", 1, None, None)
        env2.get_state()
    torch.cuda.synchronize()
    print(N, cont, env.stats().tolist(), len(log.new_games()))
print("exercise_all_kernels OK")
