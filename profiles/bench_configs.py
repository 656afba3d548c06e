#!/usr/bin/env python
"""Secondary measurements for the BASELINE.json configs that are not the bench.py headline line.
One JSON line per case (device time with CUDA events, >= 3 warm-up iterations; every line says whether its
working set exceeds the 126 MB L2).
Run on the GPU box:  python profiles/bench_configs.py > gpurun_out/configs.jsonl
Multi-GPU shapes (configs 4 and 5, the pinned-H2D ceiling) run under torchrun, rank 0 prints:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29555 \
      profiles/bench_configs.py cfg4mg cfg5mg h2d

  modeA   : 3v3 continuous, 65,536 envs, single-tick bs_step with both teams' actions from HBM (SURVEY 8d Mode A)
  cfg2    : 2v2 discrete, 4,096 envs, shipped completed_model actors (PyTorch) vs in-kernel instinct
  cfg2n   : the same with the actors in CUDA and the loop native (bs_rollout_actor)
  cfg4    : 5v5 continuous, instinct_ff vs instinct_ff (full bullet pool), 1,048,576 envs / world, fused T=32
  cfg5    : 3v3 discrete, 32,768 envs per GPU: PyTorch actor forward + bs_step_team_replay + reset of done envs
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-rl-battlespace_b200"))
from battlespace import DeviceReplay, parallel_env  # noqa: E402
from battlespace.sharding import summarize  # noqa: E402

CF = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03, lose_punishment=-0.05)
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
RANK, WORLD, LOCAL = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(LOCAL)
dev = torch.device("cuda", LOCAL)
L2_BYTES = 126 << 20


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def S(N):
    return 218 * N + 16


def mode_a_at(E, flush_l2):
    """Single-tick bs_step, both teams' actions from HBM (SURVEY 8d Mode A), device time of the step kernels only.
    Working set = state + pool + obs/rew/done outputs + the action buffers in rotation.  With `flush_l2` a
    buffer larger than L2 is rewritten between launches (outside the timed events), so every launch starts
    from DRAM like the ncu capture does."""
    N, A = 3, 3
    L = 3 * N + 2
    env = parallel_env(n_agents=N, continuous_actions=True, n_envs=E, seed=0, device=dev, **CF)
    n_act = 4
    acts = [torch.empty((E, 2 * N, A), device=dev).uniform_(-1, 1) for _ in range(n_act)]
    env.reset_batched()
    bytes_tick = 2 * N * A * 4 + 2 * N * L * 4 + 2 * N * 4 + 2 * N + 2 + 2 * S(N)
    working_set = E * (S(N) + 12 + 2 * N * L * 4 + 2 * N * 4 + 2 * N + 2) + n_act * E * 2 * N * A * 4
    trash = torch.empty(192 << 20, dtype=torch.uint8, device=dev) if flush_l2 else None
    for k in range(5):
        env.step_batched(acts[k % n_act])
    torch.cuda.synchronize()
    iters = 100                                  # few envs finish within ~100 ticks of a fresh game
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for k, (e0, e1) in enumerate(evs):
        if trash is not None:
            trash.fill_(k & 0xff)
        e0.record()
        env.step_batched(acts[k % n_act])
        e1.record()
    torch.cuda.synchronize()
    ms = float(np.median([e0.elapsed_time(e1) for e0, e1 in evs]))
    gbs = bytes_tick * E / (ms * 1e-3) / 1e9
    dram = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        dram = json.load(open(tp)).get(f"step_kernel_dram_bytes_per_launch_{E}")
    return {"case": f"modeA 3v3 continuous, {E} envs, bs_step with both teams' actions from HBM",
            "envs": E, "l2_flushed_between_launches": bool(flush_l2), "working_set_bytes": working_set,
            "working_set_exceeds_l2": working_set > L2_BYTES, "ms_step_kernel_median": ms,
            "agent_steps_per_s_kernel": E * 2 * N / (ms * 1e-3), "algorithmic_bytes_per_env_tick": bytes_tick,
            "algorithmic_gbs": gbs, "frac_algorithmic": gbs / PEAK,
            "dram_bytes_per_launch": dram, "frac_dram": (dram / (ms * 1e-3) / 1e9 / PEAK) if dram else None,
            "peak_gbs": PEAK}


def mode_a():
    """Three readings: 65,536 envs as it runs in a loop (the ~100 MB working set stays in L2:
    an L2 number, NOT an HBM fraction), the same with L2 flushed between launches, and 1,048,576 envs, whose
    ~1.1 GB working set cannot be cached -- that last line is the HBM roofline reading of Mode A."""
    return [mode_a_at(65536, False), mode_a_at(65536, True), mode_a_at(1 << 20, False)]


def mode_a_loops():
    """The single-tick game loops a caller writes around bs_step (65,536 envs): step + reset_done, and
    bs_step_team with in-kernel instinct and auto-reset."""
    N, E, A = 3, 65536, 3
    env = parallel_env(n_agents=N, continuous_actions=True, n_envs=E, seed=0, device=dev, **CF)
    acts = [torch.empty((E, 2 * N, A), device=dev).uniform_(-1, 1) for _ in range(8)]
    env.reset_batched()
    k = [0]

    def f():
        env.step_batched(acts[k[0] % 8]); env.reset_done(); k[0] += 1
    ms_loop = timed(f, 200)
    env.reset_batched()
    half = [a[:, :N].contiguous() for a in acts]

    def fa():
        env.step_team_batched(half[k[0] % 8], 0, "instinct", auto_reset=True); k[0] += 1
    ms_auto = timed(fa, 400)
    return {"case": "modeA loops, 3v3 continuous, 65536 envs (working set in L2)", "ms_step_plus_reset_done": ms_loop,
            "ms_step_team_instinct_auto_reset": ms_auto, "agent_steps_per_s_loop": E * 2 * N / (ms_loop * 1e-3),
            "agent_steps_per_s_step_team_auto_reset": E * 2 * N / (ms_auto * 1e-3)}


def cfg2():
    w = np.load(os.path.join(ROOT, "tests", "golden", "completed_model_actors.npz"))
    E, N = 4096, 2
    W = [{n: torch.as_tensor(w[f"plane{k}.{n}"], device=dev) for n in
          ("fc1.weight", "fc1.bias", "bn1.weight", "bn1.bias", "fc2.weight", "fc2.bias", "bn2.weight", "bn2.bias", "pi.weight", "pi.bias")}
         for k in range(N)]

    def actor(k, x):  # maddpg/networks.py:80-85
        g = W[k]
        x = torch.relu(torch.nn.functional.layer_norm(x @ g["fc1.weight"].T + g["fc1.bias"], (64,), g["bn1.weight"], g["bn1.bias"]))
        x = torch.relu(torch.nn.functional.layer_norm(x @ g["fc2.weight"].T + g["fc2.bias"], (64,), g["bn2.weight"], g["bn2.bias"]))
        return torch.tanh(x @ g["pi.weight"].T + g["pi.bias"])
    env = parallel_env(n_agents=N, n_envs=E, seed=0, device=dev, **CF)
    env.reset_batched()

    def f():
        obs = env.obs_buf
        logits = torch.stack([actor(k, obs[:, k]) for k in range(N)], 1).contiguous()
        env.step_team_batched(logits, 0, "instinct")
        env.reset_done()
    ms = timed(f, 600, warm=5)
    s = summarize(env.stats())
    return {"case": "cfg2 2v2 discrete, completed_model actors (PyTorch eager) vs in-kernel instinct, 4096 envs",
            "ms_per_tick": ms, "agent_steps_per_s": E * 2 * N / (ms * 1e-3), "games": s["games"],
            "red_winrate": s["red_winrate"], "red_avg_score": s["red_avg_score"], "blue_avg_score": s["blue_avg_score"],
            "reference_anchor": "README.md:30 ~80 % wins; scores.json tail red 12.9 / blue 5.5 (with OU noise 0.4)"}


def cfg2_native():
    """cfg2 with the shipped completed_model actors evaluated by the CUDA actor kernel and the whole loop
    native (bs_rollout_actor, no OU noise): the win-rate anchor through the product's own actor path."""
    from battlespace import TeamActors
    w = np.load(os.path.join(ROOT, "tests", "golden", "completed_model_actors.npz"))
    E, N, T = 4096, 2, 32
    actors = TeamActors.from_npz(w, [f"plane{k}" for k in range(N)], dev)
    env = parallel_env(n_agents=N, n_envs=E, seed=0, device=dev, **CF)
    env.reset_batched()
    ms = timed(lambda: env.rollout_actor(T, actors, 0, "instinct"), 20, warm=2) / T
    s = summarize(env.stats())
    return {"case": "cfg2 native: 2v2 discrete, completed_model actors in CUDA (bs_rollout_actor) vs in-kernel instinct, 4096 envs",
            "ms_per_tick": ms, "agent_steps_per_s": E * 2 * N / (ms * 1e-3), "games": s["games"],
            "red_winrate": s["red_winrate"], "red_avg_score": s["red_avg_score"], "blue_avg_score": s["blue_avg_score"]}


def cfg4():
    N, E, T = 5, 1 << 20, 32
    env = parallel_env(n_agents=N, continuous_actions=True, n_envs=E, seed=0, device=dev, **CF)
    ms = timed(lambda: env.rollout(T, "instinct_ff", "instinct_ff", 0, None, None), 10)
    st = env.get_state()
    live = (st["bullet_pos"][..., 0] >= 0).sum(axis=(1, 2))
    bytes_tick = 2 * S(N) / T
    return {"case": "cfg4 5v5 continuous, instinct_ff both teams (full bullet pool), 1,048,576 envs on ONE GPU, fused T=32, no replay",
            "working_set_exceeds_l2": True,
            "ms_per_launch": ms, "agent_steps_per_s": E * T * 2 * N / (ms * 1e-3), "max_live_bullets_per_env": int(live.max()),
            "mean_live_bullets_per_env": float(live.mean()), "pool_slots_per_env": 24 * N,
            "hbm_gbs_state_only": bytes_tick * E * T / (ms * 1e-3) / 1e9}


def cfg5():
    N, E, A = 3, 32768, 4
    L = 3 * N + 2
    torch.manual_seed(0)
    actors = [torch.nn.Sequential(torch.nn.Linear(L, 64), torch.nn.LayerNorm(64), torch.nn.ReLU(), torch.nn.Linear(64, 64),
                                  torch.nn.LayerNorm(64), torch.nn.ReLU(), torch.nn.Linear(64, A), torch.nn.Tanh()).to(dev)
              for _ in range(N)]  # maddpg/networks.py:54-85, random init
    env = parallel_env(n_agents=N, n_envs=E, seed=0, device=dev, **CF)
    rp = DeviceReplay(1 << 20, 512, N, L, A, dev)
    env.reset_batched()

    @torch.no_grad()
    def f():
        obs = env.obs_buf
        logits = torch.stack([actors[k](obs[:, k]) for k in range(N)], 1).contiguous()
        env.step_team_batched(logits, 0, "instinct", replay=rp, auto_reset=True)

    @torch.no_grad()
    def env_only():
        env.step_team_batched(logits0, 0, "instinct", replay=rp, auto_reset=True)
    ms = timed(f, 300, warm=5)
    logits0 = torch.zeros((E, N, A), device=dev)
    ms_env = timed(env_only, 300, warm=5)
    # the same tick captured once in a CUDA graph (the eager loop is launch-bound: ~30 small kernels per tick)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        f(); f()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    cur0 = rp.mem_cntr
    with torch.cuda.graph(graph):
        f()
    rp.mem_cntr = cur0
    ms_graph = timed(graph.replay, 300, warm=5)   # (replay rows land on the captured cursor: timing only)
    return {"case": "cfg5 3v3 discrete, 32768 envs/GPU: PyTorch actors (random init) + bs_step_team_replay with auto_reset",
            "ms_per_tick_eager": ms, "agent_steps_per_s_eager": E * 2 * N / (ms * 1e-3), "ms_env_part": ms_env,
            "env_share_eager": ms_env / ms, "ms_per_tick_cuda_graph": ms_graph,
            "agent_steps_per_s_cuda_graph": E * 2 * N / (ms_graph * 1e-3), "env_share_cuda_graph": ms_env / ms_graph}


def cfg5_native():
    """cfg5 with the learner team's actors in CUDA (bs_actor_forward) and the whole game
    loop native (bs_rollout_actor): 2 launches per tick, no PyTorch module on the path."""
    from battlespace import DeviceOUNoise, TeamActors
    N, E, A, T = 3, 32768, 4, 32
    L = 3 * N + 2
    g = torch.Generator().manual_seed(0)
    u = lambda *s, a=1.0: ((torch.rand(*s, generator=g) * 2 - 1) * a)
    params = [{"fc1.weight": u(64, L, a=.125), "fc1.bias": u(64, a=.125), "bn1.weight": torch.ones(64), "bn1.bias": torch.zeros(64),
               "fc2.weight": u(64, 64, a=.125), "fc2.bias": u(64, a=.125), "bn2.weight": torch.ones(64), "bn2.bias": torch.zeros(64),
               "pi.weight": u(A, 64, a=.003), "pi.bias": u(A, a=.003)} for _ in range(N)]   # networks.py:58-72 init
    actors = TeamActors(params, dev)
    out = {"case": "cfg5 native: 3v3 discrete, 32768 envs/GPU, bs_rollout_actor (CUDA actors + step + replay rows + auto-reset), T=32 ticks per call"}
    for rows in ("1",):
        env = parallel_env(n_agents=N, n_envs=E, seed=0, device=dev, **CF)
        rp = DeviceReplay(1 << 20, 512, N, L, A, dev)
        noise = DeviceOUNoise(E, N, A, dev, seed=1)
        env.reset_batched()
        ms = timed(lambda: env.rollout_actor(T, actors, 0, "instinct", noise=noise, replay=rp), 20, warm=3) / T
        ms_actor = timed(lambda: env.actor_actions(actors, team=0, out=env._act_buf), 200, warm=5)
        flops = 2.0 * E * N * (L * 64 + 64 * 64 + 64 * A)
        out["native"] = {"ms_per_tick": ms, "agent_steps_per_s": E * 2 * N / (ms * 1e-3), "ms_actor_kernel": ms_actor,
                             "actor_rows_per_s": E * N / (ms_actor * 1e-3), "actor_tflops_fp32": flops / (ms_actor * 1e-3) / 1e12,
                             "actor_frac_of_fp32_fma_peak": flops / (ms_actor * 1e-3) / (148 * 128 * 2 * 1.965e9)}
        out["games"] = int(env.stats()[3])
    return out


def _dist():
    import torch.distributed as dist
    if WORLD > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=RANK, world_size=WORLD, device_id=dev)
    return dist


def _max_over_ranks(x):
    dist = _dist()
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if WORLD > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _barrier():
    if WORLD > 1:
        _dist().barrier()
    torch.cuda.synchronize()


def timed_mg(fn, iters, warm=3):
    """barrier + synchronize on both sides, device time, MAX over ranks."""
    for _ in range(warm):
        fn()
    _barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    _barrier()
    return _max_over_ranks(e0.elapsed_time(e1)) / iters


def cfg4_mg():
    """BASELINE config 4 in its real shape: 5v5 continuous, instinct_ff on both sides (full 120-slot pools),
    1,048,576 envs IN TOTAL split by shard_range() over the WORLD GPUs (strong scaling: the job is fixed), global
    env index keys the Philox streams, fused T=32, with and without replay rows; one NCCL all-reduce of the
    counters at the end of the timed region."""
    from battlespace.sharding import allreduce_stats, shard_range
    N, TOTAL, T, A = 5, 1 << 20, 32, 3
    L = 3 * N + 2
    off, cnt = shard_range(TOTAL, RANK, WORLD)
    out = {"case": "cfg4 5v5 continuous instinct_ff x2, 1,048,576 envs total, fused T=32", "n_gpus": WORLD, "envs_per_gpu": cnt,
           "scaling": "strong", "nccl_ranks": WORLD}
    for rows in (False, True):
        env = parallel_env(n_agents=N, continuous_actions=True, n_envs=cnt, env_offset=off, seed=0, device=dev, **CF)
        ring_rows = cnt * 4                     # the ring keeps the last 4 ticks of each launch (capacity % envs == 0)
        rp = DeviceReplay(ring_rows, 512, N, L, A, dev) if rows else None
        ms = timed_mg(lambda: env.rollout(T, "instinct_ff", "instinct_ff", 0, None, rp), 10)
        tot = allreduce_stats(env.stats_tensor())
        torch.cuda.synchronize()
        row_bytes = 2 * N * L * 4 + N * A * 4 + N * 4 + N
        alg = (row_bytes if rows else 0) + 2 * S(N) / T
        key = "with_rows" if rows else "no_rows"
        out[key] = {"ms_per_launch": ms, "agent_steps_per_s_total": TOTAL * T * 2 * N / (ms * 1e-3),
                    "agent_steps_per_s_per_gpu": cnt * T * 2 * N / (ms * 1e-3),
                    "algorithmic_gbs_per_gpu": alg * cnt * T / (ms * 1e-3) / 1e9, "frac_of_measured_hbm_per_gpu": alg * cnt * T / (ms * 1e-3) / 1e9 / PEAK,
                    "ring_rows_per_gpu": ring_rows if rows else 0, "games_total": int(tot[3].item())}
        del env, rp
        torch.cuda.empty_cache()
    return out


def cfg5_mg():
    """BASELINE config 5 in its real shape: 3v3 discrete, 32,768 envs per GPU (262,144 on 8), per tick actor forward
    (CUDA) -> env step with in-kernel instinct + replay rows + auto-reset (bs_rollout_actor, T=32 ticks per call);
    every K=4 calls the learner rank's actor weights are broadcast over NCCL (broadcast_actors, maddpg/team.py:16-28
    after learn()) and the counters are all-reduced once per report interval."""
    from battlespace import DeviceOUNoise, TeamActors
    from battlespace.sharding import allreduce_stats, broadcast_actors
    N, E, A, T, K = 3, 32768, 4, 32, 4
    L = 3 * N + 2
    g = torch.Generator().manual_seed(RANK)
    u = lambda *s, a=1.0: ((torch.rand(*s, generator=g) * 2 - 1) * a)
    params = [{"fc1.weight": u(64, L, a=.125), "fc1.bias": u(64, a=.125), "bn1.weight": torch.ones(64), "bn1.bias": torch.zeros(64),
               "fc2.weight": u(64, 64, a=.125), "fc2.bias": u(64, a=.125), "bn2.weight": torch.ones(64), "bn2.bias": torch.zeros(64),
               "pi.weight": u(A, 64, a=.003), "pi.bias": u(A, a=.003)} for _ in range(N)]
    actors = TeamActors(params, dev)
    env = parallel_env(n_agents=N, n_envs=E, env_offset=RANK * E, seed=0, device=dev, **CF)
    rp = DeviceReplay(1 << 20, 512, N, L, A, dev)
    noise = DeviceOUNoise(E, N, A, dev, seed=1)
    env.reset_batched()
    calls = [0]

    def f():
        if calls[0] % K == 0:
            broadcast_actors(actors, src=0)
        env.rollout_actor(T, actors, 0, "instinct", noise=noise, replay=rp)
        calls[0] += 1
    ms_call = timed_mg(f, 24, warm=4)
    # the broadcast alone
    _barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        broadcast_actors(actors, src=0)
    e1.record()
    _barrier()
    bc_us = _max_over_ranks(e0.elapsed_time(e1)) / 20 * 1e3
    e0.record()
    tot = allreduce_stats(env.stats_tensor())
    e1.record()
    _barrier()
    ar_us = _max_over_ranks(e0.elapsed_time(e1)) * 1e3
    ms_tick = ms_call / T
    return {"case": "cfg5 3v3 discrete, 32768 envs/GPU, bs_rollout_actor T=32 + NCCL weight broadcast every 4 calls",
            "n_gpus": WORLD, "envs_per_gpu": E, "scaling": "weak", "nccl_ranks": WORLD, "ms_per_tick": ms_tick,
            "agent_steps_per_s_per_gpu": E * 2 * N / (ms_tick * 1e-3), "agent_steps_per_s_total": WORLD * E * 2 * N / (ms_tick * 1e-3),
            "broadcast_us": bc_us, "broadcast_bytes": int(sum(t.numel() for row in actors.tensors for t in row) * 4),
            "allreduce_stats_us": ar_us, "games_total": int(tot[3].item())}


def h2d_ceiling():
    """The box's concurrent pinned-host -> device ceiling: every rank copies 75.5 MB (the e2e leg's action block)
    with one plain cudaMemcpyAsync, all ranks at the same time, each pinned to its own share of the cores."""
    nbytes = 32 * 65536 * 3 * 3 * 4
    cpus = sorted(os.sched_getaffinity(0))
    per = max(len(cpus) // max(WORLD, 1), 1)
    os.sched_setaffinity(0, set(cpus[LOCAL * per:(LOCAL + 1) * per]) or set(cpus))
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    host.fill_(1)
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    ms = timed_mg(lambda: d.copy_(host, non_blocking=True), 20, warm=3)
    return {"case": "pinned H2D ceiling, 75.5 MB per rank, all ranks concurrently", "n_gpus": WORLD, "bytes": nbytes,
            "ms_per_copy_max_over_ranks": ms, "gbs_per_gpu": nbytes / (ms * 1e-3) / 1e9, "gbs_total": WORLD * nbytes / (ms * 1e-3) / 1e9,
            "cpus_per_rank": per}


if __name__ == "__main__":
    which = sys.argv[1:] or ["modeA", "modeAloops", "cfg2", "cfg2n", "cfg4", "cfg5", "cfg5n"]
    fns = {"modeA": mode_a, "modeAloops": mode_a_loops, "cfg2": cfg2, "cfg2n": cfg2_native, "cfg4": cfg4, "cfg5": cfg5,
           "cfg5n": cfg5_native, "cfg4mg": cfg4_mg, "cfg5mg": cfg5_mg, "h2d": h2d_ceiling}
    for name in which:
        res = fns[name]()
        if RANK == 0:
            for r in (res if isinstance(res, list) else [res]):
                print(json.dumps(r), flush=True)
    if WORLD > 1:
        _dist().destroy_process_group()
