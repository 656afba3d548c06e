#!/usr/bin/env python
"""bench.py -- agent-steps/s of the Battlespace hot path on B200 (BASELINE.json metric).

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): 3v3, continuous
actions, cf.json rewards, red = uniform[-1,1]^3 actions, blue = instinct team, 65,536 envs per GPU,
one "step" = one fused launch of T=32 ticks that also writes the MADDPG replay rows into a 2^20-row
device ring (bs_rollout_fused).  Weak scaling: every GPU owns 65,536 envs (global env index keys the
Philox streams), no collective on the step path, one NCCL all-reduce of the 8 statistics counters
per report interval (= once per timed region).

  value : kernel-path throughput, state/actions resident in HBM (red actions drawn in-kernel)
  e2e   : same metric through the host-buffer C-ABI call bs_rollout_fused_host: every step copies
          the red team's actions [T,E,3,3] float32 from pinned host memory and reads the 8 counters back
  sustained : the same launches repeated for >= 2 s with their own clock record (the timed region of
          `value` is short at the driver's --steps; `value` itself is not redefined)
  step_host_e2e : the single-tick plugin call that stands where parallel_env.step() stands
          (battle_env.py:281-381), bs_step_host with HOST buffers: actions in, obs/rew/done/env_done/winner out
  roofline / cpu_baseline : see DESIGN.md section "Measurement"

The ring holds 2^20 rows and one launch writes T*E = 2^21: the second half of a launch overwrites the
first (same env, same row, program order -- include/battlespace.h, bs_replay), i.e. the ring keeps the
last 16 ticks of each launch.  Both arms do this.

`--impl reference` times the CPU implementation of the same path on the host cores.  The reference
itself is pure Python and its tree is not present on the GPU box, so that arm runs the C port in
oracle/ (bit-exact against the reference, see tests/test_oracle_golden.py) on all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "deep-rl-battlespace_b200"))

N_AGENTS, T_TICKS, ENVS_PER_GPU, RING_ROWS = 3, 32, 65536, 1 << 20
REWARDS = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03,
               lose_punishment=-0.05)  # main.py:33-37 / models/completed_model/cf.json
METRIC, UNIT = "agent-steps/sec (3v3 vs instinct)", "agent-steps/s"


def workload_config(world):
    """The `config` object of BOTH arms (the reference arm runs a bounded env sample of it, described in its
    cpu_baseline.sample)."""
    return {"workload": "BASELINE configs[2]: 3v3 continuous, red=uniform[-1,1]^3 (Philox), blue=instinct, fused T=32 "
                        "ticks per step with auto-reset, replay rows (state,new_state,action,reward,done) written "
                        "into a ring",
            "envs_per_gpu": ENVS_PER_GPU, "ticks_per_step": T_TICKS, "replay_ring_rows": RING_ROWS, "rewards": "cf.json",
            "parallelism": f"env-sharded x{world}, no step-path collective, 1 stats all-reduce/interval",
            "l2": "inputs+outputs per launch (0.75 GB: 0.66 GB replay rows written, 44 MB state) exceed the "
                  "126 MB L2; no explicit flush"}


def python_reference_rate():
    """Pure-Python reference on this workload, measured in the BUILD container by oracle/time_reference.py
    (its tree is not on the GPU box); context for the CPU arm, which is the C port."""
    p = os.path.join(ROOT, "profiles", "python_reference_rate.json")
    if not os.path.exists(p):
        return None
    d = json.load(open(p))
    return {"value": d["agent_steps_per_s"], "unit": UNIT, "cores": d["cores"], "where": d["where"], "note": d["note"]}


def algorithmic_bytes_per_env_tick(N, A, T, actions_from_hbm):
    """SURVEY.md 8(d), Mode B: replay row + amortised state read/write (+ action read)."""
    L = 3 * N + 2
    row = 2 * N * L * 4 + N * A * 4 + N * 4 + N          # state, new_state, action, reward, done
    S = 218 * N + 16                                      # per-env state bytes
    return row + 2 * S / T + (N * A * 4 if actions_from_hbm else 0)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def gpu_numa_cpus(device_index):
    """CPUs of the NUMA node the GPU's PCIe root hangs off (None if the box does not say).  The e2e leg
    streams 75 MB of pinned host memory per step: from the far socket that copy runs at ~60 % speed."""
    try:
        bus = torch.cuda.get_device_properties(device_index).pci_bus_id  # not in every torch build
    except Exception:
        bus = None
    try:
        if bus is None:
            import subprocess
            out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(device_index)],
                                 capture_output=True, text=True, timeout=20).stdout.strip()
            bus = out.splitlines()[0].strip()
        dom, b, rest = bus.lower().split(":")
        path = f"/sys/bus/pci/devices/{dom[-4:]}:{b}:{rest}"
        node = int(open(path + "/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        return cpus or None
    except Exception:
        return None


def rank_cpu_share(cpus, local_rank, world):
    """This rank's slice of the host cores (the ranks of one box share them): every rank pins its copy thread
    and its pinned buffers to its own cores instead of all of them crowding the first ones."""
    cpus = sorted(cpus)
    if world <= 1 or len(cpus) < world:
        return set(cpus)
    per = len(cpus) // world
    return set(cpus[local_rank * per:(local_rank + 1) * per])


def cpu_port_rate(steps, warmup, envs, threads=None):
    """agent-steps/s of the CPU port (oracle/) on the same workload, all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    cores = threads or len(os.sched_getaffinity(0)) or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)   # torchrun exports OMP_NUM_THREADS=1; this arm uses every core
    from port import OracleEnv, POL_INSTINCT, POL_RANDOM
    try:                                          # libgomp may already be loaded (torch): set it at run time too
        import ctypes
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(cores)
    except OSError:
        pass
    env = OracleEnv(N_AGENTS, True, tuple(REWARDS.values()), envs, seed=0)
    # replay rows into a 2^20-row ring exactly like the GPU arm (cursor advances by T*envs per step)
    ring, cursor = None, 0
    for _ in range(warmup):
        ring = env.rollout(T_TICKS, [POL_RANDOM, POL_INSTINCT], 0, None, RING_ROWS, cursor, ring)
        cursor += T_TICKS * envs
    t0 = time.perf_counter()
    for _ in range(steps):
        ring = env.rollout(T_TICKS, [POL_RANDOM, POL_INSTINCT], 0, None, RING_ROWS, cursor, ring)
        cursor += T_TICKS * envs
    dt = time.perf_counter() - t0
    return steps * T_TICKS * envs * 2 * N_AGENTS / dt, dt, cores


def calibrated_cpu_rate(steps, warmup, target_s):
    """Size the env sample so that `steps` steps take about target_s seconds on this box's cores
    (capped at the full 65,536-env workload), then time it."""
    r0, _, _ = cpu_port_rate(1, 1, 2048)
    envs = int(target_s * r0 / (max(steps, 1) * T_TICKS * 2 * N_AGENTS))
    envs = 1 << max(11, min(16, envs.bit_length() - 1))   # a power of two in [2048, 65536]: divides the ring
    rate, dt, cores = cpu_port_rate(steps, warmup, envs)
    return rate, dt, cores, envs


def run_reference(args, rank):
    if rank != 0:
        return
    if args.ref_envs:
        envs = args.ref_envs
        rate, dt, cores = cpu_port_rate(args.steps, args.warmup, envs)
    else:
        rate, dt, cores, envs = calibrated_cpu_rate(args.steps, args.warmup, 20.0)
    sample = (f"{envs} of {ENVS_PER_GPU} envs x {T_TICKS} ticks per step, {args.steps} steps ({dt:.1f} s), replay rows "
              f"written into the {RING_ROWS}-row ring, OpenMP over envs on {cores} threads; C port of the reference "
              f"tick (the reference is pure Python and its tree is absent on this box)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(int(os.environ.get("WORLD_SIZE", 1))),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "python_reference": python_reference_rate()},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def step_host_e2e(env, N, E, world, dist, dev):
    """agent-steps/s through bs_step_host, the call that stands where parallel_env.step() stands: float32 actions
    [E,2N,3] from pinned host memory in, obs/rew/done/env_done/winner back into pinned host memory, every step;
    envs that finished are reset through bs_reset_host (mask in, first observation out) like `if env.env_done:
    env.reset()` in the reference's game loop (main.py:148,178)."""
    import ctypes as C
    import torch
    from battlespace._lib import ACT_F32, check, lib
    P, L = 2 * N, 3 * N + 2
    pin = lambda *shape, dtype: torch.empty(shape, dtype=dtype).pin_memory()
    act = pin(E, P, 3, dtype=torch.float32).uniform_(-1, 1)
    obs, rew = pin(E, P, L, dtype=torch.float32), pin(E, P, dtype=torch.float32)
    done, edone, win = pin(E, P, dtype=torch.uint8), pin(E, dtype=torch.uint8), pin(E, dtype=torch.int8)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    torch.cuda.synchronize()
    check(lib.bs_reset_host(env._h, None, None, ptr(obs)))
    steps, resets = 60, 0

    def one():
        nonlocal resets
        check(lib.bs_step_host(env._h, ptr(act), ACT_F32, None, ptr(obs), ptr(rew), ptr(done), ptr(edone), ptr(win)))
        if bool(edone.any()):
            resets += 1
            check(lib.bs_reset_host(env._h, ptr(edone), None, ptr(obs)))
    for _ in range(5):
        one()
    resets = 0
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    h2d = act.numel() * 4
    d2h = obs.numel() * 4 + rew.numel() * 4 + done.numel() + edone.numel() + win.numel()
    return {"value": steps * E * P * world / float(tt.item()), "unit": UNIT, "steps": steps, "ms_per_step": dt / steps * 1e3,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "reset_calls": resets,
            "call": "bs_step_host (+ bs_reset_host with the env_done mask when a game ended: E mask bytes in, obs out again)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU, help="envs per GPU")
    ap.add_argument("--ref-envs", type=int, default=0, help="envs per step of the CPU arm (0: sized for ~20 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sustained", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    from battlespace import DeviceReplay, allreduce_stats, parallel_env
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    N, T, E, A = N_AGENTS, T_TICKS, args.envs, 3
    env = parallel_env(n_agents=N, continuous_actions=True, n_envs=E, env_offset=rank * E, seed=0, device=dev, **REWARDS)
    replay = DeviceReplay(RING_ROWS, 512, N, 3 * N + 2, A, dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(steps, with_allreduce):
        """`steps` launches between a barrier + synchronize on both sides, device time (CUDA events on the
        launching stream), max over ranks, nvidia-smi clocks sampled meanwhile."""
        clocks = ClockSampler(local) if rank == 0 else None
        if clocks:
            clocks.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for _ in range(steps):
            env.rollout(T, "random", "instinct", 0, None, replay)     # one kernel launch per step
        stats = allreduce_stats(env.stats_tensor()) if with_allreduce else None   # the report-interval all-reduce (NCCL)
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        clk = clocks.stop() if clocks else None
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), clk, stats

    # ---- kernel path: everything resident in HBM ------------------------------------------------
    for _ in range(args.warmup):
        env.rollout(T, "random", "instinct", 0, None, replay)
    barrier()
    ms, clk, stats_dev = timed(args.steps, True)
    agent_steps = args.steps * T * E * 2 * N * world
    value = agent_steps / (ms * 1e-3)
    totals = stats_dev.cpu().numpy()

    # ---- sustained: the same launches for >= 2 s, own clock record (value keeps its definition) --
    sustained = None
    if not args.no_sustained:
        k_sus = max(args.steps, int(2500.0 / max(ms / args.steps, 1e-3)))
        ms_s, clk_s, _ = timed(k_sus, False)
        sustained = {"value": k_sus * T * E * 2 * N * world / (ms_s * 1e-3), "unit": UNIT, "steps": k_sus,
                     "seconds": ms_s * 1e-3, "ms_per_step": ms_s / k_sus, "clocks": clk_s}

    # ---- e2e: host-buffer C-ABI call, red actions copied from pinned host memory each step ------
    e2e, step_e2e = None, None
    if not args.no_e2e:
        k_e2e = max(3, min(args.steps, 40))
        # the caller's thread and its pinned buffer sit on the GPU's own NUMA node for this leg only
        # (the CPU baseline below uses every core again); with several ranks each keeps to its share of cores
        affinity0, near = os.sched_getaffinity(0), gpu_numa_cpus(dev.index)
        mine = rank_cpu_share(near or affinity0, local, world)
        if mine:
            os.sched_setaffinity(0, mine)
        host_act = torch.empty((T, E, N, A), dtype=torch.float32).uniform_(-1, 1).pin_memory()
        for _ in range(3):
            env.rollout_host(T, "external", "instinct", 0, host_act, replay)
        barrier()
        # every step: its action block H2D from pinned memory (bs_rollout_fused_host_begin returns when the block
        # has been consumed), its kernels, its counters D2H (bs_rollout_fused_host_end); one call stays in flight
        # so that the copies of step k+1 overlap the kernels of step k
        t0 = time.perf_counter()
        env.rollout_host_begin(T, "external", "instinct", 0, host_act, replay)
        for _ in range(k_e2e - 1):
            env.rollout_host_begin(T, "external", "instinct", 0, host_act, replay)
            env.rollout_host_end()
        env.rollout_host_end()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e = {"value": k_e2e * T * E * 2 * N * world / float(tt.item()), "unit": UNIT,
               "h2d_bytes_per_step": T * E * N * A * 4, "d2h_bytes_per_step": 64, "steps": k_e2e,
               "call": "bs_rollout_fused_host_begin/_end, one call in flight (red actions [T,E,3,3] f32 from pinned host "
                       "memory every step, the 8 counters back every step)",
               "host_numa_bound": bool(near), "cpus": len(mine) if mine else None}
        del host_act
        step_e2e = step_host_e2e(env, N, E, world, dist, dev)
        os.sched_setaffinity(0, affinity0)

    if rank == 0:
        peak, peak_src = measured_peak()
        bytes_per_launch = algorithmic_bytes_per_env_tick(N, A, T, False) * T * E
        launch_s = ms * 1e-3 / args.steps
        achieved = bytes_per_launch / launch_s / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("rollout_kernel_dram_bytes_per_launch")
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            rate, dt, cores, cenvs = calibrated_cpu_rate(10, 1, 12.0)
            cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{cenvs} of {E} envs x {T} ticks x 10 steps ({dt:.1f} s) of the same workload, replay rows "
                             f"written into the {RING_ROWS}-row ring, OpenMP over envs on {cores} threads",
                   "python_reference": python_reference_rate()}
        cfg = workload_config(world)
        cfg["envs_per_gpu"] = E
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg,
            "e2e": e2e, "gpu_launches": args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": "rollout_kernel<3,true>",
                         "algorithmic_bytes_per_launch": bytes_per_launch, "launch_ms": launch_s * 1e3},
            "cpu_baseline": cpu, "clocks": clk, "sustained": sustained, "step_host_e2e": step_e2e,
            "stats": {"games": int(totals[3]), "red_wins": int(totals[0]), "blue_wins": int(totals[1]),
                      "ties": int(totals[2]), "agent_steps_counted": int(totals[7])},
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
