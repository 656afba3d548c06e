"""CPU: the C-ABI library loads without a GPU and exports exactly what include/battlespace.h
declares; host-side logic (sharding, stats reduction over gloo, tie table) works; the product has
no CPU fallback (it must fail loudly without CUDA)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "battlespace.h")


def _header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bs_[a-z_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from battlespace import _lib
    declared = _header_symbols()
    assert declared == sorted(_lib.SYMBOLS)
    for name in declared:
        assert hasattr(_lib.lib, name), name
    assert _lib.lib.bs_version() == 100
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r" T (bs_[a-z_]+)", out)))
    assert exported == declared


def test_header_cites_reference_for_each_entry_point():
    src = open(HEADER).read()
    for ref in ("battle_env.py:246-279", "battle_env.py:281-381", "instinct/agent.py:10-62", "maddpg/buffer.py:25-47"):
        assert ref in src


def test_no_cpu_fallback():
    import torch
    from battlespace import _lib, parallel_env
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        parallel_env(n_agents=1)
    cfg = _lib.bs_config(1, 0, 100.0, 10.0, -1.0, -5.0, -20.0)
    h = C.c_void_p()
    rc = _lib.lib.bs_create(C.byref(cfg), 1, 0, 0, 0, C.byref(h))
    assert rc == _lib.BS_E_CUDA and b"no CUDA device" in _lib.lib.bs_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "deep-rl-battlespace_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(d, f)).read()
                assert "oracle" not in txt.lower(), os.path.join(d, f)


def test_tie_table_and_shards():
    from battlespace import shard_range, tie_tick
    assert [tie_tick(n) for n in range(1, 9)] == [121, 141, 161, 181, 200, 220, 240, 260]
    for total, world in ((1 << 20, 8), (65536, 4), (10, 3), (7, 7)):
        spans = [shard_range(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(c for _, c in spans) == total
        assert all(spans[r][0] + spans[r][1] == spans[r + 1][0] for r in range(world - 1))
    with pytest.raises(ValueError):
        shard_range(2, 0, 4)


WORKER = r"""
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
from battlespace.sharding import allreduce_stats, shard_range, summarize
from port import OracleEnv, TRAINED_REWARDS, POL_RANDOM, POL_INSTINCT
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
E = 96
off, cnt = shard_range(E, rank, world)
env = OracleEnv(3, True, TRAINED_REWARDS, cnt, env_offset=off, seed=42)   # CPU stand-in for one GPU's shard
for _ in range(8):
    env.rollout(32, [POL_RANDOM, POL_INSTINCT])
tot = allreduce_stats(env.stats())
if rank == 0:
    whole = OracleEnv(3, True, TRAINED_REWARDS, E, seed=42)
    for _ in range(8):
        whole.rollout(32, [POL_RANDOM, POL_INSTINCT])
    assert (tot == whole.stats()).all(), (tot, whole.stats())
    s = summarize(tot)
    assert s["env_ticks"] == 8 * 32 * E and s["agent_steps"] == s["env_ticks"] * 6
    print("OK", tot.tolist())
dist.destroy_process_group()
"""


def test_sharded_stats_allreduce_gloo_world2(tmp_path):
    """N>1 path on CPU: two ranks each own a contiguous env shard (global-index Philox keys), no
    collective on the step path, one all-reduce of the 8 counters -> equals the unsharded run."""
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29531", WORLD_SIZE="2", OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, str(script), os.path.join(ROOT, "deep-rl-battlespace_b200"),
                               os.path.join(ROOT, "oracle")], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "OK" in outs[0]


BCAST_WORKER = r"""
import os, sys
import numpy as np, torch
import torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from battlespace.actors import STATE_KEYS, TeamActors
from battlespace.sharding import broadcast_actors
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank = dist.get_rank()
L, H, A, N = 11, 64, 3, 3
shapes = [(H, L), (H,), (H,), (H,), (H, H), (H,), (H,), (H,), (A, H), (A,)]
def params(seed):
    rng = np.random.RandomState(seed)
    return [{k: rng.randn(*s).astype(np.float32) for k, s in zip(STATE_KEYS, shapes)} for _ in range(N)]
mine = TeamActors(params(100 + rank), "cpu")                  # every rank starts with different weights
ptrs = [t.data_ptr() for row in mine.tensors for t in row]
broadcast_actors(mine, src=0)
want = params(100)
for row, w in zip(mine.tensors, want):
    for t, k in zip(row, STATE_KEYS):
        assert np.array_equal(t.numpy(), w[k]), k
assert ptrs == [t.data_ptr() for row in mine.tensors for t in row]   # in place: bs_actor pointers stay valid
print("OK")
dist.destroy_process_group()
"""


def test_broadcast_actors_gloo_world2(tmp_path):
    """Config 5's weight broadcast: rank 1 ends with rank 0's actor parameters, in place."""
    script = tmp_path / "bcast.py"
    script.write_text(BCAST_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", WORLD_SIZE="2", OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, str(script), os.path.join(ROOT, "deep-rl-battlespace_b200")],
                              env=dict(env, RANK=str(r)), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("OK" in o for o in outs)


def test_oracle_fused_rollout_equals_unfused_loop():
    """The oracle's fused rollout (what the CUDA rollout kernel is checked against) is the same
    game loop as reset + instinct + step (main.py:148-205), row for row."""
    from port import OracleEnv, TRAINED_REWARDS, POL_EXTERNAL, POL_INSTINCT
    N, E, T = 2, 64, 150
    L, A = 3 * N + 2, 3
    rng = np.random.RandomState(0)
    ext = rng.uniform(-1.3, 1.3, (T, E, N, A)).astype(np.float32)
    f = OracleEnv(N, True, TRAINED_REWARDS, E, seed=9)
    rp = f.rollout(T, [POL_EXTERNAL, POL_INSTINCT], 0, ext, capacity=T * E)
    u = OracleEnv(N, True, TRAINED_REWARDS, E, seed=9)
    obs = u.reset()
    for t in range(T):
        blue = u.instinct(1, obs)
        act = np.concatenate([np.clip(ext[t].astype(np.float64), -1, 1), blue], 1)
        obs2, rew, done, env_done, _ = u.step(act)
        rows = slice(t * E, (t + 1) * E)
        assert np.array_equal(rp["state"][rows].reshape(E, N, L), obs[:, :N])
        assert np.array_equal(rp["new_state"][rows].reshape(E, N, L), obs2[:, :N])
        assert np.array_equal(rp["reward"][rows], rew[:, :N].astype(np.float32))
        assert np.array_equal(rp["done"][rows].astype(bool), done[:, :N])
        assert np.array_equal(rp["action"][rows], np.clip(ext[t], -1, 1))
        obs = u.reset(env_done.astype(np.uint8)) if env_done.any() else obs2
    assert np.array_equal(f.stats(), u.stats())
