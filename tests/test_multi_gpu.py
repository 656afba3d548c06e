"""GPU: the multi-GPU shapes of BASELINE configs 4 and 5 (SURVEY 8 e).

* config 4 (5v5 instinct_ff on both sides) sharded with shard_range() over two "ranks" that are two handles on
  one GPU (env_offset = the shard's first global env index): every counter and every env's state must equal the
  unsharded run -- envs never interact and Philox is keyed by the global env index (main.py has no collective at
  all; the one all-reduce of the report path is allreduce_stats()).
* config 5's actor-weight broadcast over NCCL (not gloo): needs >= 2 visible GPUs, spawns one process per GPU
  (skipped on a 1-GPU box; `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu` runs it), then a
  short bs_rollout_actor on every rank with the broadcast weights + the NCCL all-reduce of the counters, checked
  against one handle that owns all the envs."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from port import TRAINED_REWARDS

pytestmark = pytest.mark.gpu


def test_config4_sharded_over_two_ranks_equals_unsharded(cuda_device):
    from battlespace.sharding import shard_range
    from cuda_adapter import CudaEnv
    N, E, world = 5, 6145, 2                                   # odd total: the shards differ in size and are ragged
    whole = CudaEnv(N, True, TRAINED_REWARDS, E, seed=77)
    parts = []
    for r in range(world):
        off, cnt = shard_range(E, r, world)
        parts.append((off, cnt, CudaEnv(N, True, TRAINED_REWARDS, cnt, env_offset=off, seed=77)))
    for _ in range(5):
        whole.env.rollout(32, "instinct_ff", "instinct_ff")
        for _, _, p in parts:
            p.env.rollout(32, "instinct_ff", "instinct_ff")
    w = whole.get_state()["raw"]
    for off, cnt, p in parts:
        s = p.get_state()["raw"]
        for key in s:
            assert np.array_equal(s[key], w[key][off:off + cnt]), key
    assert np.array_equal(sum(p.stats() for _, _, p in parts), whole.stats())
    assert whole.stats()[6] == 5 * 32 * E


NCCL_WORKER = r"""
import os, sys
import numpy as np, torch
import torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from battlespace import DeviceReplay, parallel_env
from battlespace.actors import STATE_KEYS, TeamActors
from battlespace.sharding import allreduce_stats, broadcast_actors, shard_range
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
N, L, H, A, E, T = 3, 11, 64, 4, 4096, 24
CF = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03, lose_punishment=-0.05)
shapes = [(H, L), (H,), (H,), (H,), (H, H), (H,), (H,), (H,), (A, H), (A,)]
def params(seed):
    rng = np.random.RandomState(seed)
    return [{k: (rng.randn(*s) * 0.3).astype(np.float32) for k, s in zip(STATE_KEYS, shapes)} for _ in range(N)]
mine = TeamActors(params(100 + rank), dev)                    # every rank starts with different weights
ptrs = [t.data_ptr() for row in mine.tensors for t in row]
broadcast_actors(mine, src=0)                                 # NCCL broadcast over NVLink
torch.cuda.synchronize()
for row, w in zip(mine.tensors, params(100)):
    for t, k in zip(row, STATE_KEYS):
        assert np.array_equal(t.cpu().numpy(), w[k]), k
assert ptrs == [t.data_ptr() for row in mine.tensors for t in row]
off, cnt = shard_range(E, rank, world)
env = parallel_env(n_agents=N, n_envs=cnt, env_offset=off, seed=3, device=dev, **CF)
env.reset_batched()
env.rollout_actor(T, mine, 0, "instinct", None, DeviceReplay(T * cnt, 64, N, L, A, dev))
tot = allreduce_stats(env.stats_tensor()).cpu().numpy()       # the report-interval all-reduce
if rank == 0:
    whole = parallel_env(n_agents=N, n_envs=E, seed=3, device=dev, **CF)
    whole.reset_batched()
    whole.rollout_actor(T, TeamActors(params(100), dev), 0, "instinct", None, DeviceReplay(T * E, 64, N, L, A, dev))
    assert (tot == whole.stats()).all(), (tot, whole.stats())
print("OK", rank)
dist.destroy_process_group()
"""


def test_broadcast_actors_and_stats_over_nccl(tmp_path, cuda_device):
    import torch
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    script = tmp_path / "nccl_worker.py"
    script.write_text(NCCL_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541", WORLD_SIZE=str(world))
    procs = [subprocess.Popen([sys.executable, str(script), os.path.join(ROOT, "deep-rl-battlespace_b200")],
                              env=dict(env, RANK=str(r)), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("OK" in o for o in outs)
