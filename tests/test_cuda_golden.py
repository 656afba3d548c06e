"""GPU: the CUDA path, called through the C ABI, against the reference's own traces (tests/golden)
and against the CPU oracle on the same injected draws.

Bar (north_star): discrete events bit-exact -- hits, kills, base damage, dones, winner, bullet and
plane integer state, instinct target choice / shoot decision; float32 observations and rewards
within rtol 1e-5 (replay_trace.RTOL)."""
import os

import numpy as np
import pytest

from conftest import golden_traces, load_trace
from replay_trace import replay

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("path", golden_traces(), ids=lambda p: os.path.basename(p)[6:-4])
def test_cuda_matches_reference_trace(path, cuda_device):
    from cuda_adapter import CudaEnv
    tr = load_trace(path)
    env = CudaEnv(int(tr["n_agents"]), bool(tr["continuous"]), tr["rewards"], len(tr["ep_start"]) - 1)
    counts = replay(env, tr, exact=False)
    # fp64 libm differences may flip the last float32 bit of an observation once in ~1e8 values
    assert counts["obs_exact"] >= counts["obs_total"] - max(2, counts["obs_total"] // 100000), counts
    assert (env.stats()[:4] == tr["counters"]).all()


@pytest.mark.parametrize("name", ["1v1_cont_random", "2v2_disc_model_instinct", "3v3_cont_instinct"])
def test_cuda_host_entry_points_match_reference_trace(name, cuda_device):
    from cuda_adapter import CudaEnv
    tr = load_trace(os.path.join(os.path.dirname(golden_traces()[0]), f"trace_{name}.npz"))
    env = CudaEnv(int(tr["n_agents"]), bool(tr["continuous"]), tr["rewards"], len(tr["ep_start"]) - 1, mode="host")
    replay(env, tr, exact=False)
    assert (env.stats()[:4] == tr["counters"]).all()
