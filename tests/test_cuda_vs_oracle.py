"""GPU: CUDA path vs the CPU oracle at scale, both drawing from the same Philox streams
(include/battlespace.h), so thousands of envs can be compared step by step without the reference.

Covers: single-tick API with external actions, the in-kernel instinct team (bs_step_team ==
bs_instinct_actions + bs_step), the fused multi-tick rollout with auto-reset and replay rows,
sharding invariance, state export/import, and size-independent properties at BASELINE sizes."""
import numpy as np
import pytest

from port import (OracleEnv, TRAINED_REWARDS, DEFAULT_REWARDS, POL_EXTERNAL, POL_INSTINCT, POL_INSTINCT_FF,
                  POL_RANDOM)

pytestmark = pytest.mark.gpu
POLNAME = {POL_EXTERNAL: "external", POL_RANDOM: "random", POL_INSTINCT: "instinct", POL_INSTINCT_FF: "instinct_ff"}


def _state_equal(cs, os_, where=""):
    alive = os_["php"] > 0
    assert np.array_equal(np.maximum(cs["php"], 0), os_["php"]), f"php {where}"
    assert np.array_equal(cs["px"][alive], os_["px"][alive]) and np.array_equal(cs["py"][alive], os_["py"][alive]), f"pos {where}"
    assert np.allclose(cs["pdir"][alive], os_["pdir"][alive], rtol=0, atol=1e-9), f"dir {where}"
    assert np.array_equal(cs["bhp"], os_["bhp"]), f"bhp {where}"
    assert np.array_equal(cs["bullets"], os_["bullets"]), f"bullets {where}"
    for k in ("tick", "env_done", "winner", "episode", "base_pos"):
        assert np.array_equal(cs[k], os_[k]), f"{k} {where}"


def _rand_actions(rng, E, P, cont):
    if cont:
        return rng.uniform(-1.3, 1.3, (E, P, 3)).astype(np.float32)
    return rng.randint(0, 4, (E, P)).astype(np.int32)


@pytest.mark.parametrize("N,cont,E,steps", [(1, False, 2048, 260), (2, True, 2048, 300), (3, True, 4096, 340),
                                            (3, False, 1024, 340), (5, True, 1024, 420), (8, False, 300, 300)])
def test_step_external_actions(N, cont, E, steps, cuda_device):
    from cuda_adapter import CudaEnv
    rw = TRAINED_REWARDS
    c = CudaEnv(N, cont, rw, E, seed=1234 + N)
    o = OracleEnv(N, cont, rw, E, seed=1234 + N)
    rng = np.random.RandomState(N * 7 + cont)
    assert np.allclose(c.reset(), o.reset(), rtol=1e-5, atol=1e-7)
    for t in range(steps):
        a = _rand_actions(rng, E, 2 * N, cont)
        co, cr, cd, ced, cw = c.step(a)
        oo, or_, od, oed, ow = o.step(a)
        assert np.array_equal(cd, od) and np.array_equal(ced, oed) and np.array_equal(cw, ow), f"dones step {t}"
        assert np.allclose(co, oo, rtol=1e-5, atol=1e-7), f"obs step {t}"
        assert np.allclose(cr, or_, rtol=1e-5, atol=1e-9), f"rew step {t}"
        if t % 20 == 0 or t == steps - 1:
            _state_equal(c.get_state(), o.get_state(), f"step {t}")
        if t % 50 == 49:  # vectorised "if env_done: reset" (main.py:148)
            m = oed.astype(np.uint8)
            assert np.allclose(c.reset(m), o.reset(m), rtol=1e-5, atol=1e-7)
    assert np.array_equal(c.stats(), o.stats())


@pytest.mark.parametrize("N,cont,other", [(2, False, POL_INSTINCT), (3, True, POL_INSTINCT), (5, True, POL_INSTINCT_FF),
                                          (1, True, POL_INSTINCT)])
def test_step_team_equals_policy_plus_step(N, cont, other, cuda_device):
    """bs_step_team (scripted team in-kernel) == bs_instinct_actions + bs_step == oracle."""
    from cuda_adapter import CudaEnv
    E, P = 1024, 2 * N
    fused = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=7)
    split = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=7)
    o = OracleEnv(N, cont, TRAINED_REWARDS, E, seed=7)
    rng = np.random.RandomState(5)
    fo = fused.reset(); so = split.reset(); oo = o.reset()
    for t in range(200):
        red = _rand_actions(rng, E, N, cont)
        f = fused.step_team(red, 0, POLNAME[other])
        if other in (POL_INSTINCT, POL_INSTINCT_FF):
            blue_c = split.instinct(1, so); blue_o = o.instinct(1, oo)
            if cont:
                if other == POL_INSTINCT_FF:
                    blue_c[..., 2] = 1.0; blue_o[..., 2] = 1.0
                assert np.array_equal(blue_c[..., 2] > 0, blue_o[..., 2] > 0)          # shoot decision: exact
                assert np.allclose(blue_c, blue_o, rtol=1e-9, atol=1e-12)
                full_c = np.concatenate([red.astype(np.float64), blue_c], 1)
                full_o = np.concatenate([red.astype(np.float64), blue_o], 1)
            else:
                assert np.array_equal(blue_c, blue_o)
                full_c = full_o = np.concatenate([red, blue_c], 1).astype(np.int32)
            s = split.step(np.ascontiguousarray(full_c)); r = o.step(np.ascontiguousarray(full_o))
            so, oo = s[0], r[0]
            for x, y, z in zip(f, s, r):
                assert np.allclose(x, y, rtol=1e-6, atol=1e-7) and np.allclose(x, z, rtol=1e-5, atol=1e-7), f"step {t}"
    _state_equal(fused.get_state(), split.get_state())
    _state_equal(fused.get_state(), o.get_state())


CASES = [
    (3, True, POL_RANDOM, POL_INSTINCT, 4096, 32, 8),      # config 3 shape
    (2, False, POL_INSTINCT, POL_INSTINCT, 2048, 32, 8),
    (5, True, POL_INSTINCT_FF, POL_INSTINCT_FF, 1024, 16, 16),  # config 4 stress: full bullet pool
    (1, False, POL_RANDOM, POL_RANDOM, 4096, 64, 4),
    (3, True, POL_EXTERNAL, POL_INSTINCT, 2048, 32, 6),
    (2, False, POL_EXTERNAL, POL_INSTINCT, 1024, 16, 10),
    (4, True, POL_INSTINCT, POL_RANDOM, 512, 32, 6),
]


@pytest.mark.parametrize("N,cont,pr,pb,E,T,launches", CASES)
def test_rollout_fused_vs_oracle(N, cont, pr, pb, E, T, launches, cuda_device):
    import torch
    from battlespace import DeviceReplay, parallel_env
    from cuda_adapter import KEYS, CudaEnv
    L, A = 3 * N + 2, 3 if cont else 4
    cap = 3 * T * E + 17                                   # ring wraps inside the test
    c = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=99)
    o = OracleEnv(N, cont, TRAINED_REWARDS, E, seed=99)
    rp = DeviceReplay(cap, 512, N, L, A, c.dev)
    orp = None
    rng = np.random.RandomState(3)
    cursor = 0
    for k in range(launches):
        ext = None
        if pr == POL_EXTERNAL:
            ext = (rng.uniform(-1.3, 1.3, (T, E, N, A)) if cont else rng.uniform(-1, 1, (T, E, N, A))).astype(np.float32)
        c.env.rollout(T, POLNAME[pr], POLNAME[pb], 0, None if ext is None else torch.as_tensor(ext, device=c.dev), rp)
        orp = o.rollout(T, [pr, pb], 0, ext, cap, cursor, orp)
        cursor += T * E
    assert rp.mem_cntr == cursor
    _state_equal(c.get_state(), o.get_state())
    assert np.array_equal(c.stats(), o.stats())
    got = dict(state=rp.state_mem, new_state=rp.new_state_mem, action=rp.action_mem, reward=rp.rew_mem, done=rp.done_mem)
    for k, v in got.items():
        g = v.cpu().numpy().reshape(orp[k].shape)
        if k == "done":
            assert np.array_equal(g, orp[k]), k
        else:
            assert np.allclose(g, orp[k], rtol=1e-5, atol=1e-7), k
            assert (g == orp[k]).mean() > 0.99999, (k, (g == orp[k]).mean())


def test_shard_invariance(cuda_device):
    """Env g follows the same trajectory whichever shard holds it."""
    from cuda_adapter import CudaEnv
    N, E = 3, 1536
    whole = CudaEnv(N, True, TRAINED_REWARDS, E, seed=5)
    parts = [CudaEnv(N, True, TRAINED_REWARDS, E // 3, env_offset=k * (E // 3), seed=5) for k in range(3)]
    for _ in range(6):
        whole.env.rollout(32, "random", "instinct")
        for p in parts:
            p.env.rollout(32, "random", "instinct")
    w = whole.get_state()["raw"]
    for k, p in enumerate(parts):
        s = p.get_state()["raw"]
        sl = slice(k * (E // 3), (k + 1) * (E // 3))
        for key in s:
            assert np.array_equal(s[key], w[key][sl]), key
    assert np.array_equal(sum(p.stats() for p in parts), whole.stats())


def test_state_roundtrip(cuda_device):
    """bs_get_state -> bs_set_state into a fresh handle continues identically (checkpoint/resume)."""
    from cuda_adapter import CudaEnv
    N, E = 2, 512
    a = CudaEnv(N, False, DEFAULT_REWARDS, E, seed=11)
    for _ in range(3):
        a.env.rollout(20, "instinct", "instinct")
    b = CudaEnv(N, False, DEFAULT_REWARDS, E, seed=11)
    b.env.set_state(a.get_state()["raw"])
    for e in (a, b):
        e.env.rollout(40, "instinct", "instinct")
    sa, sb = a.get_state()["raw"], b.get_state()["raw"]
    for key in sa:
        assert np.array_equal(sa[key], sb[key]), key


def test_empty_action_dict_ties(cuda_device):
    """battle_env.py:309-313: stepping with no actions ties the game."""
    from cuda_adapter import CudaEnv
    c = CudaEnv(2, False, DEFAULT_REWARDS, 64, seed=1); o = OracleEnv(2, False, DEFAULT_REWARDS, 64, seed=1)
    c.reset(); o.reset()
    rc, ro = c.step(None), o.step(None)
    for x, y in zip(rc, ro):
        assert np.allclose(x, y)
    assert (rc[4] == 3).all() and rc[3].all() and np.array_equal(c.stats(), o.stats())


def test_full_size_properties(cuda_device):
    """BASELINE config 3 size (3v3, 65,536 envs, T=32): invariants that need no oracle."""
    import torch
    from battlespace import DeviceReplay
    from cuda_adapter import CudaEnv
    N, E, T = 3, 65536, 32
    c = CudaEnv(N, True, TRAINED_REWARDS, E, seed=2024)
    rp = DeviceReplay(1 << 20, 512, N, 3 * N + 2, 3, c.dev)
    for _ in range(12):
        c.env.rollout(T, "random", "instinct", 0, None, rp)
    st = c.get_state()["raw"]; s = c.stats()
    alive = st["plane_hp"] > 0
    assert st["plane_x"][alive].min() >= 25 and st["plane_x"][alive].max() <= 1175      # sprites.py:133-141
    assert st["plane_y"][alive].min() >= 24 and st["plane_y"][alive].max() <= 776
    assert st["plane_hp"].max() <= 4 and st["base_hp"].max() <= 15
    assert (st["plane_dir"] >= 0).all() and (st["plane_dir"] <= 360).all()              # sprites.py:100-103
    live = st["bullet_vel"][..., 2] & 1 == 1
    assert (st["bullet_pos"][..., 0][live] >= 0).all() and (st["bullet_pos"][..., 0][live] <= 1200).all()
    assert (st["bullet_pos"][..., 1][live] >= 0).all() and (st["bullet_pos"][..., 1][live] <= 800).all()
    assert not live[:, :, 11].any()                                                     # 12th update removes
    assert (st["tick"] <= 161).all()                                                    # tie table, N=3
    assert s[6] == 12 * T * E and s[7] == s[6] * 2 * N
    assert s[3] == s[0] + s[1] + s[2]                                                   # games = wins + ties
    assert (st["episode"] >= 1).all() and s[3] >= int(st["episode"].astype(np.int64).sum()) - E
    # replay rows: obs in range, new_state(t) == state(t+1) for an env unless it reset in between
    S = rp.state_mem.view(-1, N, 3 * N + 2); S2 = rp.new_state_mem.view(-1, N, 3 * N + 2)
    assert float(S.abs().max()) <= 1.0 and float(S2.abs().max()) <= 1.0
    last = (12 * T - 2) * E % (1 << 20)
    done_rows = rp.done_mem[last:last + E].bool().all(dim=1)
    same = (S2[last:last + E] == S[(last + E) % (1 << 20):(last + E) % (1 << 20) + E]).all(dim=2).all(dim=1)
    assert bool((same | done_rows).all())
    # determinism: a second handle with the same seed reproduces the counters exactly
    c2 = CudaEnv(N, True, TRAINED_REWARDS, E, seed=2024)
    for _ in range(12):
        c2.env.rollout(T, "random", "instinct")
    assert np.array_equal(c2.stats(), s)


@pytest.mark.gpu
@pytest.mark.parametrize("N,cont,E", [(3, True, 20000), (2, False, 17000), (3, True, 1000)])
def test_rollout_host_pipelined_equals_device(N, cont, E, cuda_device):
    """bs_rollout_fused_host cuts the envs into chunks and overlaps the action copy with the rollout
    (one chunk when the handle is small); rows, state and counters must equal the one-launch path."""
    import torch
    from battlespace import DeviceReplay
    from cuda_adapter import CudaEnv
    L, A, T = 3 * N + 2, 3 if cont else 4, 8
    a = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=5)
    b = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=5)
    ra, rb = DeviceReplay(2 * T * E, 512, N, L, A, a.dev), DeviceReplay(2 * T * E, 512, N, L, A, b.dev)
    rng = np.random.RandomState(11)
    for k in range(3):
        ext = torch.from_numpy(rng.uniform(-1.2, 1.2, (T, E, N, A)).astype(np.float32))
        a.env.rollout(T, "external", "instinct", 0, ext.to(a.dev), ra)
        st = b.env.rollout_host(T, "external", "instinct", 0, ext.pin_memory(), rb)
        assert np.array_equal(st, a.stats())
    _state_equal(a.get_state(), b.get_state())
    for x, y in ((ra.state_mem, rb.state_mem), (ra.new_state_mem, rb.new_state_mem), (ra.action_mem, rb.action_mem),
                 (ra.rew_mem, rb.rew_mem), (ra.done_mem, rb.done_mem)):
        assert torch.equal(x, y)


@pytest.mark.gpu
def test_rollout_host_split_phase_two_calls_in_flight(cuda_device):
    """bs_rollout_fused_host_begin/_end with two calls in flight (the copies of call k+1 overlap the kernels of
    call k, two staging buffers rotate): the host buffer is refilled right after _begin returns; counters per call,
    rows and state must equal the synchronous device path."""
    import torch
    from battlespace import DeviceReplay, _lib
    from battlespace._lib import lib
    from cuda_adapter import CudaEnv
    N, cont, E, T, calls = 3, True, 24000, 8, 5
    L, A = 3 * N + 2, 3
    a = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=6)
    b = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=6)
    ra, rb = DeviceReplay(3 * T * E, 512, N, L, A, a.dev), DeviceReplay(3 * T * E, 512, N, L, A, b.dev)
    rng = np.random.RandomState(12)
    blocks = [rng.uniform(-1.2, 1.2, (T, E, N, A)).astype(np.float32) for _ in range(calls)]
    want = []
    for blk in blocks:
        a.env.rollout(T, "external", "instinct", 0, torch.from_numpy(blk).to(a.dev), ra)
        want.append(a.stats())
    host = torch.empty((T, E, N, A), dtype=torch.float32).pin_memory()
    got = []
    host.copy_(torch.from_numpy(blocks[0]))
    b.env.rollout_host_begin(T, "external", "instinct", 0, host, rb)
    for k in range(1, calls):
        host.copy_(torch.from_numpy(blocks[k]))              # overwritten while call k-1 may still be running
        b.env.rollout_host_begin(T, "external", "instinct", 0, host, rb)
        got.append(b.env.rollout_host_end())
    assert lib.bs_rollout_fused_host_begin(b.env._h, T, 0, _lib.POL_RANDOM, _lib.POL_INSTINCT, None, None, 0) == 0
    assert lib.bs_rollout_fused_host_begin(b.env._h, T, 0, _lib.POL_RANDOM, _lib.POL_INSTINCT, None, None, 0) == _lib.BS_E_INVALID
    got.append(b.env.rollout_host_end())
    assert lib.bs_rollout_fused_host_end(b.env._h, None) == 0 and lib.bs_rollout_fused_host_end(b.env._h, None) == _lib.BS_E_INVALID
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    a.env.rollout(T, "random", "instinct")                   # the extra in-kernel call above
    _state_equal(a.get_state(), b.get_state())
    for x, y in ((ra.state_mem, rb.state_mem), (ra.new_state_mem, rb.new_state_mem), (ra.action_mem, rb.action_mem),
                 (ra.rew_mem, rb.rew_mem), (ra.done_mem, rb.done_mem)):
        assert torch.equal(x, y)


@pytest.mark.gpu
@pytest.mark.parametrize("N,cont,E,ticks,auto", [(3, True, 3000, 208, False), (2, False, 2000, 176, False),
                                                 (3, True, 3000, 208, True), (1, False, 4000, 144, True)])
def test_step_team_replay_equals_fused_rollout(N, cont, E, ticks, auto, cuda_device):
    """The reference's game loop spelled out tick by tick (reset done envs, step with the learner's
    actions and the instinct team in-kernel, store_transition -- main.py:148,179-181,205) must give the
    rows, state and counters of the fused rollout fed the same actions."""
    import torch
    from battlespace import DeviceReplay
    from cuda_adapter import CudaEnv
    L, A, T = 3 * N + 2, 3 if cont else 4, 16
    a = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=21)
    b = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=21)
    cap = ticks * E
    ra, rb = DeviceReplay(cap, 512, N, L, A, a.dev), DeviceReplay(cap, 512, N, L, A, b.dev)
    gen = torch.Generator(device="cpu").manual_seed(4)
    for k in range(ticks // T):
        ext = (torch.rand((T, E, N, A), generator=gen) * 2.4 - 1.2).to(a.dev)
        for t in range(T):
            if auto:      # finished envs are reset inside the step launch
                if k == 0 and t == 0:
                    a.env.reset_batched()
                a.env.step_team_batched(ext[t].contiguous(), 0, "instinct", replay=ra, auto_reset=True)
            else:
                a.env.reset_done()
                a.env.step_team_batched(ext[t].contiguous(), 0, "instinct", replay=ra)
        b.env.rollout(T, "external", "instinct", 0, ext, rb)
    assert ra.mem_cntr == rb.mem_cntr == cap
    assert np.array_equal(a.stats(), b.stats()) and a.stats()[3] > E  # every env finished at least one game
    if not auto:  # (with auto_reset the envs that ended on the last tick are already in their next game)
        _state_equal(a.get_state(), b.get_state())
    for x, y in ((ra.state_mem, rb.state_mem), (ra.new_state_mem, rb.new_state_mem), (ra.action_mem, rb.action_mem),
                 (ra.rew_mem, rb.rew_mem), (ra.done_mem, rb.done_mem)):
        assert torch.equal(x, y)


def test_fp32_speculated_trig_never_changes_the_result(cuda_device, monkeypatch):
    """The fp32 speculation of the plane-move / bullet-step cosine (tick.cuh cos_sin_fast, guard 2.5e-4
    against an error bound of 3.2e-5) must give the state of the exact double path, bit for bit:
    65,536 envs x 320 ticks = 1.3e8 plane-ticks, both policies shooting as often as possible."""
    from battlespace import parallel_env
    kw = dict(n_agents=3, continuous_actions=True, n_envs=65536, seed=21, hit_base_reward=1.0, hit_plane_reward=0.9,
              miss_punishment=-0.02, die_punishment=-0.03, lose_punishment=-0.05)
    fast = parallel_env(**kw)
    monkeypatch.setenv("BS_EXACT_TRIG", "1")
    exact = parallel_env(**kw)
    monkeypatch.delenv("BS_EXACT_TRIG")
    for _ in range(10):
        fast.rollout(32, "random", "instinct_ff")
        exact.rollout(32, "random", "instinct_ff")
    assert np.array_equal(fast.stats(), exact.stats()) and fast.stats()[3] > 100000
    a, b = fast.get_state(), exact.get_state()
    for k in a:
        assert np.array_equal(a[k], b[k]), k
