"""ctypes binding of libbattlespace.so (C ABI: include/battlespace.h).

The library is CUDA-only.  If it has not been built this module raises at import time;
there is no CPU or PyTorch fallback on the product path.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BATTLESPACE_LIB") or os.path.join(_HERE, "libbattlespace.so")  # override: kernel A/B runs

BS_OK, BS_E_INVALID, BS_E_CUDA, BS_E_UNSUPPORTED = 0, -1, -2, -3
ACT_I32, ACT_F32, ACT_F64, ACT_NONE = 0, 1, 2, 3
POL_EXTERNAL, POL_RANDOM, POL_INSTINCT, POL_INSTINCT_FF = 0, 1, 2, 3
WINNER_NAMES = ("none", "red", "blue", "tie")
BULLET_SLOTS = 12
STATIC_SCRATCH_BYTES = 8 << 20   # BS_STATIC_SCRATCH_BYTES
POLICIES = {"external": POL_EXTERNAL, "random": POL_RANDOM, "instinct": POL_INSTINCT,
            "instinct_ff": POL_INSTINCT_FF}

# every symbol include/battlespace.h declares (tests check the library exports exactly these)
SYMBOLS = (
    "bs_version", "bs_last_error", "bs_create", "bs_destroy", "bs_reset", "bs_step", "bs_step_team",
    "bs_step_team_replay", "bs_instinct_actions", "bs_rollout_fused", "bs_stats", "bs_stats_device", "bs_get_state",
    "bs_set_state", "bs_reset_host", "bs_step_host", "bs_step_team_host", "bs_instinct_actions_host",
    "bs_rollout_fused_host", "bs_rollout_fused_host_begin", "bs_rollout_fused_host_end", "bs_actor_forward", "bs_rollout_actor", "bs_rollout_actor_static", "bs_replay_sample", "bs_render", "bs_episode_log", "bs_episode_log_count",
)


class BattlespaceError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libbattlespace error {code}: {msg}")
        self.code = code


class bs_config(C.Structure):
    _fields_ = [("n_agents", C.c_int32), ("continuous_actions", C.c_int32),
                ("hit_base_reward", C.c_double), ("hit_plane_reward", C.c_double),
                ("miss_punishment", C.c_double), ("die_punishment", C.c_double),
                ("lose_punishment", C.c_double)]


class bs_replay(C.Structure):
    _fields_ = [("state", C.c_void_p), ("new_state", C.c_void_p), ("action", C.c_void_p),
                ("reward", C.c_void_p), ("done", C.c_void_p), ("capacity", C.c_int64)]


ACTOR_FIELDS = ("fc1_w", "fc1_b", "ln1_g", "ln1_b", "fc2_w", "fc2_b", "ln2_g", "ln2_b", "pi_w", "pi_b")


class bs_actor(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ACTOR_FIELDS]


class bs_ou_noise(C.Structure):
    _fields_ = [("state", C.c_void_p), ("normals", C.c_void_p), ("reset_mask", C.c_void_p),
                ("mu", C.c_float), ("theta", C.c_float), ("sigma", C.c_float), ("scale", C.c_float),
                ("seed", C.c_uint64), ("step", C.c_uint32)]


class bs_sprites(C.Structure):
    _fields_ = [("plane", C.c_void_p * 2), ("base", C.c_void_p * 2), ("explosion_img", C.c_void_p),
                ("explosions", C.c_void_p), ("n_explosions", C.c_int32)]


STATE_FIELDS = ("plane_x", "plane_y", "plane_dir", "plane_hp", "base_pos", "base_hp", "bullet_pos",
                "bullet_vel", "tick", "env_done", "winner", "episode", "team_return")


class bs_state_view(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in STATE_FIELDS]


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(or `make -C deep-rl-battlespace_b200/csrc`).  There is no CPU fallback.")

lib = C.CDLL(LIB_PATH)
_vp, _i, _i64, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
lib.bs_version.restype = _i
lib.bs_last_error.restype = C.c_char_p
lib.bs_create.argtypes = [C.POINTER(bs_config), _i64, _i64, _i, _u64, C.POINTER(_vp)]
lib.bs_destroy.argtypes = [_vp]
lib.bs_reset.argtypes = [_vp, _vp, _vp, _vp, _vp]
lib.bs_step.argtypes = [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_step_team.argtypes = [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_step_team_replay.argtypes = [_vp, _i, _vp, _i, _i, _vp, C.POINTER(bs_replay), _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_instinct_actions.argtypes = [_vp, _i, _vp, _vp, _vp, _vp]
lib.bs_rollout_fused.argtypes = [_vp, _i, _i, _i, _i, _vp, C.POINTER(bs_replay), _i64, _vp]
lib.bs_stats.argtypes = [_vp, _vp, _vp]
lib.bs_stats_device.argtypes = [_vp, C.POINTER(_vp)]
lib.bs_get_state.argtypes = [_vp, C.POINTER(bs_state_view)]
lib.bs_set_state.argtypes = [_vp, C.POINTER(bs_state_view)]
lib.bs_reset_host.argtypes = [_vp, _vp, _vp, _vp]
lib.bs_step_host.argtypes = [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_step_team_host.argtypes = [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_instinct_actions_host.argtypes = [_vp, _i, _vp, _vp, _vp]
lib.bs_rollout_fused_host.argtypes = [_vp, _i, _i, _i, _i, _vp, C.POINTER(bs_replay), _i64, _vp]
lib.bs_rollout_fused_host_begin.argtypes = [_vp, _i, _i, _i, _i, _vp, C.POINTER(bs_replay), _i64]
lib.bs_rollout_fused_host_end.argtypes = [_vp, _vp]
lib.bs_actor_forward.argtypes = [_vp, _i, C.POINTER(bs_actor), _i, _vp, C.POINTER(bs_ou_noise), _vp, _vp]
lib.bs_rollout_actor.argtypes = [_vp, _i, _i, C.POINTER(bs_actor), _i, _i, C.POINTER(bs_ou_noise), C.POINTER(bs_replay), _i64,
                                 _vp, _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_rollout_actor_static.argtypes = [_vp, _i, _i, C.POINTER(bs_actor), _i, _i, C.POINTER(bs_ou_noise), C.POINTER(bs_replay), _i64,
                                        _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_replay_sample.argtypes = [C.POINTER(bs_replay), _i, _i, _i, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
lib.bs_render.argtypes = [_vp, _i64, C.POINTER(bs_sprites), _vp, _vp]
lib.bs_episode_log.argtypes = [_vp, _vp, _i64]
lib.bs_episode_log_count.argtypes = [_vp, C.POINTER(_i64), _vp]
for _name in SYMBOLS:
    if _name not in ("bs_version", "bs_last_error"):
        getattr(lib, _name).restype = _i


def check(rc: int) -> None:
    if rc != BS_OK:
        raise BattlespaceError(rc, (lib.bs_last_error() or b"").decode())
