"""Design study (not a test): how many bullet updates could a "sleep counter" skip?

Idea: after a full update of a bullet, compute a conservative number of ticks during which it can neither leave the field,
reach the enemy base, reach an enemy plane (planes move <= 28 px per tick and axis, bullets <= 45) nor expire, and
only advance it until then.  The rule is evaluated on the C oracle's states (3v3, uniform-random vs instinct, the
bench workload).  Result (512 envs x 400 ticks): 54 % of the bullet updates still need the full pass; the two-pass
form breaks even at 52 %, so the idea was dropped.  usage: python tests/tools/bullet_sleep_sim.py [N]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from port import OracleEnv, TRAINED_REWARDS
N = int(sys.argv[1]) if len(sys.argv) > 1 else 3
E, T = 512, 400
VB, VP = 45, 73
env = OracleEnv(N, True, TRAINED_REWARDS, E, seed=7)
P = 2 * N
def rule(st):
    b = st["bullets"].astype(np.int32)            # [E,P,12,2]
    bx, by = b[..., 0], b[..., 1]
    present = bx >= 0
    age = np.arange(1, 13)[None, None, :]
    k = 11 - age + np.zeros_like(bx)                # expiry: update number 12 must be heavy
    edge = np.minimum(np.minimum(bx, 1200 - bx), np.minimum(by, 800 - by))
    k = np.minimum(k, np.maximum(edge, 0) * 91 >> 12)
    team = (np.arange(P) >= N).astype(int)         # owner team
    ebase = st["base_pos"].astype(np.int32)[:, 1 - team, :]     # [E,P,2] enemy base centre
    dbx = np.abs(bx - ebase[:, :, None, 0]) - 34; dby = np.abs(by - ebase[:, :, None, 1]) - 33
    k = np.minimum(k, np.maximum(np.maximum(dbx, dby), 0) * 91 >> 12)
    px, py, php = st["px"].astype(np.int32), st["py"].astype(np.int32), st["php"]
    for j in range(N):
        ej = np.where(team == 0, N + j, j)         # enemy plane index per owner
        ex, ey, eh = px[:, ej], py[:, ej], php[:, ej]   # [E,P]
        d = np.maximum(np.abs(bx - ex[:, :, None]) - 28, np.abs(by - ey[:, :, None]) - 26)
        kk = np.maximum(d, 0) * 7 >> 9
        k = np.minimum(k, np.where((eh > 0)[:, :, None], kk, 99))
    return np.maximum(k, 0), present
sleep = np.zeros((E, P, 12), np.int32)
prev = None
heavy = light = viol = 0
hist = np.zeros(13, np.int64)
for t in range(T):
    env.rollout(1, [1, 2])
    st = env.get_state()
    k, present = rule(st)
    if prev is not None:
        same = (st["episode"] == prev["episode"])[:, None, None]
        pb = prev["bullets"][..., 0] >= 0
        n_upd = pb.sum()
        was_light = pb & (sleep > 0)
        light += was_light.sum(); heavy += (pb & (sleep == 0)).sum()
        # survivors: prev slot s -> slot s+1
        surv = np.zeros_like(pb); surv[:, :, :-1] = present[:, :, 1:]
        viol += (was_light & ~surv & same).sum()
        ns = np.zeros_like(sleep)
        ns[:, :, 1:] = np.where(sleep[:, :, :-1] > 0, sleep[:, :, :-1] - 1, k[:, :, 1:])
        ns[:, :, 0] = k[:, :, 0]
        heavy += present[:, :, 0].sum()            # spawns are heavy
        ns = np.where(same, ns, k)                   # (a reset env has no bullets anyway)
        hv = np.zeros_like(present); hv[:, :, 1:] = sleep[:, :, :-1] == 0; hv[:, :, 0] = True
        hist += np.bincount(k[present & hv].ravel(), minlength=13)[:13]
        sleep = np.where(present, ns, 0)
    prev = st
print("N", N, "heavy", heavy, "light", light, "f_awake", heavy / (heavy + light), "violations", viol)
print("sleep values computed at heavy survivors:", hist.tolist())
print("bullets per env", present.sum() / E)
