"""Soak (not collected by pytest): the fused rollout against the C oracle at the BASELINE sizes for several seeds and
team sizes, longer than tests/test_full_size_oracle.py runs it.  Every counter, the whole state, every bullet slot and
the last launch's replay rows are compared bit for bit (the test module's own _compare).
usage (GPU box): python tests/tools/soak_vs_oracle.py [seeds...]"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle"))
import conftest  # noqa: F401,E402  (puts the package on sys.path)
from test_full_size_oracle import _compare  # noqa: E402
from port import POL_INSTINCT, POL_INSTINCT_FF, POL_RANDOM  # noqa: E402

seeds = [int(x) for x in sys.argv[1:]] or [101, 202, 303]
for seed in seeds:
    for N, E, T, L, pr, pb in ((3, 65536, 32, 12, POL_RANDOM, POL_INSTINCT), (1, 32768, 32, 8, POL_RANDOM, POL_RANDOM),
                               (2, 32768, 32, 8, POL_INSTINCT, POL_RANDOM), (5, 32768, 32, 6, POL_INSTINCT_FF, POL_INSTINCT_FF),
                               (8, 8192, 32, 6, POL_INSTINCT_FF, POL_RANDOM)):
        occ, st = _compare(N, E, T, L, pr, pb, ring_ticks=4, env_offset=seed * 1000, seed=seed)
        print("seed", seed, "N", N, "E", E, "ticks", T * L, "games", int(st[3]), "live bullets per env at the end", int(occ), "OK", flush=True)
