"""GPU: the headless CUDA rasteriser (bs_render) against the numpy restatement of the
draw order (oracle/raster_oracle.py), pixel for pixel.  Parity against pygame itself is unpinned: the
display and pygame.transform are unavailable on this box."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,cont", [(1, False), (3, True), (5, True)])
def test_frames_match_cpu_restatement(cuda_device, N, cont):
    import torch
    from battlespace import parallel_env
    from battlespace.render import Sprites
    import raster_oracle
    E = 64
    env = parallel_env(n_agents=N, continuous_actions=cont, n_envs=E, seed=11)
    sp = Sprites(cuda_device)
    bad = 0
    for k in range(3):
        env.rollout(25 + 20 * k, "instinct_ff" if cont else "random", "instinct")   # bullets in flight, some damage
        st = env.get_state()
        for e in (0, 17, 63):
            got = env.render_frame(e, sprites=sp).cpu().numpy()
            want = raster_oracle.render(st, e, N, sp.host[0:2], sp.host[2:4])
            assert got.shape == (800, 1200, 3) and got.dtype == np.uint8
            diff = np.any(got != want, axis=-1)
            bad += int(diff.sum())
            assert (got != 255).any()  # something was drawn
    assert bad == 0, bad


def test_recording_writes_a_video(cuda_device, tmp_path):
    """start_recording / export_video (battle_env.py:562-581) through the dict API with show=True."""
    cv2 = pytest.importorskip("cv2")
    from battlespace import parallel_env
    env = parallel_env(n_agents=2, show=True, n_envs=1, seed=3)
    path = str(tmp_path / "game.mp4")
    env.start_recording(path)
    env.reset()
    for t in range(12):
        env.step({a: 1 for a in env.possible_agents})
    env.export_video()
    assert env.video.frames == 12 and os.path.getsize(path) > 1000
    cap = cv2.VideoCapture(path)
    ok, frame = cap.read()
    assert ok and frame.shape == (800, 1200, 3)
