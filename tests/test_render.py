"""GPU: the headless CUDA rasteriser (bs_render) against oracle/raster_oracle.py, pixel for pixel.
The oracle restates pygame 2.1.2's transform.rotate (surf_rotate / rotate / rotate90 of src_c/transform.c) and its
SRCALPHA blit for the plane sprites, and the reference's draw order (battle_env.py:528-550).  Parity against pygame
itself stays unpinned: the display and pygame.transform are unavailable on this box.  With
BATTLESPACE_ASSETS=<reference>/assets the reference's own PNGs are drawn instead of the procedural stand-ins."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,cont", [(1, False), (2, False), (3, True), (5, True)])
def test_frames_match_cpu_restatement(cuda_device, N, cont):
    import torch
    from battlespace import parallel_env
    from battlespace.render import Sprites
    import raster_oracle
    E = 64
    env = parallel_env(n_agents=N, continuous_actions=cont, n_envs=E, seed=11)
    sp = Sprites(cuda_device, os.environ.get("BATTLESPACE_ASSETS"))
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


def test_explosions_are_drawn_between_bullets_and_planes(cuda_device):
    """Explosion.draw (sprites.py:405-419, battle_env.py:534-536): frames blitted from the caller's list, over
    the bullets and under bases / planes; frames outside 0..8 and lists longer than 64 are cut."""
    from battlespace import parallel_env
    from battlespace.render import Sprites
    import raster_oracle
    N, E = 3, 8
    env = parallel_env(n_agents=N, continuous_actions=True, n_envs=E, seed=4)
    sp = Sprites(cuda_device)
    env.rollout(40, "instinct_ff", "instinct")
    st = env.get_state()
    e = 5
    px, py = int(st["plane_x"][e, 0]), int(st["plane_y"][e, 0])
    bx, by = int(st["base_pos"][e, 1, 0]), int(st["base_pos"][e, 1, 1])
    ex = [(px + 10, py - 5, 2), (bx - 20, by + 20, 8), (600, 400, 0), (5, 795, 4), (1199, 3, 7), (300, 300, 9), (300, 300, -1)]
    got = env.render_frame(e, sprites=sp, explosions=ex).cpu().numpy()
    want = raster_oracle.render(st, e, N, sp.host[0:2], sp.host[2:4], sp.host_explosion, ex)
    plain = raster_oracle.render(st, e, N, sp.host[0:2], sp.host[2:4])
    assert np.array_equal(got, want)
    assert (want != plain).any(axis=-1).sum() > 500          # the animations are visible
    assert np.array_equal(env.render_frame(e, sprites=sp).cpu().numpy(), plain)   # and the list does not stick
    many = [(40 + 17 * k, 60 + 9 * k, k % 9) for k in range(80)]
    assert np.array_equal(env.render_frame(e, sprites=sp, explosions=many).cpu().numpy(),
                          raster_oracle.render(st, e, N, sp.host[0:2], sp.host[2:4], sp.host_explosion, many))


def test_render_tracks_kills_as_explosions(cuda_device):
    """render() keeps self.explosions like the reference: one animation per killed plane, 9 frames long."""
    from battlespace import parallel_env
    env = parallel_env(n_agents=2, show=True, n_envs=1, seed=3)
    env.reset()
    st = env.get_state()
    env.render()                                             # first frame: remembers the hit points
    st["plane_hp"][0, 1] = 0                                 # a plane dies between two frames
    env.set_state(st)
    env.render()
    assert env.explosions == [[int(st["plane_x"][0, 1]), int(st["plane_y"][0, 1]), 1]]
    for _ in range(8):
        env.render()
    assert env.explosions == []


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
