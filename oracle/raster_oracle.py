"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy, integer arithmetic) of the draw order of
parallel_env.render (/root/reference/envs/battle_env.py:498-560; sprites.py:155-178 Plane.draw,
:265-278 Base.draw, :353-360 Bullet.draw, blitRotate :13-33) for the headless frame the CUDA rasteriser
produces.  Parity status: **unpinned** against pygame itself -- pygame.transform.rotate and the display
are unavailable on this box; what this pins is the documented geometry (half-pixel integer coordinates,
Q14 rotation, nearest-neighbour inverse mapping, pygame's per-pixel-alpha blend formula), pixel for
pixel.  Explosions (Explosion.draw, sprites.py:405-419) are drawn from a caller-supplied list of
(x, y, frame); fonts (names, winner screen) and rounded bar corners are left out on both sides.

Input: the env state as returned by parallel_env.get_state() for ONE env, and the four RGBA sprites."""
import numpy as np

W, H, Q = 1200, 800, 14
RED, BLUE = (138, 24, 26), (0, 93, 135)


def _blend(img, ys, xs, src):
    """pygame alpha blit onto an opaque surface: d += ((s - d) * a + s) >> 8 (a == 0: untouched)."""
    a = src[..., 3].astype(np.int64)
    for c in range(3):
        d = img[ys, xs, c].astype(np.int64)
        s = src[..., c].astype(np.int64)
        img[ys, xs, c] = np.where(a == 0, d, d + (((s - d) * a + s) >> 8)).astype(np.uint8)


def _rotated(img, cx, cy, c, s, w, h, sprite=None, colour=None):
    """sprite (or a solid w x h quad) rotated counter-clockwise about (cx, cy) by the angle whose Q14
    cos / sin are c, s; every screen pixel inverse-maps its centre into the sprite."""
    r = (int(np.hypot(w, h)) + 3)
    x0, x1, y0, y1 = max(cx - r, 0), min(cx + r + 1, W), max(cy - r, 0), min(cy + r + 1, H)
    if x0 >= x1 or y0 >= y1:
        return
    ys, xs = np.mgrid[y0:y1, x0:x1].astype(np.int64)
    dx, dy = 2 * xs + 1 - 2 * cx, 2 * ys + 1 - 2 * cy
    ux, uy = dx * c - dy * s, dx * s + dy * c
    u, v = (ux + (w << Q)) >> (Q + 1), (uy + (h << Q)) >> (Q + 1)
    ok = (u >= 0) & (u < w) & (v >= 0) & (v < h)
    if sprite is None:
        img[ys[ok], xs[ok]] = colour
    else:
        _blend(img, ys[ok], xs[ok], sprite[v[ok], u[ok]])


def _bar(img, cx, cy, dy, hp, colour):
    if hp <= 0:
        return
    w, h = hp * 10, 10
    x0, y0 = cx - (w >> 1), (cy - dy) - (h >> 1)
    clip = lambda a, hi: min(max(a, 0), hi)
    img[clip(y0 - 1, H):clip(y0 + h + 1, H), clip(x0 - 1, W):clip(x0 + w + 1, W)] = 0
    img[clip(y0, H):clip(y0 + h, H), clip(x0, W):clip(x0 + w, W)] = colour


def render(state, e, n_agents, plane_sprites, base_sprites, explosion_frames=None, explosions=()):
    """state: dict of numpy arrays from get_state(); explosion_frames [9,64,64,4] RGBA and explosions =
    (x, y, frame) triples drawn after the bullets (battle_env.py:534-536); returns uint8 [800,1200,3]."""
    N, P = n_agents, 2 * n_agents
    img = np.full((H, W, 3), 255, np.uint8)
    bp, bv = state["bullet_pos"][e], state["bullet_vel"][e]          # [2N,12,2], [2N,12,3]; slot = age-1
    for age in range(11, -1, -1):                                    # oldest first, then agent order
        for i in range(P):
            if not (bv[i, age, 2] & 1):
                continue
            sx, sy = int(bv[i, age, 0]), int(bv[i, age, 1])
            ln = np.sqrt(np.float64(sx * sx + sy * sy))
            c = int(np.rint(sx / ln * (1 << Q))) if ln > 0 else 1 << Q
            s = int(np.rint(-sy / ln * (1 << Q))) if ln > 0 else 0
            _rotated(img, int(bp[i, age, 0]), int(bp[i, age, 1]), c, s, 6, 3, colour=RED if i < N else BLUE)
    for x, y, f in (explosions if explosion_frames is not None else ())[:64]:
        if not 0 <= f < 9:
            continue
        x0, y0 = int(x) - 32, int(y) - 32                               # rect of the 64 x 64 frame centred on the kill
        ys, xs = np.mgrid[max(y0, 0):min(y0 + 64, H), max(x0, 0):min(x0 + 64, W)]
        if ys.size:
            _blend(img, ys, xs, explosion_frames[f][ys - y0, xs - x0])
    for t in range(2):
        bx, by = int(state["base_pos"][e, t, 0]), int(state["base_pos"][e, t, 1])
        x0, y0 = bx - 31, by - 31
        ys, xs = np.mgrid[max(y0, 0):min(y0 + 62, H), max(x0, 0):min(x0 + 62, W)]
        _blend(img, ys, xs, base_sprites[t][ys - y0, xs - x0])
        _bar(img, bx, by, 40, int(state["base_hp"][e, t]), RED if t == 0 else BLUE)
    for i in range(P):
        hp = int(state["plane_hp"][e, i])
        if hp <= 0:
            continue
        x, y, d = int(state["plane_x"][e, i]), int(state["plane_y"][e, i]), float(state["plane_dir"][e, i])
        rad = d * (np.pi / 180.0)
        c, s = int(np.rint(np.cos(rad) * (1 << Q))), int(np.rint(np.sin(rad) * (1 << Q)))
        _rotated(img, x, y, c, s, 50, 48, sprite=plane_sprites[i >= N])
        _bar(img, x, y, 35, hp, RED if i < N else BLUE)
    return img
