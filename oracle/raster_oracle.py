"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy, integer arithmetic) of parallel_env.render
(/root/reference/envs/battle_env.py:498-560; sprites.py:155-178 Plane.draw, :265-278 Base.draw, :353-360
Bullet.draw, blitRotate :13-33) for the headless frame the CUDA rasteriser produces.

The pixel work of the reference is done by a third-party dependency that is absent here: **pygame==2.1.2**
(requirements.txt:6), call sites sprites.py:29 (transform.rotate) and :161,168,265,360,414 (Surface.blit).  Its
published algorithms are restated below, from src_c/transform.c and src_c/alphablit.c / surface.h of that release:
  * surf_rotate(): a multiple of 90 degrees goes to rotate90() (exact pixel permutation, numturns = (angle / 90) % 4);
    otherwise the destination is nx x ny with nx = (int)max(|cx+sy|, |cx-sy|, |-cx+sy|, |-cx-sy|), ny likewise from
    (sx, cy), where (cx, cy, sx, sy) = (cos*w, cos*h, sin*w, sin*h) and radangle = angle * .01745329251994329;
  * rotate(): 16.16 fixed-point nearest-neighbour inverse mapping -- isin = (int)(sin * 65536), icos = (int)(cos * 65536),
    ax = (nx << 15) - (int)(cos * ((nx - 1) << 15)), ay = (ny << 15) - (int)(sin * ((nx - 1) << 15)),
    xd = (w - nx) << 15, yd = (h - ny) << 15; row y starts at dx = ax + isin * (ny / 2 - y) + xd,
    dy = ay - icos * (ny / 2 - y) + yd and steps (icos, isin) per destination pixel; a pixel whose (dx, dy) leaves
    [0, (w << 16) - 1] x [0, (h << 16) - 1] takes the background (transparent for an SRCALPHA surface);
  * the blit of an SRCALPHA surface onto the opaque display: ALPHA_BLEND, d = ((d << 8) + (s - d) * a + s) >> 8;
  * blitRotate (sprites.py:13-33) with the even-sized 50 x 48 sprite: the rotated surface's rect is centred on the
    plane's rect centre, Rect.center setter x = cx - (w >> 1).
Parity status: **unpinned** -- pygame itself (display, transform) is unavailable on this box, so this restatement
could not be run against it; the CUDA kernel is checked against this file pixel for pixel.  Not restated: fonts
(plane names sprites.py:163-168, winner text battle_env.py:426-447: no font rasteriser here), border_radius = 3 of
the hit-point bars (square corners on both sides), and a bullet's exact direction (the env state keeps a bullet's
integer step, so its 6 x 3 quad is drawn along that step with Q14 fixed point instead of through transform.rotate).
Explosions (Explosion.draw, sprites.py:405-419) are drawn from a caller-supplied list of (x, y, frame).

Input: the env state as returned by parallel_env.get_state() for ONE env, and the four RGBA sprites."""
import math

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


def pygame_rotate(src, angle):
    """pygame.transform.rotate(surface, angle) for an RGBA uint8 image [h, w, 4] (see the module docstring)."""
    h, w = src.shape[:2]
    if math.fmod(angle, 90.0) == 0.0:                                   # rotate90()
        turns = int(angle / 90.0) % 4 if angle >= 0 else -(int(-angle / 90.0) % 4)   # C: (int)(angle / 90) % 4
        if turns < 0:
            turns += 4
        nx, ny = (h, w) if turns % 2 else (w, h)
        y, x = np.mgrid[0:ny, 0:nx]
        u = [x, w - 1 - y, w - 1 - x, y][turns]
        v = [y, x, h - 1 - y, h - 1 - x][turns]
        return src[v, u].copy()
    rad = angle * .01745329251994329
    sangle, cangle = math.sin(rad), math.cos(rad)
    cx, cy, sx, sy = cangle * w, cangle * h, sangle * w, sangle * h
    nx = int(max(abs(cx + sy), abs(cx - sy), abs(-cx + sy), abs(-cx - sy)))
    ny = int(max(abs(sx + cy), abs(sx - cy), abs(-sx + cy), abs(-sx - cy)))
    isin, icos = int(sangle * 65536), int(cangle * 65536)
    ax = (nx << 15) - int(cangle * ((nx - 1) << 15))
    ay = (ny << 15) - int(sangle * ((nx - 1) << 15))
    xd, yd = (w - nx) * 32768, (h - ny) * 32768
    y, x = np.mgrid[0:ny, 0:nx].astype(np.int64)
    cyd = ny // 2
    dx = ax + isin * (cyd - y) + xd + icos * x
    dy = ay - icos * (cyd - y) + yd + isin * x
    ok = (dx >= 0) & (dy >= 0) & (dx <= (w << 16) - 1) & (dy <= (h << 16) - 1)
    dst = np.zeros((ny, nx, 4), np.uint8)                              # bgcolor: transparent
    dst[ok] = src[dy[ok] >> 16, dx[ok] >> 16]
    return dst


def _blit(img, surf, left, top):
    """Surface.blit(surf, rect) clipped to the display."""
    ny, nx = surf.shape[:2]
    ys, xs = np.mgrid[max(top, 0):min(top + ny, H), max(left, 0):min(left + nx, W)]
    if ys.size:
        _blend(img, ys, xs, surf[ys - top, xs - left])


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
        rot = pygame_rotate(plane_sprites[i >= N], d)                 # blitRotate, sprites.py:13-33,160-161
        _blit(img, rot, x - (rot.shape[1] >> 1), y - (rot.shape[0] >> 1))
        _bar(img, x, y, 35, hp, RED if i < N else BLUE)
    return img
