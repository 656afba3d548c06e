"""Headless frames of an env: ``parallel_env.render`` (battle_env.py:498-560) rasterised
by the library's CUDA kernel into a 1200x800 RGB device tensor -- no window, no graphics API (the pygame
display, fonts and vidmaker are unavailable on this box).  ``Sprites`` holds the RGBA images on the
device: the reference's ``assets/{red,blue}_{plane,base}.png`` and ``explode0..8.png`` when a directory is
given (decoded with cv2), else procedural stand-ins of the same sizes.  ``VideoSink`` replaces ``vidmaker.Video``
(battle_env.py:562-581) with ``cv2.VideoWriter`` on the host.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from ._lib import bs_sprites

PLANE_W, PLANE_H, BASE_W, BASE_H = 50, 48, 62, 62   # sizes of the reference's PNGs (sprites.py:61,226)
RED, BLUE = (138, 24, 26), (0, 93, 135)             # sprites.py:6-7
WIDTH, HEIGHT = 1200, 800                           # sprites.py:9-10
EXPL_W, EXPL_FRAMES = 64, 9                         # sprites.py:392-398: nine frames scaled to 64 x 64


def procedural_sprites():
    """Stand-in artwork: a dart pointing along +x (direction 0) and a bordered square, team coloured."""
    out = {}
    for t, col in enumerate((RED, BLUE)):
        plane = np.zeros((PLANE_H, PLANE_W, 4), np.uint8)
        v, u = np.mgrid[0:PLANE_H, 0:PLANE_W]
        cy = (PLANE_H - 1) / 2.0
        body = np.abs(v - cy) * PLANE_W <= (PLANE_W - 1 - u) * (PLANE_H / 2.0)        # triangle, nose at u = W-1
        notch = (u < 12) & (np.abs(v - cy) < (12 - u))                                # tail notch
        mask = body & ~notch
        plane[mask, :3] = col
        plane[mask, 3] = 255
        edge = mask & ~(np.roll(mask, 1, 0) & np.roll(mask, -1, 0) & np.roll(mask, 1, 1) & np.roll(mask, -1, 1))
        plane[edge, 3] = 128                                                          # soft edge: exercises the blend
        base = np.zeros((BASE_H, BASE_W, 4), np.uint8)
        base[..., :3] = col
        base[..., 3] = 255
        base[6:-6, 6:-6, :3] = 255 - (255 - np.array(col)) // 3
        base[:2, :, 3] = base[-2:, :, 3] = base[:, :2, 3] = base[:, -2:, 3] = 96
        out[t] = (plane, base)
    return out[0][0], out[1][0], out[0][1], out[1][1]


def procedural_explosion():
    """Stand-in animation [9,64,64,4]: a disc that grows and fades from yellow to dark red."""
    v, u = np.mgrid[0:EXPL_W, 0:EXPL_W]
    r2 = (2 * u - (EXPL_W - 1)) ** 2 + (2 * v - (EXPL_W - 1)) ** 2       # integer, in half-pixels squared
    out = np.zeros((EXPL_FRAMES, EXPL_W, EXPL_W, 4), np.uint8)
    for f in range(EXPL_FRAMES):
        rad = 2 * (8 + 3 * f)
        disc, core = r2 <= rad * rad, r2 <= (rad // 2) ** 2
        out[f, disc, :3] = (255 - 20 * f, 160 - 16 * f, 16)
        out[f, core, :3] = (255, 240 - 12 * f, 96)
        out[f, disc, 3] = 255 - 24 * f
    return out


def explosion_from_assets(assets_dir):
    """Explosion.__init__ (sprites.py:392-398): explode{i}.png, colour key BLACK, transform.scale to
    64 x 64 (nearest neighbour).  Colour-keyed pixels become alpha 0."""
    import cv2
    out = np.zeros((EXPL_FRAMES, EXPL_W, EXPL_W, 4), np.uint8)
    for f in range(EXPL_FRAMES):
        im = cv2.imread(os.path.join(assets_dir, f"explode{f}.png"), cv2.IMREAD_UNCHANGED)
        if im is None:
            raise ValueError(f"explode{f}.png not found in {assets_dir}")
        if im.ndim == 2:
            im = np.repeat(im[..., None], 3, 2)
        h, w = im.shape[:2]
        ys, xs = (np.arange(EXPL_W) * h) // EXPL_W, (np.arange(EXPL_W) * w) // EXPL_W
        im = im[ys][:, xs]
        out[f, ..., :3] = im[..., [2, 1, 0]]
        alpha = im[..., 3] if im.shape[2] == 4 else np.full(im.shape[:2], 255, np.uint8)
        out[f, ..., 3] = np.where((im[..., :3] == 0).all(-1), 0, alpha)
    return out


class Sprites:
    def __init__(self, device, assets_dir=None):
        if assets_dir is not None:
            import cv2
            imgs = []
            for name, (w, h) in (("red_plane", (PLANE_W, PLANE_H)), ("blue_plane", (PLANE_W, PLANE_H)),
                                 ("red_base", (BASE_W, BASE_H)), ("blue_base", (BASE_W, BASE_H))):
                im = cv2.imread(os.path.join(assets_dir, name + ".png"), cv2.IMREAD_UNCHANGED)
                if im is None or im.shape != (h, w, 4):
                    raise ValueError(f"{name}.png must be a {w}x{h} RGBA image")
                imgs.append(np.ascontiguousarray(im[..., [2, 1, 0, 3]]))  # BGRA -> RGBA
            expl = explosion_from_assets(assets_dir)
        else:
            imgs = list(procedural_sprites())
            expl = procedural_explosion()
        self.host = imgs
        self.host_explosion = expl
        self.device = torch.device(device)
        self.dev = [torch.from_numpy(np.ascontiguousarray(i)).to(device) for i in imgs]
        self.dev_explosion = torch.from_numpy(np.ascontiguousarray(expl)).to(device)
        self.struct = bs_sprites((C.c_void_p * 2)(self.dev[0].data_ptr(), self.dev[1].data_ptr()),
                                 (C.c_void_p * 2)(self.dev[2].data_ptr(), self.dev[3].data_ptr()),
                                 self.dev_explosion.data_ptr(), None, 0)
        self._expl_list = None

    def with_explosions(self, explosions):
        """explosions: iterable of (x, y, frame); returns the struct for one bs_render call."""
        ex = [tuple(int(v) for v in e) for e in (explosions or [])][:64]
        if ex:
            self._expl_list = torch.tensor(ex, dtype=torch.int32).to(self.device)   # kept alive until the next call
            self.struct.explosions, self.struct.n_explosions = self._expl_list.data_ptr(), len(ex)
        else:
            self.struct.explosions, self.struct.n_explosions = None, 0
        return self.struct


class VideoSink:
    """vidmaker.Video(path, fps, resolution) stand-in: frames are RGB uint8 [800,1200,3]."""

    def __init__(self, path, fps):
        import cv2
        self.path = path
        self.writer = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), float(fps), (WIDTH, HEIGHT))
        self.frames = 0

    def update(self, rgb):
        self.writer.write(np.ascontiguousarray(rgb[..., ::-1]))
        self.frames += 1

    def export(self):
        self.writer.release()
