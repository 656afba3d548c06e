"""Headless frames of an env: ``parallel_env.render`` (battle_env.py:498-560) rasterised
by the library's CUDA kernel into a 1200x800 RGB device tensor -- no window, no graphics API (the pygame
display, fonts and vidmaker are unavailable on this box).  ``Sprites`` holds the four RGBA images on the
device: the reference's ``assets/{red,blue}_{plane,base}.png`` when a directory is given (decoded with
cv2), else procedural stand-ins of the same sizes.  ``VideoSink`` replaces ``vidmaker.Video``
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
        else:
            imgs = list(procedural_sprites())
        self.host = imgs
        self.dev = [torch.from_numpy(np.ascontiguousarray(i)).to(device) for i in imgs]
        self.struct = bs_sprites((C.c_void_p * 2)(self.dev[0].data_ptr(), self.dev[1].data_ptr()),
                                 (C.c_void_p * 2)(self.dev[2].data_ptr(), self.dev[3].data_ptr()))


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
