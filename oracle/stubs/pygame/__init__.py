"""TEST INFRASTRUCTURE ONLY -- stand-in for pygame==2.1.2 (reference requirements.txt:6).

The reference env (envs/battle_env.py, envs/sprites.py) touches pygame only for
integer rectangle arithmetic on the hot path; pygame itself is not installable
here (no network, no SDL).  This module restates exactly the pygame 2.1.2
behaviour those call sites rely on (rect.c / base.c of that release):

  * every Rect setter converts its value with a C cast: floats truncate toward
    zero (sprites.py:84,131,333 assign float tuples to ``rect.center``);
  * ``center`` reads ``x + (w >> 1), y + (h >> 1)`` and writes
    ``x = v - (w >> 1)``; ``right`` / ``bottom`` writes move ``x`` / ``y``;
  * ``colliderect`` is a strict-inequality overlap test, false for empty rects,
    and accepts any object carrying a ``.rect`` attribute (sprites.py:344 passes
    the Base sprite itself).

Nothing here draws.  The display/font/draw/time/event parts of pygame are only
reached with ``show=True`` and are left out on purpose (no graphics stack).
"""
import math as _math
import struct as _struct

SRCALPHA = 0x00010000


def init():
    return (0, 0)


def _ci(v):
    """C ``(int)`` cast: truncate toward zero."""
    return int(v)


class Rect:
    __slots__ = ("x", "y", "w", "h")

    def __init__(self, x, y=None, w=None, h=None):
        if y is None:  # Rect(rect) / Rect((x, y, w, h))
            src = x
            if isinstance(src, Rect):
                x, y, w, h = src.x, src.y, src.w, src.h
            else:
                x, y, w, h = src
        elif w is None:  # Rect((x, y), (w, h))
            (x, y), (w, h) = x, y
        self.x, self.y, self.w, self.h = _ci(x), _ci(y), _ci(w), _ci(h)

    # --- scalar attributes -------------------------------------------------
    width = property(lambda s: s.w, lambda s, v: setattr(s, "w", _ci(v)))
    height = property(lambda s: s.h, lambda s, v: setattr(s, "h", _ci(v)))
    left = property(lambda s: s.x, lambda s, v: setattr(s, "x", _ci(v)))
    top = property(lambda s: s.y, lambda s, v: setattr(s, "y", _ci(v)))
    right = property(lambda s: s.x + s.w, lambda s, v: setattr(s, "x", _ci(v) - s.w))
    bottom = property(lambda s: s.y + s.h, lambda s, v: setattr(s, "y", _ci(v) - s.h))
    centerx = property(lambda s: s.x + (s.w >> 1),
                       lambda s, v: setattr(s, "x", _ci(v) - (s.w >> 1)))
    centery = property(lambda s: s.y + (s.h >> 1),
                       lambda s, v: setattr(s, "y", _ci(v) - (s.h >> 1)))

    # --- pair attributes ---------------------------------------------------
    def _get_center(self):
        return (self.x + (self.w >> 1), self.y + (self.h >> 1))

    def _set_center(self, v):
        a, b = v
        self.x = _ci(a) - (self.w >> 1)
        self.y = _ci(b) - (self.h >> 1)

    center = property(_get_center, _set_center)

    def _get_topleft(self):
        return (self.x, self.y)

    def _set_topleft(self, v):
        a, b = v
        self.x, self.y = _ci(a), _ci(b)

    topleft = property(_get_topleft, _set_topleft)

    def _get_size(self):
        return (self.w, self.h)

    size = property(_get_size)

    def colliderect(self, other):
        if not isinstance(other, Rect):
            other = getattr(other, "rect", other)
            if callable(other):
                other = other()
            if not isinstance(other, Rect):
                other = Rect(other)
        if self.w == 0 or self.h == 0 or other.w == 0 or other.h == 0:
            return False
        return (self.x < other.x + other.w and self.y < other.y + other.h
                and self.x + self.w > other.x and self.y + self.h > other.y)

    def copy(self):
        return Rect(self.x, self.y, self.w, self.h)

    def __repr__(self):
        return f"<rect({self.x}, {self.y}, {self.w}, {self.h})>"


class Surface:
    def __init__(self, size, flags=0, depth=0):
        self._w, self._h = int(size[0]), int(size[1])

    def get_size(self):
        return (self._w, self._h)

    def get_width(self):
        return self._w

    def get_height(self):
        return self._h

    def get_rect(self, **attrs):
        r = Rect(0, 0, self._w, self._h)
        for k, v in attrs.items():
            setattr(r, k, v)
        return r

    def fill(self, *a, **k):
        return None

    def blit(self, *a, **k):
        return None

    def set_colorkey(self, *a, **k):
        return None


class _Image:
    @staticmethod
    def load(path):
        with open(path, "rb") as f:
            head = f.read(24)
        if head[:8] != b"\x89PNG\r\n\x1a\n":
            raise ValueError(f"not a PNG: {path}")
        w, h = _struct.unpack(">II", head[16:24])
        return Surface((w, h))


image = _Image()


class _SpriteBase:
    def __init__(self, *groups):
        pass

    def kill(self):
        pass


class _SpriteMod:
    Sprite = _SpriteBase


sprite = _SpriteMod()


class Vector2:
    __slots__ = ("x", "y")

    def __init__(self, x=0.0, y=None):
        if y is None:
            x, y = x
        self.x, self.y = float(x), float(y)

    def __sub__(self, o):
        ox, oy = (o.x, o.y) if isinstance(o, Vector2) else o
        return Vector2(self.x - ox, self.y - oy)

    def rotate(self, deg):
        r = _math.radians(deg)
        c, s = _math.cos(r), _math.sin(r)
        return Vector2(self.x * c - self.y * s, self.x * s + self.y * c)


class _MathMod:
    Vector2 = Vector2


math = _MathMod()


class _Transform:
    @staticmethod
    def rotate(surf, angle):
        # Only the output size is ever read, and it cancels in
        # ``get_rect(center=c).centerx`` (sprites.py:31,201).
        w, h = surf.get_size()
        r = _math.radians(angle)
        c, s = abs(_math.cos(r)), abs(_math.sin(r))
        return Surface((max(1, int(w * c + h * s)), max(1, int(w * s + h * c))))

    @staticmethod
    def scale(surf, size):
        return Surface(size)


transform = _Transform()


class _Display:
    @staticmethod
    def quit():
        return None


display = _Display()
