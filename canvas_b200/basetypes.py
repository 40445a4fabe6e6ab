"""The few value types the process API exchanges (fluggo.media.basetypes: v2i, box2i, rgba).

Written for this package; only the behaviour the frame path needs: tuple-like values, box2i built
either from two corners or from four integers, inclusive corners, empty <=> max < min."""
from collections import namedtuple

__all__ = ["v2i", "v2f", "box2i", "rgba"]

v2i = namedtuple("v2i", "x y")
v2f = namedtuple("v2f", "x y")
rgba = namedtuple("rgba", "r g b a")


class box2i(namedtuple("_box2i", "min max")):
    __slots__ = ()

    def __new__(cls, min=v2i(0, 0), max=v2i(-1, -1), max_x=None, max_y=None):
        if max_x is not None or max_y is not None:          # box2i(min_x, min_y, max_x, max_y)
            min, max = v2i(int(min), int(max)), v2i(int(max_x), int(max_y))
        else:
            min, max = v2i(int(min[0]), int(min[1])), v2i(int(max[0]), int(max[1]))
        return super().__new__(cls, min, max)

    @property
    def width(self):
        return builtins_max(0, self.max.x - self.min.x + 1)

    @property
    def height(self):
        return builtins_max(0, self.max.y - self.min.y + 1)

    @property
    def size(self):
        return v2i(self.width, self.height)

    @property
    def empty(self):
        return self.max.x < self.min.x or self.max.y < self.min.y

    def __bool__(self):
        return not self.empty

    def flat(self):
        return (self.min.x, self.min.y, self.max.x, self.max.y)

    def __repr__(self):
        return "box2i(%d, %d, %d, %d)" % self.flat()


import builtins as _b
builtins_max = _b.max
