"""Obstacle primitives of the collision path, flattened for the device.

Mirrors the reference's shapes (motion_planners/tkinter/viewer.py:16-17 `Box`,
`Circle`; `create_box` :108-113, `create_cylinder` :115-116) and its constants
(`STEP_SIZE`, `MIN_PROXIMITY`, viewer.py:69-70).  `World` is the flat, n-dimensional
form the C-ABI takes: an environment box plus arrays of box and sphere obstacles.
Pure host-side data; no arithmetic of the hot path happens here.
"""
from collections import namedtuple

import numpy as np

Box = namedtuple('Box', ['lower', 'upper'])
Circle = namedtuple('Circle', ['center', 'radius'])

STEP_SIZE = 1e-2       # viewer.py:69
MIN_PROXIMITY = 1e-3   # viewer.py:70


def create_box(center, extents):
    """n-D generalisation of viewer.py:108-113 (the reference's is 2-D only); the
    bounds are computed with the same two operations, c - e/2 and c + e/2."""
    center = np.asarray(center, dtype=np.float64)
    extents = np.asarray(extents, dtype=np.float64)
    return Box(center - extents / 2., center + extents / 2.)


def create_cylinder(center, radius):
    return Circle(center, radius)


class World(object):
    """Environment box + obstacle lists in the flat layout of include/mpb200.h.

    box_lo/box_hi: [n_box, D] raw (un-inflated) bounds; sph_c: [n_sph, D]; sph_r: [n_sph].
    `buffer` is the obstacle inflation of point_collides (viewer.py:91, MIN_PROXIMITY);
    the environment test uses no inflation (samplers.py:92)."""

    def __init__(self, dim, env_lo, env_hi, box_lo=None, box_hi=None, sph_c=None, sph_r=None,
                 buffer=MIN_PROXIMITY):
        self.dim = int(dim)
        f = lambda a, shape: np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(shape))
        self.env_lo = f(env_lo, (self.dim,))
        self.env_hi = f(env_hi, (self.dim,))
        self.box_lo = f(box_lo if box_lo is not None else [], (-1, self.dim))
        self.box_hi = f(box_hi if box_hi is not None else [], (-1, self.dim))
        self.sph_c = f(sph_c if sph_c is not None else [], (-1, self.dim))
        self.sph_r = f(sph_r if sph_r is not None else [], (-1,))
        if self.box_lo.shape != self.box_hi.shape:
            raise ValueError('box_lo and box_hi differ in shape')
        if self.sph_c.shape[0] != self.sph_r.shape[0]:
            raise ValueError('sph_c and sph_r differ in length')
        self.buffer = float(buffer)

    @property
    def n_box(self):
        return self.box_lo.shape[0]

    @property
    def n_sph(self):
        return self.sph_c.shape[0]

    @staticmethod
    def from_shapes(environment, obstacles, buffer=MIN_PROXIMITY):
        """From the reference's own objects: `environment` a Box, `obstacles` a list of
        Box / Circle (anything with .lower/.upper or .center/.radius)."""
        env_lo = np.asarray(environment.lower, dtype=np.float64)
        env_hi = np.asarray(environment.upper, dtype=np.float64)
        dim = env_lo.shape[0]
        blo, bhi, sc, sr = [], [], [], []
        for ob in obstacles:
            if hasattr(ob, 'lower') and hasattr(ob, 'upper'):
                blo.append(np.asarray(ob.lower, dtype=np.float64))
                bhi.append(np.asarray(ob.upper, dtype=np.float64))
            elif hasattr(ob, 'center') and hasattr(ob, 'radius'):
                sc.append(np.asarray(ob.center, dtype=np.float64))
                sr.append(float(ob.radius))
            else:
                raise NotImplementedError(ob)   # viewer.py:89
        return World(dim, env_lo, env_hi, blo or None, bhi or None, sc or None, sr or None, buffer)

    def shapes(self):
        """Back to (environment Box, [Box | Circle, ...]) for per-shape consumers."""
        obstacles = [Box(lo.copy(), hi.copy()) for lo, hi in zip(self.box_lo, self.box_hi)]
        obstacles += [Circle(c.copy(), float(r)) for c, r in zip(self.sph_c, self.sph_r)]
        return Box(self.env_lo.copy(), self.env_hi.copy()), obstacles
