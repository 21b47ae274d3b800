"""Seeded synthetic workloads for BASELINE.json's configs.

Everything is drawn from numpy's frozen legacy `RandomState`, so the same seed gives
the same bytes on every host.  `RandomState(seed).uniform(size=(N, D))` reproduces the
stream the reference's `sample_fn` consumes for a unit-box environment
(motion_planners/utils.py:246-249 via tkinter/samplers.py:32-47).
"""
import numpy as np

from .arm import ArmWorld, seven_dof_arm
from .geometry import World, create_box, create_cylinder


def unit_env(dim):
    return np.zeros(dim), np.ones(dim)


def random_boxes(dim, n_box, ext_lo, ext_hi, seed):
    """n_box axis-aligned boxes: centre ~ U(0,1)^D then per-axis extent ~ U(ext_lo, ext_hi),
    drawn box by box from RandomState(seed); bounds = c -/+ e/2."""
    rs = np.random.RandomState(seed)
    lo = np.empty((n_box, dim))
    hi = np.empty((n_box, dim))
    for b in range(n_box):
        c = rs.uniform(size=dim)
        e = rs.uniform(ext_lo, ext_hi, size=dim)
        box = create_box(c, e)
        lo[b], hi[b] = box.lower, box.upper
    return lo, hi


def problem1_world():
    """The reference's 2-D demo fixture (motion_planners/tkinter/run.py:108-129):
    three boxes and one disc in the unit square, start (0.1, 0.1), goal (0.8, 0.8)."""
    boxes = [create_box((.35, .75), (.25, .25)),
             create_box((.75, .35), (.22, .22)),
             create_box((.5, .5), (.25, .25))]
    disc = create_cylinder((.25, .25), .1)
    env = create_box((.5, .5), (1., 1.))
    world = World.from_shapes(env, boxes + [disc])
    start = np.array([0.1, 0.1])
    goal = np.average([create_box((.8, .8), (.1, .1)).lower, create_box((.8, .8), (.1, .1)).upper], axis=0)
    return world, start, goal


def boxes_world(dim, n_box, ext_lo, ext_hi, seed=0, n_sph=0, sph_r=(0.05, 0.15)):
    lo, hi = random_boxes(dim, n_box, ext_lo, ext_hi, seed)
    env_lo, env_hi = unit_env(dim)
    sc = sr = None
    if n_sph:
        rs = np.random.RandomState(seed + 7919)
        sc = rs.uniform(size=(n_sph, dim))
        sr = rs.uniform(sph_r[0], sph_r[1], size=n_sph)
    return World(dim, env_lo, env_hi, lo, hi, sc, sr)


def c2_world(seed=0):
    """Config 2: 6-DoF unit cube, 64 random boxes, extents U(0.2, 0.6) (88 % free)."""
    return boxes_world(6, 64, 0.2, 0.6, seed)


def c3_world(seed=0):
    """Config 3: 7-DoF unit cube, 64 random boxes, extents U(0.3, 0.7) (about 80 % free)."""
    return boxes_world(7, 64, 0.3, 0.7, seed)


def c4lite_world(seed=0):
    """Config 4 with reference primitives only: 7-DoF unit cube, 256 C-space boxes,
    extents U(0.1, 0.5).  (The sphere-approximated arm of config 4 has no counterpart in
    the reference; see DESIGN.md.)"""
    return boxes_world(7, 256, 0.1, 0.5, seed)


def c4_world(seed=0, n_box=256, n_sph=16):
    """Config 4 as written: a 7-DoF arm approximated by 16 spheres among 256 boxes in the 3-D workspace.
    Boxes: centre ~ U([-0.9, 0.9]^2 x [-0.1, 1.1]), per-axis extent ~ U(0.03, 0.12), drawn box by box from
    RandomState(seed); a box is redrawn while it stands in the column around the arm's fixed first links (every
    configuration would collide with it).  About half of the uniform configurations are collision-free."""
    rs = np.random.RandomState(seed)
    lo, hi = [], []
    while len(lo) < n_box:
        c = rs.uniform([-0.9, -0.9, -0.1], [0.9, 0.9, 1.1])
        e = rs.uniform(0.03, 0.12, size=3)
        if np.hypot(c[0], c[1]) < 0.25 + e[:2].max() / 2 and c[2] - e[2] / 2 < 0.45:
            continue
        box = create_box(c, e)
        lo.append(box.lower)
        hi.append(box.upper)
    return ArmWorld(seven_dof_arm(n_sph), np.array(lo), np.array(hi))


def c5_world(seed=0):
    """Config 5: 3-D unit cube, 32 boxes + 8 spheres."""
    return boxes_world(3, 32, 0.05, 0.25, seed, n_sph=8)


def uniform_samples(n, dim, seed=0):
    return np.random.RandomState(seed).uniform(size=(n, dim))


WORKLOADS = {
    # name: (world factory, dim, n_samples, k)
    'c1_problem1_2d_2k': (lambda: problem1_world()[0], 2, 2000, 10),
    'c2_lazyprm_6d_100k_64box': (c2_world, 6, 100000, 10),
    'c4lite_prm_7d_1m_256box': (c4lite_world, 7, 1000000, 10),
}
