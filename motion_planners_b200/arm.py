"""The collision model of BASELINE.json configs[3]: a 7-DoF serial arm approximated by spheres, against
axis-aligned boxes in the 3-D workspace.  Host-side data only; the arithmetic is in csrc/arm.cu.

The reference has no robot model: its collision_fn is whatever callable the user passes
(motion_planners/README.md:52-59; the demo builds one at tkinter/samplers.py:88-98 from the box / circle tests of
tkinter/viewer.py:72-82).  This module describes such a callable for a manipulator in the reference's terms --
closed environment bounds on the configuration, `norm(...) <= radius + buffer` for the obstacle test -- so that
the drop-in planners take it like any other collision_fn (primitives.GpuCollisionFn(ArmWorld(...))).

Configuration q in [0, 1]^n; joint angle i = q[i] * theta_span[i] + theta_lo[i].  Frame i is frame i-1 turned
about its local `axis[i]` (1 = y, 2 = z); link i runs `a[i]` along the turned frame's x and `d[i]` along its z.
Sphere j rides on frame `sph_link[j]` at `sph_u[j]` along x and `sph_w[j]` along z from the frame's origin.
"""
import numpy as np

from .geometry import MIN_PROXIMITY

MAX_JOINTS = 8
MAX_SPHERES = 32


class ArmModel(object):
    def __init__(self, axis, theta_lo, theta_span, a, d, base, sph_link, sph_u, sph_w, sph_r):
        f = lambda v: np.ascontiguousarray(np.asarray(v, dtype=np.float64))
        self.axis = [int(v) for v in axis]
        self.n_joints = len(self.axis)
        self.theta_lo, self.theta_span, self.a, self.d = f(theta_lo), f(theta_span), f(a), f(d)
        self.base = f(base)
        self.sph_link = [int(v) for v in sph_link]
        self.n_sph = len(self.sph_link)
        self.sph_u, self.sph_w, self.sph_r = f(sph_u), f(sph_w), f(sph_r)
        if not (1 <= self.n_joints <= MAX_JOINTS and 1 <= self.n_sph <= MAX_SPHERES):
            raise ValueError("at most %d joints and %d spheres" % (MAX_JOINTS, MAX_SPHERES))
        if any(ax not in (1, 2) for ax in self.axis):
            raise ValueError("joint axes are 1 (y) or 2 (z)")
        if any(not 0 <= l < self.n_joints for l in self.sph_link):
            raise ValueError("sphere attached to a joint that does not exist")
        for arr, n in ((self.theta_lo, self.n_joints), (self.theta_span, self.n_joints), (self.a, self.n_joints),
                       (self.d, self.n_joints), (self.base, 3), (self.sph_u, self.n_sph), (self.sph_w, self.n_sph),
                       (self.sph_r, self.n_sph)):
            if arr.shape != (n,):
                raise ValueError("model arrays have the wrong length")


def seven_dof_arm(n_sph=16):
    """A 7-joint arm of the usual z-y-z-y-z-y-z layout, 1.0 long, `n_sph` spheres spread along its links."""
    axis = [2, 1, 2, 1, 2, 1, 2]
    span = np.array([1.8, 1.2, 1.8, 1.2, 1.8, 1.2, 1.8]) * np.pi
    lo = -span / 2.0
    a = np.zeros(7)
    d = np.array([0.16, 0.0, 0.30, 0.0, 0.28, 0.0, 0.12])     # shoulder, upper arm, forearm, hand along local z
    base = np.array([0.0, 0.0, 0.14])
    # spheres along the z-links: (joint frame, offset along z, radius)
    spec = [(0, 0.00, 0.070), (0, 0.09, 0.065),
            (2, 0.00, 0.065), (2, 0.075, 0.060), (2, 0.150, 0.060), (2, 0.225, 0.058),
            (4, 0.00, 0.058), (4, 0.07, 0.055), (4, 0.14, 0.052), (4, 0.21, 0.050),
            (6, 0.00, 0.048), (6, 0.06, 0.045), (6, 0.12, 0.040), (6, 0.17, 0.035),
            (1, 0.00, 0.068), (3, 0.00, 0.060)]
    spec = spec[:n_sph]
    return ArmModel(axis, lo, span, a, d, base, [s[0] for s in spec], np.zeros(len(spec)), [s[1] for s in spec],
                    [s[2] for s in spec])


class ArmWorld(object):
    """Arm model + workspace boxes + bounds on the configuration.  `dim` is the number of joints (what the
    planners see); box_lo / box_hi are [n_box, 3] raw bounds; `buffer` inflates the sphere radii
    (viewer.py:82 `radius + buffer`, MIN_PROXIMITY)."""

    def __init__(self, model, box_lo, box_hi, env_lo=None, env_hi=None, buffer=MIN_PROXIMITY):
        self.model = model
        self.dim = model.n_joints
        f = lambda v, shape: np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(shape))
        self.box_lo, self.box_hi = f(box_lo, (-1, 3)), f(box_hi, (-1, 3))
        if self.box_lo.shape != self.box_hi.shape:
            raise ValueError("box_lo and box_hi differ in shape")
        self.env_lo = f(env_lo if env_lo is not None else np.zeros(self.dim), (self.dim,))
        self.env_hi = f(env_hi if env_hi is not None else np.ones(self.dim), (self.dim,))
        self.buffer = float(buffer)

    @property
    def n_box(self):
        return self.box_lo.shape[0]

    @property
    def n_sph(self):
        return self.model.n_sph
