"""Batched curve collision checks (motion_planners/trajectory/smooth.py:32-50 get_curve_collision_fn).

The reference's trajectory package (spline retiming, parabolic ramps) is host-side numerics outside the hot path;
what touches the path is the verdict `any(map(collision_fn, default_selector(samples)))` over the points its
discretiser takes along a curve.  That `any` is order-free, so all samples of a curve go through collision_fn in
one launch.  The discretiser is the caller's (the reference's trajectory.discretize.distance_discretize_curve, or
anything with its signature)."""
import numpy as np

from .primitives import GpuCollisionFn
from .utils import require_device_primitives


def get_curve_collision_fn(collision_fn, discretize_fn, max_velocities=None, max_accelerations=None, allow_host=False):
    """-> curve_collision_fn(curve, t0=None, t1=None) -> bool, True when the curve is None or some sample of
    discretize_fn(curve, start_t=t0, end_t=t1)[1] collides (smooth.py:34-49; the dynamical-limit tests are commented
    out in the reference and stay out)."""
    if not isinstance(collision_fn, GpuCollisionFn):
        require_device_primitives('get_curve_collision_fn', allow_host)

    def curve_collision_fn(curve, t0=None, t1=None):
        if curve is None:
            return True
        _, samples = discretize_fn(curve, start_t=t0, end_t=t1)
        samples = np.asarray(list(samples), dtype=np.float64)
        if samples.size == 0:
            return False
        if isinstance(collision_fn, GpuCollisionFn):
            return bool(collision_fn.batch(samples.reshape(len(samples), -1)).any().item())
        return any(map(collision_fn, samples))

    def batch(sample_lists):
        """Verdicts for many curves at once: sample_lists = [samples of curve 0, samples of curve 1, ...]."""
        import torch
        counts = [len(s) for s in sample_lists]
        if sum(counts) == 0:
            return [False] * len(counts)
        flat = np.concatenate([np.asarray(s, dtype=np.float64).reshape(len(s), -1) for s in sample_lists if len(s)])
        hit = collision_fn.batch(flat).to(torch.int32)
        owner = torch.repeat_interleave(torch.arange(len(counts), device=hit.device),
                                        torch.tensor(counts, device=hit.device))
        out = torch.zeros(len(counts), dtype=torch.int32, device=hit.device).index_add_(0, owner, hit)
        return [bool(v) for v in (out > 0).cpu().numpy()]

    curve_collision_fn.batch = batch
    return curve_collision_fn
