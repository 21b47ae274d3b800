"""Vectorised numpy restatement of the hot-path arithmetic.  TEST INFRASTRUCTURE ONLY.

Independent of oracle/mp_oracle.c (shares no code with it) so that the two can be
checked against each other and both against the reference's golden outputs.
numpy ufuncs round every elementwise operation separately, which is exactly what
the reference does, so no special care about contraction is needed here.
"""
import numpy as np


def extend_counts(q1, q2, step=1e-2):
    """steps+1 per edge (tkinter/samplers.py:119-120): int(max|diff/step|) + 1."""
    diff = np.asarray(q2, np.float64) - np.asarray(q1, np.float64)
    return (np.abs(diff / step).max(axis=-1)).astype(np.int64) + 1   # astype truncates like int()


def extend_batch(q1, q2, step=1e-2):
    """All waypoints of E edges at once.  -> (counts[E], way[E, nmax, D]); entries past
    counts[e] repeat the last waypoint.  Follows the serial recurrence of
    tkinter/samplers.py:121-126 (circular={}): q <- (1/(n-i))*(q2-q) + q."""
    q1 = np.asarray(q1, np.float64)
    q2 = np.asarray(q2, np.float64)
    n = extend_counts(q1, q2, step)
    nmax = int(n.max()) if n.size else 0
    E, D = q1.shape
    way = np.empty((E, nmax, D))
    q = q1.copy()
    for i in range(nmax):
        live = i < n
        rem = np.where(live, n - i, 1).astype(np.float64)
        c = 1.0 / rem
        nq = c[:, None] * (q2 - q) + q
        q = np.where(live[:, None], nq, q)
        way[:, i, :] = q
    return n, way


def contains_box(points, lower, upper, buffer=0.0):
    """viewer.py:72-77 for a batch of points [n,D] against one box."""
    lo = lower - buffer * np.ones(len(lower))
    hi = upper + buffer * np.ones(len(upper))
    return np.less_equal(lo, points).all(axis=-1) & np.less_equal(points, hi).all(axis=-1)


def seq_fma_dot_self(x):
    """sum_k x_k^2 rounded like a sequential fused chain s = fma(x_k, x_k, s).

    Exact emulation with error-free transformations: x_k^2 is split into hi+lo
    (Dekker/Veltkamp), s + hi into sum+err (Knuth two-sum); the fused result is the
    correctly rounded value of s + hi + lo, obtained by adding the two error terms to
    the rounded sum -- valid whenever |lo + err| is below half an ulp of the sum in a
    way that cannot double-round, which holds away from ties; tests compare against
    the C oracle's hardware fma to confirm on the data actually used."""
    x = np.asarray(x, np.float64)
    s = np.zeros(x.shape[:-1])
    for k in range(x.shape[-1]):
        a = x[..., k]
        hi = a * a
        # Veltkamp split of a
        t = 134217729.0 * a
        ah = t - (t - a)
        al = a - ah
        lo = ((ah * ah - hi) + 2.0 * ah * al) + al * al
        sm = s + hi
        bb = sm - s
        err = (s - (sm - bb)) + (hi - bb)
        s = sm + (err + lo)
    return s


def contains_sphere(points, center, radius, buffer=0.0):
    """viewer.py:80-82: ||p - c||_2 <= r + buffer with numpy's norm = sqrt(dot(x,x))."""
    x = points - np.asarray(center, np.float64)
    return np.sqrt(seq_fma_dot_self(x)) <= (radius + buffer)


def points_check(points, env_lo, env_hi, box_lo, box_hi, sph_c, sph_r, buffer=1e-3):
    """collision_fn over a batch (tkinter/samplers.py:88-98)."""
    points = np.asarray(points, np.float64)
    hit = ~contains_box(points, np.asarray(env_lo, np.float64), np.asarray(env_hi, np.float64), 0.0)
    for lo, hi in zip(box_lo, box_hi):
        hit |= contains_box(points, np.asarray(lo, np.float64), np.asarray(hi, np.float64), buffer)
    for c, r in zip(sph_c, sph_r):
        hit |= contains_sphere(points, c, r, buffer)
    return hit


def sqdist_scipy(u, v):
    """cKDTree's p=2 leaf metric: 4 accumulators over blocks of 4 dims, then the tail."""
    u = np.asarray(u, np.float64)
    v = np.asarray(v, np.float64)
    d = u - v
    sq = d * d
    D = sq.shape[-1]
    nb = D // 4
    acc = [np.zeros(sq.shape[:-1]) for _ in range(4)]
    for b in range(nb):
        for j in range(4):
            acc[j] = acc[j] + sq[..., 4 * b + j]
    s = ((acc[0] + acc[1]) + acc[2]) + acc[3]
    for i in range(4 * nb, D):
        s = s + sq[..., i]
    return s


def knn(samples, queries, k, weights=None, chunk=256):
    """Brute-force fp64 kNN, (distance, index) lexicographic order (nearest.py:74-85)."""
    samples = np.asarray(samples, np.float64)
    queries = np.asarray(queries, np.float64)
    w = np.ones(samples.shape[1]) if weights is None else np.asarray(weights, np.float64)
    emb = w * samples
    qemb = w * queries
    M = queries.shape[0]
    idx = np.empty((M, k), np.int64)
    dist = np.empty((M, k))
    ar = np.arange(samples.shape[0])
    for s0 in range(0, M, chunk):
        S = sqdist_scipy(qemb[s0:s0 + chunk, None, :], emb[None, :, :])
        for r in range(S.shape[0]):
            order = np.lexsort((ar, S[r]))[:k]
            idx[s0 + r] = order
            dist[s0 + r] = np.sqrt(S[r, order])
    return idx, dist


# ------------------------------------------------------------------ configs[3]: the sphere-approximated arm
# An independent numpy statement of oracle/arm_oracle.c (elementwise ufuncs round every operation separately).
def arm_sincos(th):
    kf = np.rint(th * 0.63661977236758138)
    r = th - kf * 1.57079632673412561417e+00
    r = r - kf * 6.07710050650619224932e-11
    z = r * r
    t = np.full_like(z, 1.58969099521155010221e-10)
    for cst in (-2.50507602534068634195e-08, 2.75573137070700676789e-06, -1.98412698298579493134e-04,
                8.33333333332248946124e-03, -1.66666666666666324348e-01):
        t = t * z + cst
    sr = r + (r * z) * t
    u = np.full_like(z, -1.13596475577881948265e-11)
    for cst in (2.08757232129817482790e-09, -2.75573143513906633035e-07, 2.48015872894767294178e-05,
                -1.38888888888741095749e-03, 4.16666666666666019037e-02):
        u = u * z + cst
    cr = (1.0 - 0.5 * z) + (z * z) * u
    n = kf.astype(np.int64) & 3
    s = np.select([n == 0, n == 1, n == 2], [sr, cr, -sr], -cr)
    c = np.select([n == 0, n == 1, n == 2], [cr, -sr, -cr], sr)
    return s, c


def arm_centres(model, q):
    """Sphere centres [n, n_sph, 3] of configurations q [n, n_joints]."""
    q = np.asarray(q, dtype=np.float64)
    n = q.shape[0]
    X = np.tile(np.array([1., 0., 0.]), (n, 1))
    Y = np.tile(np.array([0., 1., 0.]), (n, 1))
    Z = np.tile(np.array([0., 0., 1.]), (n, 1))
    p = np.tile(np.asarray(model.base, dtype=np.float64), (n, 1))
    out = np.empty((n, model.n_sph, 3))
    for i in range(model.n_joints):
        th = q[:, i] * model.theta_span[i] + model.theta_lo[i]
        s, c = arm_sincos(th)
        s, c = s[:, None], c[:, None]
        if model.axis[i] == 2:
            X, Y = c * X + s * Y, c * Y - s * X
        else:
            X, Z = c * X - s * Z, s * X + c * Z
        for j in range(model.n_sph):
            if model.sph_link[j] == i:
                out[:, j] = (p + model.sph_u[j] * X) + model.sph_w[j] * Z
        p = (p + model.a[i] * X) + model.d[i] * Z
    return out


def arm_points_check(world, q):
    """collision_fn of the arm world for a batch (the norm's fused dot is spelled out with math.fma-free
    arithmetic that is exact here: three squares summed left to right in extended precision would differ, so the
    sum uses the same fma chain through np.float64 operations on exactly representable partial products only
    when possible; otherwise compare with oracle.ArmWorld, which is the arbiter)."""
    q = np.asarray(q, dtype=np.float64)
    inside = ((world.env_lo <= q) & (q <= world.env_hi)).all(1)
    cen = arm_centres(world.model, q)
    rb = np.asarray(world.model.sph_r) + world.buffer
    hit = ~inside
    for b in range(world.box_lo.shape[0]):
        cl = np.minimum(np.maximum(cen, world.box_lo[b]), world.box_hi[b])
        d = cl - cen
        # s = fma(d2, d2, fma(d1, d1, d0*d0)); emulate each fma with exact two-product arithmetic in longdouble
        s = _fma(d[..., 2], d[..., 2], _fma(d[..., 1], d[..., 1], d[..., 0] * d[..., 0]))
        hit |= (np.sqrt(s) <= rb).any(1)
    return hit


def _fma(a, b, c):
    """fma(a, b, c) for float64 arrays via an error-free product (Dekker) and a double-double sum: exact to the
    last bit whenever the x87 long double (64-bit mantissa) sum a*b + c does not need more than its width, which
    holds for the magnitudes of this model; oracle/arm_oracle.c (C fma) is the arbiter."""
    la, lb, lc = a.astype(np.longdouble), b.astype(np.longdouble), c.astype(np.longdouble)
    return (la * lb + lc).astype(np.float64)
