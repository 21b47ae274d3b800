/* CPU statement of the sphere-approximated serial arm of BASELINE.json configs[3].  TEST INFRASTRUCTURE ONLY
 * (see oracle/__init__.py: only tests/, smoke() and bench.py's CPU legs may use it).
 *
 * The reference (caelan/motion-planners) has NO robot model: its collision_fn is a caller-supplied
 * callable q -> bool (README.md:52-59; the demo's is tkinter/samplers.py:88-98).  This file DEFINES the
 * callable configs[3] names -- a 7-DoF serial chain approximated by spheres, against axis-aligned boxes in
 * the 3-D workspace -- in the reference's own terms:
 *   collision(q) = not in_env(q)                                   (samplers.py:92, closed bounds, no buffer)
 *                  or any sphere k, box b:  ||clamp(c_k, lo_b, hi_b) - c_k|| <= r_k + buffer
 *                                                                   (viewer.py:80-82 contains_circle:
 *                                                                   np.linalg.norm(...) <= radius + buffer)
 * with the waypoints of an edge from the reference's extend_fn recurrence (samplers.py:113-127) in joint
 * space.  There is nothing in the reference to pin it against ("parity unpinned" for this collision model):
 * this C text, oracle/np_oracle.py's numpy text and the CUDA kernel (csrc/arm.cu) are three statements of the
 * same arithmetic and are tested bit for bit against each other.
 *
 * Every operation is one IEEE binary64 operation (compile with -ffp-contract=off); the only fused operations
 * are the explicit fma() calls of the norm (numpy's dot on this host, as in mp_oracle.c).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define ORC_EXPORT __attribute__((visibility("default")))

#define ARM_MAX_JOINTS 8
#define ARM_MAX_SPHERES 32

typedef struct {
    int n_joints;
    int axis[ARM_MAX_JOINTS];            /* 1 = y, 2 = z: the local axis joint i turns about */
    double theta_lo[ARM_MAX_JOINTS];     /* angle = fl(fl(q * span) + lo) */
    double theta_span[ARM_MAX_JOINTS];
    double a[ARM_MAX_JOINTS];            /* link i: a along the turned frame's x, d along its z */
    double d[ARM_MAX_JOINTS];
    double base[3];
    int n_sph;
    int sph_link[ARM_MAX_SPHERES];       /* the joint whose turned frame carries the sphere */
    double sph_u[ARM_MAX_SPHERES];       /* centre = frame origin + u * x-axis + w * z-axis */
    double sph_w[ARM_MAX_SPHERES];
    double sph_r[ARM_MAX_SPHERES];
} orc_arm_model;

/* sin and cos from one argument reduction and two fixed polynomials: + - * rint only, evaluated in exactly
 * this order everywhere (Cody-Waite reduction by pi/2 in two parts, Horner forms with fdlibm's coefficients;
 * |theta| stays far below 2^20 here, where the two-part reduction is exact to the last bit or two -- and
 * "the same bits everywhere" is what matters, not the last ulp). */
static void arm_sincos(double th, double *s_out, double *c_out)
{
    const double kf = rint(th * 0.63661977236758138);
    double r = th - kf * 1.57079632673412561417e+00;
    r = r - kf * 6.07710050650619224932e-11;
    const double z = r * r;
    double t = 1.58969099521155010221e-10;
    t = t * z + -2.50507602534068634195e-08;
    t = t * z + 2.75573137070700676789e-06;
    t = t * z + -1.98412698298579493134e-04;
    t = t * z + 8.33333333332248946124e-03;
    t = t * z + -1.66666666666666324348e-01;
    const double sr = r + (r * z) * t;
    double u = -1.13596475577881948265e-11;
    u = u * z + 2.08757232129817482790e-09;
    u = u * z + -2.75573143513906633035e-07;
    u = u * z + 2.48015872894767294178e-05;
    u = u * z + -1.38888888888741095749e-03;
    u = u * z + 4.16666666666666019037e-02;
    const double cr = (1.0 - 0.5 * z) + (z * z) * u;
    const int n = (int)((long long)kf & 3);
    double s, c;
    switch (n) {
        case 0: s = sr; c = cr; break;
        case 1: s = cr; c = -sr; break;
        case 2: s = -sr; c = -cr; break;
        default: s = -cr; c = sr; break;
    }
    *s_out = s;
    *c_out = c;
}

/* sphere centres of configuration q: out [n_sph][3] */
static void arm_centres(const orc_arm_model *m, const double *q, double *out)
{
    double X[3] = {1., 0., 0.}, Y[3] = {0., 1., 0.}, Z[3] = {0., 0., 1.};   /* frame axes (columns) */
    double p[3] = {m->base[0], m->base[1], m->base[2]};
    for (int i = 0; i < m->n_joints; ++i) {
        const double th = q[i] * m->theta_span[i] + m->theta_lo[i];
        double s, c;
        arm_sincos(th, &s, &c);
        if (m->axis[i] == 2) {              /* about z: x' = c x + s y, y' = c y - s x */
            for (int k = 0; k < 3; ++k) {
                const double x = X[k], y = Y[k];
                X[k] = c * x + s * y;
                Y[k] = c * y - s * x;
            }
        } else {                            /* about y: x' = c x - s z, z' = s x + c z */
            for (int k = 0; k < 3; ++k) {
                const double x = X[k], z = Z[k];
                X[k] = c * x - s * z;
                Z[k] = s * x + c * z;
            }
        }
        for (int j = 0; j < m->n_sph; ++j) {
            if (m->sph_link[j] != i) continue;
            for (int k = 0; k < 3; ++k) out[3 * j + k] = (p[k] + m->sph_u[j] * X[k]) + m->sph_w[j] * Z[k];
        }
        for (int k = 0; k < 3; ++k) p[k] = (p[k] + m->a[i] * X[k]) + m->d[i] * Z[k];
    }
}

static int sphere_hits_box(const double *c, double rb, const double *lo, const double *hi)
{
    double s = 0.;
    for (int k = 0; k < 3; ++k) {
        double cl = c[k] < lo[k] ? lo[k] : c[k];      /* clamp(c, lo, hi) */
        cl = cl > hi[k] ? hi[k] : cl;
        const double dk = cl - c[k];
        s = fma(dk, dk, s);
    }
    return sqrt(s) <= rb;
}

static int arm_collides(const orc_arm_model *m, int nbox, const double *box_lo, const double *box_hi,
                        const double *env_lo, const double *env_hi, double buffer, const double *q)
{
    for (int i = 0; i < m->n_joints; ++i)
        if (!(env_lo[i] <= q[i] && q[i] <= env_hi[i])) return 1;
    double cen[3 * ARM_MAX_SPHERES];
    arm_centres(m, q, cen);
    for (int j = 0; j < m->n_sph; ++j) {
        const double rb = m->sph_r[j] + buffer;
        for (int b = 0; b < nbox; ++b)
            if (sphere_hits_box(cen + 3 * j, rb, box_lo + 3 * b, box_hi + 3 * b)) return 1;
    }
    return 0;
}

ORC_EXPORT int orc_arm_centres(const orc_arm_model *m, const double *q, int64_t n, double *out)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) arm_centres(m, q + i * m->n_joints, out + i * 3 * m->n_sph);
    return 0;
}

ORC_EXPORT int orc_arm_points_check(const orc_arm_model *m, int nbox, const double *box_lo, const double *box_hi,
                                    const double *env_lo, const double *env_hi, double buffer, const double *q,
                                    int64_t n, uint8_t *out)
{
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < n; ++i)
        out[i] = (uint8_t)arm_collides(m, nbox, box_lo, box_hi, env_lo, env_hi, buffer, q + i * m->n_joints);
    return 0;
}

/* edges as index pairs into a sample table; waypoints by samplers.py:113-127; a negative index = no edge */
ORC_EXPORT int orc_arm_edges_check(const orc_arm_model *m, int nbox, const double *box_lo, const double *box_hi,
                                   const double *env_lo, const double *env_hi, double buffer,
                                   const double *samples, const int64_t *src, const int64_t *dst, int64_t E,
                                   double step, int skip_last, int32_t *out_n, int32_t *out_first,
                                   int32_t *out_last)
{
    const int D = m->n_joints;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t e = 0; e < E; ++e) {
        if (src[e] < 0 || dst[e] < 0) {
            out_n[e] = 0;
            out_first[e] = out_last[e] = -1;
            continue;
        }
        const double *a = samples + src[e] * D, *b = samples + dst[e] * D;
        double mx = 0.;
        for (int d = 0; d < D; ++d) {
            const double v = fabs((b[d] - a[d]) / step);
            if (v > mx) mx = v;
        }
        const int64_t n = (int64_t)mx + 1;
        const int64_t lim = n - (skip_last ? 1 : 0);
        double q[ARM_MAX_JOINTS];
        memcpy(q, a, sizeof(double) * (size_t)D);
        int32_t first = -1, last = -1;
        for (int64_t i = 0; i < n; ++i) {
            const double c = 1. / (double)(n - i);
            for (int d = 0; d < D; ++d) {
                const double t = b[d] - q[d];
                const double mm = c * t;
                q[d] = mm + q[d];
            }
            if (i < lim && arm_collides(m, nbox, box_lo, box_hi, env_lo, env_hi, buffer, q)) {
                if (first < 0) first = (int32_t)i;
                last = (int32_t)i;
            }
        }
        out_n[e] = (int32_t)n;
        out_first[e] = first;
        out_last[e] = last;
    }
    return 0;
}
