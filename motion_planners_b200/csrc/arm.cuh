// Device side of the sphere-approximated serial arm (BASELINE.json configs[3]); see arm.cu for the entry points
// and motion_planners_b200/arm.py / oracle/arm_oracle.c for the model.  Every fp64 operation is issued through
// the __d*_rn intrinsics in exactly the order of oracle/arm_oracle.c, so the three statements agree bit for bit.
#pragma once

#include "mpb_common.cuh"

constexpr int kArmMaxJoints = MPB_ARM_MAX_JOINTS;
constexpr int kArmMaxSpheres = MPB_ARM_MAX_SPHERES;
#ifndef MPB_ARM_GRID
#define MPB_ARM_GRID 32
#endif
#ifndef MPB_ARM_GRID_GLOBAL
#define MPB_ARM_GRID_GLOBAL 1      // 1: the grid's lists stay in global memory (L1 / L2) instead of shared memory
// measured on the 1M-sample build (tools/ab_arm.py), edge phase: 16^3 in shared memory 12.9 ms, 24^3 in shared memory
// 11.7, 32^3 in global memory 10.7, 48^3 in global memory 10.8: shorter lists (0.9 boxes per cell against 1.55) matter
// more than where the lists live
#endif
constexpr int kArmGrid = MPB_ARM_GRID;             // cells per workspace axis of the box grid

// kernel-parameter view of an arm world (about 1.6 KB, read through the constant cache)
struct ArmDev {
    int n_joints, n_sph, n_box, n_items;
    int axis[kArmMaxJoints];
    double theta_lo[kArmMaxJoints], theta_span[kArmMaxJoints], a[kArmMaxJoints], d[kArmMaxJoints];
    double base[3];
    double env_lo[kArmMaxJoints], env_hi[kArmMaxJoints];
    int sph_first[kArmMaxJoints + 1];              // spheres sorted by the joint that carries them
    int sph_orig[kArmMaxSpheres];                  // position of sorted sphere j in the caller's order
    double sph_u[kArmMaxSpheres], sph_w[kArmMaxSpheres];
    double sph_rb[kArmMaxSpheres];                 // radius + buffer
    double sph_lo2[kArmMaxSpheres], sph_hi2[kArmMaxSpheres];   // (radius + buffer)^2 scaled down / up by 1e-12
    double grid_lo[3], grid_inv[3];
    const double *boxes;                           // dev [n_box][3][2] {lo, hi}
    const int *cell_start;                         // dev [kArmGrid^3 + 1]
    const unsigned short *cell_items;              // dev [n_items] box indices, cell by cell
};

struct ArmSmem {
    const double2 *boxes;                          // [n_box * 3] {lo, hi}
    const int *cell_start;
    const unsigned short *cell_items;
};

inline size_t arm_smem_bytes(const ArmDev &A)
{
    size_t b = (size_t)A.n_box * 3 * sizeof(double2);
    if (MPB_ARM_GRID_GLOBAL) return b;
    b += ((size_t)(kArmGrid * kArmGrid * kArmGrid + 1) * sizeof(int) + 15) & ~(size_t)15;
    b += ((size_t)A.n_items * sizeof(unsigned short) + 15) & ~(size_t)15;
    return b;
}

__device__ __forceinline__ ArmSmem arm_stage(const ArmDev &A, unsigned char *sm)
{
    double2 *boxes = reinterpret_cast<double2 *>(sm);
    int *cs = reinterpret_cast<int *>(sm + (size_t)A.n_box * 3 * sizeof(double2));
    const size_t cs_bytes = ((size_t)(kArmGrid * kArmGrid * kArmGrid + 1) * sizeof(int) + 15) & ~(size_t)15;
    unsigned short *items = reinterpret_cast<unsigned short *>(reinterpret_cast<unsigned char *>(cs) + cs_bytes);
    for (int i = threadIdx.x; i < A.n_box * 3; i += blockDim.x) boxes[i] = make_double2(A.boxes[2 * i], A.boxes[2 * i + 1]);
    if (MPB_ARM_GRID_GLOBAL) {
        __syncthreads();
        ArmSmem g = {boxes, A.cell_start, A.cell_items};
        return g;
    }
    for (int i = threadIdx.x; i <= kArmGrid * kArmGrid * kArmGrid; i += blockDim.x) cs[i] = A.cell_start[i];
    for (int i = threadIdx.x; i < A.n_items; i += blockDim.x) items[i] = A.cell_items[i];
    __syncthreads();
    ArmSmem s = {boxes, cs, items};
    return s;
}

// oracle/arm_oracle.c arm_sincos, operation for operation
__device__ __forceinline__ void arm_sincos(double th, double &s, double &c)
{
    const double kf = rint(d_mul(th, 0.63661977236758138));
    double r = d_sub(th, d_mul(kf, 1.57079632673412561417e+00));
    r = d_sub(r, d_mul(kf, 6.07710050650619224932e-11));
    const double z = d_mul(r, r);
    double t = 1.58969099521155010221e-10;
    t = d_add(d_mul(t, z), -2.50507602534068634195e-08);
    t = d_add(d_mul(t, z), 2.75573137070700676789e-06);
    t = d_add(d_mul(t, z), -1.98412698298579493134e-04);
    t = d_add(d_mul(t, z), 8.33333333332248946124e-03);
    t = d_add(d_mul(t, z), -1.66666666666666324348e-01);
    const double sr = d_add(r, d_mul(d_mul(r, z), t));
    double u = -1.13596475577881948265e-11;
    u = d_add(d_mul(u, z), 2.08757232129817482790e-09);
    u = d_add(d_mul(u, z), -2.75573143513906633035e-07);
    u = d_add(d_mul(u, z), 2.48015872894767294178e-05);
    u = d_add(d_mul(u, z), -1.38888888888741095749e-03);
    u = d_add(d_mul(u, z), 4.16666666666666019037e-02);
    const double cr = d_add(d_sub(1.0, d_mul(0.5, z)), d_mul(d_mul(z, z), u));
    const int n = (int)((long long)kf & 3);
    s = (n == 0) ? sr : (n == 1) ? cr : (n == 2) ? -sr : -cr;
    c = (n == 0) ? cr : (n == 1) ? -sr : (n == 2) ? -cr : sr;
}

// ||clamp(c, lo, hi) - c|| <= r + buffer for some box near the centre's grid cell.  The grid lists, cell by cell,
// every box that a sphere centred anywhere in the cell could touch (built on the host with the largest inflated
// radius plus a margin), so boxes outside the list cannot change the answer.  sqrt(s) <= rb is decided on s
// wherever s is outside a 1e-12 band around rb^2 (sqrt is monotone and correctly rounded) and by the square root
// itself inside it: the same boolean as the oracle's, without a square root per pair.
__device__ __forceinline__ bool arm_sphere_hits(const ArmDev &A, const ArmSmem &S, int j, double cx, double cy, double cz)
{
    int ix = (int)floor((cx - A.grid_lo[0]) * A.grid_inv[0]);
    int iy = (int)floor((cy - A.grid_lo[1]) * A.grid_inv[1]);
    int iz = (int)floor((cz - A.grid_lo[2]) * A.grid_inv[2]);
    // a centre outside the grid cannot touch anything (the grid spans every inflated box)
    if (ix < 0 || iy < 0 || iz < 0 || ix >= kArmGrid || iy >= kArmGrid || iz >= kArmGrid) return false;
    const int cell = (ix * kArmGrid + iy) * kArmGrid + iz;
    const int e = S.cell_start[cell + 1];
    for (int it = S.cell_start[cell]; it < e; ++it) {
        const double2 *bb = S.boxes + 3 * (int)S.cell_items[it];
        double2 b = bb[0];
        double cl = cx < b.x ? b.x : cx;
        cl = cl > b.y ? b.y : cl;
        double dk = d_sub(cl, cx);
        double s = __fma_rn(dk, dk, 0.);
        b = bb[1];
        cl = cy < b.x ? b.x : cy;
        cl = cl > b.y ? b.y : cl;
        dk = d_sub(cl, cy);
        s = __fma_rn(dk, dk, s);
        b = bb[2];
        cl = cz < b.x ? b.x : cz;
        cl = cl > b.y ? b.y : cl;
        dk = d_sub(cl, cz);
        s = __fma_rn(dk, dk, s);
        if (s < A.sph_lo2[j]) return true;
        if (s <= A.sph_hi2[j] && __dsqrt_rn(s) <= A.sph_rb[j]) return true;
    }
    return false;
}

// collision_fn(q) of the arm world; centres (may be NULL) receives the sphere centres in the caller's order
template <int NJ>
__device__ __forceinline__ bool arm_collides(const ArmDev &A, const ArmSmem &S, const double (&q)[NJ], double *centres)
{
    if (!centres) {
        bool in = true;
#pragma unroll
        for (int i = 0; i < NJ; ++i) in = in && (A.env_lo[i] <= q[i]) && (q[i] <= A.env_hi[i]);
        if (!in) return true;
    }
    double X0 = 1., X1 = 0., X2 = 0., Y0 = 0., Y1 = 1., Y2 = 0., Z0 = 0., Z1 = 0., Z2 = 1.;
    double p0 = A.base[0], p1 = A.base[1], p2 = A.base[2];
    bool hit = false;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
        const double th = d_add(d_mul(q[i], A.theta_span[i]), A.theta_lo[i]);
        double s, c;
        arm_sincos(th, s, c);
        if (A.axis[i] == 2) {
            const double x0 = X0, x1 = X1, x2 = X2, y0 = Y0, y1 = Y1, y2 = Y2;
            X0 = d_add(d_mul(c, x0), d_mul(s, y0));
            X1 = d_add(d_mul(c, x1), d_mul(s, y1));
            X2 = d_add(d_mul(c, x2), d_mul(s, y2));
            Y0 = d_sub(d_mul(c, y0), d_mul(s, x0));
            Y1 = d_sub(d_mul(c, y1), d_mul(s, x1));
            Y2 = d_sub(d_mul(c, y2), d_mul(s, x2));
        } else {
            const double x0 = X0, x1 = X1, x2 = X2, z0 = Z0, z1 = Z1, z2 = Z2;
            X0 = d_sub(d_mul(c, x0), d_mul(s, z0));
            X1 = d_sub(d_mul(c, x1), d_mul(s, z1));
            X2 = d_sub(d_mul(c, x2), d_mul(s, z2));
            Z0 = d_add(d_mul(s, x0), d_mul(c, z0));
            Z1 = d_add(d_mul(s, x1), d_mul(c, z1));
            Z2 = d_add(d_mul(s, x2), d_mul(c, z2));
        }
        for (int j = A.sph_first[i]; j < A.sph_first[i + 1]; ++j) {
            const double u = A.sph_u[j], w = A.sph_w[j];
            const double cx = d_add(d_add(p0, d_mul(u, X0)), d_mul(w, Z0));
            const double cy = d_add(d_add(p1, d_mul(u, X1)), d_mul(w, Z1));
            const double cz = d_add(d_add(p2, d_mul(u, X2)), d_mul(w, Z2));
            if (centres) {
                centres[3 * A.sph_orig[j] + 0] = cx;
                centres[3 * A.sph_orig[j] + 1] = cy;
                centres[3 * A.sph_orig[j] + 2] = cz;
            } else if (arm_sphere_hits(A, S, j, cx, cy, cz)) {
                hit = true;
                break;
            }
        }
        if (hit) break;
        p0 = d_add(d_add(p0, d_mul(A.a[i], X0)), d_mul(A.d[i], Z0));
        p1 = d_add(d_add(p1, d_mul(A.a[i], X1)), d_mul(A.d[i], Z1));
        p2 = d_add(d_add(p2, d_mul(A.a[i], X2)), d_mul(A.d[i], Z2));
    }
    return hit;
}


// ---- the model's per-joint and per-sphere numbers in shared memory (lanes of a warp sit at different joints) ----
struct ArmTables {
    double theta_lo[kArmMaxJoints], theta_span[kArmMaxJoints], a[kArmMaxJoints], d[kArmMaxJoints];
    double sgn[kArmMaxJoints];                     // +1: joint turns about z (mixes x, y), -1: about y (mixes x, z)
    double sph_u[kArmMaxSpheres], sph_w[kArmMaxSpheres], sph_rb[kArmMaxSpheres], sph_lo2[kArmMaxSpheres],
        sph_hi2[kArmMaxSpheres];
    int sph_first[kArmMaxJoints + 1];
};

__device__ __forceinline__ void arm_fill_tables(const ArmDev &A, ArmTables *T)
{
    for (int i = threadIdx.x; i < kArmMaxJoints; i += blockDim.x) {
        T->theta_lo[i] = A.theta_lo[i];
        T->theta_span[i] = A.theta_span[i];
        T->a[i] = A.a[i];
        T->d[i] = A.d[i];
        T->sgn[i] = A.axis[i] == 2 ? 1.0 : -1.0;
    }
    for (int i = threadIdx.x; i <= kArmMaxJoints; i += blockDim.x) T->sph_first[i] = A.sph_first[i];
    for (int j = threadIdx.x; j < kArmMaxSpheres; j += blockDim.x) {
        T->sph_u[j] = A.sph_u[j];
        T->sph_w[j] = A.sph_w[j];
        T->sph_rb[j] = A.sph_rb[j];
        T->sph_lo2[j] = A.sph_lo2[j];
        T->sph_hi2[j] = A.sph_hi2[j];
    }
}

// the sphere test of arm_sphere_hits with the sphere's numbers from the shared tables
__device__ __forceinline__ bool arm_sphere_hits_t(const ArmDev &A, const ArmTables &T, const ArmSmem &S, int j, double cx,
                                                  double cy, double cz)
{
    const int ix = (int)floor((cx - A.grid_lo[0]) * A.grid_inv[0]);
    const int iy = (int)floor((cy - A.grid_lo[1]) * A.grid_inv[1]);
    const int iz = (int)floor((cz - A.grid_lo[2]) * A.grid_inv[2]);
    if (ix < 0 || iy < 0 || iz < 0 || ix >= kArmGrid || iy >= kArmGrid || iz >= kArmGrid) return false;
    const int cell = (ix * kArmGrid + iy) * kArmGrid + iz;
    const int e = S.cell_start[cell + 1];
    const double lo2 = T.sph_lo2[j], hi2 = T.sph_hi2[j];
    for (int it = S.cell_start[cell]; it < e; ++it) {
        const double2 *bb = S.boxes + 3 * (int)S.cell_items[it];
        double2 b = bb[0];
        double cl = cx < b.x ? b.x : cx;
        cl = cl > b.y ? b.y : cl;
        double dk = d_sub(cl, cx);
        double s = __fma_rn(dk, dk, 0.);
        b = bb[1];
        cl = cy < b.x ? b.x : cy;
        cl = cl > b.y ? b.y : cl;
        dk = d_sub(cl, cy);
        s = __fma_rn(dk, dk, s);
        b = bb[2];
        cl = cz < b.x ? b.x : cz;
        cl = cl > b.y ? b.y : cl;
        dk = d_sub(cl, cz);
        s = __fma_rn(dk, dk, s);
        if (s < lo2) return true;
        if (s <= hi2 && __dsqrt_rn(s) <= T.sph_rb[j]) return true;
    }
    return false;
}

// One joint of the chain for a lane whose frame is (X, P = Y or Z by the joint's axis, p): the turn.  Written once
// for both axes: with s' = +s (about z) or -s (about y),
//   x' = fl(c x) + fl(s' p),  p' = fl(c p) - fl(s' x)
// are bit for bit the oracle's c x + s y, c y - s x (about z) and c x - s z, s x + c z (about y): negation is exact
// and addition commutes.
struct ArmFrame {
    double X0, X1, X2, Y0, Y1, Y2, Z0, Z1, Z2, p0, p1, p2;
};

__device__ __forceinline__ void arm_turn(const ArmTables &T, int ji, double qi, ArmFrame &F)
{
    const double th = d_add(d_mul(qi, T.theta_span[ji]), T.theta_lo[ji]);
    double s, c;
    arm_sincos(th, s, c);
    const bool about_z = T.sgn[ji] > 0.;
    const double sp = about_z ? s : -s;
    const double P0 = about_z ? F.Y0 : F.Z0, P1 = about_z ? F.Y1 : F.Z1, P2 = about_z ? F.Y2 : F.Z2;
    const double x0 = F.X0, x1 = F.X1, x2 = F.X2;
    F.X0 = d_add(d_mul(c, x0), d_mul(sp, P0));
    F.X1 = d_add(d_mul(c, x1), d_mul(sp, P1));
    F.X2 = d_add(d_mul(c, x2), d_mul(sp, P2));
    const double n0 = d_sub(d_mul(c, P0), d_mul(sp, x0));
    const double n1 = d_sub(d_mul(c, P1), d_mul(sp, x1));
    const double n2 = d_sub(d_mul(c, P2), d_mul(sp, x2));
    F.Y0 = about_z ? n0 : F.Y0;
    F.Y1 = about_z ? n1 : F.Y1;
    F.Y2 = about_z ? n2 : F.Y2;
    F.Z0 = about_z ? F.Z0 : n0;
    F.Z1 = about_z ? F.Z1 : n1;
    F.Z2 = about_z ? F.Z2 : n2;
}

// to the end of link ji
__device__ __forceinline__ void arm_link(const ArmTables &T, int ji, ArmFrame &F)
{
    const double al = T.a[ji], dl = T.d[ji];
    F.p0 = d_add(d_add(F.p0, d_mul(al, F.X0)), d_mul(dl, F.Z0));
    F.p1 = d_add(d_add(F.p1, d_mul(al, F.X1)), d_mul(dl, F.Z1));
    F.p2 = d_add(d_add(F.p2, d_mul(al, F.X2)), d_mul(dl, F.Z2));
}

// sphere j of the current frame against the boxes of its grid cell
__device__ __forceinline__ bool arm_sphere(const ArmDev &A, const ArmTables &T, const ArmSmem &S, int j, const ArmFrame &F)
{
    const double u = T.sph_u[j], w = T.sph_w[j];
    const double cx = d_add(d_add(F.p0, d_mul(u, F.X0)), d_mul(w, F.Z0));
    const double cy = d_add(d_add(F.p1, d_mul(u, F.X1)), d_mul(w, F.Z1));
    const double cz = d_add(d_add(F.p2, d_mul(u, F.X2)), d_mul(w, F.Z2));
    return arm_sphere_hits_t(A, T, S, j, cx, cy, cz);
}
