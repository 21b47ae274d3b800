// Issue-rate micro-benchmarks for the roofline denominators bench.py reports against
// (the FP32 and FP64 FMA peaks "must be measured on the box").
// Not part of the planner path.
#include "mpb_common.cuh"

template <typename T>
__device__ __forceinline__ T fma_t(T a, T b, T c);
template <>
__device__ __forceinline__ float fma_t<float>(float a, float b, float c) { return __fmaf_rn(a, b, c); }
template <>
__device__ __forceinline__ double fma_t<double>(double a, double b, double c) { return __fma_rn(a, b, c); }

// 8 independent chains per thread; every chain is a dependent FMA sequence.
template <typename T>
__global__ void __launch_bounds__(1024) peak_fma_kernel(int iters, T *sink)
{
    const T b = (T)1.0000001 + (T)threadIdx.x * (T)1e-9;
    const T c = (T)1e-7;
    T a0 = (T)threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
#pragma unroll 8
    for (int i = 0; i < iters; ++i) {
        a0 = fma_t<T>(a0, b, c);
        a1 = fma_t<T>(a1, b, c);
        a2 = fma_t<T>(a2, b, c);
        a3 = fma_t<T>(a3, b, c);
        a4 = fma_t<T>(a4, b, c);
        a5 = fma_t<T>(a5, b, c);
        a6 = fma_t<T>(a6, b, c);
        a7 = fma_t<T>(a7, b, c);
    }
    sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

// The same with packed pairs: 8 independent chains of FFMA2, 2 fused multiply-adds per instruction and lane.
__global__ void __launch_bounds__(1024) peak_fma2_kernel(int iters, float *sink)
{
    const float bf = 1.0000001f + (float)threadIdx.x * 1e-9f;
    const f32x2_t b = pack_f32x2(bf, bf), c = pack_f32x2(1e-7f, 1e-7f);
    f32x2_t a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = pack_f32x2((float)threadIdx.x + j, (float)threadIdx.x - j);
#pragma unroll 8
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = fma_f32x2(a[j], b, c);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float lo, hi;
        unpack_f32x2(a[j], lo, hi);
        s += lo + hi;
    }
    sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename T>
static int run_peak(const char *who, int blocks, int threads, int iters, T *sink, void *stream)
{
    MPB_REQUIRE(blocks > 0 && threads > 0 && threads <= 1024 && iters > 0 && sink, "%s: bad argument", who);
    peak_fma_kernel<T><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, sink);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

MPB_API int mpb_peak_fma_f32(int blocks, int threads, int iters, float *sink, void *stream)
{
    return run_peak<float>("mpb_peak_fma_f32", blocks, threads, iters, sink, stream);
}

MPB_API int mpb_peak_fma_f64(int blocks, int threads, int iters, double *sink, void *stream)
{
    return run_peak<double>("mpb_peak_fma_f64", blocks, threads, iters, sink, stream);
}

MPB_API int mpb_peak_fma_f32x2(int blocks, int threads, int iters, float *sink, void *stream)
{
    MPB_REQUIRE(blocks > 0 && threads > 0 && threads <= 1024 && iters > 0 && sink, "mpb_peak_fma_f32x2: bad argument");
    peak_fma2_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, sink);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}
