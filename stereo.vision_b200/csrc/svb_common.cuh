// Shared device helpers: record packing and the path's exact float64 arithmetic.
//
// Everything that decides a bit-exact observable (site -> XYZ -> reprojected pixel, hue bin, the
// near-plane test) is written with the *_rn intrinsics so that nvcc never contracts a multiply and
// an add into an FMA: the reference evaluates these expressions operation by operation in float64
// (functions.py:191-193, 206-207; colorsys.rgb_to_hsv) and 3 % of lattice sites reproject onto the
// neighbouring pixel purely because of those roundings.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/svb200.h"

#define SVB_MAX_DIM 2048

struct CamD {            // camera constants as the kernels want them
    double f, cw, ch, fB;
};

static inline CamD make_camd(const svb_camera *c) {
    CamD k;
    k.f = c->focal_px;
    k.cw = c->centre_w;
    k.ch = c->centre_h;
    k.fB = c->focal_px * c->baseline_m;      // (f * B), product first, functions.py:191
    return k;
}

#define SVB_LAUNCH_CHECK()                                   \
    do {                                                     \
        cudaError_t e__ = cudaGetLastError();                \
        if (e__ != cudaSuccess) return (int)e__;             \
    } while (0)

// ------------------------------------------------------------------ records
__host__ __device__ __forceinline__ uint32_t pack_xyd(int x, int y, int d) {
    return (uint32_t)x | ((uint32_t)y << 11) | ((uint32_t)d << 22);
}
__host__ __device__ __forceinline__ int rec_x(uint32_t lo) { return (int)(lo & 0x7FFu); }
__host__ __device__ __forceinline__ int rec_y(uint32_t lo) { return (int)((lo >> 11) & 0x7FFu); }
__host__ __device__ __forceinline__ int rec_d(uint32_t lo) { return (int)((lo >> 22) & 0xFFu); }
__host__ __device__ __forceinline__ int rec_hue(uint32_t lo, uint32_t hi) {
    return (int)((lo >> 30) | ((hi >> 24) << 2));
}
__host__ __device__ __forceinline__ uint64_t pack_cloud(int x, int y, int d, int r, int g, int b, int hue) {
    uint32_t lo = pack_xyd(x, y, d) | ((uint32_t)(hue & 3) << 30);
    uint32_t hi = (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16) | ((uint32_t)(hue >> 2) << 24);
    return (uint64_t)lo | ((uint64_t)hi << 32);
}

// ------------------------------------------------------------------ exact float64 path arithmetic
// Z = (f*B)/d                                   functions.py:191
__device__ __forceinline__ double exact_Z(const CamD &c, int d) { return __ddiv_rn(c.fB, (double)d); }
// X = ((x - cw) * Z) / f                        functions.py:192   (Y likewise with ch, :193)
__device__ __forceinline__ double exact_XY(double centre, double f, int coord, double Z) {
    return __ddiv_rn(__dmul_rn(__dsub_rn((double)coord, centre), Z), f);
}
// x' = ((X * f) / Z) + cw                       functions.py:206   (y' likewise, :207)
__device__ __forceinline__ double exact_reproject(double centre, double f, double XY, double Z) {
    return __dadd_rn(__ddiv_rn(__dmul_rn(XY, f), Z), centre);
}
// P . abc with the products summed left to right, no FMA (functions.py:275,310 evaluate it with np.dot)
__device__ __forceinline__ double exact_dot(double X, double Y, double Z, double a, double b, double c) {
    return __dadd_rn(__dadd_rn(__dmul_rn(X, a), __dmul_rn(Y, b)), __dmul_rn(Z, c));
}

// hue bin = rint(1000 * colorsys.rgb_to_hsv(r,g,b)[0]) with np.uint8 channels
// (functions.py:73-78).  Grey -> 0.  This is the reference's float64 operation
// sequence, one rounding per operation; it is the definition, and the slow path of hue_bin().
static __device__ __noinline__ int exact_hue_bin(int r, int g, int b) {
    int maxc = max(r, max(g, b));
    int minc = min(r, min(g, b));
    if (maxc == minc) return 0;
    double range = (double)(maxc - minc);
    double h;
    if (r == maxc) {
        double bc = __ddiv_rn((double)(maxc - b), range);
        double gc = __ddiv_rn((double)(maxc - g), range);
        h = __dsub_rn(bc, gc);
    } else if (g == maxc) {
        double rc = __ddiv_rn((double)(maxc - r), range);
        double bc = __ddiv_rn((double)(maxc - b), range);
        h = __dsub_rn(__dadd_rn(2.0, rc), bc);
    } else {
        double gc = __ddiv_rn((double)(maxc - g), range);
        double rc = __ddiv_rn((double)(maxc - r), range);
        h = __dsub_rn(__dadd_rn(4.0, gc), rc);
    }
    h = __ddiv_rn(h, 6.0);
    if (h < 0.0) h = __dadd_rn(h, 1.0);          // np.remainder(h, 1.0) for h in (-1, 0)
    return __double2int_rn(__dmul_rn(h, 1000.0)); // rint, ties to even
}

// Fast path.  In real arithmetic 1000*h = 1000*hn / (6*range) with the integer
//   hn = (g-b) | 2*range + (b-r) | 4*range + (r-g)   (+ 6*range when negative),
// a rational with denominator <= 1530.  The float64 sequence above carries a relative error of a
// few 1e-16, so its rint() can differ from the exactly rounded rational only when the rational is an
// exact tie k + 1/2 (any other value is at least 1/3060 away from a tie).  Integer quotient and
// remainder give the bin; the rare exact ties (0.06 % of colours) take the float64 path, which
// decides them the way the reference's roundings do.  tests: exhaustive over all 2^24 colours.
__device__ __forceinline__ int hue_bin(int r, int g, int b) {
    int maxc = max(r, max(g, b));
    int minc = min(r, min(g, b));
    int range = maxc - minc;
    if (range == 0) return 0;
    int hn = (r == maxc) ? (g - b) : (g == maxc) ? 2 * range + (b - r) : 4 * range + (r - g);
    uint32_t den = 6u * (uint32_t)range;
    if (hn < 0) hn += (int)den;
    uint32_t num = 1000u * (uint32_t)hn;                       // < 6000*255
    // num / den (num < 2^21, den <= 1530): float estimate, off by at most one, fixed with the remainder
    uint32_t q = __float2uint_rz(__fmul_rn((float)num, __frcp_rn((float)den)));
    int rem = (int)num - (int)(q * den);
    if (rem < 0) { --q; rem += (int)den; }
    else if (rem >= (int)den) { ++q; rem -= (int)den; }
    uint32_t twice = 2u * (uint32_t)rem;
    if (twice == den) return exact_hue_bin(r, g, b);           // exact tie: let the float64 roundings decide
    return (int)(q + (twice > den ? 1u : 0u));
}

// ------------------------------------------------------------------ warp helpers
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}
