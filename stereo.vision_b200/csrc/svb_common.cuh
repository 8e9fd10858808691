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
#include "../../include/svb200_debug.h"

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
__host__ __device__ __forceinline__ uint64_t pack_cloud(int x, int y, int d, int r, int g, int b) {
    uint32_t lo = pack_xyd(x, y, d);
    uint32_t hi = (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16);
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
// exact tie k + 1/2 (any other value is at least 1/3060 away from a tie).  Quotient and remainder come
// from one multiply-high with a per-range reciprocal (exact: num * (m*den - 2^32) < 2^32 for every
// num < 1000*1530).  Exact ties (0.375 % of all colours) land on 200 of the bins, and which way the
// reference's float64 roundings push a tie depends on the bin alone -- correctly rounded division
// gives (maxc-ch)/range the same double for every colour with the same ratio -- so one bit per bin,
// taken from exact_hue_bin() / the reference over all 2^24 colours, decides them.
// tests: exhaustive over all 2^24 colours against the oracle (test_hue_bins_exhaustive).
struct HueMagic { uint32_t m[256]; };
constexpr HueMagic make_hue_magic() {
    HueMagic t{};
    for (int r = 1; r < 256; ++r) {
        unsigned long long den = 6ull * (unsigned long long)r;
        t.m[r] = (uint32_t)(((1ull << 32) + den - 1ull) / den);       // ceil(2^32 / (6*range))
    }
    return t;
}
__device__ const HueMagic d_hue_magic = make_hue_magic();
// bit q set: a tie at q + 1/2 goes to q + 1 under the reference's roundings, clear: stays at q
__constant__ uint32_t c_hue_tie_up[32] = {
    0x08400084u, 0x00008020u, 0x00802108u, 0x80200800u, 0x20080000u, 0x08421000u, 0x02008020u, 0x00040008u,
    0x80200802u, 0x20000200u, 0x08000000u, 0x02008021u, 0x10800008u, 0x84200842u, 0x20080200u, 0x00020080u,
    0x02108021u, 0x00802108u, 0x00200802u, 0x20080200u, 0x08020084u, 0x02008000u, 0x00800008u, 0x84200800u,
    0x20084200u, 0x08020084u, 0x02008020u, 0x00802000u, 0x80200802u, 0x20080200u, 0x08020080u, 0x00000020u};

__device__ __forceinline__ int hue_bin(int r, int g, int b) {
    int maxc = max(r, max(g, b));
    int minc = min(r, min(g, b));
    int range = maxc - minc;
    if (range == 0) return 0;
    int hn = (r == maxc) ? (g - b) : (g == maxc) ? 2 * range + (b - r) : 4 * range + (r - g);
    uint32_t den = 6u * (uint32_t)range;
    if (hn < 0) hn += (int)den;
    uint32_t num = 1000u * (uint32_t)hn;                       // < 6000*255
    uint32_t q = __umulhi(num, __ldg(&d_hue_magic.m[range]));  // floor(num / den)
    uint32_t twice = 2u * (num - q * den);
    if (twice == den) return (int)(q + ((c_hue_tie_up[q >> 5] >> (q & 31)) & 1u));
    return (int)(q + (twice > den ? 1u : 0u));
}
// hue bin of a cloud record's colour word  R | G << 8 | B << 16
__device__ __forceinline__ int hue_of_hi(uint32_t hi) {
    return hue_bin((int)(hi & 0xFFu), (int)((hi >> 8) & 0xFFu), (int)((hi >> 16) & 0xFFu));
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
