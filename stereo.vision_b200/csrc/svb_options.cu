// the 'mean' hole filling (fillAltDisparity, functions.py:149-162) and the road-normal glyph
// (getNormalVectorLine, functions.py:396-416).
//
//   fill_row_mean   one warp per image row: sum and number of the non-zero entries (DP4A byte sums), the row's
//                   fill value = trunc(float64 mean) = sum / count in integers (the quotient of two integers
//                   below 2^53 cannot round across an integer), then every entry below 2 is replaced.  The row is
//                   read twice, the second time from L1/L2.
//   normal_line     one thread per frame, the reference's float64 sequence operation by operation.
#include "svb_common.cuh"
#include "svb_prof.cuh"
#include <math.h>

#define FM_THREADS 256
#define FM_WARPS (FM_THREADS / 32)

// bytes below 2 -> 0xFF
__device__ __forceinline__ uint32_t below2(uint32_t w) {
    const uint32_t t = w & 0xFEFEFEFEu;                              // zero exactly for 0 and 1
    const uint32_t nz = (((t & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t) & 0x80808080u;
    uint32_t r;
    asm("prmt.b32 %0, %1, %1, 0xBA98;" : "=r"(r) : "r"(~nz));     // sign replication: 0xFF where the byte of t is zero
    return r;
}
__device__ __forceinline__ uint32_t nonzero_count4(uint32_t w) {
    return (uint32_t)__popc((((w & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | w) & 0x80808080u);
}

__global__ void __launch_bounds__(FM_THREADS)
fill_row_mean_kernel(const uint8_t *__restrict__ src, int64_t rows, int W, int vec, uint8_t *__restrict__ dst) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * FM_WARPS + (threadIdx.x >> 5);
    if (row >= rows) return;
    const uint8_t *s = src + row * W;
    uint8_t *d = dst + row * W;
    const int groups = vec ? W / 16 : 0;
    uint32_t sum = 0, cnt = 0;
    for (int g = lane; g < groups; g += 32) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(s) + g);
        sum = __dp4a(v.x, 0x01010101u, sum); sum = __dp4a(v.y, 0x01010101u, sum);
        sum = __dp4a(v.z, 0x01010101u, sum); sum = __dp4a(v.w, 0x01010101u, sum);
        cnt += nonzero_count4(v.x) + nonzero_count4(v.y) + nonzero_count4(v.z) + nonzero_count4(v.w);
    }
    for (int i = groups * 16 + lane; i < W; i += 32) {
        const uint32_t v = s[i];
        sum += v; cnt += v != 0;
    }
    sum = __reduce_add_sync(0xffffffffu, sum);
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    const uint32_t m = cnt ? sum / cnt : 0u;                         // functions.py:153-156 (nan -> 0.0), truncated by the store
    const uint32_t m4 = m * 0x01010101u;
    for (int g = lane; g < groups; g += 32) {
        uint4 v = __ldg(reinterpret_cast<const uint4 *>(s) + g);
        uint32_t h;
        h = below2(v.x); v.x = (v.x & ~h) | (m4 & h);
        h = below2(v.y); v.y = (v.y & ~h) | (m4 & h);
        h = below2(v.z); v.z = (v.z & ~h) | (m4 & h);
        h = below2(v.w); v.w = (v.w & ~h) | (m4 & h);
        reinterpret_cast<uint4 *>(d)[g] = v;
    }
    for (int i = groups * 16 + lane; i < W; i += 32) {
        const uint8_t v = s[i];
        d[i] = v < 2 ? (uint8_t)m : v;
    }
}

extern "C" int svb_fill_row_mean(const uint8_t *disparity, int F, int H, int W, uint8_t *filled, void *stream) {
    if (!disparity || !filled) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    const int64_t rows = (int64_t)F * H;
    const int vec = (W % 16 == 0) && ((((uintptr_t)disparity | (uintptr_t)filled) & 15) == 0);
    ProfScope ps(K_FILL_MEAN, (cudaStream_t)stream);
    fill_row_mean_kernel<<<(unsigned)((rows + FM_WARPS - 1) / FM_WARPS), FM_THREADS, 0, (cudaStream_t)stream>>>(disparity, rows, W, vec, filled);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ the normal glyph's end point
__global__ void normal_line_kernel(const int32_t *__restrict__ base_xy, const double *__restrict__ abc, int abc_stride,
                                   const uint8_t *__restrict__ disparity, const uint8_t *__restrict__ d_values, int F, int H, int W,
                                   CamD cam, int32_t *__restrict__ out_xy, int32_t *__restrict__ status) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    const int x = base_xy[2 * f], y = base_xy[2 * f + 1];
    int st = 0;
    double dv;
    if (disparity) {
        if (x < 0 || y < 0 || x >= W || y >= H) { out_xy[2 * f] = out_xy[2 * f + 1] = 0; status[f] = 4; return; }   // IndexError
        dv = (double)disparity[((int64_t)f * H + y) * W + x];
    } else dv = (double)d_values[f];
    const double a = abc[(int64_t)f * abc_stride], b = abc[(int64_t)f * abc_stride + 1], c = abc[(int64_t)f * abc_stride + 2];
    double Z = __ddiv_rn(cam.fB, dv);                                                     // :402
    const double X = __ddiv_rn(__dmul_rn(__dsub_rn((double)x, cam.cw), Z), cam.f);        // :403
    const double Y = __ddiv_rn(__dmul_rn(__dsub_rn((double)y, cam.ch), Z), cam.f);        // :404
    double newY = __dsub_rn(Y, 0.7), newX = __dadd_rn(X, 0.0);                            // :406-407
    const double dd = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(a, a), __dmul_rn(b, b)), __dmul_rn(c, c)));   // :409
    Z = __dsub_rn(dd, __dadd_rn(__dmul_rn(a, newX), __dmul_rn(b, newY)));                 // :411
    newX = __dadd_rn(__ddiv_rn(__dmul_rn(newX, cam.f), Z), cam.cw);                       // :413
    newY = __dadd_rn(__ddiv_rn(__dmul_rn(newY, cam.f), Z), cam.ch);                       // :414
    // int(): nan raises ValueError, inf OverflowError; the first failing conversion decides (:415)
    if (isnan(newX)) st = 1; else if (isinf(newX)) st = 2; else if (isnan(newY)) st = 1; else if (isinf(newY)) st = 2;
    int ox = 0, oy = 0;
    if (!st) {
        // Python ints are unbounded; the int32 pair saturates (no image is that large)
        ox = (int)fmax(-2147483648.0, fmin(2147483647.0, trunc(newX)));
        oy = (int)fmax(-2147483648.0, fmin(2147483647.0, trunc(newY)));
    }
    out_xy[2 * f] = ox; out_xy[2 * f + 1] = oy;
    status[f] = st;
}

extern "C" int svb_normal_line(const int32_t *base_xy, const double *abc, int abc_stride, const uint8_t *disparity,
                               const uint8_t *d_values, int F, int H, int W, const svb_camera *cam, int32_t *out_xy,
                               int32_t *status, void *stream) {
    if (!base_xy || !abc || !cam || !out_xy || !status || (!disparity && !d_values)) return SVB_E_NULL;
    if (F < 1 || abc_stride < 3 || (disparity && (H < 1 || W < 1))) return SVB_E_SHAPE;
    ProfScope ps(K_NORMAL_LINE, (cudaStream_t)stream);
    normal_line_kernel<<<(F + 127) / 128, 128, 0, (cudaStream_t)stream>>>(base_xy, abc, abc_stride, disparity, d_values, F, H, W,
                                                                          make_camd(cam), out_xy, status);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
