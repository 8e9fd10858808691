// Pre-processing front end: gammaChange, BGR -> gray, equalizeHist
// (functions.py:61-67, 80-96; stereovision.py:44,46) for F images per call.
//
//   prep_gamma_gray   thread = 16 pixels (48 BGR bytes, three 128-bit loads): byte LUT from shared memory
//                     (cv2.LUT with the reference's 256-entry gamma table), gray = (3735 B + 19235 G + 9798 R
//                     + 2^14) >> 15 of the corrected colours (OpenCV's 15-bit fixed point), 256-bin histogram of
//                     the gray image in per-warp shared-memory copies, merged with global atomics per CTA.
//   prep_equalize_lut one CTA per image: running sum of the histogram, OpenCV's float32 scale
//                     255 / (total - hist[i0]) and round-half-even, i0 = first populated bin.
//   prep_apply_lut    gray -> equalised gray through the image's own table.
#include "svb_common.cuh"
#include "svb_prof.cuh"

#define PP_THREADS 256
#define PP_WARPS (PP_THREADS / 32)

struct ByteTable { uint8_t v[256]; };

// the same through a table replicated per lane (entry i of lane l in word 32 i + l): a warp's 32 look-ups fall into
// 32 different banks whatever the bytes are
__device__ __forceinline__ uint32_t lut4_lane(const uint32_t *t, uint32_t w) {
    // one PRMT per byte index, one per pair of results: the entries are words with the value in byte 0
    const uint32_t r0 = t[__byte_perm(w, 0, 0x4440) * 32], r1 = t[__byte_perm(w, 0, 0x4441) * 32];
    const uint32_t r2 = t[__byte_perm(w, 0, 0x4442) * 32], r3 = t[__byte_perm(w, 0, 0x4443) * 32];
    return __byte_perm(__byte_perm(r0, r1, 0x1140), __byte_perm(r2, r3, 0x1140), 0x5410);
}
// four table look-ups on the bytes of w
__device__ __forceinline__ uint32_t lut4(const uint8_t *t, uint32_t w) {
    return (uint32_t)t[w & 0xFFu] | ((uint32_t)t[(w >> 8) & 0xFFu] << 8) | ((uint32_t)t[(w >> 16) & 0xFFu] << 16) | ((uint32_t)t[w >> 24] << 24);
}
__device__ __forceinline__ uint32_t gray_of(uint32_t b, uint32_t g, uint32_t r) {
    return (b * 3735u + g * 19235u + r * 9798u + 16384u) >> 15;
}

template <bool HAS_TABLE>
__global__ void __launch_bounds__(PP_THREADS)
prep_gamma_gray_kernel(const uint8_t *__restrict__ bgr, int64_t pixels_per_image, ByteTable table, uint8_t *__restrict__ bgr_out,
                       uint8_t *__restrict__ gray, uint32_t *__restrict__ hist, int vec) {
    __shared__ uint8_t s_t[256];
    __shared__ uint32_t s_rep[HAS_TABLE ? 256 * 32 : 1];
    __shared__ uint32_t s_h[PP_WARPS][256];
    const int f = blockIdx.y, tid = threadIdx.x, warp = tid >> 5;
    if (HAS_TABLE) {
        s_t[tid] = table.v[tid];
        for (int k = tid; k < 256 * 32; k += PP_THREADS) s_rep[k] = table.v[k >> 5];
    }
    const uint32_t *my_t = s_rep + (tid & 31);
    for (int k = tid; k < PP_WARPS * 256; k += PP_THREADS) (&s_h[0][0])[k] = 0;
    __syncthreads();
    const uint8_t *src = bgr + (int64_t)f * pixels_per_image * 3;
    uint8_t *dst = bgr_out ? bgr_out + (int64_t)f * pixels_per_image * 3 : nullptr;
    uint8_t *gdst = gray ? gray + (int64_t)f * pixels_per_image : nullptr;
    uint32_t *h = s_h[warp];
    const int64_t groups = pixels_per_image / 16;
    for (int64_t gidx = (int64_t)blockIdx.x * PP_THREADS + tid; gidx < groups; gidx += (int64_t)gridDim.x * PP_THREADS) {
        uint32_t w[12];
        if (vec) {
            const uint4 *p = reinterpret_cast<const uint4 *>(src + gidx * 48);
            uint4 a = __ldcs(p), b = __ldcs(p + 1), c = __ldcs(p + 2);
            w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
            w[8] = c.x; w[9] = c.y; w[10] = c.z; w[11] = c.w;
        } else {
#pragma unroll
            for (int k = 0; k < 12; ++k) {
                const uint8_t *q = src + gidx * 48 + k * 4;
                w[k] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
            }
        }
        if (HAS_TABLE) {
#pragma unroll
            for (int k = 0; k < 12; ++k) w[k] = lut4_lane(my_t, w[k]);
        }
        if (dst) {
            if (vec) {
                uint4 *p = reinterpret_cast<uint4 *>(dst + gidx * 48);
                __stcs(p, make_uint4(w[0], w[1], w[2], w[3]));
                __stcs(p + 1, make_uint4(w[4], w[5], w[6], w[7]));
                __stcs(p + 2, make_uint4(w[8], w[9], w[10], w[11]));
            } else {
#pragma unroll
                for (int k = 0; k < 48; ++k) dst[gidx * 48 + k] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
            }
        }
        // 16 pixels: pixel i sits at bytes 3i .. 3i+2.  gray = (3735 B + 19235 G + 9798 R + 2^14) >> 15 as two DP2A on the
        // 32-bit window that starts at the pixel (its fourth byte meets a zero coefficient)
        uint32_t gq[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t y4[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int o = 3 * (4 * q + r);
                const uint32_t win = (o & 3) ? __funnelshift_r(w[o >> 2], w[(o >> 2) + 1 < 12 ? (o >> 2) + 1 : 11], 8 * (o & 3)) : w[o >> 2];
                const uint32_t y = __dp2a_hi(9798u, win, __dp2a_lo(3735u | (19235u << 16), win, 16384u)) >> 15;
                atomicAdd(&h[y], 1u);
                y4[r] = y;
            }
            gq[q] = __byte_perm(__byte_perm(y4[0], y4[1], 0x1140), __byte_perm(y4[2], y4[3], 0x1140), 0x5410);
        }
        if (gdst) {
            if (vec) __stcs(reinterpret_cast<uint4 *>(gdst + gidx * 16), make_uint4(gq[0], gq[1], gq[2], gq[3]));
            else {
#pragma unroll
                for (int k = 0; k < 16; ++k) gdst[gidx * 16 + k] = (uint8_t)(gq[k >> 2] >> (8 * (k & 3)));
            }
        }
    }
    // the image's last pixels_per_image % 16 pixels
    if (blockIdx.x == 0) {
        for (int64_t i = groups * 16 + tid; i < pixels_per_image; i += PP_THREADS) {
            uint32_t b = src[i * 3], g = src[i * 3 + 1], r = src[i * 3 + 2];
            if (HAS_TABLE) { b = s_t[b]; g = s_t[g]; r = s_t[r]; }
            if (dst) { dst[i * 3] = (uint8_t)b; dst[i * 3 + 1] = (uint8_t)g; dst[i * 3 + 2] = (uint8_t)r; }
            const uint32_t y = gray_of(b, g, r);
            if (gdst) gdst[i] = (uint8_t)y;
            atomicAdd(&h[y], 1u);
        }
    }
    __syncthreads();
    if (hist) {
        uint32_t t = 0;
#pragma unroll
        for (int k = 0; k < PP_WARPS; ++k) t += s_h[k][tid];
        if (t) atomicAdd(&hist[(int64_t)f * 256 + tid], t);
    }
}

extern "C" int svb_gamma_gray_hist(const uint8_t *bgr, int F, int H, int W, const uint8_t *h_table, uint8_t *bgr_out, uint8_t *gray,
                                   uint32_t *hist, void *stream) {
    if (!bgr) return SVB_E_NULL;
    if (!bgr_out && !gray && !hist) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t pixels = (int64_t)H * W;
    if (hist) {
        cudaError_t e = cudaMemsetAsync(hist, 0, (size_t)F * 256 * sizeof(uint32_t), st);
        if (e != cudaSuccess) return (int)e;
    }
    ByteTable t;
    for (int i = 0; i < 256; ++i) t.v[i] = h_table ? h_table[i] : (uint8_t)i;
    // 16-byte loads need every image to start on a 16-byte boundary
    const int vec = ((((uintptr_t)bgr | (uintptr_t)bgr_out | (uintptr_t)gray) & 15) == 0) && (pixels % 16 == 0 || F == 1);
    int64_t blocks = (pixels / 16 + PP_THREADS - 1) / PP_THREADS;
    const int64_t want = (148 * 8 + F - 1) / F;                    // about 8 CTAs per SM over the batch
    if (blocks > want) blocks = want;
    if (blocks < 1) blocks = 1;
    dim3 grid((unsigned)blocks, F);
    ProfScope ps(K_PREP_GRAY, st);
    if (h_table) prep_gamma_gray_kernel<true><<<grid, PP_THREADS, 0, st>>>(bgr, pixels, t, bgr_out, gray, hist, vec);
    else         prep_gamma_gray_kernel<false><<<grid, PP_THREADS, 0, st>>>(bgr, pixels, t, bgr_out, gray, hist, vec);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ equalizeHist's table
__global__ void __launch_bounds__(256)
prep_equalize_lut_kernel(const uint32_t *__restrict__ hist, uint8_t *__restrict__ lut) {
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_first[8];
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t h = hist[(int64_t)f * 256 + tid];
    uint32_t inc = h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, h != 0);
    if (lane == 31) s_w[warp] = inc;
    if (lane == 0) s_first[warp] = bal ? (uint32_t)(warp * 32 + __ffs(bal) - 1) : 0xFFFFFFFFu;
    __syncthreads();
    uint32_t before = 0, total = 0, i0 = 0xFFFFFFFFu;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (k < warp) before += s_w[k];
        total += s_w[k];
        i0 = min(i0, s_first[k]);
    }
    inc += before;                                                   // sum of hist[0 .. tid]
    uint8_t out = 0;
    if (i0 != 0xFFFFFFFFu) {
        const uint32_t h0 = hist[(int64_t)f * 256 + i0];
        if (h0 == total) out = (uint8_t)i0;                          // one grey level: the image maps to itself
        else if ((uint32_t)tid > i0) {
            const float scale = __fdiv_rn(255.0f, (float)(total - h0));
            const int v = __float2int_rn(__fmul_rn((float)(inc - h0), scale));
            out = (uint8_t)min(255, max(0, v));
        }
    }
    lut[(int64_t)f * 256 + tid] = out;
}

extern "C" int svb_equalize_lut(const uint32_t *hist, int F, uint8_t *lut, void *stream) {
    if (!hist || !lut) return SVB_E_NULL;
    if (F < 1) return SVB_E_SHAPE;
    ProfScope ps(K_PREP_EQ_LUT, (cudaStream_t)stream);
    prep_equalize_lut_kernel<<<F, 256, 0, (cudaStream_t)stream>>>(hist, lut);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ per-image byte tables
__global__ void __launch_bounds__(PP_THREADS)
prep_apply_lut_kernel(const uint8_t *__restrict__ src, int64_t n, const uint8_t *__restrict__ lut, int lut_stride, uint8_t *__restrict__ dst, int vec) {
    __shared__ uint8_t s_t[256];
    const int f = blockIdx.y, tid = threadIdx.x;
    s_t[tid] = lut[(int64_t)f * lut_stride + tid];
    __syncthreads();
    const uint8_t *s = src + (int64_t)f * n;
    uint8_t *d = dst + (int64_t)f * n;
    const int64_t groups = vec ? n / 16 : 0;
    const int64_t stride = (int64_t)gridDim.x * PP_THREADS;
    for (int64_t g = (int64_t)blockIdx.x * PP_THREADS + tid; g < groups; g += 4 * stride) {        // four loads in flight
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g + u * stride < groups) v[u] = __ldcs(reinterpret_cast<const uint4 *>(s) + g + u * stride);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g + u * stride < groups)
                __stcs(reinterpret_cast<uint4 *>(d) + g + u * stride,
                       make_uint4(lut4(s_t, v[u].x), lut4(s_t, v[u].y), lut4(s_t, v[u].z), lut4(s_t, v[u].w)));
    }
    for (int64_t i = groups * 16 + (int64_t)blockIdx.x * PP_THREADS + tid; i < n; i += (int64_t)gridDim.x * PP_THREADS) d[i] = s_t[s[i]];
}

static int launch_apply(const uint8_t *src, int F, int64_t n, const uint8_t *d_lut, int lut_stride, uint8_t *dst, cudaStream_t st) {
    const int vec = ((((uintptr_t)src | (uintptr_t)dst) & 15) == 0) && (n % 16 == 0 || F == 1);
    int64_t blocks = (n / 16 + PP_THREADS - 1) / PP_THREADS;
    const int64_t want = (148 * 8 + F - 1) / F;
    if (blocks > want) blocks = want;
    if (blocks < 1) blocks = 1;
    ProfScope ps(K_PREP_APPLY, st);
    prep_apply_lut_kernel<<<dim3((unsigned)blocks, F), PP_THREADS, 0, st>>>(src, n, d_lut, lut_stride, dst, vec);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_apply_lut_frames(const uint8_t *src, int F, int64_t n_per_frame, const uint8_t *lut, uint8_t *dst, void *stream) {
    if (!src || !lut || !dst) return SVB_E_NULL;
    if (F < 1 || n_per_frame < 1) return SVB_E_SHAPE;
    return launch_apply(src, F, n_per_frame, lut, 256, dst, (cudaStream_t)stream);
}

// one table for every byte (cv2.LUT, functions.py:67): lut is a DEVICE pointer to 256 bytes
extern "C" int svb_lut_bytes(const uint8_t *src, int64_t n, const uint8_t *lut, uint8_t *dst, void *stream) {
    if (!src || !lut || !dst) return SVB_E_NULL;
    if (n < 1) return SVB_E_SHAPE;
    return launch_apply(src, 1, n, lut, 0, dst, (cudaStream_t)stream);
}
