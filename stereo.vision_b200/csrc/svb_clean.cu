// Road-mask cleaning: sanitiseRoadImage, functions.py:346-366, on bit-packed rows.
//
// The reference runs closing(9x9) -> erode(9x9) x2 -> AND road_threshold_mask -> closing(9x9) ->
// ParticleCleansing on a {0,255} byte image and then lists the non-zero pixels in row-major order.
// With OpenCV's default border (outside pixels never win a max or a min) and square kernels this is
//
//      dH9 dV9 | eV25 eH25 | & T | dH9 dV9 | eV9 eH9          (d = dilate, e = erode, H/V = 1-D pass)
//
// because two erosions by 9 after the closing's erosion by 9 are one erosion by 25, and each square
// pass splits into a row pass and a column pass in either order.  ParticleCleansing paints contours of
// area < 70 white, which cannot change a closed image (the golden fixtures keep the step).
//
// One CTA cleans a band of rows of one frame: it loads the band plus a 24-row halo (4+12+4+4), packs
// bytes to bits (32 pixels per word, LSB = leftmost), runs the nine passes between two shared-memory
// planes, and writes the cleaned bytes, the cleaned bit rows and the per-row population counts.  A
// second kernel turns bit rows + row counts into the (x,y) list in the reference's order.
#include "svb_common.cuh"
#include "svb_prof.cuh"

#define CL_THREADS 256
#define CL_HALO 24
#define CL_SEG 12               // rows per thread in a column pass
#define CL_PLANE_WORDS 6144     // target size of one shared-memory plane (24 KB)

// ------------------------------------------------------------------ bytes <-> bits
// A pixel is read by its top bit; *odd collects every byte that is neither 0 nor 255 (the caller's
// status flag: such an image is not a mask and the result is then not the reference's).
// 4 bytes -> 4 bits in the top nibble: bits 7,15,23,31 times (1 + 2^7 + 2^14 + 2^21) land on 28..31 and
// no two partial products share a position, so nothing carries.
// every byte -> 0x00 / 0xFF by its top bit (prmt with selector bit 3 set replicates the sign)
__device__ __forceinline__ uint32_t byte_signs(uint32_t v) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %1, 0xBA98;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ uint32_t pack4_top(uint32_t v, uint32_t &odd) {
    odd |= v ^ byte_signs(v);
    return (v & 0x80808080u) * 0x00204081u;
}
__device__ __forceinline__ uint32_t pack16(uint4 q, uint32_t &odd) {
    uint32_t a = pack4_top(q.x, odd) >> 28, b = pack4_top(q.y, odd) >> 28, c = pack4_top(q.z, odd) >> 28, d = pack4_top(q.w, odd) >> 28;
    return a + (b << 4) + (c << 8) + (d << 12);
}
// the same for any non-zero byte (svb_pack_mask: a cv2 mask argument is read by non-zero-ness)
__device__ __forceinline__ uint32_t pack4_any(uint32_t v) {
    uint32_t hi = (((v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | v) & 0x80808080u;      // top bit of every non-zero byte
    return (hi * 0x00204081u) >> 28;
}
__device__ __forceinline__ uint32_t pack16_any(uint4 q) {
    return pack4_any(q.x) + (pack4_any(q.y) << 4) + (pack4_any(q.z) << 8) + (pack4_any(q.w) << 12);
}
// 4 bits -> 4 bytes of 0 / 255: move bit k to bit 8k+7 with one multiply, then replicate the signs.
// `nib` holds the four bits at 0..3 (LOW) or at 4..7 (!LOW), everything else zero.
template <bool LOW>
__device__ __forceinline__ uint32_t spread4(uint32_t nib) {
    return byte_signs(nib * (LOW ? 0x10204080u : 0x01020408u));
}
__device__ __forceinline__ uint4 spread16(uint32_t h) {          // h < 65536
    uint32_t g = h >> 8;
    return make_uint4(spread4<true>(h & 0xFu), spread4<false>(h & 0xF0u), spread4<true>(g & 0xFu), spread4<false>(g & 0xF0u));
}

// 32 pixels of a row starting at x0 as one word; vec = row base and W allow 16-byte loads
template <bool ANY_NONZERO>
__device__ __forceinline__ uint32_t load_word(const uint8_t *row, int x0, int W, bool vec, uint32_t &odd) {
    if (vec && x0 + 32 <= W) {
        const uint4 *p = reinterpret_cast<const uint4 *>(row + x0);
        uint4 a = __ldg(p), b = __ldg(p + 1);
        if (ANY_NONZERO) return pack16_any(a) | (pack16_any(b) << 16);
        return pack16(a, odd) | (pack16(b, odd) << 16);
    }
    uint32_t w = 0;
    int n = min(32, W - x0);
    for (int i = 0; i < n; ++i) {
        uint32_t v = row[x0 + i];
        w |= (ANY_NONZERO ? (uint32_t)(v != 0) : (v >> 7)) << i;
        odd |= (v != 0 && v != 255) ? 1u : 0u;
    }
    return w;
}
__device__ __forceinline__ void store_word(uint8_t *row, int x0, int W, bool vec, uint32_t w) {
    if (vec && x0 + 32 <= W) {
        uint4 *p = reinterpret_cast<uint4 *>(row + x0);
        p[0] = spread16(w & 0xFFFFu);
        p[1] = spread16(w >> 16);
        return;
    }
    int n = min(32, W - x0);
    for (int i = 0; i < n; ++i) row[x0 + i] = ((w >> i) & 1u) ? 255 : 0;
}

// ------------------------------------------------------------------ row passes (bits of l | c | r)
// OR of the bits x-R .. x+R for the 32 positions of c; bits outside the row are 0
template <int R>
__device__ __forceinline__ uint32_t hdilate(uint32_t l, uint32_t c, uint32_t r) {
    if constexpr (R <= 4) {
        uint32_t o = c;
#pragma unroll
        for (int k = 1; k <= R; ++k) o |= __funnelshift_l(l, c, k) | __funnelshift_r(c, r, k);
        return o;
    } else {
        static_assert(R == 12, "row pass radius");
        // left side on the pair (hi = c, lo = l): windows of 2, 4, 8 then 8 | 4 << 8 | 1 << 12 = 13
        uint32_t h1 = c | __funnelshift_l(l, c, 1), l1 = l | (l << 1);
        uint32_t h2 = h1 | __funnelshift_l(l1, h1, 2), l2 = l1 | (l1 << 2);
        uint32_t h4 = h2 | __funnelshift_l(l2, h2, 4);
        uint32_t left = h4 | __funnelshift_l(l2, h2, 8) | __funnelshift_l(l, c, 12);
        // right side on the pair (hi = r, lo = c)
        uint32_t g1 = c | __funnelshift_r(c, r, 1), r1 = r | (r >> 1);
        uint32_t g2 = g1 | __funnelshift_r(g1, r1, 2), r2 = r1 | (r1 >> 2);
        uint32_t g4 = g2 | __funnelshift_r(g2, r2, 4);
        uint32_t right = g4 | __funnelshift_r(g2, r2, 8) | __funnelshift_r(c, r, 12);
        return left | right;
    }
}

// one row pass over a plane: dst = op(src); ERODE works on the complement (outside bits = 0 there too)
template <int R, bool ERODE>
__device__ __forceinline__ uint32_t hpass_word(const uint32_t *srow, int c, int WW, uint32_t tail) {
    uint32_t l = c > 0 ? srow[c - 1] : (ERODE ? ~0u : 0u);
    uint32_t m = srow[c];
    uint32_t r = c + 1 < WW ? srow[c + 1] : (ERODE ? ~0u : 0u);
    if (ERODE) {
        l = ~l; m = ~m; r = ~r;
        if (c + 1 == WW) m &= tail; else if (c + 2 == WW) r &= tail;
        uint32_t o = ~hdilate<R>(l, m, r);
        return c + 1 == WW ? (o & tail) : o;
    }
    uint32_t o = hdilate<R>(l, m, r);          // the tail bits of src are 0 by construction
    return c + 1 == WW ? (o & tail) : o;
}

// ------------------------------------------------------------------ column passes
struct OrOp  { static __device__ __forceinline__ uint32_t f(uint32_t a, uint32_t b) { return a | b; } static constexpr uint32_t id = 0u; };
struct AndOp { static __device__ __forceinline__ uint32_t f(uint32_t a, uint32_t b) { return a & b; } static constexpr uint32_t id = ~0u; };

// dst[y] = op over src[y-R .. y+R] (rows outside [0,NR) skipped) for CL_SEG rows of one word column
template <int R, class Op>
__device__ __forceinline__ void vpass_item(const uint32_t *src, uint32_t *dst, int c, int WW, int y0, int NR) {
    constexpr int N = CL_SEG + 2 * R;
    uint32_t s[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        int y = y0 - R + i;
        s[i] = (y >= 0 && y < NR) ? src[y * WW + c] : Op::id;
    }
    if constexpr (R == 4) {
        uint32_t t1[N - 1], t2[N - 3], t4[N - 7];
#pragma unroll
        for (int i = 0; i < N - 1; ++i) t1[i] = Op::f(s[i], s[i + 1]);
#pragma unroll
        for (int i = 0; i < N - 3; ++i) t2[i] = Op::f(t1[i], t1[i + 2]);
#pragma unroll
        for (int i = 0; i < N - 7; ++i) t4[i] = Op::f(t2[i], t2[i + 4]);
#pragma unroll
        for (int i = 0; i < CL_SEG; ++i) {
            int y = y0 + i;
            if (y < NR) dst[y * WW + c] = Op::f(t4[i], s[i + 8]);
        }
    } else {
        static_assert(R == 4 || R == 12, "column pass radius");
        uint32_t t1[N - 1], t2[N - 3], t4[N - 7], t16[CL_SEG];
#pragma unroll
        for (int i = 0; i < N - 1; ++i) t1[i] = Op::f(s[i], s[i + 1]);
#pragma unroll
        for (int i = 0; i < N - 3; ++i) t2[i] = Op::f(t1[i], t1[i + 2]);
#pragma unroll
        for (int i = 0; i < N - 7; ++i) t4[i] = Op::f(t2[i], t2[i + 4]);       // rows i .. i+7
#pragma unroll
        for (int i = 0; i < CL_SEG; ++i) t16[i] = Op::f(t4[i], t4[i + 8]);      // rows i .. i+15
#pragma unroll
        for (int i = 0; i < CL_SEG; ++i) {
            int y = y0 + i;
            // 25 = 16 + 8 + 1
            if (y < NR) dst[y * WW + c] = Op::f(Op::f(t16[i], t4[i + 16]), s[i + 24]);
        }
    }
}

template <int R, class Op>
__device__ __forceinline__ void vpass(const uint32_t *src, uint32_t *dst, int WW, int NR) {
    int segs = (NR + CL_SEG - 1) / CL_SEG;
    for (int it = threadIdx.x; it < segs * WW; it += CL_THREADS) {
        int seg = it / WW, c = it - seg * WW;
        vpass_item<R, Op>(src, dst, c, WW, seg * CL_SEG, NR);
    }
}

template <int R, bool ERODE>
__device__ __forceinline__ void hpass(const uint32_t *src, uint32_t *dst, int WW, int NR, uint32_t tail) {
    for (int it = threadIdx.x; it < NR * WW; it += CL_THREADS) {
        int y = it / WW, c = it - y * WW;
        dst[it] = hpass_word<R, ERODE>(src + y * WW, c, WW, tail);
    }
}

// ------------------------------------------------------------------ the band kernel
struct CleanGeom {
    int H, W, WW;        // WW = words per row
    int band;            // rows per band
    int vec;             // rows are 16-byte loadable
    uint32_t tail;       // valid bits of a row's last word
};

template <bool PACKED_IN>
__global__ void __launch_bounds__(CL_THREADS)
clean_band_kernel(const uint8_t *__restrict__ road, const uint32_t *__restrict__ road_bits, const uint32_t *__restrict__ tbits,
                  CleanGeom g, uint8_t *__restrict__ cleaned, uint32_t *__restrict__ cleaned_bits,
                  int32_t *__restrict__ rowcnt, int32_t *__restrict__ status) {
    extern __shared__ uint32_t s_mem[];
    const int WW = g.WW, f = blockIdx.y;
    const int y0 = blockIdx.x * g.band, y1 = min(g.H, y0 + g.band);
    const int ylo = max(0, y0 - CL_HALO), yhi = min(g.H, y1 + CL_HALO);
    const int NR = yhi - ylo;
    uint32_t *A = s_mem, *B = s_mem + (g.band + 2 * CL_HALO) * WW;
    __shared__ int s_cnt[SVB_MAX_DIM / 8];      // per-row counts of this band (band <= 256 rows)

    // ---- load + pack
    uint32_t odd = 0;
    if (PACKED_IN) {
        const uint32_t *src = road_bits + ((int64_t)f * g.H + ylo) * WW;
        for (int it = threadIdx.x; it < NR * WW; it += CL_THREADS) A[it] = __ldg(src + it);
    } else {
        const uint8_t *src = road + ((int64_t)f * g.H + ylo) * g.W;
#pragma unroll 2
        for (int it = threadIdx.x; it < NR * WW; it += CL_THREADS) {
            int y = it / WW, c = it - y * WW;
            A[it] = load_word<false>(src + (int64_t)y * g.W, c * 32, g.W, g.vec, odd);
        }
    }
    for (int i = threadIdx.x; i < g.band; i += CL_THREADS) s_cnt[i] = 0;
    __syncthreads();
    hpass<4, false>(A, B, WW, NR, g.tail);              // dH9
    __syncthreads();
    vpass<4, OrOp>(B, A, WW, NR);                       // dV9
    __syncthreads();
    vpass<12, AndOp>(A, B, WW, NR);                     // eV25
    __syncthreads();
    {                                                   // eH25, & T
        const uint32_t *t = tbits + (int64_t)ylo * WW;
        for (int it = threadIdx.x; it < NR * WW; it += CL_THREADS) {
            int y = it / WW, c = it - y * WW;
            A[it] = hpass_word<12, true>(B + y * WW, c, WW, g.tail) & __ldg(t + it);
        }
    }
    __syncthreads();
    hpass<4, false>(A, B, WW, NR, g.tail);              // dH9
    __syncthreads();
    vpass<4, OrOp>(B, A, WW, NR);                       // dV9
    __syncthreads();
    vpass<4, AndOp>(A, B, WW, NR);                      // eV9
    __syncthreads();
    // ---- eH9 on the band's own rows, store
    {
        const int r0 = y0 - ylo, rows = y1 - y0;
        const bool warp_rows = (WW & 31) == 0;          // a warp's 32 words lie in one row
        uint8_t *dst = cleaned ? cleaned + ((int64_t)f * g.H + y0) * g.W : nullptr;
        uint32_t *dbits = cleaned_bits ? cleaned_bits + ((int64_t)f * g.H + y0) * WW : nullptr;
        const int total = rows * WW, padded = (total + 31) & ~31;
        for (int it = threadIdx.x; it < padded; it += CL_THREADS) {
            uint32_t w = 0;
            int y = 0, c = 0;
            if (it < total) {
                y = it / WW; c = it - y * WW;
                w = hpass_word<4, true>(B + (r0 + y) * WW, c, WW, g.tail);
                if (dbits) dbits[it] = w;
                if (dst) store_word(dst + (int64_t)y * g.W, c * 32, g.W, g.vec, w);
            }
            int n = __popc(w);
            if (warp_rows) {
                n = __reduce_add_sync(0xFFFFFFFFu, n);
                if ((threadIdx.x & 31) == 0 && n) atomicAdd(&s_cnt[y], n);
            } else if (n) {
                atomicAdd(&s_cnt[y], n);
            }
        }
        __syncthreads();
        if (rowcnt)
            for (int i = threadIdx.x; i < rows; i += CL_THREADS) rowcnt[(int64_t)f * g.H + y0 + i] = s_cnt[i];
    }
    if (!PACKED_IN && status) {
        odd = __reduce_or_sync(0xFFFFFFFFu, odd);
        if ((threadIdx.x & 31) == 0 && odd) atomicOr(status, 1);
    }
}

// ------------------------------------------------------------------ the band kernel for 1024-pixel rows
// A row is exactly one word per lane: the row passes run in registers with the neighbours' words taken
// by shuffle (so consecutive row passes fuse, and a row is rewritten in place by the warp that owns it),
// and the planes carry rows of the pass's identity above and below the data, so the column passes load
// without bounds tests (plane A feeds the dilations: zeros, plane B the erosions: ones).
#define C32_TOP 12                       // identity rows above the data (largest column radius)
#define C32_BOT (12 + CL_SEG)            // below: radius + the overhang of the last row segment
#define C32_WW 32

template <int R, class Op, int THREADS>
__device__ __forceinline__ void vpass32(const uint32_t *src, uint32_t *dst, int NR) {
    constexpr int N = CL_SEG + 2 * R;
    const int segs = (NR + CL_SEG - 1) / CL_SEG;
    for (int it = threadIdx.x; it < segs * C32_WW; it += THREADS) {
        const int y0 = (it >> 5) * CL_SEG, c = it & 31;
        const uint32_t *s0 = src + (y0 - R + C32_TOP) * C32_WW + c;
        uint32_t *d0 = dst + (y0 + C32_TOP) * C32_WW + c;
        uint32_t s[N], o[CL_SEG];
#pragma unroll
        for (int i = 0; i < N; ++i) s[i] = s0[i * C32_WW];
        uint32_t t1[N - 1], t2[N - 3], t4[N - 7];
#pragma unroll
        for (int i = 0; i < N - 1; ++i) t1[i] = Op::f(s[i], s[i + 1]);
#pragma unroll
        for (int i = 0; i < N - 3; ++i) t2[i] = Op::f(t1[i], t1[i + 2]);
#pragma unroll
        for (int i = 0; i < N - 7; ++i) t4[i] = Op::f(t2[i], t2[i + 4]);               // rows i .. i+7
        if constexpr (R == 4) {
#pragma unroll
            for (int i = 0; i < CL_SEG; ++i) o[i] = Op::f(t4[i], s[i + 8]);
        } else {
            static_assert(R == 4 || R == 12, "column pass radius");
#pragma unroll
            for (int i = 0; i < CL_SEG; ++i) o[i] = Op::f(Op::f(Op::f(t4[i], t4[i + 8]), t4[i + 16]), s[i + 24]);
        }
#pragma unroll
        for (int i = 0; i < CL_SEG; ++i)
            if (y0 + i < NR) d0[i * C32_WW] = o[i];
    }
}

// row passes on one word per lane
template <int R>
__device__ __forceinline__ uint32_t hdilate32(uint32_t m, int lane) {
    uint32_t l = __shfl_up_sync(0xFFFFFFFFu, m, 1), r = __shfl_down_sync(0xFFFFFFFFu, m, 1);
    if (lane == 0) l = 0u;
    if (lane == 31) r = 0u;
    return hdilate<R>(l, m, r);
}
template <int R>
__device__ __forceinline__ uint32_t herode32(uint32_t m, int lane) { return ~hdilate32<R>(~m, lane); }

template <bool PACKED_IN, int THREADS>
__global__ void __launch_bounds__(THREADS)
clean_rows32_kernel(const uint8_t *__restrict__ road, const uint32_t *__restrict__ road_bits, const uint32_t *__restrict__ tbits,
                    int H, int band, uint8_t *__restrict__ cleaned, uint32_t *__restrict__ cleaned_bits,
                    int32_t *__restrict__ rowcnt, int32_t *__restrict__ status) {
    extern __shared__ uint32_t s_mem[];
    constexpr int W = 1024, NW = THREADS / 32;
    const int f = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int y0 = blockIdx.x * band, y1 = min(H, y0 + band);
    const int ylo = max(0, y0 - CL_HALO), yhi = min(H, y1 + CL_HALO);
    const int NR = yhi - ylo;
    const int plane = (band + 2 * CL_HALO + C32_TOP + C32_BOT) * C32_WW;
    uint32_t *A = s_mem, *B = s_mem + plane;
    uint32_t *Ad = A + C32_TOP * C32_WW, *Bd = B + C32_TOP * C32_WW;        // first data row
    for (int i = threadIdx.x; i < C32_TOP * C32_WW; i += THREADS) { A[i] = 0u; B[i] = ~0u; }
    for (int i = threadIdx.x; i < C32_BOT * C32_WW; i += THREADS) { Ad[NR * C32_WW + i] = 0u; Bd[NR * C32_WW + i] = ~0u; }

    // ---- load, pack, dH9 -> A
    uint32_t odd = 0;
    if (PACKED_IN) {
        const uint32_t *src = road_bits + ((int64_t)f * H + ylo) * C32_WW;
        for (int y = warp; y < NR; y += NW) Ad[y * C32_WW + lane] = hdilate32<4>(__ldg(src + y * C32_WW + lane), lane);
    } else {
        // a lane loads the row's 16-byte chunks `lane` and `32 + lane`, so that each load instruction of the warp
        // covers 512 contiguous bytes (whole sectors); two shuffles bring word `lane`'s halves (chunks 2 lane, 2 lane + 1)
        const uint4 *src = reinterpret_cast<const uint4 *>(road + ((int64_t)f * H + ylo) * W) + lane;
        const int s0 = (2 * lane) & 31, s1 = s0 + 1;
        auto gather = [&](uint32_t v) {
            const uint32_t a = __shfl_sync(0xFFFFFFFFu, v, s0), b = __shfl_sync(0xFFFFFFFFu, v, s1);
            return lane < 16 ? (a & 0xFFFFu) | (b << 16) : (a >> 16) | (b & 0xFFFF0000u);
        };
        int y = warp;
        for (; y + NW < NR; y += 2 * NW) {                                 // two rows in flight per warp
            const uint4 *p = src + (int64_t)y * (W / 16), *q = p + NW * (W / 16);
            uint4 a0 = __ldg(p), a1 = __ldg(p + 32), b0 = __ldg(q), b1 = __ldg(q + 32);
            uint32_t wa = gather(pack16(a0, odd) | (pack16(a1, odd) << 16)), wb = gather(pack16(b0, odd) | (pack16(b1, odd) << 16));
            Ad[y * C32_WW + lane] = hdilate32<4>(wa, lane);
            Ad[(y + NW) * C32_WW + lane] = hdilate32<4>(wb, lane);
        }
        for (; y < NR; y += NW) {
            const uint4 *p = src + (int64_t)y * (W / 16);
            uint4 a0 = __ldg(p), a1 = __ldg(p + 32);
            Ad[y * C32_WW + lane] = hdilate32<4>(gather(pack16(a0, odd) | (pack16(a1, odd) << 16)), lane);
        }
    }
    __syncthreads();
    vpass32<4, OrOp, THREADS>(A, B, NR);                 // dV9
    __syncthreads();
    vpass32<12, AndOp, THREADS>(B, A, NR);               // eV25
    __syncthreads();
    {                                                    // eH25, & T, dH9: in place, a warp owns its rows
        const uint32_t *t = tbits + (int64_t)ylo * C32_WW + lane;
        for (int y = warp; y < NR; y += NW) {
            uint32_t m = herode32<12>(Ad[y * C32_WW + lane], lane) & __ldg(t + y * C32_WW);
            Ad[y * C32_WW + lane] = hdilate32<4>(m, lane);
        }
    }
    __syncthreads();
    vpass32<4, OrOp, THREADS>(A, B, NR);                 // dV9
    __syncthreads();
    vpass32<4, AndOp, THREADS>(B, A, NR);                // eV9
    __syncthreads();
    {                                                    // eH9 on the band's own rows, store
        const int r0 = y0 - ylo, rows = y1 - y0;
        uint4 *dst = cleaned ? reinterpret_cast<uint4 *>(cleaned + ((int64_t)f * H + y0) * W) + lane : nullptr;
        const int sh = (lane & 1) * 16;
        uint32_t *dbits = cleaned_bits ? cleaned_bits + ((int64_t)f * H + y0) * C32_WW + lane : nullptr;
        int32_t *rc = rowcnt ? rowcnt + (int64_t)f * H + y0 : nullptr;
        for (int y = warp; y < rows; y += NW) {
            uint32_t w = herode32<4>(Ad[(r0 + y) * C32_WW + lane], lane);
            if (dbits) dbits[y * C32_WW] = w;
            if (dst) {
                // chunk `lane` is half (lane & 1) of word lane / 2, chunk 32 + lane of word 16 + lane / 2: whole sectors per store
                const uint32_t wa = __shfl_sync(0xFFFFFFFFu, w, lane >> 1), wb = __shfl_sync(0xFFFFFFFFu, w, 16 + (lane >> 1));
                uint4 *p = dst + (int64_t)y * (W / 16);
                p[0] = spread16((wa >> sh) & 0xFFFFu);
                p[32] = spread16((wb >> sh) & 0xFFFFu);
            }
            if (rc) {
                int n = __reduce_add_sync(0xFFFFFFFFu, __popc(w));
                if (lane == 0) rc[y] = n;
            }
        }
    }
    if (!PACKED_IN && status) {
        odd = __reduce_or_sync(0xFFFFFFFFu, odd);
        if (lane == 0 && odd) atomicOr(status, 1);
    }
}

static int g_clean_bands32 = 4;          // row bands per 1024-wide frame (svb_debug_clean_bands)
extern "C" int svb_debug_clean_bands(int bands) {
    if (bands < 0 || bands > 16) return SVB_E_ARG;
    g_clean_bands32 = bands;             // 0 = the general kernel
    return SVB_OK;
}

template <bool PACKED_IN, int THREADS>
static int launch_rows32(const uint8_t *road, const uint32_t *road_bits, const uint32_t *tbits, int F, int H, int bands,
                         uint8_t *cleaned, uint32_t *cleaned_bits, int32_t *rowcnt, int32_t *status, cudaStream_t st) {
    int band = (H + bands - 1) / bands;
    size_t smem = (size_t)2 * (band + 2 * CL_HALO + C32_TOP + C32_BOT) * C32_WW * sizeof(uint32_t);
    static bool once = false;
    if (!once) {
        cudaFuncSetAttribute(clean_rows32_kernel<PACKED_IN, THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        once = true;
    }
    dim3 grid((H + band - 1) / band, F);
    clean_rows32_kernel<PACKED_IN, THREADS><<<grid, THREADS, smem, st>>>(road, road_bits, tbits, H, band, cleaned, cleaned_bits, rowcnt, status);
    return SVB_OK;
}

static CleanGeom make_clean_geom(int H, int W, const void *img, const void *out) {
    CleanGeom g;
    g.H = H; g.W = W; g.WW = (W + 31) / 32;
    int room = CL_PLANE_WORDS / g.WW - 2 * CL_HALO;
    int band = room < 32 ? 32 : (room > 256 ? 256 : room);
    int bands = (H + band - 1) / band;
    g.band = (H + bands - 1) / bands;               // even split
    g.vec = (W % 16 == 0) && (((uintptr_t)img | (uintptr_t)out) % 16 == 0);
    g.tail = (W & 31) ? ((1u << (W & 31)) - 1u) : ~0u;
    return g;
}

static int launch_clean(const uint8_t *road, const uint32_t *road_bits, const uint32_t *tbits, int F, int H, int W, uint8_t *cleaned,
                        uint32_t *cleaned_bits, int32_t *rowcnt, int32_t *status, cudaStream_t st) {
    if (!tbits || (!road && !road_bits)) return SVB_E_NULL;
    if (!cleaned && !cleaned_bits) return SVB_E_NULL;
    if (F < 1 || H < 2 || W < 2 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    if (road && (const void *)road == (const void *)cleaned) return SVB_E_ARG;      // bands read each other's halo rows
    if (road_bits && road_bits == cleaned_bits) return SVB_E_ARG;
    CleanGeom g = make_clean_geom(H, W, road, cleaned);
    size_t smem = (size_t)2 * (g.band + 2 * CL_HALO) * g.WW * sizeof(uint32_t);
    dim3 grid((H + g.band - 1) / g.band, F);
    if (status && road) {
        cudaError_t e = cudaMemsetAsync(status, 0, sizeof(int32_t), st);
        if (e != cudaSuccess) return (int)e;
    }
    ProfScope ps(K_CLEAN_BAND, st);
    const bool aligned = (((uintptr_t)road | (uintptr_t)cleaned) % 16) == 0;
    if (W == 1024 && aligned && g_clean_bands32 > 0) {
        int bands = g_clean_bands32;
        while ((H + bands - 1) / bands + 2 * CL_HALO + C32_TOP + C32_BOT > 200 * 1024 / (2 * C32_WW * 4)) ++bands;
        if ((H + bands - 1) / bands > 160) {
            if (road) launch_rows32<false, 512>(road, nullptr, tbits, F, H, bands, cleaned, cleaned_bits, rowcnt, status, st);
            else launch_rows32<true, 512>(nullptr, road_bits, tbits, F, H, bands, cleaned, cleaned_bits, rowcnt, status, st);
        } else {
            if (road) launch_rows32<false, 256>(road, nullptr, tbits, F, H, bands, cleaned, cleaned_bits, rowcnt, status, st);
            else launch_rows32<true, 256>(nullptr, road_bits, tbits, F, H, bands, cleaned, cleaned_bits, rowcnt, status, st);
        }
        SVB_LAUNCH_CHECK();
        return SVB_OK;
    }
    if (road) {
        static bool once = false;
        if (!once) { cudaFuncSetAttribute(clean_band_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024); once = true; }
        clean_band_kernel<false><<<grid, CL_THREADS, smem, st>>>(road, nullptr, tbits, g, cleaned, cleaned_bits, rowcnt, status);
    } else {
        static bool once = false;
        if (!once) { cudaFuncSetAttribute(clean_band_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024); once = true; }
        clean_band_kernel<true><<<grid, CL_THREADS, smem, st>>>(nullptr, road_bits, tbits, g, cleaned, cleaned_bits, rowcnt, status);
    }
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int64_t svb_bits_words(int H, int W) {
    if (H < 1 || W < 1) return 0;
    return (int64_t)H * ((W + 31) / 32);
}

extern "C" int svb_clean_road(const uint8_t *road, const uint32_t *threshold_bits, int F, int H, int W, uint8_t *cleaned,
                              uint32_t *cleaned_bits, int32_t *rowcnt, int32_t *status, void *stream) {
    if (!road) return SVB_E_NULL;
    return launch_clean(road, nullptr, threshold_bits, F, H, W, cleaned, cleaned_bits, rowcnt, status, (cudaStream_t)stream);
}

extern "C" int svb_clean_road_bits(const uint32_t *road_bits, const uint32_t *threshold_bits, int F, int H, int W, uint8_t *cleaned,
                                   uint32_t *cleaned_bits, int32_t *rowcnt, void *stream) {
    if (!road_bits) return SVB_E_NULL;
    return launch_clean(nullptr, road_bits, threshold_bits, F, H, W, cleaned, cleaned_bits, rowcnt, nullptr, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ byte mask -> bit rows (+ row counts)
__global__ void __launch_bounds__(CL_THREADS)
pack_mask_kernel(const uint8_t *__restrict__ img, int64_t rows, int W, int WW, int vec, uint32_t *__restrict__ bits, int32_t *__restrict__ rowcnt) {
    int64_t total = rows * WW, padded = (total + 31) & ~(int64_t)31;
    const bool warp_rows = (WW & 31) == 0;
    for (int64_t it = (int64_t)blockIdx.x * CL_THREADS + threadIdx.x; it < padded; it += (int64_t)gridDim.x * CL_THREADS) {
        uint32_t w = 0, odd = 0;
        int64_t y = 0;
        if (it < total) {
            y = it / WW;
            int c = (int)(it - y * WW);
            w = load_word<true>(img + y * W, c * 32, W, vec, odd);
            bits[it] = w;
        }
        if (rowcnt) {
            int n = __popc(w);
            if (warp_rows) {
                n = __reduce_add_sync(0xFFFFFFFFu, n);
                if ((threadIdx.x & 31) == 0 && n) atomicAdd(&rowcnt[y], n);
            } else if (n) {
                atomicAdd(&rowcnt[y], n);
            }
        }
    }
}

extern "C" int svb_pack_mask(const uint8_t *img, int F, int H, int W, uint32_t *bits, int32_t *rowcnt, void *stream) {
    if (!img || !bits) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    int WW = (W + 31) / 32;
    int64_t rows = (int64_t)F * H, total = rows * WW;
    if (rowcnt) {
        cudaError_t e = cudaMemsetAsync(rowcnt, 0, sizeof(int32_t) * rows, st);
        if (e != cudaSuccess) return (int)e;
    }
    int vec = (W % 16 == 0) && ((uintptr_t)img % 16 == 0);
    int64_t blocks = (total + CL_THREADS - 1) / CL_THREADS;
    if (blocks > 148 * 16) blocks = 148 * 16;
    ProfScope ps(K_PACK_MASK, st);
    pack_mask_kernel<<<(int)blocks, CL_THREADS, 0, st>>>(img, rows, W, WW, vec, bits, rowcnt);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ bit rows -> (x,y) list, row-major
#define MP_THREADS 256
#define MP_ROWS 64              // rows per CTA, eight per warp, two per lane in the row scan

__global__ void __launch_bounds__(MP_THREADS)
mask_points_kernel(const uint32_t *__restrict__ bits, const int32_t *__restrict__ rowcnt, int H, int WW, int cap,
                   int32_t *__restrict__ pts, int32_t *__restrict__ counts) {
    const int f = blockIdx.y, row0 = blockIdx.x * MP_ROWS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int32_t *rc = rowcnt + (int64_t)f * H;
    __shared__ int s_part[MP_THREADS / 32], s_all[MP_THREADS / 32];
    __shared__ int s_row[MP_ROWS], s_cnt[MP_ROWS];
    // points before this CTA's first row, and (first CTA of the frame) the frame's total
    int before = 0, all = 0;
    const int upto = blockIdx.x == 0 ? H : row0;
    for (int y = threadIdx.x; y < upto; y += MP_THREADS) {
        int n = rc[y];
        all += n;
        if (y < row0) before += n;
    }
    before = __reduce_add_sync(0xFFFFFFFFu, before);
    all = __reduce_add_sync(0xFFFFFFFFu, all);
    if (lane == 0) { s_part[warp] = before; s_all[warp] = all; }
    __syncthreads();
    if (warp == 0) {
        int b = lane < MP_THREADS / 32 ? s_part[lane] : 0;
        int a = lane < MP_THREADS / 32 ? s_all[lane] : 0;
        b = __reduce_add_sync(0xFFFFFFFFu, b);
        a = __reduce_add_sync(0xFFFFFFFFu, a);
        // exclusive scan of the CTA's row counts, two rows per lane
        int ya = row0 + 2 * lane;
        int na = ya < H ? rc[ya] : 0, nb = ya + 1 < H ? rc[ya + 1] : 0;
        int n = na + nb, inc = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) inc += v;
        }
        s_row[2 * lane] = b + inc - n;
        s_row[2 * lane + 1] = b + inc - n + na;
        s_cnt[2 * lane] = na;
        s_cnt[2 * lane + 1] = nb;
        if (blockIdx.x == 0 && lane == 0 && counts) counts[f] = a;
    }
    __syncthreads();
    int2 *out = reinterpret_cast<int2 *>(pts) + (int64_t)f * cap;
    for (int r = warp; r < MP_ROWS; r += MP_THREADS / 32) {
        int y = row0 + r;
        if (y >= H) break;
        if (s_cnt[r] == 0) continue;                                 // most rows of a road mask are empty
        int base = s_row[r];
        const uint32_t *row = bits + ((int64_t)f * H + y) * WW;
        for (int c0 = 0; c0 < WW; c0 += 32) {
            uint32_t w = c0 + lane < WW ? __ldg(row + c0 + lane) : 0u;
            uint32_t live = __ballot_sync(0xFFFFFFFFu, w != 0);
            if (live == 0) continue;
            int n = __popc(w), inc = n;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int v = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                if (lane >= o) inc += v;
            }
            int excl = inc - n;
            const uint32_t below = (1u << lane) - 1u;
            while (live) {
                int j = __ffs(live) - 1;
                live &= live - 1;
                uint32_t wj = __shfl_sync(0xFFFFFFFFu, w, j);
                int oj = __shfl_sync(0xFFFFFFFFu, excl, j);
                if ((wj >> lane) & 1u) {
                    int idx = base + oj + __popc(wj & below);
                    if (idx < cap) out[idx] = make_int2((c0 + j) * 32 + lane, y);
                }
            }
            base += __shfl_sync(0xFFFFFFFFu, inc, 31);
        }
    }
}

extern "C" int svb_mask_points(const uint32_t *bits, const int32_t *rowcnt, int F, int H, int W, int32_t cap, int32_t *pts,
                               int32_t *counts, void *stream) {
    if (!bits || !rowcnt || (!pts && cap > 0)) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    if (cap < 0) return SVB_E_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid((H + MP_ROWS - 1) / MP_ROWS, F);
    ProfScope ps(K_MASK_POINTS, st);
    mask_points_kernel<<<grid, MP_THREADS, 0, st>>>(bits, rowcnt, H, (W + 31) / 32, cap, pts, counts);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
