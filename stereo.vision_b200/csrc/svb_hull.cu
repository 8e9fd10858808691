// the obstacle stage (stereovision.py:170-200) on the cleaned road image held as bit rows.
//
//   roadHull      = cv2.convexHull(road points)                    -> strict hull, OpenCV's orientation
//   hullMask      = drawContours(cleaned.copy(), [hull], 0, 255, -100) = cleaned OR fill(hull)
//   obstacleImage = NOT cleaned, kept where hullMask != 0          = fill(hull) AND NOT cleaned
//   centre        = int(minEnclosingCircle(hull).centre)           (functions.py:418-421)
//
// fill(hull): OpenCV draws every edge with its 8-connected line iterator and scan-fills the interior; for a convex
// polygon the union is, row by row, the span between the leftmost and the rightmost edge pixel (pinned against cv2
// on random hulls, segments and points: oracle/make_golden.py:obstacle_fixture).
//
// Hull kernel, one CTA per frame.  The hull of a mask is the hull of its rows' first and last pixels, so one warp
// gift-wraps the right side down the last pixels and another the left side back up the first pixels (each step is
// a warp-wide arg-max over the rows ahead); the other warps then turn every hull edge into per-row [lo, hi] spans
// with the line iterator's closed form while one lane finds the enclosing circle (incremental algorithm in float64
// with explicit roundings, bit for bit the oracle's; OpenCV works in float32 -- same int() centre unless a
// coordinate lies within ~1e-4 of an integer).  Masks kernel: spans + bit rows -> the two byte images.
#include "svb_common.cuh"
#include "svb_prof.cuh"
#include <math.h>

#define HL_THREADS 384
#define HL_WARPS (HL_THREADS / 32)

__device__ __forceinline__ int hl_cross(int2 o, int2 a, int2 b) {      // |coordinates| < 2048: fits int32
    return (a.x - o.x) * (b.y - o.y) - (a.y - o.y) * (b.x - o.x);
}
// 4 bits -> 4 bytes of 0 / 255 (as in svb_clean.cu)
__device__ __forceinline__ uint32_t hl_signs(uint32_t v) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %1, 0xBA98;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ uint4 hl_spread16(uint32_t h) {
    const uint32_t g = h >> 8;
    return make_uint4(hl_signs((h & 0xFu) * 0x10204080u), hl_signs((h & 0xF0u) * 0x01020408u),
                      hl_signs((g & 0xFu) * 0x10204080u), hl_signs((g & 0xF0u) * 0x01020408u));
}
__device__ __forceinline__ void hl_store_word(uint8_t *row, int x0, int W, bool vec, uint32_t w) {
    if (vec && x0 + 32 <= W) {
        uint4 *p = reinterpret_cast<uint4 *>(row + x0);
        p[0] = hl_spread16(w & 0xFFFFu);
        p[1] = hl_spread16(w >> 16);
        return;
    }
    const int n = min(32, W - x0);
    for (int i = 0; i < n; ++i) row[x0 + i] = ((w >> i) & 1u) ? 255 : 0;
}
__device__ __forceinline__ int hl_floor_div(int a, int b) {        // b > 0
    const int q = a / b;
    return q - ((a % b != 0) && (a < 0));
}
// float64 squared distance with explicit roundings (no FMA contraction): the oracle's dx*dx + dy*dy bit for bit
__device__ __forceinline__ double hl_dist(double ax, double ay, double bx, double by) {
    const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by);
    return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
}

__global__ void __launch_bounds__(HL_THREADS, 3)
road_hull_kernel(const uint32_t *__restrict__ bits, int H, int W, int cap, int32_t *__restrict__ hull_xy,
                 int32_t *__restrict__ hull_n, int32_t *__restrict__ centre_xy, uint32_t *__restrict__ spans) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    // spans lo/hi int32 [H] each; the two sides of the hull int2 [H] each; the hull int2 [2 H]
    int *s_lo = reinterpret_cast<int *>(s_raw);
    int *s_hi = s_lo + H;
    int2 *s_ma = reinterpret_cast<int2 *>(s_hi + H);
    int2 *s_mb = s_ma + H;
    int2 *s_hull = s_mb + H;
    __shared__ int s_na, s_nb, s_best;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int WW = (W + 31) / 32;
    const uint32_t *fb = bits + (int64_t)f * H * WW;
    constexpr uint32_t FULL = 0xFFFFFFFFu;

    // ---- first and last pixel of every row (four rows in flight per warp)
    for (int y0 = warp * 4; y0 < H; y0 += HL_WARPS * 4) {
        uint32_t w[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) w[r] = (y0 + r < H && lane < WW) ? __ldg(fb + (int64_t)(y0 + r) * WW + lane) : 0u;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int y = y0 + r;
            if (y >= H) break;
            int lo = w[r] ? lane * 32 + __ffs(w[r]) - 1 : 0x7FFFFFFF;
            int hi = w[r] ? lane * 32 + 31 - __clz(w[r]) : -1;
            for (int c = 32 + lane; c < WW; c += 32) {
                const uint32_t wv = __ldg(fb + (int64_t)y * WW + c);
                if (wv) { lo = min(lo, c * 32 + __ffs(wv) - 1); hi = c * 32 + 31 - __clz(wv); }
            }
            lo = __reduce_min_sync(FULL, lo);
            hi = __reduce_max_sync(FULL, hi);
            if (lane == 0) { s_lo[y] = lo; s_hi[y] = hi; }
        }
    }
    __syncthreads();
    // ---- gift wrapping, one warp per side: the right side runs down the rows' last pixels, the left side back up the
    // first pixels.  From the current vertex the next one is the point that has every other remaining point on its
    // left (cross < 0 = further right of the travel direction; collinear: the farther one, so the hull is strict).
    if (warp < 2) {
        const bool right = warp == 0;
        int2 *out = right ? s_ma : s_mb;
        int ytop = 0x7FFFFFFF, ybot = -1;
        for (int y = lane; y < H; y += 32)
            if (s_hi[y] >= 0) { ytop = min(ytop, y); ybot = max(ybot, y); }
        ytop = __reduce_min_sync(FULL, ytop);
        ybot = __reduce_max_sync(FULL, ybot);
        int k = 0;
        if (ybot >= 0) {
            int2 cur = right ? make_int2(s_hi[ytop], ytop) : make_int2(s_lo[ybot], ybot);
            const int yend = right ? ybot : ytop, dir = right ? 1 : -1;
            if (lane == 0) out[0] = cur;
            k = 1;
            while (cur.y != yend) {
                int2 best = make_int2(0, -1);
                const int cnt = (yend - cur.y) * dir;                      // rows still ahead
                for (int i = lane + 1; i <= cnt; i += 32) {
                    const int y = cur.y + dir * i;
                    const int hi = s_hi[y];
                    if (hi < 0) continue;
                    const int2 R = make_int2(right ? hi : s_lo[y], y);
                    if (best.y < 0 || hl_cross(cur, best, R) <= 0) best = R;       // a lane meets its rows nearest first
                }
#pragma unroll
                for (int d = 16; d; d >>= 1) {
                    const int2 O = make_int2(__shfl_xor_sync(FULL, best.x, d), __shfl_xor_sync(FULL, best.y, d));
                    if (O.y < 0) continue;
                    if (best.y < 0) { best = O; continue; }
                    const int c = hl_cross(cur, best, O);
                    if (c < 0 || (c == 0 && (O.y - best.y) * dir > 0)) best = O;
                }
                cur = best;
                if (lane == 0) out[k] = cur;
                ++k;
            }
        }
        if (lane == 0) { if (right) s_na = k; else s_nb = k; }
    }
    __syncthreads();
    // cycle = right side (top -> bottom) + left side (bottom -> top) without the shared end points: positive shoelace
    // sum, OpenCV's orientation; it starts at the lexicographically largest (x, y)
    const int na = s_na, nb = s_nb;
    int b0 = 0, nbe = nb;                                   // the part of the left side that is kept
    if (na > 0 && nb > 0) {
        if (s_mb[0].x == s_ma[na - 1].x && s_mb[0].y == s_ma[na - 1].y) { b0 = 1; --nbe; }
        if (nbe > 0 && s_mb[nb - 1].x == s_ma[0].x && s_mb[nb - 1].y == s_ma[0].y) --nbe;
    }
    const int n = na + nbe;
    auto cyc = [&](int i) { return i < na ? s_ma[i] : s_mb[b0 + i - na]; };
    if (warp == 0) {
        int bi = -1, bx = -1, by = -1;
        for (int i = lane; i < n; i += 32) {
            const int2 p = cyc(i);
            if (p.x > bx || (p.x == bx && p.y > by)) { bx = p.x; by = p.y; bi = i; }
        }
#pragma unroll
        for (int d = 16; d; d >>= 1) {
            const int ox = __shfl_xor_sync(FULL, bx, d), oy = __shfl_xor_sync(FULL, by, d), oi = __shfl_xor_sync(FULL, bi, d);
            if (ox > bx || (ox == bx && oy > by)) { bx = ox; by = oy; bi = oi; }
        }
        if (lane == 0) s_best = max(bi, 0);
    }
    for (int y = tid; y < H; y += HL_THREADS) { s_lo[y] = 0x7FFFFFFF; s_hi[y] = -1; }      // the spans are free again
    __syncthreads();
    {
        const int best = s_best;
        for (int i = tid; i < n; i += HL_THREADS) {
            int j = i + best;
            if (j >= n) j -= n;
            const int2 p = cyc(j);
            s_hull[i] = p;
            if (hull_xy && i < cap) {
                hull_xy[((int64_t)f * cap + i) * 2] = p.x;
                hull_xy[((int64_t)f * cap + i) * 2 + 1] = p.y;
            }
        }
        if (tid == 0 && hull_n) hull_n[f] = n;
    }
    __syncthreads();

    if (warp == 0) {
        // ---- centre of the smallest enclosing circle of the hull's vertices while the other warps draw the edges: the
        // incremental algorithm with the warp looking for the next vertex outside the circle 32 at a time (the circle
        // updates, computed by every lane alike, are exactly the sequential algorithm's)
        if (centre_xy) {
            int cx = -1, cy = -1;
            if (n >= 1) {
                double ox = s_hull[0].x, oy = s_hull[0].y, r = 0.0;                    // r = radius squared
                auto outside_from = [&](int start, int end) {                          // first vertex in [start, end) outside
                    for (int base = start; base < end; base += 32) {
                        const int idx = base + lane;
                        bool out = false;
                        if (idx < end) {
                            const int2 p = s_hull[idx];
                            out = !(hl_dist(p.x, p.y, ox, oy) <= __dadd_rn(__dmul_rn(r, 1.0 + 1e-11), 1e-9));
                        }
                        const uint32_t bal = __ballot_sync(FULL, out);
                        if (bal) return base + __ffs(bal) - 1;
                    }
                    return end;
                };
                for (int i = outside_from(0, n); i < n; i = outside_from(i + 1, n)) {
                    const int2 p = s_hull[i];
                    ox = p.x; oy = p.y; r = 0.0;
                    for (int j = outside_from(0, i); j < i; j = outside_from(j + 1, i)) {
                        const int2 q = s_hull[j];
                        ox = ((double)p.x + q.x) / 2; oy = ((double)p.y + q.y) / 2;
                        r = hl_dist(p.x, p.y, ox, oy);
                        for (int k = outside_from(0, j); k < j; k = outside_from(k + 1, j)) {
                            const int2 s = s_hull[k];
                            const double ax = p.x, ay = p.y, bx = q.x, by = q.y, sx = s.x, sy = s.y;
                            auto M = [](double u, double v) { return __dmul_rn(u, v); };
                            auto A = [](double u, double v) { return __dadd_rn(u, v); };
                            const double e0 = __dsub_rn(by, sy), e1 = __dsub_rn(sy, ay), e2 = __dsub_rn(ay, by);
                            const double d = M(2.0, A(A(M(ax, e0), M(bx, e1)), M(sx, e2)));
                            if (d == 0.0) continue;
                            const double a2 = A(M(ax, ax), M(ay, ay)), b2 = A(M(bx, bx), M(by, by)), s2 = A(M(sx, sx), M(sy, sy));
                            ox = __ddiv_rn(A(A(M(a2, e0), M(b2, e1)), M(s2, e2)), d);
                            oy = __ddiv_rn(A(A(M(a2, __dsub_rn(sx, bx)), M(b2, __dsub_rn(ax, sx))), M(s2, __dsub_rn(bx, ax))), d);
                            r = hl_dist(ax, ay, ox, oy);
                        }
                    }
                }
                cx = (int)ox; cy = (int)oy;
            }
            if (lane == 0) { centre_xy[2 * f] = cx; centre_xy[2 * f + 1] = cy; }
        }
    } else if (spans) {
        // ---- every edge as cv::LineIterator draws it (8-connected, left to right), in closed form per row: pixel i of
        // a line with major extent D and minor extent d lies m_i = floor((2 d i + D - 1) / (2 D)) minor steps from its start
        for (int e = warp - 1; e < n; e += HL_WARPS - 1) {
            int2 p1 = s_hull[e], p2 = s_hull[e + 1 < n ? e + 1 : 0];
            int dx = p2.x - p1.x, dy = p2.y - p1.y;
            if (dx < 0) { p1 = p2; dx = -dx; dy = -dy; }
            int ystep = 1;
            if (dy < 0) { dy = -dy; ystep = -1; }
            if (dy > dx) {
                for (int i = lane; i <= dy; i += 32) {
                    const int x = p1.x + (2 * dx * i + dy - 1) / (2 * dy), y = p1.y + ystep * i;
                    atomicMin(&s_lo[y], x); atomicMax(&s_hi[y], x);
                }
            } else if (dy == 0) {
                if (lane == 0) { atomicMin(&s_lo[p1.y], p1.x); atomicMax(&s_hi[p1.y], p1.x + dx); }
            } else {
                // m_i == m  <=>  2 dx m - dx + 1 <= 2 dy i < 2 dx (m + 1) - dx + 1
                for (int m = lane; m <= dy; m += 32) {
                    const int lo_num = 2 * dx * m - dx + 1;
                    const int i0 = max(0, hl_floor_div(lo_num + 2 * dy - 1, 2 * dy));
                    const int i1 = min(dx, hl_floor_div(lo_num + 2 * dx + 2 * dy - 1, 2 * dy) - 1);
                    const int y = p1.y + ystep * m;
                    if (i0 <= i1) { atomicMin(&s_lo[y], p1.x + i0); atomicMax(&s_hi[y], p1.x + i1); }
                }
            }
        }
    }
    if (!spans) return;
    __syncthreads();
    // ---- the rows' spans (lo | hi << 16; hi < lo = empty) for the mask kernel
    for (int y = tid; y < H; y += HL_THREADS) {
        const int lo = s_lo[y], hi = s_hi[y];
        spans[(int64_t)f * H + y] = hi < 0 ? 0x0000FFFFu : ((uint32_t)lo | ((uint32_t)hi << 16));
    }
}

#define HM_THREADS 256
#define HM_ITEMS 8
// hull mask = cleaned OR span, obstacle image = span AND NOT cleaned.  One item = half a bit word = 16 pixels = one
// 16-byte store per output, so that a warp's store covers 512 contiguous bytes (whole 32-byte sectors).
__device__ __forceinline__ void hm_store_half(uint8_t *row, int x0, int W, bool vec, uint32_t h) {
    if (vec) { if (x0 < W) *reinterpret_cast<uint4 *>(row + x0) = hl_spread16(h); return; }
    const int n = min(16, W - x0);
    for (int i = 0; i < n; ++i) row[x0 + i] = ((h >> i) & 1u) ? 255 : 0;
}
__global__ void __launch_bounds__(HM_THREADS)
hull_masks_kernel(const uint32_t *__restrict__ bits, const uint32_t *__restrict__ spans, int H, int W,
                  uint8_t *__restrict__ hull_mask, uint8_t *__restrict__ obstacles) {
    const int f = blockIdx.y, WW = (W + 31) / 32, total = H * WW * 2;
    const uint32_t *fb = bits + (int64_t)f * H * WW;
    const uint32_t *sp = spans + (int64_t)f * H;
    uint8_t *hm = hull_mask ? hull_mask + (int64_t)f * H * W : nullptr;
    uint8_t *ob = obstacles ? obstacles + (int64_t)f * H * W : nullptr;
    const bool vec = (W % 16 == 0) && ((((uintptr_t)hm | (uintptr_t)ob) & 15) == 0);
    const int base = blockIdx.x * HM_THREADS * HM_ITEMS + threadIdx.x;
    uint32_t cl[HM_ITEMS], sv[HM_ITEMS];
#pragma unroll
    for (int u = 0; u < HM_ITEMS; ++u) {
        const int it = base + u * HM_THREADS;
        cl[u] = it < total ? __ldg(fb + (it >> 1)) : 0u;
        sv[u] = it < total ? __ldg(sp + (it >> 1) / WW) : 0x0000FFFFu;
    }
#pragma unroll
    for (int u = 0; u < HM_ITEMS; ++u) {
        const int it = base + u * HM_THREADS;
        if (it >= total) break;
        const int wd = it >> 1, y = wd / WW, c = wd - y * WW, x0 = c * 32 + (it & 1) * 16;
        const int lo = (int)(sv[u] & 0xFFFFu), hi = (int)(sv[u] >> 16);
        uint32_t fill = 0;
        const int a = max(lo, x0), b = min(hi, x0 + 15);
        if (a <= b) fill = (0xFFFFu << (a - x0)) & (0xFFFFu >> (15 - (b - x0)));
        const uint32_t c16 = (cl[u] >> ((it & 1) * 16)) & 0xFFFFu;
        if (x0 >= W) continue;
        if (hm) hm_store_half(hm + (int64_t)y * W, x0, W, vec, c16 | fill);
        if (ob) hm_store_half(ob + (int64_t)y * W, x0, W, vec, fill & ~c16 & 0xFFFFu);
    }
}

extern "C" int64_t svb_road_hull_scratch_bytes(int F, int H) {
    return F < 1 || H < 1 ? 0 : (int64_t)F * H * (int64_t)sizeof(uint32_t);
}

extern "C" int svb_road_hull(const uint32_t *bits, int F, int H, int W, int cap, int32_t *hull_xy, int32_t *hull_n,
                             int32_t *centre_xy, uint8_t *hull_mask, uint8_t *obstacles, void *scratch,
                             int64_t scratch_bytes, void *stream) {
    if (!bits) return SVB_E_NULL;
    const bool masks = hull_mask || obstacles;
    if (!hull_xy && !hull_n && !centre_xy && !masks) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    if (cap < 0 || (hull_xy && cap < 1)) return SVB_E_ARG;
    if (masks && (!scratch || scratch_bytes < svb_road_hull_scratch_bytes(F, H) || ((uintptr_t)scratch & 3))) return SVB_E_SCRATCH;
    const size_t smem = (size_t)H * 40;
    static bool once = false;
    if (!once) { cudaFuncSetAttribute(road_hull_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); once = true; }
    uint32_t *spans = masks ? (uint32_t *)scratch : nullptr;
    {
        ProfScope ps(K_ROAD_HULL, (cudaStream_t)stream);
        road_hull_kernel<<<F, HL_THREADS, smem, (cudaStream_t)stream>>>(bits, H, W, cap, hull_xy, hull_n, centre_xy, spans);
        SVB_LAUNCH_CHECK();
    }
    if (masks) {
        const int WW = (W + 31) / 32;
        ProfScope ps(K_HULL_MASKS, (cudaStream_t)stream);
        dim3 grid((2 * H * WW + HM_THREADS * HM_ITEMS - 1) / (HM_THREADS * HM_ITEMS), F);
        hull_masks_kernel<<<grid, HM_THREADS, 0, (cudaStream_t)stream>>>(bits, spans, H, W, hull_mask, obstacles);
        SVB_LAUNCH_CHECK();
    }
    return SVB_OK;
}
