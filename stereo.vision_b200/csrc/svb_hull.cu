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
// One CTA per frame.  The hull of a mask is the hull of its rows' first and last pixels (<= 2 H candidates, already
// in (y, x) order), so two threads run the two halves of the monotone chain, one thread per hull edge walks the
// line iterator into per-row [lo, hi] spans in shared memory, the whole CTA writes the two masks, and one thread
// finds the enclosing circle (incremental algorithm, float64; OpenCV works in float32 -- same int() centre unless
// a coordinate lies within ~1e-4 of an integer).
#include "svb_common.cuh"
#include "svb_prof.cuh"
#include <math.h>

#define HL_THREADS 256
#define HL_WARPS (HL_THREADS / 32)

__device__ __forceinline__ long long hl_cross(int2 o, int2 a, int2 b) {
    return (long long)(a.x - o.x) * (b.y - o.y) - (long long)(a.y - o.y) * (b.x - o.x);
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

__global__ void __launch_bounds__(HL_THREADS)
road_hull_kernel(const uint32_t *__restrict__ bits, const int32_t *__restrict__ rowcnt, int H, int W, int cap,
                 int32_t *__restrict__ hull_xy, int32_t *__restrict__ hull_n, int32_t *__restrict__ centre_xy,
                 uint8_t *__restrict__ hull_mask, uint8_t *__restrict__ obstacles) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    // layout (H rows): spans lo/hi int32 [H] each; candidates int2 [2H]; chain stacks int32 [2H + 2] each; hull int2 [2H]
    int *s_lo = reinterpret_cast<int *>(s_raw);
    int *s_hi = s_lo + H;
    int2 *s_c = reinterpret_cast<int2 *>(s_hi + H);
    int *s_sa = reinterpret_cast<int *>(s_c + 2 * H);
    int *s_sb = s_sa + 2 * H + 2;
    int2 *s_hull = reinterpret_cast<int2 *>(s_sb + 2 * H + 2);
    __shared__ int s_nc, s_na, s_nb, s_n;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int WW = (W + 31) / 32;
    const uint32_t *fb = bits + (int64_t)f * H * WW;
    const int32_t *rc = rowcnt + (int64_t)f * H;

    // ---- first and last pixel of every row
    for (int y = warp; y < H; y += HL_WARPS) {
        int lo = 0x7FFFFFFF, hi = -1;
        if (rc[y] != 0) {
            for (int c0 = 0; c0 < WW; c0 += 32) {
                const uint32_t w = c0 + lane < WW ? __ldg(fb + (int64_t)y * WW + c0 + lane) : 0u;
                const uint32_t bal = __ballot_sync(0xffffffffu, w != 0);
                if (bal) {
                    const int jf = __ffs(bal) - 1, jl = 31 - __clz(bal);
                    const uint32_t wf = __shfl_sync(0xffffffffu, w, jf), wl = __shfl_sync(0xffffffffu, w, jl);
                    if (lo == 0x7FFFFFFF) lo = (c0 + jf) * 32 + __ffs(wf) - 1;
                    hi = (c0 + jl) * 32 + 31 - __clz(wl);
                }
            }
        }
        if (lane == 0) { s_lo[y] = lo; s_hi[y] = hi; }
    }
    __syncthreads();
    // ---- candidates in (y, x) order
    if (tid == 0) {
        int n = 0;
        for (int y = 0; y < H; ++y) {
            const int lo = s_lo[y], hi = s_hi[y];
            if (hi >= 0) {
                s_c[n++] = make_int2(lo, y);
                if (hi != lo) s_c[n++] = make_int2(hi, y);
            }
        }
        s_nc = n;
    }
    __syncthreads();
    const int nc = s_nc;
    // ---- the two halves of the monotone chain (strict: collinear points are dropped)
    if (tid == 0) {
        int k = 0;
        for (int i = 0; i < nc; ++i) {
            const int2 p = s_c[i];
            while (k >= 2 && hl_cross(s_c[s_sa[k - 2]], s_c[s_sa[k - 1]], p) <= 0) --k;
            s_sa[k++] = i;
        }
        s_na = k;
    } else if (tid == 32) {
        int k = 0;
        for (int i = nc - 1; i >= 0; --i) {
            const int2 p = s_c[i];
            while (k >= 2 && hl_cross(s_c[s_sb[k - 2]], s_c[s_sb[k - 1]], p) <= 0) --k;
            s_sb[k++] = i;
        }
        s_nb = k;
    }
    // the spans are free again: reset them for the edge walk
    __syncthreads();
    for (int y = tid; y < H; y += HL_THREADS) { s_lo[y] = 0x7FFFFFFF; s_hi[y] = -1; }
    if (tid == 0) {
        // cycle = A[:-1] + B[:-1]; orient like OpenCV (positive shoelace sum), start at the largest (x, y)
        int n = 0;
        if (nc == 1) { s_hull[0] = s_c[0]; n = 1; }
        else if (nc >= 2) {
            for (int i = 0; i + 1 < s_na; ++i) s_hull[n++] = s_c[s_sa[i]];
            for (int i = 0; i + 1 < s_nb; ++i) s_hull[n++] = s_c[s_sb[i]];
        }
        if (n >= 3) {
            long long area2 = 0;
            for (int i = 0; i < n; ++i) {
                const int2 a = s_hull[i], b = s_hull[(i + 1) % n];
                area2 += (long long)a.x * b.y - (long long)b.x * a.y;
            }
            if (area2 < 0)
                for (int i = 0, j = n - 1; i < j; ++i, --j) { const int2 t = s_hull[i]; s_hull[i] = s_hull[j]; s_hull[j] = t; }
        }
        s_n = n;
    }
    __syncthreads();
    const int n = s_n;
    if (n >= 2) {
        // rotate to the canonical start (every thread finds it, the copy goes through the candidate array)
        int best = 0;
        for (int i = 1; i < n; ++i) {
            const int2 a = s_hull[i], b = s_hull[best];
            if (a.x > b.x || (a.x == b.x && a.y > b.y)) best = i;
        }
        for (int i = tid; i < n; i += HL_THREADS) s_c[i] = s_hull[(i + best) % n];
        __syncthreads();
        for (int i = tid; i < n; i += HL_THREADS) s_hull[i] = s_c[i];
        __syncthreads();
    }
    if (hull_xy)
        for (int i = tid; i < min(n, cap); i += HL_THREADS) {
            hull_xy[((int64_t)f * cap + i) * 2] = s_hull[i].x;
            hull_xy[((int64_t)f * cap + i) * 2 + 1] = s_hull[i].y;
        }
    if (tid == 0 && hull_n) hull_n[f] = n;

    // ---- every edge through cv::LineIterator (8-connected, left to right) into the rows' spans
    for (int e = tid; e < n; e += HL_THREADS) {
        int2 p1 = s_hull[e], p2 = s_hull[(e + 1) % n];
        int dx = p2.x - p1.x, dy = p2.y - p1.y;
        if (dx < 0) { p1 = p2; dx = -dx; dy = -dy; }
        int ystep = 1;
        if (dy < 0) { dy = -dy; ystep = -1; }
        int x = p1.x, y = p1.y;
        if (dy > dx) {
            int err = dy - (dx + dx);
            const int plus = dy + dy, minus = -(dx + dx);
            for (int i = 0; i <= dy; ++i) {
                atomicMin(&s_lo[y], x); atomicMax(&s_hi[y], x);
                if (err < 0) { err += minus + plus; ++x; } else err += minus;
                y += ystep;
            }
        } else {
            int err = dx - (dy + dy);
            const int plus = dx + dx, minus = -(dy + dy);
            for (int i = 0; i <= dx; ++i) {
                atomicMin(&s_lo[y], x); atomicMax(&s_hi[y], x);
                if (err < 0) { err += minus + plus; y += ystep; } else err += minus;
                ++x;
            }
        }
    }
    if (n == 1 && tid == 0) { s_lo[s_hull[0].y] = s_hull[0].x; s_hi[s_hull[0].y] = s_hull[0].x; }
    __syncthreads();

    // ---- the two masks
    if (hull_mask || obstacles) {
        uint8_t *hm = hull_mask ? hull_mask + (int64_t)f * H * W : nullptr;
        uint8_t *ob = obstacles ? obstacles + (int64_t)f * H * W : nullptr;
        const bool vec = (W % 16 == 0) && ((((uintptr_t)hm | (uintptr_t)ob) & 15) == 0);
        for (int it = tid; it < H * WW; it += HL_THREADS) {
            const int y = it / WW, c = it - y * WW;
            const int lo = s_lo[y], hi = s_hi[y];
            uint32_t fill = 0;
            const int a = max(lo, c * 32), b = min(hi, c * 32 + 31);
            if (hi >= 0 && a <= b) fill = (0xFFFFFFFFu << (a - c * 32)) & (0xFFFFFFFFu >> (31 - (b - c * 32)));
            const uint32_t cl = __ldg(fb + it);
            if (hm) hl_store_word(hm + (int64_t)y * W, c * 32, W, vec, cl | fill);
            if (ob) hl_store_word(ob + (int64_t)y * W, c * 32, W, vec, fill & ~cl);
        }
    }

    // ---- centre of the smallest enclosing circle of the hull's vertices
    if (centre_xy && tid == 0) {
        int cx = -1, cy = -1;
        if (n >= 1) {
            double ox = s_hull[0].x, oy = s_hull[0].y, r = 0.0;
            auto inside = [&](int2 p) { return hypot((double)p.x - ox, (double)p.y - oy) <= r * (1.0 + 1e-12) + 1e-12; };
            for (int i = 0; i < n; ++i) {
                const int2 p = s_hull[i];
                if (inside(p)) continue;
                ox = p.x; oy = p.y; r = 0.0;
                for (int j = 0; j < i; ++j) {
                    const int2 q = s_hull[j];
                    if (inside(q)) continue;
                    ox = ((double)p.x + q.x) / 2; oy = ((double)p.y + q.y) / 2;
                    r = hypot((double)p.x - ox, (double)p.y - oy);
                    for (int k = 0; k < j; ++k) {
                        const int2 s = s_hull[k];
                        if (inside(s)) continue;
                        const double ax = p.x, ay = p.y, bx = q.x, by = q.y, sx = s.x, sy = s.y;
                        // explicit roundings: no FMA contraction, so the sequence is the oracle's float64 one
                        auto M = [](double u, double v) { return __dmul_rn(u, v); };
                        auto A = [](double u, double v) { return __dadd_rn(u, v); };
                        const double e0 = __dsub_rn(by, sy), e1 = __dsub_rn(sy, ay), e2 = __dsub_rn(ay, by);
                        const double d = M(2.0, A(A(M(ax, e0), M(bx, e1)), M(sx, e2)));
                        if (d == 0.0) continue;
                        const double a2 = A(M(ax, ax), M(ay, ay)), b2 = A(M(bx, bx), M(by, by)), s2 = A(M(sx, sx), M(sy, sy));
                        ox = __ddiv_rn(A(A(M(a2, e0), M(b2, e1)), M(s2, e2)), d);
                        oy = __ddiv_rn(A(A(M(a2, __dsub_rn(sx, bx)), M(b2, __dsub_rn(ax, sx))), M(s2, __dsub_rn(bx, ax))), d);
                        r = hypot(ax - ox, ay - oy);
                    }
                }
            }
            cx = (int)ox; cy = (int)oy;
        }
        centre_xy[2 * f] = cx; centre_xy[2 * f + 1] = cy;
    }
}

extern "C" int svb_road_hull(const uint32_t *bits, const int32_t *rowcnt, int F, int H, int W, int cap, int32_t *hull_xy,
                             int32_t *hull_n, int32_t *centre_xy, uint8_t *hull_mask, uint8_t *obstacles, void *stream) {
    if (!bits || !rowcnt) return SVB_E_NULL;
    if (!hull_xy && !hull_n && !centre_xy && !hull_mask && !obstacles) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    if (cap < 0 || (hull_xy && cap < 1)) return SVB_E_ARG;
    const size_t smem = (size_t)H * 8 + (size_t)H * 16 + (size_t)(2 * H + 2) * 8 + (size_t)H * 16;
    static bool once = false;
    if (!once) { cudaFuncSetAttribute(road_hull_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); once = true; }
    ProfScope ps(K_ROAD_HULL, (cudaStream_t)stream);
    road_hull_kernel<<<F, HL_THREADS, smem, (cudaStream_t)stream>>>(bits, rowcnt, H, W, cap, hull_xy, hull_n, centre_xy, hull_mask, obstacles);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
