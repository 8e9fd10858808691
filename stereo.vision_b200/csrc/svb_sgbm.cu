// Matcher stage: functions.disparity, functions.py:104-128, with the module's matcher
// cv2.StereoSGBM_create(0, 128, 21) (functions.py:26): semi-global block matching in OpenCV's MODE_SGBM,
// 3x3 median, cv2.filterSpeckles(.., 0, 4000, 123), scaling to the path's uint8 half-pixel disparity.
//
// OpenCV walks the image once, row by row, with cyclic row buffers; a row depends on the row above through
// three of the five path directions.  Here the five directions are five sets of INDEPENDENT 1-D paths
// (vertical columns, two diagonal families, rows left-to-right, rows right-to-left): a path that enters the
// image through a side border starts from the same all-zero state as one entering through the top row, and
// L_r(p,.) only depends on L_r(p-r,.).  One warp walks one path, lane l holding disparities 4l..4l+3, so
// the d-1 / d+1 neighbours are two shuffles and min_k L_r one REDUX.  Volumes live in HBM as
// [frame][y][x1][d] int16 (x1 = x - D; 256 B per pixel at D = 128):
//   hsum  horizontal block sums of the Birchfield-Tomasi pixel costs          (sgbm_hsum)
//   C     block costs = vertical running sum of hsum                            (sgbm_vsum)
//   E     per direction the excess L_r - C in [0, P2], 4 bits each (aliases hsum) (sgbm_path<dir>)
// The right-to-left kernel rebuilds S = 5 C + excesses with its own L, finds the winner, refines it, scatters the
// right-view disparity for the left-right check and emits the row (sgbm_final).  All integer; results are
// bit-identical to cv2 4.13 (held by the parity tests) as long as no block cost exceeds
// 32767 - P2 (OpenCV's own int16 stores wrap or saturate beyond that depending on its SIMD path).
#include "svb_common.cuh"
#include "svb_prof.cuh"

#define SG_MAXC 32767
#define SG_FULL 0xffffffffu

namespace {

constexpr int SG_D = 128;          // numDisparities of functions.py:26; lane l owns d = 4l..4l+3
constexpr int SG_P1 = 2, SG_P2 = 5, SG_FTZERO = 15, SG_LRDIFF = 1;   // StereoSGBM_create(0,128,21) resolved

// ------------------------------------------------------------------ prefilter
// per pixel one 16-byte record of three words (grad | raw << 16): value, lower and upper end of its half-pixel
// interval (fourth word unused).  recs: [img(2)][F][H][W] uint4; rows of the RIGHT image hold their even pixels
// first, then the odd ones (pixel x at (x & 1) * ceil(W/2) + x / 2): the lanes of sgbm_hsum read every second
// pixel, which makes their 16-byte loads contiguous.
// (gradient through the clip table) | raw << 16 of pixel xx; the first and last column hold the table's centre
__device__ __forceinline__ uint32_t prefilter_px(const uint8_t *row, const uint8_t *rn, const uint8_t *rs, int xx, int W) {
    if (xx <= 0 || xx >= W - 1) return (uint32_t)SG_FTZERO | ((uint32_t)SG_FTZERO << 16);
    int v = ((int)row[xx + 1] - (int)row[xx - 1]) * 2 + (int)rn[xx + 1] - (int)rn[xx - 1] + (int)rs[xx + 1] - (int)rs[xx - 1];
    return (uint32_t)(min(max(v, -SG_FTZERO), SG_FTZERO) + SG_FTZERO) | ((uint32_t)row[xx] << 16);
}
// grid (rows of both images of all frames, chunks of 256 columns); a lane takes its neighbours' values by shuffle,
// the first and last lane of a warp compute the pixel beyond the warp themselves.
__global__ void __launch_bounds__(256) sgbm_prefilter_kernel(const uint8_t *__restrict__ L, const uint8_t *__restrict__ R, int F, int H, int W,
                                                             uint4 *__restrict__ recs) {
    const int lane = threadIdx.x & 31;
    const int64_t r = blockIdx.x;                                   // row index over [img][frame][y]
    const int64_t nrows = (int64_t)F * H;
    const int img = r >= nrows;
    const int64_t fr = img ? r - nrows : r;
    const int y = (int)(fr % H);
    const uint8_t *row = (img ? R : L) + fr * W;
    const uint8_t *rn = y > 0 ? row - W : row, *rs = y < H - 1 ? row + W : row;
    const int x = blockIdx.y * 256 + threadIdx.x;
    uint32_t c = x < W ? prefilter_px(row, rn, rs, x, W) : 0u;
    uint32_t lft = __shfl_up_sync(SG_FULL, c, 1), rgt = __shfl_down_sync(SG_FULL, c, 1);
    if (x >= W) return;
    if (lane == 0 && x > 0) lft = prefilter_px(row, rn, rs, x - 1, W);
    if (lane == 31 && x < W - 1) rgt = prefilter_px(row, rn, rs, x + 1, W);
    // halves of (c + neighbour) / 2: both halves are below 2^9, so one shift serves both after masking the carry-out bit
    uint32_t hl = x > 0 ? ((c + lft) >> 1) & 0x00ff00ffu : c, hr = x < W - 1 ? ((c + rgt) >> 1) & 0x00ff00ffu : c;
    uint32_t lo = __vminu2(c, __vminu2(hl, hr)), hi = __vmaxu2(c, __vmaxu2(hl, hr));
    recs[r * W + (img ? (x & 1) * ((W + 1) >> 1) + (x >> 1) : x)] = make_uint4(c, lo, hi, 0u);
}

// Birchfield-Tomasi cost of the gradient half plus a quarter of the raw half (both 16-bit halves at once).
// Halves are biased by 0x8000 so that one 32-bit subtraction never borrows across them and the differences
// compare as unsigned halves: c0 = max(0, u - v1, v0 - u), c1 = max(0, v - u1, u0 - v), cost = min(c0, c1).
#define SG_BIAS 0x80008000u
// biased minimum of the two one-sided costs, both halves: BIAS + (cost_grad | cost_raw << 16)
__device__ __forceinline__ uint32_t bt_min(uint32_t u, uint32_t ub, uint32_t u0b, uint32_t u1, uint32_t v, uint32_t vb, uint32_t v0b,
                                           uint32_t v1) {
    uint32_t c0 = __vimax3_u16x2(ub - v1, v0b - u, SG_BIAS);
    uint32_t c1 = __vimax3_u16x2(vb - u1, u0b - v, SG_BIAS);
    return __vminu2(c0, c1);
}

// ------------------------------------------------------------------ horizontal block sums
// thread = (row, pair of disparities), 4 rows per CTA; it walks the row once.  The pixel cost of a column is
// computed on the fly from two 16-byte records (the right pixel of d = 2dp + 1 is the one d = 2dp used a column
// earlier, its biased words are carried); a ring of the last 32 cost pairs (one byte each) in shared memory feeds
// the sliding window, whose ends repeat the border columns (window clamped to [0, w1)).
struct RightPix { uint32_t v, vb, v0b, v1; };
__device__ __forceinline__ RightPix right_pix(const uint4 *__restrict__ p) {
    uint4 r = __ldg(p);
    RightPix q = {r.x, r.x | SG_BIAS, r.y | SG_BIAS, r.z};
    return q;
}
// cost pair of a column: cost(d = 2dp) | cost(d = 2dp + 1) << 16, cost = grad part + raw part / 4
__device__ __forceinline__ uint32_t hsum_cost(const uint4 *__restrict__ Lp, const uint4 *__restrict__ Rp, RightPix &pv) {
    uint4 u = __ldg(Lp);
    RightPix q = right_pix(Rp);
    uint32_t ub = u.x | SG_BIAS, u0b = u.y | SG_BIAS;
    uint32_t ma = bt_min(u.x, ub, u0b, u.z, q.v, q.vb, q.v0b, q.v1);
    uint32_t mb = bt_min(u.x, ub, u0b, u.z, pv.v, pv.vb, pv.v0b, pv.v1);
    pv = q;
    uint32_t g = __byte_perm(ma, mb, 0x5410) & 0x7fff7fffu;              // gradient halves of both
    uint32_t r = (__byte_perm(ma, mb, 0x7632) >> 2) & 0x1fff1fffu;       // raw halves of both, quartered
    return g + r;
}
__device__ __forceinline__ uint16_t ring_put(uint32_t pc) { return (uint16_t)__byte_perm(pc, 0, 0x4420); }   // bytes 0 and 2
__device__ __forceinline__ uint32_t ring_get(const uint16_t *slot) { return __byte_perm((uint32_t)*slot, 0, 0x4140); }
__global__ void __launch_bounds__(256) sgbm_hsum_kernel(const uint4 *__restrict__ recs, int F, int H, int W, int SW2,
                                                        uint32_t *__restrict__ hsum) {
    __shared__ uint16_t ring[32][256];
    const int w1 = W - SG_D;
    const int dp = threadIdx.x & 63, tx = threadIdx.x;
    const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 6);
    if (row >= (int64_t)F * H) return;
    const int64_t n = (int64_t)F * H * W;
    const uint4 *Lp = recs + row * W + SG_D;                      // left pixel of column c: Lp[c]
    // right pixel of (c, d = 2dp) is pixel c + D - 2dp: parity of c, index (c >> 1) + D/2 - dp in its half of the row
    const uint4 *Re = recs + n + row * W + SG_D / 2 - dp, *Ro = Re + ((W + 1) >> 1);
#define RP(c) (((c) & 1 ? Ro : Re) + ((c) >> 1))
    uint32_t *out = hsum + row * w1 * 64 + dp;
    RightPix pv = right_pix(Ro - 1);                               // pixel D - 2dp - 1, the partner of d = 2dp + 1 at column 0
    // columns 0 .. SW2: fill the window (the border column counts SW2 + 1 times), first output at c = SW2
    uint32_t pc = hsum_cost(Lp, RP(0), pv);
    const uint32_t pix0 = pc;
    uint32_t s = pc * (uint32_t)(SW2 + 1);                         // halves stay below 2^16: no carry
    ring[0][tx] = ring_put(pc);
    int c = 1;
    for (; c <= SW2; ++c) {
        pc = hsum_cost(Lp + c, RP(c), pv);
        ring[c & 31][tx] = ring_put(pc);
        s += pc;
    }
    out[0] = s;
    // output x1 = c - SW2 after adding column c; the column leaving is x1 - 1 - SW2: the border column while that is
    // negative (the first SW2 + 1 outputs after x1 = 0), a ring slot afterwards
    uint32_t *op = out + 64;                                       // running output pointer (x1 = 1 next)
    for (; c < w1 && c <= 2 * SW2; ++c) {
        s -= pix0;
        pc = hsum_cost(Lp + c, RP(c), pv);
        ring[c & 31][tx] = ring_put(pc);
        s += pc;
        *op = s; op += 64;
    }
#pragma unroll 4
    for (; c < w1; ++c) {
        s -= ring_get(&ring[(c - 2 * SW2 - 1) & 31][tx]);
        pc = hsum_cost(Lp + c, RP(c), pv);
        ring[c & 31][tx] = ring_put(pc);
        s += pc;
        *op = s; op += 64;
    }
    for (; c < w1 + SW2; ++c) {                                    // right border: the last column again
        int lv = c - 2 * SW2 - 1;
        s -= lv >= 0 ? ring_get(&ring[lv & 31][tx]) : pix0;
        s += pc;
        *op = s; op += 64;
    }
#undef RP
}

// ------------------------------------------------------------------ one step of a path
struct PathState { int l0, l1, l2, l3, mn; };

__device__ __forceinline__ void unpack4(uint2 w, int &a, int &b, int &c, int &d) {
    a = (int)(short)(w.x & 0xffffu); b = (int)(short)(w.x >> 16);
    c = (int)(short)(w.y & 0xffffu); d = (int)(short)(w.y >> 16);
}
// L_r(p,d) = C(p,d) + min(L(d), L(d-1)+P1, L(d+1)+P1, min L + P2) - min L   (see the oracle's note on the shift)
__device__ __forceinline__ void path_step(PathState &s, uint2 cw, int lane) {
    int c0, c1, c2, c3;
    unpack4(cw, c0, c1, c2, c3);
    int left = __shfl_up_sync(SG_FULL, s.l3, 1);
    int right = __shfl_down_sync(SG_FULL, s.l0, 1);
    if (lane == 0) left = SG_MAXC;
    if (lane == 31) right = SG_MAXC;
    int cap = s.mn + SG_P2;
    int n0 = c0 + min(min(s.l0, cap), min(left, s.l1) + SG_P1) - s.mn;
    int n1 = c1 + min(min(s.l1, cap), min(s.l0, s.l2) + SG_P1) - s.mn;
    int n2 = c2 + min(min(s.l2, cap), min(s.l1, s.l3) + SG_P1) - s.mn;
    int n3 = c3 + min(min(s.l3, cap), min(s.l2, right) + SG_P1) - s.mn;
    int m = __reduce_min_sync(SG_FULL, min(min(n0, n1), min(n2, n3)));
    s.l0 = n0; s.l1 = n1; s.l2 = n2; s.l3 = n3; s.mn = m;
}

// L_r - C lies in [0, P2] (the minimum term is at least min L and at most min L + P2), so a direction does not
// have to accumulate a 16-bit S volume: it leaves its four 4-bit excesses per lane as one uint16 (a quarter of a
// volume, written once, never read back by another direction) and the final kernel rebuilds
// S = 5 C + sum of the excesses.  Saturating the total once equals OpenCV's chain of saturating adds because all
// terms have one sign.
__device__ __forceinline__ uint32_t excess4(const PathState &s, uint2 cw) {
    int c0, c1, c2, c3;
    unpack4(cw, c0, c1, c2, c3);
    return (uint32_t)(s.l0 - c0) | ((uint32_t)(s.l1 - c1) << 4) | ((uint32_t)(s.l2 - c2) << 8) | ((uint32_t)(s.l3 - c3) << 12);
}
// DIR 0: rows, left to right.  DIR 1: from (x-1, y-1).  DIR 2: from (x, y-1).  DIR 3: from (x+1, y-1).
// One warp per path; 8 warps per CTA.  E: this direction's excess plane [frame][y][x1][lane] uint16.
template <int DIR>
__global__ void __launch_bounds__(256) sgbm_path_kernel(const uint2 *__restrict__ C, uint16_t *__restrict__ E, int F, int H, int w1) {
    const int lane = threadIdx.x & 31;
    const int npaths = DIR == 0 ? H : DIR == 2 ? w1 : w1 + H - 1;
    int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (wid >= (int64_t)F * npaths) return;
    int f = (int)(wid / npaths), p = (int)(wid - (int64_t)f * npaths);
    int x, y, dx, dy;
    if (DIR == 0) { x = 0; y = p; dx = 1; dy = 0; }
    else if (DIR == 2) { x = p; y = 0; dx = 0; dy = 1; }
    else if (DIR == 1) { dx = 1; dy = 1; if (p < w1) { x = p; y = 0; } else { x = 0; y = p - w1 + 1; } }
    else { dx = -1; dy = 1; if (p < w1) { x = p; y = 0; } else { x = w1 - 1; y = p - w1 + 1; } }
    int len;
    if (DIR == 0) len = w1;
    else if (DIR == 2) len = H;
    else if (DIR == 1) len = min(w1 - x, H - y);
    else len = min(x + 1, H - y);
    const int64_t stride = ((int64_t)dy * w1 + dx) * 32;                     // in lanes (4 disparities each)
    int64_t idx = (((int64_t)f * H + y) * w1 + x) * 32 + lane;
    PathState s = {0, 0, 0, 0, 0};
    int i = 0;
    // eight steps per group; the loads of the next group are issued before the current one is consumed
    uint2 c[8], cn[8];
    if (len >= 8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) c[k] = __ldg(C + idx + k * stride);
    }
    for (; i + 8 <= len; i += 8) {
        if (i + 16 <= len) {
#pragma unroll
            for (int k = 0; k < 8; ++k) cn[k] = __ldg(C + idx + (8 + k) * stride);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            path_step(s, c[k], lane);
            E[idx + k * stride] = (uint16_t)excess4(s, c[k]);
        }
        idx += 8 * stride;
#pragma unroll
        for (int k = 0; k < 8; ++k) c[k] = cn[k];
    }
    for (; i < len; ++i) {
        uint2 c = __ldg(C + idx);
        path_step(s, c, lane);
        E[idx] = (uint16_t)excess4(s, c);
        idx += stride;
    }
}

// ------------------------------------------------------------------ right-to-left sweep + winner + checks
// one warp per row, 4 warps per CTA.  Per step the warp only finds (min S, its first d, the two neighbours of the
// minimum); lane (x1 & 31) keeps that result and every 32 steps the lanes finish their columns in parallel:
// sub-pixel division, disp1, and the right-view entry disp2[x - d] = d where min S is smallest (first visited,
// i.e. largest x, on ties) as one shared-memory atomicMin on (min S, -x).  Per warp: W x (uint32 + int16).
// ------------------------------------------------------------------ block cost + vertical direction in one walk
// One warp per column (lane = 4 disparities) walks down the rows: the block cost is the running sum
// C(y) = C(y-1) + hsum(min(y+SH2, H-1)) - hsum(max(y-SH2-1, 0)) (int16 wrap-around like OpenCV's CostType stores),
// written once for the other directions, and feeds the vertical path (DIR 2) on the spot: C is not read back for
// it.  The row that leaves the window was read 2*SH2+1 rows earlier: it waits in a per-warp ring in shared memory
// (lane-private slots, no synchronisation), so hsum is read from memory exactly once.
// *status (may be NULL) gets bit 0 when a block cost exceeds 32767 - P2: the int16 cost stores of OpenCV wrap or
// saturate there depending on its SIMD path, so beyond that point no output is "the reference's".
__global__ void __launch_bounds__(256) sgbm_vsum_v_kernel(const uint2 *__restrict__ hsum, int F, int H, int w1, int SH2,
                                                          uint2 *__restrict__ C, uint16_t *__restrict__ E, int32_t *__restrict__ status) {
    extern __shared__ uint2 vring[];                               // [warp of the CTA][slot][lane]
    const int RING = 2 * SH2 + 2;
    const int lane = threadIdx.x & 31;
    int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (wid >= (int64_t)F * w1) return;
    uint2 *ring = vring + (size_t)(threadIdx.x >> 5) * RING * 32 + lane;      // slot k: ring[k * 32]
    const int f = (int)(wid / w1), x = (int)(wid - (int64_t)f * w1);
    const int64_t rs = (int64_t)w1 * 32;                           // row stride in lanes
    const int64_t base = ((int64_t)f * H * w1 + x) * 32 + lane;    // row 0 of the column
    // rows 0 .. SH2 (clamped to the image): row 0 counts SH2 + 1 times
    const uint2 h0 = __ldg(hsum + base);
    uint2 hlast = h0;
    ring[0] = h0;
    uint2 acc = make_uint2(0, 0);
    for (int k = 0; k <= SH2; ++k) acc = make_uint2(__vadd2(acc.x, h0.x), __vadd2(acc.y, h0.y));
    for (int k = 1; k <= SH2; ++k) {
        if (k <= H - 1) { hlast = __ldg(hsum + base + (int64_t)k * rs); ring[k * 32] = hlast; }
        acc.x = __vadd2(acc.x, hlast.x); acc.y = __vadd2(acc.y, hlast.y);
    }
    PathState s = {0, 0, 0, 0, 0};
    const uint32_t lim = (uint32_t)(SG_MAXC - SG_P2) * 0x00010001u;      // per 16-bit half
    uint32_t over = __vcmpgtu2(acc.x, lim) | __vcmpgtu2(acc.y, lim);
    C[base] = acc;
    path_step(s, acc, lane);
    E[base] = (uint16_t)excess4(s, acc);
    int ni = (SH2 + 1) % RING;                                     // slot of the row entering at y = 1
    int oi = 0;                                                    // slot of row y - SH2 - 1 once that is >= 0
    auto advance = [&](int y, uint2 a) {                           // a = hsum row y + SH2 (unused past the image)
        if (y + SH2 <= H - 1) { hlast = a; ring[ni * 32] = a; }
        ni = ni + 1 == RING ? 0 : ni + 1;
        uint2 b = h0;
        if (y - SH2 - 1 >= 0) { b = ring[oi * 32]; oi = oi + 1 == RING ? 0 : oi + 1; }
        acc.x = __vsub2(__vadd2(acc.x, hlast.x), b.x); acc.y = __vsub2(__vadd2(acc.y, hlast.y), b.y);
        over |= __vcmpgtu2(acc.x, lim) | __vcmpgtu2(acc.y, lim);
        C[base + (int64_t)y * rs] = acc;
        path_step(s, acc, lane);
        E[base + (int64_t)y * rs] = (uint16_t)excess4(s, acc);
    };
    int y = 1;
    // eight rows per group; the loads of the next group are issued before the current one is consumed
    uint2 a[8], an[8];
    if (y + 8 <= H) {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = __ldg(hsum + base + (int64_t)min(y + k + SH2, H - 1) * rs);
    }
    for (; y + 8 <= H; y += 8) {
        if (y + 16 <= H) {
#pragma unroll
            for (int k = 0; k < 8; ++k) an[k] = __ldg(hsum + base + (int64_t)min(y + 8 + k + SH2, H - 1) * rs);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) advance(y + k, a[k]);
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = an[k];
    }
    for (; y < H; ++y) advance(y, __ldg(hsum + base + (int64_t)min(y + SH2, H - 1) * rs));
    if (status && __any_sync(0xffffffffu, over != 0u) && lane == 0) atomicOr(status, 1);
}

// nibbles of a uint16 (each at most 10 here) to the bytes of a word
__device__ __forceinline__ uint32_t nib2bytes(uint32_t w) {
    uint32_t t = (w | (w << 8)) & 0x00ff00ffu;
    return (t | (t << 4)) & 0x0f0f0f0fu;
}
__global__ void __launch_bounds__(128) sgbm_final_kernel(const uint2 *__restrict__ C, const uint16_t *__restrict__ E0, const uint16_t *__restrict__ E,
                                                         int64_t plane, int F, int H, int W, int16_t *__restrict__ disp) {
    extern __shared__ uint32_t sh32[];
    const int w1 = W - SG_D;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    int64_t wid = blockIdx.x * 4 + wib;
    if (wid >= (int64_t)F * H) return;
    uint32_t *d2k = sh32 + (size_t)wib * W;                                   // min S << 16 | (0xffff - x)
    int16_t *d1 = (int16_t *)(sh32 + (size_t)4 * W) + (size_t)wib * W;
    for (int x = lane; x < W; x += 32) { d1[x] = -16; d2k[x] = 0xffffffffu; }
    __syncwarp();
    PathState s = {0, 0, 0, 0, 0};
    uint32_t mykey = 0xffffffffu;
    int mylo = 0, myhi = 0;
    const int64_t rowbase = wid * w1 * 32 + lane;
    // loads of one column: C and the pairwise sums of the four excess words (nibbles <= 5 each: no carry)
    auto fetch = [&](int col, uint2 &c, uint32_t &e01, uint32_t &e23) {
        int64_t i = rowbase + (int64_t)col * 32;
        c = __ldg(C + i);
        e01 = (uint32_t)__ldg(E0 + i) + __ldg(E + i);
        e23 = (uint32_t)__ldg(E + plane + i) + __ldg(E + 2 * plane + i);
    };
    auto step = [&](int x1, uint2 c, uint32_t e01, uint32_t e23) {
        path_step(s, c, lane);
        // S = 4 C + the four directions' excesses + this direction's L, saturated once at 32767.  Two disparities per
        // word: with C clamped to 8191 and L to 32747 the sum stays below 2^16 per half and still reaches 32767 exactly
        // when the unclamped sum does (L >= C; all terms are non-negative below the cost limit of this file's header).
        uint32_t eb = nib2bytes(e01) + nib2bytes(e23);
        uint32_t l01 = __vminu2(__byte_perm((uint32_t)s.l0, (uint32_t)s.l1, 0x5410), 0x7feb7febu);
        uint32_t l23 = __vminu2(__byte_perm((uint32_t)s.l2, (uint32_t)s.l3, 0x5410), 0x7feb7febu);
        uint32_t s01 = __vminu2((__vminu2(c.x, 0x1fff1fffu) << 2) + __byte_perm(eb, 0, 0x4140) + l01, 0x7fff7fffu);
        uint32_t s23 = __vminu2((__vminu2(c.y, 0x1fff1fffu) << 2) + __byte_perm(eb, 0, 0x4342) + l23, 0x7fff7fffu);
        // first minimum over ascending d: key = S << 8 | d
        const uint32_t d0 = (uint32_t)(4 * lane);
        uint32_t k0 = ((s01 << 8) & 0x00ffff00u) | d0, k1 = ((s01 >> 8) & 0x00ffff00u) | (d0 + 1);
        uint32_t k2 = ((s23 << 8) & 0x00ffff00u) | (d0 + 2), k3 = ((s23 >> 8) & 0x00ffff00u) | (d0 + 3);
        uint32_t key = __reduce_min_sync(SG_FULL, min(min(k0, k1), min(k2, k3)));
        int lo = (int)(key & 0xffu) - 1, hi = lo + 2;
        uint32_t wlo = (lo & 2) ? s23 : s01, whi = (hi & 2) ? s23 : s01;
        int pick_lo = (int)((lo & 1) ? wlo >> 16 : wlo & 0xffffu), pick_hi = (int)((hi & 1) ? whi >> 16 : whi & 0xffffu);
        int s_lo = __shfl_sync(SG_FULL, pick_lo, (lo >> 2) & 31);
        int s_hi = __shfl_sync(SG_FULL, pick_hi, (hi >> 2) & 31);
        if (lane == (x1 & 31)) { mykey = key; mylo = s_lo; myhi = s_hi; }
        if ((x1 & 31) == 0) {
            // lanes finish columns x1 + lane (those below w1); every S saturated = no winner, the pixel stays invalid
            int minS = (int)(mykey >> 8), best = (int)(mykey & 0xffu);
            int xx1 = x1 + lane;
            if (xx1 < w1 && mykey != 0xffffffffu && minS < SG_MAXC) {
                int x = xx1 + SG_D;
                atomicMin(d2k + (x - best), ((uint32_t)minS << 16) | (uint32_t)(0xffff - x));
                int d;
                if (0 < best && best < SG_D - 1) {
                    int den = max(mylo + myhi - 2 * minS, 1);
                    d = best * 16 + ((mylo - myhi) * 16 + den) / (den * 2);
                } else d = best * 16;
                d1[x] = (int16_t)d;
            }
            mykey = 0xffffffffu;
        }
    };
    constexpr int G = 4;                                    // columns whose loads are in flight while a group is computed
    int x1 = w1 - 1;
    for (int k = w1 % G; k > 0; --k, --x1) {               // leading remainder, one column at a time
        uint2 c; uint32_t e01, e23;
        fetch(x1, c, e01, e23);
        step(x1, c, e01, e23);
    }
    // group A is being computed while the loads of group B are in flight; B's four excess words are only added up
    // when B becomes A, a whole group of steps after they were issued (adding at the load would wait for it)
    uint2 cA[G], cB[G];
    uint32_t pA[G], qA[G];
    uint16_t r0[G], r1[G], r2[G], r3[G];
    // running pointers to column x1 - G of this row (the next group's first column): the group's loads then differ by
    // constant offsets only, and one 64-bit subtraction per pointer and group replaces an address per load
    const uint2 *pc = C + rowbase + (int64_t)(x1 - G) * 32;
    const uint16_t *p0 = E0 + rowbase + (int64_t)(x1 - G) * 32, *p1 = E + rowbase + (int64_t)(x1 - G) * 32;
    const uint16_t *p2 = p1 + plane, *p3 = p2 + plane;
    if (x1 >= 0) {
#pragma unroll
        for (int k = 0; k < G; ++k) fetch(x1 - k, cA[k], pA[k], qA[k]);
    }
    while (x1 >= 0) {
        const bool more = x1 - G >= 0;
        if (more) {
#pragma unroll
            for (int k = 0; k < G; ++k) {
                cB[k] = __ldg(pc - k * 32);
                r0[k] = __ldg(p0 - k * 32); r1[k] = __ldg(p1 - k * 32); r2[k] = __ldg(p2 - k * 32); r3[k] = __ldg(p3 - k * 32);
            }
        }
        pc -= G * 32; p0 -= G * 32; p1 -= G * 32; p2 -= G * 32; p3 -= G * 32;
#pragma unroll
        for (int k = 0; k < G; ++k) step(x1 - k, cA[k], pA[k], qA[k]);
        x1 -= G;
#pragma unroll
        for (int k = 0; k < G; ++k) { cA[k] = cB[k]; pA[k] = (uint32_t)r0[k] + r1[k]; qA[k] = (uint32_t)r2[k] + r3[k]; }
    }
    __syncwarp();
    int16_t *out = disp + wid * W;
    for (int x = lane; x < W; x += 32) {
        int v = d1[x];
        if (x >= SG_D && v != -16) {
            int da = v >> 4, db = (v + 15) >> 4;
            int xa = x - da, xb = x - db;                        // both inside the row: 0 <= d < D <= x
            uint32_t ka = d2k[xa], kb = d2k[xb];
            int ra = (0xffff - (int)(ka & 0xffffu)) - xa, rb = (0xffff - (int)(kb & 0xffffu)) - xb;   // right-view disparities
            if (ka != 0xffffffffu && abs(ra - da) > SG_LRDIFF && kb != 0xffffffffu && abs(rb - db) > SG_LRDIFF) v = -16;
        }
        out[x] = (int16_t)v;
    }
}

// ------------------------------------------------------------------ 3x3 median (replicated border), int16
__device__ __forceinline__ void cswap(int &a, int &b) { int t = min(a, b); b = max(a, b); a = t; }
__global__ void sgbm_median3_kernel(const int16_t *__restrict__ src, int F, int H, int W, int16_t *__restrict__ dst) {
    int64_t n = (int64_t)F * H * W;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int x = (int)(i % W), y = (int)((i / W) % H);
        const int16_t *base = src + (i - (int64_t)y * W - x);
        int v[9];
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx)
                v[(dy + 1) * 3 + dx + 1] = base[(int64_t)min(max(y + dy, 0), H - 1) * W + min(max(x + dx, 0), W - 1)];
        // median-of-9 network
        cswap(v[1], v[2]); cswap(v[4], v[5]); cswap(v[7], v[8]); cswap(v[0], v[1]); cswap(v[3], v[4]); cswap(v[6], v[7]);
        cswap(v[1], v[2]); cswap(v[4], v[5]); cswap(v[7], v[8]); cswap(v[0], v[3]); cswap(v[5], v[8]); cswap(v[4], v[7]);
        cswap(v[3], v[6]); cswap(v[1], v[4]); cswap(v[2], v[5]); cswap(v[4], v[7]); cswap(v[4], v[2]); cswap(v[6], v[4]);
        cswap(v[4], v[2]);
        dst[i] = (int16_t)v[4];
    }
}

// ------------------------------------------------------------------ cv2.filterSpeckles: union-find labelling
// label = index of the component's smallest pixel.  Rows are pre-linked into runs by one warp per row
// (ballot of the run starts), vertical links are merged with atomicMin union-find.
__device__ __forceinline__ bool sp_linked(int a, int b, int newVal, int maxDiff) {
    return a != newVal && b != newVal && abs(a - b) <= maxDiff;
}
__global__ void __launch_bounds__(256) speckle_rows_kernel(const int16_t *__restrict__ img, int F, int H, int W, int newVal, int maxDiff,
                                                           int32_t *__restrict__ label, int32_t *__restrict__ size) {
    const int lane = threadIdx.x & 31;
    int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (row >= (int64_t)F * H) return;
    const int16_t *r = img + row * W;
    int32_t *lab = label + row * W, *sz = size + row * W;
    int base = (int)((row % H) * W);                  // labels are per-frame pixel indices
    int carry = -1;
    for (int x0 = 0; x0 < W; x0 += 32) {
        int x = x0 + lane;
        int v = x < W ? r[x] : newVal;
        int pv = (x > 0 && x < W) ? r[x - 1] : newVal;
        bool valid = x < W && v != newVal;
        bool start = valid && !sp_linked(v, pv, newVal, maxDiff);
        uint32_t starts = __ballot_sync(SG_FULL, start);
        uint32_t m = starts & (0xffffffffu >> (31 - lane));
        int st = m ? x0 + 31 - __clz(m) : carry;
        if (x < W) { lab[x] = valid ? base + st : -1; sz[x] = 0; }
        int last = __shfl_sync(SG_FULL, valid ? st : -1, 31);
        carry = last;
    }
}
__device__ __forceinline__ int sp_find(const int32_t *lab, int i) {
    int p = ((volatile const int32_t *)lab)[i];
    while (p != i) { i = p; p = ((volatile const int32_t *)lab)[i]; }
    return i;
}
__device__ void sp_unite(int32_t *lab, int a, int b) {
    while (true) {
        a = sp_find(lab, a); b = sp_find(lab, b);
        if (a == b) return;
        if (a > b) { int t = a; a = b; b = t; }
        int old = atomicMin(lab + b, a);
        if (old == b) return;
        b = old;
    }
}
__global__ void speckle_merge_kernel(const int16_t *__restrict__ img, int F, int H, int W, int newVal, int maxDiff, int32_t *__restrict__ label) {
    int64_t n = (int64_t)F * H * W, hw = (int64_t)H * W;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int p = (int)(i % hw);
        if (p < W) continue;
        int v = img[i], u = img[i - W];
        if (!sp_linked(v, u, newVal, maxDiff)) continue;
        // skip when the pixel to the left has the same two links (its merge covers this one)
        int x = p % W;
        if (x > 0) {
            int vl = img[i - 1], ul = img[i - W - 1];
            if (sp_linked(v, vl, newVal, maxDiff) && sp_linked(u, ul, newVal, maxDiff) && sp_linked(vl, ul, newVal, maxDiff)) continue;
        }
        sp_unite(label + (i - p), p, p - W);
    }
}
__global__ void speckle_count_kernel(int F, int H, int W, int32_t *__restrict__ label, int32_t *__restrict__ size) {
    int64_t n = (int64_t)F * H * W, hw = (int64_t)H * W;
    const int lane = threadIdx.x & 31;
    int64_t warp0 = ((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5) * 32;
    for (int64_t i0 = warp0; i0 < n; i0 += (int64_t)gridDim.x * blockDim.x) {          // warp-uniform trip count
        int64_t i = i0 + lane;
        long long key = -1 - lane;                      // unique per lane: no match
        if (i < n) {
            int p = (int)(i % hw);
            int32_t *lab = label + (i - p);
            if (lab[p] >= 0) {
                int r = sp_find(lab, p);
                lab[p] = r;                             // compression: the root is an ancestor, concurrent finds stay valid
                key = (i - p) + r;
            }
        }
        uint32_t same = __match_any_sync(SG_FULL, key);
        if (key >= 0 && (__ffs(same) - 1) == lane) atomicAdd(size + key, __popc(same));
    }
}
__global__ void speckle_apply_kernel(int16_t *__restrict__ img, int F, int H, int W, int newVal, int maxSize, const int32_t *__restrict__ label,
                                     const int32_t *__restrict__ size) {
    int64_t n = (int64_t)F * H * W, hw = (int64_t)H * W;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int p = (int)(i % hw);
        int l = label[i];
        if (l < 0) continue;
        if (size[(i - p) + l] <= maxSize) img[i] = (int16_t)newVal;
    }
}

// ------------------------------------------------------------------ functions.py:117-127
__global__ void disparity_scale_kernel(const int16_t *__restrict__ d16, int64_t n, double factor, uint8_t *__restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int v = d16[i];
        v = v > 0 ? v : 0;                              // THRESH_TOZERO
        int s = (v >> 4) & 0xff;                        // (d / 16.).astype(uint8): truncation, then uint8 wrap
        out[i] = (uint8_t)((long long)__dmul_rn((double)s, factor) & 0xff);
    }
}

int grid_for(int64_t n, int threads) {
    int64_t b = (n + threads - 1) / threads;
    return (int)(b < 1 ? 1 : (b > 148 * 32 ? 148 * 32 : b));
}

}  // namespace

// ------------------------------------------------------------------ C ABI

extern "C" int svb_sgbm_supported(int H, int W, int num_disparities, int block) {
    return (num_disparities == SG_D && block >= 1 && block <= 31 && (block & 1) && H >= 1 && H <= SVB_MAX_DIM && W <= SVB_MAX_DIM &&
            W - SG_D > block / 2) ? 1 : 0;
}

// bytes of scratch per frame held at once: 2 images of 16-byte records + the hsum/S volume + the C volume + one int16 image
static int64_t sgbm_frame_bytes(int H, int W) {
    int64_t w1 = W - SG_D;
    int64_t vol = (int64_t)H * w1 * SG_D * 2;
    return 2 * (int64_t)H * W * 16 + 2 * vol + vol / 4 + (int64_t)H * W * 2 + 256;
}
extern "C" int64_t svb_sgbm_scratch_bytes(int F, int H, int W, int num_disparities) {
    if (F < 1 || num_disparities != SG_D || W <= SG_D) return 0;
    return (int64_t)F * sgbm_frame_bytes(H, W) + 256;
}

extern "C" int svb_sgbm_compute(const uint8_t *grayL, const uint8_t *grayR, int F, int H, int W, int num_disparities, int block,
                                int16_t *disp16, int32_t *status, void *scratch, int64_t scratch_bytes, void *stream) {
    if (!grayL || !grayR || !disp16 || !scratch) return SVB_E_NULL;
    if (F < 1) return SVB_E_SHAPE;
    if (!svb_sgbm_supported(H, W, num_disparities, block)) return SVB_E_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t per = sgbm_frame_bytes(H, W);
    int fc = (int)((scratch_bytes - 256) / per);
    if (fc < 1) return SVB_E_SCRATCH;
    if (fc > F) fc = F;
    const int w1 = W - SG_D, SW2 = block / 2;
    char *base = (char *)(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
    for (int f0 = 0; f0 < F; f0 += fc) {
        int n = F - f0 < fc ? F - f0 : fc;
        int64_t vol = (int64_t)n * H * w1 * SG_D * 2;
        uint4 *recs = (uint4 *)base;
        uint32_t *hs = (uint32_t *)(base + 2 * (int64_t)n * H * W * 16);
        uint32_t *Cv = (uint32_t *)((char *)hs + vol);
        int16_t *raw = (int16_t *)((char *)Cv + vol);
        const uint8_t *L = grayL + (int64_t)f0 * H * W, *R = grayR + (int64_t)f0 * H * W;
        const uint2 *C2 = (const uint2 *)Cv;
        uint16_t *E = (uint16_t *)hs;                          // excess planes of DIR 1, 3, 0 (a quarter volume each) over the dead hsum
        uint16_t *E0 = (uint16_t *)((char *)raw + (((int64_t)n * H * W * 2 + 255) & ~(int64_t)255));   // DIR 2's plane: hsum is still live
        const int64_t plane = (int64_t)n * H * w1 * 32;
        const unsigned diag_blocks = (unsigned)(((int64_t)n * (w1 + H - 1) + 7) / 8);
        { ProfScope ps(K_SGBM_PREFILTER, st);
          sgbm_prefilter_kernel<<<dim3((unsigned)(2 * (int64_t)n * H), (W + 255) / 256), 256, 0, st>>>(L, R, n, H, W, recs); }
        { ProfScope ps(K_SGBM_HSUM, st);
          sgbm_hsum_kernel<<<(unsigned)(((int64_t)n * H + 3) / 4), 256, 0, st>>>(recs, n, H, W, SW2, hs); }
        { ProfScope ps(K_SGBM_VSUM, st);
          size_t vs = (size_t)8 * (2 * SW2 + 2) * 32 * sizeof(uint2);
          if (vs > 48 * 1024) {
              cudaError_t e = cudaFuncSetAttribute(sgbm_vsum_v_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vs);
              if (e != cudaSuccess) return (int)e;
          }
          sgbm_vsum_v_kernel<<<(unsigned)(((int64_t)n * w1 + 7) / 8), 256, vs, st>>>((const uint2 *)hs, n, H, w1, SW2, (uint2 *)Cv, E0, status); }
        SVB_LAUNCH_CHECK();
        { ProfScope ps(K_SGBM_PATH_D1, st); sgbm_path_kernel<1><<<diag_blocks, 256, 0, st>>>(C2, E + 0 * plane, n, H, w1); }
        { ProfScope ps(K_SGBM_PATH_D3, st); sgbm_path_kernel<3><<<diag_blocks, 256, 0, st>>>(C2, E + 1 * plane, n, H, w1); }
        { ProfScope ps(K_SGBM_PATH_H, st);
          sgbm_path_kernel<0><<<(unsigned)(((int64_t)n * H + 7) / 8), 256, 0, st>>>(C2, E + 2 * plane, n, H, w1); }
        SVB_LAUNCH_CHECK();
        { ProfScope ps(K_SGBM_FINAL, st);
          if ((size_t)4 * W * 6 > 48 * 1024) {              // wide frames: more than the default dynamic shared memory
              cudaError_t e = cudaFuncSetAttribute(sgbm_final_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * W * 6);
              if (e != cudaSuccess) return (int)e;
          }
          sgbm_final_kernel<<<(unsigned)(((int64_t)n * H + 3) / 4), 128, (size_t)4 * W * 6, st>>>(C2, E0, E, plane, n, H, W, raw); }
        { ProfScope ps(K_SGBM_MEDIAN, st);
          sgbm_median3_kernel<<<grid_for((int64_t)n * H * W, 256), 256, 0, st>>>(raw, n, H, W, disp16 + (int64_t)f0 * H * W); }
        SVB_LAUNCH_CHECK();
    }
    return SVB_OK;
}

extern "C" int64_t svb_speckle_scratch_bytes(int F, int H, int W) { return F < 1 ? 0 : (int64_t)F * H * W * 8 + 256; }

extern "C" int svb_filter_speckles(int16_t *img16, int F, int H, int W, int new_val, int max_size, int max_diff, void *scratch,
                                   int64_t scratch_bytes, void *stream) {
    if (!img16 || !scratch) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1 || (int64_t)H * W >= (1 << 30)) return SVB_E_SHAPE;
    if (scratch_bytes < svb_speckle_scratch_bytes(F, H, W)) return SVB_E_SCRATCH;
    cudaStream_t st = (cudaStream_t)stream;
    int32_t *label = (int32_t *)(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
    int32_t *size = label + (int64_t)F * H * W;
    int64_t n = (int64_t)F * H * W;
    ProfScope ps(K_SPECKLE, st);
    speckle_rows_kernel<<<(unsigned)(((int64_t)F * H + 7) / 8), 256, 0, st>>>(img16, F, H, W, new_val, max_diff, label, size);
    speckle_merge_kernel<<<grid_for(n, 256), 256, 0, st>>>(img16, F, H, W, new_val, max_diff, label);
    speckle_count_kernel<<<grid_for(n, 256), 256, 0, st>>>(F, H, W, label, size);
    speckle_apply_kernel<<<grid_for(n, 256), 256, 0, st>>>(img16, F, H, W, new_val, max_size, label, size);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_disparity_scale(const int16_t *disp16, int64_t n, int max_disparity, uint8_t *out, void *stream) {
    if (!disp16 || !out) return SVB_E_NULL;
    if (n < 1 || max_disparity < 1) return SVB_E_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope ps(K_DISP_SCALE, st);
    disparity_scale_kernel<<<grid_for(n, 256), 256, 0, st>>>(disp16, n, 256.0 / (double)max_disparity, out);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
