// a5-a7: RANSAC plane fit.  One warp scores one hypothesis; persistent blocks (one per SM) walk
// contiguous ranges of (frame, trial) pairs and stage the current frame's masked cloud -- at most
// 4 bytes x lattice sites inside carmask, 128 KB at the reference size -- in shared memory with bulk
// (TMA) copies, so the S gathers of a hypothesis are shared-memory reads; the sample-index stream
// (explicit mode) is read coalesced.
//
// Arithmetic.  The score of a hypothesis is evaluated in pixel coordinates with integer
// determinants (see "Arithmetic of the score" below): no float64 point expansion, division or square
// root per hypothesis, 4 float64 operations per sample.  The "exactly collinear -> draw again" rule
// of functions.py:251-259 (an exact-zero test on float64 points) and the winning plane itself
// (abc = inv([P1;P2;P3]).1, adjugate in fp64, svb_ransac_finalize) still use the path's exact
// float64 expressions (svb_common.cuh).
#include "svb_common.cuh"
#include "svb_prof.cuh"
#include <math.h>

#define RANSAC_WARPS 8
#define RANSAC_THREADS (RANSAC_WARPS * 32)
#define MAX_TRIPLE_ROUNDS 64

// ------------------------------------------------------------------ counter-based draws (throughput mode)
__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__host__ __device__ __forceinline__ uint32_t trial_key(uint64_t seed, int frame, int trial) {
    uint32_t k = mix32((uint32_t)seed ^ 0x9E3779B9u);
    k = mix32(k ^ (uint32_t)(seed >> 32));
    k = mix32(k ^ (uint32_t)frame);
    k = mix32(k ^ (uint32_t)trial ^ 0x85EBCA6Bu);
    return k;
}
// The S sample points of a trial (throughput mode): the first S values below M of a full-period linear
// congruential sequence modulo 2^k >= M,
//   x_{n+1} = (a x_n + c) mod 2^k,   a = 5 (mod 8), c odd, x_0 uniform        (per-trial a, c, x_0 from the key)
// -- a permutation of [0, 2^k) (Hull-Dobell), so the S indices are distinct (random.sample draws without
// replacement, functions.py:286); values >= M (fewer than half) are skipped.  The scorer walks the sequence with one
// thread per trial (thread_walk); svb_ransac_dump_indices writes the same draws out with a warp per trial, lane l
// walking the sub-sequence x_l, x_{l+32}, ... with the 32-step map x -> A32 x + C32 (make_draws) and the accepted draws
// counted round by round, lane by lane -- the same order.  An earlier version drew a systematic sample (start + j * stride, stride from a
// table of 256 primes): the winner's mean error over 240 frames was then distributed differently from the
// reference's draws (two-sample KS p = 8e-17, median 0.231 against 0.271: an arithmetic progression clusters on
// image columns, and the minimum over 600 trials finds the lucky clusters).
// The state x_n is NOT the index: the low j bits of a power-of-two congruential sequence have period 2^j, so every
// run of 2^j states is exactly balanced modulo 2^j -- used directly, each sample was stratified over image columns,
// its mean error had less variance than a simple random sample's, and the winner was measurably BETTER than with
// the reference's draws (K and the angle to the true plane over 3 x 240 frames: every median on one side, KS p down
// to 0.007).  The index is the state passed through the k-bit bijection scramble() (PCG's output step: fold the
// high, long-period bits into the low ones), which keeps the S indices distinct.  tests/test_gpu_replay.py holds
// the result to random.sample's on those observables.
#define SCRAMBLE_MUL 0x9E3779B1u
struct Draws { uint32_t x, A32, C32, mask, sh; };
__device__ __forceinline__ uint32_t scramble(uint32_t x, uint32_t sh, uint32_t mask) {
    x ^= x >> sh;
    x = (x * SCRAMBLE_MUL) & mask;
    return x ^ (x >> sh);
}
__device__ __forceinline__ Draws make_draws(uint32_t key, uint32_t M, int lane) {
    Draws d;
    d.mask = M > 2u ? (0xFFFFFFFFu >> __clz(M - 1u)) : 1u;         // 2^k - 1, 2^k >= M, k >= 1
    d.sh = (33u - (uint32_t)__clz(d.mask)) >> 1;                    // ceil(k / 2)
    uint32_t A = (mix32(key ^ 0x51ED270Bu) & ~7u) | 5u;
    uint32_t C = mix32(key + 0x9E3779B9u) | 1u;
    uint32_t x = mix32(key ^ 0xC2B2AE35u);
    // x_lane = x_0 advanced by `lane` steps, then (A, C) doubled five times: the 32-step map
#pragma unroll
    for (int b = 0; b < 5; ++b) {
        if ((lane >> b) & 1) x = x * A + C;
        C = C * (A + 1u);
        A = A * A;
    }
    d.x = x & d.mask; d.A32 = A; d.C32 = C;
    return d;
}
__device__ __forceinline__ uint32_t uniform_draw(uint32_t key, uint32_t ctr, uint32_t m) {
    return __umulhi(mix32(key ^ mix32(ctr + 0x632BE5ABu)), m);
}

#include "svb_geom.cuh"

// abc = inv([P1;P2;P3]) . (1,1,1)^T = (P2xP3 + P3xP1 + P1xP2) / det      functions.py:267
__device__ __forceinline__ bool plane_of(const P3 &p1, const P3 &p2, const P3 &p3, double &a, double &b, double &c) {
    double c23x = p2.y * p3.z - p2.z * p3.y, c23y = p2.z * p3.x - p2.x * p3.z, c23z = p2.x * p3.y - p2.y * p3.x;
    double c31x = p3.y * p1.z - p3.z * p1.y, c31y = p3.z * p1.x - p3.x * p1.z, c31z = p3.x * p1.y - p3.y * p1.x;
    double c12x = p1.y * p2.z - p1.z * p2.y, c12y = p1.z * p2.x - p1.x * p2.z, c12z = p1.x * p2.y - p1.y * p2.x;
    double det = p1.x * c23x + p1.y * c23y + p1.z * c23z;
    if (det == 0.0 || !isfinite(det)) return false;             // LinAlgError -> trial skipped, :294
    a = (c23x + c31x + c12x) / det;
    b = (c23y + c31y + c12y) / det;
    c = (c23z + c31z + c12z) / det;
    return isfinite(a) && isfinite(b) && isfinite(c);
}

// The triple of trial t: explicit indices, or counter-based draws repeated while exactly collinear.
// Returns false when the trial has no plane (M < S, missing indices, retry budget exhausted).
__device__ __forceinline__ bool trial_triple(const uint32_t *__restrict__ rec, int stride, uint32_t M,
                                             const CamD &cam, const int32_t *__restrict__ triple,
                                             uint32_t key, P3 &p1, P3 &p2, P3 &p3, int idx_out[3]) {
    if (triple) {
        int i1 = triple[0], i2 = triple[1], i3 = triple[2];
        if (i1 < 0 || i2 < 0 || i3 < 0 || (uint32_t)i1 >= M || (uint32_t)i2 >= M || (uint32_t)i3 >= M) return false;
        uint32_t r1 = __ldg(rec + (int64_t)i1 * stride), r2 = __ldg(rec + (int64_t)i2 * stride),
                 r3 = __ldg(rec + (int64_t)i3 * stride);
        p1 = expand_point(cam, r1);
        p2 = expand_point(cam, r2);
        p3 = expand_point(cam, r3);
        idx_out[0] = i1; idx_out[1] = i2; idx_out[2] = i3;
        return !image_collinear(r1, r2, r3);
    }
    for (int round = 0; round < MAX_TRIPLE_ROUNDS; ++round) {
        uint32_t i1 = uniform_draw(key, 3u * round + 0u, M);
        uint32_t i2 = uniform_draw(key, 3u * round + 1u, M);
        uint32_t i3 = uniform_draw(key, 3u * round + 2u, M);
        uint32_t r1 = __ldg(rec + (int64_t)i1 * stride), r2 = __ldg(rec + (int64_t)i2 * stride),
                 r3 = __ldg(rec + (int64_t)i3 * stride);
        p1 = expand_point(cam, r1);
        p2 = expand_point(cam, r2);
        p3 = expand_point(cam, r3);
        if (!exactly_collinear(p1, p2, p3)) {
            idx_out[0] = (int)i1; idx_out[1] = (int)i2; idx_out[2] = (int)i3;
            return !image_collinear(r1, r2, r3);
        }
    }
    return false;
}

// ------------------------------------------------------------------ score: one warp per hypothesis (explicit indices)
// Arithmetic of the score.  With Z = fB/d, X = (x-cw) Z/f, Y = (y-ch) Z/f the plane a X + b Y + c Z = 1
// through three points reads, in pixel coordinates,  al x + be y + ga = d/fB  (al = a/f, be = b/f,
// ga = c - al cw - be ch): a 3x3 system with the INTEGER matrix [x_i y_i 1] and right-hand side d_i/fB.
// Cramer's rule gives  (al, be, ga) = (Na, Nb, Ng) / (fB det)  with the integer determinants
//   det = |x y 1|,  Na = |d y 1|,  Nb = |x d 1|,  Ng = |x y d|       (exact in int32 / float64)
// and for a sample point (x, y, d):   P.abc - 1 = ((Na x + Nb y + Ng)/d - det) / det,
//   mean_j |P_j.abc - 1| / |abc| = fB sum_j |(Na x_j + Nb y_j + Ng)/d_j - det| / (S sqrt(A^2 + B^2 + C^2)),
//   A = f Na, B = f Nb, C = Ng + cw Na + ch Nb                        (det cancels).
// Na x + Nb y + Ng is an exact integer (< 2^35) in float64, so the only roundings per sample are the
// product with the table value 1/d and the sum: ~1e-14 relative on the score, against the reference's
// inv()/dot() float64 chain (functions.py:267-275) which itself is only reproducible to ~1e-12 (SURVEY
// 7.1-4).  No division, square root or float64 point expansion per hypothesis; det == 0 is exactly
// "image-collinear" (the singular case).
// int -> double without the (quarter-rate) conversion instruction: 2^52 + x has x in its low mantissa bits
__device__ __forceinline__ double u2d(uint32_t x) { return __hiloint2double(0x43300000, (int)x) - 4503599627370496.0; }

// Warp-cooperative fetch of the explicit triple of one trial: lanes 0..2 fetch one record each.  Returns false when
// the trial has no plane (bad index, image-collinear triple): the decisions of trial_triple() for explicit indices.
__device__ __forceinline__ bool warp_triple(const uint32_t *__restrict__ rec, int stride, uint32_t M,
                                            const int32_t *__restrict__ triple, int lane,
                                            uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    const int i = lane < 3 ? __ldg(triple + lane) : 0;
    const bool valid = i >= 0 && (uint32_t)i < M;
    if (__any_sync(0xffffffffu, !valid)) return false;
    const uint32_t r = __ldg(rec + (int64_t)i * stride);
    r1 = __shfl_sync(0xffffffffu, r, 0); r2 = __shfl_sync(0xffffffffu, r, 1); r3 = __shfl_sync(0xffffffffu, r, 2);
    return img_det(r1, r2, r3) != 0;
}

// One warp scores one trial with explicit indices (parity mode).  STAGED: the S sample gathers read the frame's records from shared memory
// (s_rec, stride 1); otherwise from global memory (L2).  Writes the mean error (or +inf) to *err_out.
// Per sample: one index load (coalesced), one gather, and  |(Na x + Nb y + Ng)/d - det|  as two 32x32+64 integer multiply-adds whose
// 64-bit result IS the bit pattern of the double 2^52 + 2^40 + (Na x + Nb y + Ng) (the bias and the
// exponent are folded into the addend), one subtraction, one fma with the table value 1/d, one add.
__device__ __forceinline__ long long mad_wide(int a, int b, long long c) {       // a * b + c, signed 32 x 32 + 64
    long long r;
    asm("mad.wide.s32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
    return r;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ double lds_f64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

struct TrialPlane {             // the hypothesis in the integer form the sample loop consumes
    int na, nb;
    long long base;             // Ng + 2^40 with the exponent bits of 2^52 on top
    double det;
};

// |(Na x + Nb y + Ng)/d - det| for the record r
template <bool STAGED>
__device__ __forceinline__ double sample_term(const TrialPlane &h, uint32_t idx, const uint32_t *__restrict__ rec, int stride,
                                              uint32_t s_rec_addr, uint32_t s_invd_addr) {
    const uint32_t r = STAGED ? lds_u32(s_rec_addr + idx * 4u) : __ldg(rec + (int64_t)idx * stride);
    const long long L = mad_wide(h.nb, (int)((r >> 11) & 0x7FFu), mad_wide(h.na, (int)(r & 0x7FFu), h.base));
    const double lin = __longlong_as_double(L) - 4504699138998272.0;             // 2^52 + 2^40
    return fabs(fma(lin, lds_f64(s_invd_addr + ((r >> 19) & 0x7F8u)), -h.det));
}

template <bool STAGED>
__device__ __forceinline__ void score_trial(const uint32_t *__restrict__ rec, int stride, uint32_t s_rec_addr,
                                            uint32_t M, int S, const CamD &cam, double fB_over_S, uint32_t s_invd_addr,
                                            const int32_t *__restrict__ samp, const int32_t *__restrict__ trip,
                                            int lane, double *err_out) {
    uint32_t r1, r2, r3;
    if (!warp_triple(rec, stride, M, trip, lane, r1, r2, r3)) {
        if (lane == 0) *err_out = INFINITY;
        return;
    }
    const int x1 = rec_x(r1), y1 = rec_y(r1), d1 = rec_d(r1);
    const int x2 = rec_x(r2), y2 = rec_y(r2), d2 = rec_d(r2);
    const int x3 = rec_x(r3), y3 = rec_y(r3), d3 = rec_d(r3);
    const int det_i = x1 * (y2 - y3) + x2 * (y3 - y1) + x3 * (y1 - y2);          // != 0 here
    const long long ng_i = (long long)(d1 * (x2 * y3 - x3 * y2)) + (long long)(d2 * (x3 * y1 - x1 * y3)) +
                           (long long)(d3 * (x1 * y2 - x2 * y1));                 // each product < 2^31, sum < 2^33
    TrialPlane h;
    h.na = d1 * (y2 - y3) + d2 * (y3 - y1) + d3 * (y1 - y2);                      // |.| < 2^21
    h.nb = d1 * (x3 - x2) + d2 * (x1 - x3) + d3 * (x2 - x1);
    // |Na x + Nb y + Ng| < 2^35, so 2^40 + it is a positive integer below 2^52: put it in the mantissa of 2^52
    h.base = ng_i + (1ll << 40) + (0x43300000ll << 32);
    h.det = (double)det_i;
    double acc0 = 0.0;
    bool bad = false;
    for (int j = lane; j < S; j += 32) {
        int i = __ldg(samp + j);
        bad |= (i < 0 || (uint32_t)i >= M);
        acc0 += sample_term<STAGED>(h, bad ? 0u : (uint32_t)i, rec, stride, s_rec_addr, s_invd_addr);
    }
    double acc = warp_sum(acc0);
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) {
        const double Na = (double)h.na, Nb = (double)h.nb, Ng = (double)ng_i;
        const double A = cam.f * Na, B = cam.f * Nb, C = Ng + cam.cw * Na + cam.ch * Nb;
        const double e = acc * fB_over_S * rsqrt(A * A + B * B + C * C);
        *err_out = (bad || !(e < INFINITY)) ? INFINITY : e;       // NaN never beats bestError, :291
    }
}

// ------------------------------------------------------------------ score: one THREAD per hypothesis (device draws)
// With a warp per hypothesis the per-hypothesis work -- trial key, triple, the integer determinants, the draw parameters,
// the reduction and the final scaling, ~240 instructions -- is executed by all 32 lanes, and the lanes need a vote and a
// population count per round to agree on which draws count.  With device draws nothing has to be read coalesced, so a thread
// can own a hypothesis: it walks the trial's congruential sequence x_0, x_1, ... itself and keeps the first S values below
// M.  That is the sequence, in the order, that svb_ransac_dump_indices writes out with a warp per trial (lane l walks x_l,
// x_(l+32), ... and accepted draws count round by round, lane by lane): the dumped indices ARE the scorer's sample.
// Explicit indices (parity mode) keep the warp version: their loads want to be coalesced.
template <bool STAGED>
__device__ __forceinline__ uint32_t fetch_rec(uint32_t idx, const uint32_t *__restrict__ rec, int stride, const uint32_t *s_rec) {
    return STAGED ? s_rec[idx] : __ldg(rec + (int64_t)idx * stride);
}
// The sample term of the thread version.  The scorer is bound by instruction issue, the alu pipe first (shifts, masks, compares,
// selects: 61 % active against 19 % for the fma pipe that executes IMAD in the first thread version), so the term is put in the form that
// needs the fewest of those:
//     |(Na x + Nb y + Ng)/d - det| = |W| (1/d),     W = Na x + Nb y + Ng - det d        (an exact integer)
//  * one product with the table value 1/d instead of an fma that cancels: a single rounding, of a relative 2^-53;
//  * NARROW: when |Na| 2047 + |Nb| 2047 + |Ng| + 255 |det| < 2^31 (every hypothesis of a 1024 x 544 image) W fits 32 bits
//    and, with t = r >> 11 = y + 2^11 d and d = r >> 22,
//        W = Ng + Na r + (Nb - 2^11 Na) t - (2^11 Nb + det) d      (mod 2^32),
//    three IMADs on the raw record and its two shifts -- no field is masked out.  2^31 is folded into the constant and the
//    word is the low half of the double 2^52 + 2^31 + W.  Otherwise the same integer from 32 x 32 + 64 multiply-adds;
//  * STAGED: a draw >= M (skipped) reads the SENTINEL record 2^31 that the block keeps at s_rec[M]: its disparity field is
//    512, the table is zero there, its term is exactly 0 and its bit 31 counts the rejection -- a min and a shift-add
//    instead of compare, two selects and a predicated add.  From global memory the term is dropped by a select.
//  * the shared-memory reads are plain indexing: the compiler folds scaling and window base into the address operand.
#define RANSAC_SENTINEL 0x80000000u
// Draws in flight per thread.  From shared memory 4: 109.6 us, 8: 109.6, 12 and 16: 111 (not bound by latency).  A launch
// that stages nothing (clouds too large for shared memory: step 1, 2048 x 1088) gathers from L2 and wants more loads in
// flight; it has its own instantiation so that the staged kernel keeps its 80 registers.  The sum does not depend on the
// depth: even raw draws go to one accumulator, odd ones to the other, in sequence order.
#define RANSAC_UNROLL 8
#define RANSAC_UNROLL_GLOBAL 16
struct NarrowPlane { uint32_t c0, c1, c2, c3; };
struct WidePlane { int na, nb, mdet; long long base; };           // base: Ng + 2^40 with the exponent bits of 2^52 on top
template <bool NARROW>
__device__ __forceinline__ double record_term(const WidePlane &h, const NarrowPlane &n, uint32_t r, const double *s_invd) {
    const uint32_t d = r >> 22;                                                   // bits 30-31 of a cloud record are zero
    double w;
    if (NARROW) {
        const uint32_t v = n.c1 * r + (n.c2 * (r >> 11) + (n.c3 * d + n.c0));
        w = __hiloint2double(0x43300000, (int)v) - 4503601774854144.0;           // 2^52 + 2^31 (I2F.F64 on the idle xu pipe: no faster)
    } else {
        const long long L = mad_wide(h.mdet, (int)(d & 0xFFu), mad_wide(h.nb, (int)((r >> 11) & 0x7FFu), mad_wide(h.na, (int)(r & 0x7FFu), h.base)));
        w = __longlong_as_double(L) - 4504699138998272.0;                        // 2^52 + 2^40
    }
    return w * s_invd[d];                                                         // its magnitude is the term
}
// the first S values below M of the trial's scrambled congruential sequence, summed
template <bool STAGED, bool NARROW, int U>
__device__ __forceinline__ double thread_walk(const WidePlane &h, const NarrowPlane &n, uint32_t key, uint32_t M, int S,
                                              const uint32_t *__restrict__ rec, int stride, const uint32_t *s_rec, const double *s_invd) {
    // The k-bit state lives in the TOP k bits of a word (X = x << (32 - k)), so the products wrap modulo 2^k by themselves:
    // per draw one multiply-add for the state and shift, xor-and, multiply, shift, xor, multiply-high for the index --
    // make_draws' sequence through scramble(), two masks fewer.
    const uint32_t k = M > 2u ? 32u - (uint32_t)__clz(M - 1u) : 1u, up = 32u - k, sh = (k + 1u) >> 1;
    const uint32_t top = 0xFFFFFFFFu << up, down = 1u << k;       // umulhi(w, 2^k) = w >> (32 - k), on the fma pipe
    const uint32_t A = (mix32(key ^ 0x51ED270Bu) & ~7u) | 5u, C = (mix32(key + 0x9E3779B9u) | 1u) << up;
    uint32_t X = mix32(key ^ 0xC2B2AE35u) << up;
    double acc0 = 0.0, acc1 = 0.0;
    int drawn = 0;
    uint32_t rejected = 0;
    // U draws in flight; at most as many more are taken.  (drawn <= 2^k: a full period holds M >= S accepted values; the
    // bound only matters for records that break the format -- bit 31 set -- and then stops the walk instead of hanging)
    while (drawn - (int)rejected + U <= S && (uint32_t)drawn < down) {
        uint32_t is[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t Z = ((X ^ (X >> sh)) & top) * SCRAMBLE_MUL;
            is[u] = __umulhi(Z ^ (Z >> sh), down);
            X = X * A + C;
        }
        double ts[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (STAGED) {
                const uint32_t r = s_rec[min(is[u], M)];          // values >= M are skipped: the sentinel
                rejected += r >> 31;
                ts[u] = record_term<NARROW>(h, n, r, s_invd);
            } else {
                const bool keep = is[u] < M;
                const double t = record_term<NARROW>(h, n, __ldg(rec + (int64_t)min(is[u], M - 1u) * stride), s_invd);
                rejected += keep ? 0u : 1u;
                ts[u] = keep ? t : 0.0;
            }
        }
#pragma unroll
        for (int u = 0; u < U; u += 2) { acc0 += fabs(ts[u]); acc1 += fabs(ts[u + 1]); }
        drawn += U;
    }
    for (int have = drawn - (int)rejected; have < S && (uint32_t)drawn < down; ++drawn) {
        const uint32_t Z = ((X ^ (X >> sh)) & top) * SCRAMBLE_MUL;
        const uint32_t i0 = __umulhi(Z ^ (Z >> sh), down);
        X = X * A + C;
        if (i0 < M) {
            const double t = fabs(record_term<NARROW>(h, n, STAGED ? s_rec[i0] : __ldg(rec + (int64_t)i0 * stride), s_invd));
            if (drawn & 1) acc1 += t; else acc0 += t;
            ++have;
        }
    }
    return acc0 + acc1;
}
template <bool STAGED, int U>
__device__ __forceinline__ double score_trial_thread(const uint32_t *__restrict__ rec, int stride, const uint32_t *s_rec,
                                                     uint32_t M, int S, const CamD &cam, double fB_over_S, const double *s_invd,
                                                     uint32_t key) {
    // the triple: the decisions of warp_triple<false>
    uint32_t r1 = 0, r2 = 0, r3 = 0;
    bool ok = false;
    for (int round = 0; round < MAX_TRIPLE_ROUNDS; ++round) {
        r1 = fetch_rec<STAGED>(uniform_draw(key, 3u * round + 0u, M), rec, stride, s_rec);
        r2 = fetch_rec<STAGED>(uniform_draw(key, 3u * round + 1u, M), rec, stride, s_rec);
        r3 = fetch_rec<STAGED>(uniform_draw(key, 3u * round + 2u, M), rec, stride, s_rec);
        if (img_det(r1, r2, r3) != 0) { ok = true; break; }
        if (!exactly_collinear(expand_point(cam, r1), expand_point(cam, r2), expand_point(cam, r3))) break;   // rare (3.5e-4 of draws)
    }
    if (!ok) return INFINITY;
    const int x1 = rec_x(r1), y1 = rec_y(r1), d1 = rec_d(r1);
    const int x2 = rec_x(r2), y2 = rec_y(r2), d2 = rec_d(r2);
    const int x3 = rec_x(r3), y3 = rec_y(r3), d3 = rec_d(r3);
    const int det_i = x1 * (y2 - y3) + x2 * (y3 - y1) + x3 * (y1 - y2);
    const long long ng_i = (long long)(d1 * (x2 * y3 - x3 * y2)) + (long long)(d2 * (x3 * y1 - x1 * y3)) +
                           (long long)(d3 * (x1 * y2 - x2 * y1));
    WidePlane h;
    h.na = d1 * (y2 - y3) + d2 * (y3 - y1) + d3 * (y1 - y2);
    h.nb = d1 * (x3 - x2) + d2 * (x1 - x3) + d3 * (x2 - x1);
    h.mdet = -det_i;
    h.base = ng_i + (1ll << 40) + (0x43300000ll << 32);
    NarrowPlane n;
    n.c1 = (uint32_t)h.na; n.c2 = (uint32_t)h.nb - ((uint32_t)h.na << 11); n.c3 = 0u - ((uint32_t)h.nb << 11) - (uint32_t)det_i;
    n.c0 = (uint32_t)ng_i + 0x80000000u;
    const bool narrow = (long long)abs(h.na) * 2047 + (long long)abs(h.nb) * 2047 + llabs(ng_i) + (long long)abs(det_i) * 255 < (1ll << 31);
    const double acc = narrow ? thread_walk<STAGED, true, U>(h, n, key, M, S, rec, stride, s_rec, s_invd)
                              : thread_walk<STAGED, false, U>(h, n, key, M, S, rec, stride, s_rec, s_invd);
    const double Na = (double)h.na, Nb = (double)h.nb, Ng = (double)ng_i;
    const double Ap = cam.f * Na, Bp = cam.f * Nb, Cp = Ng + cam.cw * Na + cam.ch * Nb;
    const double e = acc * fB_over_S * rsqrt(Ap * Ap + Bp * Bp + Cp * Cp);
    return e < INFINITY ? e : INFINITY;                           // NaN never beats bestError, :291
}

// ---- bulk (TMA) staging helpers: one elected thread copies global -> shared and an mbarrier counts the bytes
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// Persistent kernel: block b owns the contiguous range [b, b+1) * F*T/gridDim.x of the (frame, trial) pairs,
// i.e. a few consecutive frames.  Entering a frame it stages that frame's masked cloud (<= 4 bytes x max_m)
// in shared memory with bulk copies (TMA), then its warps share it: the S gathers of a hypothesis are
// shared-memory reads (no 32-byte-sector over-fetch from L2).  Frames that do not fit (or whose records
// are strided) gather from global memory instead.
#define RANSAC_SMEM_THREADS 1024
#define RANSAC_BULK_BYTES 32768u
#define RANSAC_PT_THREADS 640                                     // thread-per-hypothesis blocks: one pass for the reference's 600 trials
template <bool PER_THREAD, int UG>                                // PER_THREAD false: explicit indices, a warp per hypothesis; UG: unroll of the global-memory walk
__global__ void __launch_bounds__(PER_THREAD ? RANSAC_PT_THREADS : RANSAC_SMEM_THREADS, 1)
ransac_score_persistent(const uint32_t *__restrict__ mcloud, int stride, int64_t slab_records,
                        const int32_t *__restrict__ counts, int cstride, int ccol,
                        int F, int T, int S, int cap, CamD cam,
                        const int32_t *__restrict__ sample_idx, const int32_t *__restrict__ triple_idx,
                        uint64_t seed, int frame0, double *__restrict__ errors) {
    extern __shared__ __align__(128) uint32_t s_rec[];
    __shared__ double s_invd[1024];                               // 1 / d; zero for d = 0 and above 255 (bits 30-31 of a record, never set by svb_project)
    __shared__ __align__(8) uint64_t s_bar;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = (int)blockDim.x >> 5;
    const int64_t total = (int64_t)F * T;
    const int64_t w0 = total * blockIdx.x / gridDim.x, w1 = total * (blockIdx.x + 1) / gridDim.x;
    for (int i = tid; i < 1024; i += (int)blockDim.x) s_invd[i] = (i > 0 && i < 256) ? __ddiv_rn(1.0, (double)i) : 0.0;
    if (tid == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    const double fB_over_S = cam.fB / (double)S;
    const uint32_t s_rec_addr = smem_u32(s_rec), s_invd_addr = smem_u32(s_invd);
    uint32_t phase = 0;
    for (int64_t w = w0; w < w1;) {
        const int f = (int)(w / T);
        const int tb = (int)(w - (int64_t)f * T);                     // first trial of this segment
        const int te = (int)min((int64_t)T, w1 - (int64_t)f * T);     // one past its last trial
        const int Mi = counts[(int64_t)f * cstride + ccol];
        const uint32_t *rec = mcloud + (int64_t)f * slab_records * stride;
        const bool none = (Mi < S || Mi < 1 || S < 1);                // random.sample raises, :286,294
        const bool staged = !none && Mi <= cap;
        if (staged) {
            const bool bulk = stride == 1 && ((((uintptr_t)rec) & 15) == 0);
            const int nb = bulk ? (Mi & ~3) : 0;                      // records moved by bulk copies
            if (bulk && tid == 0 && nb > 0) {
                const uint32_t bytes = (uint32_t)nb * 4u;
                mbar_expect_tx(&s_bar, bytes);
                for (uint32_t o = 0; o < bytes; o += RANSAC_BULK_BYTES)
                    bulk_g2s((char *)s_rec + o, (const char *)rec + o, min(RANSAC_BULK_BYTES, bytes - o), &s_bar);
            }
            for (int i = nb + tid; i < Mi; i += (int)blockDim.x) s_rec[i] = __ldg(rec + (int64_t)i * stride);
            if (PER_THREAD && tid == 0) s_rec[Mi] = RANSAC_SENTINEL;  // what a skipped draw reads (thread_walk)
            if (bulk && nb > 0) { mbar_wait(&s_bar, phase); phase ^= 1u; }
            __syncthreads();
        }
        if (PER_THREAD) {                                             // a thread per hypothesis (device draws)
            for (int t = tb + tid; t < te; t += (int)blockDim.x) {
                double e = INFINITY;
                if (!none) {
                    const uint32_t key = trial_key(seed, frame0 + f, t);
                    e = staged ? score_trial_thread<true, RANSAC_UNROLL>(rec, stride, s_rec, (uint32_t)Mi, S, cam, fB_over_S, s_invd, key)
                               : score_trial_thread<false, UG>(rec, stride, s_rec, (uint32_t)Mi, S, cam, fB_over_S, s_invd, key);
                }
                errors[(int64_t)f * T + t] = e;
            }
        } else
        for (int t = tb + warp; t < te; t += nwarps) {
            double *err_out = errors + (int64_t)f * T + t;
            if (none) {
                if (lane == 0) *err_out = INFINITY;
                continue;
            }
            const int32_t *samp = sample_idx + ((int64_t)f * T + t) * S;
            const int32_t *trip = triple_idx + ((int64_t)f * T + t) * 3;
            if (staged) score_trial<true>(rec, stride, s_rec_addr, (uint32_t)Mi, S, cam, fB_over_S, s_invd_addr, samp, trip, lane, err_out);
            else        score_trial<false>(rec, stride, s_rec_addr, (uint32_t)Mi, S, cam, fB_over_S, s_invd_addr, samp, trip, lane, err_out);
        }
        w = (int64_t)f * T + te;
        if (w < w1) {
            __syncthreads();                                          // everyone is done with this frame's records
            // the next frame's bulk copies write s_rec through the async proxy: order this frame's generic-proxy
            // reads and tail stores before them
            if (tid == 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
    }
}

extern "C" int svb_ransac_score(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                                const int32_t *counts, int counts_stride, int counts_col,
                                int F, int T, int S, int max_m, const svb_camera *cam,
                                const int32_t *sample_idx, const int32_t *triple_idx,
                                uint64_t rng_seed, int frame0, int flags, double *errors, void *stream) {
    if (!mcloud || !counts || !cam || !errors) return SVB_E_NULL;
    if (flags & ~SVB_RANSAC_SHARE_SM) return SVB_E_ARG;
    // SHARE_SM: 512-thread blocks (half of an SM's registers) so that blocks of a kernel on another stream co-reside
    const int threads = (flags & SVB_RANSAC_SHARE_SM) ? RANSAC_SMEM_THREADS / 2 : RANSAC_SMEM_THREADS;
    if ((sample_idx == nullptr) != (triple_idx == nullptr)) return SVB_E_NULL;
    if (F < 1 || T < 1 || S < 1) return SVB_E_SHAPE;
    if ((rec_stride_words != 1 && rec_stride_words != 2) || counts_stride < 1 || counts_col < 0 ||
        counts_col >= counts_stride || max_m < 0) return SVB_E_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    CamD c = make_camd(cam);
    ProfScope ps(K_RANSAC_SCORE, st);
    // shared-memory staging when the largest frame fits; otherwise cap = 0 and every gather goes to L2
    size_t smem = (((size_t)max_m + 4) & ~(size_t)3) * sizeof(uint32_t);      // the records and the sentinel behind them
    int cap = max_m;
    if (max_m == 0 || smem > 200 * 1024) { smem = 16; cap = 0; }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // one persistent block per SM, but never fewer than ~2 trials per warp in a block
    int64_t total = (int64_t)F * T;
    int blocks = (int)min((int64_t)sms, (total + 63) / 64);
    if (blocks < 1) blocks = 1;
    // the opt-in to 200 KB of dynamic shared memory is a per-device function attribute: set once per device, not per call
    static bool attr_done[64] = {false};
    if (dev < 0 || dev >= 64 || !attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(ransac_score_persistent<false, RANSAC_UNROLL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(ransac_score_persistent<true, RANSAC_UNROLL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        if (dev >= 0 && dev < 64) attr_done[dev] = true;
    }
    if (sample_idx)                                               // explicit draws: a warp per hypothesis (coalesced index loads)
        ransac_score_persistent<false, RANSAC_UNROLL><<<blocks, threads, smem, st>>>(
            mcloud, rec_stride_words, slab_stride_records, counts, counts_stride, counts_col, F, T, S, cap, c,
            sample_idx, triple_idx, rng_seed, frame0, errors);
    else {                                                        // device draws: a thread per hypothesis
        const int pt = (flags & SVB_RANSAC_SHARE_SM) ? RANSAC_PT_THREADS / 2 : RANSAC_PT_THREADS;
        if (cap > 0)
            ransac_score_persistent<true, RANSAC_UNROLL><<<blocks, pt, smem, st>>>(
                mcloud, rec_stride_words, slab_stride_records, counts, counts_stride, counts_col, F, T, S, cap, c,
                nullptr, nullptr, rng_seed, frame0, errors);
        else                                                      // nothing is staged: the deeper walk (same sums)
            ransac_score_persistent<true, RANSAC_UNROLL_GLOBAL><<<blocks, pt, smem, st>>>(
                mcloud, rec_stride_words, slab_stride_records, counts, counts_stride, counts_col, F, T, S, cap, c,
                nullptr, nullptr, rng_seed, frame0, errors);
    }
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ dump of the counter-based draws
__global__ void ransac_dump_kernel(const uint32_t *__restrict__ mcloud, int stride, int64_t slab_records,
                                   const int32_t *__restrict__ counts, int cstride, int ccol,
                                   int T, int S, CamD cam, uint64_t seed, int frame0,
                                   int32_t *__restrict__ sample_idx, int32_t *__restrict__ triple_idx) {
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * RANSAC_WARPS + (threadIdx.x >> 5);
    if (t >= T) return;
    const int Mi = counts[(int64_t)f * cstride + ccol];
    int32_t *samp = sample_idx + ((int64_t)f * T + t) * S;
    int32_t *trip = triple_idx + ((int64_t)f * T + t) * 3;
    if (Mi < S || Mi < 1) {
        for (int j = lane; j < S; j += 32) samp[j] = -1;
        if (lane < 3) trip[lane] = -1;
        return;
    }
    const uint32_t M = (uint32_t)Mi;
    const uint32_t *rec = mcloud + (int64_t)f * slab_records * stride;
    const uint32_t key = trial_key(seed, frame0 + f, t);
    const Draws dr = make_draws(key, M, lane);
    uint32_t x = dr.x;
    for (int have = 0; have < S;) {                                  // the scorer's draws, in sequence order
        const uint32_t i = scramble(x, dr.sh, dr.mask);
        const bool k = i < M;
        const uint32_t b = __ballot_sync(0xffffffffu, k);
        const int at = have + __popc(b & ((1u << lane) - 1u));
        if (k && at < S) samp[at] = (int32_t)i;
        have += __popc(b);
        x = (x * dr.A32 + dr.C32) & dr.mask;
    }
    P3 p1, p2, p3;
    int tri[3] = {-1, -1, -1};
    bool ok = trial_triple(rec, stride, M, cam, nullptr, key, p1, p2, p3, tri);
    (void)ok;                                                    // image-collinear triples are dumped too
    if (lane < 3) trip[lane] = tri[lane];
}

extern "C" int svb_ransac_dump_indices(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                                       const int32_t *counts, int counts_stride, int counts_col,
                                       int F, int T, int S, const svb_camera *cam,
                                       uint64_t rng_seed, int frame0,
                                       int32_t *sample_idx, int32_t *triple_idx, void *stream) {
    if (!mcloud || !counts || !cam || !sample_idx || !triple_idx) return SVB_E_NULL;
    if (F < 1 || T < 1 || S < 1) return SVB_E_SHAPE;
    if ((rec_stride_words != 1 && rec_stride_words != 2) || counts_stride < 1 || counts_col < 0 ||
        counts_col >= counts_stride) return SVB_E_ARG;
    dim3 grid((T + RANSAC_WARPS - 1) / RANSAC_WARPS, F);
    ProfScope ps(K_RANSAC_DUMP, (cudaStream_t)stream);
    ransac_dump_kernel<<<grid, RANSAC_THREADS, 0, (cudaStream_t)stream>>>(mcloud, rec_stride_words, slab_stride_records,
                                                                          counts, counts_stride, counts_col, T, S,
                                                                          make_camd(cam), rng_seed, frame0,
                                                                          sample_idx, triple_idx);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ arg-min + the winning plane
// smallest v with fl(v / nrm) >= thr  (so that  |s-1| < v  <=>  fl(|s-1|/nrm) < thr)
__host__ __device__ inline double plane_cutoff(double nrm, double thr) {
    if (!(thr > 0.0)) return 0.0;
    if (thr > 1.7e308) return thr;                              // +inf: everything finite is near
    double v = thr * nrm;
    if (!(v > 0.0) || v > 1.7e308) return v > 0.0 ? v : 0.0;
    union DU { double d; uint64_t u; };
    DU q;
    q.d = v;
    for (int i = 0; i < 64 && (q.d / nrm) < thr; ++i) q.u += 1;      // next double up
    for (int i = 0; i < 64; ++i) {
        DU w;
        w.u = q.u - 1;
        if (w.u != 0 && (w.d / nrm) >= thr) q.u = w.u; else break;
    }
    return q.d;
}

__host__ __device__ inline void fill_plane_row(double a, double b, double c, double thr, double best_err,
                                               double best_trial, double *row) {
#ifdef __CUDA_ARCH__
    double nrm = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(a, a), __dmul_rn(b, b)), __dmul_rn(c, c)));
#else
    volatile double aa = a * a, bb = b * b, cc = c * c;          // math.sqrt(a*a+b*b+c*c), :269/:307
    volatile double s1 = aa + bb;
    volatile double s2 = s1 + cc;
    double nrm = sqrt(s2);
#endif
    row[0] = a; row[1] = b; row[2] = c; row[3] = nrm;
    row[4] = plane_cutoff(nrm, thr);
    row[5] = best_err; row[6] = best_trial;
    row[7] = (nrm > 0.0 && nrm < 1.7e308) ? 1.0 : 0.0;
}

__global__ void ransac_finalize_kernel(const uint32_t *__restrict__ mcloud, int stride, int64_t slab_records,
                                       const int32_t *__restrict__ counts, int cstride, int ccol,
                                       int T, int S, CamD cam, const int32_t *__restrict__ triple_idx,
                                       uint64_t seed, int frame0, const double *__restrict__ errors,
                                       double thr, double *__restrict__ planes) {
    __shared__ double s_e[128];
    __shared__ int s_t[128];
    const int f = blockIdx.x;
    const double *e = errors + (int64_t)f * T;
    double be = INFINITY;
    int bt = -1;
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        double v = e[t];
        if (v < be) { be = v; bt = t; }                          // strict <, increasing t: first minimum
    }
    s_e[threadIdx.x] = be;
    s_t[threadIdx.x] = bt;
    __syncthreads();
    for (int o = blockDim.x >> 1; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            double v = s_e[threadIdx.x + o];
            int tt = s_t[threadIdx.x + o];
            bool take = tt >= 0 && (s_t[threadIdx.x] < 0 || v < s_e[threadIdx.x] ||
                                    (v == s_e[threadIdx.x] && tt < s_t[threadIdx.x]));
            if (take) { s_e[threadIdx.x] = v; s_t[threadIdx.x] = tt; }
        }
        __syncthreads();
    }
    if (threadIdx.x != 0) return;
    double *row = planes + (int64_t)f * SVB_PLANE_STRIDE;
    bt = s_t[0];
    be = s_e[0];
    bool ok = bt >= 0;
    double a = 0, b = 0, c = 0;
    if (ok) {
        const int Mi = counts[(int64_t)f * cstride + ccol];
        const uint32_t *rec = mcloud + (int64_t)f * slab_records * stride;
        const int32_t *trip = triple_idx ? triple_idx + ((int64_t)f * T + bt) * 3 : nullptr;
        P3 p1, p2, p3;
        int tri[3];
        ok = Mi >= S && trial_triple(rec, stride, (uint32_t)Mi, cam, trip, trial_key(seed, frame0 + f, bt), p1, p2, p3, tri) &&
             plane_of(p1, p2, p3, a, b, c);
    }
    if (ok) {
        fill_plane_row(a, b, c, thr, be, (double)bt, row);
    } else {
        for (int k = 0; k < SVB_PLANE_STRIDE; ++k) row[k] = 0.0;
        row[0] = row[1] = row[2] = nan("");
        row[5] = INFINITY;
        row[6] = -1.0;
    }
}

extern "C" int svb_ransac_finalize(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                                   const int32_t *counts, int counts_stride, int counts_col,
                                   int F, int T, int S, const svb_camera *cam,
                                   const int32_t *triple_idx, uint64_t rng_seed, int frame0,
                                   const double *errors, double point_threshold,
                                   double *planes, void *stream) {
    if (!mcloud || !counts || !cam || !errors || !planes) return SVB_E_NULL;
    if (F < 1 || T < 1 || S < 1) return SVB_E_SHAPE;
    if ((rec_stride_words != 1 && rec_stride_words != 2) || counts_stride < 1 || counts_col < 0 ||
        counts_col >= counts_stride) return SVB_E_ARG;
    ProfScope ps(K_RANSAC_FINAL, (cudaStream_t)stream);
    ransac_finalize_kernel<<<F, 128, 0, (cudaStream_t)stream>>>(mcloud, rec_stride_words, slab_stride_records, counts,
                                                               counts_stride, counts_col, T, S, make_camd(cam),
                                                               triple_idx, rng_seed, frame0, errors, point_threshold, planes);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_plane_from_abc(const double *h_abc, double point_threshold, double *h_plane_row) {
    if (!h_abc || !h_plane_row) return SVB_E_NULL;
    fill_plane_row(h_abc[0], h_abc[1], h_abc[2], point_threshold, 0.0, -1.0, h_plane_row);
    return SVB_OK;
}
