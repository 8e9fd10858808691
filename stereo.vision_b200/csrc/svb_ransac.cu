// a5-a7: RANSAC plane fit.  One warp scores one hypothesis; a block holds 8 hypotheses of one frame
// and shares the 255-entry table Z(d) = (f*B)/d in shared memory.  The masked cloud of a frame is
// at most 4 bytes x lattice sites (128 KB at the reference size), i.e. L2-resident: the S gathers
// of a hypothesis are 4-byte L2 reads, the sample-index stream (explicit mode) is read coalesced.
//
// Arithmetic.  The three points of a hypothesis are expanded with the path's exact float64
// expressions (svb_common.cuh) because the "exactly collinear -> draw again" rule of
// functions.py:251-259 is an exact-zero test, and the plane is abc = inv([P1;P2;P3]).1 evaluated
// with the adjugate (fp64).  The S point-plane distances of the score use the factored form
// |Z(d)*(a/f*(x-cw) + b/f*(y-ch) + c) - 1| (5 flops, no division); it differs from the reference
// expression by ~1e-16 relative, far below the gap between two hypotheses' mean errors, and only
// the arg-min (not the error value) is an observable.
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
// keyed bijection of [0, 2^bits): (odd multiply + add) and xor-shift by >= bits/2 are both invertible
// modulo 2^bits; two rounds of each, then cycle-walk into [0, m).  The three multipliers/offset are
// derived once per hypothesis from its key, so a draw costs ~10 integer instructions.
struct PermKey { uint32_t A, B, C, mask; int sh; };
__host__ __device__ __forceinline__ PermKey make_perm(uint32_t key, int bits) {
    PermKey k;
    k.A = mix32(key ^ 0x51ED270Bu) | 1u;
    k.B = mix32(key + 0x9E3779B9u);
    k.C = mix32(key + 0x3C6EF372u) | 1u;
    k.mask = (bits >= 32) ? 0xFFFFFFFFu : ((1u << bits) - 1u);
    k.sh = (bits + 1) >> 1;
    return k;
}
__host__ __device__ __forceinline__ uint32_t perm_draw(const PermKey &k, uint32_t i, uint32_t m) {
    uint32_t v = i;
    do {
        v = (v * k.A + k.B) & k.mask;
        v ^= v >> k.sh;
        v = (v * k.C) & k.mask;
        v ^= v >> k.sh;
    } while (v >= m);
    return v;
}
__host__ __device__ __forceinline__ int bits_for(uint32_t m) {       // smallest b >= 2 with 2^b >= m
    int b = 2;
    while ((1u << b) < m) ++b;
    return b;
}
__device__ __forceinline__ uint32_t uniform_draw(uint32_t key, uint32_t ctr, uint32_t m) {
    return __umulhi(mix32(key ^ mix32(ctr + 0x632BE5ABu)), m);
}

// ------------------------------------------------------------------ geometry
struct P3 { double x, y, z; };

__device__ __forceinline__ P3 expand_point(const CamD &cam, uint32_t lo) {
    P3 p;
    p.z = exact_Z(cam, rec_d(lo));
    p.x = exact_XY(cam.cw, cam.f, rec_x(lo), p.z);
    p.y = exact_XY(cam.ch, cam.f, rec_y(lo), p.z);
    return p;
}
// np.cross(P1-P2, P2-P3) == (0,0,0) exactly                     functions.py:256-259
__device__ __forceinline__ bool exactly_collinear(const P3 &a, const P3 &b, const P3 &c) {
    double ux = __dsub_rn(a.x, b.x), uy = __dsub_rn(a.y, b.y), uz = __dsub_rn(a.z, b.z);
    double vx = __dsub_rn(b.x, c.x), vy = __dsub_rn(b.y, c.y), vz = __dsub_rn(b.z, c.z);
    double cx = __dsub_rn(__dmul_rn(uy, vz), __dmul_rn(uz, vy));
    double cy = __dsub_rn(__dmul_rn(uz, vx), __dmul_rn(ux, vz));
    double cz = __dsub_rn(__dmul_rn(ux, vy), __dmul_rn(uy, vx));
    return cx == 0.0 && cy == 0.0 && cz == 0.0;
}
// Pixels collinear in the image <=> the plane through the three points contains the camera centre
// <=> [P1;P2;P3] is rank 2 in exact arithmetic.  The reference then gets LAPACK rounding noise
// (LinAlgError for about half of such triples, |abc| ~ 1e14 for the rest, machine dependent); the
// product treats every such triple as the singular case (trial skipped, functions.py:294-296).
__device__ __forceinline__ bool image_collinear(uint32_t r1, uint32_t r2, uint32_t r3) {
    int x1 = rec_x(r1), y1 = rec_y(r1);
    return (rec_x(r2) - x1) * (rec_y(r3) - y1) - (rec_x(r3) - x1) * (rec_y(r2) - y1) == 0;
}
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

// ------------------------------------------------------------------ score: one warp per hypothesis
// int -> double without the (quarter-rate) conversion instruction: 2^52 + x has x in its low mantissa bits
__device__ __forceinline__ double u2d(uint32_t x) { return __hiloint2double(0x43300000, (int)x) - 4503599627370496.0; }

__device__ __forceinline__ double bcast(double v, int src) {
    return __hiloint2double(__shfl_sync(0xffffffffu, __double2hiint(v), src), __shfl_sync(0xffffffffu, __double2loint(v), src));
}

// Warp-cooperative version of trial_triple: lanes 0..2 each fetch and expand one point (one division
// sequence per coordinate for the whole warp instead of three), then the nine values are broadcast.
__device__ __forceinline__ bool warp_triple(const uint32_t *__restrict__ rec, int stride, uint32_t M, const CamD &cam,
                                            const int32_t *__restrict__ triple, uint32_t key, int lane,
                                            P3 &p1, P3 &p2, P3 &p3) {
    const int rounds = triple ? 1 : MAX_TRIPLE_ROUNDS;
    for (int round = 0; round < rounds; ++round) {
        uint32_t idx;
        bool valid = true;
        if (triple) {
            int i = lane < 3 ? triple[lane] : 0;
            valid = i >= 0 && (uint32_t)i < M;
            idx = valid ? (uint32_t)i : 0u;
            if (__any_sync(0xffffffffu, !valid)) return false;
        } else {
            idx = uniform_draw(key, 3u * round + (lane < 3 ? lane : 0), M);
        }
        uint32_t r = __ldg(rec + (int64_t)idx * stride);
        P3 p = expand_point(cam, r | (lane < 3 ? 0u : (1u << 22)));     // idle lanes: d >= 1, no division by zero
        p1.x = bcast(p.x, 0); p1.y = bcast(p.y, 0); p1.z = bcast(p.z, 0);
        p2.x = bcast(p.x, 1); p2.y = bcast(p.y, 1); p2.z = bcast(p.z, 1);
        p3.x = bcast(p.x, 2); p3.y = bcast(p.y, 2); p3.z = bcast(p.z, 2);
        if (triple || !exactly_collinear(p1, p2, p3)) {
            uint32_t r1 = __shfl_sync(0xffffffffu, r, 0), r2 = __shfl_sync(0xffffffffu, r, 1), r3 = __shfl_sync(0xffffffffu, r, 2);
            return !image_collinear(r1, r2, r3);
        }
    }
    return false;
}

// plane as in plane_of, with one reciprocal of the determinant instead of three divisions (the result
// moves by ~1 ulp; only the arg-min of the scores and abc to 1e-5 are observables)
__device__ __forceinline__ bool plane_fast(const P3 &p1, const P3 &p2, const P3 &p3, double &a, double &b, double &c) {
    double c23x = p2.y * p3.z - p2.z * p3.y, c23y = p2.z * p3.x - p2.x * p3.z, c23z = p2.x * p3.y - p2.y * p3.x;
    double c31x = p3.y * p1.z - p3.z * p1.y, c31y = p3.z * p1.x - p3.x * p1.z, c31z = p3.x * p1.y - p3.y * p1.x;
    double c12x = p1.y * p2.z - p1.z * p2.y, c12y = p1.z * p2.x - p1.x * p2.z, c12z = p1.x * p2.y - p1.y * p2.x;
    double det = p1.x * c23x + p1.y * c23y + p1.z * c23z;
    if (det == 0.0 || !isfinite(det)) return false;
    double inv = 1.0 / det;
    a = (c23x + c31x + c12x) * inv;
    b = (c23y + c31y + c12y) * inv;
    c = (c23z + c31z + c12z) * inv;
    return isfinite(a) && isfinite(b) && isfinite(c);
}

// One warp scores trial t of one frame.  `src` = where the S sample gathers read the 4-byte records:
// global memory (stride words apart) or the frame's cloud staged in shared memory (stride 1).
template <bool EXPLICIT>
__device__ __forceinline__ void score_trial(const uint32_t *__restrict__ rec, int stride, const uint32_t *src, int src_stride,
                                            uint32_t M, int bits, int S, const CamD &cam, double inv_f, const double *s_Z,
                                            const int32_t *__restrict__ samp, const int32_t *__restrict__ trip,
                                            uint32_t key, int lane, double *err_out) {
    P3 p1, p2, p3;
    double a, b, c;
    bool ok = warp_triple(rec, stride, M, cam, trip, key, lane, p1, p2, p3);
    ok = ok && plane_fast(p1, p2, p3, a, b, c);
    if (!ok) {
        if (lane == 0) *err_out = INFINITY;
        return;
    }
    // |Z(d) * (a/f*(x-cw) + b/f*(y-ch) + c) - 1| with the constant terms folded into c2
    const double af = a * inv_f, bf = b * inv_f;
    const double c2 = c - af * cam.cw - bf * cam.ch;
    const PermKey pk = make_perm(key, bits);
    double acc = 0.0;
    bool bad = false;
#pragma unroll 4
    for (int j = lane; j < S; j += 32) {
        uint32_t idx;
        if (EXPLICIT) {
            int i = __ldg(samp + j);
            bad |= (i < 0 || (uint32_t)i >= M);
            idx = bad ? 0u : (uint32_t)i;
        } else {
            idx = perm_draw(pk, (uint32_t)j, M);
        }
        uint32_t r = src[(int64_t)idx * src_stride];
        double s = s_Z[r >> 22 & 0xFFu] * fma(af, u2d(r & 0x7FFu), fma(bf, u2d(r >> 11 & 0x7FFu), c2));
        acc += fabs(s - 1.0);
    }
    acc = warp_sum(acc);
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) {
        double e = acc / ((double)S * sqrt(a * a + b * b + c * c));
        *err_out = (bad || isnan(e)) ? INFINITY : e;              // NaN never beats bestError, :291
    }
}

// variant 1: gathers from global memory (L2); 8 hypotheses per block
template <bool EXPLICIT>
__global__ void __launch_bounds__(RANSAC_THREADS)
ransac_score_kernel(const uint32_t *__restrict__ mcloud, int stride, int64_t slab_records,
                    const int32_t *__restrict__ counts, int cstride, int ccol,
                    int T, int S, CamD cam,
                    const int32_t *__restrict__ sample_idx, const int32_t *__restrict__ triple_idx,
                    uint64_t seed, int frame0, double *__restrict__ errors) {
    __shared__ double s_Z[256];
    s_Z[threadIdx.x] = threadIdx.x ? exact_Z(cam, threadIdx.x) : 0.0;
    __syncthreads();
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * RANSAC_WARPS + (threadIdx.x >> 5);
    if (t >= T) return;
    const int Mi = counts[(int64_t)f * cstride + ccol];
    double *err_out = errors + (int64_t)f * T + t;
    if (Mi < S || Mi < 1 || S < 1) {                              // random.sample raises, :286,294
        if (lane == 0) *err_out = INFINITY;
        return;
    }
    const uint32_t *rec = mcloud + (int64_t)f * slab_records * stride;
    score_trial<EXPLICIT>(rec, stride, rec, stride, (uint32_t)Mi, bits_for((uint32_t)Mi), S, cam, 1.0 / cam.f, s_Z,
                          EXPLICIT ? sample_idx + ((int64_t)f * T + t) * S : nullptr,
                          EXPLICIT ? triple_idx + ((int64_t)f * T + t) * 3 : nullptr,
                          trial_key(seed, frame0 + f, t), lane, err_out);
}

// variant 2: the frame's masked cloud (<= 4 bytes x max_m) is staged once per block in shared memory
// and the block's warps share it for T/gridDim.x hypotheses each; the S gathers of a hypothesis are
// shared-memory reads (no 32-byte-sector over-fetch from L2).
#define RANSAC_SMEM_THREADS 1024
template <bool EXPLICIT>
__global__ void __launch_bounds__(RANSAC_SMEM_THREADS, 1)
ransac_score_smem_kernel(const uint32_t *__restrict__ mcloud, int stride, int64_t slab_records,
                         const int32_t *__restrict__ counts, int cstride, int ccol,
                         int T, int S, int cap, CamD cam,
                         const int32_t *__restrict__ sample_idx, const int32_t *__restrict__ triple_idx,
                         uint64_t seed, int frame0, double *__restrict__ errors) {
    extern __shared__ __align__(16) uint32_t s_rec[];
    __shared__ double s_Z[256];
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = RANSAC_SMEM_THREADS / 32;
    const int Mi = counts[(int64_t)f * cstride + ccol];
    const uint32_t *rec = mcloud + (int64_t)f * slab_records * stride;
    const bool none = (Mi < S || Mi < 1 || S < 1);
    const bool staged = !none && Mi <= cap;
    const int bits = bits_for((uint32_t)max(Mi, 1));
    const double inv_f = 1.0 / cam.f;
    if (threadIdx.x < 256) s_Z[threadIdx.x] = threadIdx.x ? exact_Z(cam, threadIdx.x) : 0.0;
    if (staged) {
        if (stride == 1 && ((((uintptr_t)rec) & 15) == 0)) {
            const uint4 *src = (const uint4 *)rec;
            int n4 = Mi >> 2;
            for (int i = threadIdx.x; i < n4; i += RANSAC_SMEM_THREADS) ((uint4 *)s_rec)[i] = __ldg(src + i);
            for (int i = (n4 << 2) + threadIdx.x; i < Mi; i += RANSAC_SMEM_THREADS) s_rec[i] = __ldg(rec + i);
        } else {
            for (int i = threadIdx.x; i < Mi; i += RANSAC_SMEM_THREADS) s_rec[i] = __ldg(rec + (int64_t)i * stride);
        }
    }
    __syncthreads();
    // trials of this block: t = blockIdx.x + k * gridDim.x, dealt to the warps round-robin
    for (int t = blockIdx.x + warp * gridDim.x; t < T; t += nwarps * gridDim.x) {
        double *err_out = errors + (int64_t)f * T + t;
        if (none) {
            if (lane == 0) *err_out = INFINITY;
            continue;
        }
        score_trial<EXPLICIT>(rec, stride, staged ? s_rec : rec, staged ? 1 : stride, (uint32_t)Mi, bits, S, cam, inv_f, s_Z,
                              EXPLICIT ? sample_idx + ((int64_t)f * T + t) * S : nullptr,
                              EXPLICIT ? triple_idx + ((int64_t)f * T + t) * 3 : nullptr,
                              trial_key(seed, frame0 + f, t), lane, err_out);
    }
}

extern "C" int svb_ransac_score(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                                const int32_t *counts, int counts_stride, int counts_col,
                                int F, int T, int S, int max_m, const svb_camera *cam,
                                const int32_t *sample_idx, const int32_t *triple_idx,
                                uint64_t rng_seed, int frame0, double *errors, void *stream) {
    if (!mcloud || !counts || !cam || !errors) return SVB_E_NULL;
    if ((sample_idx == nullptr) != (triple_idx == nullptr)) return SVB_E_NULL;
    if (F < 1 || T < 1 || S < 1) return SVB_E_SHAPE;
    if ((rec_stride_words != 1 && rec_stride_words != 2) || counts_stride < 1 || counts_col < 0 ||
        counts_col >= counts_stride || max_m < 0) return SVB_E_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    CamD c = make_camd(cam);
    ProfScope ps(K_RANSAC_SCORE, st);
    const size_t smem = (((size_t)max_m + 3) & ~(size_t)3) * sizeof(uint32_t);
    if (max_m > 0 && smem <= 200 * 1024 && T >= 64) {
        // blocks per frame: enough blocks for ~8 rounds over the SMs (1 block per SM), at least 32 trials each
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        int per_frame = (8 * sms + F - 1) / F;
        if (per_frame > T / 32) per_frame = T / 32;
        if (per_frame < 1) per_frame = 1;
        dim3 grid(per_frame, F);
        cudaError_t e;
        if (sample_idx) {
            e = cudaFuncSetAttribute(ransac_score_smem_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return (int)e;
            ransac_score_smem_kernel<true><<<grid, RANSAC_SMEM_THREADS, smem, st>>>(
                mcloud, rec_stride_words, slab_stride_records, counts, counts_stride, counts_col, T, S, max_m, c,
                sample_idx, triple_idx, rng_seed, frame0, errors);
        } else {
            e = cudaFuncSetAttribute(ransac_score_smem_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return (int)e;
            ransac_score_smem_kernel<false><<<grid, RANSAC_SMEM_THREADS, smem, st>>>(
                mcloud, rec_stride_words, slab_stride_records, counts, counts_stride, counts_col, T, S, max_m, c,
                nullptr, nullptr, rng_seed, frame0, errors);
        }
        SVB_LAUNCH_CHECK();
        return SVB_OK;
    }
    dim3 grid((T + RANSAC_WARPS - 1) / RANSAC_WARPS, F);
    if (sample_idx)
        ransac_score_kernel<true><<<grid, RANSAC_THREADS, 0, st>>>(mcloud, rec_stride_words, slab_stride_records, counts,
                                                                   counts_stride, counts_col, T, S, c, sample_idx,
                                                                   triple_idx, rng_seed, frame0, errors);
    else
        ransac_score_kernel<false><<<grid, RANSAC_THREADS, 0, st>>>(mcloud, rec_stride_words, slab_stride_records, counts,
                                                                    counts_stride, counts_col, T, S, c, nullptr,
                                                                    nullptr, rng_seed, frame0, errors);
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
    const int bits = bits_for(M);
    const PermKey pk = make_perm(key, bits);
    for (int j = lane; j < S; j += 32) samp[j] = (int32_t)perm_draw(pk, (uint32_t)j, M);
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
