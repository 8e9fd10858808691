// a8-a14 straight from the path's inputs: the road points of F frames from the FILLED DISPARITY and the
// colour image, without reading the 8-byte cloud records (stereovision.py:97-140).
//
// Whether a cloud point is near the plane depends on (x, y, d) only.  With the plane in pixel form,
//      P.abc - 1 = Z (af x + bf y + c2) - 1 = (s - d) / d,     s = fB (af x + bf y + c2),  Z = fB / d,
// the test |P.abc - 1| < cutoff is |s - d| < cutoff d: one float32 FMA chain per lattice site, decided for
// certain unless  | |s - d| - cutoff d |  falls below E (>= 2x the worst float32 error of that expression,
// bounded once per frame); those few sites (~1e-4) are decided by the reference's own float64 operation
// sequence (functions.py:300-312 on the X, Y, Z of functions.py:191-193).  So the kernel reads 1 byte per
// lattice site instead of an 8-byte record, and the colour -- from the image, where the projection read it
// (functions.py:195) -- only for the 16-pixel spans that hold a near point.
//
//   road_sites_near     grid (8, F): CTA p of frame f clears an eighth of the road mask and takes the units
//                       (one lattice row x 512 pixels, 16 pixels per lane) p, p+8, ...; its warps take those
//                       round-robin.  Near sites are scattered over the lanes (disparity is quantised, a
//                       third to a half of the road's sites pass), so their expensive part -- hue bin,
//                       histogram, reprojection through the displacement-bit table -- runs DENSE: the owners
//                       stage (x, y, d, colour) in the warp's shared-memory buffer, every full group of 32 is
//                       finished with all lanes on, and the result px | py << 11 | hue << 22 goes to the
//                       warp's own segment of the frame's list in HBM (coalesced, no slot atomics).
//   road_sites_scatter  grid (8, F): warp w of CTA p re-reads segment (p, w), keeps the entries whose hue bin
//                       holds more than colour_threshold points and writes 255 at (px, py).
// (A variant in which the CTA that finishes a frame last did the scatter was 30 us slower per 300 frames: one
// CTA per frame has too little memory parallelism for the 64 segments.)
#include "svb_common.cuh"
#include "svb_prof.cuh"
#include <math.h>

#define ST_THREADS 256
#define ST_WARPS (ST_THREADS / 32)
#define ST_PARTS 8                       // CTAs per frame

struct SitesGeom {
    int H, W, step;
    int LH;              // lattice rows
    int chunks;          // 512-pixel units per row
    int units;           // LH * chunks
    int dvec, cvec;      // 16-byte loads allowed on the disparity / the colour image
    int seg_cap;         // list entries per (CTA, warp) segment
};

static __device__ __noinline__ bool exact_near_site(CamD cam, int x, int y, int d, double a, double b, double c, double cutoff) {
    if (d == 0) return false;                                    // not a cloud point (functions.py:188)
    double Z = exact_Z(cam, d);
    double X = exact_XY(cam.cw, cam.f, x, Z);
    double Y = exact_XY(cam.ch, cam.f, y, Z);
    return fabs(__dsub_rn(exact_dot(X, Y, Z, a, b, c), 1.0)) < cutoff;
}

// byte k of the 16 held in four words, as a float (0x4B0000dd - 2^23: no conversion instruction)
// unit -> (lattice row, 512-pixel chunk); CSH = log2(chunks per row) when that is a power of two, else -1
template <int CSH>
__device__ __forceinline__ void unit_row_chunk(const SitesGeom &g, int u, int &ly, int &ch) {
    if (CSH >= 0) { ly = u >> CSH; ch = u & ((1 << CSH) - 1); }
    else { ly = u / g.chunks; ch = u - ly * g.chunks; }
}
__device__ __forceinline__ float byte_as_float(uint32_t w, int k) {
    return __int_as_float((int)__byte_perm(w, 0x4B000000u, 0x7440 + k)) - 8388608.0f;
}

template <int STEP, int CSH>
__global__ void __launch_bounds__(ST_THREADS, 4)
road_sites_near_kernel(const uint8_t *__restrict__ filled, const uint8_t *__restrict__ bgr, const double *__restrict__ planes,
                       SitesGeom g, CamD cam, const uint32_t *__restrict__ rbits, uint32_t *__restrict__ hist,
                       uint8_t *__restrict__ road_mask, uint32_t *__restrict__ list, int32_t *__restrict__ seg_counts,
                       int32_t *__restrict__ counts_out) {
    constexpr int SP = 16 / STEP;                                // sites per lane per unit
    constexpr int BUF = 32 + 32 * SP;                            // staged records per warp: < 32 carried + a unit's
    __shared__ uint2 s_buf[ST_WARPS][BUF];
    __shared__ uint32_t s_h[SVB_HIST_STRIDE];
    const int part = blockIdx.x, f = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double *pl = planes + (int64_t)f * SVB_PLANE_STRIDE;
    const bool valid = pl[7] != 0.0;
    const uint8_t *drow0 = filled + (int64_t)f * g.H * g.W;
    const uint8_t *crow0 = bgr + (int64_t)f * g.H * g.W * 3;
    const int my_units = (g.units - part + ST_PARTS - 1) / ST_PARTS;          // units part, part + 8, ...

    // 16 disparities of unit k of this CTA for this lane
    auto load_d = [&](int k) -> uint4 {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (k >= my_units) return v;
        int ly, ch;
        unit_row_chunk<CSH>(g, part + k * ST_PARTS, ly, ch);
        const int x0 = ch * 512 + lane * 16;
        const uint8_t *p = drow0 + ((uint32_t)(ly * STEP) * (uint32_t)g.W + (uint32_t)x0);
        if (g.dvec && x0 + 16 <= g.W) return __ldg(reinterpret_cast<const uint4 *>(p));
        uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (x0 + j < g.W) w[j >> 2] |= (uint32_t)p[j] << (8 * (j & 3));
        return make_uint4(w[0], w[1], w[2], w[3]);
    };
    uint4 dnext = make_uint4(0, 0, 0, 0);
    if (valid) dnext = load_d(warp);

    // clear my eighth of the mask
    {
        const int64_t bytes = (int64_t)g.H * g.W;
        uint8_t *mask = road_mask + (int64_t)f * bytes;
        int64_t per = (bytes / ST_PARTS) & ~(int64_t)15;
        int64_t b0 = per * part, b1 = part == ST_PARTS - 1 ? bytes : b0 + per;
        if ((((uintptr_t)mask) & 15) == 0) {
            uint4 z = make_uint4(0, 0, 0, 0);
            int64_t v1 = b0 + ((b1 - b0) & ~(int64_t)15);
            for (int64_t o = b0 + (int64_t)tid * 16; o < v1; o += ST_THREADS * 16) *(uint4 *)(mask + o) = z;
            for (int64_t o = v1 + tid; o < b1; o += ST_THREADS) mask[o] = 0;
        } else {
            for (int64_t o = b0 + tid; o < b1; o += ST_THREADS) mask[o] = 0;
        }
    }
    int32_t *my_count = seg_counts + ((int64_t)f * ST_PARTS + part) * ST_WARPS + warp;
    if (!valid) {
        if (lane == 0) *my_count = 0;
        return;
    }
    for (int k = tid; k < SVB_HIST_STRIDE; k += ST_THREADS) s_h[k] = 0;
    __syncthreads();

    // plane in pixel form, scaled by fB:  s = A x + B y + C ;  near  <=>  |s - d| < cut d
    float A, B, C, cutf, E;
    {
        const double a = pl[0], b = pl[1], c = pl[2], cutoff = pl[4];
        const double Ad = cam.fB * (a / cam.f), Bd = cam.fB * (b / cam.f);
        const double Cd = cam.fB * c - Ad * cam.cw - Bd * cam.ch;
        const double S = fabs(Ad) * (double)g.W + fabs(Bd) * (double)g.H + fabs(Cd);
        A = (float)Ad; B = (float)Bd; C = (float)Cd; cutf = (float)cutoff;
        // float32 error of |s - d| - cut d: <= 8 eps S (coefficients, two FMAs, the subtraction) + eps 256 cut + eps 256
        E = (float)(1e-6 * S + 4e-5 * fabs(cutoff) + 4e-5);
    }
    uint2 *buf = s_buf[warp];
    uint32_t *seg = list + (((int64_t)f * ST_PARTS + part) * ST_WARPS + warp) * g.seg_cap;
    int fill = 0, written = 0, spans = 0;                        // warp-uniform: staged records, list entries written, colour spans read
    const uint32_t *rb_y = rbits + (size_t)g.W * 8;                // the (y, d) rows of the displacement table
    auto dense = [&](uint2 rc, bool on) {                        // rc.x = x | y << 16, rc.y = B | G << 8 | R << 16 | d << 24
        if (on) {
            const uint32_t x = rc.x & 0xFFFFu, y = rc.x >> 16, d = rc.y >> 24, c3 = rc.y;
            const uint32_t hue = (uint32_t)hue_bin((int)((c3 >> 16) & 0xFFu), (int)((c3 >> 8) & 0xFFu), (int)(c3 & 0xFFu));
            atomicAdd(&s_h[hue], 1u);
            const uint32_t dw = d >> 5;
            const uint32_t px = x - ((__ldg(rbits + (x * 8u + dw)) >> (d & 31)) & 1u);
            const uint32_t py = y - ((__ldg(rb_y + (y * 8u + dw)) >> (d & 31)) & 1u);
            seg[written + lane] = px | (py << 11) | (hue << 22);
        }
    };
    for (int k = warp; k < my_units; k += ST_WARPS) {
        const uint4 dv = dnext;
        dnext = load_d(k + ST_WARPS);
        int ly, ch;
        unit_row_chunk<CSH>(g, part + k * ST_PARTS, ly, ch);
        const int y = ly * STEP, x0 = ch * 512 + lane * 16;

        // near test of my SP sites
        const float base = fmaf(A, (float)x0, fmaf(B, (float)y, C));
        const uint32_t dw[4] = {dv.x, dv.y, dv.z, dv.w};
        const int room = g.W - 1 - x0;                           // the site at offset o exists when o < room
        uint32_t nb = 0, unsure = 0;
#pragma unroll
        for (int j = 0; j < SP; ++j) {
            const int o = j * STEP;
            const float df = byte_as_float(dw[o >> 2], o & 3);
            const float t = fmaf(A, (float)o, base) - df;
            const float q = fmaf(-cutf, df, fabsf(t));
            if (q < -E) nb |= 1u << j;
            else if (q < E) unsure |= 1u << j;
        }
        if (room < 16) {                                         // the row's last span: sites at or beyond W - 1 do not exist
            const uint32_t okm = room <= 0 ? 0u : (1u << ((room + STEP - 1) / STEP)) - 1u;
            nb &= okm; unsure &= okm;
        }
        if (__any_sync(0xffffffffu, unsure != 0)) {              // rare: the reference's own arithmetic decides
            while (unsure) {
                const int j = __ffs(unsure) - 1;
                unsure &= unsure - 1;
                const int o = j * STEP, wi = o >> 2;
                const uint32_t word = wi == 0 ? dv.x : wi == 1 ? dv.y : wi == 2 ? dv.z : dv.w;
                const int d = (int)((word >> (8 * (o & 3))) & 0xFFu);
                if (exact_near_site(cam, x0 + o, y, d, pl[0], pl[1], pl[2], pl[4])) nb |= 1u << j;
            }
        }
        const uint32_t holders = __ballot_sync(0xffffffffu, nb != 0);
        if (!holders) continue;
        spans += __popc(holders);

        // colour of my 16 pixels (48 bytes) where I hold a near site
        uint32_t cw[12];
#pragma unroll
        for (int w = 0; w < 12; ++w) cw[w] = 0;
        if (nb) {
            const uint8_t *cp = crow0 + ((uint32_t)y * (uint32_t)g.W + (uint32_t)x0) * 3u;
            if (g.cvec && x0 + 16 <= g.W) {
                const uint4 *q = reinterpret_cast<const uint4 *>(cp);
                uint4 c0 = __ldg(q), c1 = __ldg(q + 1), c2v = __ldg(q + 2);
                cw[0] = c0.x; cw[1] = c0.y; cw[2] = c0.z; cw[3] = c0.w;
                cw[4] = c1.x; cw[5] = c1.y; cw[6] = c1.z; cw[7] = c1.w;
                cw[8] = c2v.x; cw[9] = c2v.y; cw[10] = c2v.z; cw[11] = c2v.w;
            } else {
#pragma unroll
                for (int w = 0; w < 12; ++w) {
                    uint32_t v = 0;
#pragma unroll
                    for (int b = 0; b < 4; ++b)
                        if (x0 + (w * 4 + b) / 3 < g.W) v |= (uint32_t)cp[w * 4 + b] << (8 * b);
                    cw[w] = v;
                }
            }
        }
        // stage my near sites behind the carried ones
        int cnt = __popc(nb), inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        int slot = fill + inc - cnt;
        fill += __shfl_sync(0xffffffffu, inc, 31);
        const uint32_t lo_base = (uint32_t)x0 | ((uint32_t)y << 16);
#pragma unroll
        for (int j = 0; j < SP; ++j) {
            if ((nb >> j) & 1u) {
                const int o = j * STEP, bo = 3 * o;
                // B, G, R at bytes 3o .. 3o+2 of the span, the disparity byte on top: one permute each
                const uint32_t c3 = __byte_perm(cw[bo >> 2], cw[(bo >> 2) + ((bo & 3) > 1 ? 1 : 0)], 0x3210 + 0x1111 * (bo & 3));
                const uint32_t hi = __byte_perm(c3, dw[o >> 2], 0x4210 + 0x1000 * (o & 3));
                buf[slot++] = make_uint2(lo_base + (uint32_t)o, hi);
            }
        }
        __syncwarp();
        // finish every full group of 32, carry the rest to the front
        int done = 0;
        for (; done + 32 <= fill; done += 32) {
            dense(buf[done + lane], true);
            written += 32;
        }
        if (done) {
            const int rest = fill - done;
            uint2 carry = make_uint2(0, 0);
            if (lane < rest) carry = buf[done + lane];
            __syncwarp();
            if (lane < rest) buf[lane] = carry;
            fill = rest;
            __syncwarp();
        }
    }
    if (fill) {
        dense(buf[lane], lane < fill);
        written += fill;
    }
    if (lane == 0) {
        *my_count = written;
        if (written) atomicAdd(&counts_out[f * 4 + 2], written);
        if (spans) atomicAdd(&s_h[SVB_HIST_SPANS], (uint32_t)spans);
    }
    __syncthreads();
    uint32_t *gh = hist + (int64_t)f * SVB_HIST_STRIDE;
    for (int k = tid; k < SVB_HUE_BINS; k += ST_THREADS) {
        const uint32_t v = s_h[k];
        if (v) atomicAdd(&gh[k], v);
    }
    if (tid == 0 && s_h[SVB_HIST_SPANS]) atomicAdd(&gh[SVB_HIST_SPANS], s_h[SVB_HIST_SPANS]);   // bookkeeping, not a hue bin
}

// colour filter + scatter: warp w of CTA p re-reads segment (p, w) of frame f
template <bool OVERLAY>
__global__ void __launch_bounds__(ST_THREADS)
road_sites_scatter_kernel(const uint32_t *__restrict__ list, const int32_t *__restrict__ seg_counts, SitesGeom g,
                          const uint32_t *__restrict__ hist, uint32_t colour_threshold, uint8_t *__restrict__ road_mask,
                          uint8_t *__restrict__ overlay, int32_t *__restrict__ counts_out) {
    __shared__ uint32_t s_pop[32];                               // bit b of word w: hue bin 32 w + b is populous enough
    __shared__ int s_r;
    const int part = blockIdx.x, f = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t sid = ((int64_t)f * ST_PARTS + part) * ST_WARPS + warp;
    const int n = seg_counts[sid];
    const uint32_t *sp = list + sid * g.seg_cap;
    // the first entries are on their way while the histogram is thresholded
    uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = 0;
    if (lane < n) w0 = __ldcs(sp + lane);
    if (lane + 32 < n) w1 = __ldcs(sp + lane + 32);
    if (lane + 64 < n) w2 = __ldcs(sp + lane + 64);
    if (lane + 96 < n) w3 = __ldcs(sp + lane + 96);
    if (tid == 0) s_r = 0;
    {
        const uint32_t *gh = hist + (int64_t)f * SVB_HIST_STRIDE;
        uint32_t v[32 / ST_WARPS];
#pragma unroll
        for (int i = 0; i < 32 / ST_WARPS; ++i) v[i] = __ldg(gh + (warp + i * ST_WARPS) * 32 + lane);
#pragma unroll
        for (int i = 0; i < 32 / ST_WARPS; ++i) {
            const uint32_t bal = __ballot_sync(0xffffffffu, v[i] > colour_threshold);
            if (lane == 0) s_pop[warp + i * ST_WARPS] = bal;
        }
    }
    __syncthreads();
    uint8_t *mask = road_mask + (int64_t)f * g.H * g.W;
    uint8_t *ovl = OVERLAY ? overlay + (int64_t)f * g.H * g.W * 3 : nullptr;
    int rcount = 0;
    auto place = [&](uint32_t w, bool in) {
        const uint32_t hue = w >> 22;
        if (in && ((s_pop[hue >> 5] >> (hue & 31)) & 1u)) {
            ++rcount;
            const uint32_t px = w & 0x7FFu, py = (w >> 11) & 0x7FFu;
            if (px < (uint32_t)g.W && py < (uint32_t)g.H) {
                const uint32_t at = py * (uint32_t)g.W + px;
                mask[at] = 255;
                if (OVERLAY) {
                    uint8_t *q = ovl + (size_t)at * 3;
                    q[0] = 0; q[1] = 255; q[2] = 0;               // [0,255,0], stereovision.py:133
                }
            }
        }
    };
    for (int i = lane; i < n; i += 128) {
        const uint32_t c0 = w0, c1 = w1, c2 = w2, c3 = w3;
        const int nx = i + 128;
        if (nx < n) w0 = __ldcs(sp + nx);
        if (nx + 32 < n) w1 = __ldcs(sp + nx + 32);
        if (nx + 64 < n) w2 = __ldcs(sp + nx + 64);
        if (nx + 96 < n) w3 = __ldcs(sp + nx + 96);
        place(c0, true); place(c1, i + 32 < n); place(c2, i + 64 < n); place(c3, i + 96 < n);
    }
    rcount = __reduce_add_sync(0xffffffffu, rcount);
    if (lane == 0 && rcount) atomicAdd(&s_r, rcount);
    __syncthreads();
    if (tid == 0 && s_r) atomicAdd(&counts_out[f * 4 + 3], s_r);
}

// ------------------------------------------------------------------ host side
static bool sites_geom(int H, int W, int step, const void *filled, const void *bgr, SitesGeom *out) {
    if (step != 1 && step != 2) return false;
    SitesGeom g;
    g.H = H; g.W = W; g.step = step;
    g.LH = svb_lattice_rows(H, step);
    g.chunks = (W + 511) / 512;
    g.units = g.LH * g.chunks;
    g.dvec = (W % 16 == 0) && (((uintptr_t)filled) % 16 == 0);
    g.cvec = (W % 16 == 0) && (((uintptr_t)bgr) % 16 == 0);
    const int units_per_part = (g.units + ST_PARTS - 1) / ST_PARTS;
    g.seg_cap = (units_per_part + ST_WARPS - 1) / ST_WARPS * 32 * (16 / step);
    if (out) *out = g;
    return true;
}

extern "C" int svb_road_sites_supported(int H, int W, int step) {
    if (H < 2 || W < 2 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return 0;
    return sites_geom(H, W, step, nullptr, nullptr, nullptr) ? 1 : 0;
}

extern "C" int64_t svb_road_sites_scratch_bytes(int F, int H, int W, int step) {
    SitesGeom g;
    if (F < 1 || H < 2 || W < 2 || H > SVB_MAX_DIM || W > SVB_MAX_DIM || !sites_geom(H, W, step, nullptr, nullptr, &g)) return 0;
    const int64_t segs = (int64_t)F * ST_PARTS * ST_WARPS;
    return ((segs * (int64_t)sizeof(int32_t) + 255) & ~(int64_t)255) + segs * g.seg_cap * (int64_t)sizeof(uint32_t);
}

extern "C" int svb_road_points_sites(const uint8_t *filled, const uint8_t *bgr, const double *planes, int F, int H, int W, int step,
                                     const svb_camera *cam, uint32_t colour_threshold, const uint32_t *reproj_bits, uint32_t *hist,
                                     uint8_t *road_mask, uint8_t *overlay, int32_t *counts_out, void *scratch, int64_t scratch_bytes,
                                     void *stream) {
    if (!filled || !bgr || !planes || !cam || !reproj_bits || !hist || !road_mask || !counts_out || !scratch) return SVB_E_NULL;
    if (F < 1 || H < 2 || W < 2 || H > SVB_MAX_DIM || W > SVB_MAX_DIM || step < 1) return SVB_E_SHAPE;
    SitesGeom g;
    if (!sites_geom(H, W, step, filled, bgr, &g)) return SVB_E_ARG;
    if (scratch_bytes < svb_road_sites_scratch_bytes(F, H, W, step) || (((uintptr_t)scratch) & 3)) return SVB_E_SCRATCH;
    cudaStream_t st = (cudaStream_t)stream;
    CamD c = make_camd(cam);
    const int64_t segs = (int64_t)F * ST_PARTS * ST_WARPS;
    int32_t *seg_counts = (int32_t *)scratch;
    uint32_t *list = (uint32_t *)((char *)scratch + ((segs * (int64_t)sizeof(int32_t) + 255) & ~(int64_t)255));
    cudaError_t e = cudaMemsetAsync(hist, 0, (size_t)F * SVB_HIST_STRIDE * sizeof(uint32_t), st);
    if (e != cudaSuccess) return (int)e;
    e = cudaMemset2DAsync(counts_out + 2, 4 * sizeof(int32_t), 0, 2 * sizeof(int32_t), F, st);
    if (e != cudaSuccess) return (int)e;
    dim3 grid(ST_PARTS, F);
    {
        ProfScope ps(K_ROAD_SITES, st);
#define LAUNCH_NEAR(S, C)                                                                                                   \
        road_sites_near_kernel<S, C><<<grid, ST_THREADS, 0, st>>>(filled, bgr, planes, g, c, reproj_bits, hist, road_mask, list, \
                                                                   seg_counts, counts_out)
        const int csh = (g.chunks & (g.chunks - 1)) ? -1 : (g.chunks == 1 ? 0 : g.chunks == 2 ? 1 : 2);
        if (step == 1) { if (csh == 0) LAUNCH_NEAR(1, 0); else if (csh == 1) LAUNCH_NEAR(1, 1); else if (csh == 2) LAUNCH_NEAR(1, 2); else LAUNCH_NEAR(1, -1); }
        else           { if (csh == 0) LAUNCH_NEAR(2, 0); else if (csh == 1) LAUNCH_NEAR(2, 1); else if (csh == 2) LAUNCH_NEAR(2, 2); else LAUNCH_NEAR(2, -1); }
#undef LAUNCH_NEAR
    }
    SVB_LAUNCH_CHECK();
    ProfScope ps(K_ROAD_SITES_SCATTER, st);
    if (overlay) road_sites_scatter_kernel<true><<<grid, ST_THREADS, 0, st>>>(list, seg_counts, g, hist, colour_threshold, road_mask, overlay, counts_out);
    else         road_sites_scatter_kernel<false><<<grid, ST_THREADS, 0, st>>>(list, seg_counts, g, hist, colour_threshold, road_mask, nullptr, counts_out);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
