// a8-a14: point-plane distances, threshold, hue histogram + dominant-colour filter, 3-D -> 2-D
// reprojection and the road-mask scatter -- as stand-alone helpers (one per reference function)
// and as the fused, frame-batched pair of passes the throughput path runs.
//
// Two passes are inherent: the colour filter needs the complete histogram of the near-plane points
// before any point can be kept (functions.py:215-230).  Pass 1 streams the 8-byte records once,
// tests |P.abc - 1| < cutoff (the exact-equivalent of dist < threshold, see svb_ransac_finalize),
// keeps the answer as one bit per point and builds the frame's 1001-bin histogram in shared memory
// (merged into global memory with one atomic per non-empty bin per block).  Pass 2 streams the
// records again, looks the bit and the bin count up, reprojects the survivors with the reference's
// float64 expression and stores 255 into the mask (idempotent byte stores, no atomics).
#include "svb_common.cuh"
#include "svb_prof.cuh"
#include <math.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

// test hook: 0 = automatic (displacement table given: cluster kernel when a CTA's near list fits in shared
// memory, else the streamed path; no table: legacy two-pass kernels), 1 = legacy two-pass kernels,
// 2 = cluster kernel if it fits, 3 = streamed path
int g_svb_road_mode = 0;
extern "C" int svb_debug_force_two_pass(int mode) { g_svb_road_mode = (mode < 0 || mode > 3) ? 0 : mode; return SVB_OK; }

#define ROAD_THREADS 256
#define ROAD_CHUNK 8192          // records per block in the fused passes

// ------------------------------------------------------------------ a8 calculatePointErrors
__global__ void point_errors_kernel(const uint32_t *__restrict__ rec, int stride, int n, CamD cam,
                                    double a, double b, double c, double nrm, double *__restrict__ dist) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t lo = __ldg(rec + (int64_t)i * stride);
    double Z = exact_Z(cam, rec_d(lo));
    double X = exact_XY(cam.cw, cam.f, rec_x(lo), Z);
    double Y = exact_XY(cam.ch, cam.f, rec_y(lo), Z);
    double s = exact_dot(X, Y, Z, a, b, c);
    dist[i] = fabs(__ddiv_rn(__dsub_rn(s, 1.0), nrm));           // abs((dot - 1)/d), functions.py:310
}

extern "C" int svb_point_errors(const uint32_t *records, int rec_stride_words, int n, const svb_camera *cam,
                                const double *h_abc, double *dist, void *stream) {
    if (n == 0) return SVB_OK;
    if (!records || !cam || !h_abc || !dist) return SVB_E_NULL;
    if (n < 0 || (rec_stride_words != 1 && rec_stride_words != 2)) return SVB_E_ARG;
    double row[SVB_PLANE_STRIDE];
    svb_plane_from_abc(h_abc, 0.0, row);
    ProfScope ps(K_POINT_ERR, (cudaStream_t)stream);
    point_errors_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(records, rec_stride_words, n, make_camd(cam),
                                                                            row[0], row[1], row[2], row[3], dist);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ a9 / a12 stable compaction
#define COMPACT_PER_BLOCK 2048

__device__ __forceinline__ bool keep_flag(int mode, const uint64_t *__restrict__ cloud, const double *__restrict__ dist,
                                          const uint32_t *__restrict__ hist, double thr, uint32_t uthr, int i) {
    if (mode == 0) return dist[i] < thr;                         // differences[i] < threshold, :321
    uint64_t r = cloud[i];
    return hist[hue_of_hi((uint32_t)(r >> 32))] > uthr;   // histogram[key] > threshold, :230
}

__global__ void compact_count_kernel(const uint64_t *__restrict__ cloud, int n, int mode, const double *__restrict__ dist,
                                     const uint32_t *__restrict__ hist, double thr, uint32_t uthr,
                                     uint32_t *__restrict__ block_counts) {
    __shared__ int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    int base = blockIdx.x * COMPACT_PER_BLOCK;
    int c = 0;
    for (int k = threadIdx.x; k < COMPACT_PER_BLOCK; k += blockDim.x) {
        int i = base + k;
        if (i < n && keep_flag(mode, cloud, dist, hist, thr, uthr, i)) ++c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) block_counts[blockIdx.x] = (uint32_t)s_cnt;
}

__global__ void compact_scan_kernel(uint32_t *__restrict__ block_counts, int nblocks, int32_t *__restrict__ out_count) {
    __shared__ uint32_t sw[32];
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint32_t carry = 0;
    for (int base = 0; base < nblocks; base += blockDim.x) {
        int i = base + threadIdx.x;
        int v = i < nblocks ? (int)block_counts[i] : 0;
        int inc = warp_incl_scan(v, lane);
        if (lane == 31) sw[wid] = inc;
        __syncthreads();
        uint32_t b = 0, tot = 0;
        for (int w = 0; w < nw; ++w) { if (w < wid) b += sw[w]; tot += sw[w]; }
        if (i < nblocks) block_counts[i] = carry + b + inc - v;
        carry += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) out_count[0] = (int32_t)carry;
}

__global__ void __launch_bounds__(256)
compact_write_kernel(const uint64_t *__restrict__ cloud, int n, int mode, const double *__restrict__ dist,
                     const uint32_t *__restrict__ hist, double thr, uint32_t uthr,
                     const uint32_t *__restrict__ block_offsets, uint64_t *__restrict__ out) {
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_run;
    if (threadIdx.x == 0) s_run = block_offsets[blockIdx.x];
    __syncthreads();
    int base = blockIdx.x * COMPACT_PER_BLOCK;
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int k0 = 0; k0 < COMPACT_PER_BLOCK; k0 += 256) {
        int i = base + k0 + threadIdx.x;
        bool keep = i < n && keep_flag(mode, cloud, dist, hist, thr, uthr, i);
        uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_w[wid] = __popc(bal);
        __syncthreads();
        uint32_t off = s_run, tot = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { if (w < wid) off += s_w[w]; tot += s_w[w]; }
        if (keep) out[off + __popc(bal & ((1u << lane) - 1u))] = cloud[i];
        __syncthreads();
        if (threadIdx.x == 0) s_run += tot;
        __syncthreads();
    }
}

extern "C" int64_t svb_compact_scratch_bytes(int n) {
    int nb = (n + COMPACT_PER_BLOCK - 1) / COMPACT_PER_BLOCK;
    return (int64_t)(nb + 1) * (int64_t)sizeof(uint32_t);
}

extern "C" int svb_compact_cloud(const uint64_t *cloud, int n, int mode, const double *dist,
                                 const uint32_t *hist, double threshold,
                                 uint64_t *out_cloud, int32_t *out_count,
                                 void *scratch, int64_t scratch_bytes, void *stream) {
    if (!out_count) return SVB_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        cudaError_t e = cudaMemsetAsync(out_count, 0, sizeof(int32_t), st);
        return e == cudaSuccess ? SVB_OK : (int)e;
    }
    if (!cloud || !out_cloud || !scratch) return SVB_E_NULL;
    if (n < 0 || (mode != 0 && mode != 1)) return SVB_E_ARG;
    if ((mode == 0 && !dist) || (mode == 1 && !hist)) return SVB_E_NULL;
    if (scratch_bytes < svb_compact_scratch_bytes(n)) return SVB_E_SCRATCH;
    int nb = (n + COMPACT_PER_BLOCK - 1) / COMPACT_PER_BLOCK;
    uint32_t *bc = (uint32_t *)scratch;
    // histogram[key] > threshold compares integer counts with the caller's number
    double t = threshold;
    uint32_t uthr = t < 0 ? 0u : (t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)floor(t));
    if (mode == 1 && t < 0) uthr = 0;                            // count > negative: always true (count >= 1 ... or 0 > t)
    ProfScope ps(K_COMPACT, st);
    compact_count_kernel<<<nb, 256, 0, st>>>(cloud, n, mode, dist, hist, threshold, uthr, bc);
    SVB_LAUNCH_CHECK();
    compact_scan_kernel<<<1, 256, 0, st>>>(bc, nb, out_count);
    SVB_LAUNCH_CHECK();
    compact_write_kernel<<<nb, 256, 0, st>>>(cloud, n, mode, dist, hist, threshold, uthr, bc, out_cloud);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ a9 / a12 without caller scratch (no allocation)
// The block counts of the compaction live in a file-scope device array instead of caller scratch, for callers that
// cannot allocate (CUDA graph capture, tight serving loops).  Every GPU that loads the library holds these 8 KB until
// exit.  In host code `g_compact_static` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, fetched once per device with that device current.  All calls share the array; it keeps the last
// call's contents, which is harmless here: the count kernel stores (never accumulates) every entry a call reads.
#define COMPACT_STATIC_BLOCKS (SVB_COMPACT_STATIC_MAX_N / COMPACT_PER_BLOCK)
static_assert(SVB_COMPACT_STATIC_MAX_N % COMPACT_PER_BLOCK == 0, "the maximum is a whole number of blocks");
static __device__ __align__(256) uint32_t g_compact_static[COMPACT_STATIC_BLOCKS + 1];

extern "C" int svb_compact_cloud_static(const uint64_t *cloud, int n, int mode, const double *dist,
                                        const uint32_t *hist, double threshold,
                                        uint64_t *out_cloud, int32_t *out_count, void *stream) {
    if (!out_count) return SVB_E_NULL;
    if (n < 0 || (mode != 0 && mode != 1)) return SVB_E_ARG;
    if (n > SVB_COMPACT_STATIC_MAX_N) return SVB_E_LIMIT;
    static uint32_t *dev_ptr[64] = {nullptr};                     // the array's address on each device
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return SVB_E_ARG;
    if (!dev_ptr[dev]) {                                          // set-up, once per device, that device current
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_compact_static);
        if (e != cudaSuccess) return (int)e;
        if (!p) return SVB_E_NULL;
        dev_ptr[dev] = (uint32_t *)p;
    }
    return svb_compact_cloud(cloud, n, mode, dist, hist, threshold, out_cloud, out_count, dev_ptr[dev],
                             (int64_t)sizeof(g_compact_static), stream);
}

// ------------------------------------------------------------------ a10-a11 hue histogram
__global__ void __launch_bounds__(256)
hue_hist_kernel(const uint64_t *__restrict__ cloud, int n, uint32_t *__restrict__ hist) {
    __shared__ uint32_t s_h[SVB_HIST_STRIDE];
    for (int k = threadIdx.x; k < SVB_HIST_STRIDE; k += blockDim.x) s_h[k] = 0;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint64_t r = cloud[i];
        atomicAdd(&s_h[hue_of_hi((uint32_t)(r >> 32))], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < SVB_HUE_BINS; k += blockDim.x)
        if (s_h[k]) atomicAdd(&hist[k], s_h[k]);
}

extern "C" int svb_hue_histogram(const uint64_t *cloud, int n, uint32_t *hist, void *stream) {
    if (!hist) return SVB_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(hist, 0, SVB_HIST_STRIDE * sizeof(uint32_t), st);
    if (e != cudaSuccess) return (int)e;
    if (n == 0) return SVB_OK;
    if (!cloud) return SVB_E_NULL;
    if (n < 0) return SVB_E_ARG;
    int blocks = (n + 256 * 16 - 1) / (256 * 16);
    if (blocks > 148 * 8) blocks = 148 * 8;
    ProfScope ps(K_HUE_HIST, st);
    hue_hist_kernel<<<blocks, 256, 0, st>>>(cloud, n, hist);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

__global__ void hue_bins_kernel(const uint8_t *__restrict__ rgb, int64_t n, uint16_t *__restrict__ bins) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bins[i] = (uint16_t)hue_bin(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
}

extern "C" int svb_hue_bins(const uint8_t *rgb, int64_t n, uint16_t *bins, void *stream) {
    if (n == 0) return SVB_OK;
    if (!rgb || !bins) return SVB_E_NULL;
    if (n < 0) return SVB_E_ARG;
    ProfScope ps(K_HUE_BINS, (cudaStream_t)stream);
    hue_bins_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rgb, n, bins);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ a13 reprojection
__global__ void reproject_kernel(const uint32_t *__restrict__ rec, int stride, int n, CamD cam,
                                 double *__restrict__ xy, int32_t *__restrict__ xyi) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t lo = __ldg(rec + (int64_t)i * stride);
    double Z = exact_Z(cam, rec_d(lo));
    double X = exact_XY(cam.cw, cam.f, rec_x(lo), Z);
    double Y = exact_XY(cam.ch, cam.f, rec_y(lo), Z);
    double px = exact_reproject(cam.cw, cam.f, X, Z);
    double py = exact_reproject(cam.ch, cam.f, Y, Z);
    if (xy) { xy[2 * (int64_t)i] = px; xy[2 * (int64_t)i + 1] = py; }
    if (xyi) { xyi[2 * (int64_t)i] = (int32_t)px; xyi[2 * (int64_t)i + 1] = (int32_t)py; }   // trunc toward zero
}

extern "C" int svb_reproject(const uint32_t *records, int rec_stride_words, int n, const svb_camera *cam,
                             double *xy_f64, int32_t *xy_i32, void *stream) {
    if (n == 0) return SVB_OK;
    if (!records || !cam) return SVB_E_NULL;
    if (n < 0 || (rec_stride_words != 1 && rec_stride_words != 2)) return SVB_E_ARG;
    ProfScope ps(K_REPROJECT, (cudaStream_t)stream);
    reproject_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(records, rec_stride_words, n, make_camd(cam),
                                                                         xy_f64, xy_i32);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

__global__ void reproject_table_kernel(int axis, int n_coord, CamD cam, int32_t *__restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_coord * 256) return;
    int coord = i >> 8, d = i & 255;
    if (d == 0) { out[i] = -1; return; }
    double centre = axis == 0 ? cam.cw : cam.ch;
    double Z = exact_Z(cam, d);
    double V = exact_XY(centre, cam.f, coord, Z);
    out[i] = (int32_t)exact_reproject(centre, cam.f, V, Z);
}

extern "C" int svb_reproject_table(int axis, int n_coord, const svb_camera *cam, int32_t *out, void *stream) {
    if (!cam || !out) return SVB_E_NULL;
    if (n_coord < 1 || (axis != 0 && axis != 1)) return SVB_E_ARG;
    int total = n_coord * 256;
    ProfScope ps(K_REPROJECT, (cudaStream_t)stream);
    reproject_table_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(axis, n_coord, make_camd(cam), out);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// Reprojection displacement bits.  x' = trunc(((X*f)/Z)+cw) with X = ((x-cw)*Z)/f is x or x-1
// depending only on (x, d), likewise y' on (y, d).  The table holds that one bit for
// every coordinate and every d, computed with the exact float64 sequence; 50 KB at 1024x544, so the
// lookups of pass 2 stay in L1.  Layout: word [c*8 + (d>>5)] bit (d&31), x rows first ([0,W)), then
// y rows ([W, W+H)).  *ok is cleared if some pair moves by anything other than 0 or -1 (then the
// caller must not pass the table and pass 2 evaluates the expression per point).
__global__ void reproject_bits_kernel(int H, int W, CamD cam, uint32_t *__restrict__ bits, int32_t *__restrict__ ok) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;             // one thread per (coordinate, 32 d-values)
    if (i >= (W + H) * 8) return;
    int row = i >> 3, dw = i & 7;
    bool is_x = row < W;
    int coord = is_x ? row : row - W;
    double centre = is_x ? cam.cw : cam.ch;
    uint32_t word = 0;
    for (int k = 0; k < 32; ++k) {
        int d = dw * 32 + k;
        if (d == 0) continue;
        double Z = exact_Z(cam, d);
        double V = exact_XY(centre, cam.f, coord, Z);
        int p = (int)exact_reproject(centre, cam.f, V, Z);
        int delta = coord - p;
        if (delta == 1) word |= 1u << k;
        else if (delta != 0) atomicExch(ok, 0);
    }
    bits[i] = word;
}

extern "C" int64_t svb_reproject_bits_words(int H, int W) { return (int64_t)(H + W) * 8; }

extern "C" int svb_reproject_bits(int H, int W, const svb_camera *cam, uint32_t *bits, int32_t *ok, void *stream) {
    if (!cam || !bits || !ok) return SVB_E_NULL;
    if (H < 1 || W < 1 || H > SVB_MAX_DIM || W > SVB_MAX_DIM) return SVB_E_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    int32_t one = 1;
    cudaError_t e = cudaMemcpyAsync(ok, &one, sizeof(one), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
    int total = (W + H) * 8;
    ProfScope ps(K_REPROJECT, st);
    reproject_bits_kernel<<<(total + 127) / 128, 128, 0, st>>>(H, W, make_camd(cam), bits, ok);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ a14 scatter
__global__ void scatter_points_kernel(const int32_t *__restrict__ pts, int n, uint8_t *__restrict__ img,
                                      int H, int W, int channels, uint32_t value, int32_t *__restrict__ oob) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int x = pts[2 * (int64_t)i], y = pts[2 * (int64_t)i + 1];
    if (x < 0 || y < 0 || x >= W || y >= H) {
        if (oob) atomicAdd(oob, 1);
        return;
    }
    uint8_t *p = img + ((int64_t)y * W + x) * channels;
    for (int k = 0; k < channels; ++k) p[k] = (uint8_t)(value >> (8 * k));
}

extern "C" int svb_scatter_points(const int32_t *pts, int n, uint8_t *img, int H, int W, int channels,
                                  const uint8_t *h_value, int32_t *oob, void *stream) {
    if (n == 0) return SVB_OK;
    if (!pts || !img || !h_value) return SVB_E_NULL;
    if (n < 0 || H < 1 || W < 1 || channels < 1 || channels > 4) return SVB_E_ARG;
    uint32_t v = 0;
    for (int k = 0; k < channels; ++k) v |= (uint32_t)h_value[k] << (8 * k);
    ProfScope ps(K_SCATTER, (cudaStream_t)stream);
    scatter_points_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(pts, n, img, H, W, channels, v, oob);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ fused, frame-batched road points
struct RoadGeom {
    int H, W;
    int64_t slab;        // records per frame slab (= lattice size)
    int words;           // near-bit words per frame
    int parts;           // blocks per frame
};

static RoadGeom make_road_geom(int H, int W, int step) {
    RoadGeom g;
    g.H = H; g.W = W;
    int LH = svb_lattice_rows(H, step), LW = svb_lattice_cols(W, step);
    g.slab = (int64_t)LH * LW;
    g.words = (int)((g.slab + 31) / 32);
    g.parts = (int)((g.slab + ROAD_CHUNK - 1) / ROAD_CHUNK);
    if (g.parts < 1) g.parts = 1;
    return g;
}

// the reference's operation sequence for  |P.abc - 1| < cutoff  (slow path of the filtered test)
static __device__ __noinline__ bool exact_near(CamD cam, uint32_t lo, double Z, double a, double b, double c, double cutoff) {
    double X = exact_XY(cam.cw, cam.f, rec_x(lo), Z);
    double Y = exact_XY(cam.ch, cam.f, rec_y(lo), Z);
    return fabs(__dsub_rn(exact_dot(X, Y, Z, a, b, c), 1.0)) < cutoff;
}

// pass 1: near bits + histogram of the near points
__global__ void __launch_bounds__(ROAD_THREADS)
road_pass1_kernel(const uint64_t *__restrict__ cloud, const int32_t *__restrict__ counts,
                  const double *__restrict__ planes, RoadGeom g, CamD cam,
                  uint32_t *__restrict__ nearbits, uint32_t *__restrict__ hist, int32_t *__restrict__ counts_out) {
    __shared__ uint32_t s_h[SVB_HIST_STRIDE];
    __shared__ double s_Z[256];
    __shared__ int s_k;
    const int f = blockIdx.y;
    const int n = counts[f * 4 + 0];
    const int begin = blockIdx.x * ROAD_CHUNK;
    if (begin >= n) return;
    const double *pl = planes + (int64_t)f * SVB_PLANE_STRIDE;
    const bool valid = pl[7] != 0.0;
    const int end = min(n, begin + ROAD_CHUNK);
    uint32_t *bits = nearbits + (int64_t)f * g.words;
    if (!valid) {                                                 // no plane: nothing is near (stereovision.py:117)
        for (int w = begin / 32 + threadIdx.x; w < (end + 31) / 32; w += ROAD_THREADS) bits[w] = 0u;
        return;
    }
    for (int k = threadIdx.x; k < SVB_HIST_STRIDE; k += ROAD_THREADS) s_h[k] = 0;
    s_Z[threadIdx.x] = threadIdx.x ? exact_Z(cam, threadIdx.x) : 0.0;
    if (threadIdx.x == 0) s_k = 0;
    __syncthreads();
    const double a = pl[0], b = pl[1], c = pl[2], cutoff = pl[4];
    const double af = a / cam.f, bf = b / cam.f;
    const uint64_t *rec = cloud + (int64_t)f * g.slab;
    int kcount = 0;
    for (int i0 = begin + (threadIdx.x & ~31); i0 < end; i0 += ROAD_THREADS) {
        int i = i0 + (threadIdx.x & 31);
        bool near = false;
        if (i < end) {
            uint64_t r = __ldcs((const unsigned long long *)(rec + i));
            uint32_t lo = (uint32_t)r, hi = (uint32_t)(r >> 32);
            double Z = s_Z[rec_d(lo)];
            // filtered predicate: a 5-flop estimate of P.abc decides unless it lands within a guard band
            // (~1e4 x the worst-case rounding difference) of the cutoff; then the exact sequence decides
            double p = af * ((double)rec_x(lo) - cam.cw), q = bf * ((double)rec_y(lo) - cam.ch);
            double est = fabs(fma(Z, p + q + c, -1.0));
            double band = 1e-11 * fma(Z, fabs(p) + fabs(q) + fabs(c), 1.0);
            if (fabs(est - cutoff) > band) near = est < cutoff;
            else near = exact_near(cam, lo, Z, a, b, c, cutoff);
            if (near) atomicAdd(&s_h[hue_of_hi(hi)], 1u);
        }
        uint32_t bal = __ballot_sync(0xffffffffu, near);
        if ((threadIdx.x & 31) == 0) {
            bits[i0 >> 5] = bal;
            kcount += __popc(bal);
        }
    }
    if ((threadIdx.x & 31) == 0 && kcount) atomicAdd(&s_k, kcount);
    __syncthreads();
    uint32_t *gh = hist + (int64_t)f * SVB_HIST_STRIDE;
    for (int k = threadIdx.x; k < SVB_HUE_BINS; k += ROAD_THREADS)
        if (s_h[k]) atomicAdd(&gh[k], s_h[k]);
    if (threadIdx.x == 0 && s_k) atomicAdd(&counts_out[f * 4 + 2], s_k);
}

// pass 2: colour filter, reprojection, scatter
template <bool OVERLAY, bool TABLE>
__global__ void __launch_bounds__(ROAD_THREADS)
road_pass2_kernel(const uint64_t *__restrict__ cloud, const int32_t *__restrict__ counts,
                  const double *__restrict__ planes, RoadGeom g, CamD cam, uint32_t colour_threshold,
                  const uint32_t *__restrict__ nearbits, const uint32_t *__restrict__ hist,
                  const uint32_t *__restrict__ rbits, uint8_t *__restrict__ road_mask, uint8_t *__restrict__ overlay, int32_t *__restrict__ counts_out) {
    __shared__ uint32_t s_h[SVB_HIST_STRIDE];
    __shared__ double s_Z[256];
    __shared__ int s_r;
    const int f = blockIdx.y;
    const int n = counts[f * 4 + 0];
    const int begin = blockIdx.x * ROAD_CHUNK;
    if (begin >= n) return;
    const double *pl = planes + (int64_t)f * SVB_PLANE_STRIDE;
    if (pl[7] == 0.0) return;
    const int end = min(n, begin + ROAD_CHUNK);
    const uint32_t *gh = hist + (int64_t)f * SVB_HIST_STRIDE;
    for (int k = threadIdx.x; k < SVB_HIST_STRIDE; k += ROAD_THREADS) s_h[k] = gh[k];
    s_Z[threadIdx.x] = threadIdx.x ? exact_Z(cam, threadIdx.x) : 0.0;
    if (threadIdx.x == 0) s_r = 0;
    __syncthreads();
    const uint64_t *rec = cloud + (int64_t)f * g.slab;
    const uint32_t *bits = nearbits + (int64_t)f * g.words;
    uint8_t *mask = road_mask + (int64_t)f * g.H * g.W;
    uint8_t *ovl = OVERLAY ? overlay + (int64_t)f * g.H * g.W * 3 : nullptr;
    int rcount = 0;
    for (int i0 = begin + (threadIdx.x & ~31); i0 < end; i0 += ROAD_THREADS) {
        uint32_t word = __ldg(bits + (i0 >> 5));
        int lane = threadIdx.x & 31;
        bool keep = false;
        if ((word >> lane) & 1u) {
            uint64_t r = __ldcs((const unsigned long long *)(rec + i0 + lane));
            uint32_t lo = (uint32_t)r, hi = (uint32_t)(r >> 32);
            if (s_h[hue_of_hi(hi)] > colour_threshold) {
                keep = true;
                int x = rec_x(lo), y = rec_y(lo), d = rec_d(lo), px, py;
                if (TABLE) {
                    px = x - (int)((__ldg(rbits + x * 8 + (d >> 5)) >> (d & 31)) & 1u);
                    py = y - (int)((__ldg(rbits + (g.W + y) * 8 + (d >> 5)) >> (d & 31)) & 1u);
                } else {
                    double Z = s_Z[d];
                    double X = exact_XY(cam.cw, cam.f, x, Z);
                    double Y = exact_XY(cam.ch, cam.f, y, Z);
                    px = (int)exact_reproject(cam.cw, cam.f, X, Z);
                    py = (int)exact_reproject(cam.ch, cam.f, Y, Z);
                }
                if (px >= 0 && py >= 0 && px < g.W && py < g.H) {
                    mask[(int64_t)py * g.W + px] = 255;
                    if (OVERLAY) {
                        uint8_t *q = ovl + ((int64_t)py * g.W + px) * 3;
                        q[0] = 0; q[1] = 255; q[2] = 0;           // [0,255,0], stereovision.py:133
                    }
                }
            }
        }
        uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) rcount += __popc(bal);
    }
    if ((threadIdx.x & 31) == 0 && rcount) atomicAdd(&s_r, rcount);
    __syncthreads();
    if (threadIdx.x == 0 && s_r) atomicAdd(&counts_out[f * 4 + 3], s_r);
}

// ------------------------------------------------------------------ the fused road kernel (clusters + DSMEM)
// One thread-block CLUSTER of 8 CTAs owns one frame; CTA r owns the r-th eighth of the frame's records.
//   A  clear this CTA's eighth of the road mask (replaces a memset launch)
//   B  stream the records once (128-bit loads, the next round's loads in flight while this round is
//      tested): filtered near-plane test; a near point gets its hue bin (counted in a shared-memory
//      histogram), is reprojected through the displacement-bit table and kept as one word
//      px | py<<11 | hue<<22 in a list in SHARED memory.  The list is private per thread (slot-major,
//      conflict-free, no ballots or atomics): the order of the near points is not an observable.
//      -- cluster barrier --
//   C  every CTA sums the 8 partial histograms straight out of its peers' shared memory (DSMEM)
//      -- cluster barrier arrive (peers may not exit while their shared memory is being read) --
//   D  colour filter over the thread's own near list, 255 into the mask
//      -- cluster barrier wait --
// The records are read from HBM exactly once and the near list never leaves the SM.
#define RC_THREADS 512
#define RC_PARTS 8
#define RC_PER_ROUND 4            // records per thread per round: two 128-bit loads of 2 records each
#define RC_ROUND (RC_THREADS * RC_PER_ROUND)
#define RC_WARPS (RC_THREADS / 32)

__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// int -> double without the conversion instruction: 2^52 + x has x in its low mantissa bits
__device__ __forceinline__ double road_u2d(uint32_t x) { return __hiloint2double(0x43300000, (int)x) - 4503599627370496.0; }

template <bool OVERLAY>
__global__ void __launch_bounds__(RC_THREADS, 2)
road_cluster_kernel(const uint64_t *__restrict__ cloud, const int32_t *__restrict__ counts,
                    const double *__restrict__ planes, RoadGeom g, CamD cam, uint32_t colour_threshold,
                    const uint32_t *__restrict__ rbits, uint32_t *__restrict__ hist,
                    uint8_t *__restrict__ road_mask, uint8_t *__restrict__ overlay, int32_t *__restrict__ counts_out,
                    int warp_cap) {
    extern __shared__ __align__(16) uint32_t s_list[];           // [RC_WARPS][warp_cap]: near lists, one per warp
    __shared__ uint64_t s_stage[RC_WARPS][64];                   // per-warp ring of near records awaiting the dense step
    __shared__ uint32_t s_h[SVB_HIST_STRIDE];
    __shared__ uint32_t s_tot[SVB_HIST_STRIDE];
    __shared__ double s_Z[256];
    __shared__ int s_k, s_r;
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int f = blockIdx.x / RC_PARTS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = counts[f * 4 + 0];
    const double *pl = planes + (int64_t)f * SVB_PLANE_STRIDE;
    const bool valid = pl[7] != 0.0 && n > 0;

    // records of this CTA, and the first round's loads (issued before anything else).
    // CTA r takes rounds r, r + 8, ... of RC_ROUND records each: the near points (the expensive ones) sit in
    // the lower image rows, and a contiguous split would leave the CTAs of a cluster waiting for the last one
    const uint64_t *rec = cloud + (int64_t)f * g.slab;
    const int begin = rank * RC_ROUND, end = n;
    const bool vec = ((((uintptr_t)rec) & 15) == 0);
    const int toff = warp * 128 + lane * 2;                       // warp w covers [128 w, 128 w + 128) of a round
    uint64_t nx[RC_PER_ROUND];
    auto load_round = [&](int i0) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            int i = i0 + toff + u * 64;
            nx[2 * u] = nx[2 * u + 1] = 0;
            if (vec && i + 1 < end) {
                uint4 v = __ldcs((const uint4 *)(rec + i));
                nx[2 * u] = (uint64_t)v.x | ((uint64_t)v.y << 32);
                nx[2 * u + 1] = (uint64_t)v.z | ((uint64_t)v.w << 32);
            } else {
                if (i < end) nx[2 * u] = __ldcs((const unsigned long long *)(rec + i));
                if (i + 1 < end) nx[2 * u + 1] = __ldcs((const unsigned long long *)(rec + i + 1));
            }
        }
    };
    if (valid) load_round(begin);

    // A: clear my eighth of the mask (and zero the frame's histogram when there is nothing to do)
    {
        const int64_t bytes = (int64_t)g.H * g.W;
        uint8_t *mask = road_mask + (int64_t)f * bytes;
        int64_t per = (bytes / RC_PARTS) & ~(int64_t)15;
        int64_t b0 = per * rank, b1 = rank == RC_PARTS - 1 ? bytes : b0 + per;
        if ((((uintptr_t)mask) & 15) == 0) {
            uint4 z = make_uint4(0, 0, 0, 0);
            int64_t v1 = b0 + ((b1 - b0) & ~(int64_t)15);
            for (int64_t o = b0 + (int64_t)tid * 16; o < v1; o += RC_THREADS * 16) *(uint4 *)(mask + o) = z;
            for (int64_t o = v1 + tid; o < b1; o += RC_THREADS) mask[o] = 0;
        } else {
            for (int64_t o = b0 + tid; o < b1; o += RC_THREADS) mask[o] = 0;
        }
    }
    if (!valid) {                                                // uniform over the cluster: nobody syncs
        if (rank == 0) for (int k = tid; k < SVB_HIST_STRIDE; k += RC_THREADS) hist[(int64_t)f * SVB_HIST_STRIDE + k] = 0;
        return;
    }
    for (int k = tid; k < SVB_HIST_STRIDE; k += RC_THREADS) s_h[k] = 0;
    if (tid < 256) s_Z[tid] = tid ? exact_Z(cam, tid) : 0.0;
    if (tid == 0) { s_k = 0; s_r = 0; }
    __syncthreads();

    // B: my part of the records.  Estimate e = |Z (af x + bf y + c2) - 1| (4 flops); it decides unless it
    // lands within `band` (~1e5 x the worst-case rounding difference to the reference's operation
    // sequence, bounded once per frame) of the cutoff; then the exact sequence decides.
    double af, bf, c2, cut_lo, cut_hi;
    {
        const double a = pl[0], b = pl[1], c = pl[2], cutoff = pl[4];
        af = a / cam.f; bf = b / cam.f;
        c2 = c - af * cam.cw - bf * cam.ch;
        const double band = 1e-11 * fma(s_Z[1], fabs(af) * (double)(g.W + cam.cw) + fabs(bf) * (double)(g.H + cam.ch) + fabs(c) + fabs(c2), 1.0);
        cut_lo = cutoff - band; cut_hi = cutoff + band;
    }
    // The near points are scattered over the lanes (about a third of the records), so their expensive part
    // -- hue bin, histogram, reprojection -- runs DENSE: near records are first compacted into the warp's
    // ring (ballot + popc), and whenever 32 are waiting every lane takes one.
    uint64_t *stage = s_stage[warp];
    uint32_t *wlist = s_list + (size_t)warp * warp_cap;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t fill = 0, head = 0;                                 // warp-uniform: records waiting, ring position
    int wcnt = 0;                                                // warp-uniform: near points in the warp's list
    auto dense = [&](uint64_t rc, bool on) {
        if (on) {
            const uint32_t lo = (uint32_t)rc, hi = (uint32_t)(rc >> 32);
            const uint32_t x = lo & 0x7FFu, y = (lo >> 11) & 0x7FFu, d = (lo >> 22) & 0xFFu;
            const uint32_t hue = (uint32_t)hue_of_hi(hi);
            atomicAdd(&s_h[hue], 1u);
            const uint32_t px = x - ((__ldg(rbits + x * 8 + (d >> 5)) >> (d & 31)) & 1u);
            const uint32_t py = y - ((__ldg(rbits + (g.W + y) * 8 + (d >> 5)) >> (d & 31)) & 1u);
            wlist[wcnt + lane] = (px & 0x7FFu) | ((py & 0x7FFu) << 11) | (hue << 22);
        }
    };
    for (int i0 = begin; i0 < end; i0 += RC_PARTS * RC_ROUND) {
        uint64_t r[RC_PER_ROUND];
#pragma unroll
        for (int k = 0; k < RC_PER_ROUND; ++k) r[k] = nx[k];
        if (i0 + RC_PARTS * RC_ROUND < end) load_round(i0 + RC_PARTS * RC_ROUND);
#pragma unroll
        for (int k = 0; k < RC_PER_ROUND; ++k) {
            const int i = i0 + toff + (k >> 1) * 64 + (k & 1);
            bool near = false;
            if (i < end) {
                const uint32_t lo = (uint32_t)r[k];
                const double Z = s_Z[(lo >> 22) & 0xFFu];
                const double e = fabs(fma(Z, fma(af, road_u2d(lo & 0x7FFu), fma(bf, road_u2d((lo >> 11) & 0x7FFu), c2)), -1.0));
                near = e < cut_lo;
                if (!near && e < cut_hi) near = exact_near(cam, lo, Z, pl[0], pl[1], pl[2], pl[4]);   // rare
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, near);
            if (near) stage[(head + fill + __popc(bal & lt_mask)) & 63u] = r[k];
            fill += __popc(bal);
            if (fill >= 32u) {
                __syncwarp();
                dense(stage[(head + lane) & 63u], true);
                wcnt += 32; head += 32u; fill -= 32u;
                __syncwarp();
            }
        }
    }
    __syncwarp();
    if (fill) dense(stage[(head + lane) & 63u], (uint32_t)lane < fill);
    wcnt += (int)fill;
    __syncthreads();
    cluster_arrive();
    cluster_wait();

    // C: full histogram = sum of the 8 partial ones, read through distributed shared memory
    for (int k = tid; k < SVB_HIST_STRIDE; k += RC_THREADS) {
        uint32_t t = 0;
#pragma unroll
        for (int rr = 0; rr < RC_PARTS; ++rr) t += cluster.map_shared_rank(s_h, rr)[k];
        s_tot[k] = t;
        if (rank == 0) hist[(int64_t)f * SVB_HIST_STRIDE + k] = t;
    }
    __syncthreads();
    cluster_arrive();

    // D: colour filter + scatter of the warp's near list
    uint8_t *mask = road_mask + (int64_t)f * g.H * g.W;
    uint8_t *ovl = OVERLAY ? overlay + (int64_t)f * g.H * g.W * 3 : nullptr;
    int rcount = 0;
    for (int k = lane; k < wcnt; k += 32) {
        uint32_t w = wlist[k];
        if (s_tot[w >> 22] > colour_threshold) {
            ++rcount;
            int px = (int)(w & 0x7FFu), py = (int)((w >> 11) & 0x7FFu);
            if (px < g.W && py < g.H) {
                mask[(int64_t)py * g.W + px] = 255;
                if (OVERLAY) {
                    uint8_t *q = ovl + ((int64_t)py * g.W + px) * 3;
                    q[0] = 0; q[1] = 255; q[2] = 0;               // [0,255,0], stereovision.py:133
                }
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rcount += __shfl_xor_sync(0xffffffffu, rcount, o);
    if (lane == 0) {
        if (wcnt) atomicAdd(&s_k, wcnt);
        if (rcount) atomicAdd(&s_r, rcount);
    }
    __syncthreads();
    if (tid == 0) {
        if (s_k) atomicAdd(&counts_out[f * 4 + 2], s_k);
        if (s_r) atomicAdd(&counts_out[f * 4 + 3], s_r);
    }
    cluster_wait();
}

// ------------------------------------------------------------------ the streamed road path (any lattice size)
// Three launches for all F frames, no inter-CTA synchronisation inside a kernel:
//   road_params   (one CTA per frame)   pixel-form plane + candidate-test constants; zeroes the frame's histogram
//   road_near     (one CTA per 8192 records, every warp its own pipeline) bulk-copies (TMA) 256-record slices
//                                       into shared memory; per record one integer interval test on d; the
//                                       candidates (about a third) are compacted per warp and finished DENSE:
//                                       filtered float64 plane test, hue bin + histogram, reprojection -> one
//                                       word per near point, stored coalesced into the slice's part of the list
//   road_scatter  (4 slices per warp)   colour filter of the listed points, 255 into the mask
// A record is read from HBM once; a near point costs 4 more bytes written and read back.
#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#define RS_CHUNK 2048                    // records per CTA
#define RS_PER_WARP (RS_CHUNK / RS_WARPS)

struct StreamGeom {
    int H, W;
    int64_t slab;        // records per frame slab (= lattice size)
    int chunks;          // RS_CHUNK-record chunks per frame
    int ctas;            // CTAs per frame in road_near (RS_NCH chunks each)
    int sctas;           // CTAs per frame in road_scatter
    int step;
    int64_t mask_per;    // mask bytes cleared by one road_near CTA (multiple of 16)
};
#define RS_NCH 4                         // chunks per road_near CTA (double-buffered)
#define RS_FP_STRIDE 8                   // doubles per frame in the parameter block, see road_params_kernel

static StreamGeom make_stream_geom(int H, int W, int step) {
    StreamGeom g;
    g.H = H; g.W = W; g.step = step;
    int LH = svb_lattice_rows(H, step), LW = svb_lattice_cols(W, step);
    g.slab = (int64_t)LH * LW;
    g.chunks = (int)((g.slab + RS_CHUNK - 1) / RS_CHUNK);
    if (g.chunks < 1) g.chunks = 1;
    g.ctas = (g.chunks + RS_NCH - 1) / RS_NCH;
    g.sctas = (g.chunks * RS_WARPS + RS_WARPS * 4 - 1) / (RS_WARPS * 4);      // RS_SC_SLICES slices per warp
    int64_t bytes = (int64_t)H * W;
    g.mask_per = ((bytes + g.ctas - 1) / g.ctas + 15) & ~(int64_t)15;
    return g;
}

// the frame's plane in pixel form:  P.abc - 1 = Z(d) (af x + bf y + c2) - 1
struct PixPlane { double af, bf, c2, cut_lo, cut_hi; };
__device__ __forceinline__ PixPlane pix_plane(const double *pl, const CamD &cam, int H, int W) {
    PixPlane q;
    const double a = pl[0], b = pl[1], c = pl[2], cutoff = pl[4];
    q.af = a / cam.f; q.bf = b / cam.f;
    q.c2 = c - q.af * cam.cw - q.bf * cam.ch;
    // ~1e5 x the worst-case rounding difference between this factored form and the reference's operation sequence
    const double band = 1e-11 * fma(cam.fB, fabs(q.af) * (double)(W + cam.cw) + fabs(q.bf) * (double)(H + cam.ch) + fabs(c) + fabs(q.c2), 1.0);
    q.cut_lo = cutoff - band; q.cut_hi = cutoff + band;
    return q;
}

// Per frame: the pixel-form plane and what the candidate test of road_near needs (fparam), Z(d) (ztab,
// block 0); zeroes the frame's histogram and its K / R counters.
//   fparam[f] = af, bf, c2, cut_lo, cut_hi, fB/(1+cut) (1-1e-9), fB/(1-cut) (1+1e-9), bounds_ok
// Candidate test.  From |Z t - 1| < cut with Z = fB/d, t = af x + bf y + c2:  t > 0 and
// fB t/(1+cut) < d < fB t/(1-cut).  road_near bounds t over the pixel box of a 256-record slice (t is linear,
// so its extremes sit at the box corners) and keeps the records whose d lies in the resulting interval.  Only a
// SUPERSET of the near sites is needed -- the candidates are tested again with the filtered float64 predicate --
// so the bounds are widened by 1e-9 relative and anything unusual (bounds_ok = 0) keeps every d.
__global__ void __launch_bounds__(256)
road_params_kernel(const double *__restrict__ planes, StreamGeom g, CamD cam,
                   double *__restrict__ fparam, double *__restrict__ ztab,
                   uint32_t *__restrict__ hist, int32_t *__restrict__ counts_out) {
    const int f = blockIdx.x;
    if (f == 0) ztab[threadIdx.x] = threadIdx.x ? exact_Z(cam, threadIdx.x) : 0.0;
    for (int k = threadIdx.x; k < SVB_HIST_STRIDE; k += blockDim.x) hist[(int64_t)f * SVB_HIST_STRIDE + k] = 0;
    if (threadIdx.x != 0) return;
    counts_out[f * 4 + 2] = 0; counts_out[f * 4 + 3] = 0;
    const double *pl = planes + (int64_t)f * SVB_PLANE_STRIDE;
    double *fp = fparam + (int64_t)f * RS_FP_STRIDE;
    if (pl[7] == 0.0) {
        for (int k = 0; k < RS_FP_STRIDE; ++k) fp[k] = 0.0;
        return;
    }
    const PixPlane q = pix_plane(pl, cam, g.H, g.W);
    const double cut = q.cut_hi;
    const bool ok = cut > 0.0 && cut < 0.5 && isfinite(q.af) && isfinite(q.bf) && isfinite(q.c2);
    fp[0] = q.af; fp[1] = q.bf; fp[2] = q.c2; fp[3] = q.cut_lo; fp[4] = q.cut_hi;
    fp[5] = ok ? cam.fB / (1.0 + cut) * (1.0 - 1e-9) : 0.0;
    fp[6] = ok ? cam.fB / (1.0 - cut) * (1.0 + 1e-9) : 0.0;
    fp[7] = ok ? 1.0 : 0.0;
}

// bulk (TMA) copy helpers: one elected thread copies global -> shared and an mbarrier counts the bytes
__device__ __forceinline__ uint32_t rs_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void rs_mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rs_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void rs_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rs_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void rs_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(rs_smem_u32(dst)), "l"(src), "r"(bytes), "r"(rs_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void rs_mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(rs_smem_u32(bar)), "r"(parity) : "memory");
}

// Every WARP runs its own two-stage pipeline: lane 0 bulk-copies (TMA) the warp's next 256-record slice
// into the warp's second buffer while the warp tests the current one; an mbarrier per buffer counts the
// bytes.  No block-wide barrier in the loop -- a warp that meets many near points does not hold up the
// other seven.  A CTA covers RS_NCH chunks = 32 slices; warp w takes slices w, w + 8, w + 16, w + 24.
#define RS_SLICES (RS_NCH * RS_WARPS)    // slices per CTA
__device__ __forceinline__ void rs_sts64(uint32_t addr, uint64_t v) {
    asm volatile("st.shared.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t rs_lds64(uint32_t addr) {
    uint64_t v;
    asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr) : "memory");
    return v;
}

__global__ void __launch_bounds__(RS_THREADS, 4)
road_near_kernel(const uint64_t *__restrict__ cloud, const int32_t *__restrict__ counts,
                 const double *__restrict__ planes, StreamGeom g, CamD cam,
                 const uint32_t *__restrict__ rbits,
                 const double *__restrict__ fparam, const double *__restrict__ ztab,
                 uint32_t *__restrict__ hist, uint8_t *__restrict__ road_mask,
                 uint32_t *__restrict__ list, int32_t *__restrict__ warp_counts, int32_t *__restrict__ counts_out) {
    __shared__ __align__(128) uint64_t s_rec[RS_WARPS][2][RS_PER_WARP];
    __shared__ double s_Z[256];
    __shared__ uint64_t s_stage[RS_WARPS][64];
    __shared__ uint32_t s_h[SVB_HIST_STRIDE];
    __shared__ __align__(8) uint64_t s_bar[RS_WARPS][2];
    const int f = blockIdx.y, cta = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double *pl = planes + (int64_t)f * SVB_PLANE_STRIDE;
    const bool valid = pl[7] != 0.0;
    const int n = valid ? counts[f * 4 + 0] : 0;
    const int slice0 = cta * RS_SLICES;                           // first slice of this CTA
    const int frame_slices = g.chunks * RS_WARPS;
    const uint64_t *frame = cloud + (int64_t)f * g.slab;
    const bool bulk = ((((uintptr_t)frame) & 15) == 0);           // slice starts are multiples of 2 KB from here
    const uint32_t rec_addr = rs_smem_u32(&s_rec[warp][0][0]), stage_addr = rs_smem_u32(&s_stage[warp][0]);

    // records of the warp's i-th slice -> buffer i & 1: one bulk copy (the even part), an odd straggler by a plain load
    auto issue = [&](int i) {
        const int begin = (slice0 + warp + i * RS_WARPS) * RS_PER_WARP, m = min(RS_PER_WARP, n - begin);
        if (m <= 0) return;
        const int mb = bulk ? (m & ~1) : 0;
        if (lane == 0 && mb > 0) {
            rs_mbar_expect_tx(&s_bar[warp][i & 1], (uint32_t)mb * 8u);
            rs_bulk_g2s(&s_rec[warp][i & 1][0], frame + begin, (uint32_t)mb * 8u, &s_bar[warp][i & 1]);
        }
        for (int k = mb + lane; k < m; k += 32) s_rec[warp][i & 1][k] = __ldcs((const unsigned long long *)(frame + begin + k));
    };
    if (lane == 0) {
        rs_mbar_init(&s_bar[warp][0], 1);
        rs_mbar_init(&s_bar[warp][1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    issue(0);
    for (int k = tid; k < SVB_HIST_STRIDE; k += RS_THREADS) s_h[k] = 0;
    s_Z[tid] = __ldg(ztab + tid);

    // this CTA's share of the mask clear (replaces a memset launch)
    {
        const int64_t bytes = (int64_t)g.H * g.W;
        uint8_t *mask = road_mask + (int64_t)f * bytes;
        const int64_t b0 = min(bytes, g.mask_per * cta), b1 = min(bytes, b0 + g.mask_per);
        if ((((uintptr_t)mask) & 15) == 0) {
            const uint4 z = make_uint4(0, 0, 0, 0);
            const int64_t v1 = b0 + ((b1 - b0) & ~(int64_t)15);
            for (int64_t o = b0 + (int64_t)tid * 16; o < v1; o += RS_THREADS * 16) *(uint4 *)(mask + o) = z;
            for (int64_t o = v1 + tid; o < b1; o += RS_THREADS) mask[o] = 0;
        } else {
            for (int64_t o = b0 + tid; o < b1; o += RS_THREADS) mask[o] = 0;
        }
    }
    __syncthreads();                                              // histogram zeroed, Z(d) staged

    const double *fp = fparam + (int64_t)f * RS_FP_STRIDE;
    const double af = fp[0], bf = fp[1], c2 = fp[2], cut_lo = fp[3], cut_hi = fp[4];
    const uint32_t lt_mask = (1u << lane) - 1u;
    int ktotal = 0;

    for (int i = 0; i < RS_NCH; ++i) {
        const int slice = slice0 + warp + i * RS_WARPS;
        if (slice >= frame_slices) break;
        const int begin = slice * RS_PER_WARP, m = min(RS_PER_WARP, n - begin);
        int wcnt = 0;                                             // warp-uniform: near points listed so far
        if (m > 0) {
            if (i + 1 < RS_NCH) issue(i + 1);                     // the other buffer: this warp finished with it a slice ago
            if (bulk && (m & ~1) > 0) rs_mbar_wait(&s_bar[warp][i & 1], (uint32_t)(i >> 1) & 1u);
            __syncwarp();                                         // plain-load stragglers
            const uint32_t buf = rec_addr + (uint32_t)(i & 1) * (RS_PER_WARP * 8u) + (uint32_t)lane * 8u;
            uint32_t *wlist = list + (int64_t)f * g.slab + begin;
            uint32_t fill = 0, head = 0;                          // warp-uniform: candidates waiting, ring position
            // DENSE step: every lane takes one waiting candidate
            auto dense = [&](uint64_t rc, bool on) {
                bool near = false;
                uint32_t word = 0;
                if (on) {
                    const uint32_t lo = (uint32_t)rc, hi = (uint32_t)(rc >> 32);
                    const uint32_t x = lo & 0x7FFu, y = (lo >> 11) & 0x7FFu, d = lo >> 22;
                    const double Z = s_Z[d];
                    const double e = fabs(fma(Z, fma(af, road_u2d(x), fma(bf, road_u2d(y), c2)), -1.0));
                    near = e < cut_lo;
                    if (!near && e < cut_hi) near = exact_near(cam, lo, Z, pl[0], pl[1], pl[2], pl[4]);   // rare
                    if (near) {
                        const uint32_t hue = (uint32_t)hue_of_hi(hi);
                        atomicAdd(&s_h[hue], 1u);
                        const uint32_t px = x - ((__ldg(rbits + x * 8 + (d >> 5)) >> (d & 31)) & 1u);
                        const uint32_t py = y - ((__ldg(rbits + (g.W + y) * 8 + (d >> 5)) >> (d & 31)) & 1u);
                        word = px | (py << 11) | (hue << 22);
                    }
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, near);
                if (near) __stcs(wlist + wcnt + __popc(bal & lt_mask), word);
                wcnt += __popc(bal);
            };
            // the slice's candidate interval [dlo, dlo + dw] (warp-uniform): the records are in row-major order,
            // so the first and the last one bound the slice's pixel box
            uint32_t dlo = 1u, dw = 254u;                         // keep everything unless the bounds are clean
            {
                const uint32_t la = (uint32_t)rs_lds64(buf - (uint32_t)lane * 8u);
                const uint32_t lb = (uint32_t)rs_lds64(buf - (uint32_t)lane * 8u + (uint32_t)(m - 1) * 8u);
                const uint32_t ya = (la >> 11) & 0x7FFu, yb = (lb >> 11) & 0x7FFu;
                const double xa = ya == yb ? road_u2d(la & 0x7FFu) : 0.0, xb = ya == yb ? road_u2d(lb & 0x7FFu) : (double)(g.W - 1);
                const double t1 = af * xa, t2 = af * xb, u1 = bf * road_u2d(ya), u2 = bf * road_u2d(yb);
                const double tmin = c2 + fmin(t1, t2) + fmin(u1, u2), tmax = c2 + fmax(t1, t2) + fmax(u1, u2);
                if (fp[7] != 0.0 && isfinite(tmin) && isfinite(tmax)) {
                    if (tmax <= 0.0) { dlo = 255u; dw = 0u; }     // Z t <= 0: |Z t - 1| >= 1 > cut, nothing is near
                    else {
                        const double Al = fmin(fmax(floor(fmax(tmin, 0.0) * fp[5]), 1.0), 255.0);
                        const double Bh = fmax(fmin(ceil(tmax * fp[6]), 255.0), Al);
                        dlo = (uint32_t)Al; dw = (uint32_t)Bh - dlo;
                    }
                }
            }
            auto round = [&](uint64_t r, bool cand) {
                const uint32_t bal = __ballot_sync(0xffffffffu, cand);
                if (cand) rs_sts64(stage_addr + (((head + fill + __popc(bal & lt_mask)) & 63u) << 3), r);
                fill += __popc(bal);
                if (fill >= 32u) {
                    __syncwarp();
                    dense(rs_lds64(stage_addr + (((head + lane) & 63u) << 3)), true);
                    head += 32u; fill -= 32u;
                    __syncwarp();
                }
            };
            if (m == RS_PER_WARP) {                               // full slice: no bounds to test
#pragma unroll 2
                for (int k = 0; k < RS_PER_WARP / 32; ++k) {
                    const uint64_t r = rs_lds64(buf + (uint32_t)k * 256u);
                    round(r, ((uint32_t)r >> 22) - dlo <= dw);
                }
            } else {
                for (int k = 0; k * 32 < m; ++k) {
                    uint64_t r = 0;
                    bool cand = false;
                    if (k * 32 + lane < m) {
                        r = rs_lds64(buf + (uint32_t)k * 256u);
                        cand = ((uint32_t)r >> 22) - dlo <= dw;
                    }
                    round(r, cand);
                }
            }
            __syncwarp();
            if (fill) dense(rs_lds64(stage_addr + (((head + lane) & 63u) << 3)), (uint32_t)lane < fill);
            __syncwarp();
        }
        if (lane == 0) warp_counts[(int64_t)f * frame_slices + slice] = wcnt;
        ktotal += wcnt;
    }
    if (lane == 0 && ktotal) atomicAdd(&counts_out[f * 4 + 2], ktotal);
    __syncthreads();
    if (!valid) return;
    uint32_t *gh = hist + (int64_t)f * SVB_HIST_STRIDE;
    for (int k = tid; k < SVB_HUE_BINS; k += RS_THREADS) {
        const uint32_t v = s_h[k];
        if (v) atomicAdd(&gh[k], v);
    }
}

// Colour filter + scatter.  A warp takes RS_SC_SLICES consecutive slices and has all of their list entries
// (up to 8 words per lane and slice) in flight at once: the kernel is a chain of three dependent memory
// round trips (slice counts, entries, mask stores), so the only lever is how much each trip carries.
#define RS_SC_SLICES 4
template <bool OVERLAY>
__global__ void __launch_bounds__(RS_THREADS)
road_scatter_kernel(const int32_t *__restrict__ warp_counts, const uint32_t *__restrict__ list, StreamGeom g,
                    const int32_t *__restrict__ counts_k, const uint32_t *__restrict__ hist, uint32_t colour_threshold,
                    uint8_t *__restrict__ road_mask, uint8_t *__restrict__ overlay, int32_t *__restrict__ counts_out) {
    __shared__ uint32_t s_pop[32];                                // bit per hue bin: count > colour_threshold
    const int f = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (counts_k[f * 4 + 2] == 0) return;                         // no near point in this frame
    const int slices = g.chunks * RS_WARPS;
    const int sl0 = (blockIdx.x * RS_WARPS + warp) * RS_SC_SLICES;
    int cnt[RS_SC_SLICES];
#pragma unroll
    for (int q = 0; q < RS_SC_SLICES; ++q) cnt[q] = sl0 + q < slices ? __ldg(warp_counts + (int64_t)f * slices + sl0 + q) : 0;
    const uint32_t *gh = hist + (int64_t)f * SVB_HIST_STRIDE;
#pragma unroll
    for (int k = 0; k < SVB_HIST_STRIDE / RS_THREADS; ++k) {
        const int bin = k * RS_THREADS + tid;
        const uint32_t bal = __ballot_sync(0xffffffffu, __ldg(gh + bin) > colour_threshold);
        if (lane == 0) s_pop[bin >> 5] = bal;
    }
    uint32_t w[RS_SC_SLICES][RS_PER_WARP / 32];
    const uint32_t *wl = list + (int64_t)f * g.slab + (int64_t)sl0 * RS_PER_WARP + lane;
#pragma unroll
    for (int q = 0; q < RS_SC_SLICES; ++q)
#pragma unroll
        for (int k = 0; k < RS_PER_WARP / 32; ++k)
            w[q][k] = (k * 32 + lane < cnt[q]) ? __ldcs(wl + q * RS_PER_WARP + k * 32) : 0xFFFFFFFFu;   // hue 1023: never popular
    __syncthreads();
    uint8_t *mask = road_mask + (int64_t)f * g.H * g.W;
    uint8_t *ovl = OVERLAY ? overlay + (int64_t)f * g.H * g.W * 3 : nullptr;
    int rcount = 0;
#pragma unroll
    for (int q = 0; q < RS_SC_SLICES; ++q)
#pragma unroll
        for (int k = 0; k < RS_PER_WARP / 32; ++k) {
            const uint32_t v = w[q][k], hue = v >> 22;
            if ((s_pop[hue >> 5] >> (hue & 31)) & 1u) {
                ++rcount;
                const uint32_t off = ((v >> 11) & 0x7FFu) * (uint32_t)g.W + (v & 0x7FFu);
                mask[off] = 255;
                if (OVERLAY) {
                    uint8_t *o = ovl + (int64_t)off * 3;
                    o[0] = 0; o[1] = 255; o[2] = 0;               // [0,255,0], stereovision.py:133
                }
            }
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rcount += __shfl_xor_sync(0xffffffffu, rcount, o);
    if (lane == 0 && rcount) atomicAdd(&counts_out[f * 4 + 3], rcount);
}

static int64_t stream_fp_bytes(int F) { return ((int64_t)F * RS_FP_STRIDE + 256) * 8; }            // frame parameters + Z(d) table
static int64_t stream_wc_bytes(int F, const StreamGeom &g) { return (((int64_t)F * g.chunks * RS_WARPS * 4) + 255) & ~(int64_t)255; }
static int64_t stream_list_bytes(int F, const StreamGeom &g) { return (int64_t)F * g.slab * 4; }

extern "C" int64_t svb_road_scratch_bytes(int F, int H, int W, int step) {
    if (F < 1 || H < 2 || W < 2 || step < 1) return 0;
    RoadGeom g = make_road_geom(H, W, step);
    StreamGeom sg = make_stream_geom(H, W, step);
    int64_t legacy = (int64_t)F * g.words * (int64_t)sizeof(uint32_t);
    int64_t streamed = stream_fp_bytes(F) + stream_wc_bytes(F, sg) + stream_list_bytes(F, sg);
    return legacy > streamed ? legacy : streamed;
}

extern "C" int svb_road_points(const uint64_t *cloud, const int32_t *counts, const double *planes,
                               int F, int H, int W, int step, const svb_camera *cam,
                               uint32_t colour_threshold, const uint32_t *reproj_bits,
                               uint32_t *hist, uint8_t *road_mask, uint8_t *overlay, int32_t *counts_out,
                               void *scratch, int64_t scratch_bytes, void *stream) {
    if (!cloud || !counts || !planes || !cam || !hist || !road_mask || !counts_out || !scratch) return SVB_E_NULL;
    if (F < 1 || H < 2 || W < 2 || H > SVB_MAX_DIM || W > SVB_MAX_DIM || step < 1) return SVB_E_SHAPE;
    if (scratch_bytes < svb_road_scratch_bytes(F, H, W, step)) return SVB_E_SCRATCH;
    cudaStream_t st = (cudaStream_t)stream;
    RoadGeom g = make_road_geom(H, W, step);
    CamD c = make_camd(cam);
    const int64_t rounds = (g.slab + RC_ROUND - 1) / RC_ROUND;                   // rounds of a frame, dealt round-robin
    const int64_t list_cap = (rounds + RC_PARTS - 1) / RC_PARTS * RC_ROUND;      // records one CTA can meet
    const int warp_cap = (int)(list_cap / RC_WARPS);                             // ... and one warp
    const size_t list_bytes = (size_t)list_cap * sizeof(uint32_t);
    const bool cluster_fits = list_bytes <= 160 * 1024;
    // the streamed path (needs the displacement table and 16-byte aligned scratch)
    if (reproj_bits && (((uintptr_t)scratch) & 15) == 0 && (g_svb_road_mode == 3 || (g_svb_road_mode == 0 && !cluster_fits))) {
        StreamGeom sg = make_stream_geom(H, W, step);
        double *fparam = (double *)scratch;
        double *ztab = fparam + (int64_t)F * RS_FP_STRIDE;
        int32_t *wcounts = (int32_t *)((char *)fparam + stream_fp_bytes(F));
        uint32_t *list = (uint32_t *)((char *)wcounts + stream_wc_bytes(F, sg));
        {
            ProfScope ps(K_ROAD_TABLE, st);
            road_params_kernel<<<F, 256, 0, st>>>(planes, sg, c, fparam, ztab, hist, counts_out);
        }
        SVB_LAUNCH_CHECK();
        {
            ProfScope ps(K_ROAD_NEAR, st);
            road_near_kernel<<<dim3(sg.ctas, F), RS_THREADS, 0, st>>>(cloud, counts, planes, sg, c, reproj_bits, fparam, ztab,
                                                                    hist, road_mask, list, wcounts, counts_out);
        }
        SVB_LAUNCH_CHECK();
        {
            ProfScope ps(K_ROAD_SCATTER, st);
            dim3 sgrid(sg.sctas, F);
            if (overlay) road_scatter_kernel<true><<<sgrid, RS_THREADS, 0, st>>>(wcounts, list, sg, counts_out, hist, colour_threshold, road_mask, overlay, counts_out);
            else         road_scatter_kernel<false><<<sgrid, RS_THREADS, 0, st>>>(wcounts, list, sg, counts_out, hist, colour_threshold, road_mask, nullptr, counts_out);
        }
        SVB_LAUNCH_CHECK();
        return SVB_OK;
    }
    // the cluster kernel: displacement table given and a CTA's near list fits in shared memory
    if (reproj_bits && cluster_fits && (g_svb_road_mode == 0 || g_svb_road_mode == 2)) {
        cudaError_t e = cudaMemset2DAsync(counts_out + 2, 4 * sizeof(int32_t), 0, 2 * sizeof(int32_t), F, st);
        if (e != cudaSuccess) return (int)e;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(RC_PARTS * F);
        cfg.blockDim = dim3(RC_THREADS);
        cfg.dynamicSmemBytes = list_bytes;
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = RC_PARTS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        ProfScope ps(K_ROAD_CLUSTER, st);
        if (overlay) {
            e = cudaFuncSetAttribute(road_cluster_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)list_bytes);
            if (e == cudaSuccess)
                e = cudaLaunchKernelEx(&cfg, road_cluster_kernel<true>, cloud, counts, planes, g, c, colour_threshold,
                                       reproj_bits, hist, road_mask, overlay, counts_out, warp_cap);
        } else {
            e = cudaFuncSetAttribute(road_cluster_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)list_bytes);
            if (e == cudaSuccess)
                e = cudaLaunchKernelEx(&cfg, road_cluster_kernel<false>, cloud, counts, planes, g, c, colour_threshold,
                                       reproj_bits, hist, road_mask, (uint8_t *)nullptr, counts_out, warp_cap);
        }
        if (e != cudaSuccess) return (int)e;
        SVB_LAUNCH_CHECK();
        return SVB_OK;
    }
    {
        ProfScope pc(K_ROAD_CLEAR, st);
        cudaError_t e;
        e = cudaMemsetAsync(hist, 0, (size_t)F * SVB_HIST_STRIDE * sizeof(uint32_t), st);
        if (e != cudaSuccess) return (int)e;
        e = cudaMemsetAsync(road_mask, 0, (size_t)F * H * W, st);
        if (e != cudaSuccess) return (int)e;
        e = cudaMemset2DAsync(counts_out + 2, 4 * sizeof(int32_t), 0, 2 * sizeof(int32_t), F, st);
        if (e != cudaSuccess) return (int)e;
    }
    dim3 grid(g.parts, F);
    uint32_t *nearbits = (uint32_t *)scratch;
    {
        ProfScope ps(K_ROAD_PASS1, st);
        road_pass1_kernel<<<grid, ROAD_THREADS, 0, st>>>(cloud, counts, planes, g, c, nearbits, hist, counts_out);
    }
    SVB_LAUNCH_CHECK();
    ProfScope ps2(K_ROAD_PASS2, st);
#define LAUNCH_P2(O, T)                                                                                        \
    road_pass2_kernel<O, T><<<grid, ROAD_THREADS, 0, st>>>(cloud, counts, planes, g, c, colour_threshold, nearbits, \
                                                           hist, reproj_bits, road_mask, overlay, counts_out)
    if (overlay) { if (reproj_bits) LAUNCH_P2(true, true); else LAUNCH_P2(true, false); }
    else         { if (reproj_bits) LAUNCH_P2(false, true); else LAUNCH_P2(false, false); }
#undef LAUNCH_P2
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
