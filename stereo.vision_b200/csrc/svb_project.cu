// a1-a4: temporal disparity fill, mask, disparity -> 3-D cloud with stable stream compaction.
//
// Layout.  A frame's lattice is LH x LW sites (every `step`-th pixel of every `step`-th row,
// functions.py:185-186).  One thread owns one 16-pixel chunk of one lattice row: it reads the 16
// disparity bytes (one 128-bit load), the 16 mask bytes and, for rows that carry colour, the 48
// BGR bytes (three 128-bit loads), and turns them into a 16-bit "valid" and a 16-bit "masked
// valid" bitmap.  Threads of a block are laid out row-major over a group of lattice rows, so an
// exclusive block scan of the popcounts gives every site its rank in the reference's append order.
// Records are compacted into (swizzled) shared memory first and leave the SM as fully coalesced 8-byte
// (cloud) and 4-byte (mcloud) stores.  The hue bin is not stored: only the near-plane points need it
// (functions.py:215-230), so the road kernels compute it from the record's colour word.
#include "svb_common.cuh"
#include "svb_prof.cuh"

#define PROJ_THREADS 256
#define CHUNK 16

// ------------------------------------------------------------------ a2 mask
__global__ void mask_disp_kernel(const uint8_t *__restrict__ d, const uint8_t *__restrict__ cm,
                                 uint8_t *__restrict__ out, int64_t pixels, int64_t total) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    out[i] = cm[i % pixels] != 0 ? d[i] : (uint8_t)0;
}
__global__ void mask_disp_v16(const uint4 *__restrict__ d, const uint4 *__restrict__ cm,
                              uint4 *__restrict__ out, int64_t chunks, int64_t total) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    uint4 v = __ldcs(d + i);
    uint4 m = __ldg(cm + i % chunks);
    // 0xff where mask byte != 0
    v.x &= ~__vcmpeq4(m.x, 0u);
    v.y &= ~__vcmpeq4(m.y, 0u);
    v.z &= ~__vcmpeq4(m.z, 0u);
    v.w &= ~__vcmpeq4(m.w, 0u);
    __stcs(out + i, v);
}

extern "C" int svb_mask_disparity(const uint8_t *d, const uint8_t *carmask, uint8_t *out,
                                  int F, int H, int W, void *stream) {
    if (!d || !carmask || !out) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1) return SVB_E_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t pixels = (int64_t)H * W;
    bool aligned = (pixels % 16 == 0) && (((uintptr_t)d | (uintptr_t)carmask | (uintptr_t)out) % 16 == 0);
    ProfScope ps(K_MASK, st);
    if (aligned) {
        int64_t chunks = pixels / 16, total = chunks * F;
        mask_disp_v16<<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const uint4 *)d, (const uint4 *)carmask,
                                                                      (uint4 *)out, chunks, total);
    } else {
        int64_t total = pixels * F;
        mask_disp_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d, carmask, out, pixels, total);
    }
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------ a4 projection + compaction
struct ProjGeom {
    int H, W, step, LH, LW;
    int cpr;            // 16-pixel chunks per row
    int rpb;            // lattice rows per block
    int groups;         // row groups (blocks) per frame
    int cap;            // most sites one block can emit
    int vec_ok;         // W % 16 == 0 and all bases 16-byte aligned
};

static inline int lattice_len(int n, int step) { return n <= 1 ? 0 : (n - 1 + step - 1) / step; }

static ProjGeom make_geom(int H, int W, int step, bool aligned) {
    ProjGeom g;
    g.H = H; g.W = W; g.step = step;
    g.LH = lattice_len(H, step);
    g.LW = lattice_len(W, step);
    g.cpr = (W + CHUNK - 1) / CHUNK;
    g.rpb = PROJ_THREADS / g.cpr;
    if (g.rpb < 1) g.rpb = 1;
    g.groups = (g.LH + g.rpb - 1) / g.rpb;
    g.cap = (g.rpb * g.cpr * ((CHUNK + step - 1) / step) + 31) & ~31;      // multiple of 32: the staging swizzles
    g.vec_ok = (W % CHUNK == 0) && aligned;
    return g;
}

// The 16 disparity / mask bytes of one chunk and the two site bitmaps.
struct ChunkBits {
    uint32_t d[4];
    uint32_t vm, mm;     // valid sites, valid sites inside carmask (bit j = pixel x0 + j)
};

__device__ __forceinline__ uint32_t byte_of(const uint32_t *w, int j) { return (w[j >> 2] >> ((j & 3) * 8)) & 0xFFu; }

// nonzero-byte bitmap of 4 packed bytes -> 4 bits.  Bit 7 of every byte of ((w & 0x7f..) + 0x7f..) | w says "byte != 0"
// (no carry leaves a byte); the product with 2^21 + 2^14 + 2^7 + 1 brings bits 7, 15, 23, 31 to bits 28..31 without
// two terms meeting in one bit -- five instructions, one of them on the fma pipe (the emulated __vcmpne4 and the three
// shift-ors it replaced were 14 % of proj_write's instructions, and the kernel is bound by the alu pipe).
__device__ __forceinline__ uint32_t nz4(uint32_t w) {
    const uint32_t t = (((w & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | w) & 0x80808080u;
    return (t * 0x00204081u) >> 28;
}
// ... of 16 packed bytes -> 16 bits (bit j = byte j)
__device__ __forceinline__ uint32_t nz16(const uint32_t *w) {
    return nz4(w[0]) + nz4(w[1]) * 16u + nz4(w[2]) * 256u + nz4(w[3]) * 4096u;
}

// lattice sites among the 16 pixels starting at x0 (x % step == 0 and x < W-1)
__device__ __forceinline__ uint32_t site_bits(int x0, int W, int step) {
    uint32_t site;
    if (step == 1) site = 0xFFFFu;
    else if (step == 2) site = 0x5555u;                 // x0 is a multiple of 16, hence even
    else {
        site = 0;
#pragma unroll
        for (int j = 0; j < CHUNK; ++j) if ((x0 + j) % step == 0) site |= 1u << j;
    }
    int lim = W - 1 - x0;                               // pixels of this chunk with x < W-1
    if (lim < CHUNK) site &= (lim <= 0) ? 0u : ((1u << lim) - 1u);
    return site;
}

template <bool HAS_MASK>
__device__ __forceinline__ ChunkBits load_chunk(const uint8_t *__restrict__ drow, const uint8_t *__restrict__ mrow,
                                                int x0, int W, int step, bool vec_ok) {
    ChunkBits c;
    uint32_t m[4];
    if (vec_ok) {
        uint4 v = __ldg((const uint4 *)(drow + x0));
        c.d[0] = v.x; c.d[1] = v.y; c.d[2] = v.z; c.d[3] = v.w;
        if (HAS_MASK) {
            uint4 q = __ldg((const uint4 *)(mrow + x0));
            m[0] = q.x; m[1] = q.y; m[2] = q.z; m[3] = q.w;
        }
    } else {
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            uint32_t dv = 0, mv = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int x = x0 + w * 4 + k;
                if (x < W) {
                    dv |= (uint32_t)drow[x] << (k * 8);
                    if (HAS_MASK) mv |= (uint32_t)mrow[x] << (k * 8);
                }
            }
            c.d[w] = dv;
            m[w] = mv;
        }
    }
    uint32_t nzd = nz16(c.d);
    uint32_t site = site_bits(x0, W, step);
    c.vm = nzd & site;
    if (HAS_MASK) {
        c.mm = c.vm & nz16(m);
    } else {
        c.mm = c.vm;
    }
    return c;
}

// ------------------------------------------------------------------ a1 fill
// d' = sat_add(d, (d <= 2) ? prev : 0) on 4 bytes at once           functions.py:142-146
// Written out on the word (the video intrinsics __vcmpleu4 / __vaddus4 are emulated on this part: ~20 instructions, all on
// the alu pipe the kernel is bound by; this is 12, two of them multiplies):
//   bit 7 of (d | 0x80) - 3 is clear iff (d & 0x7f) < 3, and no byte borrows (each is >= 0x80), so hb = "d <= 2" per byte;
//   in a hole byte d <= 2, so (prev & 0x7f) + d <= 0x81 does not carry into the next byte (elsewhere the addend is 0);
//   with prev's bit 7 put back the sum reaches 256 exactly when both bit 7s are set: saturate those bytes (cv2.add).
__device__ __forceinline__ uint32_t fill4(uint32_t d, uint32_t prev) {
    const uint32_t hb = ~(((d | 0x80808080u) - 0x03030303u) | d) & 0x80808080u;    // 0x80 where d <= 2
    const uint32_t m = (hb >> 7) * 0xFFu;                                          // 0xff where d <= 2
    const uint32_t s7 = (prev & m & 0x7F7F7F7Fu) + d;
    const uint32_t ph = prev & hb;
    return s7 | ph | (((s7 & ph) >> 7) * 0xFFu);
}
// One thread = 16 pixels for all F frames; the loads of the next U frames are issued back to back
// (they do not depend on the recurrence) and then consumed in order.  A rolling window of loads does
// not work here: a warp tracks outstanding loads with six scoreboards, so waiting for the oldest of
// twelve loads also waits for younger ones.
// COUNT: the kernel also leaves, per (frame, row group), the number of valid / masked-valid lattice
// sites of the FILLED frame, i.e. what proj_count_kernel would compute from `filled` -- the bytes are
// already in registers here, so the projection needs no counting pass of its own.  Site and carmask
// membership are kept as 0x80-per-byte masks, so a word's counts are two popc.
#define FILL_THREADS 64
__device__ __forceinline__ uint32_t nz80(uint32_t w) { return ((w & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | w; }   // bit 7 of a byte: byte != 0
// nonzero-byte map of 16 bytes in one word: bit 8 j + k = byte j of word k is nonzero (k = 0..3)
__device__ __forceinline__ uint32_t merge_nz(uint4 v) {
    return ((nz80(v.x) >> 7) & 0x01010101u) | ((nz80(v.y) >> 6) & 0x02020202u) |
           ((nz80(v.z) >> 5) & 0x04040404u) | ((nz80(v.w) >> 4) & 0x08080808u);
}
__device__ __forceinline__ uint4 fill16(uint4 d, uint4 p) {
    return make_uint4(fill4(d.x, p.x), fill4(d.y, p.y), fill4(d.z, p.z), fill4(d.w, p.w));
}
template <int U, bool COUNT>
__global__ void __launch_bounds__(FILL_THREADS)
fill_seq_v16(const uint4 *__restrict__ raw, const uint4 *__restrict__ prev0, uint4 *__restrict__ filled,
             int F, int64_t chunks, const uint4 *__restrict__ carmask, ProjGeom g, uint32_t *__restrict__ group_counts) {
    int64_t p = (int64_t)blockIdx.x * FILL_THREADS + threadIdx.x;
    const bool active = p < chunks;
    if (!active) p = chunks - 1;                        // keep the lane for the shuffles; it stores nothing
    bool have = prev0 != nullptr;
    uint4 prev = have ? __ldg(prev0 + p) : make_uint4(0, 0, 0, 0);

    // counting set-up: is this chunk on a lattice row, which group, which bytes are sites / inside carmask
    bool lat = false, uni = false;
    int grp = 0;
    uint4 site80 = make_uint4(0, 0, 0, 0), in80 = make_uint4(0, 0, 0, 0);
    if (COUNT) {
        int y = (int)(p / g.cpr), x0 = (int)(p % g.cpr) * CHUNK;
        int ly = y / g.step;
        lat = active && (y % g.step == 0) && ly < g.LH;
        grp = ly / g.rpb;
        if (lat) {
            // bit 8 j + k of the merged word = pixel 4 k + j of the chunk (see count below)
            const uint32_t sb = site_bits(x0, g.W, g.step);
            uint32_t sm = 0;
#pragma unroll
            for (int px = 0; px < 16; ++px)
                if (sb & (1u << px)) sm |= 1u << (8 * (px & 3) + (px >> 2));
            site80.x = sm;
            const uint4 m = carmask ? __ldg(carmask + p) : make_uint4(~0u, ~0u, ~0u, ~0u);
            in80.x = merge_nz(m) & sm;
        }
        int lat0 = __shfl_sync(0xffffffffu, (int)lat, 0), grp0 = __shfl_sync(0xffffffffu, grp, 0);
        uni = __all_sync(0xffffffffu, (int)lat == lat0 && (!lat || grp == grp0));
    }
    auto count = [&](const uint4 &v, int t) {
        if (uni && !lat) return;                        // whole warp off the lattice (odd rows at step 2)
        const uint32_t nz = merge_nz(v);
        uint32_t c = (uint32_t)__popc(nz & site80.x) | ((uint32_t)__popc(nz & in80.x) << 16);
        if (uni) {
            c = __reduce_add_sync(0xffffffffu, c);          // one REDUX; 32 x 8 sites fit the 16-bit halves
            if ((threadIdx.x & 31) == 0 && c) atomicAdd(group_counts + (int64_t)t * g.groups + grp, c);
        } else if (lat && c) {
            atomicAdd(group_counts + (int64_t)t * g.groups + grp, c);
        }
    };

    // two batches of U frames in flight: the loads of batch t + U are issued before batch t is consumed (the kernel
    // has only ~7 warps per SM -- one thread walks all F frames of its 16 pixels -- so the bytes in flight per
    // thread are what fills the memory pipe)
    int t = 0;
    uint4 d[U], e[U];
    if (U <= F) {
#pragma unroll
        for (int u = 0; u < U; ++u) d[u] = __ldcs(raw + (int64_t)u * chunks + p);
    }
    const int lane = threadIdx.x & 31;
    auto consume = [&](const uint4 (&q)[U], int t0) {
        uint32_t cpk[U];                                // my sites / masked sites of the batch's frames, 16-bit halves
#pragma unroll
        for (int u = 0; u < U; ++u) {
            uint4 v = have ? fill16(q[u], prev) : q[u];
            if (active) __stcs(filled + (int64_t)(t0 + u) * chunks + p, v);
            prev = v;
            have = true;
            if (COUNT) {
                if (uni) {
                    const uint32_t nz = lat ? merge_nz(v) : 0u;
                    cpk[u] = (uint32_t)__popc(nz & site80.x) | ((uint32_t)__popc(nz & in80.x) << 16);
                } else count(v, t0 + u);
            }
        }
        if (COUNT && uni && lat) {
            // the warp's sums of the U frames: U REDUX back to back, lane u keeps frame u's, one RED for the batch
            uint32_t mine = 0;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t r = __reduce_add_sync(0xffffffffu, cpk[u]);   // 32 x 8 sites fit the 16-bit halves
                if (lane == u) mine = r;
            }
            if (lane < U && mine) atomicAdd(group_counts + (int64_t)(t0 + lane) * g.groups + grp, mine);
        }
    };
    while (t + U <= F) {                                // d holds batch t
        const bool more = t + 2 * U <= F;
        if (more) {
#pragma unroll
            for (int u = 0; u < U; ++u) e[u] = __ldcs(raw + (int64_t)(t + U + u) * chunks + p);
        }
        consume(d, t);
        t += U;
        if (!more) break;
        const bool more2 = t + 2 * U <= F;              // e holds batch t
        if (more2) {
#pragma unroll
            for (int u = 0; u < U; ++u) d[u] = __ldcs(raw + (int64_t)(t + U + u) * chunks + p);
        }
        consume(e, t);
        t += U;
        if (!more2) break;
    }
    for (; t < F; ++t) {
        uint4 dt = __ldcs(raw + (int64_t)t * chunks + p);
        uint4 v = have ? fill16(dt, prev) : dt;
        if (active) __stcs(filled + (int64_t)t * chunks + p, v);
        prev = v;
        have = true;
        if (COUNT) count(v, t);
    }
}

// any-size fallback, one pixel per thread
__global__ void fill_seq_scalar(const uint8_t *__restrict__ raw, const uint8_t *__restrict__ prev0,
                                uint8_t *__restrict__ filled, int F, int64_t pixels) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pixels) return;
    bool have = prev0 != nullptr;
    int prev = have ? prev0[p] : 0;
    for (int t = 0; t < F; ++t) {
        int d = raw[(int64_t)t * pixels + p];
        if (have && d <= 2) d = min(d + prev, 255);
        filled[(int64_t)t * pixels + p] = (uint8_t)d;
        prev = d;
        have = true;
    }
}

#define FILL_UNROLL 16          // frames per batch; 4 / 8 / 12 / 16: 97 / 94 / 90 / 98 us per 300 frames, the step is the same (12 only wins because 300 = 25 x 12)
// ------------------------------------------------------------------ a1 fill + a4 counts, step 2, balanced
// fill_seq_v16<.., COUNT> is bound by its slowest warps: a thread walks all F frames of its pixels, so the kernel
// lasts as long as one warp's serial instruction stream (a warp issues once per ~5.6 cycles on the dependent
// SIMD-in-word chain, ncu), and the warps on lattice rows carry the counting on top of the fill.  Here the row kind
// decides the pixels per thread -- PE on lattice (even) rows, PO on odd rows -- so that more, shorter threads sit
// where the counting is.  Warps never straddle rows (W % (32 max(PE, PO)) == 0).
template <int P> struct PixVec;
template <> struct PixVec<4>  { typedef uint32_t T; };
template <> struct PixVec<8>  { typedef uint2 T; };
template <> struct PixVec<16> { typedef uint4 T; };
template <int P> struct Words { uint32_t w[P / 4]; };
template <int P>
__device__ __forceinline__ Words<P> ld_words(const uint8_t *p, bool streaming) {
    typedef typename PixVec<P>::T V;
    union { V v; Words<P> w; } u;
    u.v = streaming ? __ldcs(reinterpret_cast<const V *>(p)) : __ldg(reinterpret_cast<const V *>(p));
    return u.w;
}
template <int P>
__device__ __forceinline__ void st_words(uint8_t *p, const Words<P> &w) {
    typedef typename PixVec<P>::T V;
    union { V v; Words<P> w; } u;
    u.w = w;
    __stcs(reinterpret_cast<V *>(p), u.v);
}
// sites of a lattice row at step 2 = even x: bytes 0 and 2 of every word; bit 8 / 24 of (w & 0x00FF00FF) + 0x00FF00FF
// says non-zero.  The P/4 words' bits are interleaved into one word (shifts 0 .. 3).
template <int P>
__device__ __forceinline__ uint32_t even_nonzero_bits(const Words<P> &v) {
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < P / 4; ++i) r |= (((v.w[i] & 0x00FF00FFu) + 0x00FF00FFu) & 0x01000100u) << i;
    return r;
}

// one thread's P pixels over all F frames; COUNT: leave the (valid | masked valid << 16) site counts of every frame
template <int U, int P, bool COUNT>
__device__ __forceinline__ void fill_walk(const uint8_t *__restrict__ raw, const uint8_t *__restrict__ prev0, uint8_t *__restrict__ filled,
                                          int F, int64_t frame, int64_t off, uint32_t in_bits, bool lat, int lane,
                                          uint32_t *__restrict__ counts_row, int groups) {
    bool have = prev0 != nullptr;
    Words<P> prev;
#pragma unroll
    for (int i = 0; i < P / 4; ++i) prev.w[i] = 0;
    if (have) prev = ld_words<P>(prev0 + off, false);
    const uint8_t *src = raw + off;
    uint8_t *dst = filled + off;
    auto step = [&](Words<P> v, int t) -> uint32_t {
        if (have) {
#pragma unroll
            for (int i = 0; i < P / 4; ++i) v.w[i] = fill4(v.w[i], prev.w[i]);
        }
        st_words<P>(dst + (int64_t)t * frame, v);
        prev = v;
        have = true;
        if (!COUNT) return 0u;
        const uint32_t nz = even_nonzero_bits<P>(v);
        return (uint32_t)__popc(nz) | ((uint32_t)__popc(nz & in_bits) << 16);
    };
    auto consume = [&](const Words<P> (&q)[U], int t0) {
        uint32_t cpk[U];
#pragma unroll
        for (int u = 0; u < U; ++u) cpk[u] = step(q[u], t0 + u);
        if (COUNT && lat) {
            // the warp's sums of the U frames: U REDUX back to back, lane u keeps frame u's, one RED for the batch
            uint32_t mine = 0;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t r = __reduce_add_sync(0xffffffffu, cpk[u]);      // at most 32 x 8 sites: fits the 16-bit halves
                if (lane == u) mine = r;
            }
            if (lane < U && mine) atomicAdd(counts_row + (int64_t)(t0 + lane) * groups, mine);
        }
    };
    int t = 0;
    Words<P> d[U], e[U];
    if (U <= F) {
#pragma unroll
        for (int u = 0; u < U; ++u) d[u] = ld_words<P>(src + (int64_t)u * frame, true);
    }
    while (t + U <= F) {                                               // d holds batch t; the next batch is already in flight
        const bool more = t + 2 * U <= F;
        if (more) {
#pragma unroll
            for (int u = 0; u < U; ++u) e[u] = ld_words<P>(src + (int64_t)(t + U + u) * frame, true);
        }
        consume(d, t);
        t += U;
        if (!more) break;
        const bool more2 = t + 2 * U <= F;
        if (more2) {
#pragma unroll
            for (int u = 0; u < U; ++u) d[u] = ld_words<P>(src + (int64_t)(t + U + u) * frame, true);
        }
        consume(e, t);
        t += U;
        if (!more2) break;
    }
    // the remaining n < U frames as one partial batch: their loads are issued together (a tail of single frames pays a
    // load-to-use latency per frame)
    const int n = F - t;
    if (n > 0) {
#pragma unroll
        for (int u = 0; u < U; ++u) if (u < n) d[u] = ld_words<P>(src + (int64_t)(t + u) * frame, true);
        uint32_t mine = 0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (u < n) {
                uint32_t c = step(d[u], t + u);
                if (COUNT && lat) {
                    c = __reduce_add_sync(0xffffffffu, c);
                    if (lane == u) mine = c;
                }
            }
        }
        if (COUNT && lat && lane < n && mine) atomicAdd(counts_row + (int64_t)(t + lane) * groups, mine);
    }
}

#define FS_WARPS 4                                      // warps per CTA (warps are independent)
template <int U, int PE, int PO>
__global__ void __launch_bounds__(FS_WARPS * 32)
fill_seq_split(const uint8_t *__restrict__ raw, const uint8_t *__restrict__ prev0, uint8_t *__restrict__ filled, int F,
               int H, int W, const uint8_t *__restrict__ carmask, ProjGeom g, uint32_t *__restrict__ group_counts) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = (int64_t)blockIdx.x * FS_WARPS + (threadIdx.x >> 5);
    const int wpe = W / (32 * PE), wpo = W / (32 * PO), wpp = wpe + wpo;   // warps per even row, odd row, row pair
    const int64_t rp = wid / wpp;
    const int k = (int)(wid - rp * wpp);
    if (rp * 2 >= H) return;
    const int64_t frame = (int64_t)H * W;
    if (k < wpe) {
        const int y = (int)rp * 2, x0 = (k * 32 + lane) * PE;
        const int64_t off = (int64_t)y * W + x0;
        const int ly = y / 2;
        const bool lat = ly < g.LH;                                    // the last even row of an odd-height image is no lattice row
        uint32_t in_bits = 0xFFFFFFFFu;
        if (carmask) in_bits = even_nonzero_bits<PE>(ld_words<PE>(carmask + off, false));
        fill_walk<U, PE, true>(raw, prev0, filled, F, frame, off, in_bits, lat, lane, group_counts + ly / g.rpb, g.groups);
    } else {
        const int y = (int)rp * 2 + 1;
        if (y >= H) return;
        const int x0 = ((k - wpe) * 32 + lane) * PO;
        fill_walk<U, PO, false>(raw, prev0, filled, F, frame, (int64_t)y * W + x0, 0u, false, lane, nullptr, 0);
    }
}

static int launch_fill(const uint8_t *raw, const uint8_t *prev0, uint8_t *filled, int F, int H, int W,
                       const uint8_t *carmask, const ProjGeom *g, uint32_t *group_counts, cudaStream_t st) {
    int64_t pixels = (int64_t)H * W;
    bool aligned = (pixels % 16 == 0) && (((uintptr_t)raw | (uintptr_t)filled | (uintptr_t)prev0) % 16 == 0);
    if (!aligned) {
        if (g) return SVB_E_ARG;                        // the caller checks can_count_in_fill() first
        int threads = 256;
        int64_t blocks = (pixels + threads - 1) / threads;
        fill_seq_scalar<<<(unsigned)blocks, threads, 0, st>>>(raw, prev0, filled, F, pixels);
        return SVB_OK;
    }
    int64_t chunks = pixels / 16;
    unsigned blocks = (unsigned)((chunks + FILL_THREADS - 1) / FILL_THREADS);
    if (g && g->step == 2 && W % 512 == 0 && (carmask == nullptr || ((uintptr_t)carmask % 16) == 0)) {
#define LAUNCH_SPLIT(PE, PO)                                                                                              \
        do {                                                                                                              \
            const int64_t warps = (int64_t)((H + 1) / 2) * (W / (32 * PE) + W / (32 * PO));                               \
            fill_seq_split<FILL_UNROLL, PE, PO><<<(unsigned)((warps + FS_WARPS - 1) / FS_WARPS), FS_WARPS * 32, 0, st>>>( \
                raw, prev0, filled, F, H, W, carmask, *g, group_counts);                                                  \
        } while (0)
        // pixels per thread on lattice / odd rows.  Measured on 300 frames of 1024 x 544: (8, 8) 87.6 us, (8, 16) 94.1,
        // (4, 8) 88.6, (4, 4) 92.1
        LAUNCH_SPLIT(8, 8);
#undef LAUNCH_SPLIT
        return SVB_OK;
    }
    if (g)
        fill_seq_v16<FILL_UNROLL, true><<<blocks, FILL_THREADS, 0, st>>>((const uint4 *)raw, (const uint4 *)prev0, (uint4 *)filled,
                                                                         F, chunks, (const uint4 *)carmask, *g, group_counts);
    else
        fill_seq_v16<FILL_UNROLL, false><<<blocks, FILL_THREADS, 0, st>>>((const uint4 *)raw, (const uint4 *)prev0, (uint4 *)filled,
                                                                          F, chunks, nullptr, ProjGeom(), nullptr);
    return SVB_OK;
}

extern "C" int svb_fill_sequence(const uint8_t *raw, const uint8_t *prev0, uint8_t *filled,
                                 int F, int H, int W, void *stream) {
    if (!raw || !filled) return SVB_E_NULL;
    if (F < 1 || H < 1 || W < 1) return SVB_E_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope ps(K_FILL, st);
    int rc = launch_fill(raw, prev0, filled, F, H, W, nullptr, nullptr, nullptr, st);
    if (rc != SVB_OK) return rc;
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// exclusive scan of one packed (cloud | mcloud << 16) count per thread over the 256-thread block
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *warp_tot, uint32_t *total) {
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < PROJ_THREADS / 32; ++w) {
        uint32_t t = warp_tot[w];
        if (w < wid) base += t;
        tot += t;
    }
    *total = tot;
    return base + inc - v;
}

// pass 1: per (frame, row group) site counts
template <bool HAS_MASK>
__global__ void __launch_bounds__(PROJ_THREADS)
proj_count_kernel(const uint8_t *__restrict__ disp, const uint8_t *__restrict__ carmask, ProjGeom g,
                  uint32_t *__restrict__ group_counts) {
    __shared__ uint32_t warp_tot[PROJ_THREADS / 32];
    int f = blockIdx.y, grp = blockIdx.x;
    int r = threadIdx.x / g.cpr, ch = threadIdx.x % g.cpr;
    int ly = grp * g.rpb + r;
    uint32_t cnt = 0;
    if (r < g.rpb && ly < g.LH) {
        int y = ly * g.step;
        const uint8_t *drow = disp + ((int64_t)f * g.H + y) * g.W;
        const uint8_t *mrow = HAS_MASK ? carmask + (int64_t)y * g.W : nullptr;
        ChunkBits c = load_chunk<HAS_MASK>(drow, mrow, ch * CHUNK, g.W, g.step, g.vec_ok);
        cnt = (uint32_t)__popc(c.vm) | ((uint32_t)__popc(c.mm) << 16);
    }
    // block reduce
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < PROJ_THREADS / 32; ++w) t += warp_tot[w];
        group_counts[(int64_t)f * g.groups + grp] = t;
    }
}

// pass 2: exclusive scan of the group counts of each frame (one block per frame)
__global__ void proj_scan_kernel(uint32_t *__restrict__ group_counts, int groups, int32_t *__restrict__ counts) {
    __shared__ uint32_t sv[32], sm[32];
    int f = blockIdx.x;
    uint32_t *gc = group_counts + (int64_t)f * groups;
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint32_t carry_v = 0, carry_m = 0;
    for (int base = 0; base < groups; base += blockDim.x) {
        int i = base + threadIdx.x;
        uint32_t p = i < groups ? gc[i] : 0u;
        int v = (int)(p & 0xFFFFu), m = (int)(p >> 16);
        int iv = warp_incl_scan(v, lane), im = warp_incl_scan(m, lane);
        if (lane == 31) { sv[wid] = iv; sm[wid] = im; }
        __syncthreads();
        uint32_t bv = 0, bm = 0, tv = 0, tm = 0;
        for (int w = 0; w < nw; ++w) {
            if (w < wid) { bv += sv[w]; bm += sm[w]; }
            tv += sv[w]; tm += sm[w];
        }
        // the offsets need more than 16 bits: the cloud offset replaces the packed count in plane 0
        // of the scratch array, the mcloud offset goes to plane 1 ([F][groups] each)
        if (i < groups) {
            uint32_t *gm = group_counts + (int64_t)gridDim.x * groups + (int64_t)f * groups;
            gc[i] = carry_v + bv + iv - v;
            gm[i] = carry_m + bm + im - m;
        }
        carry_v += tv; carry_m += tm;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        counts[f * 4 + 0] = (int32_t)carry_v;
        counts[f * 4 + 1] = (int32_t)carry_m;
    }
}

// pass 3: write the records
// Shared-memory staging is swizzled: a thread appends a run of up to CHUNK/step records, so without
// it the lanes of a warp would hit the same two (8-byte list) or four (4-byte list) bank groups.
__device__ __forceinline__ int swz8(int p) { return p ^ ((p >> 4) & 15); }      // 8-byte slots, 16 per bank row
__device__ __forceinline__ int swz4(int p) { return p ^ ((p >> 5) & 31); }      // 4-byte slots, 32 per bank row

// colour word R | G << 8 | B << 16 of pixel J of a chunk whose 48 BGR bytes sit in px[12]
template <int J>
__device__ __forceinline__ uint32_t colour_of(const uint32_t *px) {
    constexpr int b0 = 3 * J, w = b0 >> 2, o = b0 & 3;            // B at byte b0, G at b0+1, R at b0+2
    constexpr int w1 = w + 1 < 12 ? w + 1 : 11;
    if constexpr (o == 0) return __byte_perm(px[w], 0u, 0x4012);                       // (0, B, G, R) <- bytes 2,1,0
    else if constexpr (o == 1) return __byte_perm(px[w], 0u, 0x4123);
    else if constexpr (o == 2) return __byte_perm(px[w], px[w1], 0x0234) & 0x00FFFFFFu;   // R = byte 0 of the next word
    else return __byte_perm(px[w], px[w1], 0x0345) & 0x00FFFFFFu;                      // G, R = bytes 0, 1 of the next word
}

template <int STEP, bool HAS_MASK, bool HAS_RGB, int J>
__device__ __forceinline__ void emit_site(const ChunkBits &c, const uint32_t *px, uint32_t base_lo,
                                          uint64_t *s_cloud, uint32_t *s_m, int &ov, int &om) {
    if (c.vm & (1u << J)) {
        uint32_t d = (c.d[J >> 2] >> ((J & 3) * 8)) & 0xFFu;
        uint32_t lo = base_lo + (uint32_t)J + (d << 22);         // x0 + J never carries into y (x0 + 15 < 2048)
        uint32_t hi = HAS_RGB ? colour_of<J>(px) : 0u;
        s_cloud[swz8(ov)] = (uint64_t)lo | ((uint64_t)hi << 32);
        ++ov;
        if (HAS_MASK) {
            if (c.mm & (1u << J)) { s_m[swz4(om)] = lo; ++om; }
        } else {
            s_m[swz4(om)] = lo; ++om;
        }
    }
}

template <int STEP, bool HAS_MASK, bool HAS_RGB>
__global__ void __launch_bounds__(PROJ_THREADS)
proj_write_kernel(const uint8_t *__restrict__ disp, const uint8_t *__restrict__ carmask,
                  const uint8_t *__restrict__ bgr, ProjGeom g, const uint32_t *__restrict__ group_counts,
                  int32_t *__restrict__ counts, uint64_t *__restrict__ cloud, uint32_t *__restrict__ mcloud) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t warp_tot[PROJ_THREADS / 32];
    __shared__ uint32_t s_base[2];                               // records of this frame in the row groups before mine
    const int cap = g.cap;                                       // sites a block can hold (multiple of 32)
    uint64_t *s_cloud = (uint64_t *)smem_raw;                    // [cap]
    uint32_t *s_m = (uint32_t *)(smem_raw + (size_t)cap * 8);    // [cap]

    int f = blockIdx.y, grp = blockIdx.x;
    // my offset in the frame's slabs = the (cloud | mcloud << 16) counts of the row groups before mine, summed here
    // (at most a few hundred words per frame) instead of by a scan kernel of its own between the count and this one
    if (threadIdx.x < 2) s_base[threadIdx.x] = 0;
    uint32_t pv = 0, pm = 0;
    for (int i = threadIdx.x; i < grp; i += PROJ_THREADS) {
        const uint32_t p = __ldg(group_counts + (int64_t)f * g.groups + i);
        pv += p & 0xFFFFu;
        pm += p >> 16;
    }
    int r = threadIdx.x / g.cpr, ch = threadIdx.x % g.cpr;
    int ly = grp * g.rpb + r;
    ChunkBits c;
    c.vm = c.mm = 0;
    int y = 0, x0 = ch * CHUNK;
    bool active = (r < g.rpb && ly < g.LH);
    uint32_t px[12];                                             // 48 BGR bytes of the chunk
    if (active) {
        y = ly * g.step;
        const uint8_t *drow = disp + ((int64_t)f * g.H + y) * g.W;
        const uint8_t *mrow = HAS_MASK ? carmask + (int64_t)y * g.W : nullptr;
        if (HAS_RGB && g.vec_ok) {                               // issue the colour loads together with d / mask
            const uint4 *q = (const uint4 *)(bgr + (((int64_t)f * g.H + y) * g.W + x0) * 3);
            uint4 a = __ldcs(q), b = __ldcs(q + 1), d = __ldcs(q + 2);
            px[0] = a.x; px[1] = a.y; px[2] = a.z; px[3] = a.w;
            px[4] = b.x; px[5] = b.y; px[6] = b.z; px[7] = b.w;
            px[8] = d.x; px[9] = d.y; px[10] = d.z; px[11] = d.w;
        }
        c = load_chunk<HAS_MASK>(drow, mrow, x0, g.W, g.step, g.vec_ok);
        if (HAS_RGB && !g.vec_ok) {
            const uint8_t *crow = bgr + (((int64_t)f * g.H + y) * g.W + x0) * 3;
#pragma unroll
            for (int w = 0; w < 12; ++w) {
                uint32_t v = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    int bi = w * 4 + k;
                    if (x0 + bi / 3 < g.W) v |= (uint32_t)crow[bi] << (k * 8);
                }
                px[w] = v;
            }
        }
    }
    uint32_t cnt = (uint32_t)__popc(c.vm) | ((uint32_t)__popc(c.mm) << 16);
    uint32_t total;
    uint32_t off = block_excl_scan(cnt, warp_tot, &total);
    int ov = (int)(off & 0xFFFFu), om = (int)(off >> 16);
    int tv = (int)(total & 0xFFFFu), tm = (int)(total >> 16);
    pv = __reduce_add_sync(0xffffffffu, pv);                     // (the barrier inside the scan orders s_base's zeroing before this)
    pm = __reduce_add_sync(0xffffffffu, pm);
    if ((threadIdx.x & 31) == 0 && (pv | pm)) { atomicAdd(&s_base[0], pv); atomicAdd(&s_base[1], pm); }

    if (c.vm) {
        const uint32_t base_lo = pack_xyd(x0, y, 0);
#define SITE(J) emit_site<STEP, HAS_MASK, HAS_RGB, J>(c, px, base_lo, s_cloud, s_m, ov, om)
        if (STEP == 2) {                                         // only even pixels are lattice sites
            SITE(0); SITE(2); SITE(4); SITE(6); SITE(8); SITE(10); SITE(12); SITE(14);
        } else {
            SITE(0); SITE(1); SITE(2); SITE(3); SITE(4); SITE(5); SITE(6); SITE(7);
            SITE(8); SITE(9); SITE(10); SITE(11); SITE(12); SITE(13); SITE(14); SITE(15);
        }
#undef SITE
    }
    __syncthreads();

    const uint32_t base_v = s_base[0], base_m = s_base[1];
    if (grp == (int)gridDim.x - 1 && threadIdx.x == 0) {        // the last row group knows the frame's totals: N, M
        counts[f * 4 + 0] = (int32_t)(base_v + (uint32_t)tv);
        counts[f * 4 + 1] = (int32_t)(base_m + (uint32_t)tm);
    }
    int64_t slab = (int64_t)f * g.LH * g.LW;
    if (cloud) {
        uint64_t *dst = cloud + slab + base_v;
        // swz8 only touches the low four bits, with bits 4-7 of the index: swz8(t + 256 k) = swz8(t) + 256 k
        const unsigned long long *src = (const unsigned long long *)s_cloud + swz8((int)threadIdx.x) - (int)threadIdx.x;
        for (int i = threadIdx.x; i < tv; i += PROJ_THREADS) __stcs((unsigned long long *)dst + i, src[i]);
    }
    if (mcloud) {
        uint32_t *dst = mcloud + slab + base_m;
        for (int i = threadIdx.x; i < tm; i += PROJ_THREADS) dst[i] = s_m[swz4(i)];
    }
}

extern "C" int svb_lattice_rows(int H, int step) { return step < 1 ? 0 : lattice_len(H, step); }
extern "C" int svb_lattice_cols(int W, int step) { return step < 1 ? 0 : lattice_len(W, step); }

extern "C" int64_t svb_project_scratch_bytes(int F, int H, int W, int step) {
    if (F < 1 || H < 2 || W < 2 || step < 1) return 0;
    ProjGeom g = make_geom(H, W, step, true);
    return (int64_t)2 * F * g.groups * (int64_t)sizeof(uint32_t);
}

static bool can_count_in_fill(int H, int W, const void *a, const void *b, const void *c, const void *d) {
    return ((int64_t)H * W) % 16 == 0 && W % CHUNK == 0 &&
           (((uintptr_t)a | (uintptr_t)b | (uintptr_t)c | (uintptr_t)d) % 16 == 0);
}

extern "C" int svb_fill_sequence_counted(const uint8_t *raw, const uint8_t *prev0, uint8_t *filled,
                                         const uint8_t *carmask, int F, int H, int W, int step,
                                         void *scratch, int64_t scratch_bytes, void *stream) {
    if (!raw || !filled || !scratch) return SVB_E_NULL;
    if (F < 1 || H < 2 || W < 2 || H > SVB_MAX_DIM || W > SVB_MAX_DIM || step < 1) return SVB_E_SHAPE;
    if (scratch_bytes < svb_project_scratch_bytes(F, H, W, step)) return SVB_E_SCRATCH;
    cudaStream_t st = (cudaStream_t)stream;
    ProjGeom g = make_geom(H, W, step, true);
    if (g.cpr > PROJ_THREADS) return SVB_E_SHAPE;
    uint32_t *gc = (uint32_t *)scratch;
    if (can_count_in_fill(H, W, raw, prev0, filled, carmask)) {
        ProfScope ps(K_FILL, st);
        cudaError_t e = cudaMemsetAsync(gc, 0, (size_t)F * g.groups * sizeof(uint32_t), st);
        if (e != cudaSuccess) return (int)e;
        int rc = launch_fill(raw, prev0, filled, F, H, W, carmask, &g, gc, st);
        if (rc != SVB_OK) return rc;
        SVB_LAUNCH_CHECK();
        return SVB_OK;
    }
    // shapes the vector kernel cannot take: plain fill, then the stand-alone counting pass
    int rc = svb_fill_sequence(raw, prev0, filled, F, H, W, stream);
    if (rc != SVB_OK) return rc;
    bool aligned = (((uintptr_t)filled | (uintptr_t)carmask) % 16 == 0);
    g = make_geom(H, W, step, aligned);
    dim3 grid(g.groups, F);
    ProfScope ps(K_PROJ_COUNT, st);
    if (carmask) proj_count_kernel<true><<<grid, PROJ_THREADS, 0, st>>>(filled, carmask, g, gc);
    else         proj_count_kernel<false><<<grid, PROJ_THREADS, 0, st>>>(filled, carmask, g, gc);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

static int project_compact_impl(const uint8_t *disp, const uint8_t *carmask, const uint8_t *bgr,
                                int F, int H, int W, int step, uint64_t *cloud, uint32_t *mcloud, int32_t *counts,
                                void *scratch, int64_t scratch_bytes, void *stream, bool counts_ready) {
    if (!disp || !counts || !scratch) return SVB_E_NULL;
    if (F < 1 || H < 2 || W < 2 || H > SVB_MAX_DIM || W > SVB_MAX_DIM || step < 1) return SVB_E_SHAPE;
    if (scratch_bytes < svb_project_scratch_bytes(F, H, W, step)) return SVB_E_SCRATCH;
    cudaStream_t st = (cudaStream_t)stream;
    bool aligned = (((uintptr_t)disp | (uintptr_t)carmask | (uintptr_t)bgr) % 16 == 0);
    ProjGeom g = make_geom(H, W, step, aligned);
    if (g.cpr > PROJ_THREADS) return SVB_E_SHAPE;
    uint32_t *gc = (uint32_t *)scratch;
    dim3 grid(g.groups, F);
    if (!counts_ready) {
        ProfScope ps(K_PROJ_COUNT, st);
        if (carmask) proj_count_kernel<true><<<grid, PROJ_THREADS, 0, st>>>(disp, carmask, g, gc);
        else         proj_count_kernel<false><<<grid, PROJ_THREADS, 0, st>>>(disp, carmask, g, gc);
        SVB_LAUNCH_CHECK();
    }
    if (!cloud && !mcloud) {                                     // counts only: N, M by the stand-alone scan
        ProfScope ps(K_PROJ_SCAN, st);
        proj_scan_kernel<<<F, 256, 0, st>>>(gc, g.groups, counts);
        SVB_LAUNCH_CHECK();
        return SVB_OK;
    }
    size_t smem = (size_t)g.cap * 12;
#define LAUNCH_W(S, M, R)                                                                           \
    do {                                                                                            \
        if (smem + 256 > 48 * 1024)                                                                 \
            cudaFuncSetAttribute(proj_write_kernel<S, M, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        proj_write_kernel<S, M, R><<<grid, PROJ_THREADS, smem, st>>>(disp, carmask, bgr, g, gc, counts, cloud, mcloud); \
    } while (0)
#define LAUNCH_WS(S)                                                                                \
    do {                                                                                            \
        if (carmask) { if (bgr) LAUNCH_W(S, true, true); else LAUNCH_W(S, true, false); }           \
        else         { if (bgr) LAUNCH_W(S, false, true); else LAUNCH_W(S, false, false); }         \
    } while (0)
    {
        ProfScope ps(K_PROJ_WRITE, st);
        if (step == 2) LAUNCH_WS(2); else LAUNCH_WS(0);
    }
#undef LAUNCH_WS
#undef LAUNCH_W
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_project_compact(const uint8_t *disp, const uint8_t *carmask, const uint8_t *bgr,
                                   int F, int H, int W, int step, const svb_camera *cam,
                                   uint64_t *cloud, uint32_t *mcloud, int32_t *counts,
                                   void *scratch, int64_t scratch_bytes, void *stream) {
    (void)cam;
    return project_compact_impl(disp, carmask, bgr, F, H, W, step, cloud, mcloud, counts, scratch, scratch_bytes, stream, false);
}

extern "C" int svb_project_compact_counted(const uint8_t *filled, const uint8_t *carmask, const uint8_t *bgr,
                                           int F, int H, int W, int step, const svb_camera *cam,
                                           uint64_t *cloud, uint32_t *mcloud, int32_t *counts,
                                           void *scratch, int64_t scratch_bytes, void *stream) {
    (void)cam;
    return project_compact_impl(filled, carmask, bgr, F, H, W, step, cloud, mcloud, counts, scratch, scratch_bytes, stream, true);
}

// ------------------------------------------------------------------ export of what the reference's list holds
__global__ void cloud_export_kernel(const uint32_t *__restrict__ rec, int stride, int n, CamD cam,
                                    double *__restrict__ xyz, float *__restrict__ xyz32, uint8_t *__restrict__ rgb,
                                    int32_t *__restrict__ sites, uint8_t *__restrict__ dout, uint16_t *__restrict__ hue) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t lo = rec[(int64_t)i * stride];
    uint32_t hi = stride > 1 ? rec[(int64_t)i * stride + 1] : 0u;
    int x = rec_x(lo), y = rec_y(lo), d = rec_d(lo);
    if (xyz || xyz32) {
        double Z = exact_Z(cam, d);
        double X = exact_XY(cam.cw, cam.f, x, Z);
        double Y = exact_XY(cam.ch, cam.f, y, Z);
        if (xyz) { xyz[3 * (int64_t)i] = X; xyz[3 * (int64_t)i + 1] = Y; xyz[3 * (int64_t)i + 2] = Z; }
        if (xyz32) { xyz32[3 * (int64_t)i] = (float)X; xyz32[3 * (int64_t)i + 1] = (float)Y; xyz32[3 * (int64_t)i + 2] = (float)Z; }
    }
    if (rgb) {
        rgb[3 * (int64_t)i] = (uint8_t)(hi & 0xFF);
        rgb[3 * (int64_t)i + 1] = (uint8_t)((hi >> 8) & 0xFF);
        rgb[3 * (int64_t)i + 2] = (uint8_t)((hi >> 16) & 0xFF);
    }
    if (sites) { sites[2 * (int64_t)i] = y; sites[2 * (int64_t)i + 1] = x; }
    if (dout) dout[i] = (uint8_t)d;
    if (hue) hue[i] = (uint16_t)hue_of_hi(hi);
}

extern "C" int svb_cloud_export(const uint32_t *records, int rec_stride_words, int n, const svb_camera *cam,
                                double *xyz, float *xyz_f32, uint8_t *rgb, int32_t *sites, uint8_t *d,
                                uint16_t *hue, void *stream) {
    if (n == 0) return SVB_OK;
    if (!records || !cam) return SVB_E_NULL;
    if (n < 0 || (rec_stride_words != 1 && rec_stride_words != 2)) return SVB_E_ARG;
    ProfScope ps(K_EXPORT, (cudaStream_t)stream);
    cloud_export_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(records, rec_stride_words, n, make_camd(cam),
                                                                            xyz, xyz_f32, rgb, sites, d, hue);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
