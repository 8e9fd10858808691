// ABI bookkeeping: version and error strings.
#include "svb_common.cuh"
#include <string.h>

extern "C" int svb_version(void) { return 15; }

extern "C" const char *svb_error_string(int code) {
    switch (code) {
        case SVB_OK: return "ok";
        case SVB_E_NULL: return "a required pointer is NULL";
        case SVB_E_SHAPE: return "invalid shape (F >= 1, 2 <= H,W <= 2048, step >= 1, T,S >= 1)";
        case SVB_E_ARG: return "invalid argument";
        case SVB_E_SCRATCH: return "scratch buffer too small";
        case SVB_E_LIMIT: return "size beyond the fixed maximum of this no-allocation entry point";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown error";
}

// ------------------------------------------------------------------ strided host -> device upload
extern "C" int svb_upload_rows(const void *h_src, void *d_dst, int64_t pitch_bytes, int64_t width_bytes, int64_t rows, void *stream) {
    if (!h_src || !d_dst) return SVB_E_NULL;
    if (pitch_bytes < width_bytes || width_bytes < 1 || rows < 1) return SVB_E_ARG;
    cudaError_t e = cudaMemcpy2DAsync(d_dst, (size_t)pitch_bytes, h_src, (size_t)pitch_bytes, (size_t)width_bytes, (size_t)rows,
                                      cudaMemcpyHostToDevice, (cudaStream_t)stream);
    return e == cudaSuccess ? SVB_OK : (int)e;
}

// The same for a block of rows of every frame of a batch: frames of `rows_per_frame` pitched rows each (the same geometry on
// both sides), rows [first, first + nrows) of each of `frames` frames, width_bytes of every row -- one 3-D copy.
extern "C" int svb_upload_row_block(const void *h_src, void *d_dst, int64_t pitch_bytes, int64_t width_bytes, int64_t rows_per_frame,
                                    int64_t first, int64_t nrows, int64_t frames, void *stream) {
    if (!h_src || !d_dst) return SVB_E_NULL;
    if (pitch_bytes < width_bytes || width_bytes < 1 || rows_per_frame < 1 || frames < 1 || first < 0 || nrows < 1 ||
        first + nrows > rows_per_frame) return SVB_E_ARG;
    cudaMemcpy3DParms p;
    memset(&p, 0, sizeof(p));
    p.srcPtr = make_cudaPitchedPtr(const_cast<void *>(h_src), (size_t)pitch_bytes, (size_t)pitch_bytes, (size_t)rows_per_frame);
    p.dstPtr = make_cudaPitchedPtr(d_dst, (size_t)pitch_bytes, (size_t)pitch_bytes, (size_t)rows_per_frame);
    p.srcPos = make_cudaPos(0, (size_t)first, 0);
    p.dstPos = make_cudaPos(0, (size_t)first, 0);
    p.extent = make_cudaExtent((size_t)width_bytes, (size_t)nrows, (size_t)frames);
    p.kind = cudaMemcpyHostToDevice;
    cudaError_t e = cudaMemcpy3DAsync(&p, (cudaStream_t)stream);
    return e == cudaSuccess ? SVB_OK : (int)e;
}

// ------------------------------------------------------------------ per-kernel event tracing
#include "svb_prof.cuh"
#include <vector>
#include <string.h>

bool g_svb_prof_on = false;
static const char *k_names[K_COUNT_] = {
    "fill_seq", "mask_disp", "proj_count", "proj_scan", "proj_write", "cloud_export", "ransac_score",
    "ransac_dump", "ransac_finalize", "point_errors", "compact_cloud", "hue_hist", "hue_bins", "reproject",
    "scatter_points", "road_pass1", "road_pass2", "road_clear", "road_cluster", "road_table", "road_near", "road_scatter",
    "clean_band", "pack_mask", "mask_points", "road_sites", "road_sites_scatter",
    "prep_gamma_gray", "prep_equalize_lut", "prep_apply_lut", "fill_row_mean", "normal_line", "road_hull", "hull_masks",
    "sgbm_prefilter", "sgbm_hsum", "sgbm_vsum", "sgbm_path_v", "sgbm_path_d1", "sgbm_path_d3", "sgbm_path_h", "sgbm_final",
    "sgbm_median", "speckle", "disparity_scale", "ransac_replay"};
const char *svb_prof_name(int id) { return (id >= 0 && id < K_COUNT_) ? k_names[id] : "svb"; }
struct EvPair { int id; cudaEvent_t a, b; };
static std::vector<EvPair> g_events;
static cudaEvent_t g_pending[K_COUNT_];

void svb_prof_begin(int id, cudaStream_t st) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, st);
    g_pending[id] = e;
}
void svb_prof_end(int id, cudaStream_t st) {
    cudaEvent_t e;
    if (!g_pending[id] || cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, st);
    g_events.push_back({id, g_pending[id], e});
    g_pending[id] = nullptr;
}

extern "C" int svb_profile_enable(int on) {
    for (auto &p : g_events) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    g_events.clear();
    memset(g_pending, 0, sizeof(g_pending));
    g_svb_prof_on = on != 0;
    return SVB_OK;
}
extern "C" int svb_profile_kernels(void) { return K_COUNT_; }
extern "C" int svb_profile_get(int id, char *name, int name_cap, double *ms_total, int *launches) {
    if (id < 0 || id >= K_COUNT_ || !name || !ms_total || !launches) return SVB_E_ARG;
    strncpy(name, k_names[id], name_cap > 0 ? name_cap - 1 : 0);
    if (name_cap > 0) name[name_cap - 1] = 0;
    double tot = 0; int n = 0;
    for (auto &p : g_events) {
        if (p.id != id) continue;
        float ms = 0;
        cudaError_t e = cudaEventElapsedTime(&ms, p.a, p.b);
        if (e != cudaSuccess) return (int)e;
        tot += ms; ++n;
    }
    *ms_total = tot; *launches = n;
    return SVB_OK;
}
