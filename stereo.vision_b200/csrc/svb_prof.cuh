// Per-kernel CUDA-event tracing (off by default).  When enabled through svb_profile_enable(1) every
// kernel launch of the library is bracketed by two events on the launching stream; svb_profile_get
// sums the elapsed times per kernel name.  bench.py uses it for the live roofline numbers.
// Every launch site is also an NVTX range named after the kernel ("NVTX ranges per stage"), so a timeline
// tool shows the path's stages without the event tracing; with no tool attached a push / pop is a few nanoseconds.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

enum SvbKernelId {
    K_FILL = 0, K_MASK, K_PROJ_COUNT, K_PROJ_SCAN, K_PROJ_WRITE, K_EXPORT, K_RANSAC_SCORE, K_RANSAC_DUMP,
    K_RANSAC_FINAL, K_POINT_ERR, K_COMPACT, K_HUE_HIST, K_HUE_BINS, K_REPROJECT, K_SCATTER, K_ROAD_PASS1,
    K_ROAD_PASS2, K_ROAD_CLEAR, K_ROAD_CLUSTER, K_ROAD_TABLE, K_ROAD_NEAR, K_ROAD_SCATTER, K_CLEAN_BAND,
    K_PACK_MASK, K_MASK_POINTS, K_ROAD_SITES, K_ROAD_SITES_SCATTER,
    K_PREP_GRAY, K_PREP_EQ_LUT, K_PREP_APPLY, K_FILL_MEAN, K_NORMAL_LINE, K_ROAD_HULL, K_HULL_MASKS,
    K_SGBM_PREFILTER, K_SGBM_HSUM, K_SGBM_VSUM, K_SGBM_PATH_V, K_SGBM_PATH_D1, K_SGBM_PATH_D3, K_SGBM_PATH_H, K_SGBM_FINAL,
    K_SGBM_MEDIAN, K_SPECKLE, K_DISP_SCALE, K_RANSAC_REPLAY, K_COUNT_
};

void svb_prof_begin(int id, cudaStream_t st);
void svb_prof_end(int id, cudaStream_t st);
const char *svb_prof_name(int id);
extern bool g_svb_prof_on;

struct ProfScope {
    int id; cudaStream_t st; bool on;
    ProfScope(int id_, cudaStream_t st_) : id(id_), st(st_), on(g_svb_prof_on) {
        nvtxRangePushA(svb_prof_name(id));
        if (on) svb_prof_begin(id, st);
    }
    ~ProfScope() {
        if (on) svb_prof_end(id, st);
        nvtxRangePop();
    }
};
