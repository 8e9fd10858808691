/*
 * svb200.h -- C ABI of libsvb200.so: the B200 (sm_100a) implementation of the per-frame hot path of
 * thien/stereo.vision (disparity -> 3-D cloud -> RANSAC plane -> road points -> road mask).
 *
 * The reference has no native layer: its "operator interface" for this path is the set of Python
 * helpers in functions.py that stereovision.py:58-140 calls.  Each entry point below states the
 * helper(s) it replaces (file:line in the reference).  The Python drop-in module
 * (stereo.vision_b200/functions.py) binds these with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - every data pointer is a CALLER-OWNED DEVICE pointer (e.g. torch.Tensor.data_ptr()) unless the
 *     parameter name starts with h_ (host memory, read during the call only);
 *   - images are C-contiguous uint8, [F,H,W] or [F,H,W,3] (BGR); F = frames in the batch;
 *   - `stream` is a cudaStream_t passed as void* (torch.cuda.current_stream().cuda_stream);
 *     all work is enqueued on it and the call returns without synchronising;
 *   - return value: 0 = ok, < 0 = argument error (SVB_E_*), > 0 = a cudaError_t from the launch;
 *   - nothing is allocated or retained: scratch comes from the caller, sizes from svb_*_bytes().
 *
 * Records
 *   cloud  : uint64 per point,  bits  0-10 x | 11-21 y | 22-29 d | 30-31 zero    (low word)
 *                               bits 32-39 R | 40-47 G | 48-55 B | 56-63 zero    (high word)
 *   mcloud : uint32 per point,  bits  0-10 x | 11-21 y | 22-29 d                 (masked cloud)
 *   (x,y) = pixel site, d = uint8 disparity.  The hue bin rint(1000*colorsys_hue) in [0,1000] is a
 *   function of (R,G,B) and is computed where it is consumed (only near-plane points need it).
 *   X,Y,Z are never stored: they are exact float64 functions of (x,y,d) (functions.py:191-193)
 *   and are recomputed where needed, which is what makes the reprojection bit-exact.
 *   Point order inside a frame is the reference's append order (row-major over the lattice).
 *   Frame f of a batch owns the slab [f*cap, f*cap + n_f) of the record arrays, cap = lattice size.
 */
#ifndef SVB200_H
#define SVB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVB_OK            0
#define SVB_E_NULL       -1   /* required pointer is NULL */
#define SVB_E_SHAPE      -2   /* H,W outside [2,2048], F < 1, step < 1, ... */
#define SVB_E_ARG        -3   /* other invalid argument */
#define SVB_E_SCRATCH    -4   /* scratch buffer too small */
#define SVB_E_LIMIT      -5   /* size beyond the fixed maximum of an entry point that allocates nothing */

#define SVB_HUE_BINS     1001 /* keys '0.0' .. '1.0' of functions.py:76-78 */
#define SVB_HIST_STRIDE  1024 /* uint32 per frame in histogram arrays */
#define SVB_HIST_SPANS   1023 /* slot of a histogram row where svb_road_points_sites leaves the number of 16-pixel colour spans
                                (48 bytes each) it read for the frame -- what crosses the link when the image is read in place */
#define SVB_PLANE_STRIDE 8    /* doubles per frame in plane arrays, see svb_ransac_finalize */

/* camera constants, functions.py:15-22 */
typedef struct svb_camera {
    double focal_px;     /* camera_focal_length_px  */
    double baseline_m;   /* stereo_camera_baseline_m */
    double centre_h;     /* image_centre_h */
    double centre_w;     /* image_centre_w */
} svb_camera;

/* ABI version (bumped on any signature change) and a static description of an error code. */
int         svb_version(void);
const char *svb_error_string(int code);

/* lattice visited by range(0,H-1,step) x range(0,W-1,step), functions.py:185-186 */
int svb_lattice_rows(int H, int step);
int svb_lattice_cols(int W, int step);

/* ---- a1  fillDisparity / disparityCleaning('previous')            functions.py:130-147
 * filled[t] = sat_add_u8(raw[t], (raw[t] <= 2) ? filled[t-1] : 0), filled[-1] = prev0.
 * prev0 == NULL: frame 0 is copied unchanged (prev_disp is None, functions.py:134).
 * The recurrence over t runs inside one launch (one thread owns 16 pixels for all F frames and keeps
 * the loads of the next 12 frames in flight). */
int svb_fill_sequence(const uint8_t *raw, const uint8_t *prev0, uint8_t *filled,
                      int F, int H, int W, void *stream);

/* a1 fused with the counting pass of a4: also leaves in `scratch` the per-row-group site counts of the
 * FILLED frames (the bytes are in registers anyway), so that svb_project_compact_counted() can skip its
 * own counting pass.  carmask may be NULL.  scratch: svb_project_scratch_bytes(F,H,W,step) bytes; it must
 * reach svb_project_compact_counted() untouched, with the same F,H,W,step and carmask. */
int svb_fill_sequence_counted(const uint8_t *raw, const uint8_t *prev0, uint8_t *filled,
                              const uint8_t *carmask, int F, int H, int W, int step,
                              void *scratch, int64_t scratch_bytes, void *stream);

/* ---- a2  maskDisparity                                             functions.py:169-172
 * out = d where carmask != 0 else 0.  (a3 capDisparity is the identity, functions.py:164-167.) */
int svb_mask_disparity(const uint8_t *d, const uint8_t *carmask, uint8_t *out,
                       int F, int H, int W, void *stream);

/* ---- a4  projectDisparityTo3d (both calls of stereovision.py:84-85)  functions.py:178-198
 * For each frame: cloud  = lattice sites with d > 0, with RGB = bgr[y,x,2..0];
 *                 mcloud = lattice sites with d > 0 and carmask != 0 (no colour).
 * Any of cloud / mcloud may be NULL (that output is skipped); bgr NULL -> colour bits are 0;
 * carmask NULL -> every site counts as inside the mask.
 * counts[f*4+0] = N_f (cloud), counts[f*4+1] = M_f (mcloud); entries 2,3 are left untouched.
 * scratch: svb_project_scratch_bytes(F,H,W,step) bytes. */
int64_t svb_project_scratch_bytes(int F, int H, int W, int step);
int svb_project_compact(const uint8_t *disp, const uint8_t *carmask, const uint8_t *bgr,
                        int F, int H, int W, int step, const svb_camera *cam,
                        uint64_t *cloud, uint32_t *mcloud, int32_t *counts,
                        void *scratch, int64_t scratch_bytes, void *stream);

/* Same, for frames that came out of svb_fill_sequence_counted() (scratch already holds the counts). */
int svb_project_compact_counted(const uint8_t *filled, const uint8_t *carmask, const uint8_t *bgr,
                                int F, int H, int W, int step, const svb_camera *cam,
                                uint64_t *cloud, uint32_t *mcloud, int32_t *counts,
                                void *scratch, int64_t scratch_bytes, void *stream);

/* Materialise what the reference's list holds: xyz float64 [n,3] (functions.py:191-193, bit-exact),
 * rgb uint8 [n,3] (functions.py:195), sites int32 [n,2] as (y,x), disparity uint8 [n].
 * Any output may be NULL.  `rec_stride_words` = 2 for a cloud, 1 for an mcloud (32-bit words). */
int svb_cloud_export(const uint32_t *records, int rec_stride_words, int n, const svb_camera *cam,
                     double *xyz, float *xyz_f32, uint8_t *rgb, int32_t *sites, uint8_t *d,
                     uint16_t *hue, void *stream);

/* ---- a5-a7  RANSAC / planarFitting / randomNonCollinearPoints       functions.py:240-298
 * Scores T hypotheses per frame: plane through the triple (abc = inv([P1;P2;P3]).1), error = mean |P.abc - 1| / |abc|
 * over the trial's S sample points (functions.py:275,289), evaluated in exact-integer form.
 * errors[f*T + t] = +inf for a singular triple or a frame with M_f < S (functions.py:286,294).
 * Records are x | y << 11 | d << 22 with bits 30-31 ZERO (what svb_project_compact writes).
 *
 * Index source
 *   sample_idx != NULL : explicit draws, sample_idx int32 [F,T,S], triple_idx int32 [F,T,3]
 *                        (the replay of Python's `random` stream) -- parity mode; one warp per hypothesis;
 *   sample_idx == NULL : counter-based on-device draws keyed by (rng_seed, frame0+f, t): S distinct indices (the first
 *                        S values below M_f of a scrambled full-period congruential sequence) and triples redrawn while
 *                        exactly collinear (functions.py:251-259) -- throughput mode; one thread per hypothesis.
 *                        svb_ransac_dump_indices writes the same draws out for the oracle.
 * max_m: an upper bound of every frame's cloud size known to the host (e.g. the number of lattice sites
 *        inside carmask), or 0.  When 4*(max_m+1) bytes fit in shared memory each block stages its frame's
 *        cloud there once and the S gathers of a hypothesis become shared-memory reads; frames that
 *        turn out larger than max_m, and max_m = 0, gather from global memory (L2).  The errors do not depend on it,
 *        nor on F or the flags: a hypothesis' terms are summed in one fixed order (tests/test_gpu_configs.py).
 * flags: SVB_RANSAC_SHARE_SM = blocks of half the size (half of an SM's registers), so that blocks of a kernel launched
 *        on another stream find room on the same SM.  Same results.  pipeline.RoadPath no longer sets it: with a thread
 *        per hypothesis the full-size block leaves that room already. */
#define SVB_RANSAC_SHARE_SM 1
int svb_ransac_score(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                     const int32_t *counts, int counts_stride, int counts_col,
                     int F, int T, int S, int max_m, const svb_camera *cam,
                     const int32_t *sample_idx, const int32_t *triple_idx,
                     uint64_t rng_seed, int frame0, int flags,
                     double *errors, void *stream);
int svb_ransac_dump_indices(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                            const int32_t *counts, int counts_stride, int counts_col,
                            int F, int T, int S, const svb_camera *cam,
                            uint64_t rng_seed, int frame0,
                            int32_t *sample_idx, int32_t *triple_idx, void *stream);
/* ---- a5 / a7, the draws: CPython's `random` stream replayed on the device      functions.py:286, 252-254
 * The reference draws, per trial, one random.sample(points, S) and then rounds of three random.sample(points, 1)
 * until np.cross(P1-P2, P2-P3) is not exactly zero, all from Python's global Mersenne Twister.  This entry point
 * replays that stream (CPython 3.12: MT19937, getrandbits/_randbelow rejection, random.sample's pool and set paths
 * with setsize = 21 + 4**ceil(log(3k,4))) on the device, one warp per generator, and writes the indices
 * svb_ransac_score / svb_ransac_finalize take as explicit draws: sample_idx int32 [F,T,S], triple_idx int32 [F,T,3]
 * (-1 everywhere for a frame with M_f < S: random.sample raises before it draws, functions.py:286,294).
 *   seeds != NULL : device uint64 [F] (chain: [1]); generator f starts as after random.seed(seeds[f]);
 *   seeds == NULL : mt_state (device uint32 [F][625], chain: [1][625]) holds random.getstate()[1] -- 624 words + index;
 *   mt_state != NULL: on return it holds the generator where the reference's RANSAC call would leave it
 *                   (random.setstate-able), whichever way it was started;
 *   chain = 0     : frame f uses generator f (the parity protocol of SURVEY 8c: random.seed(s_f) before each frame);
 *   chain = 1     : ONE generator serves frames 0..F-1 in order (random.seed(s) once before loop.py's frame loop;
 *                   sequential by construction: one warp).
 * max_m: upper bound of every frame's cloud size (bitmap / pool sizing); scratch: svb_ransac_replay_scratch_bytes. */
int64_t svb_pyrandom_setsize(int k);
int64_t svb_ransac_replay_scratch_bytes(int F, int S, int max_m);
int svb_ransac_replay_draws(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                            const int32_t *counts, int counts_stride, int counts_col,
                            int F, int T, int S, int max_m, const svb_camera *cam,
                            const uint64_t *seeds, uint32_t *mt_state, int chain,
                            int32_t *sample_idx, int32_t *triple_idx,
                            void *scratch, int64_t scratch_bytes, void *stream);

/* First strict minimum over the T errors of each frame (functions.py:291) and the winning plane.
 * planes[f*8 + ..] = { a, b, c, |abc|, cutoff, best_error, best_trial, valid }
 *   cutoff = the smallest double v with fl(v/|abc|) >= point_threshold, so that
 *            "|P.abc-1| < cutoff"  <=>  "fl(|P.abc-1| / |abc|) < point_threshold" exactly;
 *   valid  = 1.0 when a plane was found, else 0.0 (RANSAC returned (None, None)), best_trial = -1.
 * triple_idx / rng_seed as in svb_ransac_score. */
int svb_ransac_finalize(const uint32_t *mcloud, int rec_stride_words, int64_t slab_stride_records,
                        const int32_t *counts, int counts_stride, int counts_col,
                        int F, int T, int S, const svb_camera *cam,
                        const int32_t *triple_idx, uint64_t rng_seed, int frame0,
                        const double *errors, double point_threshold,
                        double *planes, void *stream);
/* planes row from host coefficients (drop-in helpers that are handed `abc` by the caller). */
int svb_plane_from_abc(const double *h_abc, double point_threshold, double *h_plane_row);

/* ---- a8  calculatePointErrors                                        functions.py:300-312
 * dist[i] = | (X*a + Y*b + Z*c - 1) / |abc| |, float64, for one cloud of n records. */
int svb_point_errors(const uint32_t *records, int rec_stride_words, int n, const svb_camera *cam,
                     const double *h_abc, double *dist, void *stream);

/* ---- a9 / a12  computePlanarThreshold, filterPointsByHistogram       functions.py:314-323, 228-230
 * Stable compaction of one cloud: keep record i when
 *   mode 0: dist[i] < threshold                       (dist float64 [n])
 *   mode 1: hist[hue(record i)] > (uint32) threshold   (hist uint32 [>=1001])
 * out_count[0] = kept.  scratch: svb_compact_scratch_bytes(n). */
int64_t svb_compact_scratch_bytes(int n);
int svb_compact_cloud(const uint64_t *cloud, int n, int mode, const double *dist,
                      const uint32_t *hist, double threshold,
                      uint64_t *out_cloud, int32_t *out_count,
                      void *scratch, int64_t scratch_bytes, void *stream);

/* ---- svb_compact_cloud without caller scratch.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same arguments, outputs and results as svb_compact_cloud, minus the scratch:
 * the block counts live in a file-scope device array of the library (8 KB, held by every GPU that loads the library
 * until exit).  Limit: n <= SVB_COMPACT_STATIC_MAX_N records, else SVB_E_LIMIT and nothing is launched.  The call
 * performs no cudaMalloc / cudaFree and calls no allocating routine; the first call on a device looks the array's
 * address up (cudaGetSymbolAddress) -- make that call once outside a capture.  Every call on a device shares the
 * array: calls that overlap on the GPU (two streams, two threads) race; serialise them.  Nothing in it has to start
 * at zero (each call stores every entry it later reads). */
#define SVB_COMPACT_STATIC_MAX_N 4194304
int svb_compact_cloud_static(const uint64_t *cloud, int n, int mode, const double *dist,
                             const uint32_t *hist, double threshold,
                             uint64_t *out_cloud, int32_t *out_count, void *stream);

/* ---- a10-a11  BGRtoHSVHue + calculateColourHistogram                 functions.py:73-78, 215-226
 * hist[bin] += 1 for every record (bin = rint(1000*hue), functions.py:76).  hist is zeroed first. */
int svb_hue_histogram(const uint64_t *cloud, int n, uint32_t *hist, void *stream);
/* hue bin of raw colours (exhaustive known-answer test of a10): bins[i] from rgb[i*3..]. */
int svb_hue_bins(const uint8_t *rgb, int64_t n, uint16_t *bins, void *stream);

/* ---- a13  project3DPointsTo2DImagePoints (+ the caller's int32 cast)  functions.py:201-209
 * xy_f64 [n,2] = (((X*f)/Z)+cw, ((Y*f)/Z)+ch); xy_i32 [n,2] = trunc toward zero
 * (stereovision.py:112).  Either output may be NULL. */
int svb_reproject(const uint32_t *records, int rec_stride_words, int n, const svb_camera *cam,
                  double *xy_f64, int32_t *xy_i32, void *stream);
/* reprojection of every (coordinate, d) pair: out[c*256 + d] for c in [0,n_coord), d in [1,255]
 * with axis 0 = x (centre_w), 1 = y (centre_h) -- the exhaustive known-answer table of SURVEY 7.1-1. */
int svb_reproject_table(int axis, int n_coord, const svb_camera *cam, int32_t *out, void *stream);

/* Displacement bits of the reprojection: x' = x - bit_x(x,d), y' = y - bit_y(y,d) (the float64
 * expression of functions.py:206-207 lands on the site itself or one pixel before it, depending only on
 * the coordinate and d).  bits: uint32 [svb_reproject_bits_words(H,W)], word [c*8 + (d>>5)] bit (d&31),
 * x rows [0,W) then y rows [W,W+H).  *ok (device int32) = 1 if every pair moves by 0 or -1, else 0 and
 * the table must not be used.  Passing the table to svb_road_points replaces four float64 divisions
 * per kept point by two L1-resident bit lookups; NULL keeps the per-point arithmetic. */
int64_t svb_reproject_bits_words(int H, int W);
int svb_reproject_bits(int H, int W, const svb_camera *cam, uint32_t *bits, int32_t *ok, void *stream);

/* ---- a14  generatePointsAsImage / the overlay loop      functions.py:339-344, stereovision.py:131-133
 * img[y,x,:] = value[:] for each (x,y) in pts int32 [n,2]; channels = 1 (mask) or 3 (overlay).
 * The image is NOT cleared (the caller copies blackImg / imgL first, as the reference does).
 * Points outside the image are reported through *oob (device int32, incremented) and skipped
 * -- the reference would raise IndexError (or wrap for small negatives). */
int svb_scatter_points(const int32_t *pts, int n, uint8_t *img, int H, int W, int channels,
                       const uint8_t *h_value, int32_t *oob, void *stream);

/* ---- f1  sanitiseRoadImage: road-mask cleaning + point list            functions.py:346-366
 * Bit rows: an H x W mask as uint32 [H, ceil(W/32)], bit (x & 31) of word (x >> 5) = pixel x, unused
 * bits of a row's last word zero.  svb_bits_words(H,W) = words of one image. */
int64_t svb_bits_words(int H, int W);
/* bits[f] = (img[f] != 0); rowcnt (int32 [F*H], may be NULL) = set pixels per row.  Used once per
 * run on road_threshold_mask (functions.py:33, only its non-zero-ness is read at :355), and to list
 * the points of any byte mask with svb_mask_points. */
int svb_pack_mask(const uint8_t *img, int F, int H, int W, uint32_t *bits, int32_t *rowcnt, void *stream);
/* cleaned[f] = closing 9x9 -> erode 9x9 twice -> keep where threshold != 0 -> closing 9x9 of road[f]
 * (functions.py:349-356; OpenCV default border: outside pixels never win).  road [F,H,W] holds 0 / 255
 * as generatePointsAsImage writes it; a pixel is read by its top bit, and *status (device int32, may
 * be NULL) gets bit 0 when a byte other than 0 / 255 was seen -- the reference would then run grey
 * morphology and its result is NOT reproduced.  ParticleCleansing (:358) provably leaves a closed image
 * unchanged and is not run.  Outputs, each optional but not both image forms NULL:
 * cleaned uint8 [F,H,W] (0 / 255, must not alias road), cleaned_bits (bit rows), rowcnt int32 [F,H]. */
int svb_clean_road(const uint8_t *road, const uint32_t *threshold_bits, int F, int H, int W, uint8_t *cleaned,
                   uint32_t *cleaned_bits, int32_t *rowcnt, int32_t *status, void *stream);
/* the same from bit rows (a producer that already holds the road image packed) */
int svb_clean_road_bits(const uint32_t *road_bits, const uint32_t *threshold_bits, int F, int H, int W,
                        uint8_t *cleaned, uint32_t *cleaned_bits, int32_t *rowcnt, void *stream);
/* pts[f] int32 [cap,2] = (x,y) of the set pixels of bits[f] in row-major order (the double loop of
 * functions.py:359-365); counts[f] = their number, also when it exceeds cap (then only the first cap
 * points are written).  rowcnt as left by svb_clean_road / svb_pack_mask. */
int svb_mask_points(const uint32_t *bits, const int32_t *rowcnt, int F, int H, int W, int32_t cap, int32_t *pts,
                    int32_t *counts, void *stream);

/* ---- f2  gammaChange / greyscale: the front end                         functions.py:61-67, 80-96
 * For F images bgr uint8 [F,H,W,3]: every byte through table (HOST pointer to 256 bytes, cv2.LUT with
 * the table of functions.py:62-65; NULL = identity) -> bgr_out; gray uint8 [F,H,W] of the corrected
 * colours (COLOR_BGR2GRAY: (3735 B + 19235 G + 9798 R + 2^14) >> 15); hist uint32 [F,256] of gray,
 * zeroed here.  Each output may be NULL, not all three. */
int svb_gamma_gray_hist(const uint8_t *bgr, int F, int H, int W, const uint8_t *h_table, uint8_t *bgr_out, uint8_t *gray,
                        uint32_t *hist, void *stream);
/* cv2.equalizeHist's table per image from its histogram: lut uint8 [F,256] (functions.py:95) */
int svb_equalize_lut(const uint32_t *hist, int F, uint8_t *lut, void *stream);
/* dst[f][i] = lut[f][src[f][i]], i < n_per_frame (lut: device, uint8 [F,256]) */
int svb_apply_lut_frames(const uint8_t *src, int F, int64_t n_per_frame, const uint8_t *lut, uint8_t *dst, void *stream);
/* dst[i] = lut[src[i]], i < n (lut: device, 256 bytes): cv2.LUT on any byte array */
int svb_lut_bytes(const uint8_t *src, int64_t n, const uint8_t *lut, uint8_t *dst, void *stream);

/* ---- f3  fillAltDisparity ('mean' option)                              functions.py:149-162
 * filled[f][y][x] = d where d >= 2, else trunc(mean of the row's non-zero entries) (0 for an empty row).
 * filled may alias disparity (a row is summed completely before the warp that owns it rewrites it). */
int svb_fill_row_mean(const uint8_t *disparity, int F, int H, int W, uint8_t *filled, void *stream);
/* ---- f3  getNormalVectorLine                                            functions.py:396-416
 * out_xy[f] = int(x'), int(y') of the reference's float64 sequence for base point base_xy[f] = (x, y), plane
 * abc[f*abc_stride .. +2] and the disparity under the base point: disparity uint8 [F,H,W], or NULL and
 * d_values uint8 [F].  status[f]: 0 ok, 1 a conversion would raise ValueError (nan), 2 OverflowError (inf),
 * 4 the base point is outside the image (IndexError).  All device pointers. */
int svb_normal_line(const int32_t *base_xy, const double *abc, int abc_stride, const uint8_t *disparity,
                    const uint8_t *d_values, int F, int H, int W, const svb_camera *cam, int32_t *out_xy,
                    int32_t *status, void *stream);

/* ---- f3  the obstacle stage                                              stereovision.py:170-200
 * From the cleaned road image as bit rows (as svb_clean_road / svb_pack_mask leave them), per frame:
 *   hull_xy [F,cap,2] / hull_n [F]  cv2.convexHull of the road points: strict hull in OpenCV's orientation, starting
 *                                   at the lexicographically largest (x, y) (OpenCV picks its own start vertex);
 *                                   hull_n is the true size also when it exceeds cap;
 *   hull_mask uint8 [F,H,W]         drawContours(cleaned.copy(), [hull], 0, 255, -100)      (:174-176)
 *   obstacles uint8 [F,H,W]         NOT cleaned, kept inside hull_mask                        (:178-180)
 *   centre_xy int32 [F,2]           int() of minEnclosingCircle(hull)'s centre, functions.py:418-421; (-1,-1) when
 *                                   the image is empty (the reference's convexHull raises and the stage is skipped).
 * Every output is optional.  Two kernels: the hull of each frame (one CTA per frame), then the masks from the rows'
 * spans, which travel through `scratch` (svb_road_hull_scratch_bytes; needed only when a mask is asked for). */
int64_t svb_road_hull_scratch_bytes(int F, int H);
int svb_road_hull(const uint32_t *bits, int F, int H, int W, int cap, int32_t *hull_xy, int32_t *hull_n,
                  int32_t *centre_xy, uint8_t *hull_mask, uint8_t *obstacles, void *scratch, int64_t scratch_bytes,
                  void *stream);

/* ---- f4  the matcher stage                                               functions.py:26, 104-128
 * svb_sgbm_compute    = stereoProcessor.compute(grayL, grayR) with the module's matcher
 *                       cv2.StereoSGBM_create(0, 128, 21) (functions.py:26,108): OpenCV's MODE_SGBM semi-global block
 *                       matcher (P1 = 2, P2 = 5, preFilterCap 15, uniquenessRatio 0, disp12MaxDiff 1 as OpenCV resolves
 *                       the zeros) followed by its 3x3 median.  grayL, grayR uint8 [F,H,W] -> disp16 int16 [F,H,W]
 *                       (disparity x 16, -16 = no match).  num_disparities must be 128 (svb_sgbm_supported),
 *                       block odd in [1,31], W - 128 > block / 2.  The cost volumes travel through `scratch`:
 *                       svb_sgbm_scratch_bytes(f, H, W, 128) is what f frames in flight need (250 MB per 1024x544
 *                       frame); any size from one frame up is accepted and the batch is processed in chunks.
 *                       *status (device int32, may be NULL; the caller zeroes it) gets bit 0 when a block cost exceeded
 *                       32767 - P2 (a mean pixel cost above 74 over the whole window): OpenCV's own int16 stores wrap
 *                       or saturate there depending on its SIMD path, so the output is then not specified.
 * svb_filter_speckles = cv2.filterSpeckles(disp16, new_val, max_size, max_diff) in place      (functions.py:111)
 * svb_disparity_scale = threshold-to-zero, (d / 16.) -> uint8, x (256. / max_disparity) -> uint8 (functions.py:117-127;
 *                       the crop of functions.py:121-123 is a view the host takes). */
int     svb_sgbm_supported(int H, int W, int num_disparities, int block);
int64_t svb_sgbm_scratch_bytes(int F, int H, int W, int num_disparities);
int     svb_sgbm_compute(const uint8_t *grayL, const uint8_t *grayR, int F, int H, int W, int num_disparities, int block,
                         int16_t *disp16, int32_t *status, void *scratch, int64_t scratch_bytes, void *stream);
int64_t svb_speckle_scratch_bytes(int F, int H, int W);
int     svb_filter_speckles(int16_t *img16, int F, int H, int W, int new_val, int max_size, int max_diff, void *scratch,
                            int64_t scratch_bytes, void *stream);
int     svb_disparity_scale(const int16_t *disp16, int64_t n, int max_disparity, uint8_t *out, void *stream);

/* ---- a8-a14 fused, batched: the road points of F frames               stereovision.py:97-140
 * reproj_bits: table of svb_reproject_bits or NULL.
 * pass 1: near = |P.abc-1| < cutoff ; per-frame hue histogram of the near points (shared memory,
 *         merged with global atomics); near bits kept;
 * pass 2: keep = near && hist[hue] > colour_threshold ; reproject (exact fp64) ; road_mask[y',x'] = 255.
 * counts[f*4+2] = K_f (near), counts[f*4+3] = R_f (kept).  hist uint32 [F,1024]; road_mask uint8
 * [F,H,W] is cleared here.  overlay (optional, uint8 [F,H,W,3], pre-filled by the caller with imgL)
 * receives BGR (0,255,0) at the same sites.  scratch: svb_road_scratch_bytes(F,H,W,step). */
int64_t svb_road_scratch_bytes(int F, int H, int W, int step);
int svb_road_points(const uint64_t *cloud, const int32_t *counts, const double *planes,
                    int F, int H, int W, int step, const svb_camera *cam,
                    uint32_t colour_threshold, const uint32_t *reproj_bits,
                    uint32_t *hist, uint8_t *road_mask, uint8_t *overlay, int32_t *counts_out,
                    void *scratch, int64_t scratch_bytes, void *stream);

/* ---- a8-a14 fused, batched, straight from the path's inputs
 * The same outputs as svb_road_points, computed from the filled disparity (a1's output) and the colour
 * image instead of the cloud records: near-plane membership depends on (x, y, d) only, so the lattice
 * sites are read at 1 byte each and the colour only where a near point sits (functions.py:195 reads the
 * same pixel).  filled uint8 [F,H,W]; bgr uint8 [F,H,W,3] (lattice rows suffice); reproj_bits is required.
 * counts[f*4+2] = K_f, counts[f*4+3] = R_f; counts[f*4+0..1] are not touched.  Applies when
 * svb_road_sites_supported(H,W,step) = 1 (step 1 or 2); otherwise SVB_E_ARG and the caller uses
 * svb_road_points.  scratch: svb_road_sites_scratch_bytes(F,H,W,step), 4-byte aligned.
 * bgr is read through ordinary loads, 48 bytes (a lane's 16 pixels) at a time and only where a near point sits, so it may
 * just as well be the caller's PINNED HOST frames (mapped into the device's address space): then only those spans cross
 * the link and no copy of the colour image is made (pipeline.SequenceStreamer(upload="pull")).  hist[f][SVB_HIST_SPANS]
 * returns the number of spans read for frame f; hist[f][0 .. SVB_HUE_BINS) are the hue bins. */
int svb_road_sites_supported(int H, int W, int step);
int64_t svb_road_sites_scratch_bytes(int F, int H, int W, int step);
int svb_road_points_sites(const uint8_t *filled, const uint8_t *bgr, const double *planes, int F, int H, int W, int step,
                          const svb_camera *cam, uint32_t colour_threshold, const uint32_t *reproj_bits, uint32_t *hist,
                          uint8_t *road_mask, uint8_t *overlay, int32_t *counts, void *scratch, int64_t scratch_bytes,
                          void *stream);

/* ---- host -> device upload of the rows the path reads
 * The projection reads the colour image only on lattice rows (every step-th row), so an end-to-end caller
 * need not send the others: this copies `rows` rows of `width_bytes` bytes, `pitch_bytes` apart in BOTH the
 * (pinned) host source and the device destination (cudaMemcpy2DAsync), i.e. the rows land where a full-frame
 * copy would have put them and the rows in between are left untouched (never read). */
int svb_upload_rows(const void *h_src, void *d_dst, int64_t pitch_bytes, int64_t width_bytes, int64_t rows, void *stream);
/* A block of rows of every frame of a batch in one 3-D copy: both sides hold `frames` frames of `rows_per_frame` rows that
 * are `pitch_bytes` apart; rows [first, first + nrows) of every frame are copied, `width_bytes` of each.  With pitch =
 * step x the image row size this addresses lattice rows: the end-to-end streamer sends the colour image only on the lattice
 * rows that can hold a near-plane point of the chunk's planes (pipeline.SequenceStreamer(upload="needed")); nothing else
 * of the colour image is ever read (functions.py:195 is only evaluated for cloud points, stereovision.py:103 for near ones). */
int svb_upload_row_block(const void *h_src, void *d_dst, int64_t pitch_bytes, int64_t width_bytes, int64_t rows_per_frame,
                         int64_t first, int64_t nrows, int64_t frames, void *stream);

/* ---- tracing: per-kernel CUDA-event timing, off by default.
 * svb_profile_enable(1) clears the records and brackets every later kernel launch of this library
 * with two events on its stream; after the caller has synchronised, svb_profile_get(id, ..) returns
 * the kernel's name, the summed device time in ms and the number of launches, id in
 * [0, svb_profile_kernels()).  Process-global; not for concurrent use from several threads. */
int svb_profile_enable(int on);
int svb_profile_kernels(void);
int svb_profile_get(int id, char *name, int name_cap, double *ms_total, int *launches);

#ifdef __cplusplus
}
#endif
#endif /* SVB200_H */
