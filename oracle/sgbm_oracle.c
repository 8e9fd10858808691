/* TEST INFRASTRUCTURE -- not part of the product.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.
 *
 * CPU restatement of the disparity stage of thien/stereo.vision (functions.py:26 and :104-128):
 *
 *     stereoProcessor = cv2.StereoSGBM_create(0, 128, 21)           functions.py:26
 *     disparity = stereoProcessor.compute(grayL, grayR)              functions.py:108
 *     cv2.filterSpeckles(disparity, 0, 4000, 128 - 5)                functions.py:111
 *     threshold TOZERO, /16. -> uint8, * (256/128) -> uint8          functions.py:117-127
 *
 * The algorithm lives in a third-party dependency that is not under /root/reference: OpenCV
 * (the reference pins no version; this container has opencv-python-headless 4.13.0).  This file
 * restates OpenCV's published semi-global block matcher in mode MODE_SGBM (Hirschmueller 2008 with
 * Birchfield-Tomasi pixel costs on an x-Sobel prefiltered image plus the raw image, a block sum, five
 * path directions in one top-to-bottom sweep, winner-takes-all with quadratic sub-pixel refinement, a
 * left-right check, then a 3x3 median) and cv2.filterSpeckles.  It is written as whole-volume array
 * code, not with OpenCV's cyclic buffers.  PINNED: oracle/make_golden_sgbm.py runs cv2 (through the
 * unmodified functions.disparity of the reference) on the same inputs and tests/test_oracle_golden.py
 * holds this file to those fixtures bit for bit.
 *
 * Parameters as StereoSGBM_create(0, D, block) resolves them: P1 = 2, P2 = 5 (0 means "default", and
 * P2 = max(P2, P1 + 1)), preFilterCap 0 -> clip 15, uniquenessRatio 0 (test never fires),
 * disp12MaxDiff 0 -> see `lrdiff`, costs are int16 and every store wraps like a C cast.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef int16_t cost_t;
#define MAX_COST 32767

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline cost_t sat16(int v) { return (cost_t)(v > 32767 ? 32767 : (v < -32768 ? -32768 : v)); }

/* x-Sobel through the clip table and the raw row; columns 0 and W-1 of BOTH planes hold the table's
 * centre value (OpenCV fills the row buffers' ends with tab[0] for every channel). */
static void prefilter_row(const uint8_t* img, int H, int W, int y, int ftzero, uint8_t* grad, uint8_t* raw)
{
    const uint8_t* r = img + (size_t)y * W;
    const uint8_t* rn = y > 0 ? r - W : r;
    const uint8_t* rs = y < H - 1 ? r + W : r;
    for (int x = 0; x < W; x++) {
        if (x == 0 || x == W - 1) { grad[x] = (uint8_t)ftzero; raw[x] = (uint8_t)ftzero; continue; }
        int g = (r[x + 1] - r[x - 1]) * 2 + rn[x + 1] - rn[x - 1] + rs[x + 1] - rs[x - 1];
        grad[x] = (uint8_t)(imin(imax(g, -ftzero), ftzero) + ftzero);
        raw[x] = r[x];
    }
}

/* half-pixel interval of a row: lo/hi = min/max of (p, (p+left)/2, (p+right)/2) */
static void interval_row(const uint8_t* p, int W, uint8_t* lo, uint8_t* hi)
{
    for (int x = 0; x < W; x++) {
        int v = p[x];
        int vl = x > 0 ? (v + p[x - 1]) / 2 : v;
        int vr = x < W - 1 ? (v + p[x + 1]) / 2 : v;
        lo[x] = (uint8_t)imin(v, imin(vl, vr));
        hi[x] = (uint8_t)imax(v, imax(vl, vr));
    }
}

/* Birchfield-Tomasi cost of one row: pix[x1*D + d], x = x1 + D, matched against x - d */
static void pixel_cost_row(const uint8_t* L, const uint8_t* R, int H, int W, int y, int D, int ftzero,
                           cost_t* pix, uint8_t* tmp)
{
    uint8_t *gl = tmp, *rl = gl + W, *gr = rl + W, *rr = gr + W;
    uint8_t *lo1 = rr + W, *hi1 = lo1 + W, *lo2 = hi1 + W, *hi2 = lo2 + W;
    int w1 = W - D;
    prefilter_row(L, H, W, y, ftzero, gl, rl);
    prefilter_row(R, H, W, y, ftzero, gr, rr);
    memset(pix, 0, sizeof(cost_t) * (size_t)w1 * D);
    for (int c = 0; c < 2; c++) {
        const uint8_t* p1 = c ? rl : gl;
        const uint8_t* p2 = c ? rr : gr;
        int shift = c ? 2 : 0;
        interval_row(p1, W, lo1, hi1);
        interval_row(p2, W, lo2, hi2);
        for (int x1 = 0; x1 < w1; x1++) {
            int x = x1 + D, u = p1[x], u0 = lo1[x], u1 = hi1[x];
            for (int d = 0; d < D; d++) {
                int v = p2[x - d], v0 = lo2[x - d], v1 = hi2[x - d];
                int c0 = imax(0, imax(u - v1, v0 - u));
                int c1 = imax(0, imax(v - u1, u0 - v));
                pix[x1 * D + d] = (cost_t)(pix[x1 * D + d] + (imin(c0, c1) >> shift));
            }
        }
    }
}

/* Path cost: L_r(p,d) = C(p,d) + min(L_r(p-r,d), L_r(p-r,d-1)+P1, L_r(p-r,d+1)+P1, min_k L_r(p-r,k)+P2) - min_k L_r(p-r,k)
 * (the paper's formula 13).  Subtracting min + P2 instead shifts every L by P2 and leaves the winner and
 * the sub-pixel term unchanged until S saturates at 32767; cv2 4.13 on saturating inputs (independent
 * noise, block 21) matches the form above and no other shift (golden case "noise").
 * lrdiff: cv2 resolves disp12MaxDiff = 0 to 1 (checked against cv2 4.13; 0 differs on ~10 % of pixels).
 * Window rows below the image repeat the last row, and column x1 = 0 of C is updated like the others
 * (both checked the same way: the alternatives differ from cv2). */
int svo_sgbm(const uint8_t* L, const uint8_t* R, int H, int W, int D, int block, int P1, int P2,
             int lrdiff, int uniq, int ftzero, int median, int16_t* disp)
{
    const int w1 = W - D, SW2 = block / 2, SH2 = block / 2;
    const int INVALID = -16;
    if (w1 <= SW2 || D <= 0 || H <= 0) {
        for (size_t i = 0; i < (size_t)H * W; i++) disp[i] = INVALID;
        return w1 <= 0 ? 0 : -1;
    }
    const size_t row = (size_t)w1 * D;
    cost_t* hsum = (cost_t*)malloc(sizeof(cost_t) * row * H);
    cost_t* pix = (cost_t*)malloc(sizeof(cost_t) * row);
    cost_t* C = (cost_t*)calloc(row, sizeof(cost_t));
    cost_t* S = (cost_t*)malloc(sizeof(cost_t) * row);
    uint8_t* tmp = (uint8_t*)malloc((size_t)W * 8);
    /* Lr[buf][dir][x1 + 1][d], zero borders in x; minLr likewise */
    const size_t lsz = (size_t)4 * (w1 + 2) * D;
    cost_t* Lr[2] = { (cost_t*)calloc(lsz, sizeof(cost_t)), (cost_t*)calloc(lsz, sizeof(cost_t)) };
    cost_t* mL[2] = { (cost_t*)calloc((size_t)4 * (w1 + 2), sizeof(cost_t)), (cost_t*)calloc((size_t)4 * (w1 + 2), sizeof(cost_t)) };
    int16_t* d2 = (int16_t*)malloc(sizeof(int16_t) * W);
    cost_t* d2c = (cost_t*)malloc(sizeof(cost_t) * W);
    int16_t* out = median ? (int16_t*)malloc(sizeof(int16_t) * (size_t)H * W) : disp;
#define LR(b, k, xi) (Lr[b] + ((size_t)(k) * (w1 + 2) + (xi)) * D)
#define ML(b, k, xi) (mL[b][(size_t)(k) * (w1 + 2) + (xi)])

    /* horizontal block sums of every row, window clamped to [0, w1) in the x1 domain */
    for (int y = 0; y < H; y++) {
        cost_t* hs = hsum + row * y;
        pixel_cost_row(L, R, H, W, y, D, ftzero, pix, tmp);
        for (int d = 0; d < D; d++) {
            int s = pix[d] * (SW2 + 1);
            for (int j = 1; j <= SW2; j++) s += pix[j * D + d];
            hs[d] = (cost_t)s;
        }
        for (int x1 = 1; x1 < w1; x1++) {
            const cost_t* add = pix + (size_t)imin(x1 + SW2, w1 - 1) * D;
            const cost_t* sub = pix + (size_t)imax(x1 - SW2 - 1, 0) * D;
            for (int d = 0; d < D; d++)
                hs[x1 * D + d] = (cost_t)(hs[(x1 - 1) * D + d] + add[d] - sub[d]);
        }
    }

    for (int y = 0; y < H; y++) {
        /* vertical block sum, incremental */
        if (y == 0) {
            for (int k = 0; k <= SH2; k++) {
                const cost_t* hs = hsum + row * imin(k, H - 1);
                int scale = k == 0 ? SH2 + 1 : 1;
                for (size_t i = 0; i < row; i++) C[i] = (cost_t)(C[i] + hs[i] * scale);
            }
        } else {
            int k = y + SH2;
            const cost_t* ha = hsum + row * imin(k, H - 1);
            const cost_t* hb = hsum + row * imax(y - SH2 - 1, 0);
            for (size_t i = 0; i < row; i++) C[i] = (cost_t)(C[i] + ha[i] - hb[i]);
        }
        memset(S, 0, sizeof(cost_t) * row);
        int cur = y & 1, prv = cur ^ 1;
        /* left-to-right sweep: directions 0 (from x-1, this row), 1 (x-1, row above), 2 (x, above), 3 (x+1, above) */
        for (int x1 = 0; x1 < w1; x1++) {
            const cost_t* src[4] = { LR(cur, 0, x1), LR(prv, 1, x1), LR(prv, 2, x1 + 1), LR(prv, 3, x1 + 2) };
            int delta[4] = { ML(cur, 0, x1) + P2, ML(prv, 1, x1) + P2, ML(prv, 2, x1 + 1) + P2, ML(prv, 3, x1 + 2) + P2 };
            const cost_t* Cp = C + (size_t)x1 * D;
            cost_t* Sp = S + (size_t)x1 * D;
            for (int k = 0; k < 4; k++) {
                cost_t* dst = LR(cur, k, x1 + 1);
                int mn = MAX_COST;
                for (int d = 0; d < D; d++) {
                    int a = src[k][d];
                    int b = (d > 0 ? src[k][d - 1] : MAX_COST) + P1;
                    int c = (d < D - 1 ? src[k][d + 1] : MAX_COST) + P1;
                    int l = Cp[d] + imin(a, imin(b, imin(c, delta[k]))) - (delta[k] - P2);
                    dst[d] = (cost_t)l;
                    mn = imin(mn, l);
                }
                ML(cur, k, x1 + 1) = (cost_t)mn;
            }
            for (int d = 0; d < D; d++)
                Sp[d] = sat16(Sp[d] + LR(cur, 0, x1 + 1)[d] + LR(cur, 1, x1 + 1)[d] + LR(cur, 2, x1 + 1)[d] + LR(cur, 3, x1 + 1)[d]);
        }
        /* right-to-left sweep (fifth direction) fused with the winner search */
        int16_t* d1 = out + (size_t)y * W;
        for (int x = 0; x < W; x++) { d1[x] = INVALID; d2[x] = INVALID; d2c[x] = MAX_COST; }
        /* direction-0 slots are reused; the cell right of the last column is a zero border */
        memset(LR(cur, 0, w1 + 1), 0, sizeof(cost_t) * D);
        ML(cur, 0, w1 + 1) = 0;
        for (int x1 = w1 - 1; x1 >= 0; x1--) {
            const cost_t* s0 = LR(cur, 0, x1 + 2);
            cost_t* dst = LR(cur, 0, x1 + 1);
            int delta0 = ML(cur, 0, x1 + 2) + P2;
            const cost_t* Cp = C + (size_t)x1 * D;
            cost_t* Sp = S + (size_t)x1 * D;
            int mn = MAX_COST, minS = MAX_COST, best = -1;
            for (int d = 0; d < D; d++) {
                int a = s0[d];
                int b = (d > 0 ? s0[d - 1] : MAX_COST) + P1;
                int c = (d < D - 1 ? s0[d + 1] : MAX_COST) + P1;
                int l = Cp[d] + imin(a, imin(b, imin(c, delta0))) - (delta0 - P2);
                dst[d] = (cost_t)l;
                mn = imin(mn, l);
                int sv = Sp[d] = sat16(Sp[d] + l);
                if (sv < minS) { minS = sv; best = d; }
            }
            ML(cur, 0, x1 + 1) = (cost_t)mn;
            int d;
            for (d = 0; d < D; d++)
                if (Sp[d] * (100 - uniq) < minS * 100 && abs(best - d) > 1) break;
            if (d < D) continue;
            d = best;
            if (d < 0) continue;            /* every S is 32767: no winner (OpenCV would index -1) */
            int x2 = x1 + D - d;
            if (d2c[x2] > minS) { d2c[x2] = (cost_t)minS; d2[x2] = (int16_t)d; }
            if (0 < d && d < D - 1) {
                int den = imax(Sp[d - 1] + Sp[d + 1] - 2 * Sp[d], 1);
                d = d * 16 + ((Sp[d - 1] - Sp[d + 1]) * 16 + den) / (den * 2);
            } else d *= 16;
            d1[x1 + D] = (int16_t)d;
        }
        for (int x = D; x < W; x++) {
            int v = d1[x];
            if (v == INVALID) continue;
            int da = v >> 4, db = (v + 15) >> 4;
            int xa = x - da, xb = x - db;
            if (0 <= xa && xa < W && d2[xa] >= 0 && abs(d2[xa] - da) > lrdiff &&
                0 <= xb && xb < W && d2[xb] >= 0 && abs(d2[xb] - db) > lrdiff)
                d1[x] = INVALID;
        }
        /* the zero border left of column 0 must stay zero for the next rows: slot (0, x=0) is never written */
    }
    if (median) {
        for (int y = 0; y < H; y++)
            for (int x = 0; x < W; x++) {
                int16_t v[9];
                int n = 0;
                for (int dy = -1; dy <= 1; dy++)
                    for (int dx = -1; dx <= 1; dx++)
                        v[n++] = out[(size_t)imin(imax(y + dy, 0), H - 1) * W + imin(imax(x + dx, 0), W - 1)];
                for (int i = 1; i < 9; i++) { int16_t t = v[i]; int j = i; while (j > 0 && v[j - 1] > t) { v[j] = v[j - 1]; j--; } v[j] = t; }
                disp[(size_t)y * W + x] = v[4];
            }
        free(out);
    }
    free(hsum); free(pix); free(C); free(S); free(tmp); free(Lr[0]); free(Lr[1]); free(mL[0]); free(mL[1]); free(d2); free(d2c);
    return 0;
#undef LR
#undef ML
}

/* cv2.filterSpeckles(img, newVal, maxSpeckleSize, maxDiff) on int16: 4-connected components of the
 * pixels != newVal whose neighbours differ by at most maxDiff; components of at most maxSpeckleSize
 * pixels become newVal. */
int svo_filter_speckles(int16_t* img, int H, int W, int newVal, int maxSize, int maxDiff)
{
    size_t n = (size_t)H * W;
    int32_t* label = (int32_t*)calloc(n, sizeof(int32_t));
    int32_t* stack = (int32_t*)malloc(sizeof(int32_t) * n);
    int32_t* members = (int32_t*)malloc(sizeof(int32_t) * n);
    int cur = 0;
    for (size_t s = 0; s < n; s++) {
        if (img[s] == newVal || label[s]) continue;
        int top = 0, cnt = 0;
        cur++;
        label[s] = cur; stack[top++] = (int32_t)s;
        while (top) {
            int p = stack[--top];
            members[cnt++] = p;
            int y = p / W, x = p % W, v = img[p];
            int nb[4] = { x + 1 < W ? p + 1 : -1, x > 0 ? p - 1 : -1, y + 1 < H ? p + W : -1, y > 0 ? p - W : -1 };
            for (int k = 0; k < 4; k++) {
                int q = nb[k];
                if (q < 0 || label[q] || img[q] == newVal || abs(v - img[q]) > maxDiff) continue;
                label[q] = cur; stack[top++] = q;
            }
        }
        if (cnt <= maxSize) label[s] = -cur;       /* mark through the seed */
        /* values are only rewritten after all components are known (they never feed another flood) */
        if (cnt <= maxSize) for (int i = 0; i < cnt; i++) label[members[i]] = -cur;
    }
    for (size_t s = 0; s < n; s++) if (label[s] < 0) img[s] = (int16_t)newVal;
    free(label); free(stack); free(members);
    return 0;
}
