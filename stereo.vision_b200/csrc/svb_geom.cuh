// Point geometry shared by the RANSAC scorer and the draw replay: a record's float64 point
// (functions.py:191-193) and the two collinearity tests.
#pragma once
#include "svb_common.cuh"

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
__device__ __forceinline__ int img_det(uint32_t r1, uint32_t r2, uint32_t r3) {
    int x1 = rec_x(r1), y1 = rec_y(r1), x2 = rec_x(r2), y2 = rec_y(r2), x3 = rec_x(r3), y3 = rec_y(r3);
    return x1 * (y2 - y3) + x2 * (y3 - y1) + x3 * (y1 - y2);
}

