/*
 * svb200_debug.h -- test and tuning hooks of libsvb200.so.  NOT part of the drop-in boundary (include/svb200.h):
 * these switch between equivalent kernel variants so that the tests can hold them to the same outputs.  They are
 * process-global and not meant for concurrent use.
 */
#ifndef SVB200_DEBUG_H
#define SVB200_DEBUG_H
#ifdef __cplusplus
extern "C" {
#endif
/* test hook: force svb_road_points onto its two-kernel path even when the fused cluster kernel applies */
int svb_debug_force_two_pass(int on);
/* test hook: row bands per frame of the 1024-wide cleaning kernel (default 4); 0 = use the general kernel */
int svb_debug_clean_bands(int bands);
#ifdef __cplusplus
}
#endif
#endif /* SVB200_DEBUG_H */
