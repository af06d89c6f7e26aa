/* rt_oracle.h -- TEST INFRASTRUCTURE ONLY.  CPU restatement ("port") of the reference's
 * per-pixel render path, in plain C, on the packed scene of include/rtb200.h.
 * Parity status: PINNED -- tests/test_oracle.py checks this file bit-for-bit against the
 * reference's own compiled C++ renderer (oracle/_ref, built by oracle/Makefile from
 * /root/reference) and against the golden vectors under tests/golden/ that were generated
 * from it (tests/golden/make_golden.py).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may link or call this.  The product (librtb200.so) never does. */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stdint.h>
#include "../include/rtb200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Full frame (all pixels, spp grid honoured), float RGB, row 0 = bottom.  threads<=1: serial. */
double rto_render(const rtb200_scene* s, const rtb200_frame* f, float* out_rgb, int threads);
/* Sub-sample lattice region [x0,x0+w) x [y0,y0+h) of the (k*W) x (k*H) lattice, no box filter. */
double rto_render_region(const rtb200_scene* s, const rtb200_frame* f, int x0, int y0, int w, int h,
                         float* out_rgb, int threads);
/* Per-pixel 64-bit signature of the whole ray tree (hit surface ids, which lights were
 * tested, quantised shadow-filter values): a change between 8-neighbours marks a silhouette or
 * shadow-terminator pixel, primary or seen through reflections/refractions.  spp grid ignored
 * (k = 1 lattice).  Also returns the primary hit order (-1 = background). */
void   rto_signatures(const rtb200_scene* s, const rtb200_frame* f, uint64_t* sig, int32_t* primary_hit, int threads);
/* Ray-level queries (the loops of Scene.cpp:1710-1715 and 1842-1861). */
void   rto_nearest(const rtb200_scene* s, int n, const float* org3, const float* dir3,
                   int32_t* out_order, float* out_t, float* out_normal3, int threads);
void   rto_shadow(const rtb200_scene* s, int n, const float* point3, const float* light3,
                  float* out_filter3, int threads);
/* Full colour of arbitrary rays treated as primaries (energy 1, bounces 0, ior 1). */
void   rto_trace(const rtb200_scene* s, const rtb200_frame* f, int n, const float* org3, const float* dir3,
                 float* out_rgb3, int threads);
/* floor(clamp(c,0,1)*255+0.5), A=255, NaN->0; returns number of NaN channels. */
uint64_t rto_quantise(const float* rgb, size_t n_pixels, uint8_t* rgba8);
/* Ray counters of the last rto_render / rto_render_region call (summed over threads):
 * primary, shadow, reflected, refracted. */
void   rto_last_ray_counts(uint64_t out[4]);
int    rto_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
