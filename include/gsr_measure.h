/* gsr_measure.h - measurement-only entry points of libgsr.so (bench.py).  NOT part of the boundary a maintainer
 * binds under GaussianRenderBase (that is include/gsr.h); they have no counterpart in the reference
 * (M-Gjerde/GaussianSplattingViewer has no timing or counter code: SURVEY.md section 6). */
#ifndef GSR_MEASURE_H
#define GSR_MEASURE_H

#include "gsr.h"

#ifdef __cplusplus
extern "C" {
#endif

/* micro-benchmarks for the blend kernel's roofline denominators ("peaks to be measured by a
 * micro-benchmark on the box"): FP32 FMA TFLOP/s, 1e12 ex2/s, and a device copy GB/s (read + write).  Synchronises. */
int gsr_measure_peaks(gsr_t* h, float* fp32_tflops, float* ex2_tops, float* copy_gbs);
/* re-blends the last frame with counters: (pixel, splat) pairs visited before the pixel terminated, and pairs that
 * contributed a fragment (the algorithmic work of SURVEY.md 8d).  Synchronises; the frame buffer passed to the last
 * draw is untouched. */
int gsr_count_pairs(gsr_t* h, unsigned long long* visited, unsigned long long* contributed);

#ifdef __cplusplus
}
#endif
#endif /* GSR_MEASURE_H */
