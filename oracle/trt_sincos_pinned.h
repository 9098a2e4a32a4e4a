/*
 * trt_sincos_pinned.h -- the pinned sin/cos of the oracle's arithmetic convention (TEST INFRASTRUCTURE).
 * WGSL leaves the accuracy of sin()/cos() to the implementation (shader.wgsl:319-320,359-360 call them);
 * the oracle fixes one implementation so that CPU and GPU agree bit for bit.  Shared by the hand-written
 * oracle (trt_oracle.c) and by the transliterated reference shader (ref_driver.c -> wgsl_rt.h), so that the
 * two differ ONLY in where the algorithm's text comes from.
 */
#ifndef TRT_SINCOS_PINNED_H
#define TRT_SINCOS_PINNED_H
#include <math.h>

/*
 * Pinned sin/cos (the arithmetic convention for WGSL sin()/cos(); see trt_oracle.h).
 *   q = rint(phi * 2/pi);  r = phi - q*pi/2 in three Cody-Waite steps (fused);
 *   sin(r) = r + r*s*(S1 + s*(S2 + s*S3)),  cos(r) = 1 + s*(C0 + s*(C1 + s*(C2 + s*C3))),  s = r*r,
 *   every multiply-add fused, then the quadrant swap.  Coefficients: own minimax fit on
 *   [-pi/4, pi/4] (oracle/fit_sincos.py).
 */
static inline void sincos_pinned(float phi, float *sn, float *cs) {
    float q = rintf(phi * 0.6366197466850281f);
    float r = fmaf(q, -1.5703125f, phi);
    r = fmaf(q, -0.0004837512969970703f, r);
    r = fmaf(q, -7.549790126404332e-08f, r);
    float s = r * r;
    float ps = fmaf(-0.00019517233886290342f, s, 0.008332177996635437f);
    ps = fmaf(ps, s, -0.16666655242443085f);
    float sr = fmaf(r * s, ps, r);
    float pc = fmaf(2.4438379114144482e-05f, s, -0.0013887366512790322f);
    pc = fmaf(pc, s, 0.04166664555668831f);
    pc = fmaf(pc, s, -0.5f);
    float cr = fmaf(pc, s, 1.0f);
    int k = (int)q & 3;
    float so = (k & 1) ? cr : sr;
    float co = (k & 1) ? sr : cr;
    if (k == 1 || k == 2) co = -co;
    if (k >= 2) so = -so;
    *sn = so;
    *cs = co;
}
#endif
