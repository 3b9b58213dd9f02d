/*
 * Deterministic sin/cos — part of the libscenesubmit specification (SS_TRIG_DETERMINISTIC).
 *
 * The reference builds its rotation from glm::sin/glm::cos of the half Euler angles
 * (GLM/detail/type_quat.inl:208-217), i.e. from the host libm.  No two libms (glibc, MSVC UCRT,
 * CUDA libdevice) agree in the last ulp, and one ulp in a matrix flips borderline frustum tests.
 * This header fixes ONE evaluation, written only with IEEE-754 binary32 +, -, * and rintf in a fixed
 * order, so that any host compiler with contraction off (-ffp-contract=off) and nvcc with
 * -fmad=false produce identical bits.  It is included by the CUDA kernels and by the CPU oracle:
 * it is a definition, like the key layout, not an implementation detail of either.
 *
 * Method: Cody-Waite reduction by pi/2 in three parts (the classic single-precision Cephes split,
 * doubled), then the Cephes minimax polynomials on |r| <= pi/4.  Max error against the exact value
 * is about 1.5 ulp for |x| <= 8192; beyond that the result is still deterministic but degrades.
 */
#ifndef SCENESUBMIT_DETTRIG_H
#define SCENESUBMIT_DETTRIG_H

#include <math.h>

#if defined(__CUDACC__)
#define SS_HD __host__ __device__ __forceinline__
#else
#define SS_HD static inline
#endif

SS_HD void ss_det_sincosf(float x, float* sin_out, float* cos_out)
{
    const float two_over_pi = 0.636619772367581343f;
    const float p1 = 1.5703125f;                 /* pi/2, leading 8 bits  */
    const float p2 = 4.837512969970703125e-4f;   /* next 11 bits          */
    const float p3 = 7.54978995489188216e-8f;    /* remainder             */

    float k = rintf(x * two_over_pi);
    float r = x - k * p1;
    r = r - k * p2;
    r = r - k * p3;

    float z = r * r;

    /* sin(r) = r + r*z*(S1 + z*(S2 + z*S3)) */
    float ps = -1.9515295891e-4f * z;
    ps = ps + 8.3321608736e-3f;
    ps = ps * z;
    ps = ps - 1.6666654611e-1f;
    ps = ps * z;
    ps = ps * r;
    float sr = ps + r;

    /* cos(r) = 1 - z/2 + z*z*(C1 + z*(C2 + z*C3)) */
    float pc = 2.443315711809948e-5f * z;
    pc = pc - 1.388731625493765e-3f;
    pc = pc * z;
    pc = pc + 4.166664568298827e-2f;
    pc = pc * z;
    pc = pc * z;
    float hz = 0.5f * z;
    pc = pc - hz;
    float cr = pc + 1.0f;

    int q = (fabsf(k) < 2.0e9f) ? (int)k : 0;
    q &= 3;
    float s = (q & 1) ? cr : sr;
    float c = (q & 1) ? sr : cr;
    if (q == 1 || q == 2) c = -c;
    if (q >= 2) s = -s;
    *sin_out = s;
    *cos_out = c;
}

#endif /* SCENESUBMIT_DETTRIG_H */
