// Packed-FP32 (two instances per thread) arithmetic shared by the instance kernel (transform_cull_x2.cu) and the fused
// submission kernel (submit.cu): f32x2 helpers with exact unfused semantics, TMA bulk copy + mbarrier wrappers,
// TRS -> affine, affine product, box-vs-planes test.  See transform_cull_x2.cu for why every operation is one fma
// with a neutral operand.
#pragma once
#include "ss_kernels.h"

namespace ss {
namespace {

typedef unsigned long long f2;            // {lo = first instance, hi = second instance}

__device__ __forceinline__ f2 pk(float lo, float hi)
{
    f2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float lo(f2 v)
{
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return a;
}
__device__ __forceinline__ float hi(f2 v)
{
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return b;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c)
{
    f2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// neutral operands, passed at run time (see header comment)
struct Neutral {
    f2 one, neg_one, neg_zero;
};

struct P2 {
    Neutral k;
    __device__ __forceinline__ f2 mul(f2 a, f2 b) const { return fma2(a, b, k.neg_zero); }
    __device__ __forceinline__ f2 add(f2 a, f2 b) const { return fma2(a, k.one, b); }
    __device__ __forceinline__ f2 sub(f2 a, f2 b) const { return fma2(b, k.neg_one, a); }
};

__device__ __forceinline__ f2 dup(float c) { return pk(c, c); }
__device__ __forceinline__ f2 abs2(f2 v) { return v & 0x7fffffff7fffffffull; }

// loads that must be ISSUED where they are written (volatile keeps ptxas from sinking them to their use)
__device__ __forceinline__ float4 ld_early_f4(const float4* p)
{
    float4 r;
    asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint64_t ld_early_u64(const uint64_t* p)
{
    uint64_t r;
    asm volatile("ld.global.nc.u64 %0, [%1];" : "=l"(r) : "l"(p));
    return r;
}

// ---- TMA bulk copy + mbarrier (PTX) --------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}

// include/scenesubmit_dettrig.h, both halves at once; same operations in the same order
__device__ __forceinline__ void det_sincos2(const P2& m, f2 x, f2& sin_out, f2& cos_out)
{
    const f2 xm = m.mul(x, dup(0.636619772367581343f));
    const float k0 = rintf(lo(xm)), k1 = rintf(hi(xm));
    const f2 k = pk(k0, k1);
    f2 r = m.sub(x, m.mul(k, dup(1.5703125f)));
    r = m.sub(r, m.mul(k, dup(4.837512969970703125e-4f)));
    r = m.sub(r, m.mul(k, dup(7.54978995489188216e-8f)));
    const f2 z = m.mul(r, r);

    f2 ps = m.mul(dup(-1.9515295891e-4f), z);
    ps = m.add(ps, dup(8.3321608736e-3f));
    ps = m.mul(ps, z);
    ps = m.sub(ps, dup(1.6666654611e-1f));
    ps = m.mul(ps, z);
    ps = m.mul(ps, r);
    const f2 sr = m.add(ps, r);

    f2 pc = m.mul(dup(2.443315711809948e-5f), z);
    pc = m.sub(pc, dup(1.388731625493765e-3f));
    pc = m.mul(pc, z);
    pc = m.add(pc, dup(4.166664568298827e-2f));
    pc = m.mul(pc, z);
    pc = m.mul(pc, z);
    const f2 hz = m.mul(dup(0.5f), z);
    pc = m.sub(pc, hz);
    const f2 cr = m.add(pc, m.k.one);

    float s[2], c[2];
    const float srr[2] = {lo(sr), hi(sr)}, crr[2] = {lo(cr), hi(cr)}, kk[2] = {k0, k1};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        int q = (fabsf(kk[h]) < 2.0e9f) ? (int)kk[h] : 0;
        q &= 3;
        float sv = (q & 1) ? crr[h] : srr[h];
        float cv = (q & 1) ? srr[h] : crr[h];
        if (q == 1 || q == 2) cv = -cv;
        if (q >= 2) sv = -sv;
        s[h] = sv;
        c[h] = cv;
    }
    sin_out = pk(s[0], s[1]);
    cos_out = pk(c[0], c[1]);
}

struct Aff2 {
    f2 r[3][4];  // row r, column c of two affine matrices
};

// trs_to_affine of ss_device.cuh on two instances
__device__ __forceinline__ void trs_to_affine2(const P2& m, const f2* t9, Aff2& out)
{
    const f2 half = dup(0.5f), two = dup(2.0f), one = m.k.one;
    f2 sx, cx, sy, cy, sz, cz;
    det_sincos2(m, m.mul(t9[3], half), sx, cx);
    det_sincos2(m, m.mul(t9[4], half), sy, cy);
    det_sincos2(m, m.mul(t9[5], half), sz, cz);
    const f2 cxcy = m.mul(cx, cy), sxsy = m.mul(sx, sy), sxcy = m.mul(sx, cy), cxsy = m.mul(cx, sy);
    const f2 qw = m.add(m.mul(cxcy, cz), m.mul(sxsy, sz));
    const f2 qx = m.sub(m.mul(sxcy, cz), m.mul(cxsy, sz));
    const f2 qy = m.add(m.mul(cxsy, cz), m.mul(sxcy, sz));
    const f2 qz = m.sub(m.mul(cxcy, sz), m.mul(sxsy, cz));

    const f2 qxx = m.mul(qx, qx), qyy = m.mul(qy, qy), qzz = m.mul(qz, qz);
    const f2 qxz = m.mul(qx, qz), qxy = m.mul(qx, qy), qyz = m.mul(qy, qz);
    const f2 qwx = m.mul(qw, qx), qwy = m.mul(qw, qy), qwz = m.mul(qw, qz);

    const f2 R00 = m.sub(one, m.mul(two, m.add(qyy, qzz))), R01 = m.mul(two, m.add(qxy, qwz)), R02 = m.mul(two, m.sub(qxz, qwy));
    const f2 R10 = m.mul(two, m.sub(qxy, qwz)), R11 = m.sub(one, m.mul(two, m.add(qxx, qzz))), R12 = m.mul(two, m.add(qyz, qwx));
    const f2 R20 = m.mul(two, m.add(qxz, qwy)), R21 = m.mul(two, m.sub(qyz, qwx)), R22 = m.sub(one, m.mul(two, m.add(qxx, qyy)));

    const f2 s0 = t9[6], s1 = t9[7], s2 = t9[8];
    out.r[0][0] = m.mul(R00, s0); out.r[0][1] = m.mul(R10, s1); out.r[0][2] = m.mul(R20, s2); out.r[0][3] = t9[0];
    out.r[1][0] = m.mul(R01, s0); out.r[1][1] = m.mul(R11, s1); out.r[1][2] = m.mul(R21, s2); out.r[1][3] = t9[1];
    out.r[2][0] = m.mul(R02, s0); out.r[2][1] = m.mul(R12, s1); out.r[2][2] = m.mul(R22, s2); out.r[2][3] = t9[2];
}

// aff_mul of ss_device.cuh: out = a * b, glm column accumulation order
__device__ __forceinline__ void aff_mul2(const P2& m, const Aff2& a, const Aff2& b, Aff2& out)
{
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            f2 acc = m.add(m.mul(a.r[r][0], b.r[0][c]), m.mul(a.r[r][1], b.r[1][c]));
            acc = m.add(acc, m.mul(a.r[r][2], b.r[2][c]));
            if (c == 3) acc = m.add(acc, a.r[r][3]);
            out.r[r][c] = acc;
        }
    }
}

// Duplicated plane constants of one axis (two planes) in shared memory:
//   n0.xyz, |n0|.xyz, d0, d1, then (unpaired axes only) n1.xyz, |n1|.xyz
constexpr int kAxisWords = 14;  // f2 words per axis
constexpr int kViewWords = 3 * kAxisWords;

// cull_box of ss_device.cuh on two instances; returns the two masks
__device__ __forceinline__ uint2 cull_box2(const P2& m, const Aff2& w, const f2* mn, const f2* mx, const f2* s_planes,
                                           const uint32_t* s_flags, uint32_t n_views)
{
    const f2 half = dup(0.5f);
    f2 cl[3], el[3], cw[3], ew[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        cl[k] = m.mul(m.add(mn[k], mx[k]), half);
        el[k] = m.mul(m.sub(mx[k], mn[k]), half);
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        cw[r] = m.add(m.add(m.mul(w.r[r][0], cl[0]), m.mul(w.r[r][1], cl[1])), m.add(m.mul(w.r[r][2], cl[2]), w.r[r][3]));
        ew[r] = m.add(m.add(m.mul(abs2(w.r[r][0]), el[0]), m.mul(abs2(w.r[r][1]), el[1])), m.mul(abs2(w.r[r][2]), el[2]));
    }
    uint2 mask = make_uint2(0u, 0u);
    for (uint32_t v = 0; v < n_views; ++v) {
        const uint32_t fl = s_flags[v];
        bool in0 = true, in1 = true;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            // 128-bit shared loads: two duplicated constants each
            const ulonglong2* c2 = reinterpret_cast<const ulonglong2*>(s_planes + v * kViewWords + ax * kAxisWords);
            const ulonglong2 n01 = c2[0], n2a0 = c2[1], a12 = c2[2], d01 = c2[3];
            const f2 t0 = m.add(m.add(m.mul(n01.x, cw[0]), m.mul(n01.y, cw[1])), m.mul(n2a0.x, cw[2]));
            const f2 r0 = m.add(m.add(m.mul(n2a0.y, ew[0]), m.mul(a12.x, ew[1])), m.mul(a12.y, ew[2]));
            const f2 s0 = m.add(t0, d01.x);  // a skipped near plane has d0 = +inf: always inside
            f2 s1, r1 = r0;
            if (fl & (1u << ax)) {  // opposite normals: the second plane's dot product is -t0 exactly
                s1 = m.sub(d01.y, t0);
            } else {
                const ulonglong2 q01 = c2[4], q2b0 = c2[5], b12 = c2[6];
                const f2 t1 = m.add(m.add(m.mul(q01.x, cw[0]), m.mul(q01.y, cw[1])), m.mul(q2b0.x, cw[2]));
                r1 = m.add(m.add(m.mul(q2b0.y, ew[0]), m.mul(b12.x, ew[1])), m.mul(b12.y, ew[2]));
                s1 = m.add(t1, d01.y);
            }
            in0 = in0 && (lo(s0) >= -lo(r0)) && (lo(s1) >= -lo(r1));
            in1 = in1 && (hi(s0) >= -hi(r0)) && (hi(s1) >= -hi(r1));
        }
        if (in0) mask.x |= (1u << v);
        if (in1) mask.y |= (1u << v);
    }
    return mask;
}

}  // namespace
}  // namespace ss
