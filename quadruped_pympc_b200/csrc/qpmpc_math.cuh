// qpmpc_math.cuh -- the rollout arithmetic written ONCE for two lane types:
//   T = float   one sample per thread (tick_kernel, tick_kernel_h, the four-lanes-per-sample kernel, the tail)
//   T = float2  two samples per thread in packed FP32 (tick_kernel_x2, qpmpc_x2.cuh): every multiply / add / fma below
//               becomes ONE FMUL2 / FADD2 / FFMA2 for both samples (sm_100a issues fma.rn.f32x2 as a single
//               instruction; block-uniform scalars ride along as `.F32` broadcast operands, no packing moves).
// Every operation is an explicit IEEE round-to-nearest intrinsic: the compiler can neither contract nor re-associate
// them, so where this header says fma it is an fma for both lane types and where it says mul + add it stays two
// roundings.  Result: sample i gets bit-identical costs from either kernel (tests/test_gpu_x2.py).
#pragma once
#include <cuda_runtime.h>

namespace qp {

// ---- lane arithmetic ------------------------------------------------------------------------------------------
__device__ __forceinline__ float qmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float qadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float qsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float qfma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float qneg(float a) { return -a; }
__device__ __forceinline__ float qmin(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ float qmax(float a, float b) { return fmaxf(a, b); }

__device__ __forceinline__ float2 bc2(float s) { return make_float2(s, s); }
__device__ __forceinline__ float2 qneg(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 qmul(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 qmul(float2 a, float b) { return __fmul2_rn(a, bc2(b)); }
__device__ __forceinline__ float2 qmul(float a, float2 b) { return __fmul2_rn(b, bc2(a)); }
__device__ __forceinline__ float2 qadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 qadd(float2 a, float b) { return __fadd2_rn(a, bc2(b)); }
__device__ __forceinline__ float2 qadd(float a, float2 b) { return __fadd2_rn(b, bc2(a)); }
__device__ __forceinline__ float2 qsub(float2 a, float2 b) { return __fadd2_rn(a, qneg(b)); }
__device__ __forceinline__ float2 qsub(float2 a, float b) { return __fadd2_rn(a, bc2(-b)); }
__device__ __forceinline__ float2 qsub(float a, float2 b) { return __fadd2_rn(qneg(b), bc2(a)); }
__device__ __forceinline__ float2 qfma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 qfma(float2 a, float b, float2 c) { return __ffma2_rn(a, bc2(b), c); }
__device__ __forceinline__ float2 qfma(float a, float2 b, float2 c) { return __ffma2_rn(b, bc2(a), c); }
__device__ __forceinline__ float2 qfma(float2 a, float2 b, float c) { return __ffma2_rn(a, b, bc2(c)); }
__device__ __forceinline__ float2 qfma(float a, float2 b, float c) { return __ffma2_rn(b, bc2(a), bc2(c)); }
__device__ __forceinline__ float2 qfma(float2 a, float b, float c) { return __ffma2_rn(a, bc2(b), bc2(c)); }
__device__ __forceinline__ float2 qmin(float2 a, float2 b) { return make_float2(fminf(a.x, b.x), fminf(a.y, b.y)); }
__device__ __forceinline__ float2 qmax(float2 a, float2 b) { return make_float2(fmaxf(a.x, b.x), fmaxf(a.y, b.y)); }
__device__ __forceinline__ float2 qmin(float2 a, float b) { return make_float2(fminf(a.x, b), fminf(a.y, b)); }
__device__ __forceinline__ float2 qmax(float2 a, float b) { return make_float2(fmaxf(a.x, b), fmaxf(a.y, b)); }

template <class T> __device__ __forceinline__ T qset(float s);
template <> __device__ __forceinline__ float qset<float>(float s) { return s; }
template <> __device__ __forceinline__ float2 qset<float2>(float s) { return make_float2(s, s); }

// a * b - c  and  c - a * b  with ONE rounding
template <class A, class B, class C> __device__ __forceinline__ auto qfms(A a, B b, C c) -> decltype(qfma(a, b, c)) { return qfma(a, b, qneg(c)); }
template <class A, class B, class C> __device__ __forceinline__ auto qfnma(A a, B b, C c) -> decltype(qfma(a, b, c)) { return qfma(qneg(a), b, c); }

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float2 rcp_approx(float2 x) { return make_float2(rcp_approx(x.x), rcp_approx(x.y)); }

// ---- the (x, y) components of one quantity as a unit --------------------------------------------------------------
// For T = float the x and y components of a leg's spline coefficients / force travel in ONE 64-bit register pair, so
// that the spline evaluation of the sample-per-thread kernels issues FFMA2 / FMUL2 / FADD2 too (the same IEEE operation
// per component: results are bit-identical to evaluating x and y separately).  For T = float2 (two samples per thread)
// the components stay two packed values.
template <class T> struct XY;
template <> struct XY<float> {
    float2 v;
    __device__ __forceinline__ float x() const { return v.x; }
    __device__ __forceinline__ float y() const { return v.y; }
};
template <> struct XY<float2> {
    float2 a, b;
    __device__ __forceinline__ float2 x() const { return a; }
    __device__ __forceinline__ float2 y() const { return b; }
};
__device__ __forceinline__ XY<float> xy_make(float x, float y) { XY<float> r; r.v = make_float2(x, y); return r; }
__device__ __forceinline__ XY<float2> xy_make(float2 x, float2 y) { XY<float2> r; r.a = x; r.b = y; return r; }
__device__ __forceinline__ XY<float> xy_add(XY<float> p, XY<float> q) { XY<float> r; r.v = __fadd2_rn(p.v, q.v); return r; }
__device__ __forceinline__ XY<float2> xy_add(XY<float2> p, XY<float2> q) { return xy_make(qadd(p.a, q.a), qadd(p.b, q.b)); }
__device__ __forceinline__ XY<float> xy_sub(XY<float> p, XY<float> q) { XY<float> r; r.v = __fadd2_rn(p.v, qneg(q.v)); return r; }
__device__ __forceinline__ XY<float2> xy_sub(XY<float2> p, XY<float2> q) { return xy_make(qsub(p.a, q.a), qsub(p.b, q.b)); }
__device__ __forceinline__ XY<float> xy_mul(XY<float> p, float s) { XY<float> r; r.v = __fmul2_rn(p.v, bc2(s)); return r; }
__device__ __forceinline__ XY<float2> xy_mul(XY<float2> p, float s) { return xy_make(qmul(p.a, s), qmul(p.b, s)); }
__device__ __forceinline__ XY<float> xy_fma(XY<float> p, float s, XY<float> c) { XY<float> r; r.v = __ffma2_rn(p.v, bc2(s), c.v); return r; }
__device__ __forceinline__ XY<float2> xy_fma(XY<float2> p, float s, XY<float2> c) { return xy_make(qfma(p.a, s, c.a), qfma(p.b, s, c.b)); }

// component-wise product / fused multiply-add of two units, and a unit of two block-uniform scalars stored next to each
// other at an 8-byte aligned address (one 64-bit load; packed operations take it as a 64-bit uniform operand)
__device__ __forceinline__ XY<float> xy_mulv(XY<float> p, XY<float> q) { XY<float> r; r.v = __fmul2_rn(p.v, q.v); return r; }
__device__ __forceinline__ XY<float2> xy_mulv(XY<float2> p, XY<float2> q) { return xy_make(qmul(p.a, q.a), qmul(p.b, q.b)); }
__device__ __forceinline__ XY<float> xy_fmav(XY<float> p, XY<float> q, XY<float> c) { XY<float> r; r.v = __ffma2_rn(p.v, q.v, c.v); return r; }
__device__ __forceinline__ XY<float2> xy_fmav(XY<float2> p, XY<float2> q, XY<float2> c) { return xy_make(qfma(p.a, q.a, c.a), qfma(p.b, q.b, c.b)); }
template <class T> __device__ __forceinline__ XY<T> xy_uniform2(float x, float y);
template <> __device__ __forceinline__ XY<float> xy_uniform2<float>(float x, float y) { return xy_make(x, y); }
template <> __device__ __forceinline__ XY<float2> xy_uniform2<float2>(float x, float y) { return xy_make(bc2(x), bc2(y)); }
template <class T> __device__ __forceinline__ XY<T> xy_uniform(const float *p);
template <> __device__ __forceinline__ XY<float> xy_uniform<float>(const float *p) { XY<float> r; r.v = *reinterpret_cast<const float2 *>(p); return r; }
template <> __device__ __forceinline__ XY<float2> xy_uniform<float2>(const float *p) { const float2 v = *reinterpret_cast<const float2 *>(p); return xy_make(bc2(v.x), bc2(v.y)); }

// ---- sine / cosine -------------------------------------------------------------------------------------------
// |x| <= pi/4 (every sane rollout): the same minimax polynomials sincosf evaluates after its range reduction, without the
// reduction and quadrant logic (~11 instructions instead of ~35, same <= 1 ulp accuracy); larger angles (diverged
// samples) take the library path.
template <class T>
__device__ __forceinline__ void sincos_poly(T x, T &sn, T &cs) {
    const T x2 = qmul(x, x);
    T ps = qfma(x2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = qfma(ps, x2, -1.6666654611e-1f);
    sn = qfma(qmul(x, x2), ps, x);
    T pc = qfma(x2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = qfma(pc, x2, 4.166664568298827e-2f);
    pc = qfma(pc, x2, -0.5f);
    cs = qfma(pc, x2, 1.0f);
}
__device__ __forceinline__ void sincos_small(float x, float *sn, float *cs) {
    if (fabsf(x) <= 0.78539816f) sincos_poly<float>(x, *sn, *cs);
    else sincosf(x, sn, cs);
}
__device__ __forceinline__ void sincos_small(float2 x, float2 *sn, float2 *cs) {
    if (fabsf(x.x) <= 0.78539816f && fabsf(x.y) <= 0.78539816f) sincos_poly<float2>(x, *sn, *cs);
    else { sincos_small(x.x, &sn->x, &cs->x); sincos_small(x.y, &sn->y, &cs->y); }
}

// Yaw is unweighted and unbounded, but every sample of a tick starts from the same yaw0: sin/cos(yaw0) are computed once
// per block (s_yaw = {yaw0, sin yaw0, cos yaw0}, stage_inputs) and each step only needs the small-angle polynomials of
// the drift d = yaw - yaw0 plus the angle-addition identities (~16 instructions instead of sincosf's ~35, <= 2 ulp).
__device__ __forceinline__ void sincos_yaw(float yaw, const float *s_yaw, float *sn, float *cs) {
    const float d = qsub(yaw, s_yaw[0]);
    if (fabsf(d) <= 0.78539816f) {
        float sd, cd;
        sincos_poly<float>(d, sd, cd);
        const float s0 = s_yaw[1], c0 = s_yaw[2];
        *sn = qfma(s0, cd, qmul(c0, sd));
        *cs = qfms(c0, cd, qmul(s0, sd));
    } else {
        sincosf(yaw, sn, cs);
    }
}
__device__ __forceinline__ void sincos_yaw(float2 yaw, const float *s_yaw, float2 *sn, float2 *cs) {
    const float2 d = qsub(yaw, s_yaw[0]);
    if (fabsf(d.x) <= 0.78539816f && fabsf(d.y) <= 0.78539816f) {
        float2 sd, cd;
        sincos_poly<float2>(d, sd, cd);
        const float s0 = s_yaw[1], c0 = s_yaw[2];
        *sn = qfma(s0, cd, qmul(c0, sd));
        *cs = qfms(c0, cd, qmul(s0, sd));
    } else {
        sincos_yaw(yaw.x, s_yaw, &sn->x, &cs->x);
        sincos_yaw(yaw.y, s_yaw, &sn->y, &cs->y);
    }
}

// One range test for the three angles of a step (roll, pitch, yaw drift): a single branch around three polynomial
// evaluations instead of three test-and-branch diamonds.  Same results: the fast path is the polynomials above, the slow
// path re-tests each angle on its own.
__device__ __forceinline__ bool angles_small(float a, float b, float c) {
    return fmaxf(fmaxf(fabsf(a), fabsf(b)), fabsf(c)) <= 0.78539816f;
}
__device__ __forceinline__ bool angles_small(float2 a, float2 b, float2 c) {
    return fmaxf(fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(b.x), fabsf(b.y))), fmaxf(fabsf(c.x), fabsf(c.y))) <= 0.78539816f;
}
// the polynomials of two angles: for one sample per thread as ONE packed evaluation (FFMA2 / FMUL2: the same IEEE
// operation per component, so the results are those of two scalar evaluations)
__device__ __forceinline__ void sincos_poly_pair(float a, float b, float &sa, float &ca, float &sb, float &cb) {
    float2 sn, cs;
    sincos_poly<float2>(make_float2(a, b), sn, cs);
    sa = sn.x; sb = sn.y; ca = cs.x; cb = cs.y;
}
__device__ __forceinline__ void sincos_poly_pair(float2 a, float2 b, float2 &sa, float2 &ca, float2 &sb, float2 &cb) {
    sincos_poly<float2>(a, sa, ca);
    sincos_poly<float2>(b, sb, cb);
}
template <class T>
__device__ __forceinline__ void sincos_rpy(T roll, T pitch, T yaw, const float *s_yaw, T &sr, T &cr, T &sp, T &cp, T &sy, T &cy) {
    const T d = qsub(yaw, s_yaw[0]);
    if (angles_small(roll, pitch, d)) {
        T sd, cd;
        sincos_poly_pair(roll, pitch, sr, cr, sp, cp);
        sincos_poly<T>(d, sd, cd);
        const float s0 = s_yaw[1], c0 = s_yaw[2];
        sy = qfma(s0, cd, qmul(c0, sd));
        cy = qfms(c0, cd, qmul(s0, sd));
    } else {
        sincos_small(roll, &sr, &cr);
        sincos_small(pitch, &sp, &cp);
        sincos_yaw(yaw, s_yaw, &sy, &cy);
    }
}

}  // namespace qp
