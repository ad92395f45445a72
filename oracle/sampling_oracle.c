/*
 * C twin of oracle/sampling_oracle.py -- TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline).
 *
 * PARITY UNPINNED (see the header of sampling_oracle.py): the reference has no golden vectors for
 * this path and cannot run in this image (needs jax/jaxlib 0.9.2, gym_quadruped, mujoco).  This file
 * restates, scalar per sample and in the reference's operation order, the same reference lines
 * (paths relative to /root/reference/quadruped_pympc/):
 *   S = controllers/sampling/centroidal_nmpc_jax.py, M = controllers/sampling/centroidal_model_jax.py,
 *   A = controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py, G = helpers/periodic_gait_generator_jax.py
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC sampling_oracle.c -o liboracle_srbd.so -lm
 * (-ffp-contract=off: no FMA contraction, so every + and * rounds once, like the NumPy oracle).
 *
 * The second half (orc_philox_*, orc_noise_*) is NOT a restatement of the reference: it is the host
 * mirror of the product's counter-based Philox4x32-10 noise generator (the reference uses JAX threefry,
 * which is not available here); it exists to check the device RNG bit-for-bit (integers) and to let the
 * timed CPU baseline do the same noise-generation work per tick as the reference tick does.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_MAX_H 64
#define ORC_MAX_NS 16
#define ORC_PHILOX_K0 0xA4093822u   /* fixed key of the product's generator (csrc/qpmpc_kernels.cuh PHILOX_K0/K1) */
#define ORC_PHILOX_K1 0x299F31D0u

enum { ORC_ZERO_ORDER = 0, ORC_LINEAR = 1, ORC_CUBIC = 2 };
enum { ORC_RANDOM_SAMPLING = 0, ORC_MPPI = 1, ORC_CEM_MPPI = 2 };

typedef struct {
    int32_t N, H, ns, param, method, adaptive;
    float mass;            /* config.py:19-72 */
    float inertia[9];      /* row-major */
    float dts[ORC_MAX_H];  /* M:42-53 */
    float bounds[ORC_MAX_NS + 1]; /* linspace(0,H,ns+1), S:187 */
    float mu, fz_min, fz_max; /* S:160-164 */
    float mg;              /* float32(mass*9.81), S:380 */
    float qdiag[24];       /* S:118-130 */
    float mpc_dt;          /* A:179 (pgg mpc_dt = dt) */
    float duty;            /* A:179 (0.65) */
    float sigma_rs[3];     /* config.py:193 */
    float sigma_mppi;      /* config.py:192 */
    float freq_avail[3];   /* config.py:109 */
    /* derived by orc_cfg_finalize */
    float inertia_inv[9];
    float inv_m;
    int32_t P_leg, P, elite_k;
} orc_cfg;

/* M:67-91 */
static void inv3(const float A[9], float out[9]) {
    float a11 = A[0], a12 = A[1], a13 = A[2], a21 = A[3], a22 = A[4], a23 = A[5], a31 = A[6], a32 = A[7], a33 = A[8];
    float DET = a11 * (a33 * a22 - a32 * a23) - a21 * (a33 * a12 - a32 * a13) + a31 * (a23 * a12 - a22 * a13);
    out[0] = (a33 * a22 - a32 * a23) / DET;  out[1] = -(a33 * a12 - a32 * a13) / DET; out[2] = (a23 * a12 - a22 * a13) / DET;
    out[3] = -(a33 * a21 - a31 * a23) / DET; out[4] = (a33 * a11 - a31 * a13) / DET;  out[5] = -(a23 * a11 - a21 * a13) / DET;
    out[6] = (a32 * a21 - a31 * a22) / DET;  out[7] = -(a32 * a11 - a31 * a12) / DET; out[8] = (a22 * a11 - a21 * a12) / DET;
}

void orc_cfg_finalize(orc_cfg *c) {
    inv3(c->inertia, c->inertia_inv);                 /* M:56 */
    c->inv_m = 1.0f / c->mass;                        /* M:118 */
    if (c->param == ORC_LINEAR) c->P_leg = (c->ns + 1) * 3;          /* S:56 */
    else if (c->param == ORC_CUBIC) c->P_leg = 4 * 3 * c->ns;        /* S:71 */
    else c->P_leg = c->H * 3;                                         /* S:84 */
    c->P = 4 * c->P_leg;
    c->elite_k = c->adaptive ? 200 : 10;              /* A:1115 | S:1075 */
}

static inline float dot3(const float *r, const float *v) { return (r[0] * v[0] + r[1] * v[1]) + r[2] * v[2]; }

/* M:93-162 ; x: 24 states, F: 12 forces, st: 4 stance flags -> d[12] */
static void fd(const orc_cfg *c, const float *x, const float *F, const float *st, float *d) {
    const float zero = 0.0f, one = 1.0f;
    float temp[3], temp2[3] = {0, 0, 0};
    for (int k = 0; k < 3; ++k)
        temp[k] = ((F[k] * st[0] + F[3 + k] * st[1]) + F[6 + k] * st[2]) + F[9 + k] * st[3];   /* M:111-116 */
    const float grav[3] = {0.0f, 0.0f, -9.81f};
    d[0] = x[3]; d[1] = x[4]; d[2] = x[5];
    for (int k = 0; k < 3; ++k) d[3 + k] = c->inv_m * temp[k] + grav[k];                         /* M:118 */
    const float *w = x + 9;
    float roll = x[6], pitch = x[7], yaw = x[8];
    float sr = sinf(roll), cr = cosf(roll), sp = sinf(pitch), cp = cosf(pitch), sy = sinf(yaw), cy = cosf(yaw);
    float E[9] = {one, zero, -sp, zero, cr, cp * sr, zero, -sr, cp * cr};                        /* M:124-130 */
    for (int l = 0; l < 4; ++l) {                                                                /* M:132-135 */
        float r[3] = {x[12 + 3 * l] - x[0], x[13 + 3 * l] - x[1], x[14 + 3 * l] - x[2]};
        const float *f = F + 3 * l;
        float sk0 = (zero * f[0] + (-r[2]) * f[1]) + r[1] * f[2];
        float sk1 = (r[2] * f[0] + zero * f[1]) + (-r[0]) * f[2];
        float sk2 = ((-r[1]) * f[0] + r[0] * f[1]) + zero * f[2];
        if (l == 0) { temp2[0] = sk0 * st[l]; temp2[1] = sk1 * st[l]; temp2[2] = sk2 * st[l]; }
        else { temp2[0] = temp2[0] + sk0 * st[l]; temp2[1] = temp2[1] + sk1 * st[l]; temp2[2] = temp2[2] + sk2 * st[l]; }
    }
    float Einv[9];
    inv3(E, Einv);                                                                               /* M:137 */
    for (int i = 0; i < 3; ++i) d[6 + i] = dot3(Einv + 3 * i, w);
    float R[9] = {cp * cy, cp * sy, -sp,                                                         /* M:141-155 */
                  sr * sp * cy - cr * sy, sr * sp * sy + cr * cy, sr * cp,
                  cr * sp * cy + sr * sy, cr * sp * sy - sr * cy, cr * cp};
    float Iw[3], wxIw[3], t1[3], Rt[3], t2[3];
    for (int i = 0; i < 3; ++i) Iw[i] = dot3(c->inertia + 3 * i, w);                             /* M:157 */
    wxIw[0] = (zero * Iw[0] + (-w[2]) * Iw[1]) + w[1] * Iw[2];
    wxIw[1] = (w[2] * Iw[0] + zero * Iw[1]) + (-w[0]) * Iw[2];
    wxIw[2] = ((-w[1]) * Iw[0] + w[0] * Iw[1]) + zero * Iw[2];
    for (int i = 0; i < 3; ++i) t1[i] = dot3(c->inertia_inv + 3 * i, wxIw);
    for (int i = 0; i < 3; ++i) Rt[i] = dot3(R + 3 * i, temp2);                                  /* M:158 */
    for (int i = 0; i < 3; ++i) t2[i] = dot3(c->inertia_inv + 3 * i, Rt);
    for (int i = 0; i < 3; ++i) d[9 + i] = -t1[i] + t2[i];
}

/* M:164-174 (feet are carried unchanged, so x[12..23] is never written) */
static void integrate(const orc_cfg *c, float *x, const float *F, const float *st, int n) {
    float d[12];
    fd(c, x, F, st, d);
    float dt = c->dts[n];
    for (int i = 0; i < 12; ++i) x[i] = x[i] + d[i] * dt;
}

/* S:187-189 */
static int chunk_index(const orc_cfg *c, float stepf) {
    int idx = 0;
    for (int i = 0; i <= c->ns; ++i) if (stepf >= c->bounds[i] && i > idx) idx = i;
    return idx;
}

static inline int wrap(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }

/* S:181-268 ; `denom` = horizon_leg/num_spline as evaluated by the caller (python float or f32 tracer) */
static void spline(const orc_cfg *c, const float *p, float stepf, float denom, float f[3]) {
    const int PL = c->P_leg;
    if (c->param == ORC_ZERO_ORDER) {                                /* S:259-268 */
        int index = (int)(int16_t)stepf;
        f[0] = p[wrap(index, PL)]; f[1] = p[wrap(index + c->H, PL)]; f[2] = p[wrap(index + 2 * c->H, PL)];
        return;
    }
    int index = chunk_index(c, stepf);
    float q = stepf / denom - (float)index;                          /* S:191-194 */
    if (c->param == ORC_LINEAR) {                                    /* S:196-199 */
        int sh = c->ns + 1;
        for (int a = 0; a < 3; ++a)
            f[a] = (1.0f - q) * p[wrap(index + sh * a, PL)] + q * p[wrap(index + sh * a + 1, PL)];
        return;
    }
    int s = 10 * index;                                              /* S:219 */
    float a = 2.0f * q * q * q - 3.0f * q * q + 1.0f;                /* S:222-225 */
    float b = (q * q * q - 2.0f * q * q + q) * 1.0f;
    float cc = -2.0f * q * q * q + 3.0f * q * q;
    float d = (q * q * q - q * q) * 1.0f;
    for (int ax = 0; ax < 3; ++ax) {                                 /* S:227-255 */
        int o = 4 * ax;
        float p0 = p[wrap(s + o, PL)], p1 = p[wrap(s + o + 1, PL)], p2 = p[wrap(s + o + 2, PL)], p3 = p[wrap(s + o + 3, PL)];
        float phi = 0.5f * (((p2 - p1) / 1.0f) + ((p1 - p0) / 1.0f));
        float phin = 0.5f * (((p3 - p2) / 1.0f) + ((p2 - p1) / 1.0f));
        f[ax] = a * p1 + b * phi + cc * p2 + d * phin;
    }
}

/* S:270-314, one leg at a time (legs are independent in the reference's where-chain) */
static void clamp_leg(const orc_cfg *c, float f[3]) {
    float fz = f[2] > c->fz_min ? f[2] : c->fz_min;
    fz = fz < c->fz_max ? fz : c->fz_max;
    float lo = -c->mu * fz, hi = c->mu * fz;
    float fx = f[0] > lo ? f[0] : lo; fx = fx < hi ? fx : hi;
    float fy = f[1] > lo ? f[1] : lo; fy = fy < hi ? fy : hi;
    f[0] = fx; f[1] = fy; f[2] = fz;
}

/* S:353-409 */
static void forces_at(const orc_cfg *c, const float *knots, const float stepf[4], const float denom[4],
                      const float cn[4], float F[12]) {
    float nst = ((cn[0] + cn[1]) + cn[2]) + cn[3];
    float fref = c->mg / nst;
    for (int l = 0; l < 4; ++l) {
        float f[3];
        spline(c, knots + l * c->P_leg, stepf[l], denom[l], f);
        f[2] = fref + f[2];
        f[0] = f[0] * cn[l] / 3.0f;
        f[1] = f[1] * cn[l] / 3.0f;
        f[2] = f[2] * cn[l];
        clamp_leg(c, f);
        F[3 * l] = f[0]; F[3 * l + 1] = f[1]; F[3 * l + 2] = f[2];
    }
}

static float error_cost(const orc_cfg *c, const float *x, const float *ref) {   /* S:450-451 */
    float acc = 0.0f;
    for (int j = 0; j < 24; ++j) {
        float e = x[j] - ref[j];
        float term = (e * c->qdiag[j]) * e;
        acc = (j == 0) ? term : acc + term;
    }
    return acc;
}

/* S:316-496 (fixed gait) and A:326-501 (adaptive), one sample */
static float rollout_one(const orc_cfg *c, const float *x0, const float *ref, const float *contact,
                         const float *timing, const float *knots, float freq) {
    float x[24];
    memcpy(x, x0, sizeof(x));
    float cost = 0.0f;
    const int H = c->H;
    if (!c->adaptive) {
        float den = (float)((double)H / (double)c->ns);
        float denom[4] = {den, den, den, den};
        for (int n = 0; n < H; ++n) {
            float cn[4] = {contact[n], contact[H + n], contact[2 * H + n], contact[3 * H + n]};
            float stepf[4] = {(float)n, (float)n, (float)n, (float)n};
            float F[12];
            forces_at(c, knots, stepf, denom, cn, F);
            integrate(c, x, F, cn, n);
            cost = cost + error_cost(c, x, ref);
        }
        return cost;
    }
    /* G:68-89,136-151 */
    float cs[4][ORC_MAX_H];
    float tt[4] = {timing[0], timing[1], timing[2], timing[3]};
    float hl[4] = {0, 0, 0, 0};
    for (int n = 0; n < H; ++n)
        for (int l = 0; l < 4; ++l) {
            tt[l] = tt[l] >= 1.0f ? 0.0f : tt[l];
            tt[l] = tt[l] + c->mpc_dt * freq;
            cs[l][n] = tt[l] < c->duty ? 1.0f : 0.0f;
        }
    float denom[4];
    for (int l = 0; l < 4; ++l) {                                    /* A:345-348 */
        float s = 0.0f;
        for (int n = 0; n < H; ++n) s = s + cs[l][n];
        hl[l] = s + 1.0f;
        denom[l] = hl[l] / (float)c->ns;
    }
    int nleg[4] = {-1, -1, -1, -1};
    for (int n = 0; n < H; ++n) {
        float cn[4] = {cs[0][n], cs[1][n], cs[2][n], cs[3][n]};
        float stepf[4];
        for (int l = 0; l < 4; ++l) { nleg[l] = (int)((float)nleg[l] + 1.0f * cn[l]); stepf[l] = (float)nleg[l]; }  /* A:353-356 */
        float F[12];
        forces_at(c, knots, stepf, denom, cn, F);
        integrate(c, x, F, cn, n);
        cost = cost + error_cost(c, x, ref);
    }
    cost = cost + (freq - 1.3f) * 100.0f * (freq - 1.3f);            /* A:500 */
    return cost;
}

/* the jit_vectorized_rollout seam (S:176-177 / A:186-187): params (N,P) -> costs (N) */
void orc_rollout(const orc_cfg *c, const float *state, const float *ref, const float *contact, const float *timing,
                 const float *params, const float *freqs, float *costs) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < c->N; ++i)
        costs[i] = rollout_one(c, state, ref, contact, timing, params + (size_t)i * c->P, freqs ? freqs[i] : 0.0f);
}

typedef struct {
    float grfs[12];
    float pred[24];
    float best_cost;
    float best_freq;
    int32_t best_index;
} orc_out;

static const float *g_sort_costs;
static int cmp_cost_idx(const void *a, const void *b) {
    int32_t ia = *(const int32_t *)a, ib = *(const int32_t *)b;
    float ca = g_sort_costs[ia], cb = g_sort_costs[ib];
    if (ca < cb) return -1;
    if (ca > cb) return 1;
    return (ia > ib) - (ia < ib);
}

/* One tick with injected noise: S:629-1094 / A twins.  `noise` (N,P) is additional_random_parameters.
 * best_new (P), costs (N), sigma_new (P, cem only), elite_idx (elite_k, cem only), scratch_params (N*P). */
void orc_compute_control(const orc_cfg *c, const float *state, const float *ref, const float *contact,
                         const float *timing, const float *best_old, const float *noise, const float *freqs,
                         float *scratch_params, float *costs, float *best_new, float *sigma_new,
                         int32_t *elite_idx, orc_out *out) {
    const int N = c->N, P = c->P;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        float *row = scratch_params + (size_t)i * P;
        const float *nz = noise + (size_t)i * P;
        for (int p = 0; p < P; ++p) row[p] = best_old[p] + nz[p];                  /* S:680 */
        float cst = rollout_one(c, state, ref, contact, timing, row, freqs ? freqs[i] : 0.0f);
        if (isnan(cst)) cst = 1000000.0f;                                          /* S:686-687 */
        if (isinf(cst)) cst = 1000000.0f;
        costs[i] = cst;
    }
    int bi = 0;
    for (int i = 1; i < N; ++i) if (costs[i] < costs[bi]) bi = i;                  /* S:690 */
    float bc = costs[bi];
    if (c->method == ORC_RANDOM_SAMPLING) {
        memcpy(best_new, scratch_params + (size_t)bi * P, sizeof(float) * P);      /* S:692 */
    } else {
        float denom = 0.0f;
        float *w = (float *)malloc(sizeof(float) * N);
        for (int i = 0; i < N; ++i) { w[i] = expf(-1.0f * (costs[i] - bc)); denom = denom + w[i]; }  /* S:828-831 */
        for (int p = 0; p < P; ++p) best_new[p] = 0.0f;
        for (int i = 0; i < N; ++i) {
            float wi = w[i] / denom;                                               /* S:832 */
            const float *nz = noise + (size_t)i * P;
            for (int p = 0; p < P; ++p) best_new[p] = best_new[p] + wi * nz[p];    /* S:833-836 */
        }
        for (int p = 0; p < P; ++p) best_new[p] = best_old[p] + best_new[p];
        free(w);
    }
    /* tail S:695-784 : spline(best_leg, 0.0, 1) */
    {
        float cn[4] = {contact[0], contact[c->H], contact[2 * c->H], contact[3 * c->H]};
        float den = (float)(1.0 / (double)c->ns);
        float stepf[4] = {0, 0, 0, 0}, denom[4] = {den, den, den, den};
        float F[12], x[24];
        forces_at(c, best_new, stepf, denom, cn, F);
        memcpy(out->grfs, F, sizeof(F));
        memcpy(x, state, sizeof(x));
        integrate(c, x, F, cn, 0);
        memcpy(out->pred, x, sizeof(x));
    }
    out->best_cost = bc;
    out->best_index = bi;
    if (c->adaptive) out->best_freq = freqs[bi];
    else out->best_freq = (c->method == ORC_CEM_MPPI) ? 1.65f : 1.4f;
    if (c->method == ORC_CEM_MPPI) {                                               /* S:1075-1081 | A:1115-1121 */
        int k = c->elite_k < N ? c->elite_k : N;
        int32_t *order = (int32_t *)malloc(sizeof(int32_t) * N);
        for (int i = 0; i < N; ++i) order[i] = i;
        g_sort_costs = costs;
        qsort(order, N, sizeof(int32_t), cmp_cost_idx);
        for (int j = 0; j < k; ++j) elite_idx[j] = order[j];
        for (int p = 0; p < P; ++p) {
            float mean = 0.0f;
            for (int j = 0; j < k; ++j) mean = mean + noise[(size_t)order[j] * P + p];
            mean = mean / (float)k;
            float ss = 0.0f;
            for (int j = 0; j < k; ++j) { float xx = noise[(size_t)order[j] * P + p] - mean; ss = ss + xx * xx; }
            float var = ss / (float)(k - 1) + 1e-8f;
            float s;
            if (c->adaptive) { s = var > 3.0f ? 3.0f : var; s = s < 0.5f ? 0.5f : s; }
            else { s = sqrtf(var); s = s > 5.0f ? 5.0f : s; s = s < 0.2f ? 0.2f : s; }
            sigma_new[p] = s;
        }
        free(order);
    }
}

/* ===================================================================================================
 * Host mirror of the product's noise generator (NOT reference code; see file header).
 * Philox4x32-10 (Salmon et al., SC'11), counter = (oct [uniform rows: quad] | stream << 8 | robot << 9, row, tick, seed), fixed key
 * (ORC_PHILOX_K0, ORC_PHILOX_K1) -- the same convention as csrc/qpmpc_kernels.cuh.
 * =================================================================================================== */
static inline void philox_round(uint32_t c[4], const uint32_t k[2]) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k[0];
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k[1];
    uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}

void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c[4] = {ctr[0], ctr[1], ctr[2], ctr[3]};
    uint32_t k[2] = {key[0], key[1]};
    for (int r = 0; r < 10; ++r) {
        philox_round(c, k);
        k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u;
    }
    memcpy(out, c, sizeof(c));
}

static inline float u01_from_bits(uint32_t b) {       /* mantissa trick of jax.random.uniform: [1,2) - 1 */
    uint32_t v = (b >> 9) | 0x3F800000u;
    float f;
    memcpy(&f, &v, 4);
    return f - 1.0f;
}

/* Box-Muller exactly as the device does it (csrc/qpmpc_kernels.cuh box_muller16), with libm in place of the SFU
 * approximations: ONE word -> two normals (radius from bits 7..22, angle from the other 16 bits = bits 7..22 of the word
 * rotated by 16), the row's scale folded into the radius constant c = -2 ln2 * s^2. */
static inline void box_muller16(uint32_t w, float c, float *z0, float *z1) {
    union { uint32_t u; float f; } ua, ub;
    ua.u = (w & 0x007fff80u) | 0x3F800000u;
    ub.u = (((w << 16) | (w >> 16)) & 0x007fff80u) | 0x40000000u;
    float u = 2.0f - ua.f;
    float th = ub.f * 3.14159265f + -9.42477796f;
    float r = sqrtf(c * log2f(u));
    *z0 = r * cosf(th);
    *z1 = r * sinf(th);
}

/* additional_random_parameters (N,P) for one tick; sigma (P) used by cem_mppi only.  Row/segment mapping:
 * random_sampling (S:647-677): rows 1..k and k+1..2k share the same normals (same key in the reference),
 * rows 2k+1.. are uniform(-s2,s2); mppi/cem (S:806-812,951-958): rows 1..N-1.
 * Normal rows: one Philox block per OCT (8 knots), counter word 0 = oct | robot << 9; uniform rows: one block per quad
 * (4 knots, 23-bit uniforms), counter word 0 = quad | robot << 9. */
void orc_noise_generate(const orc_cfg *c, uint32_t seed, uint32_t robot, uint32_t tick, const float *sigma,
                        int row_begin, int row_end, float *noise /* (row_end-row_begin, P) */) {
    const int N = c->N, P = c->P;
    const int k = N / 3;
#pragma omp parallel for schedule(static)
    for (int i = row_begin; i < row_end; ++i) {
        float *row = noise + (size_t)(i - row_begin) * P;
        if (i == 0) { for (int p = 0; p < P; ++p) row[p] = 0.0f; continue; }
        int seg = 0, j = i - 1;
        if (c->method == ORC_RANDOM_SAMPLING) {
            if (i >= 1 + 2 * k) { seg = 2; j = i - 1 - 2 * k; }
            else if (i >= 1 + k) { seg = 1; j = i - 1 - k; }
        }
        const uint32_t key[2] = {ORC_PHILOX_K0, ORC_PHILOX_K1};
        if (c->method == ORC_RANDOM_SAMPLING && seg == 2) {
            const float s2 = c->sigma_rs[2];
            for (int qd = 0; qd < P / 4; ++qd) {
                uint32_t ctr[4] = {(uint32_t)qd | (robot << 9), (uint32_t)j, tick, seed}, r[4];
                orc_philox4x32_10(ctr, key, r);
                for (int e = 0; e < 4; ++e) {
                    float v = u01_from_bits(r[e]) * (s2 - (-s2)) + (-s2);
                    row[4 * qd + e] = v > -s2 ? v : -s2;
                }
            }
            continue;
        }
        const float scale = c->method == ORC_RANDOM_SAMPLING ? c->sigma_rs[seg] : c->sigma_mppi;
        const float bc = c->method == ORC_CEM_MPPI ? -1.38629436f : (-1.38629436f * scale) * scale;
        for (int oc = 0; 8 * oc < P; ++oc) {
            uint32_t ctr[4] = {(uint32_t)oc | (robot << 9), (uint32_t)j, tick, seed}, r[4];
            orc_philox4x32_10(ctr, key, r);
            float z[8];
            for (int w = 0; w < 4; ++w) box_muller16(r[w], bc, &z[2 * w], &z[2 * w + 1]);
            for (int e = 0; e < 8 && 8 * oc + e < P; ++e) {
                const int p = 8 * oc + e;
                row[p] = c->method == ORC_CEM_MPPI ? z[e] * sigma[p] : z[e];
            }
        }
    }
}

/* per-sample choice index in {0,1,2} (stream 1): floor(bits * 3 / 2^32) */
void orc_freq_choice(uint32_t seed, uint32_t robot, uint32_t tick, int row_begin, int row_end, int32_t *choice) {
    for (int i = row_begin; i < row_end; ++i) {
        uint32_t ctr[4] = {0x100u | (robot << 9), (uint32_t)i, tick, seed}, key[2] = {ORC_PHILOX_K0, ORC_PHILOX_K1}, r[4];
        orc_philox4x32_10(ctr, key, r);
        choice[i - row_begin] = (int32_t)(((uint64_t)r[0] * 3u) >> 32);
    }
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
