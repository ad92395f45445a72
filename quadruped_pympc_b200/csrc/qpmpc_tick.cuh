// qpmpc_tick.cuh -- the fused tick kernel and its merge / finish stages (see qpmpc_kernels.cuh for the math).
#pragma once
#include "qpmpc_kernels.cuh"
#include <math_constants.h>

namespace qp {

constexpr int MAXV = 6;                 // P <= MAXV * TB knots
static_assert(MAXV * TB / 4 <= 256, "the quad index must fit bits 0-7 of Philox counter word 0 (qpmpc_kernels.cuh uniform_quad / noise_oct_nz)");
constexpr int SCRATCH_KEYS = 2048;      // u64 keys for the cross-shard elite merge (world * elite_k <= 2048)
constexpr int SCRATCH_FLOATS = 2 * SCRATCH_KEYS + 256 + 16;   // keys + histogram + misc

// shared-memory carve-up (float offsets); every region starts 16-byte aligned
struct Smem {
    float *in, *fref, *yaw, *cref, *feet4, *w, *v, *red, *tile;
    int *cmask;
    int *idx;                               // MAX_ELITE elite indices (CEM refit)
    __device__ Smem(float *base, const KParams &kp, int nt = TB) {
        in = base;
        fref = in + kp.in_stride;
        yaw = fref + MAX_H;                 // {yaw0, sin yaw0, cos yaw0, -}: directly behind fref (rollouts rely on it)
        cmask = reinterpret_cast<int *>(yaw + 4);   // stance mask per step (bit l = leg l), directly behind yaw
        cref = yaw + 4 + MAX_H;                     // -(reference) of the cost pairs (KParams::qpair), directly behind cmask
        feet4 = cref + 8;                           // the four feet as (x, y, z, mu): one 128-bit load per stance leg and step
        w = feet4 + 16;                             // 2 * nt floats: compacted (weight, sample slot) pairs of one tile
        v = w + 2 * nt;
        red = v + kp.P;
        idx = reinterpret_cast<int *>(red + 16);
        tile = red + 16 + MAX_ELITE;
    }
};
// nt: threads per block; tile_cols: tile row stride the kernel needs (nt + 1 for the sample-per-thread kernel and for
// the CEM elite refit, nt / 4 + 1 for the four-lanes-per-sample kernel without CEM)
__host__ __device__ inline size_t smem_floats(int in_stride, int P, int nt = TB, int tile_cols = TS) {
    const size_t tile = (size_t)P * tile_cols > (size_t)SCRATCH_FLOATS ? (size_t)P * tile_cols : (size_t)SCRATCH_FLOATS;
    return (size_t)in_stride + 2 * MAX_H + 4 + 8 + 16 + 2 * nt + P + 16 + MAX_ELITE + tile + 4;
}

template <int NT>
__device__ __forceinline__ void stage_inputs(const KParams &kp, int r, const Smem &s) {
    constexpr int TB = NT;
    const float *gin = kp.param_packet ? kp.packet : kp.in + (size_t)r * kp.in_stride;
    if (kp.ctrl_state == nullptr) {
        for (int k = threadIdx.x; k < kp.in_stride; k += TB) s.in[k] = gin[k];
        if (kp.contact_dev) {                                     // tables generated on the device replace the packet's
            __syncthreads();
            for (int k = threadIdx.x; k < 4 * kp.H; k += TB) s.in[IN_CONTACT + k] = __ldcg(kp.contact_dev + (size_t)r * 4 * kp.H + k);
        }
        __syncthreads();
    } else {
        // Stateful controller: the packet carries what the host knows this tick (raw state, reference, phase, frequency,
        // contact table); best, sigma, the Philox tick and the previous contact come from the device-resident record.
        // Then the host prep of the reference, S:563-627 (prepare_state_and_reference), on the shared-memory copy:
        const float *st = kp.ctrl_state + (size_t)r * kp.ctrl_stride;
        for (int k = threadIdx.x; k < kp.off_best; k += TB)
            s.in[k] = (kp.contact_dev && k >= IN_CONTACT) ? __ldcg(kp.contact_dev + (size_t)r * 4 * kp.H + (k - IN_CONTACT)) : gin[k];
        for (int k = threadIdx.x; k < 2 * kp.P; k += TB) {
            float v = __ldcg(st + k);
            if (kp.sigma_reset && k >= kp.P) v = kp.sigma0;
            s.in[kp.off_best + k] = v;
        }
        __syncthreads();
        if (kp.shift_kind) {        // shift_solution S:513-561: the spline value at `step` goes into knots [0], [2], [4] of every
            if (threadIdx.x < 4) {  // leg (sic); numpy float32 arithmetic, no contraction.  One thread per leg.
                float *p = s.in + kp.off_best + threadIdx.x * kp.P_leg;
                const int i = kp.shift_index;
                const float w0 = kp.shift_w[0], w1 = kp.shift_w[1], w2 = kp.shift_w[2], w3 = kp.shift_w[3];
                float f[3];
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    if (kp.shift_kind == 1) {                                   // zero-order, S:259-268
                        f[a] = p[i + a * kp.H];
                    } else if (kp.shift_kind == 2) {                            // linear, S:181-201
                        const int o = i + (kp.ns + 1) * a;
                        f[a] = __fadd_rn(__fmul_rn(w0, p[o]), __fmul_rn(w1, p[o + 1]));
                    } else {                                                    // cubic, S:204-257
                        const int o = 10 * i + 4 * a;
                        const float p0 = p[o], p1 = p[o + 1], p2 = p[o + 2], p3 = p[o + 3];
                        const float phi = __fmul_rn(0.5f, __fadd_rn(__fsub_rn(p2, p1), __fsub_rn(p1, p0)));
                        const float phin = __fmul_rn(0.5f, __fadd_rn(__fsub_rn(p3, p2), __fsub_rn(p2, p1)));
                        f[a] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w0, p1), __fmul_rn(w1, phi)), __fmul_rn(w2, p2)), __fmul_rn(w3, phin));
                    }
                }
                p[0] = f[0]; p[2] = f[1]; p[4] = f[2];
            }
            __syncthreads();
        }
        const unsigned int prev = __float_as_uint(__ldcg(st + 2 * kp.P + 1));
        if (threadIdx.x == 0) s.in[IN_TICK] = __ldcg(st + 2 * kp.P);
        if (threadIdx.x < 12) {                                   // swing feet <- reference footholds, S:588-595
            const int leg = threadIdx.x / 3;
            if (s.in[IN_CONTACT + leg * kp.H] == 0.0f) s.in[IN_STATE + 12 + threadIdx.x] = s.in[IN_REF + 12 + threadIdx.x];
        }
        for (int k = threadIdx.x; k < kp.P; k += TB) {            // knots of a leg that just lifted off are zeroed, S:612-625
            const int leg = k / kp.P_leg;
            if (((prev >> leg) & 1u) && s.in[IN_CONTACT + leg * kp.H] == 0.0f) s.in[kp.off_best + k] = 0.0f;
        }
        __syncthreads();
    }
    if (threadIdx.x < kp.H) {       // gravity compensation per step, S:377-380
        const float *c = s.in + IN_CONTACT;
        const int n = threadIdx.x, H = kp.H;
        const float nst = ((c[n] + c[H + n]) + c[2 * H + n]) + c[3 * H + n];
        s.fref[n] = kp.mg / nst;
        s.cmask[n] = (c[n] != 0.0f ? 1 : 0) | (c[H + n] != 0.0f ? 2 : 0) | (c[2 * H + n] != 0.0f ? 4 : 0) | (c[3 * H + n] != 0.0f ? 8 : 0);
    }
    if (threadIdx.x >= 32 && threadIdx.x < 40) {    // -(reference) in the pair order of the step cost (srbd_step)
        const int qi[8] = {3, 4, 2, 5, 6, 7, 9, 10};
        s.cref[threadIdx.x - 32] = -s.in[IN_REF + qi[threadIdx.x - 32]];
    }
    if (threadIdx.x >= 64 && threadIdx.x < 80) {    // feet padded to four floats each, the friction coefficient in the fourth (rollout_fixed_t)
        const int i = threadIdx.x - 64;
        s.feet4[i] = (i & 3) == 3 ? kp.mu : s.in[IN_STATE + 12 + 3 * (i >> 2) + (i & 3)];
    }
    if (threadIdx.x == TB - 1) {    // block-uniform yaw0 and its sine / cosine (sincos_yaw)
        const float y0 = s.in[IN_STATE + 8];
        float sn, cs;
        sincosf(y0, &sn, &cs);
        s.yaw[0] = y0; s.yaw[1] = sn; s.yaw[2] = cs; s.yaw[3] = 0.0f;
    }
    __syncthreads();
}

__device__ __forceinline__ bool lex_less(float c, int i, float oc, int oi) { return c < oc || (c == oc && i < oi); }

// lexicographic (cost, index) minimum over a warp with two REDUX instructions: the cost through the order-preserving map
// of its bits to unsigned (NaN-free, -0 folded into +0 by the caller's arithmetic: costs are sums of squares), then the
// smallest index among the lanes that hold that cost.  Indices are >= 0.  Result valid in every lane.
__device__ __forceinline__ void warp_argmin(float &c, int &i) {
    const uint32_t b = __float_as_uint(c + 0.0f);
    const uint32_t key = b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
    const uint32_t kmin = __reduce_min_sync(0xffffffffu, key);
    i = (int)__reduce_min_sync(0xffffffffu, key == kmin ? (uint32_t)i : 0xffffffffu);
    c = __uint_as_float(kmin ^ ((kmin >> 31) ? 0x80000000u : 0xffffffffu));
}

// Online-softmax merge of `n` records (m, idx, s, V[P]) spaced `stride` floats apart in global memory.
// Result: M, IDX (lexicographic min), S, and V in s.v (MPPI / CEM only).  All TB threads participate.
template <int METHOD, int NT>
__device__ void merge_records(const float *recs, int n, size_t stride, int P, const Smem &s, float &M, int &IDX, float &S,
                              const KParams *dbg = nullptr) {
    constexpr int TB = NT;
    constexpr int PF = 16;                                 // V rows per thread fetched ahead of the (min, argmin) reduction
    const int t = threadIdx.x;
    // The V rows do not depend on M: their loads go out first so that their L2 round trip overlaps the one of the
    // (m, idx) loads and the block reduction instead of following it (single tick: 79 records, one serial chain less).
    const int nc = P >> 2;
    const bool fast = METHOD != RANDOM_SAMPLING && (P & 3) == 0 && nc <= TB && (stride & 3) == 0;
    const int G = fast ? TB / nc : 1;
    const int g = fast ? t / nc : 0, c4 = fast ? t - g * nc : 0;
    const size_t stride4 = stride >> 2;
    float4 pv[PF];
    if (fast) {
        const float4 *col = reinterpret_cast<const float4 *>(recs + REC_V) + c4;
        const int cnt0 = n < TB ? n : TB;
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int k = g + u * G;
            pv[u] = (g < G && k < cnt0) ? col[(size_t)k * stride4] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    float c = CUDART_INF_F;
    int i = 0x7fffffff;
    for (int b = t; b < n; b += TB) {
        const float m = recs[b * stride + REC_M];
        const int id = __float_as_int(recs[b * stride + REC_IDX]);
        if (lex_less(m, id, c, i)) { c = m; i = id; }
    }
    block_argmin<TB>(c, i, s.red);
    M = c; IDX = i; S = 0.0f;
    if (dbg) phase_mark(*dbg, 12);
    if (METHOD == RANDOM_SAMPLING) return;
    // Fast path (P % 4 == 0, P/4 <= TB: every default configuration): float4 loads of the V rows, the records split
    // over G = TB / (P/4) thread groups (independent coalesced loads in flight instead of one serial chain per knot),
    // group partials combined in a fixed order -> deterministic.
    if (fast) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int base = 0; base < n; base += TB) {
            const int b = base + t;
            float sc = 0.0f, sv = 0.0f;
            if (b < n) {
                const float m = recs[b * stride + REC_M];
                sc = (m == CUDART_INF_F) ? 0.0f : expf(-(m - M));
                sv = recs[b * stride + REC_S] * sc;
            }
            __syncthreads();
            s.w[t] = sc;
            S += block_sum<TB>(sv, s.red);          // contains the barriers that publish s.w
            if (dbg && base == 0) phase_mark(*dbg, 13);
            const int cnt = (n - base) < TB ? (n - base) : TB;
            if (g < G) {
                const float4 *col = reinterpret_cast<const float4 *>(recs + (size_t)base * stride + REC_V) + c4;
                int k_first = g;
                if (base == 0) {                                 // the prefetched rows, same record order
#pragma unroll
                    for (int u = 0; u < PF; ++u) {
                        const int k = g + u * G;
                        const float w = k < cnt ? s.w[k] : 0.0f;
                        acc.x = fmaf(w, pv[u].x, acc.x); acc.y = fmaf(w, pv[u].y, acc.y);
                        acc.z = fmaf(w, pv[u].z, acc.z); acc.w = fmaf(w, pv[u].w, acc.w);
                    }
                    k_first = g + PF * G;
                }
                // eight independent 16-byte loads in flight per thread, then the FMAs in record order
                for (int k0 = k_first; k0 < cnt; k0 += 8 * G) {
                    float4 v[8];
                    float w[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int k = k0 + u * G;
                        const bool ok = k < cnt;
                        v[u] = ok ? col[(size_t)k * stride4] : make_float4(0.f, 0.f, 0.f, 0.f);
                        w[u] = ok ? s.w[k] : 0.0f;
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        acc.x = fmaf(w[u], v[u].x, acc.x); acc.y = fmaf(w[u], v[u].y, acc.y);
                        acc.z = fmaf(w[u], v[u].z, acc.z); acc.w = fmaf(w[u], v[u].w, acc.w);
                    }
                }
            }
        }
        if (dbg) phase_mark(*dbg, 14);
        __syncthreads();
        float4 *part = reinterpret_cast<float4 *>(s.tile);      // [G][nc] scratch (the tile is free at merge time)
        if (g < G) part[g * nc + c4] = acc;
        __syncthreads();
        if (t < nc) {
            float4 a = part[t];
            for (int gg = 1; gg < G; ++gg) {
                const float4 o = part[gg * nc + t];
                a.x += o.x; a.y += o.y; a.z += o.z; a.w += o.w;
            }
            s.v[4 * t] = a.x; s.v[4 * t + 1] = a.y; s.v[4 * t + 2] = a.z; s.v[4 * t + 3] = a.w;
        }
        __syncthreads();
        return;
    }
    float vacc[MAXV];
#pragma unroll
    for (int j = 0; j < MAXV; ++j) vacc[j] = 0.0f;
    for (int base = 0; base < n; base += TB) {
        const int b = base + t;
        float sc = 0.0f, sv = 0.0f;
        if (b < n) {
            const float m = __ldcg(recs + b * stride + REC_M);
            sc = (m == CUDART_INF_F) ? 0.0f : expf(-(m - M));
            sv = __ldcg(recs + b * stride + REC_S) * sc;
        }
        __syncthreads();
        s.w[t] = sc;
        S += block_sum<TB>(sv, s.red);          // contains the barriers that publish s.w
        const int cnt = (n - base) < TB ? (n - base) : TB;
#pragma unroll
        for (int j = 0; j < MAXV; ++j) {
            const int p = t + j * TB;
            if (p < P) {
                float acc = vacc[j];
                const float *col = recs + (size_t)base * stride + REC_V + p;
#pragma unroll 4
                for (int k = 0; k < cnt; ++k) acc = fmaf(s.w[k], __ldcg(col + k * stride), acc);
                vacc[j] = acc;
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < MAXV; ++j) { const int p = t + j * TB; if (p < P) s.v[p] = vacc[j]; }
    __syncthreads();
}

// bitonic sort of n (power of two) u64 keys in shared memory, ascending
template <int NT>
static __device__ void bitonic_sort_u64(unsigned long long *keys, int n) {
    constexpr int TB = NT;
    for (int size = 2; size <= n; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = threadIdx.x; i < n / 2; i += TB) {
                const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
                const bool up = ((lo & size) == 0);
                const unsigned long long a = keys[lo], b = keys[hi];
                if ((a > b) == up) { keys[lo] = b; keys[hi] = a; }
            }
            __syncthreads();
        }
}

// Last block of a robot: merge the block records into this shard's record (+ local top-k for CEM).
template <int METHOD, int NT>
__device__ void shard_reduce(const KParams &kp, int r, const Smem &s, float &M, int &IDX, float &S) {
    constexpr int TB = NT;
    const int t = threadIdx.x;
    merge_records<METHOD, NT>(kp.blk_recs + (size_t)r * kp.nrec * kp.blk_rec_stride, kp.nrec, kp.blk_rec_stride, kp.P, s, M, IDX, S, &kp);
    float *rec = kp.shard_send + (size_t)r * kp.shard_rec_stride;
    if (t == 0) { rec[REC_M] = M; rec[REC_IDX] = __int_as_float(IDX); rec[REC_S] = S; rec[REC_PAD] = 0.0f; }
    if (METHOD != RANDOM_SAMPLING)
        for (int p = t; p < kp.P; p += TB) rec[REC_V + p] = s.v[p];
    if (METHOD == CEM_MPPI) {
        unsigned long long *keys = reinterpret_cast<unsigned long long *>(s.tile);
        unsigned int *hist = reinterpret_cast<unsigned int *>(s.tile + 2 * SCRATCH_KEYS);
        unsigned int *misc = hist + 256;
        const int k = kp.elite_k < kp.n_local ? kp.elite_k : kp.n_local;
        block_topk_hist<NT>(kp.costs + (size_t)r * kp.n_local, kp.n_local, kp.sample_base, k, kp.cem_hist + (size_t)r * CEM_BINS,
                            keys, hist, misc);
        for (int j = t; j < kp.elite_k; j += TB) {
            const unsigned long long key = j < k ? keys[j] : ~0ull;
            rec[REC_V + kp.P + 2 * j] = __uint_as_float((unsigned int)(key >> 32));
            rec[REC_V + kp.P + 2 * j + 1] = __uint_as_float((unsigned int)(key & 0xffffffffu));
        }
        __syncthreads();
    }
}

// Merge `world` shard records, apply the control law update, GRF/predicted-state tail, CEM sigma refit; write `out`.
//   random_sampling S:690-692 | mppi S:823-836 | cem_mppi S:969-982,1075-1081 (A:1115-1121)
template <int PARAM, bool ADAPTIVE, int METHOD, int NT>
__device__ void finish(const KParams &kp, int r, const Smem &s, const float *recs, int world, size_t rec_stride,
                       bool premerged = false, float M = 0.0f, int IDX = 0, float S = 0.0f) {
    constexpr int TB = NT;
    constexpr int TS = NT + 1;
    const int t = threadIdx.x;
    // premerged: the caller (fused single-shard tick) has just produced (M, IDX, S, s.v) in shard_reduce
    if (!premerged) merge_records<METHOD, NT>(recs, world, rec_stride, kp.P, s, M, IDX, S);
    float *out = kp.out + (size_t)r * kp.out_stride;
    // the seed is read from the staged packet (IN_SEED), never from kernel-parameter space: a CUDA-graph-captured launch
    // freezes its parameters at capture time, the packet is uploaded every tick
    const uint32_t seed = __float_as_uint(s.in[IN_SEED]), robot = kp.robot_base + (uint32_t)r, tick = __float_as_uint(s.in[IN_TICK]);
    const float *s_best = s.in + kp.off_best;
    const float *s_sigma = s.in + kp.off_sigma;
    // ---- new best into s.v ----
    if (METHOD == RANDOM_SAMPLING) {                       // best = params[best_index], S:692
        if (kp.noise_mode == NOISE_INJECTED) {
            const float *row = kp.noise + ((size_t)r * kp.N + IDX) * kp.P;
            for (int p = t; p < kp.P; p += TB) s.v[p] = s_best[p] + __ldg(row + p);
        } else {
            for (int q = t; q < kp.P / 4; q += TB) {
                const float4 e = noise_quad<METHOD>(kp, seed, robot, tick, IDX, q, s_sigma);
                s.v[4 * q] = s_best[4 * q] + e.x; s.v[4 * q + 1] = s_best[4 * q + 1] + e.y;
                s.v[4 * q + 2] = s_best[4 * q + 2] + e.z; s.v[4 * q + 3] = s_best[4 * q + 3] + e.w;
            }
        }
    } else {                                               // best += sum_i w_i eps_i, S:828-836
        const float inv = 1.0f / S;
        for (int p = t; p < kp.P; p += TB) s.v[p] = fmaf(s.v[p], inv, s_best[p]);
    }
    __syncthreads();
    phase_mark(kp, 10);
    // where the carried controller state goes: the device-resident record (stateful controller) or the resident packet
    float *carry = kp.ctrl_state ? kp.ctrl_state + (size_t)r * kp.ctrl_stride
                                 : (kp.carry_best ? const_cast<float *>(kp.in) + (size_t)r * kp.in_stride + kp.off_best : nullptr);
    for (int p = t; p < kp.P; p += TB) {
        out[OUT_BEST + p] = s.v[p];
        if (carry) carry[p] = s.v[p];
    }
    // The serial read-out (GRFs at step 0, predicted state, scalars, carried key / contact) is ~300 dependent instructions of
    // ONE thread.  For CEM it runs on the LAST thread while the low threads regenerate the elite noise rows of the first
    // refit chunk (k = 10, or 200 rows on the 256-thread tail kernel: that thread has no row), otherwise right here.
    auto readout = [&]() {
        tail_grf_pred<PARAM>(kp, s.in, s.yaw, s.v, out);
        out[OUT_BESTCOST] = M;
        out[OUT_BESTIDX] = __int_as_float(IDX);
        float bf;
        if (ADAPTIVE) bf = sample_step_freq<METHOD>(kp, s.in, seed, robot, r, IDX);   // A:705,861,1022
        else bf = (METHOD == CEM_MPPI) ? 1.65f : 1.4f;                               // S:786,930,1083
        out[OUT_BESTFREQ] = bf;
        if (METHOD != CEM_MPPI) out[OUT_NELITE] = __int_as_float(0);
        // device-resident multi-tick runs: advance the Philox counter like with_newkey() does on the host (S:498-501)
        if (kp.ctrl_state) {
            carry[2 * kp.P] = __uint_as_float(tick + 1u);
            unsigned int cur = 0u;                              // I:123: previous_contact_mpc = current_contact
            for (int l = 0; l < 4; ++l) cur |= (s.in[IN_CONTACT + l * kp.H] != 0.0f ? 1u : 0u) << l;
            carry[2 * kp.P + 1] = __uint_as_float(cur);
        } else if (kp.carry_best) {
            const_cast<float *>(kp.in)[(size_t)r * kp.in_stride + IN_TICK] = __uint_as_float(tick + 1u);
        }
        if (kp.phase_ns != nullptr && blockIdx.y == 0) { unsigned long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns)); kp.phase_ns[6] = ns; }
    };
    const bool late_readout = (METHOD == CEM_MPPI) && ((kp.elite_k < kp.N ? kp.elite_k : kp.N) < NT);
    if (t == NT - 1 && !late_readout) readout();
    if (METHOD != CEM_MPPI) return;
    // ---- CEM elite refit ----
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(s.tile);
    const int K = kp.elite_k;
    const int total = world * K;
    int npow = 2;
    while (npow < total) npow <<= 1;
    __syncthreads();
    for (int j = t; j < npow; j += TB) {
        unsigned long long key = ~0ull;
        if (j < total) {
            const int w = j / K, e = j - w * K;
            const float *rec = recs + (size_t)w * rec_stride;
            const unsigned int cb = __float_as_uint(__ldcg(rec + REC_V + kp.P + 2 * e));
            const unsigned int ib = __float_as_uint(__ldcg(rec + REC_V + kp.P + 2 * e + 1));
            key = ((unsigned long long)cb << 32) | ib;
        }
        keys[j] = key;
    }
    __syncthreads();
    if (world > 1) bitonic_sort_u64<NT>(keys, npow);       // one shard: its record already holds the keys in ascending order
    const int k = K < kp.N ? K : kp.N;                     // argsort(costs)[:k]
    int *out_elite = reinterpret_cast<int *>(out + kp.off_out_elite);
    for (int j = t; j < K; j += TB) {
        const int id = j < k ? (int)(unsigned int)(keys[j] & 0xffffffffu) : -1;
        out_elite[j] = id;
        if (j < MAX_ELITE) s.idx[j] = id;
    }
    if (t == 0) out[OUT_NELITE] = __int_as_float(k);
    __syncthreads();                                       // the tile scratch (keys) is reused for the elite rows below
    phase_mark(kp, 7);
    // Per-knot mean / M2 over the elites, chunked by TB rows with an exact two-pass inside a chunk and Chan's
    // pairwise update between chunks (same result as the two-pass diag(cov) of S:1077 to rounding).
    float n_acc = 0.0f;
    float mean_acc[MAXV], m2_acc[MAXV];
#pragma unroll
    for (int j = 0; j < MAXV; ++j) { mean_acc[j] = 0.0f; m2_acc[j] = 0.0f; }
    for (int base = 0; base < k; base += TB) {
        const int cnt = (k - base) < TB ? (k - base) : TB;
        if (base > 0) __syncthreads();                         // (the first chunk follows the barrier above)
        if (base == 0 && late_readout && t == NT - 1) readout();
        const int my_idx = (t < cnt) ? s.idx[base + t] : -1;
        if (t < cnt) {
            float *tcol = s.tile + t;
            if (kp.noise_mode == NOISE_INJECTED) {
                const float *row = kp.noise + ((size_t)r * kp.N + my_idx) * kp.P;
                for (int p = 0; p < kp.P; ++p) tcol[p * TS] = __ldg(row + p);
            } else {
                const NoiseRow nr = noise_row<METHOD>(kp, my_idx);
                const int nq = kp.P >> 2, noct = (nq + 1) >> 1;
#pragma unroll 2
                for (int c = 0; c < noct; ++c) {                    // independent Philox chains: unrolled for ILP (4 warps alone on the SM)
                    float4 lo, hi;
                    noise_oct<METHOD>(nr, seed, robot, tick, c, s_sigma, lo, hi);
                    float *tp = tcol + 8 * c * TS;
                    tp[0] = lo.x; tp[TS] = lo.y; tp[2 * TS] = lo.z; tp[3 * TS] = lo.w;
                    if (2 * c + 1 < nq) { tp[4 * TS] = hi.x; tp[5 * TS] = hi.y; tp[6 * TS] = hi.z; tp[7 * TS] = hi.w; }
                }
            }
        }
        __syncthreads();
        if (kp.defer_tail && base == 0) phase_mark(kp, 2);
#pragma unroll
        for (int j = 0; j < MAXV; ++j) {
            const int p = t + j * TB;
            if (p < kp.P) {
                const float *rowp = s.tile + p * TS;
                float sum = 0.0f;
                for (int e = 0; e < cnt; ++e) sum += rowp[e];
                const float mean_c = sum / (float)cnt;
                float m2_c = 0.0f;
                for (int e = 0; e < cnt; ++e) { const float d = rowp[e] - mean_c; m2_c = fmaf(d, d, m2_c); }
                if (base == 0) { mean_acc[j] = mean_c; m2_acc[j] = m2_c; }
                else {
                    const float na = n_acc, nb = (float)cnt, nt = na + nb;
                    const float delta = mean_c - mean_acc[j];
                    mean_acc[j] = mean_acc[j] + delta * (nb / nt);
                    m2_acc[j] = m2_acc[j] + m2_c + delta * delta * (na * nb / nt);
                }
            }
        }
        n_acc += (float)cnt;
        if (kp.defer_tail && base == 0) phase_mark(kp, 3);
    }
#pragma unroll
    for (int j = 0; j < MAXV; ++j) {
        const int p = t + j * TB;
        if (p < kp.P) {
            const float var = m2_acc[j] / (float)(k - 1) + 1e-8f;          // diag(cov(elite)) + 1e-8, S:1077-1078
            float sg;
            if (ADAPTIVE) { sg = var > 3.0f ? 3.0f : var; sg = sg < 0.5f ? 0.5f : sg; }      // A:1118-1121 (no sqrt)
            else { sg = sqrtf(var); sg = sg > 5.0f ? 5.0f : sg; sg = sg < 0.2f ? 0.2f : sg; } // S:1079-1081
            out[kp.off_out_sigma + p] = sg;
            // device-resident sampling iterations (I:125-148): the next iteration draws with the refitted sigma
            if (carry) carry[kp.P + p] = sg;
        }
    }
}

// Direct I/O epilogue (last block, one controller): the finished result packet goes to mapped pinned host memory as
// self-validating 8-byte units {value, seq} (two per 16-byte store; every aligned 8-byte half arrives atomically over
// PCIe -- the scheme NCCL's LL protocol uses).  The host accepts a unit when its flag equals this tick's seq, so no
// system-scope fence and no separate stamp are needed: a fence.sys after the stores was measured at 4.9 us per tick
// (`tools/probe_latency.py phases`, slot 15 -> 11), the stores alone leave the SM in < 1 us.
template <int NT>
__device__ __forceinline__ void publish_to_host(const KParams &kp, int r) {
    if (kp.host_out == nullptr) return;
    phase_mark(kp, 15);
    __syncthreads();                                   // the packet in kp.out was written by this block
    const float2 *o = reinterpret_cast<const float2 *>(kp.out + (size_t)r * kp.out_stride);
    uint4 *dst = reinterpret_cast<uint4 *>(kp.host_out);
    const int n2 = kp.out_stride >> 1;                 // out_stride is a multiple of 4 (round4)
    constexpr int ROUNDS = 4;                          // all loads of a thread before its first store (no load/store chain)
    float2 v[ROUNDS];
    for (int base = 0; base < n2; base += ROUNDS * NT) {
#pragma unroll
        for (int j = 0; j < ROUNDS; ++j) {
            const int k = base + j * NT + (int)threadIdx.x;
            if (k < n2) v[j] = o[k];
        }
#pragma unroll
        for (int j = 0; j < ROUNDS; ++j) {
            const int k = base + j * NT + (int)threadIdx.x;
            if (k < n2)
                asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst + k), "r"(__float_as_uint(v[j].x)),
                             "r"(kp.seq), "r"(__float_as_uint(v[j].y)), "r"(kp.seq)
                             : "memory");
        }
    }
}

// Publish the tile's softmax weights for the weighted knot sum: only the non-zero ones (at T = 1 most weights underflow to
// exactly 0), as (weight, sample slot) pairs in ascending slot order -- the same addends in the same
// order as a loop over all slots that skips zeros, so the sum is bit-identical.  Returns the pair count; s_tile = sum of
// the tile's weights (fixed order).  `owner`: this thread holds the weight of sample `slot` (one lane per sample).
template <int NT>
__device__ __forceinline__ int compact_weights(float w, int slot, bool owner, const Smem &s, float &s_tile) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const bool nzb = owner && (w != 0.0f);
    const unsigned int nz = __ballot_sync(0xffffffffu, nzb);
    __shared__ int s_wcnt[NT / 32];
    if (lane == 0) s_wcnt[wid] = __popc(nz);
    s_tile = block_sum<NT>(owner ? w : 0.0f, s.red);                   // its two barriers publish s_wcnt
    int off = __popc(nz & ((1u << lane) - 1u)), total = 0;
#pragma unroll
    for (int k = 0; k < NT / 32; ++k) { const int c = s_wcnt[k]; total += c; if (k < wid) off += c; }
    if (nzb) reinterpret_cast<float2 *>(s.w)[off] = make_float2(w, __int_as_float(slot));
    __syncthreads();
    return total;
}

// ------------------------------------------------------------------------------------------------------
// Fused collective of a sample-sharded controller (one exchange step per tick): instead of a separate
// all-gather launch, the last block of each shard stores its record straight into every peer GPU's mailbox over NVLink /
// NVSwitch and reads the peers' records from its own mailbox; finish() then merges the records in rank order -- every
// rank ends the same launch with the same control.
//   Wire format: self-validating 8-byte units {value, seq} (two per 16-byte store; an aligned 8-byte unit arrives whole
//   over NVLink -- the scheme of NCCL's LL protocol, and of publish_to_host).  The receiver takes each unit as soon as its
//   flag equals this tick's seq, so there is NO system-scope fence and NO separate stamp: the r1 version (record stores,
//   __threadfence_system, st.release.sys stamps, ld.acquire.sys + nanosleep polling) cost ~16 us per tick.
//   mailbox of one rank: units [2 parity][world][R][shard_rec_stride] of 8 bytes.
//   Two parities: a rank that is one tick ahead writes the other half, and cannot get two ticks ahead because it
//   needs this rank's record of the tick in between.
// Returns false if a peer's record did not arrive within kp.peer_timeout_ns (default 2 s; peer process gone): the caller
// reports it.
// ------------------------------------------------------------------------------------------------------
__host__ __device__ inline size_t peer_mbox_units(int world, int R, int stride) { return (size_t)2 * world * R * stride; }
__host__ __device__ inline size_t peer_mbox_bytes(int world, int R, int stride) { return peer_mbox_units(world, R, stride) * 8; }

template <int NT>
__device__ bool peer_exchange(const KParams &kp, int r, const float *&recs, size_t &rec_stride) {
    const int W = kp.peer_world, R = gridDim.y, stride = kp.shard_rec_stride, t = threadIdx.x;
    const unsigned int par = kp.seq & 1u, seq = kp.seq;
    const size_t slot = ((size_t)(par * W + kp.peer_rank) * R + r);          // [parity][my rank][robot]
    const float2 *mine = reinterpret_cast<const float2 *>(kp.shard_send + (size_t)r * stride);   // written by this block
    __syncthreads();
    if (kp.phase_ns != nullptr && t == 0) { unsigned long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns)); kp.phase_ns[16] = ns; }
    // 1. push the record to every rank (own mailbox included): 16-byte stores of two {value, seq} units
    const int n2 = stride >> 1;                                               // stride is a multiple of 4
    for (int e = t; e < W * n2; e += NT) {
        const int w = e / n2, k = e - w * n2;
        const float2 v = mine[k];
        uint4 *dst = reinterpret_cast<uint4 *>(reinterpret_cast<uint2 *>(kp.peer_mbox[w]) + slot * stride) + k;
        asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(__float_as_uint(v.x)), "r"(seq),
                     "r"(__float_as_uint(v.y)), "r"(seq) : "memory");
    }
    // 2. take the W records of this tick from the local mailbox, unit by unit, into the (local) receive buffer
    __shared__ int s_ok;
    if (t == 0) s_ok = 1;
    __syncthreads();
    const uint2 *local = reinterpret_cast<const uint2 *>(kp.peer_mbox[kp.peer_rank]);
    float *recv = const_cast<float *>(kp.shard_recv);                         // [world][R][stride] floats
    unsigned long long t0 = 0ull;
    for (int e = t; e < W * stride; e += NT) {
        const int w = e / stride, k = e - w * stride;
        const uint2 *src = local + ((size_t)(par * W + w) * R + r) * stride + k;
        unsigned int val, flag, spins = 0u;
        for (;;) {
            asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(val), "=r"(flag) : "l"(src) : "memory");
            if (flag == seq) break;
            if ((++spins & 1023u) == 0u) {                                    // bounded wait: a peer that never ticks
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0ull) t0 = now;
                else if (now - t0 > kp.peer_timeout_ns) { s_ok = 0; break; }
            }
        }
        recv[((size_t)w * R + r) * stride + k] = __uint_as_float(val);
    }
    __syncthreads();
    if (kp.phase_ns != nullptr && t == 0) { unsigned long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns)); kp.phase_ns[17] = ns; }
    recs = recv + (size_t)r * stride;
    rec_stride = (size_t)R * stride;
    return s_ok != 0;
}

// what the tick kernels do after shard_reduce: finish locally (one shard), exchange with the peers and finish, or stop
// (records go to a separate collective + finish_kernel).  BESTIDX = -2 flags a peer time-out to the host.
template <int PARAM, bool ADAPTIVE, int METHOD, int NT>
__device__ __forceinline__ void after_shard_reduce(const KParams &kp, int r, const Smem &s, float M, int IDX, float S) {
    if (kp.fused_finish) {
        __syncthreads();                               // the shard record (elite keys) is read back by this block only
        finish<PARAM, ADAPTIVE, METHOD, NT>(kp, r, s, kp.shard_send + (size_t)r * kp.shard_rec_stride, 1, 0, true, M, IDX, S);
        publish_to_host<NT>(kp, r);
    } else if (kp.peer_world > 1) {
        const float *recs;
        size_t rec_stride;
        const bool ok = peer_exchange<NT>(kp, r, recs, rec_stride);
        finish<PARAM, ADAPTIVE, METHOD, NT>(kp, r, s, recs, kp.peer_world, rec_stride);
        if (!ok) {
            __syncthreads();
            if (threadIdx.x == 0) (kp.out + (size_t)r * kp.out_stride)[OUT_BESTIDX] = __int_as_float(-2);
        }
        publish_to_host<NT>(kp, r);
    }
}

// ======================================================================================================
// The tick kernel.  grid = (nblk, R), block = TB.  dynamic smem = smem_floats(...) * 4 bytes.
// ======================================================================================================
//
// NH > 1 (tick_kernel_h, one controller whose samples fit one tile per block on <= 148 blocks): the block carries NH - 1
// extra groups of TB "noise helper" threads.  At 10k samples the machine is under-filled (one warp per scheduler on half
// the SMs) and the Philox / Box-Muller phase is a serial chain of P/4 quads per thread; helper group h draws quads
// h, h + NH, ... of the same sample column into the tile, then exits (barriers after that point count the non-exited
// threads only).  Same Philox counters, same tile contents: results are bit-identical to NH = 1.
template <int PARAM, bool ADAPTIVE, int METHOD, int NH>
__device__ __forceinline__ void tick_body(const KParams &kp, int &s_last) {
    extern __shared__ __align__(16) float smem_raw[];
    const Smem s(smem_raw, kp);
    const int t = NH > 1 ? (int)(threadIdx.x & (TB - 1)) : (int)threadIdx.x, blk = blockIdx.x, r = blockIdx.y;
    const int hlp = NH > 1 ? (int)(threadIdx.x / TB) : 0;
    if (blk == 0) phase_mark(kp, 0);
    // two-stage launch: let the dependent tail_kernel be scheduled now (programmatic dependent launch); its blocks wait in
    // griddepcontrol.wait until this whole grid has completed and flushed, so only the launch latency overlaps
    if (kp.defer_tail) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    stage_inputs<TB * NH>(kp, r, s);
    if (blk == 0) phase_mark(kp, 1);
    const uint32_t seed = __float_as_uint(s.in[IN_SEED]), robot = kp.robot_base + (uint32_t)r, tick = __float_as_uint(s.in[IN_TICK]);
    const float *s_sigma = s.in + kp.off_sigma;
    float *tcol = s.tile + t;
    float m_run = CUDART_INF_F, s_run = 0.0f;
    int idx_run = 0x7fffffff;
    int grp = blk * (kp.iters / kp.iters_rec), grp_tiles = 0;     // record group of the current tile, tiles folded into it
    if (hlp == 0)
        for (int p = t; p < kp.P; p += TB) s.v[p] = 0.0f;

    for (int it = 0; it < kp.iters; ++it) {
        const int tile_base = (blk * kp.iters + it) * TB;       // local (shard) index of this tile's first sample
        if (tile_base >= kp.n_local) break;                     // block-uniform
        const int local = tile_base + t;
        const bool active = local < kp.n_local;
        const int gi = kp.sample_base + local;                  // global sample index (row of the noise tensor)
        // ---- 1. eps tile: tile[p][t] = additional_random_parameters[gi][p] ----
        if (kp.noise_mode == NOISE_INJECTED) {
            const int rows = (kp.n_local - tile_base) < TB ? (kp.n_local - tile_base) : TB;
            const float *src = kp.noise + ((size_t)r * kp.N + kp.sample_base + tile_base) * kp.P;
            for (int e = threadIdx.x; e < TB * kp.P; e += TB * NH) {
                const int row = e / kp.P, p = e - row * kp.P;
                s.tile[p * TS + row] = row < rows ? __ldg(src + e) : 0.0f;
            }
            __syncthreads();
        } else {
            if (active && gi != 0) {
                const NoiseRow nr = noise_row<METHOD>(kp, gi);
                const int nq = kp.P >> 2, noct = (nq + 1) >> 1;      // one Philox block per oct (8 knots)
                float *tp = tcol + hlp * 8 * TS;
#pragma unroll 4
                for (int c = hlp; c < noct; c += NH, tp += NH * 8 * TS) {
                    float4 lo, hi;
                    noise_oct_nz<METHOD>(nr, seed, robot, tick, c, s_sigma, lo, hi);
                    tp[0] = lo.x; tp[TS] = lo.y; tp[2 * TS] = lo.z; tp[3 * TS] = lo.w;
                    if (2 * c + 1 < nq) { tp[4 * TS] = hi.x; tp[5 * TS] = hi.y; tp[6 * TS] = hi.z; tp[7 * TS] = hi.w; }
                }
            } else {                                            // padding rows and row 0 (the previous best re-evaluated)
                for (int p = hlp; p < kp.P; p += NH) tcol[p * TS] = 0.0f;
            }
            if (NH > 1) __syncthreads();                        // the helpers' quads are in the tile
        }
        if (NH > 1 && hlp > 0) return;                          // single tile per block (host: iters == 1): helpers are done
        if (blk == 0 && it == 0) phase_mark(kp, 2);
        // ---- 2. rollout + cost ----
        float cost = CUDART_INF_F;
        if (active) {
            if (ADAPTIVE) {
                const float freq = sample_step_freq<METHOD>(kp, s.in, seed, robot, r, gi);
                cost = rollout_adaptive<PARAM>(kp, s.in, s.yaw, tcol, freq);
            } else {
                cost = rollout_fixed<PARAM>(kp, s.in, s.fref, tcol);
            }
            if (kp.mode == MODE_TICK) {                          // NaN / +-Inf -> 1e6, S:686-687
                if (isnan(cost) || isinf(cost)) cost = 1000000.0f;
            }
            kp.costs[(size_t)r * kp.n_local + local] = cost;
            if (METHOD == CEM_MPPI && kp.mode == MODE_TICK)      // first radix level of the elite selection (block_topk_hist)
                cem_count_cost(kp, r, cost);
        }
        if (kp.mode == MODE_ROLLOUT) { __syncthreads(); continue; }
        if (blk == 0 && it == 0) phase_mark(kp, 3);
        // ---- 3. tile reduction folded into the block's running record ----
        // (argmin | softmax weights, their sum and compaction offsets | compacted pairs): three block barriers; the same
        // values in the same order as block_argmin + compact_weights, which the other kernels use (seven barriers)
        const int lane = t & 31, wid = t >> 5;
        float cmin = cost;
        int imin = active ? gi : 0x7fffffff;
        warp_argmin(cmin, imin);
        if (lane == 0) { s.red[wid] = cmin; s.red[TB / 32 + wid] = __int_as_float(imin); }
        __syncthreads();
        cmin = s.red[0]; imin = __float_as_int(s.red[TB / 32]);
#pragma unroll
        for (int k = 1; k < TB / 32; ++k) {
            const float oc = s.red[k];
            const int oi = __float_as_int(s.red[TB / 32 + k]);
            if (lex_less(oc, oi, cmin, imin)) { cmin = oc; imin = oi; }
        }
        if (METHOD != RANDOM_SAMPLING) {
            const float m_new = fminf(m_run, cmin);              // cmin is finite: the tile has >= 1 active sample
            const float scale_old = (m_run == CUDART_INF_F) ? 0.0f : expf(-(m_run - m_new));
            const float w = active ? expf(-(cost - m_new)) : 0.0f;   // exp(-(c - beta)/T), T = 1, S:828-830
            // only the non-zero weights are published, as (weight, sample slot) pairs in ascending slot order
            // (compact_weights); their sum in block_sum's fixed order
            const bool nzb = w != 0.0f;
            const unsigned int nz = __ballot_sync(0xffffffffu, nzb);
            float wsum = w;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
            float *s_sum = s.red + 2 * (TB / 32);
            int *s_cnt = reinterpret_cast<int *>(s.red + 3 * (TB / 32));
            if (lane == 0) { s_sum[wid] = wsum; s_cnt[wid] = __popc(nz); }
            __syncthreads();
            float s_tile = s_sum[0];
            int off = __popc(nz & ((1u << lane) - 1u)), nnz = 0;
#pragma unroll
            for (int k = 0; k < TB / 32; ++k) {
                const int c = s_cnt[k];
                if (k > 0) s_tile += s_sum[k];
                nnz += c;
                if (k < wid) off += c;
            }
            if (nzb) reinterpret_cast<float2 *>(s.w)[off] = make_float2(w, __int_as_float(t));
            __syncthreads();
            s_run = fmaf(s_run, scale_old, s_tile);
            const float2 *wi = reinterpret_cast<const float2 *>(s.w);
            for (int p = t; p < kp.P; p += TB) {
                const float *rowp = s.tile + p * TS;
                float acc = 0.0f;
#pragma unroll 4
                for (int j = 0; j < nnz; ++j) {                  // samples whose weight underflowed to 0 add exactly nothing
                    const float2 we = wi[j];
                    acc = fmaf(we.x, rowp[__float_as_int(we.y)], acc);
                }
                s.v[p] = fmaf(s.v[p], scale_old, acc);
            }
        }
        if (lex_less(cmin, imin, m_run, idx_run)) { m_run = cmin; idx_run = imin; }   // running (min, argmin), S:690
        // one record per group of kp.iters_rec tiles (or at the shard's last tile): a block that walks several groups
        // (batched launches amortise their staging that way) leaves exactly the records one-group blocks would
        // (kp.iters is a multiple of kp.iters_rec, so a block's first tile opens a group: counters, no division)
        if (++grp_tiles == kp.iters_rec || tile_base + TB >= kp.n_local) {
            float *rec = kp.blk_recs + ((size_t)r * kp.nrec + grp) * kp.blk_rec_stride;
            ++grp; grp_tiles = 0;
            if (t == 0) { rec[REC_M] = m_run; rec[REC_IDX] = __int_as_float(idx_run); rec[REC_S] = s_run; rec[REC_PAD] = 0.0f; }
            if (METHOD != RANDOM_SAMPLING)
                for (int p = t; p < kp.P; p += TB) { rec[REC_V + p] = s.v[p]; s.v[p] = 0.0f; }
            m_run = CUDART_INF_F; s_run = 0.0f; idx_run = 0x7fffffff;
        }
        __syncthreads();
    }
    if (kp.mode == MODE_ROLLOUT) return;
    if (blk == 0) phase_mark(kp, 4);
    if (kp.defer_tail) return;                          // large launches: tail_kernel (qpmpc_tail.cuh) finishes the tick
    // release / acquire around the ticket by ONE thread between two block barriers (the shape of cooperative groups'
    // grid sync): the barrier orders the block's record stores before thread 0's fence, which is cumulative; the last
    // block's readers are ordered after thread 0's second fence by the second barrier.  128 membars per block -> 1.
    __syncthreads();
    if (t == 0) {
        __threadfence();
        const unsigned int ticket = atomicAdd(kp.counters + r, 1u);
        s_last = (ticket == (unsigned int)(kp.nblk - 1));
        if (s_last) __threadfence();
    }
    __syncthreads();
    if (blk == 0) phase_mark(kp, 5);
    if (!s_last) return;
    phase_mark(kp, 8);
    float M, S; int IDX;
    shard_reduce<METHOD, TB>(kp, r, s, M, IDX, S);
    phase_mark(kp, 9);
    after_shard_reduce<PARAM, ADAPTIVE, METHOD, TB>(kp, r, s, M, IDX, S);
    phase_mark(kp, 11);
    if (t == 0) kp.counters[r] = 0u;
}

template <int PARAM, bool ADAPTIVE, int METHOD>
__global__ void __launch_bounds__(TB, 4) tick_kernel(const __grid_constant__ KParams kp) {
    __shared__ int s_last;
    tick_body<PARAM, ADAPTIVE, METHOD, 1>(kp, s_last);
}

constexpr int NOISE_HELPERS = 4;            // thread groups per block of tick_kernel_h (one computes, three only draw noise)
template <int PARAM, bool ADAPTIVE, int METHOD>
__global__ void __launch_bounds__(TB * NOISE_HELPERS, 1) tick_kernel_h(const __grid_constant__ KParams kp) {
    __shared__ int s_last;
    tick_body<PARAM, ADAPTIVE, METHOD, NOISE_HELPERS>(kp, s_last);
}

// Step 6 alone: merges `world` gathered shard records (sample-sharded controllers).  grid = (1, R).
template <int PARAM, bool ADAPTIVE, int METHOD>
__global__ void __launch_bounds__(TB, 4) finish_kernel(const __grid_constant__ KParams kp) {
    extern __shared__ __align__(16) float smem_raw[];
    const Smem s(smem_raw, kp);
    const int r = blockIdx.y;
    stage_inputs<TB>(kp, r, s);
    finish<PARAM, ADAPTIVE, METHOD, TB>(kp, r, s, kp.shard_recv + (size_t)r * kp.shard_rec_stride, kp.world,
                                    (size_t)gridDim.y * kp.shard_rec_stride);
}

}  // namespace qp
