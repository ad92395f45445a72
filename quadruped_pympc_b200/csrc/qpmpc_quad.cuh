// qpmpc_quad.cuh -- low-latency variant of the tick kernel: FOUR LANES PER SAMPLE, one leg each.
//
// A single 10k-sample tick fills only 313 of the 592 warp schedulers of a B200 with one warp each when every thread
// owns a whole sample, and each warp then walks an ~11k-instruction dependent chain alone (ncu: stall_wait dominant,
// issue slots 25 % busy).  Here the four lanes of a quad share one sample: lane l draws a quarter of the sample's noise
// row, evaluates the spline / clamp / moment arm of leg l, the six force and torque sums are combined with two
// butterfly shuffles, and the (cheap, 12-state) body update is carried redundantly by the four lanes so that they
// stay bit-identical.  4x the warps, ~2.4x shorter chain per lane.  The thread-per-sample kernel (qpmpc_tick.cuh)
// remains the throughput path for large batches (it executes ~1.8x fewer instructions per sample); qpmpc_create picks
// by the number of samples in flight.
//
// Same tile idea as the sample-per-thread kernel, with 32 samples per tile: tile[p][slot], stride 33.
#pragma once
#include "qpmpc_tick.cuh"

namespace qp {

constexpr int QTB = 256;         // threads per block: fewer, larger blocks keep the last-block merge short
constexpr int QS = QTB / 4;     // samples per tile
constexpr int QTS = QS + 1;     // tile row stride (floats): bank = (p + slot) % 32

__device__ __forceinline__ float quad_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}

// knot j of this lane's leg: best + eps  (S:680 | S:814 | S:960)
#define QP_QKNOT(j) (s_bleg[(j)] + tleg[(j) * QTS])

// one leg's clamped force and its moment about the CoM; zero contribution when the leg is in swing
__device__ __forceinline__ void leg_wrench(const KParams &kp, const Body &b, const float *foot, float fx, float fy, float fz,
                                           float fref, float &Fx, float &Fy, float &Fz, float &tx, float &ty, float &tz) {
    constexpr float third = 1.0f / 3.0f;
    fz = fref + fz;                               // S:382-385
    fx = fx * third; fy = fy * third;             // S:388-402 with contact == 1
    clamp_leg(kp, fx, fy, fz);                    // S:405-409
    const float rx = foot[0] - b.px, ry = foot[1] - b.py, rz = foot[2] - b.pz;
    Fx = fx; Fy = fy; Fz = fz;
    tx = ry * fz - rz * fy;
    ty = rz * fx - rx * fz;
    tz = rx * fy - ry * fx;
}

// Fixed-gait rollout (S:316-496), lane `leg` of the quad.  tleg: &tile[leg * P_leg][slot].
template <int PARAM>
__device__ __forceinline__ float rollout_fixed_quad(const KParams &kp, const float *s_in, const float *s_fref, const float *tleg,
                                                    int leg) {
    const float *s_yaw = s_fref + MAX_H;
    const float *s_bleg = s_in + kp.off_best + leg * kp.P_leg;
    const float *s_contact = s_in + IN_CONTACT + leg * kp.H;
    const float *ref = s_in + IN_REF;
    const float *foot = s_in + IN_STATE + 12 + 3 * leg;
    Body b = load_body(s_in + IN_STATE);
    float cost = 0.0f;
    const int H = kp.H;
    float ka[3], kb[3], kc[3], kd[3];
    int cur = -1;
#pragma unroll 1
    for (int n = 0; n < H; ++n) {
        if (PARAM != ZERO_ORDER) {
            const int idx = kp.sp_index[n];
            if (idx != cur) {
                cur = idx;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    if (PARAM == CUBIC) {                     // S:219-255
                        const int s0 = 10 * idx + 4 * a;
                        ka[a] = QP_QKNOT(s0); kb[a] = QP_QKNOT(s0 + 1); kc[a] = QP_QKNOT(s0 + 2); kd[a] = QP_QKNOT(s0 + 3);
                    } else {                                  // S:196-199
                        const int s0 = idx + (kp.ns + 1) * a;
                        ka[a] = QP_QKNOT(s0);
                        kb[a] = QP_QKNOT(s0 + 1);
                    }
                }
            }
        }
        float Fx = 0.f, Fy = 0.f, Fz = 0.f, tx = 0.f, ty = 0.f, tz = 0.f;
        if (s_contact[n] != 0.0f) {
            const float w0 = kp.basis[n][0], w1 = kp.basis[n][1], w2 = kp.basis[n][2], w3 = kp.basis[n][3];
            float fx, fy, fz;
            if (PARAM == CUBIC) {
                fx = w0 * ka[0] + w1 * kb[0] + w2 * kc[0] + w3 * kd[0];
                fy = w0 * ka[1] + w1 * kb[1] + w2 * kc[1] + w3 * kd[1];
                fz = w0 * ka[2] + w1 * kb[2] + w2 * kc[2] + w3 * kd[2];
            } else if (PARAM == LINEAR) {
                fx = w0 * ka[0] + w1 * kb[0];
                fy = w0 * ka[1] + w1 * kb[1];
                fz = w0 * ka[2] + w1 * kb[2];
            } else {                                          // S:259-268
                fx = QP_QKNOT(n); fy = QP_QKNOT(n + H); fz = QP_QKNOT(n + 2 * H);
            }
            leg_wrench(kp, b, foot, fx, fy, fz, s_fref[n], Fx, Fy, Fz, tx, ty, tz);
        }
        Fx = quad_sum(Fx); Fy = quad_sum(Fy); Fz = quad_sum(Fz);
        tx = quad_sum(tx); ty = quad_sum(ty); tz = quad_sum(tz);
        cost += srbd_step(kp, b, Fx, Fy, Fz, tx, ty, tz, kp.dt[n], ref, s_yaw);   // S:446-451,491
    }
    return cost;
}

// spline with this lane's own clock (adaptive rollouts): S:181-268 with step = n_leg, A:353-374
template <int PARAM>
__device__ __forceinline__ void spline_dyn_quad(const KParams &kp, const float *s_bleg, const float *tleg, int step,
                                                float inv_denom, float &fx, float &fy, float &fz) {
    const int PL = kp.P_leg;
    if (PARAM == ZERO_ORDER) {
        int j0 = step, j1 = step + kp.H, j2 = step + 2 * kp.H;    // int16(step); negative indices wrap like jnp
        if (j0 < 0) j0 += PL;
        if (j1 < 0) j1 += PL;
        if (j2 < 0) j2 += PL;
        fx = QP_QKNOT(j0); fy = QP_QKNOT(j1); fz = QP_QKNOT(j2);
        return;
    }
    const float stepf = (float)step;
    int idx = 0;
    for (int i = 1; i <= kp.ns; ++i) idx = (stepf >= kp.bounds[i]) ? i : idx;      // S:187-189
    idx = idx < kp.ns ? idx : kp.ns - 1;
    const float q = fmaf(stepf, inv_denom, -(float)idx);                           // S:191-194
    if (PARAM == LINEAR) {
        const int sh = kp.ns + 1;
        fx = (1.0f - q) * QP_QKNOT(idx) + q * QP_QKNOT(idx + 1);
        fy = (1.0f - q) * QP_QKNOT(idx + sh) + q * QP_QKNOT(idx + sh + 1);
        fz = (1.0f - q) * QP_QKNOT(idx + 2 * sh) + q * QP_QKNOT(idx + 2 * sh + 1);
        return;
    }
    const float q2 = q * q, q3 = q2 * q;
    const float a = 2.0f * q3 - 3.0f * q2 + 1.0f, bb = q3 - 2.0f * q2 + q, c = -2.0f * q3 + 3.0f * q2, d = q3 - q2;
    const int s0 = 10 * idx;
    float f[3];
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        const float p0 = QP_QKNOT(s0 + 4 * ax), p1 = QP_QKNOT(s0 + 4 * ax + 1), p2 = QP_QKNOT(s0 + 4 * ax + 2),
                    p3 = QP_QKNOT(s0 + 4 * ax + 3);
        const float phi = 0.5f * ((p2 - p1) + (p1 - p0)), phin = 0.5f * ((p3 - p2) + (p2 - p1));
        f[ax] = a * p1 + bb * phi + c * p2 + d * phin;
    }
    fx = f[0]; fy = f[1]; fz = f[2];
}

// Gait-adaptive rollout (A:326-501), lane `leg`: own phase clock (G:68-89), stance counter and spline clock.
template <int PARAM>
__device__ __forceinline__ float rollout_adaptive_quad(const KParams &kp, const float *s_in, const float *s_yaw,
                                                       const float *tleg, int leg, float freq) {
    const float *s_bleg = s_in + kp.off_best + leg * kp.P_leg;
    const float *ref = s_in + IN_REF;
    const float *foot = s_in + IN_STATE + 12 + 3 * leg;
    const int H = kp.H;
    const float adv = kp.mpc_dt * freq;                       // G:79
    float t = s_in[IN_PHASE + leg];
    float cnt = 0.0f;
    for (int n = 0; n < H; ++n) {                             // horizon_leg = sum_n contact + 1, A:345-348
        t = t >= 1.0f ? 0.0f : t;                             // G:73-76
        t = t + adv;                                          // G:79-82
        cnt += (t < kp.duty) ? 1.0f : 0.0f;                   // G:84-87
    }
    const float inv_denom = 1.0f / ((cnt + 1.0f) * kp.inv_ns);
    t = s_in[IN_PHASE + leg];
    int nleg = -1;                                            // A:339
    Body b = load_body(s_in + IN_STATE);
    float cost = 0.0f;
#pragma unroll 1
    for (int n = 0; n < H; ++n) {
        t = t >= 1.0f ? 0.0f : t;
        t = t + adv;
        const float c = (t < kp.duty) ? 1.0f : 0.0f;
        nleg += (int)c;                                       // A:353-356
        const float nst = quad_sum(c);                        // small integers: exact in any order
        const float fref = kp.mg / nst;                       // inf when no leg is in stance
        float Fx = 0.f, Fy = 0.f, Fz = 0.f, tx = 0.f, ty = 0.f, tz = 0.f;
        if (c != 0.0f) {
            float fx, fy, fz;
            spline_dyn_quad<PARAM>(kp, s_bleg, tleg, nleg, inv_denom, fx, fy, fz);
            leg_wrench(kp, b, foot, fx, fy, fz, fref, Fx, Fy, Fz, tx, ty, tz);
        }
        Fx = quad_sum(Fx); Fy = quad_sum(Fy); Fz = quad_sum(Fz);
        tx = quad_sum(tx); ty = quad_sum(ty); tz = quad_sum(tz);
        cost += srbd_step(kp, b, Fx, Fy, Fz, tx, ty, tz, kp.dt[n], ref, s_yaw);
    }
    const float df = freq - 1.3f;
    cost += df * 100.0f * df;                                 // A:500
    return cost;
}

// ======================================================================================================
// grid = (nblk, R), block = QTB threads = QS samples per tile.  Same records / last-block merge / finish as tick_kernel.
// ======================================================================================================
template <int PARAM, bool ADAPTIVE, int METHOD>
__global__ void __launch_bounds__(QTB, 2) tick_kernel_quad(const __grid_constant__ KParams kp) {
    extern __shared__ __align__(16) float smem_raw[];
    const Smem s(smem_raw, kp, QTB);
    const int t = threadIdx.x, blk = blockIdx.x, r = blockIdx.y;
    const int leg = t & 3, slot = t >> 2;
    __shared__ int s_last;
    if (blk == 0) phase_mark(kp, 0);
    if (kp.defer_tail) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // see tick_body
    stage_inputs<QTB>(kp, r, s);
    if (blk == 0) phase_mark(kp, 1);
    const uint32_t seed = __float_as_uint(s.in[IN_SEED]), robot = kp.robot_base + (uint32_t)r, tick = __float_as_uint(s.in[IN_TICK]);
    const float *s_sigma = s.in + kp.off_sigma;
    float *tcol = s.tile + slot;
    const float *tleg = tcol + leg * kp.P_leg * QTS;
    float m_run = CUDART_INF_F, s_run = 0.0f;
    int idx_run = 0x7fffffff;
    for (int p = t; p < kp.P; p += QTB) s.v[p] = 0.0f;
    const int nq = kp.P >> 2, noct = (nq + 1) >> 1;
    const int c_lo = (leg * noct) >> 2, c_hi = ((leg + 1) * noct) >> 2;   // this lane's share of the sample's noise octs (Philox blocks)

    int grp = blk * (kp.iters / kp.iters_rec), grp_tiles = 0;     // record group of the current tile, tiles folded into it
    for (int it = 0; it < kp.iters; ++it) {
        const int tile_base = (blk * kp.iters + it) * QS;
        if (tile_base >= kp.n_local) break;                     // block-uniform
        const int local = tile_base + slot;
        const bool active = local < kp.n_local;
        const int gi = kp.sample_base + local;
        // ---- 1. eps tile: tile[p][slot] = additional_random_parameters[gi][p] ----
        if (kp.noise_mode == NOISE_INJECTED) {
            const int rows = (kp.n_local - tile_base) < QS ? (kp.n_local - tile_base) : QS;
            const float *src = kp.noise + ((size_t)r * kp.N + kp.sample_base + tile_base) * kp.P;
            for (int e = t; e < QS * kp.P; e += QTB) {
                const int row = e / kp.P, p = e - row * kp.P;
                s.tile[p * QTS + row] = row < rows ? __ldg(src + e) : 0.0f;
            }
            __syncthreads();
        } else {
            NoiseRow nr = noise_row<METHOD>(kp, gi);
            nr.zero = nr.zero || !active;
#pragma unroll 2
            for (int c = c_lo; c < c_hi; ++c) {
                float4 lo, hi;
                noise_oct<METHOD>(nr, seed, robot, tick, c, s_sigma, lo, hi);
                float *tp = tcol + 8 * c * QTS;
                tp[0] = lo.x; tp[QTS] = lo.y; tp[2 * QTS] = lo.z; tp[3 * QTS] = lo.w;
                if (2 * c + 1 < nq) { tp[4 * QTS] = hi.x; tp[5 * QTS] = hi.y; tp[6 * QTS] = hi.z; tp[7 * QTS] = hi.w; }
            }
            __syncwarp();                                        // the four lanes of a quad live in one warp
        }
        if (blk == 0 && it == 0) phase_mark(kp, 2);
        // ---- 2. rollout + cost (all four lanes end with the same bits) ----
        float cost = CUDART_INF_F;
        if (ADAPTIVE) {
            // inactive quads run along (shuffles need the whole warp) on zero noise; their cost is discarded
            const float freq = active ? sample_step_freq<METHOD>(kp, s.in, seed, robot, r, gi) : 1.3f;
            cost = rollout_adaptive_quad<PARAM>(kp, s.in, s.yaw, tleg, leg, freq);
        } else {
            cost = rollout_fixed_quad<PARAM>(kp, s.in, s.fref, tleg, leg);
        }
        if (!active) cost = CUDART_INF_F;
        else {
            if (kp.mode == MODE_TICK) {                          // NaN / +-Inf -> 1e6, S:686-687
                if (isnan(cost) || isinf(cost)) cost = 1000000.0f;
            }
            if (leg == 0) {
                kp.costs[(size_t)r * kp.n_local + local] = cost;
                if (METHOD == CEM_MPPI && kp.mode == MODE_TICK)  // first radix level of the elite selection (block_topk_hist)
                    cem_count_cost(kp, r, cost);
            }
        }
        if (blk == 0 && it == 0) phase_mark(kp, 3);
        if (kp.mode == MODE_ROLLOUT) { __syncthreads(); continue; }
        // ---- 3. tile reduction folded into the block's running record ----
        float cmin = cost;
        int imin = active ? gi : 0x7fffffff;
        block_argmin<QTB>(cmin, imin, s.red);
        if (METHOD != RANDOM_SAMPLING) {
            const float m_new = fminf(m_run, cmin);
            const float scale_old = (m_run == CUDART_INF_F) ? 0.0f : expf(-(m_run - m_new));
            const float w = active ? expf(-(cost - m_new)) : 0.0f;   // exp(-(c - beta)/T), T = 1, S:828-830
            float s_tile;
            const int nnz = compact_weights<QTB>(w, slot, leg == 0, s, s_tile);   // its barriers also publish the tile
            s_run = fmaf(s_run, scale_old, s_tile);
            const float2 *wi = reinterpret_cast<const float2 *>(s.w);
            for (int p = t; p < kp.P; p += QTB) {
                const float *rowp = s.tile + p * QTS;
                float acc = 0.0f;
#pragma unroll 4
                for (int j = 0; j < nnz; ++j) {
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
        if (++grp_tiles == kp.iters_rec || (blk * kp.iters + it + 1) * QS >= kp.n_local) {
            float *rec = kp.blk_recs + ((size_t)r * kp.nrec + grp) * kp.blk_rec_stride;
            ++grp; grp_tiles = 0;
            if (t == 0) { rec[REC_M] = m_run; rec[REC_IDX] = __int_as_float(idx_run); rec[REC_S] = s_run; rec[REC_PAD] = 0.0f; }
            if (METHOD != RANDOM_SAMPLING)
                for (int p = t; p < kp.P; p += QTB) { rec[REC_V + p] = s.v[p]; s.v[p] = 0.0f; }
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
    shard_reduce<METHOD, QTB>(kp, r, s, M, IDX, S);
    phase_mark(kp, 9);
    after_shard_reduce<PARAM, ADAPTIVE, METHOD, QTB>(kp, r, s, M, IDX, S);
    phase_mark(kp, 11);
    if (t == 0) kp.counters[r] = 0u;
}

}  // namespace qp
