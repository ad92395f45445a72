// qpmpc_x2.cuh -- throughput variant of the tick kernel: TWO SAMPLES PER THREAD in packed FP32.
//
// The sample-per-thread kernel is issue-bound (ncu r1m, 512 robots x 10k: issue slots 74 % busy, half of the issued
// instructions FP32, none packed).  sm_100a issues fma.rn.f32x2 / mul.f32x2 / add.f32x2 as ONE instruction (FFMA2 /
// FMUL2 / FADD2) for two lanes of a 64-bit register pair, and block-uniform scalars ride along as `.F32` broadcast
// operands.  Here thread t owns samples (2t, 2t + 1) of a 256-sample tile: the whole rollout (spline, clamp, moment
// arms, SRBD step, cost; qpmpc_math.cuh / rollout_fixed_t<PARAM, float2>) runs once for both, so the FP32 half of the
// instruction stream halves, and so do the block-uniform loads (reference, feet, basis, stance masks) and the loop
// control.  Philox / Box-Muller, the friction clamps (FMNMX per lane) and the SFU reciprocal stay per sample.
// Same Philox counters, same explicit round-to-nearest arithmetic: every sample's cost is bit-identical to
// tick_kernel's (tests/test_gpu_x2.py).  Fixed gait only: in the gait-adaptive rollout the two samples of a pair
// follow different contact sequences (per-sample step frequency), which would turn the block-uniform stance branches
// into per-lane selects.
#pragma once
#include "qpmpc_tick.cuh"

namespace qp {

constexpr int XTB = 128;          // threads per block
constexpr int XS = 2 * XTB;       // samples per tile
constexpr int XTS = XS + 2;       // tile row stride (floats): even (8-byte aligned pairs), bank = (2p + slot) % 32

// knot p of this thread's two samples: best[p] + (eps[p][2t], eps[p][2t + 1])
struct Knots2 {
    const float *s_best, *tpair;
    __device__ __forceinline__ float2 operator()(int p) const {
        return qadd(*reinterpret_cast<const float2 *>(tpair + p * XTS), s_best[p]);
    }
    __device__ __forceinline__ XY<float2> pair(int px, int py) const { return xy_make((*this)(px), (*this)(py)); }
};

// compact_weights for two samples per thread: (weight, slot) pairs of the non-zero weights in ascending slot order
// (slot = 2 * thread + lane half), the same addends in the same order as a loop over all slots.
__device__ __forceinline__ int compact_weights_x2(float w0, float w1, int t, const Smem &s, float &s_tile) {
    const int lane = t & 31, wid = t >> 5;
    const bool nz0 = (w0 != 0.0f), nz1 = (w1 != 0.0f);
    const unsigned int b0 = __ballot_sync(0xffffffffu, nz0), b1 = __ballot_sync(0xffffffffu, nz1);
    __shared__ int s_wcnt[XTB / 32];
    if (lane == 0) s_wcnt[wid] = __popc(b0) + __popc(b1);
    s_tile = block_sum<XTB>(w0 + w1, s.red);                            // its two barriers publish s_wcnt
    const unsigned int below = (1u << lane) - 1u;
    int off = __popc(b0 & below) + __popc(b1 & below), total = 0;
#pragma unroll
    for (int k = 0; k < XTB / 32; ++k) { const int c = s_wcnt[k]; total += c; if (k < wid) off += c; }
    float2 *wp = reinterpret_cast<float2 *>(s.w);
    if (nz0) wp[off] = make_float2(w0, __int_as_float(2 * t));
    if (nz1) wp[off + (nz0 ? 1 : 0)] = make_float2(w1, __int_as_float(2 * t + 1));
    __syncthreads();
    return total;
}

// ======================================================================================================
// grid = (nblk, R), block = XTB threads = XS samples per tile.  Same records / last-block merge / finish as tick_kernel.
// ======================================================================================================
template <int PARAM, int METHOD>
__global__ void __launch_bounds__(XTB, 2) tick_kernel_x2(const __grid_constant__ KParams kp) {
    extern __shared__ __align__(16) float smem_raw[];
    const Smem s(smem_raw, kp, XS);                              // s.w holds up to XS (weight, slot) pairs
    const int t = threadIdx.x, blk = blockIdx.x, r = blockIdx.y;
    __shared__ int s_last;
    if (blk == 0) phase_mark(kp, 0);
    if (kp.defer_tail) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // see tick_body
    stage_inputs<XTB>(kp, r, s);
    if (blk == 0) phase_mark(kp, 1);
    const uint32_t seed = __float_as_uint(s.in[IN_SEED]), robot = kp.robot_base + (uint32_t)r, tick = __float_as_uint(s.in[IN_TICK]);
    const float *s_sigma = s.in + kp.off_sigma;
    float *tpair = s.tile + 2 * t;
    float m_run = CUDART_INF_F, s_run = 0.0f;
    int idx_run = 0x7fffffff;
    for (int p = t; p < kp.P; p += XTB) s.v[p] = 0.0f;

    int grp = blk * (kp.iters / kp.iters_rec), grp_tiles = 0;     // record group of the current tile, tiles folded into it
    for (int it = 0; it < kp.iters; ++it) {
        const int tile_base = (blk * kp.iters + it) * XS;       // local (shard) index of this tile's first sample
        if (tile_base >= kp.n_local) break;                     // block-uniform
        const int local0 = tile_base + 2 * t;
        const bool act0 = local0 < kp.n_local, act1 = local0 + 1 < kp.n_local;
        const int gi0 = kp.sample_base + local0, gi1 = gi0 + 1; // global sample indices (rows of the noise tensor)
        // ---- 1. eps tile: tile[p][slot] = additional_random_parameters[gi(slot)][p] ----
        if (kp.noise_mode == NOISE_INJECTED) {
            const int rows = (kp.n_local - tile_base) < XS ? (kp.n_local - tile_base) : XS;
            const float *src = kp.noise + ((size_t)r * kp.N + kp.sample_base + tile_base) * kp.P;
            for (int e = t; e < XS * kp.P; e += XTB) {
                const int row = e / kp.P, p = e - row * kp.P;
                s.tile[p * XTS + row] = row < rows ? __ldg(src + e) : 0.0f;
            }
            __syncthreads();
        } else {
            NoiseRow nr0 = noise_row<METHOD>(kp, gi0), nr1 = noise_row<METHOD>(kp, gi1);
            nr0.zero = nr0.zero || !act0;                       // row 0 (the previous best re-evaluated) and padding rows
            nr1.zero = nr1.zero || !act1;
            float *tp = tpair;
            const int nq = kp.P >> 2, noct = (nq + 1) >> 1;
            for (int c = 0; c < noct; ++c, tp += 8 * XTS) {
                float4 a0, b0, a1, b1;
                noise_oct<METHOD>(nr0, seed, robot, tick, c, s_sigma, a0, b0);
                noise_oct<METHOD>(nr1, seed, robot, tick, c, s_sigma, a1, b1);
                *reinterpret_cast<float2 *>(tp) = make_float2(a0.x, a1.x);
                *reinterpret_cast<float2 *>(tp + XTS) = make_float2(a0.y, a1.y);
                *reinterpret_cast<float2 *>(tp + 2 * XTS) = make_float2(a0.z, a1.z);
                *reinterpret_cast<float2 *>(tp + 3 * XTS) = make_float2(a0.w, a1.w);
                if (2 * c + 1 < nq) {
                    *reinterpret_cast<float2 *>(tp + 4 * XTS) = make_float2(b0.x, b1.x);
                    *reinterpret_cast<float2 *>(tp + 5 * XTS) = make_float2(b0.y, b1.y);
                    *reinterpret_cast<float2 *>(tp + 6 * XTS) = make_float2(b0.z, b1.z);
                    *reinterpret_cast<float2 *>(tp + 7 * XTS) = make_float2(b0.w, b1.w);
                }
            }
        }
        if (blk == 0 && it == 0) phase_mark(kp, 2);
        // ---- 2. rollout + cost of both samples ----
        float c0 = CUDART_INF_F, c1 = CUDART_INF_F;
        if (act0) {
            const Knots2 knot{s.in + kp.off_best, tpair};
            const float2 c = rollout_fixed_t<PARAM, float2>(kp, s.in, s.fref, knot);
            c0 = c.x; c1 = act1 ? c.y : CUDART_INF_F;
            if (kp.mode == MODE_TICK) {                          // NaN / +-Inf -> 1e6, S:686-687
                if (isnan(c0) || isinf(c0)) c0 = 1000000.0f;
                if (act1 && (isnan(c1) || isinf(c1))) c1 = 1000000.0f;
            }
            float *cdst = kp.costs + (size_t)r * kp.n_local + local0;
            cdst[0] = c0;
            if (act1) cdst[1] = c1;
            if (METHOD == CEM_MPPI && kp.mode == MODE_TICK) {    // first radix level of the elite selection (block_topk_hist)
                cem_count_cost(kp, r, c0);
                if (act1) cem_count_cost(kp, r, c1);
            }
        }
        if (kp.mode == MODE_ROLLOUT) { __syncthreads(); continue; }
        if (blk == 0 && it == 0) phase_mark(kp, 3);
        // ---- 3. tile reduction folded into the block's running record ----
        float cmin = c0;
        int imin = act0 ? gi0 : 0x7fffffff;
        if (c1 < cmin) { cmin = c1; imin = gi1; }               // lexicographic: gi0 < gi1 wins ties
        block_argmin<XTB>(cmin, imin, s.red);
        if (METHOD != RANDOM_SAMPLING) {
            const float m_new = fminf(m_run, cmin);              // cmin is finite: the tile has >= 1 active sample
            const float scale_old = (m_run == CUDART_INF_F) ? 0.0f : expf(-(m_run - m_new));
            const float w0 = act0 ? expf(-(c0 - m_new)) : 0.0f;  // exp(-(c - beta)/T), T = 1, S:828-830
            const float w1 = act1 ? expf(-(c1 - m_new)) : 0.0f;
            float s_tile;
            const int nnz = compact_weights_x2(w0, w1, t, s, s_tile);
            s_run = fmaf(s_run, scale_old, s_tile);
            const float2 *wi = reinterpret_cast<const float2 *>(s.w);
            for (int p = t; p < kp.P; p += XTB) {
                const float *rowp = s.tile + p * XTS;
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
        if (++grp_tiles == kp.iters_rec || (blk * kp.iters + it + 1) * XS >= kp.n_local) {
            float *rec = kp.blk_recs + ((size_t)r * kp.nrec + grp) * kp.blk_rec_stride;
            ++grp; grp_tiles = 0;
            if (t == 0) { rec[REC_M] = m_run; rec[REC_IDX] = __int_as_float(idx_run); rec[REC_S] = s_run; rec[REC_PAD] = 0.0f; }
            if (METHOD != RANDOM_SAMPLING)
                for (int p = t; p < kp.P; p += XTB) { rec[REC_V + p] = s.v[p]; s.v[p] = 0.0f; }
            m_run = CUDART_INF_F; s_run = 0.0f; idx_run = 0x7fffffff;
        }
        __syncthreads();
    }
    if (kp.mode == MODE_ROLLOUT) return;
    if (blk == 0) phase_mark(kp, 4);
    if (kp.defer_tail) return;                          // large launches: tail_kernel (qpmpc_tail.cuh) finishes the tick
    // release / acquire around the ticket by ONE thread between two block barriers (see tick_body)
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
    shard_reduce<METHOD, XTB>(kp, r, s, M, IDX, S);
    phase_mark(kp, 9);
    after_shard_reduce<PARAM, false, METHOD, XTB>(kp, r, s, M, IDX, S);
    phase_mark(kp, 11);
    if (t == 0) kp.counters[r] = 0u;
}

}  // namespace qp
