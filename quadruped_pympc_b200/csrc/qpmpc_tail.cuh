// qpmpc_tail.cuh -- the reduction tail of a LARGE launch as a second, multi-block kernel (CUDA-graph-captured pair /
// back-to-back launches on one stream).
//
// In the fused tick kernel the last block of a robot does everything after the rollouts alone: merge of the block
// records, CEM elite selection, update, read-out.  That is the right shape while there are few records (one 10k-sample
// controller: 79), but at 50 000 samples (BASELINE config 3: 391 records, 200 elites) the lone block spent 15 us merging
// records and 54 us selecting elites while 147 SMs idled, and at 1 M samples 1 184 records.  From QPMPC_TAIL_MIN_BLOCKS
// records on, the tick kernel stops after its block records (kp.defer_tail) and this kernel finishes the tick with
// kp.tail_nb blocks per robot:
//   1. (CEM) every block finds the threshold bin of the fine cost histogram the rollouts filled (65 536 bins of the cost
//      bits >> 16, only the occupied window is scanned) -- the same bin in every block -- and filters ITS SLICE of the costs: keys below the bin go to the `selected` list, keys inside it to
//      the `candidate` list (global memory, atomic append; the order inside the lists does not matter, the result is the
//      exact sorted top-k);
//   2. the block records are merged into TAIL_NPART = 32 partial records (online softmax, fixed slices and order -- the
//      same whatever the number of blocks, so results do not depend on how robots are grouped into launches);
//   3. last-block-done: the last block merges the tail_nb partial records, refines the candidates to the exact k
//      smallest (cost, index) keys (shared-memory radix select; a second 12-bit histogram level first when the threshold
//      bin is crowded), and continues exactly like the fused kernel's last block (shard record, peer exchange or finish,
//      publication).
// Results are identical to the fused path: the elite set is the exact stable argsort, the record merge is the same
// associative online-softmax merge in a fixed (two-level) order.
#pragma once
#include "qpmpc_tick.cuh"

namespace qp {

constexpr int TAIL_MAX_NB = 32;
constexpr int TAIL_NPART = 32;    // partial records per robot: FIXED, so that the merge order does not depend on how many blocks share the work
constexpr int TAIL_TB = 256;      // threads per tail block: the CEM elite refit stages up to 256 noise rows in one pass

// exact `need` smallest of `n` 64-bit keys (cost bits << 32 | index) that all share the cost prefix above bit
// (32 + pshift): appended to s_keys[*s_count ...].  keys in global memory.  s_cand: CEM_CAND u64, s_hist: 4096 u32
// (second level) -- both inside the tile scratch; s_misc: >= 8 u32.
template <int NT>
static __device__ void refine_candidates(const unsigned long long *keys, unsigned int n, int need, int pshift,
                                         unsigned long long *s_keys, unsigned int *s_count, unsigned long long *s_cand,
                                         unsigned int *s_h2, unsigned int *s_hist, unsigned int *s_misc) {
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    unsigned int m = n;
    unsigned int thr2 = 0xffffffffu;                     // second-level prefix (12 more cost bits), if used
    if (n > (unsigned int)CEM_CAND && pshift >= 12) {
        for (int i = t; i < CEM_BINS; i += NT) s_h2[i] = 0u;
        __syncthreads();
        for (unsigned int i = t; i < n; i += NT) {
            const unsigned int cb = (unsigned int)(__ldcg(keys + i) >> 32);
            atomicAdd(&s_h2[(cb >> (pshift - 12)) & (CEM_BINS - 1)], 1u);
        }
        __syncthreads();
        cem_find_bin<NT>([&](int i) { return s_h2[i]; }, (unsigned int)need, s_hist, s_misc);
        thr2 = s_misc[0];
        const unsigned int below2 = s_misc[1];
        m = s_misc[2];
        __syncthreads();
        need -= (int)below2;
        pshift -= 12;
    }
    // gather: keys below the (second-level) threshold are selected, keys inside it become shared-memory candidates
    if (t == 0) s_misc[4] = 0u;
    __syncthreads();
    const bool overflow = m > (unsigned int)CEM_CAND;    // thousands of (nearly) identical costs: exact slow path below
    if (!overflow) {
        for (unsigned int i = t; i < n; i += NT) {
            const unsigned long long key = __ldcg(keys + i);
            if (thr2 == 0xffffffffu) { const unsigned int slot = atomicAdd(&s_misc[4], 1u); s_cand[slot] = key; continue; }
            const unsigned int b2 = ((unsigned int)(key >> 32) >> pshift) & (CEM_BINS - 1);
            if (b2 < thr2) { const unsigned int slot = atomicAdd(s_count, 1u); if (slot < MAX_ELITE) s_keys[slot] = key; }
            else if (b2 == thr2) { const unsigned int slot = atomicAdd(&s_misc[4], 1u); if (slot < CEM_CAND) s_cand[slot] = key; }
        }
        __syncthreads();
    } else {
        // 8-bit MSB-first radix select straight over the global list (only keys inside the threshold prefix take part)
        unsigned long long prefix = 0ull, above = 0ull;
        const int low_bits = 32 + pshift;
        for (int shift = low_bits - 8; ; shift -= 8) {
            const int sh = shift < 0 ? 0 : shift;
            const int width = shift < 0 ? 8 + shift : 8;
            const unsigned int dmask = (1u << width) - 1u;
            for (int b = t; b < 256; b += NT) s_hist[b] = 0u;
            __syncthreads();
            for (unsigned int i = t; i < n; i += NT) {
                const unsigned long long key = __ldcg(keys + i);
                const bool inside = thr2 == 0xffffffffu || (((unsigned int)(key >> 32) >> pshift) & (CEM_BINS - 1)) == thr2;
                if (inside && (key & above) == prefix) atomicAdd(&s_hist[(unsigned int)(key >> sh) & dmask], 1u);
            }
            __syncthreads();
            if (t == 0) {
                unsigned int acc = 0u; int d = 0;
                for (; d < (int)dmask; ++d) { if (acc + s_hist[d] >= (unsigned int)need) break; acc += s_hist[d]; }
                s_misc[5] = (unsigned int)d; s_misc[6] = acc;
            }
            __syncthreads();
            prefix |= (unsigned long long)s_misc[5] << sh;
            above |= (unsigned long long)dmask << sh;
            need -= (int)s_misc[6];
            __syncthreads();
            if (sh == 0) break;
        }
        // `prefix` = the low bits of the need-th smallest key inside the threshold prefix: take everything <= it
        const unsigned long long low_mask = ~(~0ull << low_bits);
        for (unsigned int i = t; i < n; i += NT) {
            const unsigned long long key = __ldcg(keys + i);
            const unsigned int b2 = ((unsigned int)(key >> 32) >> pshift) & (CEM_BINS - 1);
            const bool take = thr2 == 0xffffffffu ? (key & low_mask) <= prefix
                                                  : (b2 < thr2 || (b2 == thr2 && (key & low_mask) <= prefix));
            if (take) { const unsigned int slot = atomicAdd(s_count, 1u); if (slot < MAX_ELITE) s_keys[slot] = key; }
        }
        __syncthreads();
        return;
    }
    // radix select of `need` among the m shared-memory candidates over the low (32 + pshift) key bits (early exit)
    const int low_bits = 32 + pshift;
    const unsigned long long low_mask = ~(~0ull << low_bits);
    unsigned long long thr_low = low_mask;
    if ((unsigned int)need < m) {
        unsigned long long prefix = 0ull, above = 0ull;
        for (int shift = low_bits - 8; ; shift -= 8) {
            const int sh = shift < 0 ? 0 : shift;
            const int width = shift < 0 ? 8 + shift : 8;
            const unsigned int dmask = (1u << width) - 1u;
            for (int b = t; b < 256; b += NT) s_hist[b] = 0u;
            __syncthreads();
            for (unsigned int c = t; c < m; c += NT) {
                const unsigned long long key = s_cand[c];
                if ((key & above) == prefix) atomicAdd(&s_hist[(unsigned int)(key >> sh) & dmask], 1u);
            }
            __syncthreads();
            if (wid == 0) {                           // 256 bins: 8 per lane, warp scan
                unsigned int h8[8], sum = 0u;
#pragma unroll
                for (int j = 0; j < 8; ++j) { h8[j] = s_hist[lane * 8 + j]; sum += h8[j]; }
                unsigned int inc = sum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned int v = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += v;
                }
                const unsigned int ex = inc - sum;
                if (ex < (unsigned int)need && ex + sum >= (unsigned int)need) {
                    unsigned int acc = ex;
                    int d = -1;
                    unsigned int accd = 0u, cntd = 0u;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (d < 0 && acc + h8[j] >= (unsigned int)need) { d = lane * 8 + j; accd = acc; cntd = h8[j]; }
                        acc += h8[j];
                    }
                    s_misc[5] = (unsigned int)d; s_misc[6] = accd; s_misc[7] = cntd;
                }
            }
            __syncthreads();
            const unsigned int d = s_misc[5], accd = s_misc[6], cntd = s_misc[7];
            prefix |= (unsigned long long)d << sh;
            above |= (unsigned long long)dmask << sh;
            need -= (int)accd;
            __syncthreads();
            if ((unsigned int)need == cntd || sh == 0) { thr_low = prefix | (low_mask & ~above); break; }
        }
    }
    for (unsigned int c = t; c < m; c += NT) {
        const unsigned long long key = s_cand[c];
        if ((key & low_mask) <= thr_low) { const unsigned int slot = atomicAdd(s_count, 1u); if (slot < MAX_ELITE) s_keys[slot] = key; }
    }
    __syncthreads();
}

// Block-wide search of the bin where the running count over `nb` consecutive bins of a GLOBAL histogram reaches `need`:
// s_misc[0..2] = (bin offset | 0xffffffff, count below, count inside).  s_warp: >= NT/32 u32.
template <int NT>
static __device__ void find_bin_window(const unsigned int *bins, int nb, unsigned int need, unsigned int *s_warp, unsigned int *s_misc) {
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    const int per = (nb + NT - 1) / NT, lo = t * per, hi = lo + per < nb ? lo + per : nb;
    unsigned int mysum = 0u;
    for (int i = lo; i < hi; ++i) mysum += __ldcg(bins + i);
    unsigned int incl = mysum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    __syncthreads();
    if (lane == 31) s_warp[wid] = incl;
    if (t == 0) s_misc[0] = 0xffffffffu;
    __syncthreads();
    unsigned int excl = incl - mysum;
    for (int w = 0; w < wid; ++w) excl += s_warp[w];
    if (excl < need && excl + mysum >= need) {                                    // exactly one thread
        unsigned int acc = excl;
        for (int i = lo; i < hi; ++i) {
            const unsigned int c = __ldcg(bins + i);
            if (acc + c >= need) { s_misc[0] = (unsigned int)i; s_misc[1] = acc; s_misc[2] = c; break; }
            acc += c;
        }
    }
    __syncthreads();
}

// per-robot scratch of the tail kernel (global memory): counters [8] = {selected, candidates, ticket, -, ...}, then the
// two key lists (n_local u64 each; the selected list holds < k keys, the candidate list at most n_local)
__host__ __device__ inline size_t tail_cnt_stride() { return 8; }

// grid = (tail_nb, R), block = TAIL_TB.
template <int PARAM, bool ADAPTIVE, int METHOD>
__global__ void __launch_bounds__(TAIL_TB, 1) tail_kernel(const __grid_constant__ KParams kp) {
    extern __shared__ __align__(16) float smem_raw[];
    constexpr int TB = TAIL_TB;                          // (shadows the tick kernels' block size inside this kernel)
    const Smem s(smem_raw, kp, TB);
    const int t = threadIdx.x, blk = blockIdx.x, r = blockIdx.y, NB = gridDim.x;
    __shared__ int s_last;
    unsigned int *cnt = kp.tail_cnt + (size_t)r * tail_cnt_stride();
    unsigned long long *sel = kp.tail_keys + (size_t)r * 2 * kp.n_local, *cand = sel + kp.n_local;
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(s.tile);
    unsigned int *hist = reinterpret_cast<unsigned int *>(s.tile + 2 * SCRATCH_KEYS);
    unsigned int *misc = hist + 256;
    const int k = kp.elite_k < kp.n_local ? kp.elite_k : kp.n_local;
    // Launched as a programmatic dependent of the tick kernel.  The tick's inputs (packet, carried controller state, device
    // contact tables) are not written by the tick kernel, so every block stages them NOW, while that kernel is still
    // running; then wait until its grid has completed and its stores are visible.
    stage_inputs<TB>(kp, r, s);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (blk == 0) phase_mark(kp, 8);
    // ---- 1. CEM: threshold bin of the (fine, windowed) cost histogram, then this block's slice of the costs ----
    // CEM: the first half of the blocks filters cost slices, the second half merges block records -- the two jobs overlap
    const int NBF = (METHOD == CEM_MPPI && NB >= 2) ? NB / 2 : 0;         // filter blocks
    const int NBM = NB - NBF;                                             // merge blocks = partial records
    unsigned int win_lo = 0u, win_hi = 0u;
    if (METHOD == CEM_MPPI) { win_lo = __ldcg(&cnt[4]); win_hi = __ldcg(&cnt[5]); }          // occupied bins (cem_count_cost)
    if (METHOD == CEM_MPPI && (blk < NBF || NB < 2)) {
        const int FB = NB < 2 ? 1 : NBF;
        const unsigned int *g_hist = kp.cem_hist + (size_t)r * CEM_BINS_FINE;
        find_bin_window<TB>(g_hist + win_lo, (int)(win_hi - win_lo + 1u), (unsigned int)k, hist, misc);
        const unsigned int thr = misc[0] == 0xffffffffu ? 0xffffffffu : win_lo + misc[0];
        __syncthreads();
        const int per = (((kp.n_local + FB - 1) / FB) + 3) & ~3;           // 16-byte aligned slices
        const int lo = blk * per, n_slice = kp.n_local - lo < per ? kp.n_local - lo : per;
        const float *costs = kp.costs + (size_t)r * kp.n_local;
        if (n_slice > 0 && thr != 0xffffffffu) {
            const int idx_base = kp.sample_base + lo;
            cem_for_each_cost<TB>(costs + lo, n_slice, [&](unsigned int cb, int i) {
                const unsigned int b = cb >> CEM_SHIFT_FINE;
                if (b <= thr) {
                    const unsigned long long key = ((unsigned long long)cb << 32) | (unsigned int)(idx_base + i);
                    if (b < thr) sel[atomicAdd(&cnt[0], 1u)] = key;
                    else cand[atomicAdd(&cnt[1], 1u)] = key;
                }
            });
        }
        if (blk == 0) phase_mark(kp, 4);
    }
    // ---- 2. this block's slice of the block records -> one partial record ----
    if (blk >= NBF || NB < 2) {
        const int mb = NB < 2 ? 0 : blk - NBF;
        const int per = (kp.nrec + TAIL_NPART - 1) / TAIL_NPART;           // block records per partial record
        for (int j = mb; j < TAIL_NPART; j += (NB < 2 ? 1 : NBM)) {
            const int lo = j * per, n = kp.nrec - lo < per ? kp.nrec - lo : per;
            float *part = kp.part_recs + ((size_t)r * TAIL_NPART + j) * kp.blk_rec_stride;
            float M = CUDART_INF_F, S = 0.0f;
            int IDX = 0x7fffffff;
            __syncthreads();
            if (n > 0) merge_records<METHOD, TB>(kp.blk_recs + ((size_t)r * kp.nrec + lo) * kp.blk_rec_stride, n, kp.blk_rec_stride, kp.P, s, M, IDX, S);
            if (t == 0) { part[REC_M] = M; part[REC_IDX] = __int_as_float(IDX); part[REC_S] = S; part[REC_PAD] = 0.0f; }
            if (METHOD != RANDOM_SAMPLING)
                for (int p = t; p < kp.P; p += TB) part[REC_V + p] = n > 0 ? s.v[p] : 0.0f;
        }
        if (mb == 0) phase_mark(kp, 5);
    }
    __syncthreads();
    if (t == 0) {
        __threadfence();
        const unsigned int ticket = atomicAdd(&cnt[2], 1u);
        s_last = (ticket == (unsigned int)(NB - 1));
        if (s_last) __threadfence();
    }
    __syncthreads();
    if (!s_last) return;
    // ---- 3. last block: the tick's inputs, the merged record, the exact elite set, then as the fused kernel ----
    phase_mark(kp, 1);
    float M, S; int IDX;
    merge_records<METHOD, TB>(kp.part_recs + (size_t)r * TAIL_NPART * kp.blk_rec_stride, TAIL_NPART, kp.blk_rec_stride, kp.P, s, M, IDX, S, &kp);
    float *rec = kp.shard_send + (size_t)r * kp.shard_rec_stride;
    if (t == 0) { rec[REC_M] = M; rec[REC_IDX] = __int_as_float(IDX); rec[REC_S] = S; rec[REC_PAD] = 0.0f; }
    if (METHOD != RANDOM_SAMPLING)
        for (int p = t; p < kp.P; p += TB) rec[REC_V + p] = s.v[p];
    if (METHOD == CEM_MPPI) {
        __syncthreads();
        const unsigned int nsel = __ldcg(&cnt[0]), m = __ldcg(&cnt[1]);
        unsigned int *g_hist = kp.cem_hist + (size_t)r * CEM_BINS_FINE;
        for (unsigned int i = win_lo + t; i <= win_hi; i += TB) g_hist[i] = 0u;                    // clean for the next tick
        if (t == 0) misc[3] = 0u;
        __syncthreads();
        if (m > 0 && (int)nsel < k)
            refine_candidates<TB>(cand, m, k - (int)nsel, CEM_SHIFT_FINE, keys, &misc[3], keys + MAX_ELITE,
                                  reinterpret_cast<unsigned int *>(keys), hist, misc);
        __syncthreads();
        for (unsigned int i = t; i < nsel; i += TB) { const unsigned int slot = atomicAdd(&misc[3], 1u); if (slot < MAX_ELITE) keys[slot] = __ldcg(sel + i); }
        __syncthreads();
        const unsigned int got = misc[3] < (unsigned int)MAX_ELITE ? misc[3] : (unsigned int)MAX_ELITE;   // == k
        // rank sort (keys are unique: the index is part of the key): thread t places keys[t] at the number of smaller keys.
        // 2 barriers instead of the 36 barrier-separated stages of a 256-key bitonic network.
        static_assert(TAIL_TB >= MAX_ELITE, "one thread per elite key");
        unsigned long long *sorted = keys + MAX_ELITE;    // (the candidate list is dead by now)
        const unsigned long long mine = t < (int)got ? keys[t] : ~0ull;
        int rank = 0;
        for (unsigned int j = 0; j < got; ++j) rank += (keys[j] < mine) ? 1 : 0;
        __syncthreads();
        if (t < (int)got) sorted[rank] = mine;
        __syncthreads();
        for (int j = t; j < kp.elite_k; j += TB) {
            const unsigned long long key = j < k && j < (int)got ? sorted[j] : ~0ull;
            rec[REC_V + kp.P + 2 * j] = __uint_as_float((unsigned int)(key >> 32));
            rec[REC_V + kp.P + 2 * j + 1] = __uint_as_float((unsigned int)(key & 0xffffffffu));
        }
        __syncthreads();
    }
    if (t == 0) { cnt[0] = 0u; cnt[1] = 0u; cnt[2] = 0u; cnt[4] = 0xffffffffu; cnt[5] = 0u; }
    phase_mark(kp, 9);
    after_shard_reduce<PARAM, ADAPTIVE, METHOD, TB>(kp, r, s, M, IDX, S);
    phase_mark(kp, 11);
}

}  // namespace qp
