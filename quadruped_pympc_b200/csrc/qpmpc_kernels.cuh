// qpmpc_kernels.cuh -- device code of libqpmpc_b200 (sm_100a).
//
// One fused launch per MPC tick:  noise (Philox4x32-10, or an injected tensor)  ->  perturb  ->  spline  ->
// clamp -> SRBD explicit-Euler rollout + cost with the horizon loop in registers -> block-level
// (min, argmin) / online-softmax (m, s, sum w*eps) partials -> last-block-done merge -> MPPI/CEM update ->
// GRF read-out + predicted state.   grid = (blocks per robot, robots), one sample per thread per tile.
//
// Reference lines restated here (paths relative to the reference's quadruped_pympc/):
//   S = controllers/sampling/centroidal_nmpc_jax.py, M = controllers/sampling/centroidal_model_jax.py,
//   A = controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py, G = helpers/periodic_gait_generator_jax.py
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "qpmpc_math.cuh"

namespace qp {

constexpr int TB = 128;          // threads per block = samples per tile
constexpr int TS = TB + 1;       // tile row stride (floats): bank = (p + t) % 32 for both access patterns
constexpr int MAX_H = 64;
constexpr int MAX_NS = 16;
constexpr int MAX_ELITE = 256;
constexpr int CEM_SHIFT = 20;    // first-level digit of the elite selection: sign + exponent + 3 mantissa bits of the cost
constexpr int CEM_BINS = 4096;
constexpr int CEM_CAND = 2048 - MAX_ELITE;   // candidate keys (threshold bin) held in shared memory during the refinement
constexpr int MAX_PEERS = 8;     // ranks of one sample-sharded controller that exchange records through peer memory (one NVSwitch node)
constexpr int PACKET_MAX = 448;  // floats of one tick packet that may travel in kernel-parameter space (cubic 300, zero-order 396)

enum { ZERO_ORDER = 0, LINEAR = 1, CUBIC = 2 };
enum { RANDOM_SAMPLING = 0, MPPI = 1, CEM_MPPI = 2 };
enum { MODE_TICK = 0, MODE_ROLLOUT = 2 };             // MODE_ROLLOUT: explicit params in, raw costs out
enum { NOISE_PHILOX = 0, NOISE_INJECTED = 1 };

// ---- per-robot input record (floats), written by the host ------------------------------------------------
enum { IN_STATE = 0, IN_REF = 24, IN_PHASE = 48, IN_NOMFREQ = 52, IN_OPTSWING = 53, IN_SEED = 54, IN_TICK = 55,
       IN_ROBOT = 56, IN_CONTACT = 60 };              // contact: 4 x H ; then best (P) ; then sigma (P)
// ---- per-robot output record -------------------------------------------------------------------------------
enum { OUT_GRF = 0, OUT_PRED = 12, OUT_BESTCOST = 36, OUT_BESTFREQ = 37, OUT_BESTIDX = 38, OUT_NELITE = 39,
       OUT_BEST = 40 };                               // then sigma_new (P) ; then elite indices (elite_k ints)
// ---- block / shard record -------------------------------------------------------------------------------------
enum { REC_M = 0, REC_IDX = 1, REC_S = 2, REC_PAD = 3, REC_V = 4 };   // then V (P) ; shard: then top-k (cost, idx) pairs

struct KParams {
    int N;             // samples of the whole controller (global)
    int n_local;       // samples evaluated by this launch (this shard)
    int sample_base;   // global index of this shard's first sample
    int H, ns, P, P_leg;
    int iters;         // tiles per block
    int nblk;          // blocks per robot
    int nrec;          // block records per robot: one per group of `iters_rec` tiles (depends on n_local only, so that the
    int iters_rec;     //   merge order -- and with it every result bit -- is the same however many tiles a block walks)
    int elite_k;
    int mode;          // MODE_TICK | MODE_ROLLOUT
    int noise_mode;    // NOISE_PHILOX | NOISE_INJECTED
    int freq_injected; // 1: per-sample step frequencies come from `freqs`
    int fused_finish;  // 1: the last block also merges the (single) shard record and writes `out`
    int world;         // shard records to merge in finish
    int carry_best;    // 1: write best_new back into the input record (device-resident multi-tick runs)
    int off_best, off_sigma, in_stride;
    int off_out_sigma, off_out_elite, out_stride;
    int blk_rec_stride, shard_rec_stride;
    float inv_m, mg, fz_min, fz_max, mu;    // (fz_min, fz_max): an aligned pair, one 64-bit uniform load per clamped leg
    alignas(16) float I[12], Iinv[12];      // row-major 3 x 3 (+ padding: 128-bit uniform loads)
    float qd[12];
    int q_xyyaw;                            // 1: q_diag weights x, y or yaw (the reference's Q does not)
    uint32_t bm_bits[2];                    // 0x3F800000, 0x40000000: exponent patterns of box_muller16 (see there)
    alignas(16) float qpair[8];              // q_diag as the pairs srbd_step sums: (vx,vy) (z,vz) (roll,pitch) (wx,wy)
    float mpc_dt, duty;
    float sigma_rs[3], sigma_mppi;
    float freq_avail[3];
    float inv_ns, denom_fixed_inv;          // 1/ns ; 1/(H/ns)
    float dt[MAX_H];
    alignas(16) float basis[MAX_H][4];      // fixed gait, per step: cubic Hermite weights of the chunk's knots p0..p3 | linear (1-q, q, -, -)
    float bounds[MAX_NS + 1];               // linspace(0,H,ns+1), S:187
    unsigned char sp_index[MAX_H];          // fixed gait: spline chunk index per step, S:189
    const float *in;                        // [R][in_stride]
    float *out;                             // [R][out_stride]
    float *costs;                           // [R][n_local]
    float *blk_recs;                        // [R][nblk][blk_rec_stride]
    float *shard_send;                      // [R][shard_rec_stride]  (this shard's record)
    const float *shard_recv;                // [world][R][shard_rec_stride]
    unsigned int *counters;                 // [R]
    unsigned int *cem_hist;                 // [R][CEM_BINS] histogram of the cost bits >> CEM_SHIFT, filled by every sample (CEM only)
    const float *noise;                     // injected: [R][N][P] (additional noise, or explicit params in MODE_ROLLOUT)
    const float *freqs;                     // injected: [R][N]
    // ---- direct I/O (one controller, host buffers): no copy-engine hops around the kernel ----
    int param_packet;                       // 1: robot 0's tick packet is `packet` below (kernel-parameter space), not kp.in
    unsigned int seq;                       // completion stamp stored after the result packet
    float *host_out;                        // mapped pinned host alias: [out_stride floats][stamp], or NULL
    float packet[PACKET_MAX];
    // ---- peer exchange of a sample-sharded controller (fused collective, qpmpc_tick.cuh peer_exchange) ----
    // ---- stateful controller (qpmpc_step): best / sigma / Philox tick / previous contact live on the
    //      device across ticks; the host uploads only state, reference, phase, frequency and the contact table ----
    const float *contact_dev;               // [R][4][H] contact tables made on the device (qpmpc_pgg_contact_sequence), or NULL
    float *ctrl_state;                      // [R][ctrl_stride] = {best P, sigma P, tick u32, prev-contact mask u32, pad}, or NULL
    int ctrl_stride;
    int sigma_reset;                        // 1: this launch starts from sigma0 (I:128-129: reset at iteration 0 of every tick)
    float sigma0;
    int shift_kind;                         // 0: off | 1 + PARAM: this launch first shifts the carried best (shift_solution,
    int shift_index;                        //   S:513-561 from prepare_state_and_reference S:571-573): spline chunk / step
    float shift_w[4];                       //   index and basis at step = 1/mpc_frequency, evaluated on the host
    uint32_t seed;                          // Philox counter word 3 of this launch (all robots of a batch share it)
    uint32_t robot_base;                    // global robot id = robot_base + blockIdx.y (< 2^23), part of counter word 0
    int peer_world;                         // > 1: the last block pushes its shard record into every peer's mailbox over
    int peer_rank;                          //      NVLink, waits for the peers' records and finishes in the same launch
    float *peer_mbox[MAX_PEERS];            // mailbox base of every rank (own entry = local memory): see peer_exchange
    unsigned long long peer_timeout_ns;     // the in-kernel wait for the peers' records gives up after this long
    unsigned long long *phase_ns;           // debug (qpmpc_debug_phase_times): %globaltimer at phase boundaries, or NULL
    // ---- two-stage launches (qpmpc_tail.cuh): the tick kernel stops after its block records, tail_kernel finishes ----
    int defer_tail;                         // 1: no last-block stage in the tick kernel
    int tail_nb;                            // blocks per robot of tail_kernel
    float *part_recs;                       // [R][TAIL_NPART][blk_rec_stride] partial records
    unsigned int *tail_cnt;                 // [R][8] {selected, candidates, ticket, ...}, zero between launches
    unsigned long long *tail_keys;          // [R][2][n_local] selected / candidate keys of the elite selection (CEM only)
};

static_assert(sizeof(KParams) <= 4096, "KParams must fit the classic 4 KB kernel-parameter space");

// phase slots: block 0 of robot 0 -> [0..5] = start, staged, noise done, rollout done, tile reduced, record written ;
//              the last block of robot 0 -> [8..11] = ticket taken, blocks merged, finish done (best), tail written
__device__ __forceinline__ void phase_mark(const KParams &kp, int slot) {
    if (kp.phase_ns != nullptr && threadIdx.x == 0 && blockIdx.y == 0) {
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        kp.phase_ns[slot] = ns;
    }
}

// ======================================================================================================
// RNG: Philox4x32-10 in counter mode under a FIXED key (host mirror: oracle/sampling_oracle.c):
//   counter = (quad | stream << 8 | robot << 9,  row,  tick,  seed),   key = (PHILOX_K0, PHILOX_K1)
// Everything that varies -- per thread, per robot, per tick, per seed -- sits in the 128-bit counter, which enters once,
// in round 0; the key is a compile-time constant, so the 2 x 10 round keys are immediates and cost no instruction
// (keyed by (seed, robot) they cost 2 integer adds per round per thread: 6 % of the whole tick, ncu r1d).  A fixed-key
// Philox is a fixed bijection of the counter space: distinct (quad, stream, robot, row, tick, seed) -> distinct blocks.
// ======================================================================================================
constexpr uint32_t PHILOX_K0 = 0xA4093822u, PHILOX_K1 = 0x299F31D0u;   // digits of pi (nothing up the sleeve)
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float u01_from_bits(uint32_t b) { return __uint_as_float((b >> 9) | 0x3F800000u) - 1.0f; }

// Two normals from ONE 32-bit word (Box-Muller on the SFU: lg2 / sqrt / sin / cos approximations): a Philox block of
// four words yields EIGHT normals (r2: it yielded four; the generator was 22 % of the executed instructions).
//   radius from bits 7..22 (in place: they ARE the 16 leading mantissa bits):  u = 2 - [1,2) (step 2^-16) in [2^-16, 1]
//                                 ->  r = sqrt(c * log2 u),  c = -2 ln2 * s^2 (the row's scale s folded into the
//                                 constant);  |z| <= 4.71 s
//   angle  from the other 16 bits (0..6 and 23..31: the word rotated by 16 puts them on the mantissa):
//                                 th = pi * ([2,4) - 3)  in [-pi, pi), 65 536 directions
// Per word: one LOP3 (mask | exponent) for the radius, one rotate + one LOP3 for the angle.  `one` / `two` are the
// exponent patterns 0x3F800000 / 0x40000000 handed in as kernel parameters (KParams::bm_bits): LOP3 encodes a single
// immediate, so as literals the compiler splits each (w & mask) | exponent into two instructions.
// (The reference draws with jax.random.normal = sqrt(2) * erfinv(uniform), resolution 2^-23, |z| <= 5.4; only the
//  distribution is part of the contract -- parity tests inject or export the tensor; tests/test_gpu_parity.py checks
//  moments, tail mass and a Kolmogorov-Smirnov distance of the device's own draw.)
__device__ __forceinline__ uint32_t mask_or(uint32_t a, uint32_t mask, uint32_t bits) {      // (a & mask) | bits
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(mask), "r"(bits));
    return d;
}
__device__ __forceinline__ void box_muller16(uint32_t w, float c, uint32_t one, uint32_t two, float &z0, float &z1) {
    const float u = 2.0f - __uint_as_float(mask_or(w, 0x007fff80u, one));
    const float th = fmaf(__uint_as_float(mask_or(__funnelshift_l(w, w, 16), 0x007fff80u, two)), 3.14159265f, -9.42477796f);
    float l2, r, sn, cs;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(u));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(c * l2));
    asm("sin.approx.ftz.f32 %0, %1;" : "=f"(sn) : "f"(th));
    asm("cos.approx.ftz.f32 %0, %1;" : "=f"(cs) : "f"(th));
    const float2 z = __fmul2_rn(make_float2(cs, sn), make_float2(r, r));   // one FMUL2, r as the broadcast operand
    z0 = z.x;
    z1 = z.y;
}
constexpr float BM_C1 = -1.38629436f;         // -2 ln 2: unit-variance normals

// One quad (4 consecutive knots) / one oct (8 consecutive knots = one Philox block) of additional_random_parameters
// for global row i.
//  random_sampling S:647-677: rows [1,1+k) s0*N, [1+k,1+2k) s1*N (same key => same normals), [1+2k,N) U(-s2,s2)
//  mppi S:806-812: s*N on rows 1..N-1 ; cem_mppi S:951-958: N*sigma[p] ; row 0 is always zero.
// NoiseRow: what depends on the row only (computed once per sample, outside the loop over quads).
struct NoiseRow { uint32_t j; int seg; float scale, c; bool zero; uint32_t one, two; };   // c = -2 ln2 * scale^2 (cem_mppi: -2 ln2); one/two: box_muller16

template <int METHOD>
__device__ __forceinline__ NoiseRow noise_row(const KParams &kp, int i) {
    NoiseRow nr;
    nr.zero = (i == 0); nr.seg = 0; nr.j = (uint32_t)(i - 1); nr.scale = kp.sigma_mppi;
    if (METHOD == RANDOM_SAMPLING) {
        const int k = kp.N / 3;
        if (i >= 1 + 2 * k) { nr.seg = 2; nr.j = (uint32_t)(i - 1 - 2 * k); }
        else if (i >= 1 + k) { nr.seg = 1; nr.j = (uint32_t)(i - 1 - k); }
        nr.scale = kp.sigma_rs[nr.seg];
    }
    nr.c = (METHOD == CEM_MPPI) ? BM_C1 : (BM_C1 * nr.scale) * nr.scale;
    nr.one = kp.bm_bits[0]; nr.two = kp.bm_bits[1];
    return nr;
}

// uniform rows of random_sampling: one Philox block per quad, 23-bit uniforms (counter word 0 = quad index)
__device__ __forceinline__ float4 uniform_quad(const NoiseRow &nr, uint32_t seed, uint32_t robot, uint32_t tick, int q) {
    const uint4 r = philox4x32_10((uint32_t)q | (robot << 9), nr.j, tick, seed, PHILOX_K0, PHILOX_K1);
    const float s2 = nr.scale, span = s2 - (-s2);
    float4 v;
    v.x = fmaxf(fmaf(u01_from_bits(r.x), span, -s2), -s2);
    v.y = fmaxf(fmaf(u01_from_bits(r.y), span, -s2), -s2);
    v.z = fmaxf(fmaf(u01_from_bits(r.z), span, -s2), -s2);
    v.w = fmaxf(fmaf(u01_from_bits(r.w), span, -s2), -s2);
    return v;
}
// four normals of a row from two words of its Philox block; knots 4q .. 4q+3 (cem_mppi: times sigma[p])
template <int METHOD>
__device__ __forceinline__ float4 normal_quad(const NoiseRow &nr, uint32_t w0, uint32_t w1, int q, const float *s_sigma) {
    float4 v;
    box_muller16(w0, nr.c, nr.one, nr.two, v.x, v.y);
    box_muller16(w1, nr.c, nr.one, nr.two, v.z, v.w);
    if (METHOD == CEM_MPPI) {
        v.x *= s_sigma[4 * q + 0]; v.y *= s_sigma[4 * q + 1]; v.z *= s_sigma[4 * q + 2]; v.w *= s_sigma[4 * q + 3];
    }
    return v;
}

// Oct c of a row >= 1 (the caller has dealt with the all-zero row 0): quads 2c (lo) and 2c+1 (hi) from ONE Philox block,
// counter = (c | robot << 9, row, tick, seed).  `hi` is meaningless when 2c+1 is not a quad of the row (P/4 odd).
template <int METHOD>
__device__ __forceinline__ void noise_oct_nz(const NoiseRow &nr, uint32_t seed, uint32_t robot, uint32_t tick, int c,
                                             const float *s_sigma, float4 &lo, float4 &hi) {
    if (METHOD == RANDOM_SAMPLING && nr.seg == 2) {
        lo = uniform_quad(nr, seed, robot, tick, 2 * c);
        hi = uniform_quad(nr, seed, robot, tick, 2 * c + 1);
        return;
    }
    const uint4 r = philox4x32_10((uint32_t)c | (robot << 9), nr.j, tick, seed, PHILOX_K0, PHILOX_K1);
    lo = normal_quad<METHOD>(nr, r.x, r.y, 2 * c, s_sigma);
    hi = normal_quad<METHOD>(nr, r.z, r.w, 2 * c + 1, s_sigma);
}
template <int METHOD>
__device__ __forceinline__ void noise_oct(const NoiseRow &nr, uint32_t seed, uint32_t robot, uint32_t tick, int c,
                                          const float *s_sigma, float4 &lo, float4 &hi) {
    if (nr.zero) { lo = make_float4(0.f, 0.f, 0.f, 0.f); hi = lo; return; }
    noise_oct_nz<METHOD>(nr, seed, robot, tick, c, s_sigma, lo, hi);
}

// one quad alone (finish stage, export kernels): half of its oct
template <int METHOD>
__device__ __forceinline__ float4 noise_quad(const NoiseRow &nr, uint32_t seed, uint32_t robot, uint32_t tick, int q,
                                             const float *s_sigma) {
    if (nr.zero) return make_float4(0.f, 0.f, 0.f, 0.f);
    if (METHOD == RANDOM_SAMPLING && nr.seg == 2) return uniform_quad(nr, seed, robot, tick, q);
    const uint4 r = philox4x32_10((uint32_t)(q >> 1) | (robot << 9), nr.j, tick, seed, PHILOX_K0, PHILOX_K1);
    return (q & 1) ? normal_quad<METHOD>(nr, r.z, r.w, q, s_sigma) : normal_quad<METHOD>(nr, r.x, r.y, q, s_sigma);
}
// convenience form for one-off rows (finish stage, export kernels)
template <int METHOD>
__device__ __forceinline__ float4 noise_quad(const KParams &kp, uint32_t seed, uint32_t robot, uint32_t tick, int i,
                                             int q, const float *s_sigma) {
    return noise_quad<METHOD>(noise_row<METHOD>(kp, i), seed, robot, tick, q, s_sigma);
}

// Per-sample step frequency (row a15).  A:688-692 (random_sampling) | A:835-838 (mppi) | A:994-1000 (cem_mppi)
template <int METHOD>
__device__ __forceinline__ float sample_step_freq(const KParams &kp, const float *s_in, uint32_t seed, uint32_t robot,
                                                  int robot_slot, int i) {
    if (kp.freq_injected) return __ldg(kp.freqs + (size_t)robot_slot * kp.N + i);
    const uint32_t tick = __float_as_uint(s_in[IN_TICK]);
    uint4 r = philox4x32_10(0x100u | (robot << 9), (uint32_t)i, tick, seed, PHILOX_K0, PHILOX_K1);
    const int c = (int)__umulhi(r.x, 3u);
    const float nominal = s_in[IN_NOMFREQ];
    const int opt = __float_as_int(s_in[IN_OPTSWING]);
    if (METHOD == RANDOM_SAMPLING) return opt ? kp.freq_avail[c] : nominal;
    if (METHOD == MPPI) return kp.freq_avail[c];
    const float inc = (c == 0) ? 0.0f : (c == 1 ? 0.2f : 0.4f);
    return fmaf(inc, (float)opt, nominal);
}

// ======================================================================================================
// SRBD: M:93-174 (fd + explicit Euler) and the cost S:450-451, one step, state in registers.
// Written once for T = float (one sample per thread) and T = float2 (two samples per thread, packed FP32):
// see qpmpc_math.cuh -- explicit round-to-nearest operations only, so both lane types give the same bits.
// ======================================================================================================
// 0, but not to the compiler (see rollout_fixed_t)
__device__ __forceinline__ int opaque_zero() {
    int z;
    asm volatile("mov.u32 %0, 0;" : "=r"(z));
    return z;
}

template <class T> struct BodyT { T px, py, pz, vx, vy, vz, roll, pitch, yaw, wx, wy, wz; };
typedef BodyT<float> Body;

// S:270-314 (box friction pyramid, same comparison order).  where(x > lo, x, lo) == fmaxf(x, lo) and where(x < hi, x, hi)
// == fminf(x, hi) for every input the reference can produce, NaN included (a NaN force takes the bound, as `where`
// does; only the sign of a zero can differ, which no later product can see): one FMNMX instead of FSETP + FSEL.
template <class T>
__device__ __forceinline__ void clamp_leg(const KParams &kp, float mu, T &fx, T &fy, T &fz) {
    fz = qmin(qmax(fz, kp.fz_min), kp.fz_max);
    const T hi = qmul(fz, mu), lo = qneg(hi);
    fx = qmin(qmax(fx, lo), hi);
    fy = qmin(qmax(fy, lo), hi);
}
template <class T>
__device__ __forceinline__ void clamp_leg(const KParams &kp, T &fx, T &fy, T &fz) { clamp_leg(kp, kp.mu, fx, fy, fz); }

// Fx..: sum of stance forces; tx..: sum of (foot - com) x F (world frame), M:111-118,132-135
template <class T>
__device__ __forceinline__ void add_leg(const KParams &kp, const BodyT<T> &b, float footx, float footy, float footz, T fx, T fy, T fz,
                                        T &Fx, T &Fy, T &Fz, T &tx, T &ty, T &tz) {
    const T rx = qsub(footx, b.px), ry = qsub(footy, b.py), rz = qsub(footz, b.pz);
    Fx = qadd(Fx, fx); Fy = qadd(Fy, fy); Fz = qadd(Fz, fz);
    tx = qfnma(rz, fy, qfma(ry, fz, tx));               // the moment accumulates term by term: two FMAs per component
    ty = qfnma(rx, fz, qfma(rz, fx, ty));
    tz = qfnma(ry, fx, qfma(rx, fy, tz));
}
template <class T>
__device__ __forceinline__ void add_leg(const KParams &kp, const BodyT<T> &b, const float *foot, T fx, T fy, T fz,
                                        T &Fx, T &Fy, T &Fz, T &tx, T &ty, T &tz) {
    add_leg(kp, b, foot[0], foot[1], foot[2], fx, fy, fz, Fx, Fy, Fz, tx, ty, tz);
}

// one term of e^T Q e on the state after the step, S:450-451 (the reference's Q also weights x, y, yaw with 0: kept, so
// that a non-default q_diag stays exact; a test-and-branch per term costs more than the three instructions it skips)
template <class T>
__device__ __forceinline__ void cost_term(T &c, T x, float ref, float q) {
    const T e = qsub(x, ref);
    c = qfma(qmul(e, q), e, c);
}

// two terms of e^T Q e as one unit: nref = -(reference pair) (shared memory, stage_inputs), q = weight pair (KParams::qpair)
template <class T>
__device__ __forceinline__ void cost_pair(XY<T> &acc, T x, T y, XY<T> nref, XY<T> q) {
    const XY<T> e = xy_add(xy_make(x, y), nref);
    acc = xy_fmav(xy_mulv(e, q), e, acc);
}

template <class T>
__device__ __forceinline__ T srbd_step(const KParams &kp, BodyT<T> &b, T Fx, T Fy, T Fz, T tx, T ty, T tz, float dt,
                                       const float *ref, const float *s_yaw) {
    T sr, cr, sp, cp, sy, cy;
    sincos_rpy(b.roll, b.pitch, b.yaw, s_yaw, sr, cr, sp, cp, sy, cy);
    // Euler rates = inverse of the conjugate Euler-rate matrix times w, M:124-130,137.  The matrix is
    // [[1, 0, -sp], [0, cr, cp sr], [0, -sr, cp cr]]; its inverse in closed form is [[1, sr tp, cr tp], [0, cr, -sr],
    // [0, sr / cp, cr / cp]] (determinant cp), so the rates need one reciprocal on the SFU (<= 1 ulp) and six FMAs
    const T swcw = qfma(cr, b.wz, qmul(sr, b.wy));
    const T dpitch = qfnma(sr, b.wz, qmul(cr, b.wy));
    const T dyaw = qmul(swcw, rcp_approx(cp));
    const T droll = qfma(sp, dyaw, b.wx);
    // b_R_w tau, M:141-155: b_R_w = Rx(roll)^T Ry(pitch)^T Rz(yaw)^T applied as three plane rotations (12 operations;
    // forming the nine matrix entries first takes 19)
    const T ux = qfma(sy, ty, qmul(cy, tx)), uy = qfnma(sy, tx, qmul(cy, ty));
    const T bx = qfnma(sp, tz, qmul(cp, ux)), uz = qfma(cp, tz, qmul(sp, ux));
    const T by = qfma(sr, uz, qmul(cr, uy)), bz = qfnma(sr, uy, qmul(cr, uz));
    // angular acceleration Iinv * (b_R_w tau - w x (I w)), M:157-159
    const T Iwx = qfma(b.wz, kp.I[2], qfma(b.wy, kp.I[1], qmul(b.wx, kp.I[0])));
    const T Iwy = qfma(b.wz, kp.I[5], qfma(b.wy, kp.I[4], qmul(b.wx, kp.I[3])));
    const T Iwz = qfma(b.wz, kp.I[8], qfma(b.wy, kp.I[7], qmul(b.wx, kp.I[6])));
    const T hx = qfma(b.wz, Iwy, qfnma(b.wy, Iwz, bx));
    const T hy = qfma(b.wx, Iwz, qfnma(b.wz, Iwx, by));
    const T hz = qfma(b.wy, Iwx, qfnma(b.wx, Iwy, bz));
    const T dwx = qfma(hz, kp.Iinv[2], qfma(hy, kp.Iinv[1], qmul(hx, kp.Iinv[0])));
    const T dwy = qfma(hz, kp.Iinv[5], qfma(hy, kp.Iinv[4], qmul(hx, kp.Iinv[3])));
    const T dwz = qfma(hz, kp.Iinv[8], qfma(hy, kp.Iinv[7], qmul(hx, kp.Iinv[6])));
    // explicit Euler, M:171-172 (all derivatives from the old state); linear acceleration F / m + g (M:118) with
    // dt / m and g dt as per-step uniforms
    const float dtm = dt * kp.inv_m, gdt = -9.81f * dt;
    b.px = qfma(b.vx, dt, b.px); b.py = qfma(b.vy, dt, b.py); b.pz = qfma(b.vz, dt, b.pz);
    b.vx = qfma(Fx, dtm, b.vx); b.vy = qfma(Fy, dtm, b.vy); b.vz = qfma(Fz, dtm, qadd(b.vz, gdt));
    b.roll = qfma(droll, dt, b.roll); b.pitch = qfma(dpitch, dt, b.pitch); b.yaw = qfma(dyaw, dt, b.yaw);
    b.wx = qfma(dwx, dt, b.wx); b.wy = qfma(dwy, dt, b.wy); b.wz = qfma(dwz, dt, b.wz);
    // e^T Q e on the state after the step, S:450-451 (diagonal Q).  The terms are summed as two chains of four -- (vx, pz,
    // roll, wx) and (vy, vz, pitch, wy) -- that travel as one unit (FADD2 / FMUL2 / FFMA2 for one sample per thread: three
    // instructions per two terms instead of six), then wz; the same association for every lane type.
    const float *nref = s_yaw + 4 + MAX_H;      // -(ref) of the four pairs, behind the stance masks (Smem::cref)
    XY<T> acc;
    { const XY<T> e = xy_add(xy_make(b.vx, b.vy), xy_uniform<T>(nref));
      acc = xy_mulv(xy_mulv(e, xy_uniform<T>(kp.qpair)), e); }
    cost_pair(acc, b.pz, b.vz, xy_uniform<T>(nref + 2), xy_uniform<T>(kp.qpair + 2));
    cost_pair(acc, b.roll, b.pitch, xy_uniform<T>(nref + 4), xy_uniform<T>(kp.qpair + 4));
    cost_pair(acc, b.wx, b.wy, xy_uniform<T>(nref + 6), xy_uniform<T>(kp.qpair + 6));
    T c = qadd(acc.x(), acc.y());
    cost_term(c, b.wz, ref[11], kp.qd[11]);
    // the reference's Q weights x, y and yaw with 0 (S:118-130): (e * 0) * e adds exactly nothing for every finite e, so
    // the three terms sit behind ONE block-uniform branch (taken only for a non-default q_diag, which stays exact)
    if (kp.q_xyyaw) {
        cost_term(c, b.px, ref[0], kp.qd[0]);
        cost_term(c, b.py, ref[1], kp.qd[1]);
        cost_term(c, b.yaw, ref[8], kp.qd[8]);
    }
    return c;
}

template <class T>
__device__ __forceinline__ BodyT<T> load_body_t(const float *s) {
    BodyT<T> b;
    b.px = qset<T>(s[0]); b.py = qset<T>(s[1]); b.pz = qset<T>(s[2]); b.vx = qset<T>(s[3]); b.vy = qset<T>(s[4]);
    b.vz = qset<T>(s[5]); b.roll = qset<T>(s[6]); b.pitch = qset<T>(s[7]); b.yaw = qset<T>(s[8]); b.wx = qset<T>(s[9]);
    b.wy = qset<T>(s[10]); b.wz = qset<T>(s[11]);
    return b;
}
__device__ __forceinline__ Body load_body(const float *s) { return load_body_t<float>(s); }

// knot j of leg l for the thread whose tile column is `tcol`: best + eps  (S:680 | S:814 | S:960)
#define QP_KNOT(l, j) (s_best[(l) * kp.P_leg + (j)] + tcol[((l) * kp.P_leg + (j)) * TS])

// Knot sources of the fixed-gait rollout: knot p of this thread's sample(s) = best[p] + eps[p]
struct Knots1 {                       // one sample per thread: tile column `tcol`, row stride TS
    const float *s_best, *tcol;
    __device__ __forceinline__ float operator()(int p) const { return qadd(tcol[p * TS], s_best[p]); }
    // knots px (of the x axis) and py (of the y axis) as one packed add
    __device__ __forceinline__ XY<float> pair(int px, int py) const {
        return xy_add(xy_make(tcol[px * TS], tcol[py * TS]), xy_make(s_best[px], s_best[py]));
    }
};

// Fixed-gait rollout (S:316-496): every leg's spline clock is the step n, contact table shared by all samples,
// so the swing/stance branch is block-uniform and swing legs are skipped (their force is multiplied by 0 twice,
// S:388-402 and M:111-116).  s_fref[n] = mass*9.81 / n_stance(n) (S:377-380).
template <int PARAM, class T, class KNOTS>
__device__ __forceinline__ T rollout_fixed_t(const KParams &kp, const float *s_in, const float *s_fref, const KNOTS &knot) {
    const int *s_cmask = reinterpret_cast<const int *>(s_fref + MAX_H + 4);   // stance mask per step (stage_inputs)
    // The block-uniform operands of the step (feet, reference, yaw0) are read through a pointer the compiler cannot prove
    // uniform (+ an opaque zero): one base register and LDS [R + imm].  As uniform addresses every group of loads
    // re-derives the CTA's shared-memory window (S2UR + UMOV + ULEA, ~8 issue slots per step: the uniform registers are
    // taken by the constants of the step).  The stance masks stay uniform: their branches must not look divergent.
    const int zero = opaque_zero();
    const float *s_inv = s_in + zero, *s_frefv = s_fref + zero;
    const float *s_yaw = s_frefv + MAX_H;
    const float *ref = s_inv + IN_REF;
    const float4 *feet4 = reinterpret_cast<const float4 *>(s_frefv + 2 * MAX_H + 4 + 8);     // Smem::feet4
    BodyT<T> b = load_body_t<T>(s_in + IN_STATE);
    T cost = qset<T>(0.0f);
    const int H = kp.H, PL = kp.P_leg;
    constexpr float third = 1.0f / 3.0f;   // forces_z / forces_x = 30/10, S:388
    // hoisted knots of the current spline chunk (cubic: k1, phi, k2, phi_next ; linear: k0, k1)
    // (x and y axes as XY units: one packed operation for both when T = float, see qpmpc_math.cuh; z on its own)
    XY<T> kaxy[4], kbxy[4], kcxy[4], kdxy[4];
    T kaz[4], kbz[4], kcz[4], kdz[4];
    int cur = -1;
#pragma unroll 1
    for (int n = 0; n < H; ++n) {
        if (PARAM != ZERO_ORDER) {
            const int idx = kp.sp_index[n];
            if (idx != cur) {
                cur = idx;
#pragma unroll
                for (int l = 0; l < 4; ++l) {
                    if (PARAM == CUBIC) {                     // S:219-255
                        const int sx = l * PL + 10 * idx, sy = sx + 4, sz = sx + 8;
                        // the knots of the chunk as they are: kp.basis holds the Hermite weights per KNOT (host)
                        kaxy[l] = knot.pair(sx, sy); kbxy[l] = knot.pair(sx + 1, sy + 1);
                        kcxy[l] = knot.pair(sx + 2, sy + 2); kdxy[l] = knot.pair(sx + 3, sy + 3);
                        kaz[l] = knot(sz); kbz[l] = knot(sz + 1); kcz[l] = knot(sz + 2); kdz[l] = knot(sz + 3);
                    } else {                                  // S:196-199
                        const int sx = l * PL + idx, sy = sx + (kp.ns + 1), sz = sy + (kp.ns + 1);
                        kaxy[l] = knot.pair(sx, sy);
                        kbxy[l] = knot.pair(sx + 1, sy + 1);
                        kaz[l] = knot(sz);
                        kbz[l] = knot(sz + 1);
                    }
                }
            }
        }
        const float w0 = kp.basis[n][0], w1 = kp.basis[n][1], w2 = kp.basis[n][2], w3 = kp.basis[n][3];
        const float fref = s_fref[n];
        T Fx = qset<T>(0.f), Fy = qset<T>(0.f), Fz = qset<T>(0.f), tx = qset<T>(0.f), ty = qset<T>(0.f), tz = qset<T>(0.f);
        const int cm = s_cmask[n];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            if (cm & (1 << l)) {
                XY<T> fxy;
                T fz;
                // spline (weights per knot) with the gravity compensation as the start of the z chain (S:382-385);
                // x and y forces / 3 (S:388-402 with contact == 1)
                if (PARAM == CUBIC) {
                    fxy = xy_mul(xy_fma(kdxy[l], w3, xy_fma(kcxy[l], w2, xy_fma(kbxy[l], w1, xy_mul(kaxy[l], w0)))), third);
                    fz = qfma(kdz[l], w3, qfma(kcz[l], w2, qfma(kbz[l], w1, qfma(kaz[l], w0, qset<T>(fref)))));
                } else if (PARAM == LINEAR) {
                    fxy = xy_mul(xy_fma(kbxy[l], w1, xy_mul(kaxy[l], w0)), third);
                    fz = qfma(kbz[l], w1, qfma(kaz[l], w0, qset<T>(fref)));
                } else {                                      // S:259-268
                    fxy = xy_mul(knot.pair(l * PL + n, l * PL + n + H), third);
                    fz = qadd(knot(l * PL + n + 2 * H), fref);
                }
                T fx = fxy.x(), fy = fxy.y();
                const float4 ft = feet4[l];                   // (foot x, y, z, mu)
                clamp_leg(kp, ft.w, fx, fy, fz);              // S:405-409
                add_leg(kp, b, ft.x, ft.y, ft.z, fx, fy, fz, Fx, Fy, Fz, tx, ty, tz);
            }
        }
        cost = qadd(cost, srbd_step(kp, b, Fx, Fy, Fz, tx, ty, tz, kp.dt[n], ref, s_yaw));   // S:446-451,491
    }
    return cost;
}

template <int PARAM>
__device__ __forceinline__ float rollout_fixed(const KParams &kp, const float *s_in, const float *s_fref, const float *tcol) {
    const Knots1 knot{s_in + kp.off_best, tcol};
    return rollout_fixed_t<PARAM, float>(kp, s_in, s_fref, knot);
}

// Spline with a per-thread clock (adaptive rollouts and nothing else): S:181-268 with step = n_leg, A:353-374.
template <int PARAM>
__device__ __forceinline__ void spline_dyn(const KParams &kp, const float *s_best, const float *tcol, int l, int step,
                                           float inv_denom, float &fx, float &fy, float &fz) {
    const int PL = kp.P_leg;
    if (PARAM == ZERO_ORDER) {
        int j0 = step, j1 = step + kp.H, j2 = step + 2 * kp.H;    // int16(step); negative indices wrap like jnp
        if (j0 < 0) j0 += PL;
        if (j1 < 0) j1 += PL;
        if (j2 < 0) j2 += PL;
        fx = QP_KNOT(l, j0); fy = QP_KNOT(l, j1); fz = QP_KNOT(l, j2);
        return;
    }
    const float stepf = (float)step;
    int idx = 0;
    for (int i = 1; i <= kp.ns; ++i) idx = (stepf >= kp.bounds[i]) ? i : idx;      // S:187-189
    idx = idx < kp.ns ? idx : kp.ns - 1;                                           // unreachable guard (step < H)
    const float q = fmaf(stepf, inv_denom, -(float)idx);                           // S:191-194
    if (PARAM == LINEAR) {
        const int sh = kp.ns + 1;
        fx = (1.0f - q) * QP_KNOT(l, idx) + q * QP_KNOT(l, idx + 1);
        fy = (1.0f - q) * QP_KNOT(l, idx + sh) + q * QP_KNOT(l, idx + sh + 1);
        fz = (1.0f - q) * QP_KNOT(l, idx + 2 * sh) + q * QP_KNOT(l, idx + 2 * sh + 1);
        return;
    }
    const float q2 = q * q, q3 = q2 * q;
    const float a = 2.0f * q3 - 3.0f * q2 + 1.0f, bb = q3 - 2.0f * q2 + q, c = -2.0f * q3 + 3.0f * q2, d = q3 - q2;
    const int s0 = 10 * idx;
    float f[3];
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        const float p0 = QP_KNOT(l, s0 + 4 * ax), p1 = QP_KNOT(l, s0 + 4 * ax + 1), p2 = QP_KNOT(l, s0 + 4 * ax + 2),
                    p3 = QP_KNOT(l, s0 + 4 * ax + 3);
        const float phi = 0.5f * ((p2 - p1) + (p1 - p0)), phin = 0.5f * ((p3 - p2) + (p2 - p1));
        f[ax] = a * p1 + bb * phi + c * p2 + d * phin;
    }
    fx = f[0]; fy = f[1]; fz = f[2];
}

// The same cubic spline with the chunk's knots kept in registers: a leg's chunk index changes at most ns - 1 times over the
// horizon, so p1, phi, p2, phi_next of the three axes (the same sums as spline_dyn) are rebuilt only when it does --
// 24 shared loads and 27 additions per stance leg and step less, identical results.
struct ChunkKnots { float v[12]; int idx; };   // per leg: [axis][p1, phi, p2, phi_next]
__device__ __forceinline__ void spline_dyn_cubic_cached(const KParams &kp, const float *s_best, const float *tcol, int l, int step,
                                                        float inv_denom, ChunkKnots &ck, float &fx, float &fy, float &fz) {
    const float stepf = (float)step;
    int idx = 0;
    for (int i = 1; i <= kp.ns; ++i) idx = (stepf >= kp.bounds[i]) ? i : idx;      // S:187-189
    idx = idx < kp.ns ? idx : kp.ns - 1;
    const float q = fmaf(stepf, inv_denom, -(float)idx);                           // S:191-194
    if (idx != ck.idx) {
        ck.idx = idx;
        const int s0 = 10 * idx;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const float p0 = QP_KNOT(l, s0 + 4 * ax), p1 = QP_KNOT(l, s0 + 4 * ax + 1), p2 = QP_KNOT(l, s0 + 4 * ax + 2),
                        p3 = QP_KNOT(l, s0 + 4 * ax + 3);
            ck.v[4 * ax] = p1; ck.v[4 * ax + 1] = 0.5f * ((p2 - p1) + (p1 - p0));
            ck.v[4 * ax + 2] = p2; ck.v[4 * ax + 3] = 0.5f * ((p3 - p2) + (p2 - p1));
        }
    }
    const float q2 = q * q, q3 = q2 * q;
    const float a = 2.0f * q3 - 3.0f * q2 + 1.0f, bb = q3 - 2.0f * q2 + q, c = -2.0f * q3 + 3.0f * q2, d = q3 - q2;
    fx = a * ck.v[0] + bb * ck.v[1] + c * ck.v[2] + d * ck.v[3];
    fy = a * ck.v[4] + bb * ck.v[5] + c * ck.v[6] + d * ck.v[7];
    fz = a * ck.v[8] + bb * ck.v[9] + c * ck.v[10] + d * ck.v[11];
}

// Gait-adaptive rollout (A:326-501): per-sample contact sequence from the phase clock (G:68-89,136-151),
// per-leg stance counters as spline clocks, plus the step-frequency cost A:500.
template <int PARAM>
__device__ __forceinline__ float rollout_adaptive(const KParams &kp, const float *s_in, const float *s_yaw, const float *tcol, float freq) {
    const float *s_best = s_in + kp.off_best;
    const float *ref = s_in + IN_REF;
    const float *feet = s_in + IN_STATE + 12;
    const int H = kp.H;
    constexpr float third = 1.0f / 3.0f;
    const float adv = kp.mpc_dt * freq;                       // G:79
    float tt[4], inv_denom[4];
    int nleg[4];
    // pass 1: horizon_leg = sum_n contact + 1, A:345-348
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        float t = s_in[IN_PHASE + l];
        float cnt = 0.0f;
        for (int n = 0; n < H; ++n) {
            t = t >= 1.0f ? 0.0f : t;                          // G:73-76
            t = t + adv;                                       // G:79-82
            cnt += (t < kp.duty) ? 1.0f : 0.0f;                // G:84-87
        }
        inv_denom[l] = 1.0f / ((cnt + 1.0f) * kp.inv_ns);      // 1 / (horizon_leg / num_spline), S:191
        tt[l] = s_in[IN_PHASE + l];
        nleg[l] = -1;                                          // A:339
    }
    Body b = load_body(s_in + IN_STATE);
    float cost = 0.0f;
    ChunkKnots ck[4];
#pragma unroll
    for (int l = 0; l < 4; ++l) ck[l].idx = -1;
#pragma unroll 1
    for (int n = 0; n < H; ++n) {
        float c[4];
        float nst = 0.0f;
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            float t = tt[l];
            t = t >= 1.0f ? 0.0f : t;
            t = t + adv;
            tt[l] = t;
            c[l] = (t < kp.duty) ? 1.0f : 0.0f;
            nleg[l] += (int)c[l];                              // A:353-356
            nst += c[l];
        }
        const float fref = kp.mg / nst;                        // A (=S:377-380); inf when no leg is in stance
        float Fx = 0.f, Fy = 0.f, Fz = 0.f, tx = 0.f, ty = 0.f, tz = 0.f;
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            if (c[l] != 0.0f) {                                // swing legs contribute exactly zero (see rollout_fixed)
                float fx, fy, fz;
                if (PARAM == CUBIC) spline_dyn_cubic_cached(kp, s_best, tcol, l, nleg[l], inv_denom[l], ck[l], fx, fy, fz);
                else spline_dyn<PARAM>(kp, s_best, tcol, l, nleg[l], inv_denom[l], fx, fy, fz);
                fz = fref + fz;
                fx = fx * third; fy = fy * third;
                clamp_leg(kp, fx, fy, fz);
                add_leg(kp, b, feet + 3 * l, fx, fy, fz, Fx, Fy, Fz, tx, ty, tz);
            }
        }
        cost += srbd_step(kp, b, Fx, Fy, Fz, tx, ty, tz, kp.dt[n], ref, s_yaw);
    }
    const float df = freq - 1.3f;
    cost += df * 100.0f * df;                                  // A:500
    return cost;
}

// ======================================================================================================
// block-level helpers
// ======================================================================================================
// lexicographic (cost, index) minimum over the block; result valid in every thread. s_red: >= 2*(TB/32) floats
template <int NT>
__device__ __forceinline__ void block_argmin(float &c, int &i, float *s_red) {
    constexpr int TB = NT;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oc = __shfl_xor_sync(0xffffffffu, c, o);
        const int oi = __shfl_xor_sync(0xffffffffu, i, o);
        if (oc < c || (oc == c && oi < i)) { c = oc; i = oi; }
    }
    const int w = threadIdx.x >> 5, nw = TB / 32;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { s_red[w] = c; s_red[nw + w] = __int_as_float(i); }
    __syncthreads();
    c = s_red[0]; i = __float_as_int(s_red[nw]);
#pragma unroll
    for (int k = 1; k < nw; ++k) {
        const float oc = s_red[k];
        const int oi = __float_as_int(s_red[nw + k]);
        if (oc < c || (oc == c && oi < i)) { c = oc; i = oi; }
    }
}

// fixed-order block sum (deterministic); result valid in every thread
template <int NT>
__device__ __forceinline__ float block_sum(float v, float *s_red) {
    constexpr int TB = NT;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, nw = TB / 32;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[w] = v;
    __syncthreads();
    float s = s_red[0];
#pragma unroll
    for (int k = 1; k < nw; ++k) s += s_red[k];
    return s;
}

// Tail shared by the three control laws: GRF read-out at step 0 + one Euler step (S:695-784 = :839-928 = :985-1073).
// spline(best_leg, 0.0, 1): zero-order -> p[0],p[H],p[2H]; linear -> p[0],p[ns+1],p[2(ns+1)]; cubic (q=0) -> p[1],p[5],p[9].
// Executed by one thread.  best: P floats (shared or global).
template <int PARAM>
__device__ void tail_grf_pred(const KParams &kp, const float *s_in, const float *s_yaw, const float *best, float *out) {
    const float *s_contact = s_in + IN_CONTACT;
    const int H = kp.H;
    float c[4], nst;
#pragma unroll
    for (int l = 0; l < 4; ++l) c[l] = s_contact[l * H];
    nst = ((c[0] + c[1]) + c[2]) + c[3];                       // S:714-716
    const float fref = kp.mg / nst;                            // S:717
    Body b = load_body(s_in + IN_STATE);
    float Fx = 0.f, Fy = 0.f, Fz = 0.f, tx = 0.f, ty = 0.f, tz = 0.f;
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        const float *p = best + l * kp.P_leg;
        float fx, fy, fz;
        if (PARAM == ZERO_ORDER) { fx = p[0]; fy = p[H]; fz = p[2 * H]; }
        else if (PARAM == LINEAR) { fx = p[0]; fy = p[kp.ns + 1]; fz = p[2 * (kp.ns + 1)]; }
        else { fx = p[1]; fy = p[5]; fz = p[9]; }
        fz = fref + fz;                                        // S:719-722
        fx = fx * c[l] / 3.0f; fy = fy * c[l] / 3.0f; fz = fz * c[l];   // S:724-738
        clamp_leg(kp, fx, fy, fz);                             // S:741-745
        out[OUT_GRF + 3 * l] = fx; out[OUT_GRF + 3 * l + 1] = fy; out[OUT_GRF + 3 * l + 2] = fz;
        if (c[l] != 0.0f) {                                    // M:111-116,132-135 multiply by stance again
            const float *foot = s_in + IN_STATE + 12 + 3 * l;
            add_leg(kp, b, foot, fx * c[l], fy * c[l], fz * c[l], Fx, Fy, Fz, tx, ty, tz);
        }
    }
    const float zero_ref[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    (void)srbd_step(kp, b, Fx, Fy, Fz, tx, ty, tz, kp.dt[0], zero_ref, s_yaw);   // S:784
    out[OUT_PRED + 0] = b.px; out[OUT_PRED + 1] = b.py; out[OUT_PRED + 2] = b.pz;
    out[OUT_PRED + 3] = b.vx; out[OUT_PRED + 4] = b.vy; out[OUT_PRED + 5] = b.vz;
    out[OUT_PRED + 6] = b.roll; out[OUT_PRED + 7] = b.pitch; out[OUT_PRED + 8] = b.yaw;
    out[OUT_PRED + 9] = b.wx; out[OUT_PRED + 10] = b.wy; out[OUT_PRED + 11] = b.wz;
    for (int k = 12; k < 24; ++k) out[OUT_PRED + k] = s_in[IN_STATE + k];  // M:174 feet unchanged
}

// ------------------------------------------------------------------------------------------------------
// Exact top-k by (cost, index) over `n` costs in global memory: 64-bit composite keys (cost bits << 32 | idx),
// MSB-first 8-bit radix select with shared-memory histograms, then a bitonic sort of the k winners.
// costs are >= 0 (sum of squares; NaN/Inf already saturated to 1e6), so their bit patterns order like the floats.
// Result: s_keys[0..k) sorted ascending = argsort(costs, stable)[:k] (S:1075).  s_keys: >= MAX_ELITE u64,
// s_hist: >= 256 u32.  Called by all TB threads of one block.
// ------------------------------------------------------------------------------------------------------
template <int NT>
static __device__ void block_topk(const float *costs, int n, int idx_base, int k, unsigned long long *s_keys, unsigned int *s_hist,
                           unsigned int *s_misc) {
    constexpr int TB = NT;
    const int t = threadIdx.x;
    unsigned long long prefix = 0ull;      // selected high bits so far
    int need = k;                          // how many we still need among keys matching the prefix
    int decided_bits = 0;
    // After the passes: every key < prefix(threshold) is in, plus all keys == threshold path ... composite keys are
    // unique, so after 8 passes exactly one key matches the full prefix and `need` == 1.
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = 56 - 8 * pass;
        for (int b = t; b < 256; b += TB) s_hist[b] = 0u;
        __syncthreads();
        const unsigned long long mask = decided_bits == 0 ? 0ull : (~0ull << (64 - decided_bits));
        for (int i = t; i < n; i += TB) {
            const unsigned long long key = ((unsigned long long)__float_as_uint(__ldcg(costs + i)) << 32) | (unsigned int)(idx_base + i);
            if ((key & mask) == prefix) atomicAdd(&s_hist[(unsigned int)(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (t == 0) {
            unsigned int acc = 0u; int d = 0;
            for (; d < 256; ++d) { if (acc + s_hist[d] >= (unsigned int)need) break; acc += s_hist[d]; }
            s_misc[0] = (unsigned int)d; s_misc[1] = acc;
        }
        __syncthreads();
        prefix |= ((unsigned long long)s_misc[0]) << shift;
        need -= (int)s_misc[1];
        decided_bits += 8;
        __syncthreads();
    }
    // prefix is now the k-th smallest composite key; gather every key <= prefix (exactly k of them)
    if (t == 0) s_misc[2] = 0u;
    for (int i = t; i < MAX_ELITE; i += TB) s_keys[i] = ~0ull;
    __syncthreads();
    for (int i = t; i < n; i += TB) {
        const unsigned long long key = ((unsigned long long)__float_as_uint(__ldcg(costs + i)) << 32) | (unsigned int)(idx_base + i);
        if (key <= prefix) { const unsigned int slot = atomicAdd(&s_misc[2], 1u); if (slot < MAX_ELITE) s_keys[slot] = key; }
    }
    __syncthreads();
    // bitonic sort of MAX_ELITE keys (padding = ~0)
    for (int size = 2; size <= MAX_ELITE; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = t; i < MAX_ELITE / 2; i += TB) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool up = ((lo & size) == 0);
                const unsigned long long a = s_keys[lo], b = s_keys[hi];
                if ((a > b) == up) { s_keys[lo] = b; s_keys[hi] = a; }
            }
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Exact top-k by (cost, index), histogram-assisted (the CEM elite selection of one shard, S:1075 | A:1115).
// Every sample has already counted itself into g_hist[cost bits >> CEM_SHIFT] (warp-aggregated REDs in the rollout
// phase, cem_count_cost), so the last block starts with the first radix level for free:
//   1. scan the 4096 bins (each thread owns CEM_BINS / NT consecutive bins) -> threshold bin b*, #keys below it; zero
//      the bins for the next tick;
//   1b. only if b* holds more keys than the shared-memory candidate list (cost distributions narrower than 1/8 octave):
//      one pass over the costs builds a second 4096-bin histogram of the next 12 cost bits inside b* (shared memory);
//   2. ONE pass over the costs (8 independent loads in flight per thread): prefix < threshold -> selected, prefix ==
//      threshold -> candidate list in shared memory;
//   3. radix-select the remaining (k - below) among the candidates in shared memory, 8-bit digits over the low key
//      bits, warp-parallel bin scan, early exit as soon as a bucket is taken whole (costs are nearly always distinct,
//      so this ends inside the remaining cost bits);
//   4. bitonic sort of the k winners.
// Falls back to block_topk (8 passes over global memory) if even the 24-bit prefix leaves more than CEM_CAND candidates
// (thousands of bit-identical costs).  Same contract as block_topk; s_keys: >= MAX_ELITE + CEM_CAND u64.
// ------------------------------------------------------------------------------------------------------
// Two-stage launches (kp.defer_tail, qpmpc_tail.cuh) count into a FINE histogram instead: cost bits >> CEM_SHIFT_FINE
// (65 536 bins, 128 per octave: narrow cost distributions -- HyQReal at 50k samples sits inside one octave -- put a few
// hundred keys into the threshold bin instead of half of all samples) and track the occupied bin window
// (tail_cnt[4] = min bin, [5] = max bin) so that the tail scans a few hundred bins, not 65 536.
constexpr int CEM_SHIFT_FINE = 16;
constexpr int CEM_BINS_FINE = 65536;
__device__ __forceinline__ void cem_count_cost(const KParams &kp, int r, float cost) {
    const unsigned int active = __activemask();
    if (!kp.defer_tail) {
        const unsigned int bin = __float_as_uint(cost) >> CEM_SHIFT;
        const unsigned int peers = __match_any_sync(active, bin);               // one RED per distinct bin per warp
        if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(kp.cem_hist + (size_t)r * CEM_BINS + bin, (unsigned int)__popc(peers));
        return;
    }
    const unsigned int bin = __float_as_uint(cost) >> CEM_SHIFT_FINE;
    const unsigned int peers = __match_any_sync(active, bin);
    if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(kp.cem_hist + (size_t)r * CEM_BINS_FINE + bin, (unsigned int)__popc(peers));
    const unsigned int lo = __reduce_min_sync(active, bin), hi = __reduce_max_sync(active, bin);
    if ((threadIdx.x & 31) == __ffs(active) - 1) {
        unsigned int *cnt = kp.tail_cnt + (size_t)r * 8;
        atomicMin(cnt + 4, lo);
        atomicMax(cnt + 5, hi);
    }
}

// Block-wide search of the bin where the running count reaches `need`; bins[] are read through LOAD(i), i in
// [0, CEM_BINS).  Result in s_misc[0..2] = (bin | 0xffffffff, count below, count inside).  s_warp: >= NT/32 u32.
template <int NT, class LOAD>
__device__ __forceinline__ void cem_find_bin(LOAD load, unsigned int need, unsigned int *s_warp, unsigned int *s_misc) {
    constexpr int BPT = CEM_BINS / NT;
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    unsigned int mysum = 0u;
#pragma unroll 8
    for (int j = 0; j < BPT; ++j) mysum += load(t * BPT + ((j + t) & (BPT - 1)));   // rotated: conflict-free in shared memory
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
        for (int j = 0; j < BPT; ++j) {
            const unsigned int c = load(t * BPT + j);
            if (acc + c >= need) { s_misc[0] = (unsigned int)(t * BPT + j); s_misc[1] = acc; s_misc[2] = c; break; }
            acc += c;
        }
    }
    __syncthreads();
}

// Visit every cost of one shard from a single block with many loads in flight: 8 x 16-byte loads per thread when the
// slice is 16-byte aligned (32 costs per thread per round), scalar loads otherwise.  f(cost bits, local index).
template <int NT, class F>
__device__ __forceinline__ void cem_for_each_cost(const float *costs, int n, F f) {
    const int t = threadIdx.x;
    if ((reinterpret_cast<uintptr_t>(costs) & 15) == 0) {
        const int n4 = n >> 2;
        const uint4 *c4 = reinterpret_cast<const uint4 *>(costs);
        for (int base = 0; base < n4; base += 8 * NT) {
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = base + u * NT + t;
                v[u] = i < n4 ? __ldcg(c4 + i) : make_uint4(~0u, ~0u, ~0u, ~0u);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = base + u * NT + t;
                if (i < n4) { f(v[u].x, 4 * i); f(v[u].y, 4 * i + 1); f(v[u].z, 4 * i + 2); f(v[u].w, 4 * i + 3); }
            }
        }
        for (int i = (n4 << 2) + t; i < n; i += NT) f(__float_as_uint(__ldcg(costs + i)), i);
        return;
    }
    for (int base = 0; base < n; base += 8 * NT) {
        unsigned int cb[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int i = base + u * NT + t;
            cb[u] = i < n ? __float_as_uint(__ldcg(costs + i)) : ~0u;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int i = base + u * NT + t;
            if (i < n) f(cb[u], i);
        }
    }
}

template <int NT>
static __device__ void block_topk_hist(const float *costs, int n, int idx_base, int k, unsigned int *g_hist,
                                       unsigned long long *s_keys, unsigned int *s_hist, unsigned int *s_misc) {
    static_assert((CEM_BINS / NT) >= 4 && ((CEM_BINS / NT) & (CEM_BINS / NT - 1)) == 0, "bins per thread: power of two");
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    // ---- 1. threshold bin of the first 12 key bits ----
    {   // stage the global histogram in shared memory (coalesced 16-byte loads), leave it zeroed for the next tick
        uint4 *s4 = reinterpret_cast<uint4 *>(s_keys);
        uint4 *g4 = reinterpret_cast<uint4 *>(g_hist);
        for (int i = t; i < CEM_BINS / 4; i += NT) { s4[i] = __ldcg(g4 + i); g4[i] = make_uint4(0u, 0u, 0u, 0u); }
        __syncthreads();
    }
    {
        const unsigned int *s_h1 = reinterpret_cast<const unsigned int *>(s_keys);
        cem_find_bin<NT>([&](int i) { return s_h1[i]; }, (unsigned int)k, s_hist, s_misc);
    }
    unsigned int thr_prefix = s_misc[0], below = s_misc[1], m = s_misc[2];
    int pshift = CEM_SHIFT;                           // the prefix is cost bits >> pshift
    __syncthreads();
    if (thr_prefix != 0xffffffffu && m > (unsigned int)CEM_CAND) {
        // ---- 1b. second level inside the threshold bin: cost bits [CEM_SHIFT-1 : CEM_SHIFT-12] ----
        unsigned int *s_h2 = reinterpret_cast<unsigned int *>(s_keys);            // 4096 u32 = the first 2048 u64 slots
        for (int i = t; i < CEM_BINS; i += NT) s_h2[i] = 0u;
        __syncthreads();
        {
            const unsigned int bin1 = thr_prefix;
            cem_for_each_cost<NT>(costs, n, [&](unsigned int cb, int) {
                if ((cb >> CEM_SHIFT) == bin1) atomicAdd(&s_h2[(cb >> (CEM_SHIFT - 12)) & (CEM_BINS - 1)], 1u);
            });
        }
        __syncthreads();
        const unsigned int bin1 = thr_prefix, below1 = below;
        cem_find_bin<NT>([&](int i) { return s_h2[i]; }, (unsigned int)k - below1, s_hist, s_misc);
        thr_prefix = s_misc[0] == 0xffffffffu ? 0xffffffffu : ((bin1 << 12) | s_misc[0]);
        below = below1 + s_misc[1]; m = s_misc[2];
        pshift = CEM_SHIFT - 12;
        __syncthreads();
    }
    if (thr_prefix == 0xffffffffu || m > (unsigned int)CEM_CAND) {                 // histogram unusable: exact slow path
        block_topk<NT>(costs, n, idx_base, k, s_keys, s_hist, s_misc);
        return;
    }
    if (t == 0) { s_misc[3] = 0u; s_misc[4] = 0u; }
    __syncthreads();
    unsigned long long *s_cand = s_keys + MAX_ELITE;
    // ---- 2. one pass over the costs ----
    cem_for_each_cost<NT>(costs, n, [&](unsigned int cb, int i) {
        const unsigned int b = cb >> pshift;
        if (b <= thr_prefix) {
            const unsigned long long key = ((unsigned long long)cb << 32) | (unsigned int)(idx_base + i);
            if (b < thr_prefix) { const unsigned int slot = atomicAdd(&s_misc[3], 1u); if (slot < MAX_ELITE) s_keys[slot] = key; }
            else { const unsigned int slot = atomicAdd(&s_misc[4], 1u); if (slot < CEM_CAND) s_cand[slot] = key; }
        }
    });
    __syncthreads();
    // ---- 3. refine inside the threshold prefix: radix select over the low (32 + pshift) key bits ----
    const int low_bits = 32 + pshift;
    const unsigned long long low_mask = ~(~0ull << low_bits);
    int need = k - (int)below;                        // 1 <= need <= m
    unsigned long long thr_low = low_mask;            // candidates with (key & low_mask) <= thr_low are selected
    if ((unsigned int)need < m) {
        unsigned long long prefix = 0ull, above = 0ull;   // decided low bits
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
            if ((unsigned int)need == cntd || sh == 0) {          // the bucket is taken whole (always true at the last digit)
                thr_low = prefix | (low_mask & ~above);
                break;
            }
        }
    }
    for (unsigned int c = t; c < m; c += NT) {
        const unsigned long long key = s_cand[c];
        if ((key & low_mask) <= thr_low) {
            const unsigned int slot = atomicAdd(&s_misc[3], 1u);
            if (slot < MAX_ELITE) s_keys[slot] = key;
        }
    }
    __syncthreads();
    const unsigned int got = s_misc[3];               // == k
    __syncthreads();
    for (int i = (int)got + t; i < MAX_ELITE; i += NT) s_keys[i] = ~0ull;
    __syncthreads();
    // ---- 4. sort the winners ----
    for (int size = 2; size <= MAX_ELITE; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = t; i < MAX_ELITE / 2; i += NT) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool up = ((lo & size) == 0);
                const unsigned long long a = s_keys[lo], b = s_keys[hi];
                if ((a > b) == up) { s_keys[lo] = b; s_keys[hi] = a; }
            }
            __syncthreads();
        }
    }
}

}  // namespace qp
