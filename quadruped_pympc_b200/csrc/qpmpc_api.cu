// qpmpc_api.cu -- host side of libqpmpc_b200.so: the C ABI declared in include/qpmpc.h.
// No CPU fallback anywhere: every compute entry point needs a CUDA device and fails with QPMPC_ERR_CUDA otherwise.
#include "../../include/qpmpc.h"
#include "qpmpc_inst.cuh"
#include "qpmpc_quad.cuh"
#include "qpmpc_x2.cuh"
#include "qpmpc_tail.cuh"

#include <dlfcn.h>
#include <unistd.h>
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <mutex>
#include <string>
#include <vector>

using namespace qp;

// samples per launch (all robots) up to which the four-lanes-per-sample kernel is chosen automatically (measured
// crossover on B200)
static const long long QPMPC_QUAD_MAX_SAMPLES = 6144;    // r1g sweep: 4k -> quad 22.6 vs 24.1 us, 10k -> 26.7 vs 24.6 us
// block records per robot from which the reduction tail runs as a second, multi-block kernel (qpmpc_tail.cuh) instead of
// in the tick kernel's last block: 192 records = 24 576 samples per controller (one 10k..18k-sample controller keeps the
// single fused launch -- its tail is short and a second launch would cost more than it saves)
static const int QPMPC_TAIL_MIN_BLOCKS = 192;
// The two-samples-per-thread packed-FP32 kernel (qpmpc_x2.cuh) is opt-in (lanes_per_sample = 2), never chosen
// automatically: measured on B200 (profiles/r2c_*, 128 robots x 10k) it executes 24 % fewer instructions per sample
// than the sample-per-thread kernel (5 316 vs 7 024) but its 208 registers and 104 KB tile leave 8 warps per SM, issue
// slots 45 % busy instead of 73 % -- 1.26x slower.  See DESIGN.md section 3.

// ------------------------------------------------------------------------------------------------------
// error plumbing
// ------------------------------------------------------------------------------------------------------
static thread_local std::string g_err = "";
static int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define CK(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e__ = (call);                                                                         \
        if (e__ != cudaSuccess) return fail(QPMPC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

// ------------------------------------------------------------------------------------------------------
// small kernels: finish stage of a sharded tick, noise / frequency export, FP32 peak probe
// ------------------------------------------------------------------------------------------------------
template <int METHOD>
__global__ void noise_export_kernel(const __grid_constant__ KParams kp, uint32_t seed, uint32_t robot, uint32_t tick,
                                    const float *sigma, float *dst) {
    const int nq = kp.P / 4;
    const size_t total = (size_t)kp.N * nq;
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(w / nq), q = (int)(w - (size_t)i * nq);
        const float4 e = noise_quad<METHOD>(kp, seed, robot, tick, i, q, sigma);
        reinterpret_cast<float4 *>(dst)[w] = e;      // (N,P) row-major, float4-vectorised and coalesced
    }
}

template <int METHOD>
__global__ void freq_export_kernel(const __grid_constant__ KParams kp, const float *in_rec, float *dst) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kp.N; i += gridDim.x * blockDim.x)
        dst[i] = sample_step_freq<METHOD>(kp, in_rec, kp.seed, kp.robot_base, 0, i);
}

// Batched epilogue of the tick: what the callers do next with the GRFs of every robot --
//   nmpc_GRFs.leg = GRFs[3 leg : 3 leg + 3] * current_contact[leg]            (srbd_controller_interface.py:225-230)
//   tau.leg = -J_leg^T nmpc_GRFs.leg,  J_leg = feet_jac.leg[:, legs_qvel_idx.leg] (3 x 3)   (wb_interface.py:372-375)
// read straight from the device-resident result packets of the last tick.  One thread per (robot, leg, joint).
__global__ void stance_torque_kernel(const float *out_packets, int out_stride, const float *jac, const float *contact,
                                     float *grf_masked, float *tau, int R) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R * 12) return;
    const int r = i / 12, lj = i - r * 12, leg = lj / 3, j = lj - leg * 3;
    const float c = contact[r * 4 + leg];
    const float *F = out_packets + (size_t)r * out_stride + OUT_GRF + 3 * leg;
    const float *J = jac + (size_t)r * 36 + leg * 9;                 // row-major 3 x 3: J[k][j] = d foot_k / d q_j
    const float f0 = F[0] * c, f1 = F[1] * c, f2 = F[2] * c;
    grf_masked[i] = (j == 0 ? f0 : (j == 1 ? f1 : f2));
    tau[i] = -((J[0 * 3 + j] * f0 + J[1 * 3 + j] * f1) + J[2 * 3 + j] * f2);
}

// Contact tables of R robots from their gait phase clocks (first half): the horizon look-ahead of the
// reference's PeriodicGaitGenerator.compute_contact_sequence (helpers/periodic_gait_generator.py:93-118) with run()
// (:48-76) unrolled per leg, in float64 like the numpy original (the phase comparisons are discontinuities).
// One thread per (robot, leg).  table: [R][4][H] float 0/1.
struct PggParams {
    double duty, offset[4], dts[QPMPC_MAX_HORIZON];
    int H, R;
};
__global__ void pgg_contact_kernel(const __grid_constant__ PggParams pp, const double *phase, const unsigned char *init,
                                   const double *step_freq, float *table) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pp.R * 4) return;
    const int r = i >> 2, leg = i & 3;
    double t = phase[i];
    bool in_init = init ? (init[i] != 0) : false;
    const double f = step_freq[r];
    for (int n = 0; n < pp.H; ++n) {
        t = __dadd_rn(t, __dmul_rn(pp.dts[n], f));        // self._phase_signal[leg] += dt * new_step_freq   (:53)
        t = t - floor(t);                                  // % 1.0 (exact for t >= 0)                        (:56)
        float c;
        if (in_init) {                                     // :61-67
            if (t <= pp.offset[leg]) c = 1.0f;
            else { in_init = false; c = 1.0f; t = 0.0; }
        } else {
            c = (t < pp.duty) ? 1.0f : 0.0f;               // :71-74
        }
        table[((size_t)r * 4 + leg) * pp.H + n] = c;
    }
}

// Reference footholds of R robots (second half): FootholdReferenceGenerator.
// compute_footholds_reference (helpers/foothold_reference_generator.py:53-157) -- hip-centric heuristic with the desired-
// velocity offset, the moving-average velocity-error compensation and the manual CoM offset -- in float64 like the numpy
// original, one thread per robot (the per-leg work is 4 x a 2 x 2 rotation).  The generator's only state, the deque of
// the last 20 horizontal-frame base velocities (:27,98-99), lives in a device ring per robot.
constexpr int FRG_HIST = 20;                                 // vel_moving_average_length, :16
struct FrgParams {
    double com_height_nominal, hip_height, hip_offset, gravity, com_off_b[3];
    int R;
};
// hist: [R][FRG_HIST][2] ring, meta: [R][2] = {count, next slot}; feet out: [R][4][3]
__global__ void foothold_reference_kernel(const __grid_constant__ FrgParams fp, const double *base_pos, const double *euler,
                                          const double *vel_xy, const double *ref_vel_xy, const double *hips,
                                          const double *lift_off_z, const double *stance_time, double *hist, int *meta,
                                          double *feet) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= fp.R) return;
    const double yaw = euler[3 * r + 2];
    const double cy = cos(yaw), sy = sin(yaw);               // R_W2H = [[cy, sy], [-sy, cy]], :88-90
    const double vx = vel_xy[2 * r], vy = vel_xy[2 * r + 1], rvx = ref_vel_xy[2 * r], rvy = ref_vel_xy[2 * r + 1];
    const double vHx = __dadd_rn(__dmul_rn(cy, vx), __dmul_rn(sy, vy)), vHy = __dadd_rn(__dmul_rn(-sy, vx), __dmul_rn(cy, vy));     // :93
    const double rHx = __dadd_rn(__dmul_rn(cy, rvx), __dmul_rn(sy, rvy)), rHy = __dadd_rn(__dmul_rn(-sy, rvx), __dmul_rn(cy, rvy)); // :94
    // moving average over the deque (append, then mean oldest -> newest), :97-99
    double *hr = hist + (size_t)r * FRG_HIST * 2;
    int count = meta[2 * r], next = meta[2 * r + 1];
    hr[2 * next] = vHx; hr[2 * next + 1] = vHy;
    next = (next + 1) % FRG_HIST;
    if (count < FRG_HIST) ++count;
    meta[2 * r] = count; meta[2 * r + 1] = next;
    double sx = 0.0, sy2 = 0.0;
    const int oldest = (next - count + FRG_HIST) % FRG_HIST;
    for (int k = 0; k < count; ++k) { const int j = (oldest + k) % FRG_HIST; sx = __dadd_rn(sx, hr[2 * j]); sy2 = __dadd_rn(sy2, hr[2 * j + 1]); }
    const double mx = sx / (double)count, my = sy2 / (double)count;
    // desired-velocity offset, clipped to +-1.5 hip heights, :104-106
    const double half_stance = stance_time[r] / 2.0, lim = __dmul_rn(fp.hip_height, 1.5);
    double dx = __dmul_rn(half_stance, rHx), dy = __dmul_rn(half_stance, rHy);
    dx = fmin(fmax(dx, -lim), lim); dy = fmin(fmax(dy, -lim), lim);
    // velocity-error compensation, clipped to +-0.05, :109-112
    const double kk = sqrt(fp.com_height_nominal / fp.gravity);
    double ex = __dmul_rn(kk, __dadd_rn(mx, -rHx)), ey = __dmul_rn(kk, __dadd_rn(my, -rHy));
    ex = ex > 0.05 ? 0.05 : ex; ex = ex < -0.05 ? -0.05 : ex;
    ey = ey > 0.05 ? 0.05 : ey; ey = ey < -0.05 ? -0.05 : ey;
    const double ox = __dadd_rn(dx, ex), oy = __dadd_rn(dy, ey);         // vel_offset + error_compensation, :131
    // manual CoM offset in world frame: R_B2W = Rotation.from_euler("xyz", euler) = Rz Ry Rx, :140-142
    const double roll = euler[3 * r], pitch = euler[3 * r + 1];
    const double cr = cos(roll), sr = sin(roll), cp = cos(pitch), sp = sin(pitch);
    const double bx = fp.com_off_b[0], by = fp.com_off_b[1], bz = fp.com_off_b[2];
    const double cwx = (cy * cp) * bx + (cy * sp * sr - sy * cr) * by + (cy * sp * cr + sy * sr) * bz;
    const double cwy = (sy * cp) * bx + (sy * sp * sr + cy * cr) * by + (sy * sp * cr - cy * sr) * bz;
    const double px = base_pos[3 * r], py = base_pos[3 * r + 1];
    for (int leg = 0; leg < 4; ++leg) {
        const double hx = __dadd_rn(hips[(r * 4 + leg) * 3], -px), hy = __dadd_rn(hips[(r * 4 + leg) * 3 + 1], -py);
        double fx = __dadd_rn(__dmul_rn(cy, hx), __dmul_rn(sy, hy));      // R_W2H @ (hip - base), :118-121
        double fy = __dadd_rn(__dmul_rn(-sy, hx), __dmul_rn(cy, hy));
        fy = (leg == 0 || leg == 2) ? __dadd_rn(fy, fp.hip_offset) : __dadd_rn(fy, -fp.hip_offset);   // :126-129
        fx = __dadd_rn(fx, ox); fy = __dadd_rn(fy, oy);                   // :131
        double wx = __dadd_rn(__dadd_rn(__dmul_rn(cy, fx), __dmul_rn(-sy, fy)), px);   // R_W2H.T @ f + base, :134-137
        double wy = __dadd_rn(__dadd_rn(__dmul_rn(sy, fx), __dmul_rn(cy, fy)), py);
        wx = __dadd_rn(wx, cwx); wy = __dadd_rn(wy, cwy);                 // :143-146
        const double wz = lift_off_z[r * 4 + leg];                        // :150-151
        double *o = feet + ((size_t)r * 4 + leg) * 3;
        o[0] = wx; o[1] = wy; o[2] = wz;
    }
}

__global__ void fp32_peak_kernel(float *out, int iters, float a, float b) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

// ------------------------------------------------------------------------------------------------------
// kernel dispatch tables (parametrisation x adaptive x method)
// ------------------------------------------------------------------------------------------------------
static KernelFn g_tick_table[3][2][3], g_finish_table[3][2][3], g_quad_table[3][2][3], g_tick_h_table[3][2][3], g_x2_table[3][3], g_tail_table[3][2][3];
static void fill_kernel_tables_once() {
    qp_inst_tick_zero(g_tick_table[ZERO_ORDER], g_finish_table[ZERO_ORDER], g_tick_h_table[ZERO_ORDER], g_tail_table[ZERO_ORDER]);
    qp_inst_tick_linear(g_tick_table[LINEAR], g_finish_table[LINEAR], g_tick_h_table[LINEAR], g_tail_table[LINEAR]);
    qp_inst_tick_cubic(g_tick_table[CUBIC], g_finish_table[CUBIC], g_tick_h_table[CUBIC], g_tail_table[CUBIC]);
    qp_inst_quad_zero(g_quad_table[ZERO_ORDER]);
    qp_inst_quad_linear(g_quad_table[LINEAR]);
    qp_inst_quad_cubic(g_quad_table[CUBIC]);
    qp_inst_x2_zero(g_x2_table[ZERO_ORDER]);
    qp_inst_x2_linear(g_x2_table[LINEAR]);
    qp_inst_x2_cubic(g_x2_table[CUBIC]);
}
static void fill_kernel_tables() {             // handles may be created concurrently (ROS2 thread mode)
    static std::once_flag once;
    std::call_once(once, fill_kernel_tables_once);
}

// ------------------------------------------------------------------------------------------------------
// NCCL through dlopen (no link-time dependency; the torch-bundled libnccl.so.2 is passed in by the host code)
// ------------------------------------------------------------------------------------------------------
struct NcclId { char internal[128]; };
typedef void *NcclComm;
struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(NcclId *) = nullptr;
    int (*CommInitRank)(NcclComm *, int, NcclId, int) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, NcclComm, cudaStream_t) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*CommAbort)(NcclComm) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
static int load_nccl(const char *path) {
    if (g_nccl.lib) return QPMPC_OK;
    void *lib = dlopen(path && path[0] ? path : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return fail(QPMPC_ERR_NCCL, "dlopen(%s) failed: %s", path ? path : "libnccl.so.2", dlerror());
    g_nccl.GetUniqueId = (int (*)(NcclId *))dlsym(lib, "ncclGetUniqueId");
    g_nccl.CommInitRank = (int (*)(NcclComm *, int, NcclId, int))dlsym(lib, "ncclCommInitRank");
    g_nccl.AllGather = (int (*)(const void *, void *, size_t, int, NcclComm, cudaStream_t))dlsym(lib, "ncclAllGather");
    g_nccl.CommDestroy = (int (*)(NcclComm))dlsym(lib, "ncclCommDestroy");
    g_nccl.CommAbort = (int (*)(NcclComm))dlsym(lib, "ncclCommAbort");
    g_nccl.GetErrorString = (const char *(*)(int))dlsym(lib, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllGather || !g_nccl.CommDestroy)
        return fail(QPMPC_ERR_NCCL, "libnccl is missing required symbols");
    g_nccl.lib = lib;
    return QPMPC_OK;
}

// ------------------------------------------------------------------------------------------------------
// the handle
// ------------------------------------------------------------------------------------------------------
struct qpmpc_handle {
    qpmpc_config cfg;
    KParams kp;                       // template: constants + device pointers for MODE_TICK
    int R, N, H, P, n_local;
    size_t smem_bytes, finish_smem_bytes, tail_smem_bytes = 0;
    cudaStream_t stream = nullptr;
    float *d_in = nullptr, *d_out = nullptr, *d_costs = nullptr, *d_blk = nullptr, *d_send = nullptr, *d_recv = nullptr;
    float *d_noise = nullptr, *d_freqs = nullptr, *d_flush = nullptr, *d_sigma = nullptr;
    unsigned int *d_counters = nullptr, *d_cem_hist = nullptr;
    float *h_in = nullptr, *h_out = nullptr, *h_rec = nullptr;   // pinned (mapped when io_mode == MAPPED)
    float *kin = nullptr, *kout = nullptr;                       // what the kernel dereferences for in / out
    size_t noise_capacity = 0, freqs_capacity = 0;
    cudaGraphExec_t graph_exec = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // chunk-pipelined batched tick (pipelined_batch): second compute stream, upload / download streams, per-chunk events
    cudaStream_t stream2 = nullptr, stream_in = nullptr, stream_out = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_chunk[8] = {}, ev_in[8] = {}, ev_k[8] = {};
    uint64_t launches = 0;
    bool resident_uploaded = false, partial_done = false;
    NcclComm comm = nullptr;
    KernelFn tick_fn = nullptr, finish_fn = nullptr, tail_fn = nullptr;
    int tail_nb = 0;                  // > 0: two-stage launches (tick kernel + tail_kernel with tail_nb blocks per robot)
    float *d_part = nullptr;
    unsigned int *d_tail_cnt = nullptr;
    unsigned long long *d_tail_keys = nullptr;
    int threads = TB;                 // threads per block of tick_fn (finish_fn always runs TB)
    bool helpers = false;             // tick_fn = tick_kernel_h (sample per thread + noise helper threads)
    bool direct = false;              // QPMPC_IO_DIRECT in effect
    float *h_dout = nullptr, *d_dout = nullptr;   // mapped pinned result packet as {value, seq} units (host / device alias)
    float *h_dec = nullptr;                       // the same packet decoded to plain floats (direct_tick)
    bool host_times_on = false;                   // qpmpc_debug_host_times: CLOCK_MONOTONIC ns at entry / launched / stamp seen / unpacked
    uint64_t host_ns[4] = {0, 0, 0, 0};
    float shift_step = 0.0f;                      // stateful controller: shift_solution step (1 / mpc_frequency), 0 = off
    mutable const float *last_out = nullptr;      // pinned host packets of the last unpacked tick (qpmpc_publish_shm)
    unsigned int seq = 0;
    int lanes = 1;                    // 1: sample per thread ; 4: four lanes per sample (qpmpc_quad.cuh)
    unsigned long long *d_phase = nullptr;
    KParams partial_kp;               // noise / frequency injection state carried from shard_partial to shard_finish
    // peer exchange (fused collective of a sample-sharded controller, qpmpc_tick.cuh peer_exchange)
    float *d_mbox = nullptr;          // this rank's mailbox (cudaMalloc: exportable with cudaIpcGetMemHandle)
    float *d_sink = nullptr;
    float *d_contact = nullptr;       // [R][4][H] contact tables generated on the device (qpmpc_pgg_contact_sequence)
    void *d_pgg = nullptr;            // its inputs: R x 4 phases (f64), R step frequencies (f64), R x 4 init flags
    bool contacts_on_device = false;  // armed by qpmpc_pgg_contact_sequence, consumed by the next tick / step
    void *d_frg = nullptr;            // foothold reference: inputs, velocity-history rings, outputs (qpmpc_footholds_reference)
    float *d_epi = nullptr;           // stance-torque epilogue scratch: [R][36 jac + 4 contact + 12 grf + 12 tau]
    float *d_ctrl = nullptr;          // stateful controller records (qpmpc_step): [R][ctrl_stride]
    int ctrl_stride = 0;          // loopback test hook: where the records pushed to absent peers go
    float *peer_mbox[MAX_PEERS] = {}; // every rank's mailbox as seen from this device
    bool peer_ipc[MAX_PEERS] = {};    // opened with cudaIpcOpenMemHandle (to be closed)
    bool peer_attached = false;
    bool peer_timing = false;         // qpmpc_debug_peer_wait
    unsigned long long peer_timeout_ns = 2000000000ull;   // qpmpc_shard_peer_timeout
    uint64_t peer_wait_ns[2] = {0, 0};
};

static size_t round4(size_t x) { return (x + 3) & ~(size_t)3; }

static int chunk_index_host(const qpmpc_config *c, float stepf, const float *bounds) {
    int idx = 0;
    for (int i = 0; i <= c->num_splines; ++i)
        if (stepf >= bounds[i] && i > idx) idx = i;
    return idx;
}

// 3x3 inverse by cofactors in float32, same expression as M:67-91
static void inv3f(const float A[9], float out[9]) {
    const float a11 = A[0], a12 = A[1], a13 = A[2], a21 = A[3], a22 = A[4], a23 = A[5], a31 = A[6], a32 = A[7], a33 = A[8];
    const float DET = a11 * (a33 * a22 - a32 * a23) - a21 * (a33 * a12 - a32 * a13) + a31 * (a23 * a12 - a22 * a13);
    out[0] = (a33 * a22 - a32 * a23) / DET;  out[1] = -(a33 * a12 - a32 * a13) / DET; out[2] = (a23 * a12 - a22 * a13) / DET;
    out[3] = -(a33 * a21 - a31 * a23) / DET; out[4] = (a33 * a11 - a31 * a13) / DET;  out[5] = -(a23 * a11 - a21 * a13) / DET;
    out[6] = (a32 * a21 - a31 * a22) / DET;  out[7] = -(a32 * a11 - a31 * a12) / DET; out[8] = (a22 * a11 - a21 * a12) / DET;
}

static int validate_config(const qpmpc_config *c) {
    if (!c) return fail(QPMPC_ERR_INVALID, "config is NULL");
    if (c->abi_version != QPMPC_ABI_VERSION) return fail(QPMPC_ERR_INVALID, "abi_version %d != %d", c->abi_version, QPMPC_ABI_VERSION);
    if (c->num_samples < 1) return fail(QPMPC_ERR_INVALID, "num_samples must be >= 1");
    if (c->horizon < 1 || c->horizon > QPMPC_MAX_HORIZON) return fail(QPMPC_ERR_INVALID, "horizon must be in [1,%d]", QPMPC_MAX_HORIZON);
    if (c->parametrization < 0 || c->parametrization > 2) return fail(QPMPC_ERR_INVALID, "unknown control parametrization %d", c->parametrization);
    if (c->method < 0 || c->method > 2) return fail(QPMPC_ERR_INVALID, "Error: sampling method not recognized (%d)", c->method);  // S:107-110
    if (c->parametrization != QPMPC_ZERO_ORDER && (c->num_splines < 1 || c->num_splines > QPMPC_MAX_SPLINES))
        return fail(QPMPC_ERR_INVALID, "num_splines must be in [1,%d]", QPMPC_MAX_SPLINES);
    if (c->num_robots < 1) return fail(QPMPC_ERR_INVALID, "num_robots must be >= 1");
    if (c->robot_base < 0) return fail(QPMPC_ERR_INVALID, "robot_base must be >= 0");
    // the global robot id shares Philox counter word 0 with the quad index (bits 0-7) and the stream flag (bit 8)
    if ((long long)c->robot_base + c->num_robots > (1ll << 23))
        return fail(QPMPC_ERR_INVALID, "robot_base + num_robots = %lld exceeds 2^23 (robot id field of the Philox counter)", (long long)c->robot_base + c->num_robots);
    if (c->shard_world < 1 || c->shard_rank < 0 || c->shard_rank >= c->shard_world) return fail(QPMPC_ERR_INVALID, "bad shard rank/world");
    if (!(c->mass > 0.f)) return fail(QPMPC_ERR_INVALID, "mass must be positive");
    if (c->elite_k < 0 || c->elite_k > QPMPC_MAX_ELITE) return fail(QPMPC_ERR_INVALID, "elite_k must be in [0,%d]", QPMPC_MAX_ELITE);
    return QPMPC_OK;
}

extern "C" int qpmpc_abi_version(void) { return QPMPC_ABI_VERSION; }
extern "C" const char *qpmpc_last_error(void) { return g_err.c_str(); }

extern "C" int qpmpc_device_count(int *count) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { if (count) *count = 0; return fail(QPMPC_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e)); }
    if (count) *count = n;
    return QPMPC_OK;
}

extern "C" void qpmpc_config_defaults(qpmpc_config *c) {
    memset(c, 0, sizeof(*c));
    c->abi_version = QPMPC_ABI_VERSION;
    c->num_samples = 10000; c->horizon = 12; c->num_splines = 2;                 // config.py:92,186,187
    c->parametrization = QPMPC_CUBIC_SPLINE; c->method = QPMPC_RANDOM_SAMPLING;  // config.py:184-185
    c->adaptive = 0; c->num_robots = 1; c->device = 0; c->io_mode = QPMPC_IO_DIRECT; c->use_graph = 1;
    c->shard_rank = 0; c->shard_world = 1; c->elite_k = 0; c->seed = 42u;        // S:167
    c->mass = 15.019f;                                                            // go2, config.py:25-29
    const float I[9] = {1.58460467e-01f, 1.21660000e-04f, -1.55444692e-02f, 1.21660000e-04f, 4.68645637e-01f,
                        -3.12000000e-05f, -1.55444692e-02f, -3.12000000e-05f, 5.24474661e-01f};
    memcpy(c->inertia, I, sizeof(I));
    for (int i = 0; i < QPMPC_MAX_HORIZON; ++i) c->dts[i] = 0.02f;               // config.py:93
    c->mu = 0.5f; c->grf_min = 0.f; c->grf_max = (float)(15.019 * 9.81);         // config.py:96-98
    const float q[12] = {0, 0, 1500, 200, 200, 200, 500, 500, 0, 20, 20, 50};    // S:118-130
    memcpy(c->q_diag, q, sizeof(q));
    c->sigma_random_sampling[0] = 0.2f; c->sigma_random_sampling[1] = 3.f; c->sigma_random_sampling[2] = 10.f;
    c->sigma_mppi = 3.f;
    c->step_freq_available[0] = 1.4f; c->step_freq_available[1] = 2.0f; c->step_freq_available[2] = 2.4f;
    c->pgg_duty_factor = 0.65f;                                                   // A:179
    c->robot_base = 0; c->lanes_per_sample = 0;
}

static void free_handle(qpmpc_handle *h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    if (h->comm) {
        // Every tick is synchronous, so nothing is in flight on this communicator.  ncclCommAbort releases it locally;
        // ncclCommDestroy's finalize step was seen to hang when NCCL-exchange handles are created and destroyed
        // repeatedly in a process that also runs torch.distributed's own communicator (tools/multi_gpu_check.py).
        if (h->stream) cudaStreamSynchronize(h->stream);
        if (g_nccl.CommAbort) g_nccl.CommAbort(h->comm);
        else if (g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    }
    for (int w = 0; w < MAX_PEERS; ++w)
        if (h->peer_ipc[w] && h->peer_mbox[w]) cudaIpcCloseMemHandle(h->peer_mbox[w]);
    cudaFree(h->d_part); cudaFree(h->d_tail_cnt); cudaFree(h->d_tail_keys);
    cudaFree(h->d_mbox); cudaFree(h->d_sink); cudaFree(h->d_ctrl); cudaFree(h->d_epi); cudaFree(h->d_contact); cudaFree(h->d_pgg); cudaFree(h->d_frg);
    if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
    if (h->stream2) cudaStreamDestroy(h->stream2);
    if (h->stream_in) cudaStreamDestroy(h->stream_in);
    if (h->stream_out) cudaStreamDestroy(h->stream_out);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    for (int c = 0; c < 8; ++c) {
        if (h->ev_chunk[c]) cudaEventDestroy(h->ev_chunk[c]);
        if (h->ev_in[c]) cudaEventDestroy(h->ev_in[c]);
        if (h->ev_k[c]) cudaEventDestroy(h->ev_k[c]);
    }
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    cudaFree(h->d_in); cudaFree(h->d_out); cudaFree(h->d_costs); cudaFree(h->d_blk); cudaFree(h->d_send); cudaFree(h->d_recv);
    cudaFree(h->d_noise); cudaFree(h->d_freqs); cudaFree(h->d_flush); cudaFree(h->d_counters); cudaFree(h->d_cem_hist); cudaFree(h->d_sigma); cudaFree(h->d_phase);
    if (h->h_in) cudaFreeHost(h->h_in);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->h_rec) cudaFreeHost(h->h_rec);
    if (h->h_dout) cudaFreeHost(h->h_dout);
    free(h->h_dec);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

extern "C" void qpmpc_destroy(qpmpc_handle *h) { free_handle(h); }

static int samples_per_tile(int lanes) { return lanes == 4 ? QS : (lanes == 2 ? XS : TB); }
static void plan_blocks(int n_local, int lanes, int *iters, int *nblk) {
    const int per_tile = samples_per_tile(lanes);      // samples per tile
    const int tiles = (n_local + per_tile - 1) / per_tile;
    const int cap = lanes == 2 ? 148 * 2 * 2 : 148 * 4 * 2;   // two full waves of resident CTAs on 148 SMs
    int it = (tiles + cap - 1) / cap;
    if (it < 1) it = 1;
    *iters = it;
    *nblk = (tiles + it - 1) / it;
}

extern "C" int qpmpc_create(const qpmpc_config *cfg, qpmpc_handle **out) {
    if (!out) return fail(QPMPC_ERR_INVALID, "out handle pointer is NULL");
    *out = nullptr;
    int rc = validate_config(cfg);
    if (rc) return rc;
    int ndev = 0;
    rc = qpmpc_device_count(&ndev);
    if (rc) return rc;
    if (ndev == 0) return fail(QPMPC_ERR_CUDA, "no CUDA device: libqpmpc_b200 has no CPU fallback");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(QPMPC_ERR_INVALID, "device %d out of range (%d devices)", cfg->device, ndev);
    CK(cudaSetDevice(cfg->device));

    qpmpc_handle *h = new qpmpc_handle();
    h->cfg = *cfg;
    const int H = cfg->horizon, ns = cfg->parametrization == QPMPC_ZERO_ORDER ? 1 : cfg->num_splines;
    h->H = H; h->N = cfg->num_samples; h->R = cfg->num_robots;
    int P_leg;
    if (cfg->parametrization == QPMPC_LINEAR_SPLINE) P_leg = (ns + 1) * 3;          // S:56
    else if (cfg->parametrization == QPMPC_CUBIC_SPLINE) P_leg = 4 * 3 * ns;        // S:71
    else P_leg = H * 3;                                                              // S:84
    h->P = 4 * P_leg;
    if (h->P > MAXV * TB) { delete h; return fail(QPMPC_ERR_INVALID, "P = %d control parameters exceed the kernel limit %d", 4 * P_leg, MAXV * TB); }
    // contiguous sample slice of this shard (global index = base + local keeps argmin / elite indices global)
    const int W = cfg->shard_world, rk = cfg->shard_rank;
    const int per = (h->N + W - 1) / W;
    const int base = std::min(per * rk, h->N);
    h->n_local = std::min(per, h->N - base);
    if (h->n_local < 1) { delete h; return fail(QPMPC_ERR_INVALID, "shard %d of %d owns no samples (N=%d)", rk, W, h->N); }
    int elite_k = cfg->elite_k > 0 ? cfg->elite_k : (cfg->adaptive ? 200 : 10);     // A:1115 | S:1075
    if (cfg->method != QPMPC_CEM_MPPI) elite_k = 0;
    if (W * elite_k > SCRATCH_KEYS) { delete h; return fail(QPMPC_ERR_INVALID, "shard_world * elite_k = %d exceeds %d", W * elite_k, SCRATCH_KEYS); }

    KParams &kp = h->kp;
    memset(&kp, 0, sizeof(kp));
    kp.N = h->N; kp.n_local = h->n_local; kp.sample_base = base;
    kp.H = H; kp.ns = ns; kp.P = h->P; kp.P_leg = P_leg;
    // Kernel choice: the four-lanes-per-sample kernel wins while the GPU is under-filled (few samples in flight: one
    // controller at 10k samples), the sample-per-thread kernel executes ~1.8x fewer instructions per sample and wins
    // once its warps fill the machine.  lanes_per_sample = 0 picks by the number of samples of one launch.
    const int lps = cfg->lanes_per_sample;
    if (lps != 0 && lps != 1 && lps != 2 && lps != 4 && lps != 14) { delete h; return fail(QPMPC_ERR_INVALID, "lanes_per_sample must be 0 (auto), 1, 2, 4 or 14"); }
    h->lanes = lps == 4 ? 4 : (lps != 0 ? 1 : ((long long)h->n_local * h->R <= QPMPC_QUAD_MAX_SAMPLES ? 4 : 1));
    // two samples per thread in packed FP32 (tick_kernel_x2): fixed gait only, and only while two 256-sample tiles fit one
    // SM (otherwise the plain kernel keeps more warps resident).  lanes_per_sample = 2 asks for it; requests that do not
    // qualify fall back to the sample-per-thread kernel.
    {
        const size_t x2_in_stride = round4((size_t)IN_CONTACT + 4 * H + 2 * (size_t)h->P);
        const bool x2_ok = !cfg->adaptive && smem_floats((int)x2_in_stride, h->P, XS, XTS) * sizeof(float) <= 113 * 1024;
        if (x2_ok && lps == 2) h->lanes = 2;
    }
    plan_blocks(h->n_local, h->lanes, &kp.iters_rec, &kp.nrec);          // record structure: a function of n_local only
    {   // Batched launches with many waves of blocks let a block walk k record groups: the packet staging, the yaw sine
        // and the ticket are paid once per block (3-5 % of a one-tile block's life).  The RECORDS stay those of k = 1, so
        // results do not depend on it (tests/test_gpu_batched.py::test_result_does_not_depend_on_the_partition).
        const long long blocks = (long long)h->R * kp.nrec;
        const int k = h->R > 1 ? (blocks / 4 >= 2 * 592 ? 4 : (blocks / 2 >= 2 * 592 ? 2 : 1)) : 1;
        const int tiles = (h->n_local + samples_per_tile(h->lanes) - 1) / samples_per_tile(h->lanes);
        kp.iters = kp.iters_rec * k;
        kp.nblk = (tiles + kp.iters - 1) / kp.iters;
    }
    // sample per thread + noise helper threads (tick_kernel_h): one controller, one tile per block, no more blocks than SMs
    h->helpers = (h->lanes == 1 && (lps == 0 || lps == 14) && h->R == 1 && kp.iters == 1 && kp.nblk <= 148);
    kp.elite_k = elite_k;
    kp.mode = MODE_TICK; kp.noise_mode = NOISE_PHILOX; kp.freq_injected = 0;
    kp.fused_finish = (W == 1); kp.world = W; kp.carry_best = 0;
    kp.off_best = IN_CONTACT + 4 * H;
    kp.off_sigma = kp.off_best + h->P;
    kp.in_stride = (int)round4(kp.off_sigma + h->P);
    kp.off_out_sigma = OUT_BEST + h->P;
    kp.off_out_elite = OUT_BEST + 2 * h->P;
    kp.out_stride = (int)round4(kp.off_out_elite + elite_k);
    kp.blk_rec_stride = (int)round4(REC_V + h->P);
    kp.shard_rec_stride = (int)round4(REC_V + h->P + 2 * elite_k);
    kp.inv_m = 1.0f / cfg->mass;                                                    // M:118
    kp.mg = (float)((double)cfg->mass * 9.81);                                      // S:380 (folded in double, then cast)
    kp.mu = cfg->mu; kp.fz_min = cfg->grf_min; kp.fz_max = cfg->grf_max;
    memcpy(kp.I, cfg->inertia, 9 * sizeof(float));
    inv3f(cfg->inertia, kp.Iinv);                                                   // M:56
    memcpy(kp.qd, cfg->q_diag, sizeof(kp.qd));
    kp.q_xyyaw = (cfg->q_diag[0] != 0.f || cfg->q_diag[1] != 0.f || cfg->q_diag[8] != 0.f) ? 1 : 0;
    kp.bm_bits[0] = 0x3F800000u; kp.bm_bits[1] = 0x40000000u;
    { const int qi[8] = {3, 4, 2, 5, 6, 7, 9, 10}; for (int k = 0; k < 8; ++k) kp.qpair[k] = cfg->q_diag[qi[k]]; }
    kp.mpc_dt = cfg->dts[H - 1];                                                    // A:179: pgg mpc_dt = self.dt
    kp.duty = cfg->pgg_duty_factor;
    memcpy(kp.sigma_rs, cfg->sigma_random_sampling, sizeof(kp.sigma_rs));
    kp.sigma_mppi = cfg->sigma_mppi;
    memcpy(kp.freq_avail, cfg->step_freq_available, sizeof(kp.freq_avail));
    kp.inv_ns = 1.0f / (float)ns;
    const float denom_fixed = (float)((double)H / (double)ns);                     // horizon_leg / num_spline, S:191
    kp.denom_fixed_inv = 1.0f / denom_fixed;
    for (int i = 0; i <= ns; ++i) kp.bounds[i] = (float)((double)H * i / (double)ns);   // linspace(0,H,ns+1), S:187
    for (int n = 0; n < H; ++n) {
        kp.dt[n] = cfg->dts[n];
        const float stepf = (float)n;
        const int idx = cfg->parametrization == QPMPC_ZERO_ORDER ? 0 : chunk_index_host(cfg, stepf, kp.bounds);
        kp.sp_index[n] = (unsigned char)idx;
        const float q = stepf / denom_fixed - (float)idx;                           // S:191-194
        if (cfg->parametrization == QPMPC_CUBIC_SPLINE) {                           // S:222-225
            const double a = 2.0f * q * q * q - 3.0f * q * q + 1.0f, b = (q * q * q - 2.0f * q * q + q) * 1.0f;
            const double c = -2.0f * q * q * q + 3.0f * q * q, d = (q * q * q - q * q) * 1.0f;
            // f = a p1 + b phi + c p2 + d phi_next with phi = (p2 - p0) / 2, phi_next = (p3 - p1) / 2 (S:227-255),
            // collected per knot: the kernels evaluate f = sum_i w_i p_i on the four knots of the chunk directly
            kp.basis[n][0] = (float)(-0.5 * b);
            kp.basis[n][1] = (float)(a - 0.5 * d);
            kp.basis[n][2] = (float)(c + 0.5 * b);
            kp.basis[n][3] = (float)(0.5 * d);
        } else if (cfg->parametrization == QPMPC_LINEAR_SPLINE) {                   // S:197
            kp.basis[n][0] = 1.0f - q; kp.basis[n][1] = q;
        }
    }
    h->threads = h->lanes == 4 ? QTB : (h->lanes == 2 ? XTB : (h->helpers ? TB * NOISE_HELPERS : TB));
    h->finish_smem_bytes = smem_floats(kp.in_stride, h->P) * sizeof(float);
    if (h->lanes == 4)      // the CEM elite refit of finish() stages up to `threads` noise rows in the tile
        h->smem_bytes = smem_floats(kp.in_stride, h->P, QTB, cfg->method == QPMPC_CEM_MPPI ? QTB + 1 : QTS) * sizeof(float);
    else if (h->lanes == 2)
        h->smem_bytes = smem_floats(kp.in_stride, h->P, XS, XTS) * sizeof(float);
    else
        h->smem_bytes = h->finish_smem_bytes;
    if (h->smem_bytes > 227 * 1024) { delete h; return fail(QPMPC_ERR_INVALID, "tile of %d knots needs %zu B shared memory (> 227 KB)", h->P, h->smem_bytes); }

    fill_kernel_tables();
    h->tick_fn = h->lanes == 2 ? g_x2_table[cfg->parametrization][cfg->method]
                               : (h->lanes == 4 ? g_quad_table : (h->helpers ? g_tick_h_table : g_tick_table))[cfg->parametrization][cfg->adaptive ? 1 : 0][cfg->method];
    h->finish_fn = g_finish_table[cfg->parametrization][cfg->adaptive ? 1 : 0][cfg->method];
    h->tail_fn = g_tail_table[cfg->parametrization][cfg->adaptive ? 1 : 0][cfg->method];
    if (kp.nrec >= QPMPC_TAIL_MIN_BLOCKS) h->tail_nb = std::max(1, std::min(TAIL_MAX_NB, 296 / h->R));
#define CKH(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { free_handle(h); return fail(QPMPC_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__)); } } while (0)
    CKH(cudaFuncSetAttribute((const void *)h->tick_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_bytes));
    CKH(cudaFuncSetAttribute((const void *)h->finish_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->finish_smem_bytes));
    h->tail_smem_bytes = smem_floats(kp.in_stride, h->P, TAIL_TB, TAIL_TB + 1) * sizeof(float);
    if (h->tail_smem_bytes > 227 * 1024) h->tail_nb = 0;                  // (P > 200 knots: keep the fused path)
    CKH(cudaFuncSetAttribute((const void *)h->tail_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->tail_smem_bytes));
    CKH(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CKH(cudaEventCreate(&h->ev0));
    CKH(cudaEventCreate(&h->ev1));
    const size_t R = h->R;
    const unsigned host_flags = cfg->io_mode == QPMPC_IO_MAPPED ? cudaHostAllocMapped : cudaHostAllocDefault;
    CKH(cudaHostAlloc((void **)&h->h_in, R * kp.in_stride * sizeof(float), host_flags));
    CKH(cudaHostAlloc((void **)&h->h_out, R * kp.out_stride * sizeof(float), host_flags));
    CKH(cudaHostAlloc((void **)&h->h_rec, (size_t)W * R * kp.shard_rec_stride * sizeof(float), cudaHostAllocDefault));
    memset(h->h_in, 0, R * kp.in_stride * sizeof(float));
    memset(h->h_out, 0, R * kp.out_stride * sizeof(float));
    CKH(cudaMalloc((void **)&h->d_in, R * kp.in_stride * sizeof(float)));
    CKH(cudaMalloc((void **)&h->d_out, R * kp.out_stride * sizeof(float)));
    CKH(cudaMalloc((void **)&h->d_costs, R * (size_t)h->N * sizeof(float)));
    CKH(cudaMalloc((void **)&h->d_blk, R * (size_t)std::max(kp.nrec, ((h->N + samples_per_tile(h->lanes) - 1) / samples_per_tile(h->lanes))) * kp.blk_rec_stride * sizeof(float)));
    CKH(cudaMalloc((void **)&h->d_send, R * kp.shard_rec_stride * sizeof(float)));
    CKH(cudaMalloc((void **)&h->d_recv, (size_t)W * R * kp.shard_rec_stride * sizeof(float)));
    CKH(cudaMalloc((void **)&h->d_counters, R * sizeof(unsigned int)));
    CKH(cudaMalloc((void **)&h->d_sigma, (size_t)h->P * sizeof(float)));
    CKH(cudaMemset(h->d_counters, 0, R * sizeof(unsigned int)));
    if (h->tail_nb > 0) {
        CKH(cudaMalloc((void **)&h->d_part, R * (size_t)TAIL_NPART * kp.blk_rec_stride * sizeof(float)));
        CKH(cudaMalloc((void **)&h->d_tail_cnt, R * tail_cnt_stride() * sizeof(unsigned int)));
        {
            std::vector<unsigned int> init(R * tail_cnt_stride(), 0u);
            for (size_t r = 0; r < R; ++r) init[r * tail_cnt_stride() + 4] = 0xffffffffu;     // empty bin window
            CKH(cudaMemcpy(h->d_tail_cnt, init.data(), init.size() * sizeof(unsigned int), cudaMemcpyHostToDevice));
        }
        if (cfg->method == QPMPC_CEM_MPPI) CKH(cudaMalloc((void **)&h->d_tail_keys, R * 2 * (size_t)h->n_local * sizeof(unsigned long long)));
        kp.tail_nb = h->tail_nb; kp.part_recs = h->d_part; kp.tail_cnt = h->d_tail_cnt; kp.tail_keys = h->d_tail_keys;
    }
    if (cfg->method == QPMPC_CEM_MPPI) {       // per-robot cost histogram: zero between ticks (the last block re-zeroes it)
        const size_t bins = h->tail_nb > 0 ? CEM_BINS_FINE : CEM_BINS;      // two-stage launches use the fine histogram
        CKH(cudaMalloc((void **)&h->d_cem_hist, R * bins * sizeof(unsigned int)));
        CKH(cudaMemset(h->d_cem_hist, 0, R * bins * sizeof(unsigned int)));
    }
    CKH(cudaMemset(h->d_in, 0, R * kp.in_stride * sizeof(float)));
    CKH(cudaMemset(h->d_out, 0, R * kp.out_stride * sizeof(float)));
    if (cfg->io_mode == QPMPC_IO_MAPPED) {
        CKH(cudaHostGetDevicePointer((void **)&h->kin, h->h_in, 0));
        CKH(cudaHostGetDevicePointer((void **)&h->kout, h->h_out, 0));
    } else {
        h->kin = h->d_in; h->kout = h->d_out;
    }
    kp.in = h->kin; kp.out = h->kout; kp.costs = h->d_costs; kp.blk_recs = h->d_blk;
    kp.shard_send = h->d_send; kp.shard_recv = h->d_recv; kp.counters = h->d_counters; kp.cem_hist = h->d_cem_hist;
    kp.noise = nullptr; kp.freqs = nullptr;
    kp.seed = cfg->seed; kp.robot_base = (uint32_t)cfg->robot_base;
    h->direct = (cfg->io_mode == QPMPC_IO_DIRECT && h->R == 1 && kp.in_stride <= PACKET_MAX);   // W > 1: peer-exchange ticks only
    if (h->direct) {
        CKH(cudaHostAlloc((void **)&h->h_dout, (size_t)kp.out_stride * 2 * sizeof(float), cudaHostAllocMapped));
        memset(h->h_dout, 0, (size_t)kp.out_stride * 2 * sizeof(float));      // flag 0 is never a valid seq (next_seq)
        h->h_dec = (float *)calloc((size_t)kp.out_stride, sizeof(float));
        if (!h->h_dec) { free_handle(h); return fail(QPMPC_ERR_INVALID, "out of host memory"); }
        CKH(cudaHostGetDevicePointer((void **)&h->d_dout, h->h_dout, 0));
    }
    *out = h;
    return QPMPC_OK;
}

extern "C" int qpmpc_num_parameters(const qpmpc_handle *h) { return h ? h->P : QPMPC_ERR_INVALID; }
extern "C" int qpmpc_num_local_samples(const qpmpc_handle *h) { return h ? h->n_local : QPMPC_ERR_INVALID; }
extern "C" int qpmpc_shard_record_floats(const qpmpc_handle *h) { return h ? h->R * h->kp.shard_rec_stride : QPMPC_ERR_INVALID; }

// ------------------------------------------------------------------------------------------------------
// packing / unpacking of the per-robot records
// ------------------------------------------------------------------------------------------------------
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static int pack_one(qpmpc_handle *h, int r, const qpmpc_tick_in *in, bool explicit_params, bool stateful = false) {
    const KParams &kp = h->kp;
    float *rec = h->h_in + (size_t)r * kp.in_stride;
    if (!in->state || !in->reference) return fail(QPMPC_ERR_INVALID, "state / reference must not be NULL");
    if (!in->contact_sequence && !h->contacts_on_device) return fail(QPMPC_ERR_INVALID, "contact_sequence is NULL and no device-generated table is armed (qpmpc_pgg_contact_sequence)");
    if (in->contact_sequence && in->contact_ld < h->H) return fail(QPMPC_ERR_INVALID, "contact_ld %d < horizon %d", in->contact_ld, h->H);
    memcpy(rec + IN_STATE, in->state, 24 * sizeof(float));
    memcpy(rec + IN_REF, in->reference, 24 * sizeof(float));
    for (int l = 0; l < 4; ++l) rec[IN_PHASE + l] = in->phase_signal ? in->phase_signal[l] : 0.f;
    if (h->cfg.adaptive && !in->phase_signal) return fail(QPMPC_ERR_INVALID, "phase_signal is required by the gait-adaptive controller");
    rec[IN_NOMFREQ] = in->nominal_step_freq;
    rec[IN_OPTSWING] = u2f((uint32_t)in->optimize_swing);
    // Philox counter word 3 (the seed) travels in the packet and is read from there by the kernels: a graph-captured
    // launch freezes kernel-parameter space at capture time, the packet is uploaded every tick.
    if (stateful) { /* keys are ignored: the record holds the counter, the seed is the handle's */ }
    else if (r == 0) h->kp.seed = in->key[0];
    else if (in->key[0] != h->kp.seed) return fail(QPMPC_ERR_INVALID, "all robots of one batch must share key[0] (robot %d differs); streams differ by robot id", r);
    rec[IN_SEED] = u2f(stateful ? h->cfg.seed : in->key[0]);
    rec[IN_TICK] = u2f(in->key[1]);
    rec[IN_ROBOT] = u2f((uint32_t)(h->cfg.robot_base + r));   // global robot id = Philox key word 1
    for (int l = 0; l < 4 && in->contact_sequence; ++l)
        for (int n = 0; n < h->H; ++n) {
            const float c = in->contact_sequence[(size_t)l * in->contact_ld + n];
            if (c != 0.0f && c != 1.0f) return fail(QPMPC_ERR_INVALID, "contact_sequence[%d][%d] = %g is not 0 or 1", l, n, (double)c);
            rec[IN_CONTACT + l * h->H + n] = c;
        }
    float *best = rec + kp.off_best, *sigma = rec + kp.off_sigma;
    if (explicit_params || !in->best) memset(best, 0, h->P * sizeof(float));
    else memcpy(best, in->best, h->P * sizeof(float));
    if (stateful) return QPMPC_OK;                    // best / sigma / key live in the device-resident record
    if (h->cfg.method == QPMPC_CEM_MPPI && !explicit_params) {
        if (!in->sigma) return fail(QPMPC_ERR_INVALID, "sigma is required by cem_mppi");
        memcpy(sigma, in->sigma, h->P * sizeof(float));
    } else memset(sigma, 0, h->P * sizeof(float));
    return QPMPC_OK;
}

static void unpack_one(const qpmpc_handle *h, int r, const float *base, qpmpc_tick_out *out) {
    const KParams &kp = h->kp;
    const float *rec = base + (size_t)r * kp.out_stride;
    h->last_out = base;
    memcpy(out->grfs, rec + OUT_GRF, 12 * sizeof(float));
    memcpy(out->predicted_state, rec + OUT_PRED, 24 * sizeof(float));
    out->best_cost = rec[OUT_BESTCOST];
    out->best_freq = rec[OUT_BESTFREQ];
    out->best_index = (int32_t)f2u(rec[OUT_BESTIDX]);
    out->num_elite = (int32_t)f2u(rec[OUT_NELITE]);
    if (out->best) memcpy(out->best, rec + OUT_BEST, h->P * sizeof(float));
    if (out->sigma_new && h->cfg.method == QPMPC_CEM_MPPI) memcpy(out->sigma_new, rec + kp.off_out_sigma, h->P * sizeof(float));
    if (out->elite_indices && kp.elite_k > 0) memcpy(out->elite_indices, rec + kp.off_out_elite, kp.elite_k * sizeof(int32_t));
}

static int ensure_noise(qpmpc_handle *h, size_t floats) {
    if (h->noise_capacity >= floats) return QPMPC_OK;
    cudaFree(h->d_noise); h->d_noise = nullptr; h->noise_capacity = 0;
    CK(cudaMalloc((void **)&h->d_noise, floats * sizeof(float)));
    h->noise_capacity = floats;
    return QPMPC_OK;
}
static int ensure_freqs(qpmpc_handle *h, size_t floats) {
    if (h->freqs_capacity >= floats) return QPMPC_OK;
    cudaFree(h->d_freqs); h->d_freqs = nullptr; h->freqs_capacity = 0;
    CK(cudaMalloc((void **)&h->d_freqs, floats * sizeof(float)));
    h->freqs_capacity = floats;
    return QPMPC_OK;
}

static int launch(qpmpc_handle *h, KernelFn fn, const KParams &kp, dim3 grid, cudaStream_t stream = nullptr) {
    const bool is_finish = (fn == h->finish_fn);
    cudaStream_t st = stream ? stream : h->stream;
    if (!is_finish && h->tail_nb > 0 && kp.mode == MODE_TICK) {
        // large launch: the tick kernel stops after its block records, tail_kernel (multi-block) finishes the tick
        KParams k2 = kp;
        k2.defer_tail = 1;
        void *args[] = {(void *)&k2};
        CK(cudaLaunchKernel((const void *)fn, grid, dim3(h->threads), args, h->smem_bytes, st));
        // programmatic dependent launch: the tail's blocks are scheduled while the tick kernel drains (it triggers at its
        // start) and wait in griddepcontrol.wait for its completion -- hides the ~5 us launch latency between the two
        cudaLaunchConfig_t lc;
        memset(&lc, 0, sizeof(lc));
        lc.gridDim = dim3(h->tail_nb, grid.y); lc.blockDim = dim3(TAIL_TB); lc.dynamicSmemBytes = h->tail_smem_bytes; lc.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        lc.attrs = attr; lc.numAttrs = 1;
        CK(cudaLaunchKernelExC(&lc, (const void *)h->tail_fn, args));
        h->launches += 2;
        return QPMPC_OK;
    }
    void *args[] = {(void *)&kp};
    CK(cudaLaunchKernel((const void *)fn, grid, dim3(is_finish ? TB : h->threads), args,
                        is_finish ? h->finish_smem_bytes : h->smem_bytes, st));
    h->launches++;
    return QPMPC_OK;
}

// upload injected noise / freqs of all robots if any is present; fills kp accordingly
static int stage_injection(qpmpc_handle *h, const qpmpc_tick_in *in, int count, KParams &kp) {
    bool any_noise = false, any_freq = false;
    for (int r = 0; r < count; ++r) { any_noise |= (in[r].noise != nullptr); any_freq |= (in[r].step_freqs != nullptr); }
    if (any_noise) {
        const size_t per = (size_t)h->N * h->P;
        int rc = ensure_noise(h, per * count);
        if (rc) return rc;
        for (int r = 0; r < count; ++r) {
            if (!in[r].noise) return fail(QPMPC_ERR_INVALID, "noise injected for some robots but not robot %d", r);
            CK(cudaMemcpyAsync(h->d_noise + per * r, in[r].noise, per * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        }
        kp.noise_mode = NOISE_INJECTED; kp.noise = h->d_noise;
    }
    if (any_freq) {
        int rc = ensure_freqs(h, (size_t)h->N * count);
        if (rc) return rc;
        for (int r = 0; r < count; ++r) {
            if (!in[r].step_freqs) return fail(QPMPC_ERR_INVALID, "step_freqs injected for some robots but not robot %d", r);
            CK(cudaMemcpyAsync(h->d_freqs + (size_t)h->N * r, in[r].step_freqs, (size_t)h->N * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        }
        kp.freq_injected = 1; kp.freqs = h->d_freqs;
    }
    return QPMPC_OK;
}

// Direct I/O tick of one controller: the packet travels in kernel-parameter space, the result packet comes back through
// mapped pinned memory and is detected by its stamp -- no copy-engine hop and no stream synchronisation on the way.
static inline unsigned int next_seq(qpmpc_handle *h) {
    if (++h->seq == 0) ++h->seq;                       // 0 is the initial content of every stamp
    return h->seq;
}
static inline uint64_t mono_ns() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}
// The {value, seq} units rely on the GPU's 16-byte PCIe stores arriving as atomic aligned 8-byte halves and on the host
// reading an aligned uint64_t with one load: true on x86-64 and on AArch64 (both guarantee single-copy atomicity of
// naturally aligned 8-byte accesses); any other host must use QPMPC_IO_MEMCPY.
#if !defined(__x86_64__) && !defined(__aarch64__)
#error "direct_tick: the {value, seq} publication needs single-copy-atomic aligned 8-byte loads on the host"
#endif
static_assert(sizeof(uint64_t) == 8 && alignof(uint64_t) == 8, "8-byte units");
static const uint64_t DIRECT_TICK_DEADLINE_NS = 4000000000ull;   // > the in-kernel peer wait (2 s), so that one reports first
static int direct_tick(qpmpc_handle *h, KParams &kp, dim3 grid, qpmpc_tick_out *out) {
    const bool timed = h->host_times_on;
    if (timed) h->host_ns[0] = mono_ns();
    memcpy(kp.packet, h->h_in, (size_t)kp.in_stride * sizeof(float));
    kp.param_packet = 1; kp.host_out = h->d_dout;
    if (kp.seq == 0) kp.seq = next_seq(h);
    int rc = launch(h, h->tick_fn, kp, grid);
    if (rc) return rc;
    if (timed) h->host_ns[1] = mono_ns();
    // every 8-byte unit {value, seq} validates itself: take each one as soon as its flag is this tick's seq
    const volatile uint64_t *unit = reinterpret_cast<const volatile uint64_t *>(h->h_dout);
    uint32_t *dec = reinterpret_cast<uint32_t *>(h->h_dec);
    unsigned long long spins = 0;
    uint64_t t_spin0 = 0;
    for (int u = 0; u < kp.out_stride; ++u) {
        uint64_t w;
        while ((uint32_t)((w = unit[u]) >> 32) != kp.seq) {
            if ((++spins & 0xffff) == 0) {       // a failed launch would never publish: surface the error instead
                cudaError_t q = cudaStreamQuery(h->stream);
                if (q != cudaSuccess && q != cudaErrorNotReady) return fail(QPMPC_ERR_CUDA, "tick kernel failed: %s", cudaGetErrorString(q));
                if (q == cudaSuccess && (uint32_t)(unit[u] >> 32) != kp.seq) return fail(QPMPC_ERR_CUDA, "tick kernel finished without publishing its result");
                // a hung kernel (or a peer wait that never ends) must not block a real-time loop for ever
                const uint64_t now = mono_ns();
                if (t_spin0 == 0) t_spin0 = now;
                else if (now - t_spin0 > DIRECT_TICK_DEADLINE_NS)
                    return fail(QPMPC_ERR_CUDA, "tick kernel did not publish its result within %.1f s", 1e-9 * (double)DIRECT_TICK_DEADLINE_NS);
            }
        }
        dec[u] = (uint32_t)w;
    }
    if (timed) h->host_ns[2] = mono_ns();
    unpack_one(h, 0, h->h_dec, &out[0]);
    if (timed) h->host_ns[3] = mono_ns();
    return QPMPC_OK;
}

extern "C" int qpmpc_debug_host_times(qpmpc_handle *h, int32_t enable, uint64_t out[4]) {
    if (!h) return fail(QPMPC_ERR_INVALID, "NULL handle");
    h->host_times_on = enable != 0;
    if (out) memcpy(out, h->host_ns, sizeof(h->host_ns));
    return QPMPC_OK;
}

// Device-generated contact tables (qpmpc_pgg_contact_sequence) are used by the next tick / step whose inputs carry no
// host table, once.
static void bind_device_contacts(qpmpc_handle *h, const qpmpc_tick_in *in, KParams &kp) {
    if (h->contacts_on_device && in[0].contact_sequence == nullptr) kp.contact_dev = h->d_contact;
    h->contacts_on_device = false;
}

extern "C" int qpmpc_pgg_contact_sequence(qpmpc_handle *h, const double *phase_signal, const uint8_t *init,
                                          const double *step_freq, double duty_factor, const double phase_offset[4],
                                          const double *dts, const int32_t *lengths, int32_t n_dts, float *tables) {
    if (!h || !phase_signal || !step_freq || !phase_offset || !dts || !lengths || n_dts < 1) return fail(QPMPC_ERR_INVALID, "NULL / bad argument");
    CK(cudaSetDevice(h->cfg.device));
    const size_t R = h->R;
    const int H = h->H;
    PggParams pp;
    memset(&pp, 0, sizeof(pp));
    pp.duty = duty_factor; pp.H = H; pp.R = (int)R;
    memcpy(pp.offset, phase_offset, 4 * sizeof(double));
    pp.dts[0] = 0.0;                                       // contact_sequence[:, 0] = self.run(0.0, self.step_freq)   (:106)
    for (int i = 1, j = 0; i < H; ++i) {                   // :110-115
        if (i >= lengths[j]) { if (j + 1 >= n_dts) return fail(QPMPC_ERR_INVALID, "contact_sequence_lenghts do not cover the horizon"); ++j; }
        pp.dts[i] = dts[j];
    }
    const size_t in_bytes = R * 4 * sizeof(double) + R * sizeof(double) + R * 4;
    if (!h->d_pgg) CK(cudaMalloc(&h->d_pgg, in_bytes));
    if (!h->d_contact) CK(cudaMalloc((void **)&h->d_contact, R * 4 * H * sizeof(float)));
    double *d_phase = (double *)h->d_pgg, *d_freq = d_phase + R * 4;
    unsigned char *d_init = (unsigned char *)(d_freq + R);
    CK(cudaMemcpyAsync(d_phase, phase_signal, R * 4 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_freq, step_freq, R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (init) CK(cudaMemcpyAsync(d_init, init, R * 4, cudaMemcpyHostToDevice, h->stream));
    const int threads = 128, blocks = (int)((R * 4 + threads - 1) / threads);
    pgg_contact_kernel<<<blocks, threads, 0, h->stream>>>(pp, d_phase, init ? d_init : nullptr, d_freq, h->d_contact);
    h->launches++;
    CK(cudaGetLastError());
    if (tables) {
        CK(cudaMemcpyAsync(tables, h->d_contact, R * 4 * H * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    h->contacts_on_device = true;                          // the tick kernels of this stream run after the table kernel
    return QPMPC_OK;
}

// MPC -> WBC output record: the 75-double seqlock payload of ros2/run_controller.py:72-92, written
// as the MPC process does at :373-388 (sequence odd while writing, even when stable), read at :587-602.
enum { SHM_GRF = 0, SHM_FH = 12, SHM_JP = 24, SHM_JV = 36, SHM_JA = 48, SHM_PRED = 60, SHM_BSF = 72, SHM_LAST = 73, SHM_STAMP = 74 };
extern "C" int qpmpc_shm_pack(double *record, uint64_t *seq, const float *grfs, const float *current_contact,
                              const double *footholds, const float *predicted_state, double best_sample_freq,
                              double last_mpc_loop_time) {
    if (!record || !seq || !grfs || !predicted_state) return fail(QPMPC_ERR_INVALID, "NULL argument");
    const uint64_t s0 = __atomic_load_n(seq, __ATOMIC_RELAXED);
    __atomic_store_n(seq, s0 | 1, __ATOMIC_RELAXED);                       // :374-376 (even -> s+1, odd stays)
    __atomic_thread_fence(__ATOMIC_RELEASE);
    const double nan = __builtin_nan("");
    for (int leg = 0; leg < 4; ++leg) {
        const double c = current_contact ? (double)current_contact[leg] : 1.0;     // interface :225-230 (float64 product)
        for (int k = 0; k < 3; ++k) record[SHM_GRF + 3 * leg + k] = (double)grfs[3 * leg + k] * c;
    }
    for (int k = 0; k < 12; ++k) {
        record[SHM_FH + k] = footholds ? footholds[k] : 0.0;               // ref_state feet, interface :170-175
        // the sampling controller returns no joint trajectories (interface :178-180); numpy stores None as NaN (:380-382)
        record[SHM_JP + k] = nan; record[SHM_JV + k] = nan; record[SHM_JA + k] = nan;
        record[SHM_PRED + k] = (double)predicted_state[k];                 // first 12 of the 24, :383
    }
    record[SHM_BSF] = best_sample_freq;
    record[SHM_LAST] = last_mpc_loop_time;
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);                                   // time.monotonic(), :386
    record[SHM_STAMP] = (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
    __atomic_store_n(seq, (s0 | 1) + 1, __ATOMIC_RELEASE);                 // :388
    return QPMPC_OK;
}

extern "C" int qpmpc_publish_shm(qpmpc_handle *h, int robot, double *record, uint64_t *seq, const float *current_contact,
                                 const double *footholds, double last_mpc_loop_time) {
    if (!h) return fail(QPMPC_ERR_INVALID, "NULL handle");
    if (robot < 0 || robot >= h->R) return fail(QPMPC_ERR_INVALID, "robot %d out of range", robot);
    if (!h->last_out) return fail(QPMPC_ERR_INVALID, "no tick result on the host yet");
    const float *rec = h->last_out + (size_t)robot * h->kp.out_stride;
    return qpmpc_shm_pack(record, seq, rec + OUT_GRF, current_contact, footholds, rec + OUT_PRED, (double)rec[OUT_BESTFREQ],
                          last_mpc_loop_time);
}

// layout of d_frg (doubles): base_pos 3R | euler 3R | vel 2R | ref vel 2R | hips 12R | lift_off_z 4R | stance R | feet 12R |
// hist 40R | then (ints) meta 2R
extern "C" int qpmpc_footholds_reference(qpmpc_handle *h, const double *base_position, const double *base_ori_euler_xyz,
                                         const double *base_xy_lin_vel, const double *ref_base_xy_lin_vel,
                                         const double *hips_position, const double *lift_off_z, const double *stance_time,
                                         double com_height_nominal, double hip_height, double hip_offset, double gravity,
                                         const double *com_pos_offset_b, int32_t reset_history, double *ref_feet) {
    if (!h || !base_position || !base_ori_euler_xyz || !base_xy_lin_vel || !ref_base_xy_lin_vel || !hips_position ||
        !lift_off_z || !stance_time) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    const size_t R = h->R;
    const size_t n_d = R * (3 + 3 + 2 + 2 + 12 + 4 + 1 + 12 + 2 * FRG_HIST);
    if (!h->d_frg) {
        CK(cudaMalloc(&h->d_frg, n_d * sizeof(double) + R * 2 * sizeof(int)));
        CK(cudaMemset(h->d_frg, 0, n_d * sizeof(double) + R * 2 * sizeof(int)));
    }
    double *d = (double *)h->d_frg;
    double *d_pos = d, *d_eul = d_pos + 3 * R, *d_vel = d_eul + 3 * R, *d_rvel = d_vel + 2 * R, *d_hips = d_rvel + 2 * R,
           *d_loz = d_hips + 12 * R, *d_st = d_loz + 4 * R, *d_feet = d_st + R, *d_hist = d_feet + 12 * R;
    int *d_meta = (int *)(d_hist + 2 * FRG_HIST * R);
    if (reset_history) CK(cudaMemsetAsync(d_meta, 0, R * 2 * sizeof(int), h->stream));
    CK(cudaMemcpyAsync(d_pos, base_position, 3 * R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_eul, base_ori_euler_xyz, 3 * R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_vel, base_xy_lin_vel, 2 * R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_rvel, ref_base_xy_lin_vel, 2 * R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_hips, hips_position, 12 * R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_loz, lift_off_z, 4 * R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_st, stance_time, R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    FrgParams fp;
    memset(&fp, 0, sizeof(fp));
    fp.com_height_nominal = com_height_nominal; fp.hip_height = hip_height; fp.hip_offset = hip_offset; fp.gravity = gravity;
    if (com_pos_offset_b) memcpy(fp.com_off_b, com_pos_offset_b, 3 * sizeof(double));
    fp.R = (int)R;
    const int threads = 64, blocks = (int)((R + threads - 1) / threads);
    foothold_reference_kernel<<<blocks, threads, 0, h->stream>>>(fp, d_pos, d_eul, d_vel, d_rvel, d_hips, d_loz, d_st, d_hist,
                                                                 d_meta, d_feet);
    h->launches++;
    CK(cudaGetLastError());
    if (ref_feet) {
        CK(cudaMemcpyAsync(ref_feet, d_feet, 12 * R * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return QPMPC_OK;
}

// ------------------------------------------------------------------------------------------------------
// Chunk-pipelined batched tick (host arrays in, host arrays out; R >= PIPE_MIN_ROBOTS).  The robots are independent, so
// the batch is cut into up to 8 chunks: an upload stream copies chunk c+1 (packed on the host meanwhile) while chunk c
// computes, a download stream returns chunk c-1 (unpacked on the host meanwhile), and the chunk kernels alternate
// between two compute streams so that the tail wave of one chunk is filled by the next -- the copy engines and the host
// work hide behind the kernels instead of bracketing one big launch.  Every per-robot array is addressed as
// base + robot * stride in the kernels, so a chunk is the same kernel on offset base pointers (and robot_base, which
// keeps the Philox streams).
// ------------------------------------------------------------------------------------------------------
static const int PIPE_MIN_ROBOTS = 64;

static int validate_one(const qpmpc_handle *h, int r, const qpmpc_tick_in *in, uint32_t seed0, bool stateful) {
    if (!in->state || !in->reference) return fail(QPMPC_ERR_INVALID, "state / reference must not be NULL");
    if (!in->contact_sequence && !h->contacts_on_device) return fail(QPMPC_ERR_INVALID, "contact_sequence is NULL and no device-generated table is armed (qpmpc_pgg_contact_sequence)");
    if (in->contact_sequence && in->contact_ld < h->H) return fail(QPMPC_ERR_INVALID, "contact_ld %d < horizon %d", in->contact_ld, h->H);
    if (h->cfg.adaptive && !in->phase_signal) return fail(QPMPC_ERR_INVALID, "phase_signal is required by the gait-adaptive controller");
    if (!stateful && in->key[0] != seed0) return fail(QPMPC_ERR_INVALID, "all robots of one batch must share key[0] (robot %d differs); streams differ by robot id", r);
    if (!stateful && h->cfg.method == QPMPC_CEM_MPPI && !in->sigma) return fail(QPMPC_ERR_INVALID, "sigma is required by cem_mppi");
    for (int l = 0; l < 4 && in->contact_sequence; ++l)
        for (int n = 0; n < h->H; ++n) {
            const float c = in->contact_sequence[(size_t)l * in->contact_ld + n];
            if (c != 0.0f && c != 1.0f) return fail(QPMPC_ERR_INVALID, "contact_sequence[%d][%d] = %g is not 0 or 1", l, n, (double)c);
        }
    return QPMPC_OK;
}

static void offset_robots(KParams &kp, int r0) {
    kp.in += (size_t)r0 * kp.in_stride; kp.out += (size_t)r0 * kp.out_stride;
    kp.costs += (size_t)r0 * kp.n_local;
    kp.blk_recs += (size_t)r0 * kp.nrec * kp.blk_rec_stride;
    kp.shard_send += (size_t)r0 * kp.shard_rec_stride;
    kp.counters += r0;
    if (kp.cem_hist) kp.cem_hist += (size_t)r0 * (kp.tail_nb > 0 ? CEM_BINS_FINE : CEM_BINS);
    if (kp.contact_dev) kp.contact_dev += (size_t)r0 * 4 * kp.H;
    if (kp.ctrl_state) kp.ctrl_state += (size_t)r0 * kp.ctrl_stride;
    if (kp.part_recs) kp.part_recs += (size_t)r0 * TAIL_NPART * kp.blk_rec_stride;
    if (kp.tail_cnt) kp.tail_cnt += (size_t)r0 * tail_cnt_stride();
    if (kp.tail_keys) kp.tail_keys += (size_t)r0 * 2 * kp.n_local;
    kp.robot_base += (uint32_t)r0;
}

static void shift_basis(const qpmpc_handle *h, KParams &kp);

// stateful: qpmpc_step semantics (iterations launches per chunk, sigma reset / shift at iteration 0)
static int pipelined_batch(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out, bool stateful, int iterations,
                           float sigma0) {
    const int R = h->R;
    const bool timed = h->host_times_on;                   // qpmpc_debug_host_times: entered / all chunks issued / last chunk back / unpacked
    if (timed) h->host_ns[0] = mono_ns();
    for (int r = 0; r < R; ++r) { int rc = validate_one(h, r, &in[r], in[0].key[0], stateful); if (rc) return rc; }
    if (!h->stream2) {
        CK(cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&h->stream_in, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&h->stream_out, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
        for (int c = 0; c < 8; ++c) {
            CK(cudaEventCreateWithFlags(&h->ev_chunk[c], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_in[c], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_k[c], cudaEventDisableTiming));
        }
    }
    const int per = std::max(32, (R + 7) / 8), C = (R + per - 1) / per;
    CK(cudaEventRecord(h->ev_fork, h->stream));            // after whatever the handle queued before (device tables, resets)
    CK(cudaStreamWaitEvent(h->stream2, h->ev_fork, 0));
    CK(cudaStreamWaitEvent(h->stream_in, h->ev_fork, 0));
    auto pack_chunk = [&](int c) {
        for (int r = c * per; r < std::min(R, (c + 1) * per); ++r) { int rc = pack_one(h, r, &in[r], false, stateful); if (rc) return rc; }
        return (int)QPMPC_OK;
    };
    int rc = pack_chunk(0);                                 // (cannot fail after validate_one; sets kp.seed from robot 0)
    if (rc) return rc;
    if (stateful) h->kp.seed = h->cfg.seed;
    KParams base = h->kp;
    if (h->contacts_on_device && in[0].contact_sequence == nullptr) base.contact_dev = h->d_contact;   // bind_device_contacts,
    base.in = h->d_in; base.out = h->d_out;                // (the armed flag is dropped after the last chunk is packed)
    if (stateful) { base.ctrl_state = h->d_ctrl; base.ctrl_stride = h->ctrl_stride; base.sigma0 = sigma0; }
    for (int c = 0; c < C && rc == QPMPC_OK; ++c) {
        const int r0 = c * per, rn = std::min(per, R - r0);
        cudaStream_t st = (c & 1) ? h->stream2 : h->stream;
        KParams kc = base;
        offset_robots(kc, r0);
        const size_t in_off = (size_t)r0 * base.in_stride, out_off = (size_t)r0 * base.out_stride;
        // CKL: a CUDA error inside the loop must not return with work in flight on four streams -- record it, leave the
        // loop and drain through the common path below
#define CKL(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess && rc == QPMPC_OK) rc = fail(QPMPC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); } while (0)
        CKL(cudaMemcpyAsync(h->d_in + in_off, h->h_in + in_off, (size_t)rn * base.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream_in));
        CKL(cudaEventRecord(h->ev_in[c], h->stream_in));
        CKL(cudaStreamWaitEvent(st, h->ev_in[c], 0));
        for (int it = 0; it < iterations && rc == QPMPC_OK; ++it) {
            if (stateful) {
                kc.sigma_reset = (it == 0 && h->cfg.method == QPMPC_CEM_MPPI) ? 1 : 0;
                kc.shift_kind = 0;
                if (it == 0 && h->shift_step > 0.0f) shift_basis(h, kc);
            }
            rc = launch(h, h->tick_fn, kc, dim3(kc.nblk, rn), st);
        }
        if (rc == QPMPC_OK) {
            CKL(cudaEventRecord(h->ev_k[c], st));
            CKL(cudaStreamWaitEvent(h->stream_out, h->ev_k[c], 0));
            CKL(cudaMemcpyAsync(h->h_out + out_off, h->d_out + out_off, (size_t)rn * base.out_stride * sizeof(float), cudaMemcpyDeviceToHost, h->stream_out));
            CKL(cudaEventRecord(h->ev_chunk[c], h->stream_out));
            if (rc == QPMPC_OK && c + 1 < C) rc = pack_chunk(c + 1);   // host packing of the next chunk overlaps this chunk's kernels
        }
#undef CKL
    }
    h->contacts_on_device = false;
    if (rc != QPMPC_OK) {
        // drain every stream; in stateful mode the chunks already launched have advanced their device records (best,
        // key counter, previous contact) while this call fails: the batch is half-advanced and the caller must
        // qpmpc_step_reset before the next step (documented in include/qpmpc.h)
        cudaStreamSynchronize(h->stream_in); cudaStreamSynchronize(h->stream); cudaStreamSynchronize(h->stream2); cudaStreamSynchronize(h->stream_out);
        const std::string first = g_err;
        (void)cudaGetLastError();
        g_err = first + (stateful ? " [stateful batch half-advanced: call qpmpc_step_reset]" : "");
        return rc;
    }
    if (timed) h->host_ns[1] = mono_ns();
    for (int c = 0; c < C; ++c) {
        CK(cudaEventSynchronize(h->ev_chunk[c]));
        if (timed && c == C - 1) h->host_ns[2] = mono_ns();
        for (int r = c * per; r < std::min(R, (c + 1) * per); ++r) unpack_one(h, r, h->h_out, &out[r]);
    }
    if (timed) h->host_ns[3] = mono_ns();
    // every download waited for its chunk's kernels and those for their upload: all four streams are idle here
    return QPMPC_OK;
}

static int run_tick(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out) {
    if (!h || !in || !out) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (h->cfg.shard_world != 1) return fail(QPMPC_ERR_STATE, "handle is a shard: use qpmpc_tick_sharded / qpmpc_shard_partial");
    CK(cudaSetDevice(h->cfg.device));
    const int R = h->R;
    if (R >= PIPE_MIN_ROBOTS && h->cfg.io_mode != QPMPC_IO_MAPPED) {
        bool injected = false;
        for (int r = 0; r < R; ++r) injected |= (in[r].noise != nullptr || in[r].step_freqs != nullptr);
        if (!injected) return pipelined_batch(h, in, out, false, 1, 0.0f);
    }
    for (int r = 0; r < R; ++r) { int rc = pack_one(h, r, &in[r], false); if (rc) return rc; }
    KParams kp = h->kp;
    bind_device_contacts(h, in, kp);
    int rc = stage_injection(h, in, R, kp);
    if (rc) return rc;
    const bool plain = (kp.noise_mode == NOISE_PHILOX && !kp.freq_injected);
    const size_t in_bytes = (size_t)R * kp.in_stride * sizeof(float), out_bytes = (size_t)R * kp.out_stride * sizeof(float);
    const dim3 grid(kp.nblk, R);
    const bool mapped = h->cfg.io_mode == QPMPC_IO_MAPPED;
    if (plain && h->direct) return direct_tick(h, kp, grid, out);
    if (plain && h->cfg.use_graph && !mapped && kp.contact_dev == nullptr) {   // (the graph is captured without device tables)
        if (!h->graph_exec) {                      // capture (H2D, kernel, D2H) once; the packet addresses never change
            cudaGraph_t g = nullptr;
            CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
            cudaMemcpyAsync(h->d_in, h->h_in, in_bytes, cudaMemcpyHostToDevice, h->stream);
            const int lrc = launch(h, h->tick_fn, kp, grid);       // one kernel, or the (tick, tail) pair of a large launch
            h->launches -= (h->tail_nb > 0 ? 2 : 1);               // counted per replay below
            cudaMemcpyAsync(h->h_out, h->d_out, out_bytes, cudaMemcpyDeviceToHost, h->stream);
            CK(cudaStreamEndCapture(h->stream, &g));
            if (lrc) { if (g) cudaGraphDestroy(g); return lrc; }
            CK(cudaGraphInstantiate(&h->graph_exec, g, 0));
            CK(cudaGraphDestroy(g));
        }
        CK(cudaGraphLaunch(h->graph_exec, h->stream));
        h->launches += (h->tail_nb > 0 ? 2 : 1);
    } else {
        if (!mapped) CK(cudaMemcpyAsync(h->d_in, h->h_in, in_bytes, cudaMemcpyHostToDevice, h->stream));
        rc = launch(h, h->tick_fn, kp, grid);
        if (rc) return rc;
        if (!mapped) CK(cudaMemcpyAsync(h->h_out, h->d_out, out_bytes, cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < R; ++r) unpack_one(h, r, h->h_out, &out[r]);
    return QPMPC_OK;
}

// The sampling-iteration loop of one MPC tick (I:125-168) kept on the device: iteration i+1 starts from iteration i's
// best (and, for cem_mppi, its refitted sigma) with the next Philox counter -- exactly what the interface does with
// with_newkey() / best_control_parameters / with_newsigma() between two jitted_compute_control calls, minus the
// D2H + H2D round trip and the host bookkeeping in between.  One H2D, `iterations` launches, one D2H.
extern "C" int qpmpc_tick_iterations(qpmpc_handle *h, const qpmpc_tick_in *in, int32_t iterations, qpmpc_tick_out *out) {
    if (!h || !in || !out) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (iterations < 1) return fail(QPMPC_ERR_INVALID, "iterations must be >= 1");
    if (h->cfg.shard_world != 1) return fail(QPMPC_ERR_STATE, "handle is a shard: sampling iterations run on unsharded handles");
    if (iterations == 1) return run_tick(h, in, out);
    CK(cudaSetDevice(h->cfg.device));
    const int R = h->R;
    for (int r = 0; r < R; ++r) {
        if (in[r].noise || in[r].step_freqs) return fail(QPMPC_ERR_INVALID, "injected noise / frequencies describe ONE iteration: use qpmpc_tick");
        int rc = pack_one(h, r, &in[r], false);
        if (rc) return rc;
    }
    KParams kp = h->kp;
    bind_device_contacts(h, in, kp);
    kp.in = h->d_in; kp.out = h->d_out;                      // the carried state lives in the device copy of the packet
    kp.carry_best = 1;
    CK(cudaMemcpyAsync(h->d_in, h->h_in, (size_t)R * kp.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    for (int it = 0; it < iterations; ++it) {
        int rc = launch(h, h->tick_fn, kp, dim3(kp.nblk, R));
        if (rc) return rc;
    }
    CK(cudaMemcpyAsync(h->h_out, h->d_out, (size_t)R * kp.out_stride * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < R; ++r) unpack_one(h, r, h->h_out, &out[r]);
    return QPMPC_OK;
}

// ------------------------------------------------------------------------------------------------------
// Stateful controller: everything the interface keeps between ticks -- best_control_parameters,
// sigma, the key counter, previous_contact_mpc (I:19,123-168) -- lives in a device record; one call does what the block
// I:118-168 does: prepare_state_and_reference (in the kernel's staging), the sampling-iteration loop, the result.
// ------------------------------------------------------------------------------------------------------
static int ensure_ctrl(qpmpc_handle *h) {
    if (h->d_ctrl) return QPMPC_OK;
    h->ctrl_stride = (int)round4(2 * (size_t)h->P + 4);
    const size_t n = (size_t)h->R * h->ctrl_stride;
    CK(cudaMalloc((void **)&h->d_ctrl, n * sizeof(float)));
    std::vector<float> init(n, 0.0f);
    for (int r = 0; r < h->R; ++r) {
        float *rec = init.data() + (size_t)r * h->ctrl_stride;
        rec[2 * h->P] = u2f(1u);                       // master_key counter after the first with_newkey()
        rec[2 * h->P + 1] = u2f(0xFu);                 // previous_contact_mpc = [1, 1, 1, 1], I:19
    }
    // on the handle's (non-blocking) stream: ordered behind a tick whose tail may still be running (direct_tick returns
    // as soon as the stamped packet is on the host)
    CK(cudaMemcpyAsync(h->d_ctrl, init.data(), n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->kp.seed = h->cfg.seed;
    return QPMPC_OK;
}

extern "C" int qpmpc_step_reset(qpmpc_handle *h, const uint32_t key[2], const float *best, const int32_t *previous_contact) {
    if (!h || !key) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    int rc = ensure_ctrl(h);
    if (rc) return rc;
    const size_t n = (size_t)h->R * h->ctrl_stride;
    std::vector<float> init(n, 0.0f);
    for (int r = 0; r < h->R; ++r) {
        float *rec = init.data() + (size_t)r * h->ctrl_stride;
        if (best) memcpy(rec, best + (size_t)r * h->P, (size_t)h->P * sizeof(float));
        rec[2 * h->P] = u2f(key[1]);
        uint32_t m = 0xFu;
        if (previous_contact) { m = 0u; for (int l = 0; l < 4; ++l) m |= (previous_contact[4 * r + l] != 0 ? 1u : 0u) << l; }
        rec[2 * h->P + 1] = u2f(m);
    }
    CK(cudaMemcpyAsync(h->d_ctrl, init.data(), n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->kp.seed = key[0];
    h->cfg.seed = key[0];
    return QPMPC_OK;
}

extern "C" int qpmpc_step_state(qpmpc_handle *h, int robot, float *best, float *sigma, int32_t previous_contact[4], uint32_t key[2]) {
    if (!h || robot < 0 || robot >= h->R) return fail(QPMPC_ERR_INVALID, "bad argument");
    CK(cudaSetDevice(h->cfg.device));
    int rc = ensure_ctrl(h);
    if (rc) return rc;
    std::vector<float> rec((size_t)h->ctrl_stride);
    CK(cudaMemcpyAsync(rec.data(), h->d_ctrl + (size_t)robot * h->ctrl_stride, rec.size() * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (best) memcpy(best, rec.data(), (size_t)h->P * sizeof(float));
    if (sigma) memcpy(sigma, rec.data() + h->P, (size_t)h->P * sizeof(float));
    if (previous_contact) { const uint32_t m = f2u(rec[2 * h->P + 1]); for (int l = 0; l < 4; ++l) previous_contact[l] = (m >> l) & 1u; }
    if (key) { key[0] = h->cfg.seed; key[1] = f2u(rec[2 * h->P]); }
    return QPMPC_OK;
}

// shift_solution for the stateful controller: index and basis of the spline at `step`, in numpy float32 order
// (S:186-196 linear, S:209-225 cubic, S:265 zero-order); the kernel's staging applies them to the carried best.
static void shift_basis(const qpmpc_handle *h, KParams &kp) {
    const float step = h->shift_step;
    const int H = h->H, ns = kp.ns;
    kp.shift_kind = 1 + (int)h->cfg.parametrization;
    if (h->cfg.parametrization == QPMPC_ZERO_ORDER) { kp.shift_index = (int)(int16_t)step; return; }
    int index = 0;                                            // max(where(step >= linspace(0, H, ns + 1), arange, 0))
    for (int k = 0; k <= ns; ++k) if (step >= kp.bounds[k] && k > index) index = k;
    volatile float q = step / (float)((double)H / (double)ns) - (float)index;
    kp.shift_index = index;
    if (h->cfg.parametrization == QPMPC_LINEAR_SPLINE) { kp.shift_w[0] = 1.0f - q; kp.shift_w[1] = q; return; }
    volatile float q2 = q * q, q3 = q2 * q;                   // q*q*q
    volatile float tq = 2.0f * q, tq2 = tq * q, tq3 = tq2 * q;        // 2*q*q*q
    volatile float hq = 3.0f * q, hq2 = hq * q;               // 3*q*q
    volatile float mq = -2.0f * q, mq2 = mq * q, mq3 = mq2 * q;       // -2*q*q*q
    volatile float a = tq3 - hq2, b = q3 - tq2;
    kp.shift_w[0] = a + 1.0f;                                 // 2q^3 - 3q^2 + 1
    kp.shift_w[1] = b + q;                                    // q^3 - 2q^2 + q
    kp.shift_w[2] = mq3 + hq2;                                // -2q^3 + 3q^2
    kp.shift_w[3] = q3 - q2;                                  // q^3 - q^2
}

extern "C" const char *qpmpc_kernel_name(const qpmpc_handle *h) {
    if (!h) return "";
    return h->lanes == 4 ? "tick_kernel_quad" : (h->lanes == 2 ? "tick_kernel_x2" : (h->helpers ? "tick_kernel_h" : "tick_kernel"));
}

extern "C" int qpmpc_step_set_shift(qpmpc_handle *h, float step) {
    if (!h) return fail(QPMPC_ERR_INVALID, "NULL handle");
    if (!(step >= 0.0f)) return fail(QPMPC_ERR_INVALID, "step must be >= 0");
    if (step > 0.0f && h->cfg.adaptive) return fail(QPMPC_ERR_INVALID, "shift_solution: the gait-adaptive reference raises here (A:540-552)");
    if (step > 0.0f && h->cfg.parametrization == QPMPC_ZERO_ORDER && (int)(int16_t)step >= h->H) return fail(QPMPC_ERR_INVALID, "step beyond the horizon");
    if (step > 0.0f && h->cfg.parametrization != QPMPC_ZERO_ORDER && step >= (float)h->H) return fail(QPMPC_ERR_INVALID, "step beyond the horizon");
    h->shift_step = step;
    return QPMPC_OK;
}

extern "C" int qpmpc_step(qpmpc_handle *h, const qpmpc_tick_in *in, int32_t iterations, float sigma0, qpmpc_tick_out *out) {
    if (!h || !in || !out) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (iterations < 1) return fail(QPMPC_ERR_INVALID, "iterations must be >= 1");
    if (h->cfg.shard_world != 1) return fail(QPMPC_ERR_STATE, "handle is a shard: the stateful controller runs on unsharded handles");
    CK(cudaSetDevice(h->cfg.device));
    int rc = ensure_ctrl(h);
    if (rc) return rc;
    const int R = h->R;
    const uint32_t seed = h->cfg.seed;
    for (int r = 0; r < R; ++r)
        if (in[r].noise || in[r].step_freqs) return fail(QPMPC_ERR_INVALID, "the stateful controller draws its own noise: use qpmpc_tick for injection");
    if (R >= PIPE_MIN_ROBOTS && h->cfg.io_mode != QPMPC_IO_MAPPED) return pipelined_batch(h, in, out, true, iterations, sigma0);
    for (int r = 0; r < R; ++r) {
        rc = pack_one(h, r, &in[r], false, true);
        if (rc) return rc;
    }
    h->kp.seed = seed;                                 // pack_one took key[0] of the (ignored) input keys
    KParams kp = h->kp;
    bind_device_contacts(h, in, kp);
    kp.ctrl_state = h->d_ctrl; kp.ctrl_stride = h->ctrl_stride; kp.sigma0 = sigma0;
    const dim3 grid(kp.nblk, R);
    const bool direct = h->direct;
    if (direct) {
        memcpy(kp.packet, h->h_in, (size_t)kp.in_stride * sizeof(float));
        kp.param_packet = 1;
    } else {
        kp.in = h->d_in; kp.out = h->d_out;
        CK(cudaMemcpyAsync(h->d_in, h->h_in, (size_t)R * kp.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    }
    for (int it = 0; it < iterations; ++it) {
        kp.sigma_reset = (it == 0 && h->cfg.method == QPMPC_CEM_MPPI) ? 1 : 0;
        kp.shift_kind = 0;
        if (it == 0 && h->shift_step > 0.0f) shift_basis(h, kp);
        if (direct && it == iterations - 1) { kp.seq = 0; return direct_tick(h, kp, grid, out); }
        rc = launch(h, h->tick_fn, kp, grid);
        if (rc) return rc;
    }
    CK(cudaMemcpyAsync(h->h_out, h->d_out, (size_t)R * kp.out_stride * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < R; ++r) unpack_one(h, r, h->h_out, &out[r]);
    return QPMPC_OK;
}

extern "C" int qpmpc_tick(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out) {
    if (h && h->R != 1) return fail(QPMPC_ERR_INVALID, "handle batches %d robots: use qpmpc_tick_batched", h->R);
    return run_tick(h, in, out);
}
extern "C" int qpmpc_tick_batched(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out) { return run_tick(h, in, out); }

// Structure-of-arrays form of qpmpc_tick_batched for host code that keeps R robots in flat arrays (one row per robot):
// no per-robot struct marshalling.  sigma / phase_signal / nominal_step_freq / optimize_swing may be NULL when unused.
extern "C" int qpmpc_tick_batched_flat(qpmpc_handle *h, const float *states, const float *references, const float *contacts,
                                       int32_t contact_ld, const float *best, const float *sigma, const float *phase_signals,
                                       const float *nominal_step_freqs, const int32_t *optimize_swing, const uint32_t key[2],
                                       float *grfs, float *predicted_states, float *best_out, float *best_costs,
                                       float *best_freqs, int32_t *best_indices, float *sigma_new) {
    if (!h || !states || !references || !key) return fail(QPMPC_ERR_INVALID, "NULL argument");
    const int R = h->R, P = h->P;
    std::vector<qpmpc_tick_in> in((size_t)R);
    std::vector<qpmpc_tick_out> out((size_t)R);
    for (int r = 0; r < R; ++r) {
        qpmpc_tick_in &i = in[r];
        memset(&i, 0, sizeof(i));
        i.state = states + (size_t)r * 24; i.reference = references + (size_t)r * 24;
        i.contact_sequence = contacts ? contacts + (size_t)r * 4 * contact_ld : nullptr; i.contact_ld = contact_ld;
        i.best = best ? best + (size_t)r * P : nullptr;
        i.sigma = sigma ? sigma + (size_t)r * P : nullptr;
        i.phase_signal = phase_signals ? phase_signals + (size_t)r * 4 : nullptr;
        i.nominal_step_freq = nominal_step_freqs ? nominal_step_freqs[r] : 0.f;
        i.optimize_swing = optimize_swing ? optimize_swing[r] : 0;
        i.key[0] = key[0]; i.key[1] = key[1];
        qpmpc_tick_out &o = out[r];
        memset(&o, 0, sizeof(o));
        o.best = best_out ? best_out + (size_t)r * P : nullptr;
        o.sigma_new = sigma_new ? sigma_new + (size_t)r * P : nullptr;
    }
    int rc = run_tick(h, in.data(), out.data());
    if (rc) return rc;
    for (int r = 0; r < R; ++r) {
        if (grfs) memcpy(grfs + (size_t)r * 12, out[r].grfs, 12 * sizeof(float));
        if (predicted_states) memcpy(predicted_states + (size_t)r * 24, out[r].predicted_state, 24 * sizeof(float));
        if (best_costs) best_costs[r] = out[r].best_cost;
        if (best_freqs) best_freqs[r] = out[r].best_freq;
        if (best_indices) best_indices[r] = out[r].best_index;
    }
    return QPMPC_OK;
}

// Structure-of-arrays form of qpmpc_step for R robots kept in flat host arrays (states are RAW, see qpmpc_step).
extern "C" int qpmpc_step_flat(qpmpc_handle *h, const float *states, const float *references, const float *contacts,
                               int32_t contact_ld, const float *phase_signals, const float *nominal_step_freqs,
                               const int32_t *optimize_swing, int32_t iterations, float sigma0, float *grfs,
                               float *predicted_states, float *best_out, float *best_costs, float *best_freqs,
                               int32_t *best_indices, float *sigma_new) {
    if (!h || !states || !references) return fail(QPMPC_ERR_INVALID, "NULL argument");
    const int R = h->R, P = h->P;
    std::vector<qpmpc_tick_in> in((size_t)R);
    std::vector<qpmpc_tick_out> out((size_t)R);
    for (int r = 0; r < R; ++r) {
        qpmpc_tick_in &i = in[r];
        memset(&i, 0, sizeof(i));
        i.state = states + (size_t)r * 24; i.reference = references + (size_t)r * 24;
        i.contact_sequence = contacts ? contacts + (size_t)r * 4 * contact_ld : nullptr; i.contact_ld = contact_ld;
        i.phase_signal = phase_signals ? phase_signals + (size_t)r * 4 : nullptr;
        i.nominal_step_freq = nominal_step_freqs ? nominal_step_freqs[r] : 0.f;
        i.optimize_swing = optimize_swing ? optimize_swing[r] : 0;
        qpmpc_tick_out &o = out[r];
        memset(&o, 0, sizeof(o));
        o.best = best_out ? best_out + (size_t)r * P : nullptr;
        o.sigma_new = sigma_new ? sigma_new + (size_t)r * P : nullptr;
    }
    int rc = qpmpc_step(h, in.data(), iterations, sigma0, out.data());
    if (rc) return rc;
    for (int r = 0; r < R; ++r) {
        if (grfs) memcpy(grfs + (size_t)r * 12, out[r].grfs, 12 * sizeof(float));
        if (predicted_states) memcpy(predicted_states + (size_t)r * 24, out[r].predicted_state, 24 * sizeof(float));
        if (best_costs) best_costs[r] = out[r].best_cost;
        if (best_freqs) best_freqs[r] = out[r].best_freq;
        if (best_indices) best_indices[r] = out[r].best_index;
    }
    return QPMPC_OK;
}

extern "C" int qpmpc_stance_torques(qpmpc_handle *h, const float *feet_jac, const float *current_contact,
                                    float *grfs_contact, float *tau) {
    if (!h || !feet_jac || !current_contact) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    const size_t R = h->R;
    if (!h->d_epi) CK(cudaMalloc((void **)&h->d_epi, R * 64 * sizeof(float)));
    float *d_jac = h->d_epi, *d_con = d_jac + R * 36, *d_grf = d_con + R * 4, *d_tau = d_grf + R * 12;
    CK(cudaMemcpyAsync(d_jac, feet_jac, R * 36 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_con, current_contact, R * 4 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    const int threads = 128, blocks = (int)((R * 12 + threads - 1) / threads);
    stance_torque_kernel<<<blocks, threads, 0, h->stream>>>(h->kout, h->kp.out_stride, d_jac, d_con, d_grf, d_tau, (int)R);
    h->launches++;
    CK(cudaGetLastError());
    if (grfs_contact) CK(cudaMemcpyAsync(grfs_contact, d_grf, R * 12 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (tau) CK(cudaMemcpyAsync(tau, d_tau, R * 12 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

// ---- qpmpc_stance_torques without any allocation ---------------------------------------------------------------------
// (no counterpart in the reference: added for callers that may not call cudaMalloc
// -- code under CUDA graph capture, tight serving loops, harnesses that forbid it.)
// The staging area of stance_torque_kernel (the Jacobians and contact flags going in, the masked GRFs and torques coming
// out: 64 floats per robot) lives in this file-scope device array instead of the handle's lazily cudaMalloc'ed d_epi:
// QPMPC_NO_ALLOC_MAX_ROBOTS * 64 floats = 256 KiB that every GPU loading the library holds until the process exits.
// In host code `g_static_epi` is only a host-side stand-in for the symbol, NOT a device address: static_epi_address()
// resolves it with cudaGetSymbolAddress once per device (that device current) and the kernel is given that pointer.
// Every call on a device shares the one array: it keeps the previous call's contents (the outputs are cleared before each
// launch; the inputs are overwritten), and calls that overlap on the GPU (two streams, two host threads, two handles on
// one device) race on it -- serialise them.
static __device__ __align__(256) float g_static_epi[QPMPC_NO_ALLOC_MAX_ROBOTS * 64];
constexpr int STATIC_EPI_MAX_DEVICES = 64;
static float *g_static_epi_ptr[STATIC_EPI_MAX_DEVICES] = {};
static std::mutex g_static_epi_mutex;

static int static_epi_address(int device, float **ptr) {          // the caller has made `device` current
    if (device < 0 || device >= STATIC_EPI_MAX_DEVICES) return fail(QPMPC_ERR_INVALID, "device %d: the static scratch is tracked for devices 0..%d", device, STATIC_EPI_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_static_epi_mutex);
    if (!g_static_epi_ptr[device]) {
        float *p = nullptr;
        CK(cudaGetSymbolAddress((void **)&p, g_static_epi));
        if (!p) return fail(QPMPC_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch on device %d", device);
        g_static_epi_ptr[device] = p;
    }
    *ptr = g_static_epi_ptr[device];
    return QPMPC_OK;
}

extern "C" int qpmpc_stance_torques_static(qpmpc_handle *h, const float *feet_jac, const float *current_contact,
                                           float *grfs_contact, float *tau) {
    if (!h || !feet_jac || !current_contact) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (h->R > QPMPC_NO_ALLOC_MAX_ROBOTS)
        return fail(QPMPC_ERR_INVALID, "qpmpc_stance_torques_static: %d robots exceed the static scratch (QPMPC_NO_ALLOC_MAX_ROBOTS = %d); use qpmpc_stance_torques",
                    h->R, QPMPC_NO_ALLOC_MAX_ROBOTS);
    CK(cudaSetDevice(h->cfg.device));
    float *scratch = nullptr;
    const int rc = static_epi_address(h->cfg.device, &scratch);
    if (rc) return rc;
    const size_t R = h->R;
    float *d_jac = scratch, *d_con = d_jac + R * 36, *d_grf = d_con + R * 4, *d_tau = d_grf + R * 12;    // d_epi's layout
    CK(cudaMemcpyAsync(d_jac, feet_jac, R * 36 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_con, current_contact, R * 4 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemsetAsync(d_grf, 0, R * 24 * sizeof(float), h->stream));       // the last call's results do not survive
    const int threads = 128, blocks = (int)((R * 12 + threads - 1) / threads);
    stance_torque_kernel<<<blocks, threads, 0, h->stream>>>(h->kout, h->kp.out_stride, d_jac, d_con, d_grf, d_tau, (int)R);
    h->launches++;
    CK(cudaGetLastError());
    if (grfs_contact) CK(cudaMemcpyAsync(grfs_contact, d_grf, R * 12 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (tau) CK(cudaMemcpyAsync(tau, d_tau, R * 12 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_get_costs(qpmpc_handle *h, int robot, float *dst) {
    if (!h || !dst || robot < 0 || robot >= h->R) return fail(QPMPC_ERR_INVALID, "bad argument");
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaMemcpyAsync(dst, h->d_costs + (size_t)robot * h->n_local, (size_t)h->n_local * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_rollout(qpmpc_handle *h, const float *state, const float *reference, const float *contact_sequence,
                             int32_t contact_ld, const float *phase_signal, const float *params, const float *step_freqs,
                             float *costs) {
    if (!h || !params || !costs) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (h->R != 1) return fail(QPMPC_ERR_INVALID, "qpmpc_rollout needs a single-robot handle");
    if (h->cfg.adaptive && !step_freqs) return fail(QPMPC_ERR_INVALID, "step_freqs is required by the gait-adaptive rollout (A:326)");
    CK(cudaSetDevice(h->cfg.device));
    qpmpc_tick_in in;
    memset(&in, 0, sizeof(in));
    in.state = state; in.reference = reference; in.contact_sequence = contact_sequence; in.contact_ld = contact_ld;
    in.phase_signal = phase_signal;
    int rc = pack_one(h, 0, &in, true);
    if (rc) return rc;
    KParams kp = h->kp;
    kp.mode = MODE_ROLLOUT; kp.n_local = h->N; kp.sample_base = 0; kp.fused_finish = 0;
    plan_blocks(h->N, h->lanes, &kp.iters, &kp.nblk);
    kp.iters_rec = kp.iters; kp.nrec = kp.nblk;
    const size_t per = (size_t)h->N * h->P;
    rc = ensure_noise(h, per);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h->d_noise, params, per * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    kp.noise_mode = NOISE_INJECTED; kp.noise = h->d_noise;
    if (step_freqs) {
        rc = ensure_freqs(h, h->N);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->d_freqs, step_freqs, (size_t)h->N * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        kp.freq_injected = 1; kp.freqs = h->d_freqs;
    }
    float *din = h->d_in;
    CK(cudaMemcpyAsync(din, h->h_in, (size_t)kp.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    kp.in = din;
    rc = launch(h, h->tick_fn, kp, dim3(kp.nblk, 1));
    if (rc) return rc;
    CK(cudaMemcpyAsync(costs, h->d_costs, (size_t)h->N * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_generate_noise(qpmpc_handle *h, int robot, const uint32_t key[2], const float *sigma, float *dst) {
    if (!h || !key || !dst) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (h->cfg.method == QPMPC_CEM_MPPI && !sigma) return fail(QPMPC_ERR_INVALID, "sigma is required by cem_mppi");
    CK(cudaSetDevice(h->cfg.device));
    const size_t per = (size_t)h->N * h->P;
    int rc = ensure_noise(h, per);
    if (rc) return rc;
    if (sigma) CK(cudaMemcpyAsync(h->d_sigma, sigma, (size_t)h->P * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    const int threads = 256;
    const int blocks = (int)std::min<size_t>((per / 4 + threads - 1) / threads, 148 * 8);
    switch (h->cfg.method) {
        case QPMPC_RANDOM_SAMPLING: noise_export_kernel<RANDOM_SAMPLING><<<blocks, threads, 0, h->stream>>>(h->kp, key[0], (uint32_t)robot, key[1], h->d_sigma, h->d_noise); break;
        case QPMPC_MPPI: noise_export_kernel<MPPI><<<blocks, threads, 0, h->stream>>>(h->kp, key[0], (uint32_t)robot, key[1], h->d_sigma, h->d_noise); break;
        default: noise_export_kernel<CEM_MPPI><<<blocks, threads, 0, h->stream>>>(h->kp, key[0], (uint32_t)robot, key[1], h->d_sigma, h->d_noise); break;
    }
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dst, h->d_noise, per * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_generate_step_freqs(qpmpc_handle *h, int robot, const uint32_t key[2], float nominal_step_freq,
                                         int32_t optimize_swing, float *dst) {
    if (!h || !key || !dst) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    int rc = ensure_freqs(h, h->N);
    if (rc) return rc;
    float *rec = h->h_in;                        // scratch use of robot 0's packet (overwritten by the next tick anyway)
    rec[IN_NOMFREQ] = nominal_step_freq; rec[IN_OPTSWING] = u2f((uint32_t)optimize_swing);
    rec[IN_SEED] = u2f(key[0]); rec[IN_TICK] = u2f(key[1]); rec[IN_ROBOT] = u2f((uint32_t)robot);
    CK(cudaMemcpyAsync(h->d_in, rec, (size_t)h->kp.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    KParams kp = h->kp;
    kp.freq_injected = 0;
    kp.seed = key[0]; kp.robot_base = (uint32_t)robot;
    const int threads = 256, blocks = std::min((h->N + threads - 1) / threads, 148 * 8);
    switch (h->cfg.method) {
        case QPMPC_RANDOM_SAMPLING: freq_export_kernel<RANDOM_SAMPLING><<<blocks, threads, 0, h->stream>>>(kp, h->d_in, h->d_freqs); break;
        case QPMPC_MPPI: freq_export_kernel<MPPI><<<blocks, threads, 0, h->stream>>>(kp, h->d_in, h->d_freqs); break;
        default: freq_export_kernel<CEM_MPPI><<<blocks, threads, 0, h->stream>>>(kp, h->d_in, h->d_freqs); break;
    }
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dst, h->d_freqs, (size_t)h->N * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

// ------------------------------------------------------------------------------------------------------
// sample-sharded controller: partial -> (NCCL all-gather | caller-side gather) -> finish
// ------------------------------------------------------------------------------------------------------
static int shard_partial_async(qpmpc_handle *h, const qpmpc_tick_in *in, KParams &kp) {
    const int R = h->R;
    for (int r = 0; r < R; ++r) { int rc = pack_one(h, r, &in[r], false); if (rc) return rc; }
    kp = h->kp;
    kp.fused_finish = 0;
    int rc = stage_injection(h, in, R, kp);
    if (rc) return rc;
    if (h->cfg.io_mode != QPMPC_IO_MAPPED)
        CK(cudaMemcpyAsync(h->d_in, h->h_in, (size_t)R * kp.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    return launch(h, h->tick_fn, kp, dim3(kp.nblk, R));
}

static int shard_finish_async(qpmpc_handle *h, const KParams &kp_in) {
    KParams kp = kp_in;
    int rc = launch(h, h->finish_fn, kp, dim3(1, h->R));
    if (rc) return rc;
    if (h->cfg.io_mode != QPMPC_IO_MAPPED)
        CK(cudaMemcpyAsync(h->h_out, h->d_out, (size_t)h->R * kp.out_stride * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_partial(qpmpc_handle *h, const qpmpc_tick_in *in) {
    if (!h || !in) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    int rc = shard_partial_async(h, in, h->partial_kp);
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    h->partial_done = true;
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_record_device_ptr(qpmpc_handle *h, void **send, void **recv) {
    if (!h) return fail(QPMPC_ERR_INVALID, "NULL handle");
    if (send) *send = h->d_send;
    if (recv) *recv = h->d_recv;
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_get_record(qpmpc_handle *h, float *host_rec) {
    if (!h || !host_rec) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaMemcpy(host_rec, h->d_send, (size_t)h->R * h->kp.shard_rec_stride * sizeof(float), cudaMemcpyDeviceToHost));
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_copy_record(qpmpc_handle *h, int slot, const float *host_rec) {
    if (!h || !host_rec || slot < 0 || slot >= h->cfg.shard_world) return fail(QPMPC_ERR_INVALID, "bad argument");
    CK(cudaSetDevice(h->cfg.device));
    const size_t n = (size_t)h->R * h->kp.shard_rec_stride;
    CK(cudaMemcpy(h->d_recv + n * slot, host_rec, n * sizeof(float), cudaMemcpyHostToDevice));
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_finish(qpmpc_handle *h, qpmpc_tick_out *out) {
    if (!h || !out) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (!h->partial_done) return fail(QPMPC_ERR_STATE, "qpmpc_shard_finish without a preceding qpmpc_shard_partial");
    CK(cudaSetDevice(h->cfg.device));
    int rc = shard_finish_async(h, h->partial_kp);
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < h->R; ++r) unpack_one(h, r, h->h_out, &out[r]);
    h->partial_done = false;
    return QPMPC_OK;
}

extern "C" int qpmpc_nccl_unique_id(const char *libnccl_path, unsigned char id[128]) {
    int rc = load_nccl(libnccl_path);
    if (rc) return rc;
    NcclId nid;
    int e = g_nccl.GetUniqueId(&nid);
    if (e != 0) return fail(QPMPC_ERR_NCCL, "ncclGetUniqueId: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(e) : "error");
    memcpy(id, nid.internal, 128);
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_init_nccl(qpmpc_handle *h, const char *libnccl_path, const unsigned char id[128]) {
    if (!h || !id) return fail(QPMPC_ERR_INVALID, "NULL argument");
    int rc = load_nccl(libnccl_path);
    if (rc) return rc;
    CK(cudaSetDevice(h->cfg.device));
    NcclId nid;
    memcpy(nid.internal, id, 128);
    int e = g_nccl.CommInitRank(&h->comm, h->cfg.shard_world, nid, h->cfg.shard_rank);
    if (e != 0) return fail(QPMPC_ERR_NCCL, "ncclCommInitRank: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(e) : "error");
    return QPMPC_OK;
}

// ---- peer exchange: the collective fused into the tick kernel (one launch per tick on every rank) ----
struct PeerBlob {                         // QPMPC_PEER_BLOB_BYTES = 128
    cudaIpcMemHandle_t ipc;               // 64 bytes
    uint64_t pid, raw_ptr;
    int32_t device, world, rank, rec_stride, robots;
    unsigned char uuid[16];               // physical GPU: two shards of one controller may not share a GPU (see attach)
    unsigned char pad[128 - 64 - 16 - 20 - 16];
};
static_assert(sizeof(PeerBlob) == QPMPC_PEER_BLOB_BYTES, "peer blob layout");

extern "C" int qpmpc_shard_peer_export(qpmpc_handle *h, unsigned char blob[QPMPC_PEER_BLOB_BYTES]) {
    if (!h || !blob) return fail(QPMPC_ERR_INVALID, "NULL argument");
    const int W = h->cfg.shard_world;
    if (W < 2 || W > MAX_PEERS) return fail(QPMPC_ERR_INVALID, "peer exchange needs 2..%d shards (got %d); use the NCCL path otherwise", MAX_PEERS, W);
    CK(cudaSetDevice(h->cfg.device));
    if (!h->d_mbox) {
        const size_t bytes = peer_mbox_bytes(W, h->R, h->kp.shard_rec_stride);
        CK(cudaMalloc((void **)&h->d_mbox, bytes));
        CK(cudaMemset(h->d_mbox, 0, bytes));
        CK(cudaDeviceSynchronize());
    }
    PeerBlob b;
    memset(&b, 0, sizeof(b));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, h->cfg.device));
    memcpy(b.uuid, &prop.uuid, 16);
    CK(cudaIpcGetMemHandle(&b.ipc, h->d_mbox));
    b.pid = (uint64_t)getpid(); b.raw_ptr = (uint64_t)(uintptr_t)h->d_mbox;
    b.device = h->cfg.device; b.world = W; b.rank = h->cfg.shard_rank; b.rec_stride = h->kp.shard_rec_stride; b.robots = h->R;
    memcpy(blob, &b, sizeof(b));
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_peer_attach(qpmpc_handle *h, const unsigned char *blobs) {
    if (!h || !blobs) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (!h->d_mbox) return fail(QPMPC_ERR_STATE, "qpmpc_shard_peer_export has not been called on this handle");
    const int W = h->cfg.shard_world, me = h->cfg.shard_rank;
    CK(cudaSetDevice(h->cfg.device));
    // Every shard's kernel waits for its peers' stamps, so the W kernels must be able to run at the same time: one
    // physical GPU per shard.  (Two waiting kernels on one GPU are not guaranteed to be co-scheduled.)
    for (int a = 0; a < W; ++a)
        for (int b = a + 1; b < W; ++b)
            if (memcmp(blobs + (size_t)a * QPMPC_PEER_BLOB_BYTES + offsetof(PeerBlob, uuid),
                       blobs + (size_t)b * QPMPC_PEER_BLOB_BYTES + offsetof(PeerBlob, uuid), 16) == 0)
                return fail(QPMPC_ERR_INVALID, "shards %d and %d are on the same GPU: the peer exchange needs one GPU per shard", a, b);
    for (int w = 0; w < W; ++w) {
        PeerBlob b;
        memcpy(&b, blobs + (size_t)w * QPMPC_PEER_BLOB_BYTES, sizeof(b));
        if (b.world != W || b.rank != w || b.rec_stride != h->kp.shard_rec_stride || b.robots != h->R)
            return fail(QPMPC_ERR_INVALID, "peer blob %d does not describe shard %d of this controller", w, w);
        if (w == me) { h->peer_mbox[w] = h->d_mbox; continue; }
        if (b.pid == (uint64_t)getpid()) {              // same process (one host thread per GPU): plain peer access
            if (b.device != h->cfg.device) {
                int can = 0;
                CK(cudaDeviceCanAccessPeer(&can, h->cfg.device, b.device));
                if (!can) return fail(QPMPC_ERR_CUDA, "device %d cannot access device %d", h->cfg.device, b.device);
                cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(QPMPC_ERR_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
                (void)cudaGetLastError();
            }
            h->peer_mbox[w] = (float *)(uintptr_t)b.raw_ptr;
        } else {
            void *p = nullptr;
            CK(cudaIpcOpenMemHandle(&p, b.ipc, cudaIpcMemLazyEnablePeerAccess));
            h->peer_mbox[w] = (float *)p; h->peer_ipc[w] = true;
        }
    }
    h->peer_attached = true;
    return QPMPC_OK;
}

// Test hooks (single-GPU coverage of peer_exchange without two kernels waiting on each other): `loopback` points every
// absent peer at a sink on this device; `inject` plays a peer's part from the host BEFORE the tick is launched -- it
// writes that peer's record and the stamp of the upcoming tick into the local mailbox.
extern "C" int qpmpc_shard_peer_loopback(qpmpc_handle *h) {
    if (!h) return fail(QPMPC_ERR_INVALID, "NULL handle");
    const int W = h->cfg.shard_world;
    if (W < 2 || W > MAX_PEERS) return fail(QPMPC_ERR_INVALID, "peer exchange needs 2..%d shards (got %d)", MAX_PEERS, W);
    CK(cudaSetDevice(h->cfg.device));
    const size_t bytes = peer_mbox_bytes(W, h->R, h->kp.shard_rec_stride);
    if (!h->d_mbox) { CK(cudaMalloc((void **)&h->d_mbox, bytes)); CK(cudaMemset(h->d_mbox, 0, bytes)); }
    if (!h->d_sink) { CK(cudaMalloc((void **)&h->d_sink, bytes)); CK(cudaMemset(h->d_sink, 0, bytes)); }
    CK(cudaDeviceSynchronize());
    for (int w = 0; w < W; ++w) h->peer_mbox[w] = (w == h->cfg.shard_rank) ? h->d_mbox : h->d_sink;
    h->peer_attached = true;
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_peer_inject(qpmpc_handle *h, int rank, const float *host_rec) {
    if (!h || !host_rec) return fail(QPMPC_ERR_INVALID, "NULL argument");
    const int W = h->cfg.shard_world, R = h->R, stride = h->kp.shard_rec_stride;
    if (!h->d_mbox || rank < 0 || rank >= W || rank == h->cfg.shard_rank) return fail(QPMPC_ERR_INVALID, "bad rank / no mailbox");
    CK(cudaSetDevice(h->cfg.device));
    unsigned int seq = h->seq + 1u;
    if (seq == 0u) seq = 1u;
    const unsigned int par = seq & 1u;
    const size_t slot = (size_t)(par * W + rank) * R;
    // the peer's record as the {value, seq} units its kernel would have stored (qpmpc_tick.cuh peer_exchange)
    std::vector<uint32_t> units((size_t)R * stride * 2);
    for (size_t k = 0; k < (size_t)R * stride; ++k) { units[2 * k] = f2u(host_rec[k]); units[2 * k + 1] = seq; }
    CK(cudaMemcpy(reinterpret_cast<uint32_t *>(h->d_mbox) + slot * stride * 2, units.data(), units.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    return QPMPC_OK;
}

static int tick_peer(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out) {
    const int R = h->R;
    for (int r = 0; r < R; ++r) { int rc = pack_one(h, r, &in[r], false); if (rc) return rc; }
    KParams kp = h->kp;
    kp.fused_finish = 0;
    kp.peer_world = h->cfg.shard_world; kp.peer_rank = h->cfg.shard_rank;
    kp.peer_timeout_ns = h->peer_timeout_ns;
    memcpy(kp.peer_mbox, h->peer_mbox, sizeof(kp.peer_mbox));
    kp.seq = next_seq(h);                              // every rank ticks in lockstep: same stamp everywhere
    if (h->peer_timing) {                              // qpmpc_debug_peer_wait: %globaltimer around the exchange (slots 6, 7)
        if (!h->d_phase) CK(cudaMalloc((void **)&h->d_phase, 24 * sizeof(unsigned long long)));
        kp.phase_ns = h->d_phase;
    }
    int rc = stage_injection(h, in, R, kp);
    if (rc) return rc;
    const dim3 grid(kp.nblk, R);
    const bool plain = (kp.noise_mode == NOISE_PHILOX && !kp.freq_injected);
    if (plain && h->direct) {
        rc = direct_tick(h, kp, grid, out);
        if (rc) return rc;
    } else {
        const bool mapped = h->cfg.io_mode == QPMPC_IO_MAPPED;
        if (!mapped) CK(cudaMemcpyAsync(h->d_in, h->h_in, (size_t)R * kp.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        rc = launch(h, h->tick_fn, kp, grid);
        if (rc) return rc;
        if (!mapped) CK(cudaMemcpyAsync(h->h_out, h->d_out, (size_t)R * kp.out_stride * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        for (int r = 0; r < R; ++r) unpack_one(h, r, h->h_out, &out[r]);
    }
    if (h->peer_timing) {
        unsigned long long ph[24];
        CK(cudaMemcpyAsync(ph, h->d_phase, sizeof(ph), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        h->peer_wait_ns[0] = ph[16]; h->peer_wait_ns[1] = ph[17];
    }
    for (int r = 0; r < R; ++r)
        if (out[r].best_index == -2) return fail(QPMPC_ERR_NCCL, "peer exchange timed out: a shard of this controller did not tick (robot %d)", r);
    return QPMPC_OK;
}

extern "C" int qpmpc_shard_peer_timeout(qpmpc_handle *h, double seconds) {
    if (!h || !(seconds > 0.0) || seconds > 3.5) return fail(QPMPC_ERR_INVALID, "timeout must be in (0, 3.5] s (the host-side deadline is 4 s)");
    h->peer_timeout_ns = (unsigned long long)(seconds * 1e9);
    return QPMPC_OK;
}

// Debug: %globaltimer (this GPU's clock, ns) when the last block of the last sharded tick started pushing its record and
// when it had taken every peer's record: the difference is what the exchange costs this rank (transfer + waiting for
// the slowest rank).
extern "C" int qpmpc_debug_peer_wait(qpmpc_handle *h, int32_t enable, uint64_t out[2]) {
    if (!h) return fail(QPMPC_ERR_INVALID, "NULL handle");
    h->peer_timing = enable != 0;
    if (out) { out[0] = h->peer_wait_ns[0]; out[1] = h->peer_wait_ns[1]; }
    return QPMPC_OK;
}

extern "C" int qpmpc_tick_sharded(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out) {
    if (!h || !in || !out) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (!h->comm && !h->peer_attached) return fail(QPMPC_ERR_STATE, "neither qpmpc_shard_peer_attach nor qpmpc_shard_init_nccl has been called");
    CK(cudaSetDevice(h->cfg.device));
    if (h->peer_attached) return tick_peer(h, in, out);
    KParams kp;
    int rc = shard_partial_async(h, in, kp);
    if (rc) return rc;
    const size_t n = (size_t)h->R * kp.shard_rec_stride;
    int e = g_nccl.AllGather(h->d_send, h->d_recv, n, /*ncclFloat32*/ 7, h->comm, h->stream);   // one small collective per tick
    if (e != 0) return fail(QPMPC_ERR_NCCL, "ncclAllGather: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(e) : "error");
    rc = shard_finish_async(h, kp);
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < h->R; ++r) unpack_one(h, r, h->h_out, &out[r]);
    return QPMPC_OK;
}

// ------------------------------------------------------------------------------------------------------
// device-resident benchmarking helpers
// ------------------------------------------------------------------------------------------------------
extern "C" int qpmpc_upload_resident(qpmpc_handle *h, const qpmpc_tick_in *in) {
    if (!h || !in) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    for (int r = 0; r < h->R; ++r) { int rc = pack_one(h, r, &in[r], false); if (rc) return rc; }
    CK(cudaMemcpyAsync(h->d_in, h->h_in, (size_t)h->R * h->kp.in_stride * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->resident_uploaded = true;
    return QPMPC_OK;
}

static int bench_resident(qpmpc_handle *h, int steps, int flush_l2, float *ms, float *wall_ms);
extern "C" int qpmpc_bench_resident(qpmpc_handle *h, int steps, int flush_l2, float *ms) { return bench_resident(h, steps, flush_l2, ms, nullptr); }
// the same loop timed with the HOST wall clock instead of events (launch call -> cudaStreamSynchronize returns; the L2 flush
// has finished before the clock starts; ms[] is zeroed): sees launch latency and whatever the ranks of one box contend for
extern "C" int qpmpc_bench_resident_wall(qpmpc_handle *h, int steps, int flush_l2, float *ms, float *wall_ms) {
    if (!wall_ms) return fail(QPMPC_ERR_INVALID, "bad argument");
    return bench_resident(h, steps, flush_l2, ms, wall_ms);
}
static int bench_resident(qpmpc_handle *h, int steps, int flush_l2, float *ms, float *wall_ms) {
    if (!h || !ms || steps < 1) return fail(QPMPC_ERR_INVALID, "bad argument");
    if (!h->resident_uploaded) return fail(QPMPC_ERR_STATE, "qpmpc_upload_resident has not been called");
    if (h->cfg.shard_world != 1) return fail(QPMPC_ERR_STATE, "resident benchmarking needs an unsharded handle");
    CK(cudaSetDevice(h->cfg.device));
    const size_t flush_bytes = (size_t)256 << 20;            // > 126 MB L2
    if (flush_l2 && !h->d_flush) CK(cudaMalloc((void **)&h->d_flush, flush_bytes));
    KParams kp = h->kp;
    kp.in = h->d_in; kp.out = h->d_out;                      // inputs / outputs stay in HBM
    kp.carry_best = 1;                                       // tick i+1 starts from tick i's best with the next Philox counter
    const dim3 grid(kp.nblk, h->R);
    for (int i = 0; i < steps; ++i) {
        if (flush_l2) CK(cudaMemsetAsync(h->d_flush, i & 0xff, flush_bytes, h->stream));
        if (wall_ms) {                                       // host clock: launch call -> cudaStreamSynchronize returns (no events in
            CK(cudaStreamSynchronize(h->stream));            // the way; the flush has completed before the clock starts)
            const uint64_t t0 = mono_ns();
            int rc = launch(h, h->tick_fn, kp, grid);
            if (rc) return rc;
            CK(cudaStreamSynchronize(h->stream));
            wall_ms[i] = (float)(1e-6 * (double)(mono_ns() - t0));
            ms[i] = 0.0f;
            continue;
        }
        CK(cudaEventRecord(h->ev0, h->stream));
        int rc = launch(h, h->tick_fn, kp, grid);
        if (rc) return rc;
        CK(cudaEventRecord(h->ev1, h->stream));
        CK(cudaEventSynchronize(h->ev1));
        CK(cudaEventElapsedTime(&ms[i], h->ev0, h->ev1));
    }
    return QPMPC_OK;
}

extern "C" int qpmpc_fetch_resident_out(qpmpc_handle *h, qpmpc_tick_out *out) {
    if (!h || !out) return fail(QPMPC_ERR_INVALID, "NULL argument");
    CK(cudaSetDevice(h->cfg.device));
    std::vector<float> tmp((size_t)h->R * h->kp.out_stride);
    CK(cudaMemcpy(tmp.data(), h->d_out, tmp.size() * sizeof(float), cudaMemcpyDeviceToHost));
    for (int r = 0; r < h->R; ++r) unpack_one(h, r, tmp.data(), &out[r]);
    return QPMPC_OK;
}

extern "C" int qpmpc_debug_phase_times(qpmpc_handle *h, uint64_t out[16]) {
    if (!h || !out) return fail(QPMPC_ERR_INVALID, "NULL argument");
    if (!h->resident_uploaded) return fail(QPMPC_ERR_STATE, "qpmpc_upload_resident has not been called");
    if (h->cfg.shard_world != 1) return fail(QPMPC_ERR_STATE, "needs an unsharded handle");
    CK(cudaSetDevice(h->cfg.device));
    if (!h->d_phase) CK(cudaMalloc((void **)&h->d_phase, 24 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(h->d_phase, 0, 16 * sizeof(unsigned long long), h->stream));
    KParams kp = h->kp;
    kp.in = h->d_in; kp.out = h->d_out; kp.carry_best = 1; kp.phase_ns = h->d_phase;
    if (h->direct) { kp.host_out = h->d_dout; kp.seq = next_seq(h); }     // include the direct-I/O result publication (slot 15 -> 11)
    int rc = launch(h, h->tick_fn, kp, dim3(kp.nblk, h->R));
    if (rc) return rc;
    CK(cudaMemcpyAsync(out, h->d_phase, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_kernel_launch_count(const qpmpc_handle *h, uint64_t *count) {
    if (!h || !count) return fail(QPMPC_ERR_INVALID, "NULL argument");
    *count = h->launches;
    return QPMPC_OK;
}

extern "C" int qpmpc_measure_fp32_peak(int device, float *tflops) {
    if (!tflops) return fail(QPMPC_ERR_INVALID, "NULL argument");
    int ndev = 0;
    int rc = qpmpc_device_count(&ndev);
    if (rc) return rc;
    if (device < 0 || device >= ndev) return fail(QPMPC_ERR_INVALID, "device %d out of range", device);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    const int threads = 256, blocks = prop.multiProcessorCount * 8, iters = 4096;
    float *d = nullptr;
    CK(cudaMalloc((void **)&d, (size_t)threads * blocks * sizeof(float)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 0.f;
    for (int rep = 0; rep < 6; ++rep) {
        CK(cudaEventRecord(e0));
        fp32_peak_kernel<<<blocks, threads>>>(d, iters, 0.999f, 1e-3f);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double flop = (double)threads * blocks * (double)iters * 16.0 * 8.0 * 2.0;
        const float tf = (float)(flop / (ms * 1e-3) / 1e12);
        if (rep > 0 && tf > best) best = tf;
    }
    CK(cudaGetLastError());
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    *tflops = best;
    return QPMPC_OK;
}
