/*
 * qpmpc.h -- C ABI of libqpmpc_b200.so: the B200-native (sm_100a, hand-written CUDA) replacement for the
 * sampling-based centroidal MPC hot path of iit-DLSLab/Quadruped-PyMPC.
 *
 * Reference surface replaced (paths relative to the reference's quadruped_pympc/):
 *   S = controllers/sampling/centroidal_nmpc_jax.py            (Sampling_MPC, fixed gait)
 *   A = controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py (Sampling_MPC, gait adaptive)
 *   M = controllers/sampling/centroidal_model_jax.py           (Centroidal_Model_JAX)
 *   G = helpers/periodic_gait_generator_jax.py                 (PeriodicGaitGeneratorJax)
 *   I = interfaces/srbd_controller_interface.py                (the caller; unchanged)
 *
 * Plain C: pointers + sizes, no torch / C++ types.  All functions return QPMPC_OK (0) or a negative
 * qpmpc_status; qpmpc_last_error() gives the message of the last failure on the calling thread.
 * There is NO CPU fallback: without a CUDA device every compute entry point fails with QPMPC_ERR_CUDA.
 * A handle is not thread-safe (one tick in flight per controller, like the reference, I:118-168) and must be
 * created in the process that ticks it (CUDA contexts do not survive fork(); ros2/run_controller.py:225-243).
 */
#ifndef QPMPC_H
#define QPMPC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QPMPC_ABI_VERSION 2
#define QPMPC_MAX_HORIZON 64
#define QPMPC_MAX_SPLINES 16
#define QPMPC_MAX_ELITE 256
#define QPMPC_STATE_DIM 24  /* S:35 */

typedef enum {
    QPMPC_OK = 0,
    QPMPC_ERR_INVALID = -1,   /* bad argument / configuration (reference: sys.exit(1) on unknown method, S:107-110) */
    QPMPC_ERR_CUDA = -2,      /* CUDA runtime failure or no device: the product path has no CPU fallback */
    QPMPC_ERR_NCCL = -3,      /* NCCL unavailable or failed (sample-sharded controllers only) */
    QPMPC_ERR_STATE = -4      /* call sequence error (e.g. finish without partial) */
} qpmpc_status;

/* mpc_params['control_parametrization'], config.py:185; S:52-93 */
typedef enum { QPMPC_ZERO_ORDER = 0, QPMPC_LINEAR_SPLINE = 1, QPMPC_CUBIC_SPLINE = 2 } qpmpc_parametrization;
/* mpc_params['sampling_method'], config.py:184; S:96-110 */
typedef enum { QPMPC_RANDOM_SAMPLING = 0, QPMPC_MPPI = 1, QPMPC_CEM_MPPI = 2 } qpmpc_method;

/* Host<->device transport of the per-tick packet (inputs ~1 KB, outputs ~1 KB). */
typedef enum {
    QPMPC_IO_MEMCPY = 0,  /* pinned staging + cudaMemcpyAsync H2D / D2H around the kernel */
    QPMPC_IO_MAPPED = 1,  /* kernel reads/writes mapped pinned host memory directly (zero-copy over PCIe) */
    QPMPC_IO_DIRECT = 2   /* default for one controller: the tick packet travels in the kernel's parameter space, the last
                             block stores the result packet to mapped pinned host memory as self-validating 8-byte
                             {value, tick sequence} units and the host takes each unit when its sequence matches -- no
                             copy-engine hop, no system-scope fence and no stream synchronisation around the kernel.
                             Falls back to QPMPC_IO_MEMCPY for batched / sharded handles and oversized packets. */
} qpmpc_io_mode;

/* Everything Sampling_MPC.__init__ (S:23-177 / A:24-187) and Centroidal_Model_JAX.__init__ (M:20-65) freeze. */
typedef struct qpmpc_config {
    int32_t abi_version;          /* = QPMPC_ABI_VERSION */
    int32_t num_samples;          /* N = mpc_params['num_parallel_computations'], config.py:187 */
    int32_t horizon;              /* H, config.py:92 (<= QPMPC_MAX_HORIZON) */
    int32_t num_splines;          /* config.py:186 */
    int32_t parametrization;      /* qpmpc_parametrization */
    int32_t method;               /* qpmpc_method */
    int32_t adaptive;             /* mpc_params['optimize_step_freq'], config.py:108 -> A instead of S (I:77-83) */
    int32_t num_robots;           /* R >= 1: independent controllers batched in one launch (simulation/batched_simulations.py) */
    int32_t device;               /* CUDA device ordinal */
    int32_t io_mode;              /* qpmpc_io_mode */
    int32_t use_graph;            /* 1: capture (H2D, kernel, D2H) in a CUDA graph and replay it per tick */
    int32_t shard_rank;           /* sample-split controller: this handle owns a slice of the N samples */
    int32_t shard_world;          /* 1 = not sharded */
    int32_t elite_k;              /* 0 = reference default (10 fixed gait S:1075, 200 adaptive A:1115) */
    uint32_t seed;                /* Philox key word 0 (the reference seeds PRNGKey(42), S:167) */
    float mass;                   /* config.mass, M:38 */
    float inertia[9];             /* config.inertia row-major, M:39 */
    float dts[QPMPC_MAX_HORIZON]; /* M:42-53 (uniform dt or fine-grained head) */
    float mu;                     /* S:160 */
    float grf_min, grf_max;       /* S:163-164 */
    float q_diag[12];             /* S:118-130 (entries 12..23 of Q are zero) */
    float sigma_random_sampling[3]; /* config.py:193 */
    float sigma_mppi;             /* config.py:192 */
    float step_freq_available[3]; /* config.py:109 */
    float pgg_duty_factor;        /* 0.65 hard-coded in A:179 */
    int32_t robot_base;           /* global id of this handle's robot 0 (robot-partitioned multi-GPU runs; Philox key word 1) */
    int32_t lanes_per_sample;     /* kernel variant: 1 = one sample per thread (throughput), 4 = four lanes per sample, one
                                     leg each (latency: one small controller), 14 = one sample per thread plus three
                                     noise-helper threads per sample (latency: one controller of up to 148 x 128 samples;
                                     falls back to 1 when the launch does not qualify), 2 = two samples per thread in
                                     packed FP32 (FFMA2; throughput: batches and large sample counts, fixed gait only;
                                     falls back to 1 for gait-adaptive controllers or when two of its 256-sample tiles do
                                     not fit one SM), 0 = choose by samples per launch */
} qpmpc_config;

/* One robot's inputs for one tick = the arguments of jitted_compute_control (I:140-168) */
typedef struct qpmpc_tick_in {
    const float *state;            /* 24: [pos, lin vel, rpy, ang vel (base), foot FL, FR, RL, RR], S:575-586 */
    const float *reference;        /* 24, same order, S:597-608 */
    const float *contact_sequence; /* 4 rows (FL,FR,RL,RR) x >=H columns, values exactly 0 or 1 */
    int32_t contact_ld;            /* row stride of contact_sequence in floats (caller may pass a (4,2H) table) */
    const float *best;             /* P: best_control_parameters going in (S:634) */
    const float *sigma;            /* P: cem_mppi sigma (S:941); NULL otherwise */
    const float *phase_signal;     /* 4: pgg_phase_signal ("timing", A:326); may be NULL when not adaptive */
    float nominal_step_freq;       /* A:637 */
    int32_t optimize_swing;        /* A:638 */
    uint32_t key[2];               /* (seed, counter): the controller's master_key (S:498-501) */
    /* parity hooks (north_star: "with the reference's noise tensor injected"); NULL in production */
    const float *noise;            /* N x P additional_random_parameters (S:647-677), row-major, host memory */
    const float *step_freqs;       /* N per-sample step frequencies (A:688-692), host memory */
} qpmpc_tick_in;

/* One robot's outputs = the tuple jitted_compute_control returns (S:786-787, S:1085-1094, A:798-806) */
typedef struct qpmpc_tick_out {
    float grfs[12];               /* nmpc_GRFs [FL xyz, FR, RL, RR], S:747 */
    float predicted_state[24];    /* nmpc_predicted_state, S:784 */
    float best_cost;              /* S:691 */
    float best_freq;              /* 1.4 / 1.65 constants (S:786,1083) or the best sample's frequency (A:705) */
    int32_t best_index;           /* nanargmin(costs), S:690 (global sample index) */
    int32_t num_elite;            /* cem_mppi: number of elite indices written */
    float *best;                  /* P, caller-allocated: best_control_parameters coming out */
    float *sigma_new;             /* P, caller-allocated or NULL: new_sigma_cem_mppi, S:1077-1081 */
    int32_t *elite_indices;       /* elite_k, caller-allocated or NULL: argsort(costs)[:k], S:1075 */
} qpmpc_tick_out;

typedef struct qpmpc_handle qpmpc_handle;

/* ---- lifecycle -------------------------------------------------------------------------------------- */
int qpmpc_abi_version(void);
const char *qpmpc_last_error(void);
int qpmpc_device_count(int *count);                       /* QPMPC_ERR_CUDA when no driver / device */
void qpmpc_config_defaults(qpmpc_config *cfg);            /* config.py defaults for robot 'go2' */
int qpmpc_create(const qpmpc_config *cfg, qpmpc_handle **out);   /* replaces Sampling_MPC.__init__, S:23-177 */
void qpmpc_destroy(qpmpc_handle *h);
int qpmpc_num_parameters(const qpmpc_handle *h);          /* P = num_control_parameters, S:52-93 */
int qpmpc_num_local_samples(const qpmpc_handle *h);       /* samples evaluated by this shard */
const char *qpmpc_kernel_name(const qpmpc_handle *h);     /* tick kernel variant the handle launches (lanes_per_sample) */

/* ---- the tick: replaces Sampling_MPC.jitted_compute_control (S:112, called at I:140/159) ------------- */
int qpmpc_tick(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out);
/* The sampling-iteration loop of one MPC tick (srbd_controller_interface.py:125-168, num_sampling_iterations > 1) on the
 * device: iteration i+1 starts from iteration i's best_control_parameters (cem_mppi: and its new sigma) with key[1] + i
 * as the Philox counter.  `in` describes iteration 0; `out` is the last iteration's result.  All robots of the handle
 * (R >= 1).  Equal, bit for bit, to `iterations` calls of qpmpc_tick[_batched] chained by the host. */
int qpmpc_tick_iterations(qpmpc_handle *h, const qpmpc_tick_in *in, int32_t iterations, qpmpc_tick_out *out);
/* ---- Stateful controller: the whole sampling block of SRBDControllerInterface.compute_control
 * (srbd_controller_interface.py:118-168) in one call.  best_control_parameters, sigma, the key counter and
 * previous_contact_mpc persist in a device record between calls (initially zeros, -, key = (cfg.seed, 1),
 * [1,1,1,1] like srbd_controller_interface.py:19).  `in[r].state` is the RAW 24-state (no foothold substitution);
 * in[r].best / sigma / key are ignored.  Per call, on the device: prepare_state_and_reference (swing feet <- reference
 * footholds, knots of a leg that just lifted off zeroed, centroidal_nmpc_jax.py:563-627), then `iterations` sampling
 * iterations (with_newkey / best / with_newsigma carried; cem_mppi restarts from `sigma0` at iteration 0 like
 * srbd_controller_interface.py:128-129), then previous_contact <- contact_sequence[:, 0].  `out` = last iteration.
 * shift_solution: see qpmpc_step_set_shift.
 * Error behaviour: invalid inputs are rejected before anything is launched (no robot advances).  A CUDA error in the
 * middle of a chunk-pipelined batch (R >= 64) drains all streams and returns QPMPC_ERR_CUDA; the device records of the
 * chunks already launched HAVE advanced, so the caller must qpmpc_step_reset before stepping again (the error string
 * says so). */
int qpmpc_step(qpmpc_handle *h, const qpmpc_tick_in *in, int32_t iterations, float sigma0, qpmpc_tick_out *out);
/* shift_solution (mpc_params['shift_solution'], config.py:194) for the stateful controller: step > 0 makes every
 * qpmpc_step* first write the carried best's spline value at `step` (= 1 / mpc_frequency, S:571-573) into knots [0], [2],
 * [4] of every leg exactly as S:513-561 does, before the lift-off zeroing; 0 = off (default).  Refused for gait-adaptive
 * handles (the reference's adaptive shift_solution raises, A:540-552). */
int qpmpc_step_set_shift(qpmpc_handle *h, float step);
/* (re)initialise the records: key = (seed, counter of the NEXT iteration), best (R x P) or NULL = zeros,
 * previous_contact (R x 4) or NULL = ones */
int qpmpc_step_reset(qpmpc_handle *h, const uint32_t key[2], const float *best, const int32_t *previous_contact);
/* read one robot's record (any pointer may be NULL) */
int qpmpc_step_state(qpmpc_handle *h, int robot, float *best, float *sigma, int32_t previous_contact[4], uint32_t key[2]);
/* R robots in one launch (replaces "one process per robot", simulation/batched_simulations.py:43-58).
 * in / out are arrays of num_robots structs; parity hooks noise/step_freqs are honoured per robot. */
int qpmpc_tick_batched(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out);
/* The same for callers that keep the R robots in flat row-major arrays (states R x 24, references R x 24, contacts
 * R x 4 x contact_ld, best / sigma R x P, phase_signals R x 4, per-robot nominal_step_freqs / optimize_swing; outputs
 * grfs R x 12, predicted_states R x 24, best_out R x P, best_costs / best_freqs / best_indices R, sigma_new R x P).
 * Optional pointers may be NULL.  One key (seed, counter) for the whole batch; robots differ in Philox key word 1. */
int qpmpc_tick_batched_flat(qpmpc_handle *h, const float *states, const float *references, const float *contacts,
                            int32_t contact_ld, const float *best, const float *sigma, const float *phase_signals,
                            const float *nominal_step_freqs, const int32_t *optimize_swing, const uint32_t key[2],
                            float *grfs, float *predicted_states, float *best_out, float *best_costs, float *best_freqs,
                            int32_t *best_indices, float *sigma_new);
/* qpmpc_step for R robots in flat arrays (states RAW, R x 24; outputs like qpmpc_tick_batched_flat) */
int qpmpc_step_flat(qpmpc_handle *h, const float *states, const float *references, const float *contacts,
                    int32_t contact_ld, const float *phase_signals, const float *nominal_step_freqs,
                    const int32_t *optimize_swing, int32_t iterations, float sigma0, float *grfs,
                    float *predicted_states, float *best_out, float *best_costs, float *best_freqs,
                    int32_t *best_indices, float *sigma_new);
/* Contact tables from the gait phase clocks, on the device (first half): the horizon look-ahead of
 * PeriodicGaitGenerator.compute_contact_sequence (helpers/periodic_gait_generator.py:93-118; run() :48-76 per leg, in
 * float64 like the numpy original) for all R robots.  phase_signal R x 4 (pgg._phase_signal), init R x 4 or NULL
 * (pgg._init), step_freq R, duty_factor / phase_offset[4] of the gait, (dts, lengths, n_dts) =
 * contact_sequence_dts / contact_sequence_lenghts.  The R x 4 x H tables stay on the device and are used by the NEXT
 * qpmpc_tick* / qpmpc_step* call whose inputs pass contact_sequence = NULL (once); `tables` (host, R x 4 x H) receives a
 * copy unless NULL.  FULL_STANCE gaits need no call: pass tables of ones. */
int qpmpc_pgg_contact_sequence(qpmpc_handle *h, const double *phase_signal, const uint8_t *init, const double *step_freq,
                               double duty_factor, const double phase_offset[4], const double *dts,
                               const int32_t *lengths, int32_t n_dts, float *tables);
/* Reference footholds of the R robots, on the device (second half): FootholdReferenceGenerator.
 * compute_footholds_reference (helpers/foothold_reference_generator.py:53-157; call site interfaces/wb_interface.py:
 * 231-238) in float64: base_position R x 3, base_ori_euler_xyz R x 3, base_xy_lin_vel / ref_base_xy_lin_vel R x 2 (world
 * frame), hips_position R x 4 x 3 (FL FR RL RR), lift_off_z R x 4 (frg.lift_off_positions[leg][2]), stance_time R
 * (frg.stance_time); com_height_nominal = simulation_params['ref_z'], hip_height / hip_offset (0.1, :44) / gravity
 * (config.gravity_constant) of the generator, com_pos_offset_b (3) or NULL = zeros.  The generator's state -- the deque
 * of the last 20 horizontal-frame base velocities (:16,27,98-99) -- is kept per robot on the device; reset_history != 0
 * empties it before this call (= a freshly constructed generator).  ref_feet: R x 4 x 3 (host), world frame. */
int qpmpc_footholds_reference(qpmpc_handle *h, const double *base_position, const double *base_ori_euler_xyz,
                              const double *base_xy_lin_vel, const double *ref_base_xy_lin_vel,
                              const double *hips_position, const double *lift_off_z, const double *stance_time,
                              double com_height_nominal, double hip_height, double hip_offset, double gravity,
                              const double *com_pos_offset_b, int32_t reset_history, double *ref_feet);
/* MPC -> WBC output record: the 75-double shared-memory payload of ros2/run_controller.py:72-92
 * (GRF 0..11 | footholds 12..23 | joints pos / vel / acc 24..59 | predicted state 60..71 | best_sample_freq 72 |
 * last_mpc_loop_time 73 | stamp_mono 74) written under the reference's seqlock (:373-388: *seq odd while the payload is
 * being written, even when stable; reader :587-602).  `record` = the SharedMemory buffer, `seq` = the 64-bit
 * multiprocessing.Value('Q').  qpmpc_shm_pack packs caller-supplied values (no GPU, no handle): grfs (12) are multiplied
 * by current_contact (4; NULL = ones) as interfaces/srbd_controller_interface.py:225-230 does, footholds (12 doubles;
 * NULL = zeros) are the reference footholds the interface passes through (:170-175), the joint slots are NaN (the
 * sampling controller returns None there and numpy stores None as NaN), predicted_state's first 12 entries are kept,
 * the stamp is CLOCK_MONOTONIC seconds.  qpmpc_publish_shm does the same for robot `robot` of the handle straight from
 * the pinned host packet the last tick / step landed in -- no Python objects between the kernel's result and the record. */
#define QPMPC_SHM_DOUBLES 75
int qpmpc_shm_pack(double *record, uint64_t *seq, const float *grfs, const float *current_contact,
                   const double *footholds, const float *predicted_state, double best_sample_freq,
                   double last_mpc_loop_time);
int qpmpc_publish_shm(qpmpc_handle *h, int robot, double *record, uint64_t *seq, const float *current_contact,
                      const double *footholds, double last_mpc_loop_time);
/* Batched epilogue on the device-resident results of the last tick / step (all R robots of the handle):
 *   grfs_contact[r][3 leg + k] = GRFs * current_contact[leg]                  (srbd_controller_interface.py:225-230)
 *   tau[r][3 leg + j] = -(J_leg^T grfs_contact_leg)[j]                        (wb_interface.py:372-375)
 * feet_jac: R x 4 x 3 x 3, leg order FL FR RL RR, J_leg = feet_jac.leg[:, legs_qvel_idx.leg] row-major;
 * current_contact: R x 4.  Outputs R x 12 each (either may be NULL). */
int qpmpc_stance_torques(qpmpc_handle *h, const float *feet_jac, const float *current_contact, float *grfs_contact,
                         float *tau);
/* qpmpc_stance_torques for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arguments, same results; the kernel's staging area (64 floats per robot) is a file-scope static device array of
 * QPMPC_NO_ALLOC_MAX_ROBOTS robots (256 KiB held by every GPU that loads the library, until exit) instead of the handle's
 * lazily allocated buffer, and the call issues no cudaMalloc / cudaFree.  A handle with more robots is rejected with
 * QPMPC_ERR_INVALID.  All calls on one device share the array: calls that overlap on the GPU race on it. */
#define QPMPC_NO_ALLOC_MAX_ROBOTS 1024
int qpmpc_stance_torques_static(qpmpc_handle *h, const float *feet_jac, const float *current_contact, float *grfs_contact,
                                float *tau);
/* costs (N) of the last tick of robot r, fetched lazily (the interface ignores them, I:138,158) */
int qpmpc_get_costs(qpmpc_handle *h, int robot, float *dst);

/* ---- the RNG-free seam: replaces jit_vectorized_rollout (S:176-177 / A:186-187) ----------------------- */
/* params: N x P explicit control parameters (host); step_freqs: N or NULL; costs: N (raw, not saturated) */
int qpmpc_rollout(qpmpc_handle *h, const float *state, const float *reference, const float *contact_sequence,
                  int32_t contact_ld, const float *phase_signal, const float *params, const float *step_freqs,
                  float *costs);

/* ---- noise generator export (debug / parity): the (rows x P) additional_random_parameters the tick with
 * key (seed,counter) would draw for robot r; sigma (P) only for cem_mppi.  Same device code as the tick. */
int qpmpc_generate_noise(qpmpc_handle *h, int robot, const uint32_t key[2], const float *sigma, float *dst);
/* per-sample step frequencies (N) the adaptive tick would draw (A:688-692 | A:835-838 | A:994-1000) */
int qpmpc_generate_step_freqs(qpmpc_handle *h, int robot, const uint32_t key[2], float nominal_step_freq,
                              int32_t optimize_swing, float *dst);

/* ---- sample-sharded controller, one shard per GPU.  partial = rollouts + local reduction into a
 * device-resident record; finish = merge `world` records (gathered by NCCL or by the caller) + tail. -------- */
int qpmpc_shard_record_floats(const qpmpc_handle *h);     /* floats per shard record */
int qpmpc_shard_partial(qpmpc_handle *h, const qpmpc_tick_in *in);           /* async on the handle's stream */
int qpmpc_shard_record_device_ptr(qpmpc_handle *h, void **send, void **recv); /* device buffers: 1 / world records */
int qpmpc_shard_finish(qpmpc_handle *h, qpmpc_tick_out *out);                /* merges recv[0..world) */
int qpmpc_shard_copy_record(qpmpc_handle *h, int slot, const float *host_rec); /* tests: fill recv[slot] from host */
int qpmpc_shard_get_record(qpmpc_handle *h, float *host_rec);                /* tests: read send record */
/* NCCL plumbing (dlopen of libnccl.so.2; no link-time dependency).  id: 128-byte ncclUniqueId. */
int qpmpc_nccl_unique_id(const char *libnccl_path, unsigned char id[128]);
int qpmpc_shard_init_nccl(qpmpc_handle *h, const char *libnccl_path, const unsigned char id[128]);
/* Peer exchange (2..8 shards on one NVSwitch node): the collective is fused into the tick kernel.  Each rank exports
 * a 128-byte blob describing its mailbox (CUDA IPC handle; or the raw pointer when the ranks are threads of one
 * process), the launcher all-gathers the blobs (torch.distributed / MPI / a pipe), every rank attaches all of them in
 * rank order.  From then on qpmpc_tick_sharded is ONE launch per rank: rollouts + local reduction, the last block
 * stores its record into every peer's mailbox over NVLink, waits for the peers' stamps, merges in rank order and
 * writes the result -- identical on every rank.  All ranks must tick in lockstep with identical inputs; a rank whose
 * peers stay silent for ~2 s gets QPMPC_ERR_NCCL. */
#define QPMPC_PEER_BLOB_BYTES 128
int qpmpc_shard_peer_export(qpmpc_handle *h, unsigned char blob[QPMPC_PEER_BLOB_BYTES]);
int qpmpc_shard_peer_attach(qpmpc_handle *h, const unsigned char *blobs /* world x QPMPC_PEER_BLOB_BYTES */);
/* Test hooks: cover the fused exchange on ONE GPU without two kernels waiting on each other.  loopback: absent peers
 * are a sink on this device.  inject: play peer `rank` from the host before the tick is launched (its record and the
 * stamp of the upcoming tick go into the local mailbox). */
int qpmpc_shard_peer_loopback(qpmpc_handle *h);
int qpmpc_shard_peer_inject(qpmpc_handle *h, int rank, const float *host_rec);
/* peer exchange if attached (one launch), else partial + ncclAllGather + finish */
int qpmpc_tick_sharded(qpmpc_handle *h, const qpmpc_tick_in *in, qpmpc_tick_out *out);
/* How long the in-kernel exchange waits for the peers' records before the tick fails with QPMPC_ERR_NCCL (default 2 s). */
int qpmpc_shard_peer_timeout(qpmpc_handle *h, double seconds);
/* Debug: with enable != 0 every following peer-exchange tick records this GPU's %globaltimer (ns) when its last block
 * starts pushing the shard record (out[0]) and when it has taken all peers' records (out[1]); out = the last tick's pair. */
int qpmpc_debug_peer_wait(qpmpc_handle *h, int32_t enable, uint64_t out[2]);

/* ---- device-resident benchmarking helpers (inputs already in HBM; CUDA-event timing on the handle's stream) */
/* Upload one packet per robot and keep it resident; subsequent qpmpc_bench_resident calls replay it. */
int qpmpc_upload_resident(qpmpc_handle *h, const qpmpc_tick_in *in);
/* Runs `steps` resident ticks; ms[i] = device time of tick i (events around the kernel(s) only).  If
 * flush_l2 != 0 a buffer larger than L2 is rewritten before every timed tick (outside the events). */
int qpmpc_bench_resident(qpmpc_handle *h, int steps, int flush_l2, float *ms);
/* The same loop timed with the host wall clock instead of events: wall_ms[i] = launch call of tick i until
 * cudaStreamSynchronize returns (the L2 flush has completed before the clock starts; ms[] is zeroed): includes launch
 * latency and anything the ranks of one box contend for. */
int qpmpc_bench_resident_wall(qpmpc_handle *h, int steps, int flush_l2, float *ms, float *wall_ms);
int qpmpc_fetch_resident_out(qpmpc_handle *h, qpmpc_tick_out *out);  /* D2H of the last resident tick's outputs */
int qpmpc_kernel_launch_count(const qpmpc_handle *h, uint64_t *count); /* kernels launched by this handle so far */
/* Debug: runs ONE resident tick with %globaltimer stamps (ns) taken at phase boundaries.  out[0..5]: block 0 of robot 0
 * (start, inputs staged, noise drawn, rollouts done, tile reduced, record published); out[8..11]: the last block of robot
 * 0 (ticket taken, block records merged, -, result packet written).  Slots never reached stay 0. */
int qpmpc_debug_phase_times(qpmpc_handle *h, uint64_t out[16]);
/* host-side timeline of the last direct-I/O tick (CLOCK_MONOTONIC ns): call entered, kernel launched, result stamp
 * seen, result unpacked.  enable != 0 switches the four clock reads on for the following ticks; out may be NULL. */
int qpmpc_debug_host_times(qpmpc_handle *h, int32_t enable, uint64_t out[4]);
/* FP32 FMA peak of the device (dependent-free FFMA streams on every SM), TFLOP/s; sm_mhz may be NULL */
int qpmpc_measure_fp32_peak(int device, float *tflops);

#ifdef __cplusplus
}
#endif
#endif /* QPMPC_H */
