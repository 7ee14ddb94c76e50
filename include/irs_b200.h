/*
 * irs_b200.h - C ABI of the B200-native batched `CommunicationEnv.step` path.
 *
 * The reference (lasseufpa/intent_radio_sched_multi_slice) is pure Python: there is no FFI
 * today.  This header is the boundary a maintainer binds (ctypes) in
 * place of the Python call chain
 *
 *   CommunicationEnv.step      sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:271-373
 *   CommunicationEnv.reset     sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:375-482
 *   UEs.step / Buffer          sixg_radio_mgmt/sixg_radio_mgmt/ues.py:149-243, buffer.py:53-130
 *   IBSched.action_format      agents/ib_sched.py:223-349   (+ agents/common.py:442-701)
 *   IBSched.obs_space_format   agents/ib_sched.py:63-204    (+ agents/common.py:9-378)
 *   IBSched.calculate_reward   agents/ib_sched.py:206-221   (+ agents/common.py:381-439)
 *   MARR.step / MAPF.step      agents/marr.py:40-47, agents/mapf.py:41-111
 *   MultSliceTraffic.step      traffics/mult_slice.py:15-34
 *   MimicQuadriga/FixedSE.step channels/mimic_quadriga.py:30-58, channels/fixed_se.py:28-41
 *   Metrics.step               sixg_radio_mgmt/sixg_radio_mgmt/metrics.py:62-100 (as accumulators)
 *
 * Conventions: plain pointers and sizes, no C++ or torch types; every function returns an
 * int status (0 = OK, <0 = irs_status) and never throws; all device memory is owned by the
 * caller (one state blob whose size irs_state_layout reports); every device call is
 * asynchronous on the given CUDA stream (a cudaStream_t passed as void*), allocates nothing
 * and is CUDA-graph capturable.  A handle is not thread-safe (one handle per process/device,
 * like one reference env per Ray worker).
 */
#ifndef IRS_B200_H
#define IRS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IRS_MAX_UES 25        /* env_config/mult_slice.yml:14 max_number_ues */
#define IRS_MAX_SLICES 5      /* env_config/mult_slice.yml:13 */
#define IRS_NUM_RBS 135       /* env_config/mult_slice.yml:5 */
#define IRS_WINDOW 10         /* agents/ib_sched.py:49 max_obs_memory */
#define IRS_MAX_UES_SLICE 5   /* agents/ib_sched.py:50 */
#define IRS_NOALLOC_MAX_ENVS 1048576 /* irs_check_errors_noalloc: 4,096 scratch slots (16 KB) x 256 envs */
#define IRS_RBS_PER_RBG 5     /* agents/ib_sched.py:56 */
#define IRS_OBS0 50           /* agents/ib_sched.py:413-470: player_0 observations */
#define IRS_OBSI 19           /* player_{s+1} observations */
#define IRS_NUM_METRICS 16
#define IRS_POISSON_K 512      /* Poisson CDF table length: lambda + 12 sqrt(lambda) + 10 <= 512, i.e. lambda <= 310 Mbit per step */

typedef enum {
    IRS_OK = 0,
    IRS_ERR_ARG = -1,        /* bad argument / config */
    IRS_ERR_UNBOUND = -2,    /* irs_bind / irs_load_scenarios not called yet */
    IRS_ERR_CUDA = -3,       /* a CUDA runtime call failed (message in irs_last_error) */
    IRS_ERR_ENV = -4         /* irs_check_errors: a device-side invariant was violated */
} irs_status;

/* slice flags */
enum { IRS_F_ACTIVE = 1, IRS_F_REQ = 2, IRS_F_THR = 4, IRS_F_REL = 8, IRS_F_LAT = 16,
       IRS_F_EMBB = 32, IRS_F_URLLC = 64 /* use case of the slice type, sched_colran.py:352-363 */ };

/* One slice slot of a scenario: the dense form of one `slice_req["slice_s"]` entry
 * (associations/mult_slice.py:58-347) plus its UE count.  80 bytes. */
typedef struct {
    double priority, traffic, buffer_size, buffer_latency, message_size;
    double thr_value, rel_value, lat_value;
    int32_t flags, n_ues, lam_idx, pad;
} irs_slice;

/* One network scenario (what one ep_N.npz of gen_assoc_mult_slice.py:229-237 holds).  432 bytes. */
typedef struct {
    irs_slice sl[IRS_MAX_SLICES];
    int8_t ue_slice[32]; /* slice slot of each UE, -1 = not associated */
} irs_scenario;

typedef struct {
    int32_t num_envs;
    int32_t depth_cap;       /* max buffer depth (buffer_latency + 1) of any scenario row; <= 16383 (the live extent of a
                                row is kept in 14 bits of the per-UE record) */
    int32_t channel_mode;    /* 0 SE injected per call, 1 fixed_se (2.0), 2 native mimic_quadriga,
                                3 device-resident SE trace bank (irs_set_se_trace; channels/quadriga.py) */
    int32_t traffic_mode;    /* 0 traffic (bits) injected per call, 1 native Poisson */
    int32_t sort_slices;     /* IBSched.enable_sort_slices (ib_sched.py:57); MARR/MAPF: 0 */
    int32_t fixed_intra;     /* -1 per-slice action (player_1..5), 0 rr, 1 pf, 2 mt */
    int32_t policy;          /* 0 actions supplied, 1 MARR, 2 MAPF, 3 random (device RNG) */
    int32_t assoc_mode;      /* 0 scenario = episode % scen_mod (mult_slice.py:444-452),
                                1 scenario = episode / scen_div (mult_slice_seq.py:38-46),
                                2 generated on the device per (env, episode) with the distribution of the
                                  reference's generator mode (associations/mult_slice.py:359-416): the n_scen
                                  bank rows are then slice-TYPE templates (sl[0] = the type's parameters,
                                  sl[0].n_ues = max UEs, sl[0].pad = min UEs; 5 <= n_scen <= 32) */
    int32_t scen_mod, scen_div, n_scen;
    int32_t max_steps, max_episodes, initial_episode;
    int32_t auto_reset;      /* step() resets a finished env in the same call (VecEnv style) */
    int32_t env_id_base;     /* global id of env 0 of this shard: RNG key, so results do not
                                depend on how envs are sharded over GPUs */
    int32_t surface;         /* agent surface written to obs0 / reward[0]:
                                0 IBSched family (ib_sched.py:63-221; MARR, MAPF, SB3 adapters),
                                1 SchedTWC (agents/sched_twc.py:166-413), 2 SchedColORAN (agents/sched_colran.py:166-417):
                                  50-value per-slice metric vector, weighted-negative-drift / eMBB-URLLC reward, and
                                  every frame pushed TWICE into the 10-entry window, as those agents do
                                  (sched_twc.py:176-179) */
    int32_t random_episodes; /* enable_random_episodes (comm_env.py:219-226, 434-444): each env plays a shuffled cycle
                                of [initial_episode, max_episodes) instead of counting up; the order is a counter-based
                                pseudo-random permutation keyed by (seed, global env id, permutation number), or the
                                table given to irs_set_episode_order */
    uint64_t seed;
    double bandwidth;        /* Hz, env_config/mult_slice.yml:3 */
} irs_config;

/* Byte offsets of every array inside the caller-owned device state blob.  The caller may
 * view them (e.g. as torch tensors) but must treat all but `out_*` as read-only. */
typedef struct {
    size_t total_bytes;
    int32_t row_stride;      /* ints per buffer row: depth_cap rounded up to 4 */
    int32_t reserved;
    size_t buf;              /* i32 [E][25][row_stride]  age histograms (buffer.py:47) kept as RINGS: the packets of age a
                                of UE u (depth D_u = its slice's buffer_latency + 1) live in slot (k - a) mod D_u, k = the
                                env's step counter, so a step moves no packets: it reads the expiring slot, the oldest
                                slots the drain reaches, and writes slot k mod D_u.  irs_read_buffers returns the
                                age-indexed view of buffer.py */
    size_t ue_i;             /* i32 [E][25][8]  per-UE record: packets in buffer, window sums of
                                sent / dropped, newest frame's sent / dropped / incoming,
                                isclose(win[1].occ, 0), row word = path taken (0 no older packets, 1 all older packets
                                left: window cleared, 2 drain inside the lane's 4-slot look-ahead, 3 longer drain by the
                                warp) | live extent << 2 (14 bits: every age at or beyond it is empty) | 4-byte slot
                                accesses of the step << 16 */
    size_t ue_d;             /* f64 [E][25][4]  newest window frame: occupancy, mean delay, mean SE,
                                pkt_throughputs */
    size_t ring;             /* i32 [E][25][10][2] (sent, dropped) of the last 10 frames */
    size_t rb;               /* i16 [E][25][2] (rb_start, rb_count): contiguous RB run of each UE */
    size_t step, episode, scen, win_len, win_head, err; /* i32 [E] */
    size_t visit;            /* i32 [E] resets this env has seen: part of the generators' key, so that an episode NUMBER
                                played again draws fresh channel / traffic noise (the reference keeps ONE np_random stream
                                running across episodes, comm_env.py:631-679) */
    size_t ep_pos, ep_gen, ep_base; /* i32 [E] random_episodes: position in the current permutation, permutations drawn so
                                far, first episode of the permuted range (comm_env.py:219-226, 434-444) */
    size_t wsum;             /* i64 [E][25] sum(age * packets) of each buffer, updated incrementally (buffer.py:116-130) */
    size_t sorted;           /* i8  [E][5] IBSched.sorted_slices */
    size_t acc;              /* f64 [E][16] episode-metric accumulators */
    size_t bank;             /* irs_scenario [n_scen] */
    size_t cdf;              /* f64 [n_lambda][IRS_POISSON_K] Poisson CDF tables */
    size_t in_a0;            /* f64 [E][5]  staging for irs_step_host */
    size_t in_intra;         /* i32 [E][5] */
    /* the env surface written by every step/reset */
    size_t out_obs0;         /* f64 [E][50] */
    size_t out_obsi;         /* f64 [E][5][19] */
    size_t out_mask0;        /* i8  [E][5] */
    size_t out_maski;        /* i8  [E][5][5] */
    size_t out_reward;       /* f64 [E][6] */
    size_t out_done;         /* u8  [E] */
    size_t metrics_total;    /* f64 [16] result of irs_episode_metrics */
    size_t scratch;          /* i32 [64] small device-side results (irs_check_errors: first failing env, its error word) */
    size_t out32;            /* f32 [E][151] staging of the binary32 host surface (obs0 | obsi | reward per env
                                block: [E][50], then [E][95], then [E][6]) */
} irs_layout;

typedef struct {
    const double* a0;        /* [E][5] inter-slice scores in [-1,1] (player_0) */
    const int32_t* intra;    /* [E][5] intra-slice algorithm per slice (player_1..5), may be NULL */
    const uint32_t* sched_bits; /* [E][25][5] or NULL.  When given, the step takes this scheduling matrix as it is
                                (bit r of UE u's 135-bit row = RB r, words little-end first) instead of running the
                                built-in action_format: the slot for an ARBITRARY agent callback, which the caller
                                evaluates on the host (comm_env.py:300, 484-512).  a0 / intra are then ignored.  An RB
                                given to two UEs sets error bit 16 (comm_env.py:593-596).  irs_step only. */
} irs_actions;

typedef struct {
    const double* se;        /* [E][25][135] spectral efficiencies (channel_mode 0) */
    const double* traffic;   /* [E][25] offered bits (traffic_mode 0) */
} irs_inject;

/* Optional extra outputs (raw per-UE step metrics of ues.py:236-243 and the inputs actually
 * used).  Any pointer may be NULL.  Device pointers for irs_step/irs_reset. */
typedef struct {
    double* pkt_incoming;    /* [E][25] */
    double* pkt_throughputs; /* [E][25] */
    int32_t* sent;           /* [E][25] pkt_effective_thr */
    int32_t* dropped;        /* [E][25] */
    double* occ;             /* [E][25] buffer_occupancies */
    double* lat;             /* [E][25] buffer_latencies */
    int16_t* rb_start;       /* [E][25] */
    int16_t* rb_count;       /* [E][25] */
    int8_t* sorted_slices;   /* [E][5] */
    double* a0_used;         /* [E][5] */
    int32_t* intra_used;     /* [E][5] */
    double* se_used;         /* [E][25][135] */
    double* traffic_used;    /* [E][25] */
} irs_raw_outputs;

/* Host-side surface for irs_step_host / irs_reset_host (pinned memory recommended). */
typedef struct {
    double* obs0;            /* [E][50] */
    double* obsi;            /* [E][5][19] */
    int8_t* mask0;           /* [E][5] */
    int8_t* maski;           /* [E][5][5] */
    double* reward;          /* [E][6] */
    uint8_t* done;           /* [E] */
    /* Optional binary32 copies of the float64 surfaces, converted on the device (round to nearest) so
     * that half the bytes cross PCIe: RL front ends cast observations to float32 anyway, and 2^-24 is
     * inside the 1e-5 tolerance of the path.  A NULL pointer (float64 or float32) skips that copy. */
    float* obs0_f32;         /* [E][50] */
    float* obsi_f32;         /* [E][5][19] */
    float* reward_f32;       /* [E][6] */
} irs_host_surface;

typedef struct irs_env irs_env;

/* Replaces CommunicationEnv.__init__ (comm_env.py:79-269) for a batch of envs. */
int irs_create(const irs_config* cfg, int32_t n_lambda, irs_env** out);
int irs_destroy(irs_env* env);
/* Size and layout of the state blob the caller must allocate on the device (256-B aligned). */
int irs_state_layout(const irs_env* env, irs_layout* out);
/* Hand the blob to the handle and zero-initialise it (comm_env.py:616-682 create_scenario). */
int irs_bind(irs_env* env, void* state_dev, size_t bytes, void* stream);
/* Upload the scenario bank and Poisson tables (host pointers). Replaces
 * MultSliceAssociation.load_episode_data (associations/mult_slice.py:490-508). */
int irs_load_scenarios(irs_env* env, const irs_scenario* bank_host, int32_t n_scen,
                       const double* poisson_cdf_host, int32_t n_lambda, void* stream);
/* CommunicationEnv.reset (comm_env.py:375-482).  mask_dev: u8[E] or NULL (= all);
 * initial_episode: options["initial_episode"] (-1 = keep counting);
 * initial_episode_dev: optional i32[E] per-env override (NULL = use the scalar). */
int irs_reset(irs_env* env, const uint8_t* mask_dev, int32_t initial_episode,
              const int32_t* initial_episode_dev, const irs_inject* inject,
              const irs_raw_outputs* raw, void* stream);
/* CommunicationEnv.step (comm_env.py:271-373) for every env of the batch. */
int irs_step(irs_env* env, const irs_actions* actions_dev, const irs_inject* inject_dev,
             const irs_raw_outputs* raw_dev, void* stream);
/* irs_step with CUDA events recorded around each of its kernels; waits for the step and returns
 * their durations in ms: [0] action->RB, [1] per-UE row kernel, [2] obs/reward/rollover.
 * Measurement aid for bench.py (roofline of the dominant kernel); not graph capturable. */
int irs_step_profiled(irs_env* env, const irs_actions* actions_dev, const irs_inject* inject_dev,
                      const irs_raw_outputs* raw_dev, void* stream, float* kernel_ms_host);
/* gym's reset(seed=...) (comm_env.py:398-413): re-key the native generators.  Takes effect for
 * the next reset/step; does not touch the state. */
int irs_set_seed(irs_env* env, uint64_t seed);

/* channel_mode 3 - QuadrigaChannel / QuadrigaChannelSeq (channels/quadriga.py:38-87,
 * channels/quadriga_seq.py:28-39) as a device-resident trace: `se_bank_dev` is f64
 * [n_episodes][n_steps][25][135], owned by the caller and kept alive while the env uses it.  Slot i of the bank
 * holds the file of env episode first_episode + i (the reference opens a new target_cell_power.mat per episode,
 * quadriga.py:46-55, 78-87).  An env in episode `ep` at step counter `k` reads frame
 * [ep - first_episode][k % n_steps]: like the reference, reset and the first step of an episode both read frame 0
 * (comm_env.py:303-305,449-451 call channel.step with the step counter before it is incremented).  An episode
 * outside the bank sets the env's error word (bit 8, irs_check_errors) unless wrap != 0, which maps episodes onto
 * slots modulo n_episodes (synthetic banks: benchmarks). */
int irs_set_se_trace(irs_env* env, const double* se_bank_dev, int32_t n_episodes, int32_t n_steps, int32_t first_episode,
                     int32_t wrap);

/* The per-frame transform of QuadrigaChannel.step (channels/quadriga.py:56-74):
 *   SE = log2(1 + ((tx_power / 135) * p) / (0 + noise_power)),
 * over `n_frames` frames of 25 x 135 powers.  layout 0: each input frame is [135][25] (UE fastest:
 * the order h5py sees in target_cell_power.mat, which the reference transposes), layout 1: [25][135].
 * Output frames are always [25][135].  Device pointers; asynchronous on `stream`. */
int irs_power_to_se(const double* power_dev, double* se_dev, int64_t n_frames, int32_t layout, double tx_power,
                    double noise_power, void* stream);
/* n steps back to back with the on-device policy (cfg.policy != 0) and native generators. */
int irs_run(irs_env* env, int32_t n_steps, void* stream);
/* Same as irs_step but actions come from host memory and the env surface is copied back to
 * host memory; returns after the stream has drained (the call a single-process user makes). */
int irs_step_host(irs_env* env, const irs_actions* actions_host, const irs_host_surface* out_host,
                  void* stream);
int irs_reset_host(irs_env* env, int32_t initial_episode, const irs_host_surface* out_host, void* stream);
/* Sum the per-env accumulators in env order into layout.metrics_total (f64[16]) and, when
 * host_out != NULL, copy them to the host (synchronises).  Replaces Metrics.save
 * (metrics.py:73-100) as the source of episode statistics. */
int irs_episode_metrics(irs_env* env, double* host_out, void* stream);
/* debug=True of the reference (comm_env.py:313-319): the error words are scanned ON THE DEVICE (first env with a
 * non-zero word) and two integers come back; synchronises; returns IRS_ERR_ENV, the env index and its error word
 * (bits: 1 invalid intra-slice action, 2 episode out of range, 4 RB sum, 8 episode outside the SE bank, 16 RB given
 * to two UEs). */
int irs_check_errors(irs_env* env, int32_t* first_env, int32_t* code, void* stream);
/* random_episodes with a GIVEN order (parity runs replay the permutation the reference drew from its np_random):
 * order_dev is i32 [E][n], caller-owned device memory kept alive while in use; env e plays order_dev[e][pos % n].
 * The call also moves every env to the table's entry at its current position (position 0 right after irs_bind:
 * comm_env.py:225).  A reset with initial_episode != -1 restarts at position 0 (the reference draws a new
 * permutation there, comm_env.py:435-442: upload it before that reset).  NULL returns to the built-in permutation. */
int irs_set_episode_order(irs_env* env, const int32_t* order_dev, int32_t n, void* stream);
/* Buffer.buffer of every UE as the reference indexes it (buffer.py:47: index = age in steps): i32
 * [E][25][row_stride] into caller-owned device memory, de-rotating the ring rows.  Asynchronous on `stream`. */
int irs_read_buffers(irs_env* env, int32_t* out_dev, void* stream);
/* No counterpart in the reference: added for callers that cannot allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The error scan
 * of irs_check_errors on caller-passed arrays, with its scratch in a file-scope static __device__ array of the library
 * (16 KB: one i32 slot per 256 envs; every GPU that loads the library holds it until exit) instead of the state blob:
 * err_dev = i32 [num_envs] error words (layout.err of a bound state blob, or any array of that shape); out_dev = i32 [2]
 * in device memory <- (index of the first env with a non-zero word, that word), or (-1, 0) when there is none.
 * Asynchronous on `stream`; allocates nothing (no cudaMalloc / cudaFree / allocating library call) and does not
 * synchronise.  num_envs > IRS_NOALLOC_MAX_ENVS is rejected with IRS_ERR_ARG and a message in errbuf (optional, may be
 * NULL).  The scratch is shared by every call on the device and keeps the last call's contents; a call overwrites every
 * slot it reads, so nothing needs clearing between calls, but calls that OVERLAP on the GPU (different streams) race:
 * serialise them. */
int irs_check_errors_noalloc(const int32_t* err_dev, int32_t num_envs, int32_t* out_dev, void* stream, char* errbuf,
                             size_t errbuf_len);
const char* irs_last_error(const irs_env* env);
int irs_version(void);

#ifdef __cplusplus
}
#endif
#endif /* IRS_B200_H */
