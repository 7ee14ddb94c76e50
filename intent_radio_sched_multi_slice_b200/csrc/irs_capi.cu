// irs_capi.cu - the C ABI declared in include/irs_b200.h: host-side handle, state-blob layout
// and kernel launches.  No allocation, no synchronisation and no host<->device copy happens
// in irs_step / irs_reset / irs_run (they are CUDA-graph capturable); the *_host variants add
// the copies a single-process caller with host buffers needs.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <limits.h>
#include <mutex>
#include <new>

#include <nvtx3/nvToolsExt.h> // header-only; a no-op unless a profiler is attached
#include <time.h>
static double wall_ms() { timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec * 1e-6; }

#include "irs_action.cuh"
#include "irs_env.cuh"
#include "irs_ue.cuh"
#include "irs_fused.cuh"
#include "irs_serial.cuh"
#include "irs_lane8.cuh"
#include "irs_alt.cuh"

#define IRS_MAX_HOST_CHUNKS 16

struct irs_env {
    irs_config cfg;
    irs_layout lay;
    int32_t n_lambda;
    bool bound, loaded;
    DevState st;
    char* blob;
    cudaEvent_t ev[4];
    bool have_ev;
    cudaStream_t copy_stream;      // irs_step_host: device->host copies of finished chunks
    cudaStream_t side_stream;      // irs_step_host: every other chunk computes here, so that the kernels of two chunks overlap
    cudaEvent_t fork_ev, join_ev;
    int host_streams;              // IRS_HOST_STREAMS (1 or 2 compute streams; default 1)

    cudaEvent_t chunk_ev[IRS_MAX_HOST_CHUNKS];
    bool have_copy;
    int ue_split;                  // IRS_UE_SPLIT: log2 of the warps per env in k_ue_env (-1 = by batch size)
    int host_chunks;               // IRS_HOST_CHUNKS (0 = four equal chunks from 16,384 envs)
    int fused_max;                 // IRS_FUSED_MAX_ENVS (default 2,368): launches of at most this many envs run the whole
                                   // step in ONE kernel (k_env_step) - small batches are bound by launch and dependency
                                   // latency, not by throughput
    int per_env;                   // IRS_PER_ENV: how the per-env stages of a launch above fused_max run - 0 / "lane8": eight
                                   // lanes per env (measured fastest from 2,368 to 65,536 envs per launch), "warp": one warp
                                   // per env, "thread": one thread per env
    const double* se_bank;         // channel_mode 3 (irs_set_se_trace)
    int32_t bank_eps, bank_steps, bank_first, bank_wrap;
    const int32_t* ep_order;       // random_episodes with a given order (irs_set_episode_order)
    int32_t ep_order_n;
    char err[256];
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static void compute_layout(const irs_config& c, int32_t n_lambda, irs_layout* L)
{
    memset(L, 0, sizeof(*L));
    const size_t E = (size_t)c.num_envs;
    L->row_stride = (c.depth_cap + 3) / 4 * 4;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    L->buf = take(E * U * L->row_stride * 4);
    L->ue_i = take(E * U * UI_N * 4);
    L->ue_d = take(E * U * UD_N * 8);
    L->ring = take(E * U * W * 8);
    L->rb = take(E * U * 4);
    L->step = take(E * 4); L->episode = take(E * 4); L->scen = take(E * 4);
    L->win_len = take(E * 4); L->win_head = take(E * 4); L->err = take(E * 4);
    L->visit = take(E * 4); L->ep_pos = take(E * 4); L->ep_gen = take(E * 4); L->ep_base = take(E * 4);
    L->wsum = take(E * U * 8);
    L->sorted = take(E * S);
    L->acc = take(E * IRS_NUM_METRICS * 8);
    // assoc_mode 2: the n_scen slice-type templates are followed by two private scenario rows per env
    L->bank = take(((size_t)c.n_scen + (c.assoc_mode == 2 ? 2 * E : 0)) * sizeof(irs_scenario));
    L->cdf = take((size_t)(n_lambda > 0 ? n_lambda : 1) * IRS_POISSON_K * 8);
    L->in_a0 = take(E * S * 8); L->in_intra = take(E * S * 4);
    L->out_obs0 = take(E * IRS_OBS0 * 8);
    L->out_obsi = take(E * S * IRS_OBSI * 8);
    L->out_mask0 = take(E * S);
    L->out_maski = take(E * S * NS);
    L->out_reward = take(E * 6 * 8);
    L->out_done = take(E);
    L->metrics_total = take(IRS_NUM_METRICS * 8);
    L->scratch = take(64 * 4);
    L->out32 = take(E * (IRS_OBS0 + S * IRS_OBSI + 6) * 4);
    L->total_bytes = off;
}

static int fail(irs_env* env, int code, const char* msg)
{
    if (env) snprintf(env->err, sizeof(env->err), "%s", msg);
    return code;
}
static int cuda_fail(irs_env* env, cudaError_t e, const char* where)
{
    if (env) snprintf(env->err, sizeof(env->err), "%s: %s", where, cudaGetErrorString(e));
    return IRS_ERR_CUDA;
}
#define CK(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return cuda_fail(env, _e, #call); } while (0)

static StepArgs make_args(const irs_env* env)
{
    StepArgs A;
    memset(&A, 0, sizeof(A));
    A.st = env->st;
    A.cfg = env->cfg;
    A.bw_rb = env->cfg.bandwidth / (double)R; // ues.py:110
    A.initial_episode = -1;
    A.se_bank = env->se_bank;
    A.bank_eps = env->bank_eps;
    A.bank_steps = env->bank_steps;
    A.bank_first = env->bank_first;
    A.bank_wrap = env->bank_wrap;
    A.ep_order = env->ep_order;
    A.ep_order_n = env->ep_order_n;
    A.win_cap = env->cfg.surface ? W / 2 : W;
    A.win_mul = env->cfg.surface ? 2 : 1;
    A.defer_rollover = 0;
    A.ue_split_log2 = 0;
    A.env_begin = 0;
    A.env_end = env->cfg.num_envs;
    return A;
}

// channels/quadriga.py:56-74: SE = log2(1 + ((P_tx / R) * p) / (0 + N0)); one block per frame, the
// [R][U] file order transposed to [U][R] through shared memory so both sides stay coalesced.
__global__ void __launch_bounds__(256) k_power_to_se(const double* __restrict__ power, double* __restrict__ se, int layout,
                                                     double tx_per_rb, double noise)
{
    __shared__ double tile[U * R];
    const double* in = power + (size_t)blockIdx.x * (U * R);
    double* out = se + (size_t)blockIdx.x * (U * R);
    const double den = 0.0 + noise; // intercell_interference (zeros) + thermal_noise_power
    for (int i = threadIdx.x; i < U * R; i += blockDim.x) {
        const double v = log2(1.0 + (tx_per_rb * in[i]) / den);
        if (layout == 0) { const int r = i / U, u = i - r * U; tile[u * R + r] = v; }
        else tile[i] = v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < U * R; i += blockDim.x) out[i] = tile[i];
}

extern "C" {

int irs_version(void) { return 1; }

int irs_create(const irs_config* cfg, int32_t n_lambda, irs_env** out)
{
    if (!cfg || !out) return IRS_ERR_ARG;
    *out = nullptr;
    if (cfg->num_envs <= 0 || cfg->depth_cap <= 0 || cfg->depth_cap > 16383 /* ROW_LIVE: 14 bits */ ||
        cfg->random_episodes < 0 || cfg->random_episodes > 1 || cfg->n_scen <= 0 || cfg->max_steps <= 0 ||
        cfg->channel_mode < 0 || cfg->channel_mode > 3 || cfg->traffic_mode < 0 || cfg->traffic_mode > 1 ||
        cfg->policy < 0 || cfg->policy > 3 || cfg->surface < 0 || cfg->surface > 2 || cfg->fixed_intra < -1 || cfg->fixed_intra > 2 ||
        (cfg->assoc_mode == 0 && cfg->scen_mod <= 0) || (cfg->assoc_mode == 1 && cfg->scen_div <= 0) ||
        cfg->assoc_mode < 0 || cfg->assoc_mode > 2 || (cfg->assoc_mode == 2 && (cfg->n_scen < S || cfg->n_scen > 32)) || !(cfg->bandwidth > 0) ||
        (cfg->traffic_mode == 1 && n_lambda <= 0) || (long long)cfg->num_envs * U >= (1ll << 31))
        return IRS_ERR_ARG;
    irs_env* env = new (std::nothrow) irs_env;
    if (!env) return IRS_ERR_ARG;
    memset(env, 0, sizeof(*env));
    env->cfg = *cfg;
    env->n_lambda = n_lambda;
    const char* us_ = getenv("IRS_UE_SPLIT");
    env->ue_split = us_ ? atoi(us_) : -1;
    if (env->ue_split > 2) env->ue_split = 2;
    const char* hc = getenv("IRS_HOST_CHUNKS");
    env->host_chunks = hc ? atoi(hc) : 0;
    const char* hs_ = getenv("IRS_HOST_STREAMS");
    env->host_streams = hs_ ? atoi(hs_) : 1;
    const char* fm_ = getenv("IRS_FUSED_MAX_ENVS");
    env->fused_max = fm_ ? atoi(fm_) : 2368;
    const char* pe_ = getenv("IRS_PER_ENV");
    env->per_env = !pe_ ? 0 : (!strcmp(pe_, "warp") ? 1 : (!strcmp(pe_, "lane8") ? 2 : (!strcmp(pe_, "thread") ? 3 : 0)));
    compute_layout(env->cfg, n_lambda, &env->lay);
    *out = env;
    return IRS_OK;
}

int irs_destroy(irs_env* env)
{
    if (env && env->have_ev)
        for (int i = 0; i < 4; ++i) cudaEventDestroy(env->ev[i]);
    if (env && env->have_copy) {
        for (int i = 0; i < IRS_MAX_HOST_CHUNKS; ++i) cudaEventDestroy(env->chunk_ev[i]);
        cudaStreamDestroy(env->copy_stream);
        cudaStreamDestroy(env->side_stream);
        cudaEventDestroy(env->fork_ev);
        cudaEventDestroy(env->join_ev);
    }
    delete env;
    return IRS_OK;
}

int irs_state_layout(const irs_env* env, irs_layout* out)
{
    if (!env || !out) return IRS_ERR_ARG;
    *out = env->lay;
    return IRS_OK;
}

const char* irs_last_error(const irs_env* env) { return env ? env->err : "null handle"; }

int irs_bind(irs_env* env, void* state_dev, size_t bytes, void* stream)
{
    if (!env || !state_dev) return IRS_ERR_ARG;
    if (bytes < env->lay.total_bytes) return fail(env, IRS_ERR_ARG, "state blob smaller than irs_state_layout().total_bytes");
    if (((uintptr_t)state_dev & 255) != 0) return fail(env, IRS_ERR_ARG, "state blob must be 256-byte aligned");
    char* b = (char*)state_dev;
    const irs_layout& L = env->lay;
    DevState& s = env->st;
    env->blob = b;
    s.buf = (int32_t*)(b + L.buf);
    s.ue_i = (int32_t*)(b + L.ue_i);
    s.ue_d = (double*)(b + L.ue_d);
    s.ring = (int2*)(b + L.ring);
    s.rb = (short2*)(b + L.rb);
    s.step = (int32_t*)(b + L.step); s.episode = (int32_t*)(b + L.episode); s.scen = (int32_t*)(b + L.scen);
    s.win_len = (int32_t*)(b + L.win_len); s.win_head = (int32_t*)(b + L.win_head); s.err = (int32_t*)(b + L.err);
    s.visit = (int32_t*)(b + L.visit); s.ep_pos = (int32_t*)(b + L.ep_pos); s.ep_gen = (int32_t*)(b + L.ep_gen);
    s.ep_base = (int32_t*)(b + L.ep_base);
    s.wsum = (long long*)(b + L.wsum);
    s.sorted = (int8_t*)(b + L.sorted);
    s.acc = (double*)(b + L.acc);
    s.bank = (const irs_scenario*)(b + L.bank);
    s.cdf = (const double*)(b + L.cdf);
    s.in_a0 = (double*)(b + L.in_a0); s.in_intra = (int32_t*)(b + L.in_intra);
    s.out_obs0 = (double*)(b + L.out_obs0); s.out_obsi = (double*)(b + L.out_obsi);
    s.out_mask0 = (int8_t*)(b + L.out_mask0); s.out_maski = (int8_t*)(b + L.out_maski);
    s.out_reward = (double*)(b + L.out_reward); s.out_done = (uint8_t*)(b + L.out_done);
    s.metrics_total = (double*)(b + L.metrics_total);
    s.scratch = (int32_t*)(b + L.scratch);
    s.row_stride = L.row_stride;
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaMemsetAsync(b, 0, L.total_bytes, st));
    StepArgs A = make_args(env);
    k_init<<<(env->cfg.num_envs + 255) / 256, 256, 0, st>>>(A);
    CK(cudaGetLastError());
    env->bound = true;
    env->loaded = false;
    return IRS_OK;
}

int irs_load_scenarios(irs_env* env, const irs_scenario* bank_host, int32_t n_scen, const double* cdf_host,
                       int32_t n_lambda, void* stream)
{
    if (!env || !bank_host) return IRS_ERR_ARG;
    if (!env->bound) return fail(env, IRS_ERR_UNBOUND, "irs_bind first");
    if (n_scen != env->cfg.n_scen || n_lambda != env->n_lambda) return fail(env, IRS_ERR_ARG, "bank size differs from irs_create");
    for (int i = 0; i < n_scen; ++i) {
        for (int s = 0; s < S; ++s) {
            const irs_slice& sl = bank_host[i].sl[s];
            if (sl.n_ues > NS) return fail(env, IRS_ERR_ARG, "more than 5 UEs in a slice (agents/ib_sched.py:50)");
            if (env->cfg.assoc_mode == 2 && s == 0 && (sl.pad < 1 || sl.pad > sl.n_ues))
                return fail(env, IRS_ERR_ARG, "assoc_mode 2: template row needs 1 <= min UEs (sl[0].pad) <= max UEs (sl[0].n_ues)");
            if (sl.n_ues > 0 && ((int)sl.buffer_latency + 1 > env->cfg.depth_cap || sl.buffer_size < 1 || sl.message_size < 1))
                return fail(env, IRS_ERR_ARG, "slice buffer parameters out of range (depth_cap too small?)");
            if ((sl.flags & IRS_F_REQ) && env->cfg.traffic_mode == 1 && (sl.lam_idx < 0 || sl.lam_idx >= n_lambda))
                return fail(env, IRS_ERR_ARG, "lam_idx outside the Poisson tables");
        }
        for (int u = 0; u < U; ++u)
            if (bank_host[i].ue_slice[u] < -1 || bank_host[i].ue_slice[u] >= S) return fail(env, IRS_ERR_ARG, "ue_slice out of range");
    }
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaMemcpyAsync(env->blob + env->lay.bank, bank_host, (size_t)n_scen * sizeof(irs_scenario), cudaMemcpyHostToDevice, st));
    if (n_lambda > 0 && cdf_host)
        CK(cudaMemcpyAsync(env->blob + env->lay.cdf, cdf_host, (size_t)n_lambda * IRS_POISSON_K * 8, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st)); // the host arrays may be freed on return
    env->loaded = true;
    return IRS_OK;
}

static int check_ready(irs_env* env, const irs_inject* inj)
{
    if (!env) return IRS_ERR_ARG;
    if (!env->bound || !env->loaded) return fail(env, IRS_ERR_UNBOUND, "irs_bind and irs_load_scenarios first");
    if (env->cfg.channel_mode == 0 && !(inj && inj->se)) return fail(env, IRS_ERR_ARG, "channel_mode 0 needs inject.se");
    if (env->cfg.channel_mode == 3 && !env->se_bank) return fail(env, IRS_ERR_ARG, "channel_mode 3 needs irs_set_se_trace");
    return IRS_OK;
}

int irs_reset(irs_env* env, const uint8_t* mask_dev, int32_t initial_episode, const int32_t* initial_episode_dev,
              const irs_inject* inject, const irs_raw_outputs* raw, void* stream)
{
    int rc = check_ready(env, inject);
    if (rc) return rc;
    StepArgs A = make_args(env);
    if (inject) { A.se = inject->se; A.traffic = inject->traffic; }
    if (raw) A.raw = *raw;
    A.reset_mask = mask_dev;
    A.initial_episode = initial_episode;
    A.initial_episode_dev = initial_episode_dev;
    const int E = env->cfg.num_envs;
    k_reset<<<(E + ENV_WARPS - 1) / ENV_WARPS, ENV_WARPS * 32, 0, (cudaStream_t)stream>>>(A);
    if (env->cfg.surface != 0) k_surface_alt<<<(E + ENV_WARPS - 1) / ENV_WARPS, ENV_WARPS * 32, 0, (cudaStream_t)stream>>>(A, 2);
    CK(cudaGetLastError());
    return IRS_OK;
}

struct NvtxRange { // host-side range around the enqueue of one phase (shows the launch sequence on a timeline)
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

static int launch_step(irs_env* env, const StepArgs& A, cudaStream_t st, bool profile = false)
{
    const int E = A.env_end - A.env_begin;
    if (E <= 0) return IRS_OK;
    NvtxRange whole("irs_step");
    const bool native_src = env->cfg.channel_mode == 2 && env->cfg.traffic_mode == 1, mem_src = env->cfg.channel_mode != 2;
    if (profile) CK(cudaEventRecord(env->ev[0], st));
    if (E <= env->fused_max && env->cfg.surface == 0 && !A.sched_bits && !A.raw.se_used) {
        if (profile) CK(cudaEventRecord(env->ev[1], st));
        if (native_src) k_env_step<<<E, STEP_WARPS * 32, 0, st>>>(A); // one block per env
        else k_env_step_any<<<E, STEP_WARPS * 32, 0, st>>>(A);
        if (profile) { CK(cudaEventRecord(env->ev[2], st)); CK(cudaEventRecord(env->ev[3], st)); }
        CK(cudaGetLastError());
        return IRS_OK;
    }
    const int per_env = env->per_env ? env->per_env : 2;
    const bool serial = per_env == 3, lane8 = per_env == 2;
    const unsigned l8_blocks = (unsigned)(((size_t)E * 8 + L8_THREADS - 1) / L8_THREADS);
    nvtxRangePushA("action->RB");
    if (A.sched_bits) k_sched_runs<<<(E + 7) / 8, 256, 0, st>>>(A);
    else if (lane8) k_action_8<<<l8_blocks, L8_THREADS, 0, st>>>(A);
    else if (serial) k_action_t<<<(E + SERIAL_THREADS - 1) / SERIAL_THREADS, SERIAL_THREADS, 0, st>>>(A);
    else k_action<<<(E + 7) / 8, 256, 0, st>>>(A);
    nvtxRangePop();
    if (profile) CK(cudaEventRecord(env->ev[1], st));
    nvtxRangePushA("channel+traffic+buffers");
    StepArgs B2 = A; // small batches: 2 or 4 warps per env, so that the launch still puts ~32 warps on every SM
    B2.ue_split_log2 = env->ue_split >= 0 ? env->ue_split : (E <= 1184 ? 2 : (E <= 2368 ? 1 : 0));
    const unsigned env_warps = (unsigned)E << B2.ue_split_log2;
    const unsigned ue_blocks = (env_warps + FUSED_WARPS - 1) / FUSED_WARPS;
    if (A.raw.se_used || A.sched_bits || !(native_src || mem_src)) k_ue_env_raw<<<ue_blocks, FUSED_WARPS * 32, 0, st>>>(B2);
    else if (native_src) k_ue_env<<<ue_blocks, FUSED_WARPS * 32, 0, st>>>(B2);
    else k_ue_env_mem<<<ue_blocks, FUSED_WARPS * 32, 0, st>>>(B2);
    nvtxRangePop();
    if (profile) CK(cudaEventRecord(env->ev[2], st));
    NvtxRange tail("obs+reward+rollover");
    const bool alt = env->cfg.surface != 0; // SchedTWC / SchedColORAN: their surface reads the step's records before
                                            // finished envs roll over, then once more for the reset observation
    const unsigned env_blocks = (E + ENV_WARPS - 1) / ENV_WARPS;
    if (lane8) {
        k_env_tail_8<<<l8_blocks, L8_THREADS, 0, st>>>(A);
    } else if (serial) {
        k_env_tail_t<<<(E + SERIAL_THREADS - 1) / SERIAL_THREADS, SERIAL_THREADS, 0, st>>>(A);
    } else {
        StepArgs B = A;
        B.defer_rollover = alt ? 1 : 0;
        k_env_tail<<<env_blocks, ENV_WARPS * 32, 0, st>>>(B);
    }
    if (alt) k_surface_alt<<<env_blocks, ENV_WARPS * 32, 0, st>>>(A, 0);
    if ((serial || lane8 || alt) && env->cfg.auto_reset) {
        k_rollover<<<(E + ENV_WARPS * ROLL_ENVS - 1) / (ENV_WARPS * ROLL_ENVS), ENV_WARPS * 32, 0, st>>>(A);
        if (alt) k_surface_alt<<<env_blocks, ENV_WARPS * 32, 0, st>>>(A, 1);
    }
    if (profile) CK(cudaEventRecord(env->ev[3], st));
    CK(cudaGetLastError());
    return IRS_OK;
}

static int step_args(irs_env* env, const irs_actions* actions, const irs_inject* inject, const irs_raw_outputs* raw, StepArgs* A)
{
    int rc = check_ready(env, inject);
    if (rc) return rc;
    const bool sched = actions && actions->sched_bits; // a scheduling matrix replaces the built-in action_format
    const bool given = actions && actions->a0 && !sched; // explicit actions override the on-device policy
    if (env->cfg.policy == 0 && !given && !sched) return fail(env, IRS_ERR_ARG, "policy 0 needs actions.a0");
    if (given && env->cfg.fixed_intra < 0 && !actions->intra) return fail(env, IRS_ERR_ARG, "fixed_intra -1 needs actions.intra");
    if (env->cfg.traffic_mode == 0 && !(inject && inject->traffic)) return fail(env, IRS_ERR_ARG, "traffic_mode 0 needs inject.traffic");
    *A = make_args(env);
    if (given) { A->a0 = actions->a0; A->intra = actions->intra; A->cfg.policy = 0; }
    if (sched) A->sched_bits = actions->sched_bits;
    if (inject) { A->se = inject->se; A->traffic = inject->traffic; }
    if (raw) A->raw = *raw;
    return IRS_OK;
}

int irs_step(irs_env* env, const irs_actions* actions, const irs_inject* inject, const irs_raw_outputs* raw, void* stream)
{
    StepArgs A;
    int rc = step_args(env, actions, inject, raw, &A);
    if (rc) return rc;
    return launch_step(env, A, (cudaStream_t)stream);
}

int irs_step_profiled(irs_env* env, const irs_actions* actions, const irs_inject* inject, const irs_raw_outputs* raw,
                      void* stream, float* kernel_ms_host)
{
    StepArgs A;
    int rc = step_args(env, actions, inject, raw, &A);
    if (rc) return rc;
    if (!kernel_ms_host) return IRS_ERR_ARG;
    if (!env->have_ev) {
        for (int i = 0; i < 4; ++i) CK(cudaEventCreate(&env->ev[i]));
        env->have_ev = true;
    }
    rc = launch_step(env, A, (cudaStream_t)stream, true);
    if (rc) return rc;
    CK(cudaEventSynchronize(env->ev[3]));
    for (int i = 0; i < 3; ++i) CK(cudaEventElapsedTime(&kernel_ms_host[i], env->ev[i], env->ev[i + 1]));
    return IRS_OK;
}

int irs_set_seed(irs_env* env, uint64_t seed)
{
    if (!env) return IRS_ERR_ARG;
    env->cfg.seed = seed;
    return IRS_OK;
}

int irs_set_se_trace(irs_env* env, const double* se_bank_dev, int32_t n_episodes, int32_t n_steps, int32_t first_episode,
                     int32_t wrap)
{
    if (!env) return IRS_ERR_ARG;
    if (!se_bank_dev || n_episodes <= 0 || n_steps <= 0) return fail(env, IRS_ERR_ARG, "irs_set_se_trace: empty bank");
    if (((uintptr_t)se_bank_dev & 7) != 0) return fail(env, IRS_ERR_ARG, "irs_set_se_trace: bank must be 8-byte aligned");
    env->se_bank = se_bank_dev;
    env->bank_eps = n_episodes;
    env->bank_steps = n_steps;
    env->bank_first = first_episode;
    env->bank_wrap = wrap ? 1 : 0;
    return IRS_OK;
}

int irs_set_episode_order(irs_env* env, const int32_t* order_dev, int32_t n, void* stream)
{
    if (!env) return IRS_ERR_ARG;
    if (!env->bound) return fail(env, IRS_ERR_UNBOUND, "irs_bind first");
    if (!env->cfg.random_episodes) return fail(env, IRS_ERR_ARG, "irs_set_episode_order needs cfg.random_episodes");
    if (order_dev && n <= 0) return fail(env, IRS_ERR_ARG, "irs_set_episode_order: empty order");
    env->ep_order = order_dev;
    env->ep_order_n = order_dev ? n : 0;
    if (order_dev) {
        StepArgs A = make_args(env);
        k_apply_order<<<(env->cfg.num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(A);
        CK(cudaGetLastError());
    }
    return IRS_OK;
}

int irs_read_buffers(irs_env* env, int32_t* out_dev, void* stream)
{
    if (!env || !out_dev) return IRS_ERR_ARG;
    if (!env->bound || !env->loaded) return fail(env, IRS_ERR_UNBOUND, "irs_bind and irs_load_scenarios first");
    StepArgs A = make_args(env);
    const size_t rows = (size_t)env->cfg.num_envs * U;
    k_read_buffers<<<(unsigned)((rows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(A, out_dev);
    CK(cudaGetLastError());
    return IRS_OK;
}

int irs_power_to_se(const double* power_dev, double* se_dev, int64_t n_frames, int32_t layout, double tx_power,
                    double noise_power, void* stream)
{
    if (!power_dev || !se_dev || n_frames < 0 || layout < 0 || layout > 1 || !(noise_power > 0)) return IRS_ERR_ARG;
    if (n_frames == 0) return IRS_OK;
    const long long per_launch = 1ll << 30;
    for (long long f0 = 0; f0 < n_frames; f0 += per_launch) {
        const long long nf = n_frames - f0 < per_launch ? n_frames - f0 : per_launch;
        k_power_to_se<<<(unsigned)nf, 256, 0, (cudaStream_t)stream>>>(power_dev + f0 * U * R, se_dev + f0 * U * R, layout,
                                                                        tx_power / (double)R, noise_power);
    }
    return cudaGetLastError() == cudaSuccess ? IRS_OK : IRS_ERR_CUDA;
}

int irs_run(irs_env* env, int32_t n_steps, void* stream)
{
    int rc = check_ready(env, nullptr);
    if (rc) return rc;
    if (env->cfg.policy == 0 || env->cfg.traffic_mode == 0) return fail(env, IRS_ERR_ARG, "irs_run needs an on-device policy and native generators");
    StepArgs A = make_args(env);
    for (int i = 0; i < n_steps; ++i) {
        rc = launch_step(env, A, (cudaStream_t)stream);
        if (rc) return rc;
    }
    return IRS_OK;
}

// float64 -> float32 of one surface array for envs [e0, e1) (the binary32 host surface)
__global__ void k_to_f32(const double* __restrict__ src, float* __restrict__ dst, size_t n)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (float)src[i];
}

// Device->host copies of the env surface for envs [e0, e1), enqueued on `st` (no synchronisation).
static int copy_surface_range(irs_env* env, const irs_host_surface* o, bool with_reward, size_t e0, size_t e1, cudaStream_t st,
                              int which = 3 /* 1: observation vectors, 2: masks / rewards / done, 3: all */)
{
    const size_t n = e1 - e0;
    const irs_layout& L = env->lay;
    char* b = env->blob;
    const size_t E = (size_t)env->cfg.num_envs;
    auto d2h32 = [&](float* host, size_t dev_off64, size_t stage_off, size_t per_env) -> cudaError_t {
        if (!host) return cudaSuccess;
        const double* src = (const double*)(b + dev_off64) + e0 * per_env;
        float* stage = (float*)(b + L.out32) + stage_off + e0 * per_env;
        const size_t cnt = n * per_env;
        k_to_f32<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(src, stage, cnt);
        return cudaMemcpyAsync(host + e0 * per_env, stage, cnt * 4, cudaMemcpyDeviceToHost, st);
    };
    auto d2h = [&](void* host, size_t dev_off, size_t per_env) -> cudaError_t {
        if (!host) return cudaSuccess;
        return cudaMemcpyAsync((char*)host + e0 * per_env, b + dev_off + e0 * per_env, n * per_env, cudaMemcpyDeviceToHost, st);
    };
    if (which & 1) {
        CK(d2h32(o->obs0_f32, L.out_obs0, 0, IRS_OBS0));
        CK(d2h32(o->obsi_f32, L.out_obsi, E * IRS_OBS0, S * IRS_OBSI));
        CK(d2h(o->obs0, L.out_obs0, IRS_OBS0 * 8));
        CK(d2h(o->obsi, L.out_obsi, S * IRS_OBSI * 8));
    }
    if (which & 2) {
        if (with_reward) CK(d2h32(o->reward_f32, L.out_reward, E * (IRS_OBS0 + S * IRS_OBSI), 6));
        CK(d2h(o->mask0, L.out_mask0, S));
        CK(d2h(o->maski, L.out_maski, S * NS));
        if (with_reward) {
            CK(d2h(o->reward, L.out_reward, 6 * 8));
            CK(d2h(o->done, L.out_done, 1));
        }
    }
    return IRS_OK;
}

static int copy_surface(irs_env* env, const irs_host_surface* o, bool with_reward, cudaStream_t st)
{
    int rc = copy_surface_range(env, o, with_reward, 0, (size_t)env->cfg.num_envs, st);
    if (rc) return rc;
    CK(cudaStreamSynchronize(st));
    return IRS_OK;
}

// How irs_step_host splits the batch: the surface of a finished chunk crosses PCIe while the next chunk computes.
// Measured on B200 (profiles/r2_notes.md): the float64 surface of 65,536 envs needs 1.42 ms on the wire (57 GB/s),
// the unchunked step 0.71 ms - but a chunk is three dependent kernels that each end in a partly filled wave and run the
// warp-per-env action / tail kernels, 17-20 ns per env against 10.8 unchunked, so chunked compute (1.26 ms in four
// chunks) only just keeps ahead of the copy (21.7 ns per env).  The call is then copy-bound at
// (first chunk's kernels) + (whole copy): FOUR equal chunks measured best (1.82 ms; 3, 5, 6, 8, 12, 16 chunks, a small
// first chunk with doubling sizes, and two alternating compute streams all measured 1.87-2.1 ms: fewer chunks start the
// copy later, more chunks compute slower than the wire).  Small batches stay whole (launch latency dominates).
// IRS_HOST_CHUNKS=n forces n equal chunks, IRS_HOST_STREAMS=2 alternates the chunks between two compute streams.
static int host_chunk_plan(irs_env* env, size_t* bounds /* [IRS_MAX_HOST_CHUNKS + 1] */)
{
    const size_t E = (size_t)env->cfg.num_envs;
    static const char* plan = getenv("IRS_HOST_PLAN");
    if (env->host_chunks <= 0 && E >= 16384 && plan && !strcmp(plan, "geom")) { // small first chunk, doubling sizes
        size_t s = env->fused_max >= 1024 ? (size_t)env->fused_max : 2368, done = 0;
        int n = 0;
        bounds[0] = 0;
        while (E - done > 2 * s && n < IRS_MAX_HOST_CHUNKS - 1) { done += s; bounds[++n] = done; s *= 2; }
        bounds[++n] = E;
        return n;
    }
    int chunks = env->host_chunks > 0 ? env->host_chunks : (E < 16384 ? 1 : 4);
    if (chunks > IRS_MAX_HOST_CHUNKS) chunks = IRS_MAX_HOST_CHUNKS;
    size_t per = (E + chunks - 1) / chunks;
    if (env->host_chunks <= 0 && chunks > 1) per = (per + 31) / 32 * 32; // whole blocks of every kernel
    bounds[0] = 0;
    for (int c = 1; c <= chunks; ++c) bounds[c] = (size_t)c * per < E ? (size_t)c * per : E;
    return chunks;
}

int irs_step_host(irs_env* env, const irs_actions* actions_host, const irs_host_surface* out_host, void* stream)
{
    int rc = check_ready(env, nullptr);
    if (rc) return rc;
    if (env->cfg.channel_mode == 0 || env->cfg.traffic_mode == 0) return fail(env, IRS_ERR_ARG, "irs_step_host needs device-side channel and traffic (channel_mode 1/2/3, traffic_mode 1)");
    if (!out_host) return IRS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t E = (size_t)env->cfg.num_envs;
    StepArgs A = make_args(env);
    const bool given = actions_host && actions_host->a0;
    if (env->cfg.policy == 0 && !given) return fail(env, IRS_ERR_ARG, "policy 0 needs actions.a0");
    if (given) {
        A.a0 = env->st.in_a0;
        A.cfg.policy = 0;
        if (actions_host->intra) A.intra = env->st.in_intra;
        else if (env->cfg.fixed_intra < 0) return fail(env, IRS_ERR_ARG, "fixed_intra -1 needs actions.intra");
    }
    // the actions of envs [e0, e1) to the device, right before the kernels that read them (a chunk's slice is small,
    // so the first kernel does not wait for the whole batch's actions)
    auto upload_actions = [&](size_t e0, size_t e1, cudaStream_t s_) -> cudaError_t {
        if (!given) return cudaSuccess;
        cudaError_t rc_ = cudaMemcpyAsync(env->st.in_a0 + e0 * S, actions_host->a0 + e0 * S, (e1 - e0) * S * 8, cudaMemcpyHostToDevice, s_);
        if (rc_ == cudaSuccess && actions_host->intra)
            rc_ = cudaMemcpyAsync(env->st.in_intra + e0 * S, actions_host->intra + e0 * S, (e1 - e0) * S * 4, cudaMemcpyHostToDevice, s_);
        return rc_;
    };
    size_t bounds[IRS_MAX_HOST_CHUNKS + 1];
    const int chunks = host_chunk_plan(env, bounds);
    if (chunks == 1) {
        CK(upload_actions(0, E, st));
        rc = launch_step(env, A, st);
        if (rc) return rc;
        return copy_surface(env, out_host, true, st);
    }
    if (!env->have_copy) {
        CK(cudaStreamCreateWithFlags(&env->copy_stream, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&env->side_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&env->fork_ev, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&env->join_ev, cudaEventDisableTiming));
        for (int i = 0; i < IRS_MAX_HOST_CHUNKS; ++i) CK(cudaEventCreateWithFlags(&env->chunk_ev[i], cudaEventDisableTiming));
        env->have_copy = true;
    }
    // IRS_HOST_STREAMS=2: chunks alternate between two compute streams, so that the partly filled last waves of one
    // chunk's kernels are filled by the next chunk's (chunks touch disjoint env ranges: no hazard).  Measured no better
    // than one stream (a chunk must FINISH early for its copy to start early), kept as an option.
    const bool two = env->host_streams >= 2;
    if (two) { // the side stream starts after everything the caller queued on `st`
        CK(cudaEventRecord(env->fork_ev, st));
        CK(cudaStreamWaitEvent(env->side_stream, env->fork_ev, 0));
    }
    // IRS_HOST_TRACE=1: timeline of one call on stderr (ms since the call's first enqueue) - a debugging aid
    static const bool trace = getenv("IRS_HOST_TRACE") != nullptr;
    static double last_return = 0;
    const double w_entry = wall_ms();
    cudaEvent_t t0 = nullptr, tk[IRS_MAX_HOST_CHUNKS] = {}, tc[IRS_MAX_HOST_CHUNKS] = {};
    if (trace) {
        CK(cudaEventCreate(&t0));
        for (int c = 0; c < chunks; ++c) { CK(cudaEventCreate(&tk[c])); CK(cudaEventCreate(&tc[c])); }
        CK(cudaEventRecord(t0, st));
    }
    for (int c = 0; c < chunks; ++c) {
        const size_t e0 = bounds[c], e1 = bounds[c + 1];
        if (e0 >= e1) break;
        A.env_begin = (int32_t)e0;
        A.env_end = (int32_t)e1;
        const cudaStream_t cs = (two && (c & 1)) ? env->side_stream : st;
        CK(upload_actions(e0, e1, cs));
        rc = launch_step(env, A, cs);
        if (rc) return rc;
        CK(cudaEventRecord(env->chunk_ev[c], cs));
        if (trace) CK(cudaEventRecord(tk[c], cs));
        CK(cudaStreamWaitEvent(env->copy_stream, env->chunk_ev[c], 0));
        // the observation vectors (94 % of the bytes) chunk by chunk; the four small arrays once for the whole batch
        // after the last chunk: per-copy set-up is what they cost (16 small copies measured 0.06 ms of a 1.78 ms call)
        rc = copy_surface_range(env, out_host, true, e0, e1, env->copy_stream, 1);
        if (rc) return rc;
        if (trace) CK(cudaEventRecord(tc[c], env->copy_stream));
    }
    rc = copy_surface_range(env, out_host, true, 0, E, env->copy_stream, 2);
    if (rc) return rc;
    if (two) { // the caller's stream continues after both compute streams
        CK(cudaEventRecord(env->join_ev, env->side_stream));
        CK(cudaStreamWaitEvent(st, env->join_ev, 0));
    }
    const double w_enq = wall_ms();
    CK(cudaStreamSynchronize(env->copy_stream));
    const double w_sync = wall_ms();
    if (trace) {
        fprintf(stderr, "irs_step_host trace: host [since last return %.3f, enqueue %.3f, wait %.3f]", last_return ? w_entry - last_return : 0.0,
                w_enq - w_entry, w_sync - w_enq);
        for (int c = 0; c < chunks; ++c) {
            float a = 0, b = 0;
            cudaEventElapsedTime(&a, t0, tk[c]);
            cudaEventElapsedTime(&b, t0, tc[c]);
            fprintf(stderr, " [%zu envs: kernels done %.3f, copied %.3f]", bounds[c + 1] - bounds[c], a, b);
            cudaEventDestroy(tk[c]); cudaEventDestroy(tc[c]);
        }
        fprintf(stderr, "\n");
        cudaEventDestroy(t0);
        last_return = wall_ms();
    }
    return IRS_OK;
}

int irs_reset_host(irs_env* env, int32_t initial_episode, const irs_host_surface* out_host, void* stream)
{
    int rc = irs_reset(env, nullptr, initial_episode, nullptr, nullptr, nullptr, stream);
    if (rc) return rc;
    if (!out_host) return IRS_ERR_ARG;
    return copy_surface(env, out_host, false, (cudaStream_t)stream);
}

int irs_episode_metrics(irs_env* env, double* host_out, void* stream)
{
    if (!env || !env->bound) return IRS_ERR_UNBOUND;
    cudaStream_t st = (cudaStream_t)stream;
    StepArgs A = make_args(env);
    k_metrics<<<IRS_NUM_METRICS, 256, 0, st>>>(A);
    CK(cudaGetLastError());
    if (host_out) {
        CK(cudaMemcpyAsync(host_out, env->st.metrics_total, IRS_NUM_METRICS * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    return IRS_OK;
}

int irs_check_errors(irs_env* env, int32_t* first_env, int32_t* code, void* stream)
{
    if (!env || !env->bound) return IRS_ERR_UNBOUND;
    cudaStream_t st = (cudaStream_t)stream;
    const int E = env->cfg.num_envs;
    StepArgs A = make_args(env);
    CK(cudaMemsetAsync(env->st.scratch, 0x7f, 4, st)); // 0x7f7f7f7f: larger than any env index
    k_err_scan<<<(E + 255) / 256, 256, 0, st>>>(A);
    k_err_word<<<1, 1, 0, st>>>(A);
    CK(cudaGetLastError());
    int32_t h[2] = {0, 0};
    CK(cudaMemcpyAsync(h, env->st.scratch, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (h[0] < 0 || h[0] >= E) return IRS_OK;
    if (first_env) *first_env = h[0];
    if (code) *code = h[1];
    snprintf(env->err, sizeof(env->err), "env %d error word %d", h[0], h[1]);
    return IRS_ERR_ENV;
}

// ---- irs_check_errors_noalloc: the error scan on a library-owned static scratch (no counterpart in the reference; added
// for callers that cannot allocate: graph capture, serving loops) ------------------
// One int per 256-env block (the block's first env with a non-zero error word, INT_MAX: none), then one block reduces the
// slots.  Every slot a call reads is written by the same call, so nothing has to be cleared between calls; the array is
// shared by every call on the device and keeps the last call's contents - calls that overlap on the GPU race.
#define NOALLOC_SLOTS (IRS_NOALLOC_MAX_ENVS / 256)
static __device__ __align__(256) int32_t g_noalloc_scratch[NOALLOC_SLOTS];

__global__ void __launch_bounds__(256) k_err_scan_blocks(const int32_t* __restrict__ err, int num_envs, int32_t* __restrict__ slots)
{
    const int e = blockIdx.x * 256 + threadIdx.x;
    const bool bad = e < num_envs && err[e] != 0;
    const uint32_t b = __ballot_sync(FULL, bad);
    __shared__ int first[8];
    if ((threadIdx.x & 31) == 0) first[threadIdx.x >> 5] = b ? blockIdx.x * 256 + (int)(threadIdx.x & ~31u) + __ffs(b) - 1 : INT_MAX;
    __syncthreads();
    if (threadIdx.x == 0) {
        int m = first[0];
#pragma unroll
        for (int w = 1; w < 8; ++w) m = min(m, first[w]);
        slots[blockIdx.x] = m;
    }
}
__global__ void __launch_bounds__(256) k_err_first(const int32_t* __restrict__ err, int n_slots, const int32_t* __restrict__ slots,
                                                  int32_t* __restrict__ out)
{
    __shared__ int part[256];
    int m = INT_MAX;
    for (int i = threadIdx.x; i < n_slots; i += 256) m = min(m, slots[i]);
    part[threadIdx.x] = m;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if (threadIdx.x < d) part[threadIdx.x] = min(part[threadIdx.x], part[threadIdx.x + d]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const int e = part[0];
        out[0] = e == INT_MAX ? -1 : e;
        out[1] = e == INT_MAX ? 0 : err[e];
    }
}

// device address of the static scratch: the symbol's name is a host-side stand-in in host code, so the address is asked
// for once per device, with that device current
static int32_t* noalloc_scratch_ptr(cudaError_t* ce)
{
    static std::mutex mu;
    static int32_t* ptr[64] = {};
    int dev = -1;
    *ce = cudaGetDevice(&dev);
    if (*ce != cudaSuccess) return nullptr;
    if (dev < 0 || dev >= 64) { *ce = cudaErrorInvalidDevice; return nullptr; }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void* q = nullptr;
        *ce = cudaGetSymbolAddress(&q, g_noalloc_scratch);
        if (*ce != cudaSuccess) return nullptr;
        ptr[dev] = (int32_t*)q;
    }
    return ptr[dev];
}

int irs_check_errors_noalloc(const int32_t* err_dev, int32_t num_envs, int32_t* out_dev, void* stream, char* errbuf,
                             size_t errbuf_len)
{
    auto say = [&](int code, const char* fmt, long long a, long long b) {
        if (errbuf && errbuf_len) snprintf(errbuf, errbuf_len, fmt, a, b);
        return code;
    };
    if (!err_dev || !out_dev) return say(IRS_ERR_ARG, "irs_check_errors_noalloc: null err_dev / out_dev", 0, 0);
    if (num_envs < 1) return say(IRS_ERR_ARG, "irs_check_errors_noalloc: num_envs %lld < 1", num_envs, 0);
    if (num_envs > IRS_NOALLOC_MAX_ENVS)
        return say(IRS_ERR_ARG, "irs_check_errors_noalloc: num_envs %lld exceeds IRS_NOALLOC_MAX_ENVS (%lld): the static scratch "
                                "holds one slot per 256 envs", num_envs, IRS_NOALLOC_MAX_ENVS);
    cudaError_t ce;
    int32_t* slots = noalloc_scratch_ptr(&ce);
    if (!slots) {
        if (errbuf && errbuf_len) snprintf(errbuf, errbuf_len, "irs_check_errors_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(ce));
        return IRS_ERR_CUDA;
    }
    const int n_slots = (num_envs + 255) / 256;
    cudaStream_t st = (cudaStream_t)stream;
    k_err_scan_blocks<<<n_slots, 256, 0, st>>>(err_dev, num_envs, slots);
    k_err_first<<<1, 256, 0, st>>>(err_dev, n_slots, slots, out_dev);
    ce = cudaGetLastError();
    if (ce != cudaSuccess) {
        if (errbuf && errbuf_len) snprintf(errbuf, errbuf_len, "irs_check_errors_noalloc: launch: %s", cudaGetErrorString(ce));
        return IRS_ERR_CUDA;
    }
    return IRS_OK;
}

#ifdef IRS_PHASE_CLOCK
// debug builds only: the SM-clock stamps of the last k_env_step launch, [block][warp][stamp] (scripts/phase_clock.py)
int irs_debug_phase_clock(long long* out_host, int n_blocks)
{
    cudaDeviceSynchronize();
    return (int)cudaMemcpyFromSymbol(out_host, g_phase_clock, sizeof(long long) * 32 * (size_t)n_blocks);
}
#endif

} // extern "C"
