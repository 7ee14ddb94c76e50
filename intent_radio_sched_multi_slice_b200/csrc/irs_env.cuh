// irs_env.cuh - per-env tail of a step (reward, episode metrics, done, rollover) and reset.
// One warp per env.  See irs_obs.cuh for the reference symbols replaced.
#pragma once
#include "irs_obs.cuh"

#define ENV_WARPS 4 /* envs per block in k_env_tail / k_reset */

// assoc_mode 2 - MultSliceAssociation in GENERATOR mode (associations/mult_slice.py:359-416, 454-466): a fresh
// network scenario per (env, episode) instead of a replayed file: 3..5 slices, slice slots and slice types drawn
// without replacement, UEs per slice uniform in the type's [min, max], random UE ids.  The first n_scen rows of the
// bank are the slice-TYPE templates (sl[0] = the type's parameters, sl[0].n_ues = max UEs, sl[0].pad = min UEs);
// the env's own rows follow (two per env, alternating with the episode's parity so that a row is never rewritten in
// the launch that read it through the read-only cache).  Draws come from the counter-based generator, keyed by
// (seed, global env id, episode): same distribution as the reference, not its PCG64 stream.  Serial, by one lane.
struct ScenWords {
    uint64_t seed;
    int episode, gid, i;
    uint4 cur;
    __device__ uint32_t next()
    {
        if ((i & 3) == 0) cur = rng_quad(seed, K_SCEN, 0, i >> 2, 0, episode, gid);
        const uint32_t w = (i & 3) == 0 ? cur.x : ((i & 3) == 1 ? cur.y : ((i & 3) == 2 ? cur.z : cur.w));
        ++i;
        return w;
    }
    __device__ int below(int n) { return (int)(((uint64_t)next() * (uint32_t)n) >> 32); } // uniform in [0, n)
};
__device__ __noinline__ void generate_scenario_row(const StepArgs& A, uint64_t key, int e, int ep, irs_scenario* row)
{
    const irs_scenario* types = A.st.bank;
    const int T = A.cfg.n_scen;
    ScenWords g;
    g.seed = key; g.episode = ep; g.gid = A.cfg.env_id_base + e; g.i = 0;
    int slots[S], tys[32], counts[S], ues[U];
    const int k = 3 + g.below(S - 3 + 1);
    for (int j = 0; j < S; ++j) slots[j] = j;
    for (int j = 0; j < k; ++j) { const int r = j + g.below(S - j); const int t = slots[j]; slots[j] = slots[r]; slots[r] = t; }
    for (int j = 0; j < T && j < 32; ++j) tys[j] = j;
    for (int j = 0; j < k; ++j) { const int r = j + g.below(T - j); const int t = tys[j]; tys[j] = tys[r]; tys[r] = t; }
    int total = 0;
    for (int j = 0; j < k; ++j) {
        const int mn = types[tys[j]].sl[0].pad, mx = types[tys[j]].sl[0].n_ues;
        counts[j] = mn + g.below(mx - mn + 1);
        total += counts[j];
    }
    for (int j = 0; j < U; ++j) ues[j] = j;
    for (int j = 0; j < total; ++j) { const int r = j + g.below(U - j); const int t = ues[j]; ues[j] = ues[r]; ues[r] = t; }
    irs_slice empty;
    memset(&empty, 0, sizeof(empty));
    for (int s = 0; s < S; ++s) row->sl[s] = empty;
    for (int u = 0; u < 32; ++u) row->ue_slice[u] = -1;
    int off = 0;
    for (int j = 0; j < k; ++j) {
        irs_slice sl = types[tys[j]].sl[0];
        sl.n_ues = counts[j];
        sl.pad = 0;
        row->sl[slots[j]] = sl;
        for (int c = 0; c < counts[j]; ++c) row->ue_slice[ues[off++]] = (int8_t)slots[j];
    }
}

// CommunicationEnv.reset (comm_env.py:375-482) for env e, by the whole warp.  `se_row` is a
// 136-double shared-memory scratch row owned by the warp.  The observation window is NOT
// cleared (ib_sched.py:51,64: the deque lives in the agent) - the zero frame of the reset is
// pushed into it (comm_env.py:474-482), quirk Q1 of SURVEY 7.4.
__device__ __noinline__ void reset_env_warp(const StepArgs& A, int e, int initial_episode, double* se_row)
{
    const DevState& st = A.st;
    const irs_config& c = A.cfg;
    const int lane = lane_id();
    const int gid = c.env_id_base + e;
    int step = st.step[e], ep = st.episode[e];
    if (c.random_episodes) { // comm_env.py:434-444: a shuffled cycle over [initial_episode, max_episodes)
        int pos = st.ep_pos[e], gen = st.ep_gen[e], base = st.ep_base[e];
        bool moved = false;
        if (initial_episode != -1) { base = initial_episode; gen += 1; pos = 0; moved = true; } // :435-442 new permutation
        if (step == c.max_steps) { pos += 1; moved = true; }                                    // :443-444 next(cycle)
        const int n = A.ep_order ? A.ep_order_n : c.max_episodes - base;
        if (n <= 0) {
            if (lane == 0) atomicOr(&st.err[e], ERR_EPISODE_RANGE);
            return;
        }
        pos %= n;
        if (moved) ep = A.ep_order ? A.ep_order[(size_t)e * n + pos] : shuffled_episode(c.seed, gid, gen, base, n, pos);
        __syncwarp();
        if (lane == 0) { st.ep_pos[e] = pos; st.ep_gen[e] = gen; st.ep_base[e] = base; }
    } else {
        const int init = initial_episode != -1 ? initial_episode : c.initial_episode;
        if ((step == 0 && ep == 0) || (step == c.max_steps && ep == c.max_episodes - 1) || initial_episode != -1) {
            ep = init;                                              // comm_env.py:416-421
        } else if (step == 0 && ep <= c.max_episodes - 1) {         // :422-425
        } else if (ep < c.max_episodes - 1) {
            ep += 1;                                                // :426-427
        } else {
            if (lane == 0) atomicOr(&st.err[e], ERR_EPISODE_RANGE); // :428-433
            return;
        }
    }
    const int visit = st.visit[e] + 1;
    const uint64_t key = env_key(c.seed, visit);
    // fresh UEs: empty buffers (comm_env.py:620-625, associations/mult_slice.py:468-488).  Only the live window of a
    // ring row can hold packets: ages [0, live) = slots (step - age) mod depth, with the depths of the OLD scenario.
    {
        const irs_scenario* old = st.bank + st.scen[e];
        for (int u = 0; u < U; ++u) {
            const size_t eu = (size_t)e * U + u;
            const int live = ROW_LIVE(st.ue_i[eu * UI_N + UI_PATH]);
            const int sl = ((const volatile int8_t*)old->ue_slice)[u];
            int depth = sl < 0 ? 101 : (int)((const volatile double*)&old->sl[sl].buffer_latency)[0] + 1;
            if (depth > st.row_stride) depth = st.row_stride;
            int* row = st.buf + eu * st.row_stride;
            const int head = step % depth;
            for (int a = lane; a < live; a += 32) row[head - a + (head < a ? depth : 0)] = 0;
        }
    }
    __syncwarp();
    int scen;
    if (c.assoc_mode == 2) { // the env's other private row: the one in use was read through the read-only cache
        const int row0 = c.n_scen + 2 * e;
        scen = st.scen[e] == row0 ? row0 + 1 : row0;
        if (lane == 0) generate_scenario_row(A, key, e, ep, const_cast<irs_scenario*>(st.bank) + scen);
        __syncwarp();
    } else {
        scen = scen_of_episode(c, ep);
    }
    const irs_scenario* sc = st.bank + scen;
    __syncwarp();
    if (lane == 0) { st.step[e] = 0; st.episode[e] = ep; st.scen[e] = scen; st.visit[e] = visit; }
    __syncwarp();
    // channel frame 0 of the new episode -> mean SE per UE (comm_env.py:449-451)
    double sebar = 0;
    for (int u = 0; u < U; ++u) {
        channel_row(A, key, e, u, sc->ue_slice[u] >= 0, 0, ep, se_row);
        __syncwarp();
        if (A.raw.se_used)
            for (int r = lane; r < R; r += 32) A.raw.se_used[((size_t)e * U + u) * R + r] = se_row[r];
        const double sum = __shfl_sync(FULL, np_sum135_pair(se_row, se_row), 0);
        if (lane == u) sebar = DIVC(sum, R);
        __syncwarp();
    }
    // push the all-zero metrics frame (comm_env.py:474-478)
    const int win_len = st.win_len[e];
    const int p = (st.win_head[e] + 1) % A.win_cap;
    UeFrame f;
    f.occ = 0; f.lat = 0; f.sebar = sebar; f.sent = 0; f.rb_count = 0;
    f.sum_sent = 0; f.sum_drop = 0; f.occ1z = false;
    if (lane < U) {
        const size_t eu = (size_t)e * U + lane;
        int* ui = st.ue_i + eu * UI_N;
        double* ud = st.ue_d + eu * UD_N;
        int ss = ui[UI_SUM_SENT], sd = ui[UI_SUM_DROP];
        if (win_len == A.win_cap) { const int2 ev = st.ring[eu * W + p]; ss -= ev.x; sd -= ev.y; }
        st.ring[eu * W + p] = make_int2(0, 0);
        f.sum_sent = ss; f.sum_drop = sd;
        f.occ1z = isclose0(ud[UD_OCC]);
        ui[UI_TOT] = 0; ui[UI_SUM_SENT] = ss; ui[UI_SUM_DROP] = sd; ui[UI_SENT] = 0; ui[UI_DROP] = 0;
        ui[UI_INC] = 0; ui[UI_OCC1Z] = f.occ1z ? 1 : 0; ui[UI_PATH] = 0;
        ud[UD_OCC] = 0; ud[UD_LAT] = 0; ud[UD_SEBAR] = sebar; ud[UD_THR] = 0;
        st.rb[eu] = make_short2(0, 0);
        st.wsum[eu] = 0;
    }
    const int win_len_new = win_len < A.win_cap ? win_len + 1 : A.win_cap;
    __syncwarp();
    if (lane == 0) { st.win_len[e] = win_len_new; st.win_head[e] = p; }
    format_obs_warp(A, e, sc, f, win_len_new);
}

// Everything of CommunicationEnv.step after UEs.step (comm_env.py:339-373).
// Core on register inputs: `f` = this lane's newest frame (already pushed into the per-UE
// window), dropped / inc of that frame, the env's window length / head / step BEFORE this step.
__device__ __forceinline__ void env_tail_core(const StepArgs& A, int e, const irs_scenario* sc, UeFrame f,
                                              int dropped, int inc, int win_len_old, int head_old, int step_old,
                                              double* se_row, bool block_sync = false)
{
    const DevState& st = A.st;
    const irs_config& c = A.cfg;
    const int lane = lane_id();
    PHASE_STAMP(3, 0);
    if (lane >= U) { f.sent = 0; f.rb_count = 0; dropped = 0; inc = 0; }
    const int win_len = win_len_old < A.win_cap ? win_len_old + 1 : A.win_cap;
    const int head = (head_old + 1) % A.win_cap;
    const int step = step_old + 1; // comm_env.py:339
    __syncwarp();
    if (lane == 0) { st.win_len[e] = win_len; st.win_head[e] = head; st.step[e] = step; }

    const SliceView v = format_obs_warp(A, e, sc, f, win_len); // comm_env.py:340
    PHASE_STAMP(3, 4);

    if (block_sync) __syncthreads(); // keeps the block's warps in the same code region (I-cache)
    // ---- calculate_reward (ib_sched.py:206-221, common.py:381-439) ----------------------
    const double prio_raw = __ldg(&sc->sl[lane < S ? lane : 0].priority);
    double a_s = 0, p_s = 0;
    if (v.active) { // common.py:400-408; inactive metrics are already 0 (quirk Q4)
        p_s = prio_raw;
        double mn = v.dsl[0];
        mn = v.dsl[1] < mn ? v.dsl[1] : mn;
        mn = v.dsl[2] < mn ? v.dsl[2] : mn;
        a_s = mn;
    }
    double Aa[S], Pp[S];
    int nneg = 0, npneg = 0;
#pragma unroll
    for (int s = 0; s < S; ++s) {
        Aa[s] = __shfl_sync(FULL, a_s, s);
        Pp[s] = __shfl_sync(FULL, p_s, s);
        nneg += Aa[s] < 0;
        npneg += (Pp[s] * Aa[s]) < 0;
    }
    double r0;
    {
        double res = 0.;
        int m = 0;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const bool sel = nneg == 0 ? true : (npneg ? (Aa[s] * Pp[s] < 0) : (Aa[s] < 0));
            if (sel) { res += Aa[s]; ++m; }
        }
        r0 = div_small(0. + res, m) - ((nneg != 0 && npneg) ? 1 : 0); // common.py:410-426
    }
    double r_intra = 0; // common.py:428-437
    {
        bool any = false;
        double mn = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k)
            if (v.act[k] != 0) { mn = any ? (v.dsl[k] < mn ? v.dsl[k] : mn) : v.dsl[k]; any = true; }
        if (any) r_intra = mn;
    }
    if (lane == 0) st.out_reward[(size_t)e * 6] = r0;
    if (lane < S) st.out_reward[(size_t)e * 6 + 1 + lane] = r_intra;

    PHASE_STAMP(3, 5);
    if (block_sync) __syncthreads();
    // ---- episode metrics (metrics.py:62-71 as running sums; definitions from
    //      results/gen_results.py:791-809, 874-970, 973-1022) -----------------------------
    long long t_sent = f.sent, t_drop = dropped, t_inc = inc;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        t_sent += __shfl_xor_sync(FULL, t_sent, d);
        t_drop += __shfl_xor_sync(FULL, t_drop, d);
        t_inc += __shfl_xor_sync(FULL, t_inc, d);
    }
    bool viol = false;
    double vmn = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k)
        if (v.act[k] != 0 && v.dsl[k] < 0) { vmn = viol ? (v.dsl[k] < vmn ? v.dsl[k] : vmn) : v.dsl[k]; viol = true; }
    double add10 = 0, add11 = 0;
    double ri[S];
#pragma unroll
    for (int s = 0; s < S; ++s) ri[s] = __shfl_sync(FULL, r_intra, s);
    double* acc = st.acc + (size_t)e * IRS_NUM_METRICS;
    double a9 = acc[9], a12 = acc[12], a13 = acc[13];
#pragma unroll
    for (int pos = 0; pos < S; ++pos) { // sorted order, like the observation vector
        const uint32_t b = __ballot_sync(FULL, lane < S && v.pos == pos);
        const int sl = __ffs(b) - 1;
        const bool act_s = __shfl_sync(FULL, (int)v.active, sl);
        const long long snt = __shfl_sync(FULL, v.snt, sl);
        const double msg = __shfl_sync(FULL, v.msg, sl);
        const double pr = __shfl_sync(FULL, prio_raw, sl);
        const bool vi = __shfl_sync(FULL, (int)viol, sl);
        const double mn = __shfl_sync(FULL, vmn, sl);
        if (!act_s) continue;
        a9 += DIVC((double)snt * msg, 1e6);
        if (vi) {
            add10 += 1; a12 += mn;
            if (pr != 0) { add11 += 1; a13 += mn; }
        }
    }
    const bool done = step == c.max_steps; // comm_env.py:370
    if (lane == 0) {
        acc[0] += r0;
#pragma unroll
        for (int s = 0; s < S; ++s) acc[1 + s] += ri[s];
        acc[6] += (double)t_sent; acc[7] += (double)t_drop; acc[8] += (double)t_inc;
        acc[9] = a9; acc[10] += add10; acc[11] += add11; acc[12] = a12; acc[13] = a13;
        acc[14] += 1;
        if (done) acc[15] += 1;
        st.out_done[e] = done ? 1 : 0;
    }
    PHASE_STAMP(3, 6);
    if (done && c.auto_reset && !A.defer_rollover) {
        __syncwarp();
        reset_env_warp(A, e, -1, se_row);
    }
}

__device__ __forceinline__ void env_tail_warp(const StepArgs& A, int e, double* se_row)
{
    const DevState& st = A.st;
    const int lane = lane_id();
    const size_t eu = (size_t)e * U + (lane < U ? lane : 0);
    const int* ui = st.ue_i + eu * UI_N;
    const double* ud = st.ue_d + eu * UD_N;
    UeFrame f;
    f.occ = ud[UD_OCC]; f.lat = ud[UD_LAT]; f.sebar = ud[UD_SEBAR];
    f.sent = ui[UI_SENT]; f.sum_sent = ui[UI_SUM_SENT]; f.sum_drop = ui[UI_SUM_DROP];
    f.rb_count = st.rb[eu].y; f.occ1z = ui[UI_OCC1Z] != 0;
    // the per-UE part of the window push was done by k_ue_step; the core finishes the per-env part
    env_tail_core(A, e, st.bank + st.scen[e], f, ui[UI_DROP], ui[UI_INC], st.win_len[e], st.win_head[e], st.step[e],
                  se_row);
}

__global__ void __launch_bounds__(ENV_WARPS * 32) k_env_tail(const __grid_constant__ StepArgs A)
{
    __shared__ double sm[ENV_WARPS][136];
    const int e = A.env_begin + blockIdx.x * ENV_WARPS + (threadIdx.x >> 5);
    if (e >= A.env_end) return;
    env_tail_warp(A, e, sm[threadIdx.x >> 5]);
}

__global__ void __launch_bounds__(ENV_WARPS * 32) k_reset(const __grid_constant__ StepArgs A)
{
    __shared__ double sm[ENV_WARPS][136];
    const int e = blockIdx.x * ENV_WARPS + (threadIdx.x >> 5);
    if (e >= A.cfg.num_envs) return;
    if (A.reset_mask && !A.reset_mask[e]) return;
    const int init = A.initial_episode_dev ? A.initial_episode_dev[e] : A.initial_episode;
    reset_env_warp(A, e, init, sm[threadIdx.x >> 5]);
}

// state right after irs_bind: episode counter at its initial value, identity slice order
// (ib_sched.py:58), everything else zero.
__global__ void k_init(const __grid_constant__ StepArgs A)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= A.cfg.num_envs) return;
    int ep = A.cfg.initial_episode;
    if (A.cfg.random_episodes) { // comm_env.py:219-226: the first permutation is drawn at construction
        const int n = A.cfg.max_episodes - ep;
        A.st.ep_base[e] = ep;
        if (n > 0) ep = shuffled_episode(A.cfg.seed, A.cfg.env_id_base + e, 0, ep, n, 0);
    }
    A.st.episode[e] = ep;
    for (int s = 0; s < S; ++s) A.st.sorted[(size_t)e * S + s] = (int8_t)s;
}

// A scheduling matrix given directly (irs_actions.sched_bits) instead of the built-in action_format: RBs per UE =
// its row sum (what ib_sched.py:188-190 reads from sched_decision), first RB for the raw outputs; an RB given to two
// UEs is the reference's "same RB for more than one UE" exception (comm_env.py:593-596).  One warp per env.
__global__ void __launch_bounds__(256) k_sched_runs(const __grid_constant__ StepArgs A)
{
    const int e = A.env_begin + (int)((blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5);
    if (e >= A.env_end) return;
    const int lane = lane_id();
    int cnt = 0, first = 0;
    uint32_t w[5] = {0, 0, 0, 0, 0};
    if (lane < U) {
        const uint32_t* sb = A.sched_bits + ((size_t)e * U + lane) * 5;
#pragma unroll
        for (int i = 4; i >= 0; --i) {
            w[i] = sb[i] & (i == 4 ? 0x7fu : 0xffffffffu); // 135 = 4 * 32 + 7 RBs
            cnt += __popc(w[i]);
            if (w[i]) first = 32 * i + __ffs(w[i]) - 1;
        }
        A.st.rb[(size_t)e * U + lane] = make_short2((short)first, (short)cnt);
        if (A.raw.rb_start) A.raw.rb_start[(size_t)e * U + lane] = (int16_t)first;
        if (A.raw.rb_count) A.raw.rb_count[(size_t)e * U + lane] = (int16_t)cnt;
    }
    int total = cnt, covered = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        uint32_t o = w[i];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) o |= __shfl_xor_sync(FULL, o, d);
        covered += __popc(o);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) total += __shfl_xor_sync(FULL, total, d);
    if (lane == 0 && total != covered) atomicOr(&A.st.err[e], ERR_RB_SHARED);
}

// irs_set_episode_order: every env moves to the given table's entry at its current position
__global__ void k_apply_order(const __grid_constant__ StepArgs A)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= A.cfg.num_envs || !A.ep_order) return;
    A.st.episode[e] = A.ep_order[(size_t)e * A.ep_order_n + A.st.ep_pos[e] % A.ep_order_n];
}

// irs_read_buffers: the age-indexed view of the ring rows (buffer.py:47), one warp per (env, UE) row.
__global__ void __launch_bounds__(256) k_read_buffers(const __grid_constant__ StepArgs A, int32_t* __restrict__ out)
{
    const DevState& st = A.st;
    const size_t row_id = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row_id >= (size_t)A.cfg.num_envs * U) return;
    const int e = (int)(row_id / U), u = (int)(row_id - (size_t)e * U);
    const irs_scenario* sc = st.bank + st.scen[e];
    UeParams up = ue_params(sc, sc->ue_slice[u]);
    if (up.depth > st.row_stride) up.depth = st.row_stride;
    const int head = st.step[e] % up.depth;
    const int* row = st.buf + row_id * st.row_stride;
    int32_t* dst = out + row_id * st.row_stride;
    for (int a = lane_id(); a < st.row_stride; a += 32)
        dst[a] = a < up.depth ? row[head - a + (head < a ? up.depth : 0)] : 0;
}

// irs_check_errors: index of the first env with a non-zero error word into scratch[0] (INT_MAX: none), its word into
// scratch[1] - two launches on the stream, so that 8 bytes instead of the whole error array cross PCIe.
__global__ void __launch_bounds__(256) k_err_scan(const __grid_constant__ StepArgs A)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < A.cfg.num_envs && A.st.err[e] != 0) atomicMin(&A.st.scratch[0], e);
}
__global__ void k_err_word(const __grid_constant__ StepArgs A)
{
    const int e = A.st.scratch[0];
    A.st.scratch[1] = (e >= 0 && e < A.cfg.num_envs) ? A.st.err[e] : 0;
}

// Sum of the per-env accumulators in a FIXED order (reproducible bits for a given batch size): block k sums
// metric k; thread t adds its contiguous range of envs serially, then a shared-memory tree over the threads.
__global__ void __launch_bounds__(256) k_metrics(const __grid_constant__ StepArgs A)
{
    __shared__ double part[256];
    const int k = blockIdx.x, t = threadIdx.x;
    const int E = A.cfg.num_envs, per = (E + 255) / 256;
    const int e0 = t * per, e1 = e0 + per < E ? e0 + per : E;
    double acc = 0;
    for (int e = e0; e < e1; ++e) acc += A.st.acc[(size_t)e * IRS_NUM_METRICS + k];
    part[t] = acc;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if (t < d) part[t] += part[t + d];
        __syncthreads();
    }
    if (t == 0) A.st.metrics_total[k] = part[0];
}
