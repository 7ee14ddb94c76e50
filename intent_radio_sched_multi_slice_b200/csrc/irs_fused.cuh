// irs_fused.cuh - k_ue_env: the per-UE middle of CommunicationEnv.step (channel, traffic, throughput,
// buffers, per-UE window push; comm_env.py:302-326, ues.py:149-243) of one env in ONE warp, between the
// action and the tail kernels; and k_env_step: the whole step of a small batch in one launch.  One body
// (ue_env_body), instantiated per source of the SE rows: k_ue_env / k_env_step native generators, k_ue_env_mem rows
// from memory or constant, k_ue_env_raw / k_env_step_any chosen at run time.
//
// Lane u = UE u for everything that is per-UE scalar work (closed-form buffer arithmetic, ring-row
// update, window push, and - native generators - the per-episode mean, the serial tail of the SE row and
// the Poisson arrival).  The warp-wide stage is the 128 leading SE values of every associated UE: pair
// passes (lanes 0..15 one UE, lanes 16..31 another; a lane draws one Philox quad = eight table normals =
// the 8 elements of one accumulator of NumPy's pairwise sum, and adds them up in registers; the 8-lane tree
// of each row half leaves its sums in shared memory), then a lane = UE epilogue that adds the 7-element
// serial tail to the second half, the halves, and the reduction identity.
// Unassociated UEs are skipped entirely.  Buffer rows are rings (irs_ue.cuh): a UE lane loads the expiring
// slot and its 4 oldest slots in the prologue, before the generators run, and resolves the step's drain on
// them after the closed form; only a drain that reaches further, or a window of more than 4 slots that
// empties, is handed to the whole warp.
#pragma once
#include "irs_action.cuh"
#include "irs_env.cuh"
#include "irs_ue.cuh"

#ifndef FUSED_WARPS
#define FUSED_WARPS 4 /* warps per block */
#endif
#define GRP 4         /* UE rows generated per group */

// The 4 KB normal table of the generators goes to shared memory with cp.async at kernel entry and is waited for right
// before the generators run, so the copy overlaps the prologue's record loads.
__device__ __forceinline__ void normal_tab_request(float2* ntab)
{
    const char* src = reinterpret_cast<const char*>(g_normal_tab);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(ntab);
    for (int i = threadIdx.x; i < IRS_NORMAL_TAB_N / 2; i += blockDim.x)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16 * i), "l"(src + 16 * i));
    asm volatile("cp.async.commit_group;");
}
__device__ __forceinline__ void normal_tab_wait()
{
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
}

template <typename T>
__device__ __forceinline__ T sel4(const T (&a)[GRP], int j)
{
    T v = a[0];
#pragma unroll
    for (int k = 1; k < GRP; ++k) v = (j == k) ? a[k] : v;
    return v;
}

// FULL_STEP = false: the per-UE middle of the step, between k_action and k_env_tail (large batches: every warp of the
//                    GPU runs the same small code region).
// FULL_STEP = true : the whole step of ONE env in ONE block of NW warps and one launch - for SMALL batches
//                    (BASELINE config 2: 1,024 envs), where the step is bound by launch and dependency latency.
//                    Warp 0 turns the action into RB runs (lane = UE) and hands them over through shared memory; the
//                    block's warps share the env's UEs (u mod NW) for the middle; the per-UE frames go back through
//                    shared memory and warp 0 finishes with observation / reward / rollover.
struct StepXchg { // hand-over between the warps of a k_env_step block
    short2 rb[32];
    double occ[32], lat[32], sebar[32];
    int sent[32], sum_sent[32], sum_drop[32], dropped[32], inc[32], occ1z[32];
    double mu[32], traffic[32]; // pass A of every UE, by the block's last warp
};
// SRC = where the frame's SE rows and arrivals come from, fixed at compile time for the hot instantiations (each kernel
//       then holds one of the three row paths): 1 native generators (channel_mode 2 + traffic_mode 1), 2 rows from
//       memory or constant (channel_mode 0 / 1 / 3), 0 decided at run time (any combination; the raw-output kernel).
template <bool FULL_STEP, bool RAWSE, int NW, int SRC>
__device__ __forceinline__ void ue_env_body(const StepArgs& A, double (*sm)[GRP][136], const float2* __restrict__ ntab,
                                            StepXchg* xc = nullptr)
{
    const DevState& st = A.st;
    const irs_config& cfg = A.cfg;
    const int lane = lane_id();
    const int wib = threadIdx.x >> 5;
    const int split = FULL_STEP ? (NW == 4 ? 2 : (NW == 2 ? 1 : 0)) : A.ue_split_log2, gmask = (1 << split) - 1;
    const int wglob = blockIdx.x * NW + wib;
    const int e = A.env_begin + (wglob >> split), grp = wglob & gmask; // this warp serves the UEs u with (u & gmask) == grp
    if (e >= A.env_end) { normal_tab_wait(); return; } // every warp of the block meets the table barrier exactly once
    PHASE_CLOCK(0);

    // ---- phase 0: env scalars and this lane's UE record -----------------------------------------
    const int step = st.step[e], episode = st.episode[e];
    const int win_len = st.win_len[e], head_w = st.win_head[e];
    const int p = (head_w + 1) % A.win_cap; // window slot of this frame (deque.appendleft)
    const uint64_t key = env_key(cfg.seed, st.visit[e]);
    const irs_scenario* sc = st.bank + st.scen[e];
    const bool ue_lane = lane < U && (lane & gmask) == grp;
    const int s_lane = lane < U ? (int)sc->ue_slice[lane] : -1; // this lane's UE, whichever warp serves it
    const int s_mine = ue_lane ? s_lane : -1;
    const size_t eu = (size_t)e * U + (ue_lane ? lane : 0);
    const int4 rec_a = *reinterpret_cast<const int4*>(st.ue_i + eu * UI_N); // tot, sum_sent, sum_drop, sent
    const double2 rec_d0 = *reinterpret_cast<const double2*>(st.ue_d + eu * UD_N); // occ, lat
    int2 evict = make_int2(0, 0);
    if (ue_lane && win_len == A.win_cap) evict = st.ring[eu * W + p];
    const int tot_prev = ue_lane ? rec_a.x : 0;
    const int frame = step + 1, gid = cfg.env_id_base + e;
    const bool native = SRC == 1 ? true : (SRC == 2 ? false : (cfg.channel_mode == 2 && cfg.traffic_mode == 1));
    const bool mem_rows = SRC == 2 ? true : (SRC == 1 ? false : cfg.channel_mode != 2);
    // In the one-launch step the generators' per-UE pass (pass A below) does not wait for the action: the block's last
    // warp runs it for ALL of the env's UEs while warp 0 turns the action into RB runs.
    constexpr bool SPLIT_A = FULL_STEP && NW > 1;
    double mu_l = 0.0, traffic = 0.0;
    int rb_s = 0, rb_c = 0;

    // pass A, lane = UE: the per-episode mean (mimic_quadriga.py:37-41), the serial tail of the row (RBs 128..134, SE
    // quad 16) into the serving warp's scratch at ws[14 u ..] - (SE, throughput term) pairs, or just the seven SE values
    // when the RB runs are not known yet (SPLIT_A) - and the Poisson arrival (traffics/mult_slice.py:28-31)
    auto pass_a = [&](int s_a, double* ws_a) {
        const int fl_a = __ldg(&sc->sl[s_a].flags);
        const uint4 wm = rng_quad(key, K_MU, lane, 0, 0, episode, gid);
        mu_l = fabs(__fma_rn(7.5, (double)normal_lo(wm.x, ntab), 10.0));
        const uint4 wt = rng_quad(key, K_SE, lane, 16, frame, episode, gid);
        const uint32_t ww[4] = {wt.x, wt.y, wt.z, wt.w};
        double* tp = ws_a + 14 * lane;
        const uint32_t* sb = (RAWSE && A.sched_bits) ? A.sched_bits + ((size_t)e * U + lane) * 5 : nullptr;
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const float z = (k & 1) ? normal_hi(ww[k >> 1], ntab) : normal_lo(ww[k >> 1], ntab);
            const double x = fabs(__fma_rn(1.5, (double)z, mu_l));
            if (SPLIT_A) tp[k] = x;
            else reinterpret_cast<double2*>(tp)[k] = make_double2(x, sched_at(sb, rb_s, rb_c, 128 + k) ? A.bw_rb * x : 0.0);
            if (RAWSE && A.raw.se_used) A.raw.se_used[((size_t)e * U + lane) * R + 128 + k] = x;
        }
        if (fl_a & IRS_F_REQ) {
            const int lam_a = __ldg(&sc->sl[s_a].lam_idx);
            const uint4 wq = rng_quad(key, K_TRAFFIC, lane, 0, frame, episode, gid);
            const double u01 = ((double)(wq.x >> 5) * 67108864.0 + (double)(wq.y >> 6)) * (1.0 / 9007199254740992.0);
            const double* cdf = st.cdf + (size_t)lam_a * IRS_POISSON_K;
            int lo = 0;
#pragma unroll
            for (int stp = IRS_POISSON_K / 2; stp > 0; stp >>= 1)
                if (__ldg(&cdf[lo + stp - 1]) <= u01) lo += stp;
            if (__ldg(&cdf[lo]) <= u01) lo += 1;
            traffic = (double)lo * 1e6;
        }
    };

    if (FULL_STEP) { // ib_sched.py:223-349 on every UE's previous frame (occupancy, mean SE, window sum of sent packets)
        if (SPLIT_A) normal_tab_wait();
        if (wib == 0) {
            const size_t ea = (size_t)e * U + (lane < U ? lane : 0);
            int rs_a, rc_a;
            action_to_rb_core(A, e, sc, st.ue_d[ea * UD_N + UD_OCC], st.ue_d[ea * UD_N + UD_SEBAR],
                              st.ue_i[ea * UI_N + UI_SUM_SENT], win_len, step, episode, rs_a, rc_a);
            xc->rb[lane] = make_short2((short)rs_a, (short)rc_a);
        } else if (SPLIT_A && wib == NW - 1 && native) {
            if (s_lane >= 0) pass_a(s_lane, &sm[lane & gmask][0][0]); // an unassociated UE has SE 0 and no traffic
            xc->mu[lane] = mu_l; xc->traffic[lane] = traffic;
        }
    }
    // ---- the ring slots the step can touch (after the action on warp 0: the loads hang on a chain of three dependent
    //      loads - scenario, slice, depth - that the action must not wait for) -----------------------------------------
    UeParams up = ue_params(sc, s_mine);
    if (up.depth > st.row_stride) up.depth = st.row_stride;
    int* row = st.buf + eu * (size_t)st.row_stride;
    const int live_prev = ue_lane ? ROW_LIVE(st.ue_i[eu * UI_N + UI_PATH]) : 0; // ages [0, live_prev) may hold packets
    const long long wsum_prev = ue_lane ? st.wsum[eu] : 0;
    const int head = frame % up.depth;                 // slot of age 0 after this step
    const int a_old = min(live_prev, up.depth - 1);    // oldest age that may hold packets after the shift (0: none)
    // buffer.py:60: the packets that reach age `depth` now sit in the slot the new arrivals will take
    const int expired = (ue_lane && live_prev == up.depth) ? row[head] : 0;
    int look[RING_LOOK]; // ages a_old, a_old - 1, ...: consecutive slots from ring_slot(head, a_old)
    const int p_old = ring_slot(head, a_old, up.depth);
#pragma unroll
    for (int k = 0; k < RING_LOOK; ++k) {
        const int q = p_old + k;
        look[k] = (ue_lane && a_old - k >= 1) ? row[q >= up.depth ? q - up.depth : q] : 0;
    }
    if (FULL_STEP) {
        __syncthreads();
        rb_s = xc->rb[lane].x; rb_c = xc->rb[lane].y;
        if (SPLIT_A && native) { mu_l = xc->mu[lane]; traffic = xc->traffic[lane]; }
    } else {
        const short2 run = st.rb[eu];
        rb_s = run.x; rb_c = run.y;
    }

    PHASE_CLOCK(1);
    if (!SPLIT_A) normal_tab_wait(); // the generators' table, requested at kernel entry, is in shared memory from here on
    PHASE_CLOCK(2);
    // ---- phase 2: channel + traffic of this frame, and the two row sums per UE --------------
    const int fl_mine = s_mine >= 0 ? __ldg(&sc->sl[s_mine].flags) : 0;
    const bool has_traffic = (fl_mine & IRS_F_REQ) != 0;
    const int lam_idx = has_traffic ? __ldg(&sc->sl[s_mine].lam_idx) : 0;
    double sum_t = 0.0, sum_se = 0.0;
    if (!native) {
        if (cfg.traffic_mode == 0) traffic = ue_lane ? A.traffic[eu] : 0.0;
        else if (has_traffic) traffic = native_traffic(key, st.cdf, lane, lam_idx, frame, episode, gid);
    }
    if (native) {
        // Per-warp scratch in shared memory: the tail of UE u at ws[14 u .. 14 u + 13], per-UE results
        // (sum_t, sum_se) at ws[352 + 4 u ..].
        double* ws = &sm[wib][0][0];
        const bool gen = ue_lane && s_mine >= 0; // an unassociated UE has SE 0 and no traffic in native mode
        if (!SPLIT_A && gen) pass_a(s_mine, ws);
        __syncwarp();
        PHASE_CLOCK(3);
        // pair passes: lanes 0..15 serve one associated UE, lanes 16..31 the next.  Lane m of a side generates
        // SE quad m = the 8 elements accumulator m & 7 of row half m >> 3 visits (se_index), and adds them up in
        // registers in NumPy's order: both 135-term sums (ues.py:108-116 throughput terms, np.mean(SE)) without a
        // shared-memory row.
        uint32_t todo = __ballot_sync(FULL, gen);
#pragma unroll 1
        while (todo != 0) { // warp-uniform
            const int u0 = __ffs(todo) - 1;
            todo &= todo - 1;
            const bool live1 = todo != 0;                 // odd count: side 1 of the last pass shadows side 0
            const int u1 = live1 ? __ffs(todo) - 1 : u0;
            todo &= todo - 1;
            const int side = lane >> 4;
            const int uj = side ? u1 : u0;
            const bool live = side ? live1 : true;
            const double mu = __shfl_sync(FULL, mu_l, uj);
            const int rs = __shfl_sync(FULL, rb_s, uj), rc = __shfl_sync(FULL, rb_c, uj);
            const int l16 = lane & 15, base = ((l16 >> 3) << 6) + (l16 & 7);
            const unsigned d0 = (unsigned)(base - rs);
            double acc_se = 0.0, acc_t = 0.0;
            double* se_out = (RAWSE && A.raw.se_used) ? A.raw.se_used + ((size_t)e * U + uj) * R : nullptr;
            const uint4 wb = rng_quad(key, K_SE, uj, l16, frame, episode, gid);
            const uint32_t ww[4] = {wb.x, wb.y, wb.z, wb.w};
            const uint32_t* sb = (RAWSE && A.sched_bits) ? A.sched_bits + ((size_t)e * U + uj) * 5 : nullptr;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float y = (k & 1) ? normal_hi(ww[k >> 1], ntab) : normal_lo(ww[k >> 1], ntab);
                const double x = fabs(__fma_rn(1.5, (double)y, mu)); // channels/mimic_quadriga.py:49-56
                const bool in_run = (RAWSE && sb) ? sched_at(sb, 0, 0, base + 8 * k)
                                                 : d0 + 8u * k < (unsigned)rc; // ((BW/R) * sched) * SE; adding the 0.0 of an
                if (k == 0) { acc_se = x; acc_t = in_run ? A.bw_rb * x : 0.0; } // unscheduled RB changes nothing (terms >= 0)
                else { acc_se += x; if (in_run) acc_t += A.bw_rb * x; }
                if (RAWSE && se_out) se_out[base + 8 * k] = x;
            }
            acc_se = acc_se + __shfl_down_sync(FULL, acc_se, 1);
            acc_t = acc_t + __shfl_down_sync(FULL, acc_t, 1);
            acc_se = acc_se + __shfl_down_sync(FULL, acc_se, 2);
            acc_t = acc_t + __shfl_down_sync(FULL, acc_t, 2);
            acc_se = acc_se + __shfl_down_sync(FULL, acc_se, 4);
            acc_t = acc_t + __shfl_down_sync(FULL, acc_t, 4);
            // the two half-row trees of the UE: (SE, term) of half l16 >> 3 at ws[352 + 4 u + 2 half]
            if ((l16 & 7) == 0 && live) *reinterpret_cast<double2*>(ws + 352 + 4 * uj + (l16 >> 2)) = make_double2(acc_se, acc_t);
        }
        __syncwarp();
        // lane = UE again: the 7-element tail is added serially to the second half's tree, then the halves, then the
        // reduction identity - once for all UEs instead of once per pass on one lane of each side
        if (gen) {
            const double2 h0 = *reinterpret_cast<const double2*>(ws + 352 + 4 * lane);
            double2 h1 = *reinterpret_cast<const double2*>(ws + 352 + 4 * lane + 2);
            const double* tp = ws + 14 * lane;
            const uint32_t* sb = (RAWSE && A.sched_bits) ? A.sched_bits + eu * 5 : nullptr;
#pragma unroll
            for (int k = 0; k < 7; ++k) {
                if (SPLIT_A) {
                    const double xt = tp[k];
                    h1.x += xt;
                    h1.y += sched_at(sb, rb_s, rb_c, 128 + k) ? A.bw_rb * xt : 0.0;
                } else {
                    const double2 xt = reinterpret_cast<const double2*>(tp)[k];
                    h1.x += xt.x;
                    h1.y += xt.y;
                }
            }
            sum_se = 0. + (h0.x + h1.x);
            sum_t = 0. + (h0.y + h1.y);
        }
    } else if (mem_rows) {
        // Rows that come from memory (injected frame, SE trace bank) or are constant (fixed_se):
        // the same two-UEs-per-pass register reduction, each lane loading the 8 elements of its
        // accumulator straight from global memory (every 32-byte sector fully used).  Rows of
        // unassociated UEs are not read: their RB run is empty, so their throughput is 0 whatever
        // the row holds (ues.py:108-116), and their mean SE is never observed (ib_sched.py:111,147).
        const double* frame_base = nullptr;
        if (cfg.channel_mode == 0) frame_base = A.se + (size_t)e * U * R;
        else if (cfg.channel_mode == 3) frame_base = trace_frame(A, e, episode, step);
        double* ws = &sm[wib][0][0];
        const bool all_rows = RAWSE && A.raw.se_used;
        uint32_t todo = __ballot_sync(FULL, ue_lane && (s_mine >= 0 || all_rows));
        // L2 prefetch of the rows of the pair after next (lanes 0..9 / 16..25: the <= 10 128-byte lines a
        // 1080-byte row touches): the loads of a pass then find their rows in L2 instead of waiting on HBM
        auto prefetch_pair = [&](uint32_t m, int skip_pairs) {
            if (!frame_base) return;
            for (int i = 0; i < 2 * skip_pairs; ++i) m &= m - 1;
            if (!m) return;
            const int ua = __ffs(m) - 1;
            m &= m - 1;
            const int ub = m ? __ffs(m) - 1 : -1;
            const int up_ = (lane >> 4) ? ub : ua, line = lane & 15;
            if (up_ >= 0 && line < 10) {
                const char* row0 = reinterpret_cast<const char*>(frame_base + (size_t)up_ * R);
                const char* pl = reinterpret_cast<const char*>(reinterpret_cast<uintptr_t>(row0) & ~(uintptr_t)127) + 128 * line;
                if (pl < row0 + R * 8) asm volatile("prefetch.global.L2 [%0];" ::"l"(pl));
            }
        };
#ifndef SE_PREFETCH_AHEAD
#define SE_PREFETCH_AHEAD 2 /* pairs of rows requested ahead of the pass that sums them */
#endif
#pragma unroll
        for (int k = 0; k < SE_PREFETCH_AHEAD; ++k) prefetch_pair(todo, k);
        while (todo != 0) {
            prefetch_pair(todo, SE_PREFETCH_AHEAD);
            const int u0 = __ffs(todo) - 1;
            todo &= todo - 1;
            const bool live1 = todo != 0;
            const int u1 = live1 ? __ffs(todo) - 1 : u0;
            todo &= todo - 1;
            const int side = lane >> 4;
            const int uj = side ? u1 : u0;
            const bool live = side ? live1 : true;
            const int rs = __shfl_sync(FULL, rb_s, uj), rc = __shfl_sync(FULL, rb_c, uj);
            const int l16 = lane & 15, base = ((l16 >> 3) << 6) + (l16 & 7);
            const double* rowp = frame_base ? frame_base + (size_t)uj * R : nullptr;
            double x[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) x[k] = rowp ? __ldg(rowp + base + 8 * k) : 2.0; // channels/fixed_se.py:26
            const uint32_t* sb = (RAWSE && A.sched_bits) ? A.sched_bits + ((size_t)e * U + uj) * 5 : nullptr;
            // ((BW/R) * sched) * SE, ues.py:108-116; the + 0.0 of an unscheduled RB changes nothing (terms >= 0)
            double acc_se = x[0];
            double acc_t = sched_at(sb, rs, rc, base) ? A.bw_rb * x[0] : 0.0;
#pragma unroll
            for (int k = 1; k < 8; ++k) {
                acc_se += x[k];
                if (sched_at(sb, rs, rc, base + 8 * k)) acc_t += A.bw_rb * x[k];
            }
            acc_se = acc_se + __shfl_down_sync(FULL, acc_se, 1);
            acc_t = acc_t + __shfl_down_sync(FULL, acc_t, 1);
            acc_se = acc_se + __shfl_down_sync(FULL, acc_se, 2);
            acc_t = acc_t + __shfl_down_sync(FULL, acc_t, 2);
            acc_se = acc_se + __shfl_down_sync(FULL, acc_se, 4);
            acc_t = acc_t + __shfl_down_sync(FULL, acc_t, 4);
            // the two half-row trees of the UE: (SE, term) of half l16 >> 3 at ws[64 + 4 u + 2 half]
            if ((l16 & 7) == 0 && live) *reinterpret_cast<double2*>(ws + 64 + 4 * uj + (l16 >> 2)) = make_double2(acc_se, acc_t);
            if (all_rows && live) {
                double* se_out = A.raw.se_used + ((size_t)e * U + uj) * R;
#pragma unroll
                for (int k = 0; k < 8; ++k) se_out[base + 8 * k] = x[k];
            }
        }
        __syncwarp();
        // lane = UE: the 7-element tail (RBs 128..134; its sectors were just read by the UE's pass) is added serially to
        // the second half's tree, then the halves, then the reduction identity
        if (ue_lane && (s_mine >= 0 || all_rows)) {
            const double2 h0 = *reinterpret_cast<const double2*>(ws + 64 + 4 * lane);
            double2 h1 = *reinterpret_cast<const double2*>(ws + 64 + 4 * lane + 2);
            const double* rowp = frame_base ? frame_base + (size_t)lane * R : nullptr;
            const uint32_t* sb = (RAWSE && A.sched_bits) ? A.sched_bits + eu * 5 : nullptr;
#pragma unroll
            for (int k = 0; k < 7; ++k) {
                const double xt = rowp ? __ldg(rowp + 128 + k) : 2.0;
                h1.x += xt;
                if (sched_at(sb, rb_s, rb_c, 128 + k)) h1.y += A.bw_rb * xt;
                if (all_rows) A.raw.se_used[eu * R + 128 + k] = xt;
            }
            sum_se = 0. + (h0.x + h1.x);
            sum_t = 0. + (h0.y + h1.y);
        }
    } else {
    uint32_t todo = __ballot_sync(FULL, ue_lane);
    while (todo != 0) { // warp-uniform
        int ul[GRP], n = 0;
#pragma unroll
        for (int j = 0; j < GRP; ++j) {
            ul[j] = todo ? (__ffs(todo) - 1) : 0;
            if (todo) { ++n; todo &= todo - 1; }
        }
#pragma unroll 1
        for (int jj = 0; jj < n; ++jj) {
            const int uj = sel4(ul, jj);
            const int assoc = __shfl_sync(FULL, (int)(s_mine >= 0), uj);
            channel_row(A, key, e, uj, assoc != 0, frame, episode, sm[wib][jj]);
        }
        __syncwarp();
#pragma unroll 1
        for (int jj = 0; jj < n; ++jj) {
            const int uj = sel4(ul, jj);
            const int rs = __shfl_sync(FULL, rb_s, uj), rc = __shfl_sync(FULL, rb_c, uj);
            const double sums = np_sum135_thr_se(sm[wib][jj], rs, rs + rc, A.bw_rb,
                                                 (RAWSE && A.sched_bits) ? A.sched_bits + ((size_t)e * U + uj) * 5 : nullptr);
            const double a = __shfl_sync(FULL, sums, 0), b = __shfl_sync(FULL, sums, 16);
            if (lane == uj) { sum_t = a; sum_se = b; }
            if (RAWSE && A.raw.se_used)
                for (int r = lane; r < R; r += 32) A.raw.se_used[((size_t)e * U + uj) * R + r] = sm[wib][jj][r];
        }
        __syncwarp();
    }
    }
    if (RAWSE && native && A.raw.se_used) { // unassociated UEs have SE 0 in native mode
        for (int u = 0; u < U; ++u) {
            if ((u & gmask) != grp || __shfl_sync(FULL, s_mine, u) >= 0) continue;
            for (int r = lane; r < R; r += 32) A.raw.se_used[((size_t)e * U + u) * R + r] = 0.0;
        }
    }


    PHASE_CLOCK(4);
    // ---- phase 3: throughput, arrivals, closed-form buffer arithmetic - one UE per lane -------
    const double thr = floor(sum_t / up.pkt_bits);   // ues.py:108-116
    const double inc = floor(traffic / up.pkt_bits); // ues.py:199
    const double sebar = DIVC(sum_se, R);         // np.mean(SE[0,u,:])
    const int thr_i = (int)fmin(fmax(thr, 0.0), 2147483647.0);
    const int total = tot_prev - expired;            // occupancy after the age shift (buffer.py:60-62)
    int dropped = expired, kept;
    if ((double)total + inc <= (double)up.max_pkts) { // buffer.py:62-80
        kept = (int)inc;
    } else {
        dropped += (int)(long long)(inc - (double)(up.max_pkts - total));
        kept = up.max_pkts - total;
    }
    const int sent_old = min(thr_i, total);          // oldest first (buffer.py:97-103)
    const int a_left = kept - min(kept, thr_i - sent_old);
    const int sent = sent_old + (kept - a_left);
    const int rem = total + kept - sent;

    // ---- phase 4: the ring rows (buffer.py:53-104) --------------------------------------------
    int path = ROW_NONE, live = a_left != 0;
    int acc = (live_prev == up.depth) + min(RING_LOOK, a_old); // 4-byte slot accesses (reads so far)
    long long wdec = 0;
    int want = sent_old;
    int a_next = a_old; // first age the warp has to continue from (ROW_WARP)
    if (ue_lane && total > 0) {
        if (sent_old == total) { // all older packets leave: ages [1, a_old] become empty
            path = ROW_CLEAR;
            if (a_old <= RING_LOOK) {
#pragma unroll
                for (int k = 0; k < RING_LOOK; ++k)
                    if (look[k] != 0) { const int q = p_old + k; row[q >= up.depth ? q - up.depth : q] = 0; ++acc; }
            }
        } else { // the oldest `sent_old` packets leave; some older packet stays
            path = ROW_LANE;
            live = 0;
#pragma unroll
            for (int k = 0; k < RING_LOOK; ++k) {
                const int a = a_old - k;
                if (live == 0 && a >= 1) {
                    const int n = look[k], take = min(n, want);
                    if (take) {
                        const int q = p_old + k;
                        row[q >= up.depth ? q - up.depth : q] = n - take;
                        wdec += (long long)a * take;
                        want -= take;
                        ++acc;
                    }
                    if (n - take > 0) live = a + 1;
                    a_next = a - 1;
                }
            }
            if (live == 0) {
                if (want > 0 && a_next >= 1) path = ROW_WARP;
                else live = max(a_next, 0) + 1; // ages above a_next are empty
            }
        }
    }
    if (ue_lane && (a_left != 0 || expired != 0)) { row[head] = a_left; ++acc; }
    uint32_t clr = __ballot_sync(FULL, path == ROW_CLEAR && a_old > RING_LOOK);
    uint32_t wrp = __ballot_sync(FULL, path == ROW_WARP);
    while (clr) {
        const int u = __ffs(clr) - 1;
        clr &= clr - 1;
        const int ao = __shfl_sync(FULL, a_old, u);
        ring_clear_warp(st.buf + ((size_t)e * U + u) * st.row_stride, __shfl_sync(FULL, up.depth, u),
                        __shfl_sync(FULL, head, u), ao);
        if (lane == u) acc += ao;
    }
    while (wrp) {
        const int u = __ffs(wrp) - 1;
        wrp &= wrp - 1;
        int lv = 0, ac = 0;
        const long long w = ring_drain_warp(st.buf + ((size_t)e * U + u) * st.row_stride, __shfl_sync(FULL, up.depth, u),
                                            __shfl_sync(FULL, head, u), __shfl_sync(FULL, a_next, u),
                                            __shfl_sync(FULL, want, u), lv, ac);
        if (lane == u) { wdec += w; live = lv; acc += ac; }
    }
    // sum(age * packets) after the step: every remaining packet one step older, the drained ones gone
    const long long wsum = (rem == 0 || path == ROW_CLEAR || path == ROW_NONE)
                               ? 0 : wsum_prev + (long long)tot_prev - (long long)up.depth * expired - wdec;

    // ---- phase 5: occupancy, mean delay, window push, record stores ---------------------------
    const double occ = (double)rem / (double)up.max_pkts;               // buffer.py:106-114
    const double lat = rem != 0 ? (double)wsum / (double)rem : 0.0;      // buffer.py:116-130
    const int sum_sent = rec_a.y + sent - evict.x;
    const int sum_drop = rec_a.z + dropped - evict.y;
    const bool occ1z = isclose0(rec_d0.x); // win[1].occ of the new window
    if (ue_lane) {
        *reinterpret_cast<int4*>(st.ue_i + eu * UI_N) = make_int4(rem, sum_sent, sum_drop, sent);
        *reinterpret_cast<int4*>(st.ue_i + eu * UI_N + 4) = make_int4(dropped, (int)inc, occ1z ? 1 : 0, ROW_WORD(path, live, min(acc, 0xffff)));
        *reinterpret_cast<double2*>(st.ue_d + eu * UD_N) = make_double2(occ, lat);
        *reinterpret_cast<double2*>(st.ue_d + eu * UD_N + 2) = make_double2(sebar, thr);
        st.ring[eu * W + p] = make_int2(sent, dropped);
        st.wsum[eu] = wsum;
        if (A.raw.pkt_incoming) A.raw.pkt_incoming[eu] = inc;
        if (A.raw.pkt_throughputs) A.raw.pkt_throughputs[eu] = thr;
        if (A.raw.sent) A.raw.sent[eu] = sent;
        if (A.raw.dropped) A.raw.dropped[eu] = dropped;
        if (A.raw.occ) A.raw.occ[eu] = occ;
        if (A.raw.lat) A.raw.lat[eu] = lat;
        if (A.raw.traffic_used) A.raw.traffic_used[eu] = traffic;
        if (FULL_STEP) st.rb[eu] = make_short2((short)rb_s, (short)rb_c);
    }
    if (!FULL_STEP) return;
    // ---- phase 6: observation, reward, metrics, done, rollover (comm_env.py:339-373), by warp 0 on every UE's frame
    if (ue_lane) {
        xc->occ[lane] = occ; xc->lat[lane] = lat; xc->sebar[lane] = sebar;
        xc->sent[lane] = sent; xc->sum_sent[lane] = sum_sent; xc->sum_drop[lane] = sum_drop;
        xc->dropped[lane] = dropped; xc->inc[lane] = (int)inc; xc->occ1z[lane] = occ1z ? 1 : 0;
    }
    __syncthreads();
    PHASE_CLOCK(5);
    if (wib != 0) return;
    UeFrame f;
    const int lu = lane < U ? lane : 0;
    f.occ = xc->occ[lu]; f.lat = xc->lat[lu]; f.sebar = xc->sebar[lu]; f.sent = xc->sent[lu];
    f.sum_sent = xc->sum_sent[lu]; f.sum_drop = xc->sum_drop[lu]; f.rb_count = xc->rb[lu].y; f.occ1z = xc->occ1z[lu] != 0;
    env_tail_core(A, e, sc, f, xc->dropped[lu], xc->inc[lu], win_len, head_w, step, sm[0][0]);
    PHASE_CLOCK(6);
}

#ifndef UE_ENV_MINB
#define UE_ENV_MINB 8 /* 64 registers: 32 warps per SM measured fastest (profiles/r1_notes.md) */
#endif
// native generators (the default of the batched env)
__global__ void __launch_bounds__(FUSED_WARPS * 32, UE_ENV_MINB) k_ue_env(const __grid_constant__ StepArgs A)
{
    __shared__ __align__(16) double sm[FUSED_WARPS][GRP][136];
    __shared__ __align__(16) float2 ntab[IRS_NORMAL_TAB_N];
    normal_tab_request(ntab);
    ue_env_body<false, false, FUSED_WARPS, 1>(A, sm, ntab);
}
// rows from memory (SE trace bank, injected frames) or constant: no generator table in shared memory
__global__ void __launch_bounds__(FUSED_WARPS * 32, UE_ENV_MINB) k_ue_env_mem(const __grid_constant__ StepArgs A)
{
    __shared__ __align__(16) double sm[FUSED_WARPS][GRP][136];
    ue_env_body<false, false, FUSED_WARPS, 2>(A, sm, nullptr);
}
// every combination, with the optional per-RB SE output (irs_raw_outputs.se_used; tests and trace dumps) and given
// scheduling matrices
__global__ void __launch_bounds__(FUSED_WARPS * 32, UE_ENV_MINB) k_ue_env_raw(const __grid_constant__ StepArgs A)
{
    __shared__ __align__(16) double sm[FUSED_WARPS][GRP][136];
    __shared__ __align__(16) float2 ntab[IRS_NORMAL_TAB_N];
    normal_tab_request(ntab);
    ue_env_body<false, true, FUSED_WARPS, 0>(A, sm, ntab);
}

// the whole step in one launch, one block of STEP_WARPS warps per env (small batches; IBSched-family surface, no per-RB
// raw output): native generators | any source
#ifndef STEP_WARPS
#define STEP_WARPS 2
#endif
#ifndef ENV_STEP_MINB
#define ENV_STEP_MINB 8
#endif
__global__ void __launch_bounds__(STEP_WARPS * 32, ENV_STEP_MINB) k_env_step(const __grid_constant__ StepArgs A)
{
    __shared__ __align__(16) double sm[STEP_WARPS][GRP][136];
    __shared__ __align__(16) float2 ntab[IRS_NORMAL_TAB_N];
    __shared__ StepXchg xc;
    normal_tab_request(ntab);
    ue_env_body<true, false, STEP_WARPS, 1>(A, sm, ntab, &xc);
}
__global__ void __launch_bounds__(STEP_WARPS * 32, ENV_STEP_MINB) k_env_step_any(const __grid_constant__ StepArgs A)
{
    __shared__ __align__(16) double sm[STEP_WARPS][GRP][136];
    __shared__ __align__(16) float2 ntab[IRS_NORMAL_TAB_N];
    __shared__ StepXchg xc;
    normal_tab_request(ntab);
    ue_env_body<true, false, STEP_WARPS, 0>(A, sm, ntab, &xc);
}
