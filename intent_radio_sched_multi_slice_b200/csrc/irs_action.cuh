// irs_action.cuh - K3: action -> resource-block runs, one warp per env, lane u = UE u.
//
// Replaces IBSched.action_format (agents/ib_sched.py:223-349) with its helpers
// scores_to_rbs / round_int_equal_sum / round_robin / proportional_fairness /
// max_throughput / distribute_rbs_ues (agents/common.py:442-701) and, optionally, the
// trivial policies that drive it (agents/marr.py:40-47, agents/mapf.py:41-111).
// The dense 1 x 25 x 135 0/1 matrix of the reference is never materialised: a UE's RBs are
// one contiguous run, stored as (rb_start, rb_count).
#pragma once
#include "irs_common.cuh"
#include "irs_rng.cuh"

// agents/common.py:481-505 on 5 values held by one thread (fully unrolled): the per-slice rounding of the thread- and
// lane8-per-env stages.
__device__ __forceinline__ void round_int_equal_sum5(const double (&x)[S], int target, int (&out)[S])
{
    double res = 0.;
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < S; ++i)
        if (x[i] != 0) { res += x[i]; ++cnt; }
    const double sum = 0. + res;
    const Recip rsum = recip_of(sum); // all x zero: sum = 0 and no quotient is used
    int tot = 0;
#pragma unroll
    for (int i = 0; i < S; ++i) {
        out[i] = (x[i] != 0) ? (int)floor(div_by((double)target * x[i], rsum)) : 0;
        tot += out[i];
    }
    const int adj = target - tot;
#pragma unroll
    for (int i = 0; i < S; ++i) {
        if (x[i] == 0) continue;
        int rank = 0; // descending value, ties -> higher index first (stable argsort reversed)
#pragma unroll
        for (int j = 0; j < S; ++j)
            if (x[j] != 0 && (x[j] > x[i] || (x[j] == x[i] && j > i))) ++rank;
        if (adj > rank) out[i] += idiv_small(adj - rank - 1, cnt) + 1;
    }
}

// agents/common.py:442-461 by one thread (the thread-per-env stage)
__device__ __forceinline__ void scores_to_rbs5(const double (&action)[S], int total, const bool (&active)[S],
                                               int (&out)[S])
{
    double a1[S], x[S];
    double res = 0.;
#pragma unroll
    for (int i = 0; i < S; ++i) { a1[i] = action[i] + 1; res += a1[i]; }
    const double sum = 0. + res;
    if (sum != 0) {
#pragma unroll
        for (int i = 0; i < S; ++i) x[i] = ((double)total * a1[i]) / sum;
    } else {
        double sa = 0.;
#pragma unroll
        for (int i = 0; i < S; ++i) sa += active[i] ? 1.0 : 0.0;
        sa = 0. + sa;
#pragma unroll
        for (int i = 0; i < S; ++i) x[i] = ((double)total / sa) * (active[i] ? 1.0 : 0.0);
    }
    round_int_equal_sum5(x, total, out);
}

// agents/common.py:442-461 scores_to_rbs + :481-505 round_int_equal_sum on the five scores every lane holds.  The serial
// sums are evaluated by every lane in the reference's order; each of the two float64 divisions per score runs on the
// score's own lane (lane s = slice s, lanes >= 5 shadow slice 0) and the five results are shared by shuffles: two
// dependent divisions on the warp's critical path instead of ten.
__device__ __forceinline__ void scores_to_rbs5_lanes(const double (&action)[S], int total, const bool (&active)[S],
                                                     int (&out)[S])
{
    const int lane = lane_id();
    const int s = lane < S ? lane : 0;
    double a1_mine = 0, act_mine = 0;
    double res = 0.;
#pragma unroll
    for (int i = 0; i < S; ++i) {
        const double a1 = action[i] + 1;
        res += a1;
        if (i == s) { a1_mine = a1; act_mine = active[i] ? 1.0 : 0.0; }
    }
    const double sum = 0. + res;
    double x_mine;
    if (sum != 0) {
        x_mine = ((double)total * a1_mine) / sum;
    } else {
        double sa = 0.;
#pragma unroll
        for (int i = 0; i < S; ++i) sa += active[i] ? 1.0 : 0.0;
        sa = 0. + sa;
        x_mine = ((double)total / sa) * act_mine;
    }
    double x[S];
    double res2 = 0.;
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < S; ++i) {
        x[i] = __shfl_sync(FULL, x_mine, i);
        if (x[i] != 0) { res2 += x[i]; ++cnt; }
    }
    const double sum2 = 0. + res2;
    int out_mine = (x_mine != 0) ? (int)floor(((double)total * x_mine) / sum2) : 0;
    int tot = 0, rank = 0; // rank: descending value, ties -> higher index first (stable argsort reversed)
#pragma unroll
    for (int i = 0; i < S; ++i) {
        tot += __shfl_sync(FULL, out_mine, i);
        if (x[i] != 0 && (x[i] > x_mine || (x[i] == x_mine && i > s))) ++rank;
    }
    const int adj = total - tot;
    if (x_mine != 0 && adj > rank) out_mine += idiv_small(adj - rank - 1, cnt) + 1;
#pragma unroll
    for (int i = 0; i < S; ++i) out[i] = __shfl_sync(FULL, out_mine, i);
}

// agents/common.py:508-555: share of this lane when `rbs_s` RBs go round-robin to `elmask`.
__device__ __forceinline__ int rr_share(uint32_t elmask, int rbs_s)
{
    const int lane = lane_id();
    if (!((elmask >> lane) & 1u)) return 0;
    const int m = __popc(elmask);
    const int i = __popc(elmask & ((1u << lane) - 1u));
    const int q = idiv_small(rbs_s, m);
    return q + (i < rbs_s - q * m ? 1 : 0);
}

template <int N>
__device__ __forceinline__ double sel5(const double (&a)[N], int i)
{
    double v = a[0];
#pragma unroll
    for (int k = 1; k < N; ++k) v = (i == k) ? a[k] : v;
    return v;
}
template <int N>
__device__ __forceinline__ int sel5i(const int (&a)[N], int i)
{
    int v = a[0];
#pragma unroll
    for (int k = 1; k < N; ++k) v = (i == k) ? a[k] : v;
    return v;
}

// Core of K3 on register inputs (this lane's UE: previous-frame occupancy, mean SE, window sum
// of sent packets); returns this lane's RB run.  Shared by k_action and the fused k_env_step.
__device__ __forceinline__ void action_to_rb_core(const StepArgs& A, int e, const irs_scenario* sc, double occ,
                                                  double sebar, int sum_sent, int win_len, int step, int episode,
                                                  int& start_out, int& count_out)
{
    const DevState& st = A.st;
    const irs_config& cfg = A.cfg;
    const int lane = lane_id();
    PHASE_STAMP(2, 0);
    const int s_mine = lane < U ? (int)sc->ue_slice[lane] : -1;
    uint32_t smask[S], gmask = 0;
#pragma unroll
    for (int s = 0; s < S; ++s) {
        smask[s] = __ballot_sync(FULL, s_mine == s);
        if (s_mine == s) gmask = smask[s];
    }
    const GroupIdx G = group_index(gmask);
    bool active[S];
    int nact = 0;
#pragma unroll
    for (int s = 0; s < S; ++s) { active[s] = (__ldg(&sc->sl[s].flags) & IRS_F_ACTIVE) != 0; nact += active[s]; }

    const UeParams up = ue_params(sc, s_mine);

    // ---- where the action comes from ---------------------------------------------------
    double a0_in[S];
    int intra[S] = {0, 0, 0, 0, 0};
    if (cfg.policy == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) a0_in[s] = A.a0[(size_t)e * S + s];
        if (A.intra) {
#pragma unroll
            for (int s = 0; s < S; ++s) intra[s] = A.intra[(size_t)e * S + s];
        }
    } else if (cfg.policy == 1) { // agents/marr.py:40-47
#pragma unroll
        for (int s = 0; s < S; ++s) a0_in[s] = __ldg(&sc->sl[s].n_ues) > 0 ? 1.0 : -1.0;
    } else if (cfg.policy == 2) { // agents/mapf.py:41-111
        const int n = G.n;
        const double occ_mean = div_small(group_sum_ordered(occ, G), n);
        const double snt_mean = div_small(group_sum_ordered(div_small((double)sum_sent, win_len), G), n);
        double occ_mbps[S], sent_mbps[S], w[S];
        double mx = 0;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const int src = smask[s] ? (__ffs(smask[s]) - 1) : lane;
            double om = __shfl_sync(FULL, occ_mean, src);
            double sm = __shfl_sync(FULL, snt_mean, src);
            if (!smask[s]) { om = 0.0 / 0.0; sm = 0.0 / 0.0; } // np.mean of an empty selection
            const double bsz = __ldg(&sc->sl[s].buffer_size), msg = __ldg(&sc->sl[s].message_size);
            occ_mbps[s] = active[s] ? DIVC((om * bsz) * msg, 1e6) : 0.0;
            sent_mbps[s] = active[s] ? DIVC(sm * msg, 1e6) : 0.0;
            mx = (s == 0 || occ_mbps[s] > mx) ? occ_mbps[s] : mx;
        }
        double res = 0.;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            w[s] = !isclose0(sent_mbps[s]) ? occ_mbps[s] / sent_mbps[s] : 2 * mx;
            if (!active[s]) w[s] = 0;
            res += w[s];
        }
        const double sw = 0. + res;
#pragma unroll
        for (int s = 0; s < S; ++s) a0_in[s] = (sw > 0 ? w[s] / sw : 2.0) - 1;
    } else { // random actions from the counter-based generator (BASELINE config 5)
        uint32_t w[12];
        const int gid = cfg.env_id_base + e;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const uint4 v = rng_quad(env_key(cfg.seed, st.visit[e]), K_POLICY, 0, q, step + 1, episode, gid);
            w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int s = 0; s < S; ++s) {
            a0_in[s] = (double)w[s] * (1.0 / 2147483648.0) - 1.0;
            intra[s] = (int)(((uint64_t)w[5 + s] * 3u) >> 32);
        }
    }
    if (lane < S) {
        if (A.raw.a0_used) A.raw.a0_used[(size_t)e * S + lane] = sel5(a0_in, lane);
        if (A.raw.intra_used) A.raw.intra_used[(size_t)e * S + lane] = sel5i(intra, lane);
    }

    // ---- ib_sched.py:223-349 -----------------------------------------------------------
    int count = 0, start = 0;
    PHASE_STAMP(2, 1);
    if (nact != 0) { // warp-uniform
        double a0[S];
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const int src = st.sorted[(size_t)e * S + s];
            a0[s] = active[s] ? sel5(a0_in, src) : -1.0; // :247 (quirk Q3), :248-255
        }
        int rbs_per_slice[S];
        scores_to_rbs5_lanes(a0, R / IRS_RBS_PER_RBG, active, rbs_per_slice);
        PHASE_STAMP(2, 2);
        int base_s = 0, rbs_s = 0, algo = 0;
        {
            int run = 0;
#pragma unroll
            for (int s = 0; s < S; ++s) {
                rbs_per_slice[s] *= IRS_RBS_PER_RBG;
                if (s_mine == s) { base_s = run; rbs_s = rbs_per_slice[s]; algo = intra[s]; }
                run += rbs_per_slice[s];
            }
        }
        if (cfg.fixed_intra >= 0) algo = cfg.fixed_intra;
        const bool member = s_mine >= 0;
        const int n = G.n;
        // round robin over UEs with a non-empty buffer (all of them if none) - common.py:517-524
        uint32_t elmask = gmask & __ballot_sync(FULL, member && !isclose0(occ));
        if (elmask == 0) elmask = gmask;
        const int rr_buf = rr_share(elmask, rbs_s);
        const int rr_all = rr_share(gmask, rbs_s); // fallback: account_buffer=False
        // throughput each UE could use - common.py:567-583, 648-664
        const double ta_a = div_small(sebar * DIVC((double)rbs_s * cfg.bandwidth, R), n);
        const double ta_b = (occ * (double)up.max_pkts) * up.pkt_bits;
        const double ta = ta_a < ta_b ? ta_a : ta_b;
        PHASE_STAMP(2, 3);
        // proportional fairness (common.py:558-636) weighs by ta / mean sent throughput, max throughput (:639-701) by ta:
        // the same normalise-and-round over the slice's UEs.  A slice runs ONE of the two, so each lane feeds the value
        // of its own slice's algorithm through a single pass.
        double val = ta;
        if (__any_sync(FULL, member && algo == 1)) {
            const double tamax = group_max(ta, G);
            double snt = div_small((double)sum_sent, win_len) * up.pkt_bits;
            if (isclose0(ta)) snt = 1;
            const double w = !isclose0(snt) ? ta / snt : 2 * tamax;
            if (algo == 1) val = w;
        }
        PHASE_STAMP(2, 4);
        const double sv = group_sum_ordered(val, G);
        const int share_r = group_round_int_equal_sum(((double)rbs_s * val) / sv, G, rbs_s);
        const int share = sv != 0 ? share_r : rr_all;
        PHASE_STAMP(2, 5);
        if (member) {
            if (algo == 0) count = rr_buf;
            else if (algo == 1 || algo == 2) count = share;
            else atomicOr(&st.err[e], ERR_INTRA_ACTION); // ValueError, ib_sched.py:341-344
        }
        // common.py:464-478 distribute_rbs_ues: consecutive runs inside the slice's block
        start = base_s + group_prefix_excl(count, G);
        if (count <= 0) start = 0;
        // ib_sched.py:345-347 assert sum == 135
        int total = count;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) total += __shfl_xor_sync(FULL, total, d);
        int expect = 0;
#pragma unroll
        for (int s = 0; s < S; ++s) expect += smask[s] ? rbs_per_slice[s] : 0;
        if (lane == 0 && total != expect) atomicOr(&st.err[e], ERR_RB_SUM);
    }
    PHASE_STAMP(2, 6);
    start_out = start;
    count_out = count;
    if (lane < U) {
        const size_t eu = (size_t)e * U + lane;
        if (A.raw.rb_start) A.raw.rb_start[eu] = (int16_t)start;
        if (A.raw.rb_count) A.raw.rb_count[eu] = (int16_t)count;
    }
}

__device__ __forceinline__ void action_to_rb_warp(const StepArgs& A, int e)
{
    const DevState& st = A.st;
    const int lane = lane_id();
    const size_t eu = (size_t)e * U + (lane < U ? lane : 0);
    int start, count;
    action_to_rb_core(A, e, st.bank + st.scen[e], st.ue_d[eu * UD_N + UD_OCC], st.ue_d[eu * UD_N + UD_SEBAR],
                      st.ue_i[eu * UI_N + UI_SUM_SENT], st.win_len[e], st.step[e], st.episode[e], start, count);
    if (lane < U) st.rb[eu] = make_short2((short)start, (short)count);
}

__global__ void __launch_bounds__(256) k_action(const __grid_constant__ StepArgs A)
{
    const int e = A.env_begin + (int)((blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5);
    if (e >= A.env_end) return;
    action_to_rb_warp(A, e);
}
