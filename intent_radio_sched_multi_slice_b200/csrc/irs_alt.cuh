// irs_alt.cuh - the baseline agents' surfaces: SchedTWC (agents/sched_twc.py:166-413) and SchedColORAN
// (agents/sched_colran.py:166-417), cfg.surface 1 / 2.  One warp per env, lane s < 5 = slice s, which walks its
// (at most 5) UEs serially in ascending order - NumPy's order for these short means.
//
// Both agents return the same 50-value observation (sched_twc.py:189-345): per slice the three declared
// requirement values, then for each of seven metrics one value per slice - mean SE, mean served throughput (Mbps),
// mean effective throughput (Mbps), mean buffer occupancy, mean buffer latency, mean packet-loss rate
// (get_metric_value "reliability", common.py:32-53) and the requested throughput.  Rewards:
//   TWC     weighted mean of the negative slice intent drifts, weight 2 for prioritary slices (:347-413);
//   ColORAN sum over slices of  +throughput/200 Mbps (eMBB)  -buffered Mbps/2000 (URLLC)  (sched_colran.py:347-417).
// They push every frame twice into the 10-entry window (sched_twc.py:176-179): the state keeps 5 distinct frames
// and A.win_mul = 2 doubles the window sums (see intent_drift_ue).
// Written into out_obs0[E][50] and out_reward[E][0]; runs after the tail of a step (phase 0), after the rollover of
// finished envs (phase 1: observation only, for the envs that were reset) and after irs_reset (phase 2).
#pragma once
#include "irs_env.cuh"

// np.sum of n <= 15 doubles (numpy pairwise_sum: serial below 8 elements, else 8 accumulators + serial remainder)
__device__ __forceinline__ double np_sum_small(const double* a, int n)
{
    double res;
    if (n < 8) {
        res = 0.;
        for (int i = 0; i < n; ++i) res += a[i];
    } else {
        res = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
        for (int i = 8; i < n; ++i) res += a[i];
    }
    return 0. + res;
}

__global__ void __launch_bounds__(ENV_WARPS * 32) k_surface_alt(const __grid_constant__ StepArgs A, int phase)
{
    const DevState& st = A.st;
    const irs_config& c = A.cfg;
    const int lane = lane_id();
    const int e = A.env_begin + blockIdx.x * ENV_WARPS + (threadIdx.x >> 5);
    if (e >= A.env_end) return;
    if (phase == 1 && !st.out_done[e]) return;
    if (phase == 2 && A.reset_mask && !A.reset_mask[e]) return;
    const irs_scenario* sc = st.bank + st.scen[e];
    const int win_len = st.win_len[e];
    const int s = lane < S ? lane : 0;
    const irs_slice* sl = &sc->sl[s];
    const int fl = sl->flags;
    const double pkt = sl->message_size, buf = sl->buffer_size;
    int n = 0;
    double sum_se = 0., sum_thr_mbps = 0., sum_eff = 0., sum_occ = 0., sum_lat = 0., sum_loss = 0., sum_thr = 0.;
    double sum_d[3] = {0., 0., 0.};
    if (lane < S) {
        for (int u = 0; u < U; ++u) {
            if (sc->ue_slice[u] != s) continue;
            const size_t eu = (size_t)e * U + u;
            const int* ui = st.ue_i + eu * UI_N;
            const double* ud = st.ue_d + eu * UD_N;
            UeFrame f;
            f.occ = ud[UD_OCC]; f.lat = ud[UD_LAT]; f.sebar = ud[UD_SEBAR];
            f.sent = ui[UI_SENT]; f.sum_sent = ui[UI_SUM_SENT]; f.sum_drop = ui[UI_SUM_DROP];
            f.rb_count = 0; f.occ1z = ui[UI_OCC1Z] != 0;
            const double thr = ud[UD_THR];
            ++n;
            sum_se += f.sebar;
            sum_thr += thr;
            sum_thr_mbps += DIVC(thr * pkt, 1e6);                               // pkts_to_mbps, sched_twc.py:430-431
            sum_eff += DIVC((double)((long long)f.sent * (long long)pkt), 1e6);
            sum_occ += f.occ;
            sum_lat += f.lat;
            const double dw = (double)(f.sum_drop * A.win_mul), tw = (double)(f.sum_sent * A.win_mul);
            const double den = (f.occ * buf + dw) + tw;                     // common.py:40-47
            sum_loss += den != 0 ? dw / den : 0.0;                          // :48-53
            if (fl & IRS_F_REQ) {
                double drift[3];
                intent_drift_ue(sl, fl, f, win_len, A.win_mul, drift);
                sum_d[0] += drift[0]; sum_d[1] += drift[1]; sum_d[2] += drift[2];
            }
        }
        double* o = st.out_obs0 + (size_t)e * IRS_OBS0;
        const bool has = n != 0;
        const double dn = (double)n;
        o[3 * s + 0] = (has && (fl & IRS_F_REL)) ? sl->rel_value : 0.0;  // requirements: reliability, latency, throughput
        o[3 * s + 1] = (has && (fl & IRS_F_LAT)) ? sl->lat_value : 0.0;
        o[3 * s + 2] = (has && (fl & IRS_F_THR)) ? sl->thr_value : 0.0;
        o[15 + s] = has ? (0. + sum_se) / dn : 0.0;
        o[20 + s] = has ? (0. + sum_thr_mbps) / dn : 0.0;
        o[25 + s] = has ? (0. + sum_eff) / dn : 0.0;
        o[30 + s] = has ? (0. + sum_occ) / dn : 0.0;
        o[35 + s] = has ? (0. + sum_lat) / dn : 0.0;
        o[40 + s] = has ? (0. + sum_loss) / dn : 0.0;
        o[45 + s] = (fl & IRS_F_ACTIVE) ? sl->traffic : 0.0;
    }
    if (phase != 0) return;

    // ---- rewards (lane 0 gathers the slices in ascending order) ------------------------------------
    double v[3];
    bool neg[3];
    const int fbit[3] = {IRS_F_THR, IRS_F_REL, IRS_F_LAT};
#pragma unroll
    for (int k = 0; k < 3; ++k) { // calculate_slice_ue_obs (common.py:343-378): mean drift of declared intents, else -2
        const double d = (lane < S && n > 0 && (fl & IRS_F_REQ) && (fl & fbit[k])) ? (0. + sum_d[k]) / (double)n : -2.0;
        const bool valid = !(fabs(d - (-2.0)) <= 1e-8 + 1e-5 * 2.0);
        v[k] = d > 0 ? 0.0 : d;          // sched_twc.py:392-394: positive drifts do not count
        neg[k] = valid && v[k] < 0;
    }
    const double wgt = (sl->priority != 0) ? 2.0 : 1.0; // :382-390
    const double e_s = ((((0. + sum_thr) / (double)n) * pkt) / 1e6) / 200.0;             // sched_colran.py:381-392
    const double u_s = (((((0. + sum_occ) / (double)n) * buf) * pkt) / 1e6) / 2000.0;    // :393-413
    double terms[15], wsum = 0;
    int m = 0;
    double r_col = 0;
#pragma unroll
    for (int ss = 0; ss < S; ++ss) {
        const int n_s = __shfl_sync(FULL, n, ss), fl_s = __shfl_sync(FULL, fl, ss);
        const double w_s = __shfl_sync(FULL, wgt, ss);
        const double es = __shfl_sync(FULL, e_s, ss), us = __shfl_sync(FULL, u_s, ss);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double vk = __shfl_sync(FULL, v[k], ss);
            const int nk = __shfl_sync(FULL, (int)neg[k], ss);
            if (n_s > 0 && nk) { terms[m] = vk * w_s; wsum += w_s; ++m; }
        }
        if ((fl_s & IRS_F_ACTIVE) && n_s > 0) {
            if (fl_s & IRS_F_EMBB) r_col += es;
            if (fl_s & IRS_F_URLLC) r_col -= us;
        }
    }
    if (lane == 0) {
        double r;
        if (c.surface == 1) {
            for (int i = 0; i < m; ++i) terms[i] = terms[i] / wsum;
            r = !isclose0(wsum) ? np_sum_small(terms, m) : 0.0; // sched_twc.py:398-411
        } else {
            r = r_col;
        }
        double* acc = st.acc + (size_t)e * IRS_NUM_METRICS;
        acc[0] += r - st.out_reward[(size_t)e * 6]; // the episode sum of player_0's reward follows the surface
        st.out_reward[(size_t)e * 6] = r;
    }
}
