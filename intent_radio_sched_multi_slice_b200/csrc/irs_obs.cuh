// irs_obs.cuh - K4: observation, action masks, reward, episode metrics and episode rollover,
// one warp per env: lane u = UE u, and lanes 0..4 double as slice 0..4.
//
// Replaces
//   intent_drift_calc + get_metric_value     agents/common.py:9-340
//   calculate_slice_ue_obs                   agents/common.py:343-378
//   IBSched.obs_space_format / sort_slices   agents/ib_sched.py:63-204, 351-370
//   IBSched.calculate_reward / unsort_slices agents/ib_sched.py:206-221, 372-392
//   calculate_reward_no_mask                 agents/common.py:381-439
//   Metrics.step (as running sums)           sixg_radio_mgmt/sixg_radio_mgmt/metrics.py:62-71
//   CommunicationEnv.reset                   sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:375-482
#pragma once
#include "irs_common.cuh"
#include "irs_rng.cuh"
#include "irs_ue.cuh"

struct UeFrame { // newest window frame of this lane's UE (+ window sums)
    double occ, lat, sebar;
    int sent, sum_sent, sum_drop, rb_count;
    bool occ1z;
};
struct SliceView { // valid on lanes 0..4 (slice = lane)
    double dsl[3], act[3];
    double prio, msg;
    long long snt;
    int n, pos;
    bool active;
};

// common.py:9-65 + 68-340 for one UE; r = intent_overfulfillment_rate 0.2 (ib_sched.py:53)
// `mul` = 2 when the agent pushes every frame twice into the window (win[1] is then the newest frame again and
// every window sum counts each kept frame twice), else 1.
// Each metric ends in ONE division: the reference's two alternatives (intent met: distance over the over-fulfilment
// band; intent violated: distance over the requirement) are a choice of numerator and denominator, so neighbouring
// lanes that take different sides still share the division (float64 division is ~35 instructions and these were
// two thirds of the per-UE cost of the tail kernels).  Same operands, same operation: bit-identical results.
// reliability target t = (100 - v) / 100 of a slice (common.py:188-233): a per-slice constant that a caller looping over
// UEs computes once per slice
__device__ __forceinline__ double rel_target(const irs_slice* sl) { return DIVC(100 - __ldg(&sl->rel_value), 100); }
// The divisors that are constants of a slice - v r | v (throughput), t r (reliability met), v r (latency met) - as
// reciprocals: a caller that walks several UEs of ONE slice (the eight-lanes-per-env tail) divides once per divisor and
// takes the exact sequence of div_by per UE.  The violated branches of reliability and latency (rare) keep the division.
struct DriftRecips { double thr_ok, thr_bad, rel_ok, lat_ok; }; // RN(1 / divisor); the divisor itself is recomputed per UE (one multiply)
__device__ __forceinline__ DriftRecips drift_recips(const irs_slice* sl, int flags, double rel_t)
{
    const double r = 0.2;
    DriftRecips d;
    d.thr_ok = d.thr_bad = d.rel_ok = d.lat_ok = 1.0;
    if (flags & IRS_F_THR) { const double v = __ldg(&sl->thr_value); d.thr_ok = 1.0 / (v * r); d.thr_bad = 1.0 / v; }
    if (flags & IRS_F_REL) d.rel_ok = 1.0 / (rel_t * r);
    if (flags & IRS_F_LAT) d.lat_ok = 1.0 / (__ldg(&sl->lat_value) * r);
    return d;
}
__device__ __forceinline__ void intent_drift_ue(const irs_slice* sl, int flags, const UeFrame& f, int win_len, int mul,
                                                double (&drift)[3], const double* rel_t = nullptr,
                                                const DriftRecips* dr = nullptr)
{
    const double r = 0.2;
    drift[0] = drift[1] = drift[2] = 0;
    if (flags & IRS_F_THR) {
        const double v = __ldg(&sl->thr_value);
        double m = DIVC((double)((long long)f.sent * (long long)__ldg(&sl->message_size)), 1e6); // :26-31
        bool zero = isclose0(f.occ);                                                        // :104-107
        if (win_len * mul > 1) zero = zero || (mul == 2 ? isclose0(f.occ) : f.occ1z);          // :108-116
        if (zero) m = v * (1.1 + r);                                                        // :117-119
        const bool ok = m >= v;
        const double num = ok ? m - v : v - m; // :154-162 (m - v) / (v r) | :172-181 (v - m) / v
        const double den = ok ? v * r : v;
        const double q = dr ? div_const(num, den, ok ? dr->thr_ok : dr->thr_bad) : num / den;
        drift[0] = ok ? (m > v * (1 + r) ? 1.0 : 0.0 + q) : 0.0 - q; // :164-168 capped at 1 (0 +- q: the reference's own signed zeros)
    }
    if (flags & IRS_F_REL) {
        const double v = __ldg(&sl->rel_value);
        const double den = (f.occ * __ldg(&sl->buffer_size) + (double)(f.sum_drop * mul)) + (double)(f.sum_sent * mul); // :40-47
        const double m = den != 0 ? (double)(f.sum_drop * mul) / den : 0.0;                                      // :48-53
        const double t = rel_t ? *rel_t : rel_target(sl);
        const bool ok = 100 * (1 - m) >= v;                        // :124-126
        double q; // :188-220 (t - m) / (t r) | :223-233 (m - t) / (v / 100)
        if (dr && ok) q = div_const(t - m, t * r, dr->rel_ok);
        else {
            double den2 = t * r;
            if (!ok) den2 = DIVC(v, 100); // violated (rare): the other denominator
            q = (ok ? t - m : m - t) / den2;
        }
        drift[1] = ok ? (m < t * (1 - r) ? 1.0 : 0.0 + q) : 0.0 - q;
    }
    if (flags & IRS_F_LAT) {
        const double v = __ldg(&sl->lat_value);
        const double m = f.lat;
        const bool ok = m <= v;
        double q; // :290-319 (v - m) / (v r) | :322-335 (m - v) / (buffer_latency - v)
        if (dr && ok) q = div_const(v - m, v * r, dr->lat_ok);
        else q = (ok ? v - m : m - v) / (ok ? v * r : __ldg(&sl->buffer_latency) - v);
        drift[2] = ok ? (m < v * (1 - r) ? 1.0 : 0.0 + q) : 0.0 - q;
    }
}

// ib_sched.py:63-204.  Writes obs0/obs_i/masks/sorted for env e; returns the slice view the
// reward needs.  `sc` is the scenario the env is playing; all 32 lanes must call.
__device__ __noinline__ SliceView format_obs_warp(const StepArgs& A, int e, const irs_scenario* sc,
                                                     const UeFrame& f, int win_len)
{
    const DevState& st = A.st;
    const int lane = lane_id();
    const int s_mine = lane < U ? (int)sc->ue_slice[lane] : -1;
    uint32_t smask[S], gmask = 0;
#pragma unroll
    for (int s = 0; s < S; ++s) {
        smask[s] = __ballot_sync(FULL, s_mine == s);
        if (s_mine == s) gmask = smask[s];
    }
    const bool member = s_mine >= 0;
    const GroupIdx G = group_index(gmask);
    const int n_mine = G.n;
    const irs_slice* sl_u = &sc->sl[member ? s_mine : 0];
    const int fl_u = member ? __ldg(&sl_u->flags) : 0;

    // per-UE drift, then the slice means (common.py:371-373 np.mean over the slice's UEs)
    double drift[3] = {0, 0, 0};
    if (member && (fl_u & IRS_F_REQ)) intent_drift_ue(sl_u, fl_u, f, win_len, A.win_mul, drift);
    PHASE_STAMP(3, 1);
    double dsl_u[3];
    const int fbit[3] = {IRS_F_THR, IRS_F_REL, IRS_F_LAT};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double mean = div_small(group_sum_ordered(drift[k], G), n_mine);
        dsl_u[k] = ((fl_u & IRS_F_REQ) && (fl_u & fbit[k])) ? mean : -2.0;
    }
    const double se_slice_u = div_small(group_sum_ordered(f.sebar, G), n_mine); // ib_sched.py:146-157
    const int rbs_alloc_u = group_sum_int(f.rb_count, G);                      // :176-181
    const int snt_u = group_sum_int(f.sent, G);

    PHASE_STAMP(3, 2);
    // ---- slice lanes: fetch the slice-level values from the slice's first UE lane -------
    const int sl_id = lane < S ? lane : 0;
    const uint32_t my_smask = lane < S ? smask[0] * (lane == 0) + smask[1] * (lane == 1) + smask[2] * (lane == 2) +
                                             smask[3] * (lane == 3) + smask[4] * (lane == 4)
                                       : 0u;
    const int src = my_smask ? (__ffs(my_smask) - 1) : lane;
    SliceView v;
    v.n = __popc(my_smask);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double d = __shfl_sync(FULL, dsl_u[k], src);
        v.dsl[k] = v.n > 0 ? d : -2.0;
    }
    double se_slice = __shfl_sync(FULL, se_slice_u, src);
    int rbs_alloc = __shfl_sync(FULL, rbs_alloc_u, src);
    int snt = __shfl_sync(FULL, snt_u, src);
    if (v.n == 0) { se_slice = 0; rbs_alloc = 0; snt = 0; }
    const irs_slice* sl = &sc->sl[sl_id];
    const int fl = __ldg(&sl->flags);
    v.active = (fl & IRS_F_ACTIVE) != 0;
    v.msg = __ldg(&sl->message_size);
    v.snt = snt;
#pragma unroll
    for (int k = 0; k < 3; ++k) { // ib_sched.py:142-145  (np.isclose(x, -2): atol 1e-8 + rtol 1e-5 * 2)
        const bool close = fabs(v.dsl[k] - (-2.0)) <= 1e-8 + 1e-5 * 2.0;
        v.act[k] = close ? 0 : 1;
        if (close) v.dsl[k] = 0;
    }
    const double traffic = v.active ? __ldg(&sl->traffic) : 0; // :125-134
    v.prio = v.n != 0 ? __ldg(&sl->priority) : 0;              // :135-141

    // ---- sort_slices (ib_sched.py:351-370): stable ascending argsort of n_ues * traffic --
    int pos = 0;
    if (A.cfg.sort_slices) {
        const double key = (double)__ldg(&sl->n_ues) * ((fl & IRS_F_REQ) ? __ldg(&sl->traffic) : 0.0);
#pragma unroll
        for (int j = 0; j < S; ++j) {
            const double kj = __shfl_sync(FULL, key, j);
            if (kj < key || (kj == key && j < lane)) ++pos;
        }
        if (lane < S) st.sorted[(size_t)e * S + pos] = (int8_t)lane;
    } else {
#pragma unroll
        for (int j = 0; j < S; ++j)
            if (st.sorted[(size_t)e * S + j] == lane) pos = j;
    }
    v.pos = pos;
    PHASE_STAMP(3, 3);

    if (lane < S) {
        double* p0 = st.out_obs0 + (size_t)e * IRS_OBS0 + pos * 10; // :160-173
        p0[0] = v.dsl[0]; p0[1] = v.dsl[1]; p0[2] = v.dsl[2];
        p0[3] = v.act[0]; p0[4] = v.act[1]; p0[5] = v.act[2];
        p0[6] = v.prio; p0[7] = DIVC(traffic, 120.0); p0[8] = DIVC((double)v.n, 5.0); p0[9] = DIVC(se_slice, 40.0);
        double* pi = st.out_obsi + ((size_t)e * S + lane) * IRS_OBSI; // :186-200
        pi[0] = v.dsl[0]; pi[1] = v.dsl[1]; pi[2] = v.dsl[2];
        pi[3] = v.act[0]; pi[4] = v.act[1]; pi[5] = v.act[2];
        pi[6] = DIVC((double)rbs_alloc, R); pi[7] = DIVC(traffic, 120.0); pi[8] = DIVC((double)v.n, 5.0);
        for (int j = v.n; j < NS; ++j) { pi[9 + j] = 0; pi[14 + j] = 0; }
        st.out_mask0[(size_t)e * S + lane] = v.active ? 1 : 0; // :76-81 (unsorted order)
        for (int j = 0; j < NS; ++j) st.out_maski[((size_t)e * S + lane) * NS + j] = j < v.n ? 1 : 0;
        if (A.raw.sorted_slices) A.raw.sorted_slices[(size_t)e * S + pos] = (int8_t)lane;
    }
    if (member) { // per-UE tail of player_{s+1}'s observation
        const int j = __popc(gmask & ((1u << lane) - 1u));
        double* pi = st.out_obsi + ((size_t)e * S + s_mine) * IRS_OBSI;
        pi[9 + j] = f.occ;
        pi[14 + j] = DIVC(f.sebar, 40.0);
    }
    return v;
}
