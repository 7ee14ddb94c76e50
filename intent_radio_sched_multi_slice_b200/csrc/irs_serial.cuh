// irs_serial.cuh - thread-per-env variants of the per-env stages for LARGE batches.
//
// k_action / k_env_tail give one env to a warp (lane = UE): fine for a few thousand envs, but most
// of their ~2.4 k warp-instructions per env are shuffle plumbing and work replicated over lanes.
// With tens of thousands of envs there is enough parallelism to give every env its own THREAD:
// the code becomes the plain serial algorithm (NumPy's serial sums are just loops), a warp
// advances 32 envs at once, and the issue slots per env drop by ~10x.  Same reference symbols,
// same arithmetic order, same state layout as irs_action.cuh.  Round 1's default for launches of
// >= 49,152 envs; since round 2 the eight-lanes-per-env kernels (irs_lane8.cuh) are faster at every
// size (0.078 + 0.101 ms against 0.099 + 0.192 at 65,536 envs) and these are selected with
// IRS_PER_ENV=thread (every parity test runs them too).
#pragma once
#include "irs_action.cuh"

#define SERIAL_THREADS 64
#ifndef TAIL_UNROLL
#define TAIL_UNROLL 5 /* UE iterations whose record loads are in flight together */
#endif
#define IRS_STR(x) #x
#define IRS_PRAGMA(x) _Pragma(IRS_STR(x))
#define IRS_UNROLL(n) IRS_PRAGMA(unroll n)
#ifndef SERIAL_PREFETCH
#define SERIAL_PREFETCH 1
#endif

// The env's per-UE records (ue_i 800 B, ue_d 800 B, rb 100 B, contiguous per env) pulled into L2 ahead of
// the serial loop: a thread's loads are uncoalesced (one env per lane), so each would otherwise pay a full
// HBM round trip with only a handful of warps per SM to hide it.
__device__ __forceinline__ void prefetch_env_records(const DevState& st, size_t eu0)
{
#if SERIAL_PREFETCH
    const char* a = reinterpret_cast<const char*>(st.ue_i + eu0 * UI_N);
    const char* b = reinterpret_cast<const char*>(st.ue_d + eu0 * UD_N);
    const char* c = reinterpret_cast<const char*>(st.rb + eu0);
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a + 128 * k));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(b + 128 * k));
    }
    asm volatile("prefetch.global.L2 [%0];" ::"l"(c));
#endif
}

struct SliceUes { // UE indices of one slice, ascending (<= 5: agents/ib_sched.py:50)
    int u[NS];
    int n;
};

__device__ __forceinline__ int ue_slice_of(const uint32_t (&w)[8], int u)
{
    return (int)(int8_t)((w[u >> 2] >> (8 * (u & 3))) & 0xffu);
}

// agents/common.py:508-555 share of the i-th eligible UE
__device__ __forceinline__ int rr_share_i(int rbs_s, int m, int i)
{
    const int q = idiv_small(rbs_s, m);
    return q + (i < rbs_s - q * m ? 1 : 0);
}

__global__ void __launch_bounds__(SERIAL_THREADS) k_action_t(const __grid_constant__ StepArgs A)
{
    const DevState& st = A.st;
    const irs_config& cfg = A.cfg;
    const int e = A.env_begin + blockIdx.x * SERIAL_THREADS + threadIdx.x;
    if (e >= A.env_end) return;
    const irs_scenario* sc = st.bank + st.scen[e];
    uint32_t us[8];
    {
        const uint4* p = reinterpret_cast<const uint4*>(sc->ue_slice);
        const uint4 a = __ldg(p), b = __ldg(p + 1);
        us[0] = a.x; us[1] = a.y; us[2] = a.z; us[3] = a.w; us[4] = b.x; us[5] = b.y; us[6] = b.z; us[7] = b.w;
    }
    const size_t eu0 = (size_t)e * U;
    prefetch_env_records(st, eu0);
    const int win_len = st.win_len[e];
    bool active[S];
    int nact = 0;
#pragma unroll
    for (int s = 0; s < S; ++s) { active[s] = (__ldg(&sc->sl[s].flags) & IRS_F_ACTIVE) != 0; nact += active[s]; }

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
        double occ_mbps[S], sent_mbps[S], w[S];
        double mx = 0;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            double so = 0., ss = 0.;
            int n = 0;
            for (int u = 0; u < U; ++u) {
                if (ue_slice_of(us, u) != s) continue;
                so += st.ue_d[(eu0 + u) * UD_N + UD_OCC];
                ss += div_small((double)st.ue_i[(eu0 + u) * UI_N + UI_SUM_SENT], win_len);
                ++n;
            }
            double om = div_small(0. + so, n), sm = div_small(0. + ss, n); // n == 0: 0/0 = nan like np.mean([])
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
        const int step = st.step[e], episode = st.episode[e];
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
#pragma unroll
    for (int s = 0; s < S; ++s) {
        if (A.raw.a0_used) A.raw.a0_used[(size_t)e * S + s] = a0_in[s];
        if (A.raw.intra_used) A.raw.intra_used[(size_t)e * S + s] = intra[s];
    }

    // ---- ib_sched.py:223-349: every UE starts with no RBs -----------------------------------
    for (int u = 0; u < U; ++u) {
        st.rb[eu0 + u] = make_short2(0, 0);
        if (A.raw.rb_start) A.raw.rb_start[eu0 + u] = 0;
        if (A.raw.rb_count) A.raw.rb_count[eu0 + u] = 0;
    }
    if (nact == 0) return;
    double a0[S];
    {
        int srt[S];
#pragma unroll
        for (int s = 0; s < S; ++s) srt[s] = st.sorted[(size_t)e * S + s];
#pragma unroll
        for (int s = 0; s < S; ++s) a0[s] = active[s] ? sel5(a0_in, srt[s]) : -1.0; // :247 (quirk Q3), :248-255
    }
    int rbs_per_slice[S];
    scores_to_rbs5(a0, R / IRS_RBS_PER_RBG, active, rbs_per_slice);
    int base = 0, total = 0, expect = 0;
#pragma unroll 1
    for (int s = 0; s < S; ++s) {
        const int rbs_s = sel5i(rbs_per_slice, s) * IRS_RBS_PER_RBG;
        SliceUes g;
        g.n = 0;
#pragma unroll
        for (int k = 0; k < NS; ++k) g.u[k] = 0;
        for (int u = 0; u < U; ++u)
            if (ue_slice_of(us, u) == s && g.n < NS) {
#pragma unroll
                for (int k = 0; k < NS; ++k)
                    if (k == g.n) g.u[k] = u;
                ++g.n;
            }
        if (g.n > 0) {
            expect += rbs_s;
            int algo = cfg.fixed_intra >= 0 ? cfg.fixed_intra : sel5i(intra, s);
            const int n = g.n;
            const double max_pkts = __ldg(&sc->sl[s].buffer_size), pkt_bits = __ldg(&sc->sl[s].message_size);
            double occ[NS], ta[NS];
            int cnt[NS];
            bool elig[NS];
            int m_el = 0;
#pragma unroll
            for (int i = 0; i < NS; ++i) {
                occ[i] = i < n ? st.ue_d[(eu0 + g.u[i]) * UD_N + UD_OCC] : 0.0;
                elig[i] = i < n && !isclose0(occ[i]);
                m_el += elig[i];
                cnt[i] = 0;
            }
            if (algo == 0) { // round robin over UEs with a non-empty buffer (all if none) - common.py:517-524
                const bool all = m_el == 0;
                const int m = all ? n : m_el;
                int k = 0;
#pragma unroll
                for (int i = 0; i < NS; ++i)
                    if (i < n && (all || elig[i])) { cnt[i] = rr_share_i(rbs_s, m, k); ++k; }
            } else if (algo == 1 || algo == 2) {
                // throughput each UE could use - common.py:567-583, 648-664
                double tamax = 0, stt = 0.;
#pragma unroll
                for (int i = 0; i < NS; ++i) {
                    ta[i] = 0;
                    if (i < n) {
                        const double sebar = st.ue_d[(eu0 + g.u[i]) * UD_N + UD_SEBAR];
                        const double ta_a = div_small(sebar * DIVC((double)rbs_s * cfg.bandwidth, R), n);
                        const double ta_b = (occ[i] * max_pkts) * pkt_bits;
                        ta[i] = ta_a < ta_b ? ta_a : ta_b;
                        tamax = (i == 0 || ta[i] > tamax) ? ta[i] : tamax;
                        stt += ta[i];
                    }
                }
                double x[S] = {0, 0, 0, 0, 0};
                bool ok;
                if (algo == 1) { // proportional fairness - common.py:558-636
                    double w[NS], sw = 0.;
#pragma unroll
                    for (int i = 0; i < NS; ++i) {
                        w[i] = 0;
                        if (i < n) {
                            double snt = div_small((double)st.ue_i[(eu0 + g.u[i]) * UI_N + UI_SUM_SENT], win_len) * pkt_bits;
                            if (isclose0(ta[i])) snt = 1;
                            w[i] = !isclose0(snt) ? ta[i] / snt : 2 * tamax;
                            sw += w[i];
                        }
                    }
                    sw = 0. + sw;
                    ok = sw != 0;
                    const Recip rsw = recip_of(sw);
#pragma unroll
                    for (int i = 0; i < NS; ++i)
                        if (i < n) x[i] = div_by((double)rbs_s * w[i], rsw);
                } else { // max throughput - common.py:639-701
                    stt = 0. + stt;
                    ok = stt != 0;
                    const Recip rst = recip_of(stt);
#pragma unroll
                    for (int i = 0; i < NS; ++i)
                        if (i < n) x[i] = div_by((double)rbs_s * ta[i], rst);
                }
                if (ok) {
                    round_int_equal_sum5(x, rbs_s, cnt);
                } else { // fallback: round robin without the buffer filter
#pragma unroll
                    for (int i = 0; i < NS; ++i)
                        if (i < n) cnt[i] = rr_share_i(rbs_s, n, i);
                }
            } else {
                atomicOr(&st.err[e], ERR_INTRA_ACTION); // ValueError, ib_sched.py:341-344
            }
            // common.py:464-478 distribute_rbs_ues: consecutive runs inside the slice's block
            int rb = base;
#pragma unroll
            for (int i = 0; i < NS; ++i) {
                if (i >= n) continue;
                const int c = cnt[i], start = c > 0 ? rb : 0;
                rb += c;
                total += c;
                const size_t eu = eu0 + g.u[i];
                st.rb[eu] = make_short2((short)start, (short)c);
                if (A.raw.rb_start) A.raw.rb_start[eu] = (int16_t)start;
                if (A.raw.rb_count) A.raw.rb_count[eu] = (int16_t)c;
            }
        }
        base += rbs_s;
    }
    if (total != expect) atomicOr(&st.err[e], ERR_RB_SUM); // ib_sched.py:345-347
}

// ---- thread-per-env tail: observation, masks, reward, episode metrics, done -------------------
// Same reference symbols as irs_obs.cuh / irs_env.cuh (agents/common.py:9-439, agents/ib_sched.py:63-221,
// 351-392, metrics.py:62-71, comm_env.py:339-373).  The rollover of finished envs needs a whole warp
// (buffer zeroing, the reset channel frame) and runs in k_rollover right after.
#include "irs_env.cuh"

#ifndef TAIL_T_MINB
#define TAIL_T_MINB 8 /* 128 registers: 16 warps per SM, one wave at 65,536 envs (measured fastest) */
#endif
__global__ void __launch_bounds__(SERIAL_THREADS, TAIL_T_MINB) k_env_tail_t(const __grid_constant__ StepArgs A)
{
    const DevState& st = A.st;
    const irs_config& c = A.cfg;
    const int e = A.env_begin + blockIdx.x * SERIAL_THREADS + threadIdx.x;
    if (e >= A.env_end) return;
    const irs_scenario* sc = st.bank + st.scen[e];
    uint32_t us[8];
    {
        const uint4* p = reinterpret_cast<const uint4*>(sc->ue_slice);
        const uint4 a = __ldg(p), b = __ldg(p + 1);
        us[0] = a.x; us[1] = a.y; us[2] = a.z; us[3] = a.w; us[4] = b.x; us[5] = b.y; us[6] = b.z; us[7] = b.w;
    }
    const size_t eu0 = (size_t)e * U;
    prefetch_env_records(st, eu0);
    const int win_len_old = st.win_len[e], head_old = st.win_head[e], step_old = st.step[e];
    const int win_len = win_len_old < A.win_cap ? win_len_old + 1 : A.win_cap;
    const int step = step_old + 1; // comm_env.py:339
    st.win_len[e] = win_len; st.win_head[e] = (head_old + 1) % A.win_cap; st.step[e] = step;

    // ---- one pass over the UEs (ascending, so every per-slice sum runs in NumPy's order): totals for the
    //      metrics, per-UE drift (common.py:68-340) accumulated into the UE's slice, per-UE observation tail.
    //      A single loop keeps the warp's 32 envs converged: ~2/3 of the lanes are busy in every iteration,
    //      whereas a loop per slice would run each body with ~1/6 of them.
    double sum_d[S][3], sum_se[S];
    long long snt[S];
    int n_s[S], rbs_alloc[S], flags[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        sum_d[s][0] = sum_d[s][1] = sum_d[s][2] = 0.; sum_se[s] = 0.;
        snt[s] = 0; n_s[s] = 0; rbs_alloc[s] = 0;
        flags[s] = __ldg(&sc->sl[s].flags);
    }
    double rel_t[S]; // the reliability target of each slice: one division per slice instead of per UE
#pragma unroll
    for (int s = 0; s < S; ++s) rel_t[s] = rel_target(&sc->sl[s]);
    long long t_sent = 0, t_drop = 0, t_inc = 0;
    IRS_UNROLL(TAIL_UNROLL)
    for (int u = 0; u < U; ++u) {
        const size_t eu = eu0 + u;
        const int4 lo = *reinterpret_cast<const int4*>(st.ue_i + eu * UI_N);     // tot, sum_sent, sum_drop, sent
        const int4 hi = *reinterpret_cast<const int4*>(st.ue_i + eu * UI_N + 4); // drop, inc, occ1z, path
        t_sent += lo.w; t_drop += hi.x; t_inc += hi.y;
        const int s = ue_slice_of(us, u);
        if (s < 0) continue;
        const double2 d0 = *reinterpret_cast<const double2*>(st.ue_d + eu * UD_N); // occ, lat
        const double sebar = st.ue_d[eu * UD_N + UD_SEBAR];
        const irs_slice* sl = &sc->sl[s];
        const int fl = sel5i(flags, s);
        UeFrame f;
        f.occ = d0.x; f.lat = d0.y; f.sebar = sebar; f.sent = lo.w; f.sum_sent = lo.y; f.sum_drop = lo.z;
        f.rb_count = st.rb[eu].y; f.occ1z = hi.z != 0;
        double drift[3] = {0, 0, 0};
        const double rt = sel5(rel_t, s);
        if (fl & IRS_F_REQ) intent_drift_ue(sl, fl, f, win_len, A.win_mul, drift, &rt);
        int j = 0;
#pragma unroll
        for (int ss = 0; ss < S; ++ss) {
            if (ss != s) continue;
            j = n_s[ss];
            sum_d[ss][0] += drift[0]; sum_d[ss][1] += drift[1]; sum_d[ss][2] += drift[2];
            sum_se[ss] += sebar;
            rbs_alloc[ss] += f.rb_count;
            snt[ss] += f.sent;
            n_s[ss] = j + 1;
        }
        if (j < NS) { // per-UE tail of player_{s+1}'s observation (ib_sched.py:186-200)
            double* pi = st.out_obsi + ((size_t)e * S + s) * IRS_OBSI;
            pi[9 + j] = f.occ;
            pi[14 + j] = DIVC(sebar, 40.0);
        }
    }
    double dsl[S][3], act[S][3], se_slice[S], prio[S], traffic[S], msg[S];
    const int fbit[3] = {IRS_F_THR, IRS_F_REL, IRS_F_LAT};
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const irs_slice* sl = &sc->sl[s];
        const int fl = flags[s], n = n_s[s];
#pragma unroll
        for (int k = 0; k < 3; ++k) { // common.py:371-373 np.mean over the slice's UEs; undeclared / empty -> -2
            const double mean = div_small(0. + sum_d[s][k], n);
            const double d = (n > 0 && (fl & IRS_F_REQ) && (fl & fbit[k])) ? mean : -2.0;
            const bool close = fabs(d - (-2.0)) <= 1e-8 + 1e-5 * 2.0; // ib_sched.py:142-145 np.isclose(x, -2)
            act[s][k] = close ? 0 : 1;
            dsl[s][k] = close ? 0 : d;
        }
        se_slice[s] = n > 0 ? div_small(0. + sum_se[s], n) : 0.0; // ib_sched.py:146-157
        traffic[s] = (fl & IRS_F_ACTIVE) ? __ldg(&sl->traffic) : 0; // :125-134
        prio[s] = n != 0 ? __ldg(&sl->priority) : 0;               // :135-141
        msg[s] = __ldg(&sl->message_size);
        double* pi = st.out_obsi + ((size_t)e * S + s) * IRS_OBSI;
        for (int j = n; j < NS; ++j) { pi[9 + j] = 0; pi[14 + j] = 0; }
    }

    // ---- sort_slices (ib_sched.py:351-370): stable ascending argsort of n_ues * traffic ---------
    int pos[S];
    if (c.sort_slices) {
        double key[S];
#pragma unroll
        for (int s = 0; s < S; ++s)
            key[s] = (double)__ldg(&sc->sl[s].n_ues) * ((flags[s] & IRS_F_REQ) ? __ldg(&sc->sl[s].traffic) : 0.0);
#pragma unroll
        for (int s = 0; s < S; ++s) {
            int p = 0;
#pragma unroll
            for (int j = 0; j < S; ++j)
                if (key[j] < key[s] || (key[j] == key[s] && j < s)) ++p;
            pos[s] = p;
            st.sorted[(size_t)e * S + p] = (int8_t)s;
        }
    } else {
#pragma unroll
        for (int s = 0; s < S; ++s) pos[s] = 0;
#pragma unroll
        for (int j = 0; j < S; ++j) {
            const int sj = st.sorted[(size_t)e * S + j];
#pragma unroll
            for (int s = 0; s < S; ++s)
                if (sj == s) pos[s] = j;
        }
    }
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const bool active = (flags[s] & IRS_F_ACTIVE) != 0;
        double* p0 = st.out_obs0 + (size_t)e * IRS_OBS0 + pos[s] * 10; // :160-173
        p0[0] = dsl[s][0]; p0[1] = dsl[s][1]; p0[2] = dsl[s][2];
        p0[3] = act[s][0]; p0[4] = act[s][1]; p0[5] = act[s][2];
        p0[6] = prio[s]; p0[7] = DIVC(traffic[s], 120.0); p0[8] = DIVC((double)n_s[s], 5.0); p0[9] = DIVC(se_slice[s], 40.0);
        double* pi = st.out_obsi + ((size_t)e * S + s) * IRS_OBSI; // :186-200
        pi[0] = dsl[s][0]; pi[1] = dsl[s][1]; pi[2] = dsl[s][2];
        pi[3] = act[s][0]; pi[4] = act[s][1]; pi[5] = act[s][2];
        pi[6] = DIVC((double)rbs_alloc[s], R); pi[7] = DIVC(traffic[s], 120.0); pi[8] = DIVC((double)n_s[s], 5.0);
        st.out_mask0[(size_t)e * S + s] = active ? 1 : 0; // :76-81 (unsorted order)
#pragma unroll
        for (int j = 0; j < NS; ++j) st.out_maski[((size_t)e * S + s) * NS + j] = j < n_s[s] ? 1 : 0;
        if (A.raw.sorted_slices) A.raw.sorted_slices[(size_t)e * S + pos[s]] = (int8_t)s;
    }

    // ---- calculate_reward (ib_sched.py:206-221, common.py:381-439) ----------------------------
    double Aa[S], Pp[S], r_intra[S], vmn[S];
    bool viol[S];
    int nneg = 0, npneg = 0;
#pragma unroll
    for (int s = 0; s < S; ++s) {
        Aa[s] = 0; Pp[s] = 0;
        if (flags[s] & IRS_F_ACTIVE) { // common.py:400-408; inactive metrics are already 0 (quirk Q4)
            Pp[s] = __ldg(&sc->sl[s].priority);
            double mn = dsl[s][0];
            mn = dsl[s][1] < mn ? dsl[s][1] : mn;
            mn = dsl[s][2] < mn ? dsl[s][2] : mn;
            Aa[s] = mn;
        }
        nneg += Aa[s] < 0;
        npneg += (Pp[s] * Aa[s]) < 0;
        bool any = false, vi = false;
        double mn = 0, vm = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (act[s][k] != 0) { mn = any ? (dsl[s][k] < mn ? dsl[s][k] : mn) : dsl[s][k]; any = true; } // common.py:428-437
            if (act[s][k] != 0 && dsl[s][k] < 0) { vm = vi ? (dsl[s][k] < vm ? dsl[s][k] : vm) : dsl[s][k]; vi = true; }
        }
        r_intra[s] = any ? mn : 0;
        viol[s] = vi; vmn[s] = vm;
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
    st.out_reward[(size_t)e * 6] = r0;
#pragma unroll
    for (int s = 0; s < S; ++s) st.out_reward[(size_t)e * 6 + 1 + s] = r_intra[s];

    // ---- episode metrics (metrics.py:62-71 as running sums; results/gen_results.py:791-809, 874-1022) ---
    double* acc = st.acc + (size_t)e * IRS_NUM_METRICS;
    double a9 = acc[9], a12 = acc[12], a13 = acc[13], add10 = 0, add11 = 0;
#pragma unroll
    for (int p = 0; p < S; ++p) { // sorted order, like the observation vector
#pragma unroll
        for (int s = 0; s < S; ++s) {
            if (pos[s] != p || !(flags[s] & IRS_F_ACTIVE)) continue;
            a9 += DIVC((double)snt[s] * msg[s], 1e6);
            if (viol[s]) {
                add10 += 1; a12 += vmn[s];
                if (__ldg(&sc->sl[s].priority) != 0) { add11 += 1; a13 += vmn[s]; }
            }
        }
    }
    const bool done = step == c.max_steps; // comm_env.py:370
    acc[0] += r0;
#pragma unroll
    for (int s = 0; s < S; ++s) acc[1 + s] += r_intra[s];
    acc[6] += (double)t_sent; acc[7] += (double)t_drop; acc[8] += (double)t_inc;
    acc[9] = a9; acc[10] += add10; acc[11] += add11; acc[12] = a12; acc[13] = a13;
    acc[14] += 1;
    if (done) acc[15] += 1;
    st.out_done[e] = done ? 1 : 0;
}

// Rollover of the envs that finished in this step (CommunicationEnv.reset, comm_env.py:375-482) by a whole warp per
// finished env.  A warp looks at ROLL_ENVS envs (in all but one step of an episode none of them is done, so the kernel
// is launched with 1/8 of the warps a warp-per-env grid would need) and resets the finished ones one after the other.
#define ROLL_ENVS 8
__global__ void __launch_bounds__(ENV_WARPS * 32) k_rollover(const __grid_constant__ StepArgs A)
{
    __shared__ double sm[ENV_WARPS][136];
    const int lane = lane_id();
    const int e0 = A.env_begin + (blockIdx.x * ENV_WARPS + (threadIdx.x >> 5)) * ROLL_ENVS;
    if (e0 >= A.env_end) return;
    const bool done = lane < ROLL_ENVS && e0 + lane < A.env_end && A.st.out_done[e0 + lane];
    uint32_t todo = __ballot_sync(FULL, done);
    while (todo) {
        const int e = e0 + __ffs(todo) - 1;
        todo &= todo - 1;
        reset_env_warp(A, e, -1, sm[threadIdx.x >> 5]);
        __syncwarp();
    }
}
