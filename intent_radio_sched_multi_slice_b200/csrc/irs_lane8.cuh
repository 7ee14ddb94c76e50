// irs_lane8.cuh - the per-env stages with EIGHT LANES PER ENV (lane g < 5 = slice g), four envs per warp.
//
// The per-env stages (action -> RB runs; observation / masks / reward / metrics) are per-SLICE work over at most five
// UEs each, joined by a little cross-slice arithmetic (27 RBGs over 5 scores, the slice sort, the inter-slice reward).
//   * one WARP per env (irs_action.cuh, irs_obs.cuh / irs_env.cuh: lane = UE) spends most of its ~2.4 k warp
//     instructions per env on shuffle plumbing and on work replicated over 32 lanes;
//   * one THREAD per env (irs_serial.cuh) executes the plain serial algorithm - 17-19 k dependent instructions per
//     thread - and a 65,536-env batch is a single wave of 14 warps per SM: latency-bound (issue-active 20-33 %), and
//     slower still for the chunks of irs_step_host, which cannot fill the GPU with one thread per env.
// Here a slice's lane walks ITS OWN UEs (ascending, so every per-slice sum runs in NumPy's serial order), the eight
// lanes of an env share loads of the same 800-byte record block, and the cross-slice steps are a handful of width-8
// shuffles.  The dependent chain per lane is ~1/5 of the serial one and there are 8x the threads to hide it.
// Same reference symbols, same arithmetic order, same state layout as irs_serial.cuh; the rollover of finished envs
// runs in k_rollover right after the tail.
#pragma once
#include "irs_serial.cuh"

#define L8_THREADS 128 /* 16 envs per block */
#ifndef L8_TAIL_MINB
#define L8_TAIL_MINB 7 /* 72 registers: measured fastest (6: 0.100 ms, 7: 0.097, 8: 0.100 at 65,536 envs) */
#endif
#ifndef L8_ACTION_MINB
#define L8_ACTION_MINB 10 /* 48 registers (8: 0.062 ms, 10: 0.058, 12: 0.062) */
#endif

#ifndef L8_PREFETCH
#define L8_PREFETCH 1 /* measured: k_env_tail_8 0.097 -> 0.093 ms; 1: the eight lanes request the env's record block (ue_i, ue_d: 2 x 800 B) into L2 at kernel entry */
#endif
__device__ __forceinline__ void l8_prefetch_records(const DevState& st, size_t eu0, int g)
{
#if L8_PREFETCH
    if (g < 7) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(st.ue_i + eu0 * UI_N) + 128 * g));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(st.ue_d + eu0 * UD_N) + 128 * g));
    }
#endif
}

// UE indices of slice s (ascending, <= 5), 5 bits each; n = how many.  The 25 slice ids are compared four at a time
// (per-byte SIMD compare), the matches gathered into a 25-bit mask and its lowest five bits listed.
__device__ __forceinline__ uint32_t slice_ue_list(const uint32_t (&us)[8], int s, int& n)
{
    const uint32_t pat = (uint32_t)(s & 0xff) * 0x01010101u;
    uint32_t bits = 0;
#pragma unroll
    for (int w = 0; w < 7; ++w) {
        const uint32_t m = __vcmpeq4(us[w], pat) & 0x01010101u;       // byte b of the word = UE 4w + b
        bits |= (((m * 0x00204081u) >> 21) & 0xfu) << (4 * w);        // one bit per byte, in order
    }
    bits &= 0x01ffffffu;
    uint32_t list = 0;
    n = 0;
#pragma unroll
    for (int i = 0; i < NS; ++i)
        if (bits) { list |= (uint32_t)(__ffs(bits) - 1) << (5 * i); bits &= bits - 1; ++n; }
    return list;
}
__device__ __forceinline__ int group_sum8(int v)
{
    v += __shfl_xor_sync(FULL, v, 1);
    v += __shfl_xor_sync(FULL, v, 2);
    v += __shfl_xor_sync(FULL, v, 4);
    return v;
}
__device__ __forceinline__ long long group_sum8(long long v)
{
    v += __shfl_xor_sync(FULL, v, 1);
    v += __shfl_xor_sync(FULL, v, 2);
    v += __shfl_xor_sync(FULL, v, 4);
    return v;
}

// ---- action -> RB runs (agents/ib_sched.py:223-349, agents/common.py:442-701) ----------------------------------
__global__ void __launch_bounds__(L8_THREADS, L8_ACTION_MINB) k_action_8(const __grid_constant__ StepArgs A)
{
    const DevState& st = A.st;
    const irs_config& cfg = A.cfg;
    const int tid = blockIdx.x * L8_THREADS + threadIdx.x;
    const int g = tid & 7;
    int e = A.env_begin + (tid >> 3);
    const bool live = e < A.env_end; // the last warp may hold groups past the end: they run along (shuffles) and store nothing
    if (!live) e = A.env_end - 1;
    l8_prefetch_records(st, (size_t)e * U, g);
    const irs_scenario* sc = st.bank + st.scen[e];
    uint32_t us[8];
    {
        const uint4* p = reinterpret_cast<const uint4*>(sc->ue_slice);
        const uint4 a = __ldg(p), b = __ldg(p + 1);
        us[0] = a.x; us[1] = a.y; us[2] = a.z; us[3] = a.w; us[4] = b.x; us[5] = b.y; us[6] = b.z; us[7] = b.w;
    }
    const size_t eu0 = (size_t)e * U;
    const int win_len = st.win_len[e];
    const bool mine = g < S;
    const int s = mine ? g : 0;
    bool active[S];
    int active_i[S];
    int nact = 0;
#pragma unroll
    for (int j = 0; j < S; ++j) {
        active[j] = (__ldg(&sc->sl[j].flags) & IRS_F_ACTIVE) != 0;
        active_i[j] = active[j];
        nact += active[j];
    }
    int n;
    const uint32_t list = slice_ue_list(us, mine ? g : -2, n);

    // ---- where the action comes from ---------------------------------------------------
    double a0_in[S];
    int intra[S] = {0, 0, 0, 0, 0};
    if (cfg.policy == 0) {
#pragma unroll
        for (int j = 0; j < S; ++j) a0_in[j] = A.a0[(size_t)e * S + j];
        if (A.intra) {
#pragma unroll
            for (int j = 0; j < S; ++j) intra[j] = A.intra[(size_t)e * S + j];
        }
    } else if (cfg.policy == 1) { // agents/marr.py:40-47
#pragma unroll
        for (int j = 0; j < S; ++j) a0_in[j] = __ldg(&sc->sl[j].n_ues) > 0 ? 1.0 : -1.0;
    } else if (cfg.policy == 2) { // agents/mapf.py:41-111: the slice's lane averages over its UEs, the five results are shared
        double so = 0., ss = 0.;
        for (int i = 0; i < n; ++i) {
            const int u = (list >> (5 * i)) & 31;
            so += st.ue_d[(eu0 + u) * UD_N + UD_OCC];
            ss += div_small((double)st.ue_i[(eu0 + u) * UI_N + UI_SUM_SENT], win_len);
        }
        const double om = div_small(0. + so, n), sm = div_small(0. + ss, n); // n == 0: 0/0 = nan like np.mean([])
        const double bsz = __ldg(&sc->sl[s].buffer_size), msg = __ldg(&sc->sl[s].message_size);
        const double occ_m = active[s] ? DIVC((om * bsz) * msg, 1e6) : 0.0;
        const double sent_m = active[s] ? DIVC(sm * msg, 1e6) : 0.0;
        double occ_mbps[S], sent_mbps[S], w[S];
        double mx = 0;
#pragma unroll
        for (int j = 0; j < S; ++j) {
            occ_mbps[j] = __shfl_sync(FULL, occ_m, j, 8);
            sent_mbps[j] = __shfl_sync(FULL, sent_m, j, 8);
            mx = (j == 0 || occ_mbps[j] > mx) ? occ_mbps[j] : mx;
        }
        double res = 0.;
#pragma unroll
        for (int j = 0; j < S; ++j) {
            w[j] = !isclose0(sent_mbps[j]) ? occ_mbps[j] / sent_mbps[j] : 2 * mx;
            if (!active[j]) w[j] = 0;
            res += w[j];
        }
        const double sw = 0. + res;
#pragma unroll
        for (int j = 0; j < S; ++j) a0_in[j] = (sw > 0 ? w[j] / sw : 2.0) - 1;
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
        for (int j = 0; j < S; ++j) {
            a0_in[j] = (double)w[j] * (1.0 / 2147483648.0) - 1.0;
            intra[j] = (int)(((uint64_t)w[5 + j] * 3u) >> 32);
        }
    }
    if (live && mine) {
        if (A.raw.a0_used) A.raw.a0_used[(size_t)e * S + s] = sel5(a0_in, s);
        if (A.raw.intra_used) A.raw.intra_used[(size_t)e * S + s] = sel5i(intra, s);
    }

    // ---- ib_sched.py:223-349: every UE starts with no RBs -----------------------------------
    if (live)
        for (int u = g; u < U; u += 8) {
            st.rb[eu0 + u] = make_short2(0, 0);
            if (A.raw.rb_start) A.raw.rb_start[eu0 + u] = 0;
            if (A.raw.rb_count) A.raw.rb_count[eu0 + u] = 0;
        }
    __syncwarp(); // the zeroes above are written by other lanes than the runs below
    int total = 0, expect = 0;
    // scores -> RBGs (common.py:442-505 scores_to_rbs + round_int_equal_sum), one score per lane: the serial sums are
    // evaluated by every lane in the reference's order, each of the two divisions per score by the score's own lane, the
    // five results shared by width-8 shuffles.  Outside the nact branch: the four envs of a warp shuffle together.
    int rbs_per_slice[S];
    {
        double a0[S];
        int srt[S];
#pragma unroll
        for (int j = 0; j < S; ++j) srt[j] = st.sorted[(size_t)e * S + j];
#pragma unroll
        for (int j = 0; j < S; ++j) a0[j] = active[j] ? sel5(a0_in, srt[j]) : -1.0; // :247 (quirk Q3), :248-255
        const int target = R / IRS_RBS_PER_RBG;
        double res = 0.;
#pragma unroll
        for (int j = 0; j < S; ++j) res += a0[j] + 1;
        const double sum = 0. + res;
        double x_mine;
        if (sum != 0) {
            x_mine = ((double)target * (sel5(a0, s) + 1)) / sum;
        } else {
            double sa = 0.;
#pragma unroll
            for (int j = 0; j < S; ++j) sa += active[j] ? 1.0 : 0.0;
            sa = 0. + sa;
            x_mine = ((double)target / sa) * (sel5i(active_i, s) ? 1.0 : 0.0);
        }
        double xs[S];
        double res2 = 0.;
        int cnt = 0;
#pragma unroll
        for (int j = 0; j < S; ++j) {
            xs[j] = __shfl_sync(FULL, x_mine, j, 8);
            if (xs[j] != 0) { res2 += xs[j]; ++cnt; }
        }
        const double sum2 = 0. + res2;
        int out_mine = (x_mine != 0) ? (int)floor(((double)target * x_mine) / sum2) : 0;
        int tot = 0, rank = 0; // rank: descending value, ties -> higher index first (stable argsort reversed)
#pragma unroll
        for (int j = 0; j < S; ++j) {
            tot += __shfl_sync(FULL, out_mine, j, 8);
            if (xs[j] != 0 && (xs[j] > x_mine || (xs[j] == x_mine && j > s))) ++rank;
        }
        const int adj = target - tot;
        if (x_mine != 0 && adj > rank) out_mine += idiv_small(adj - rank - 1, cnt) + 1;
#pragma unroll
        for (int j = 0; j < S; ++j) rbs_per_slice[j] = __shfl_sync(FULL, out_mine, j, 8);
    }
    if (nact != 0) {
        int base = 0;
#pragma unroll
        for (int j = 0; j < S; ++j)
            if (j < s) base += rbs_per_slice[j] * IRS_RBS_PER_RBG;
        const int rbs_s = sel5i(rbs_per_slice, s) * IRS_RBS_PER_RBG;
        if (mine && n > 0) {
            expect = rbs_s;
            const int algo = cfg.fixed_intra >= 0 ? cfg.fixed_intra : sel5i(intra, s);
            const double max_pkts = __ldg(&sc->sl[s].buffer_size), pkt_bits = __ldg(&sc->sl[s].message_size);
            double occ[NS], ta[NS];
            int cnt[NS];
            bool elig[NS];
            int m_el = 0;
#pragma unroll
            for (int i = 0; i < NS; ++i) {
                const int u = (list >> (5 * i)) & 31;
                occ[i] = i < n ? st.ue_d[(eu0 + u) * UD_N + UD_OCC] : 0.0;
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
                        const int u = (list >> (5 * i)) & 31;
                        const double sebar = st.ue_d[(eu0 + u) * UD_N + UD_SEBAR];
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
                            const int u = (list >> (5 * i)) & 31;
                            double snt = div_small((double)st.ue_i[(eu0 + u) * UI_N + UI_SUM_SENT], win_len) * pkt_bits;
                            if (isclose0(ta[i])) snt = 1;
                            w[i] = !isclose0(snt) ? ta[i] / snt : 2 * tamax;
                            sw += w[i];
                        }
                    }
                    sw = 0. + sw;
                    ok = sw != 0;
                    const Recip rsw = recip_of(sw); // one division for the slice's shares; sw = 0: shares not used
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
            } else if (live) {
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
                const size_t eu = eu0 + ((list >> (5 * i)) & 31);
                if (live) {
                    st.rb[eu] = make_short2((short)start, (short)c);
                    if (A.raw.rb_start) A.raw.rb_start[eu] = (int16_t)start;
                    if (A.raw.rb_count) A.raw.rb_count[eu] = (int16_t)c;
                }
            }
        }
    }
    total = group_sum8(total);
    expect = group_sum8(expect);
    if (live && g == 0 && total != expect) atomicOr(&st.err[e], ERR_RB_SUM); // ib_sched.py:345-347
}

// ---- observation, masks, reward, episode metrics, done (agents/common.py:9-439, agents/ib_sched.py:63-221, 351-392,
//      metrics.py:62-71, comm_env.py:339-373) ----------------------------------------------------------------------
__global__ void __launch_bounds__(L8_THREADS, L8_TAIL_MINB) k_env_tail_8(const __grid_constant__ StepArgs A)
{
    const DevState& st = A.st;
    const irs_config& c = A.cfg;
    const int tid = blockIdx.x * L8_THREADS + threadIdx.x;
    const int g = tid & 7;
    int e = A.env_begin + (tid >> 3);
    const bool live = e < A.env_end;
    if (!live) e = A.env_end - 1;
    l8_prefetch_records(st, (size_t)e * U, g);
    const irs_scenario* sc = st.bank + st.scen[e];
    uint32_t us[8];
    {
        const uint4* p = reinterpret_cast<const uint4*>(sc->ue_slice);
        const uint4 a = __ldg(p), b = __ldg(p + 1);
        us[0] = a.x; us[1] = a.y; us[2] = a.z; us[3] = a.w; us[4] = b.x; us[5] = b.y; us[6] = b.z; us[7] = b.w;
    }
    const size_t eu0 = (size_t)e * U;
    const int win_len_old = st.win_len[e], head_old = st.win_head[e], step_old = st.step[e];
    const int win_len = win_len_old < A.win_cap ? win_len_old + 1 : A.win_cap;
    const int step = step_old + 1; // comm_env.py:339
    __syncwarp();                  // every lane of the env has read the old values
    if (live && g == 0) { st.win_len[e] = win_len; st.win_head[e] = (head_old + 1) % A.win_cap; st.step[e] = step; }

    const bool mine = g < S;
    const int s = mine ? g : 0;
    const irs_slice* sl = &sc->sl[s];
    const int fl = mine ? __ldg(&sl->flags) : 0;
    int n;
    const uint32_t list = slice_ue_list(us, mine ? g : -2, n);

    // totals over ALL 25 UEs (also those without a slice): lane g takes UEs g, g + 8, g + 16, g + 24
    long long t_sent = 0, t_drop = 0, t_inc = 0;
    for (int u = g; u < U; u += 8) {
        const int* ui = st.ue_i + (eu0 + u) * UI_N;
        t_sent += ui[UI_SENT]; t_drop += ui[UI_DROP]; t_inc += ui[UI_INC];
    }
    t_sent = group_sum8(t_sent); t_drop = group_sum8(t_drop); t_inc = group_sum8(t_inc);

    // ---- my slice's UEs, ascending: drift (common.py:68-340), sums in NumPy's serial order, per-UE observation tail
    double sum_d[3] = {0., 0., 0.}, sum_se = 0.;
    long long snt = 0;
    int rbs_alloc = 0;
    double* pi = st.out_obsi + ((size_t)e * S + s) * IRS_OBSI;
    const double rel_t = (fl & IRS_F_REL) ? rel_target(sl) : 0.0;
    const DriftRecips dr = drift_recips(sl, (fl & IRS_F_REQ) ? fl : 0, rel_t); // the slice's divisors, once for its UEs
#pragma unroll
    for (int j = 0; j < NS; ++j) {
        if (j >= n) continue;
        const size_t eu = eu0 + ((list >> (5 * j)) & 31);
        const int4 lo = *reinterpret_cast<const int4*>(st.ue_i + eu * UI_N); // tot, sum_sent, sum_drop, sent
        const int occ1z = st.ue_i[eu * UI_N + UI_OCC1Z];
        const double2 d0 = *reinterpret_cast<const double2*>(st.ue_d + eu * UD_N); // occ, lat
        const double sebar = st.ue_d[eu * UD_N + UD_SEBAR];
        UeFrame f;
        f.occ = d0.x; f.lat = d0.y; f.sebar = sebar; f.sent = lo.w; f.sum_sent = lo.y; f.sum_drop = lo.z;
        f.rb_count = st.rb[eu].y; f.occ1z = occ1z != 0;
        double drift[3] = {0, 0, 0};
        if (fl & IRS_F_REQ) intent_drift_ue(sl, fl, f, win_len, A.win_mul, drift, &rel_t, &dr);
        sum_d[0] += drift[0]; sum_d[1] += drift[1]; sum_d[2] += drift[2];
        sum_se += sebar;
        rbs_alloc += f.rb_count;
        snt += f.sent;
        if (live) { // per-UE tail of player_{s+1}'s observation (ib_sched.py:186-200)
            pi[9 + j] = f.occ;
            pi[14 + j] = DIVC(sebar, 40.0);
        }
    }
    double dsl[3], act[3];
    const int fbit[3] = {IRS_F_THR, IRS_F_REL, IRS_F_LAT};
#pragma unroll
    for (int k = 0; k < 3; ++k) { // common.py:371-373 np.mean over the slice's UEs; undeclared / empty -> -2
        const double mean = div_small(0. + sum_d[k], n);
        const double d = (n > 0 && (fl & IRS_F_REQ) && (fl & fbit[k])) ? mean : -2.0;
        const bool close = fabs(d - (-2.0)) <= 1e-8 + 1e-5 * 2.0; // ib_sched.py:142-145 np.isclose(x, -2)
        act[k] = close ? 0 : 1;
        dsl[k] = close ? 0 : d;
    }
    const double se_slice = n > 0 ? div_small(0. + sum_se, n) : 0.0;        // ib_sched.py:146-157
    const double traffic = (fl & IRS_F_ACTIVE) ? __ldg(&sl->traffic) : 0;    // :125-134
    const double prio_raw = __ldg(&sl->priority);
    const double prio = n != 0 ? prio_raw : 0;                               // :135-141
    const double msg = __ldg(&sl->message_size);
    if (live && mine)
        for (int j = n; j < NS; ++j) { pi[9 + j] = 0; pi[14 + j] = 0; }

    // ---- sort_slices (ib_sched.py:351-370): stable ascending argsort of n_ues * traffic ---------
    int pos = 0;
    if (c.sort_slices) {
        const double key = mine ? (double)__ldg(&sl->n_ues) * ((fl & IRS_F_REQ) ? __ldg(&sl->traffic) : 0.0) : 0.0;
#pragma unroll
        for (int j = 0; j < S; ++j) {
            const double kj = __shfl_sync(FULL, key, j, 8);
            if (kj < key || (kj == key && j < s)) ++pos;
        }
        if (live && mine) st.sorted[(size_t)e * S + pos] = (int8_t)s;
    } else {
#pragma unroll
        for (int j = 0; j < S; ++j)
            if (st.sorted[(size_t)e * S + j] == s) pos = j;
    }
    if (live && mine) {
        const bool is_active = (fl & IRS_F_ACTIVE) != 0;
        double* p0 = st.out_obs0 + (size_t)e * IRS_OBS0 + pos * 10; // :160-173
        p0[0] = dsl[0]; p0[1] = dsl[1]; p0[2] = dsl[2];
        p0[3] = act[0]; p0[4] = act[1]; p0[5] = act[2];
        p0[6] = prio; p0[7] = DIVC(traffic, 120.0); p0[8] = DIVC((double)n, 5.0); p0[9] = DIVC(se_slice, 40.0);
        pi[0] = dsl[0]; pi[1] = dsl[1]; pi[2] = dsl[2]; // :186-200
        pi[3] = act[0]; pi[4] = act[1]; pi[5] = act[2];
        pi[6] = DIVC((double)rbs_alloc, R); pi[7] = DIVC(traffic, 120.0); pi[8] = DIVC((double)n, 5.0);
        st.out_mask0[(size_t)e * S + s] = is_active ? 1 : 0; // :76-81 (unsorted order)
#pragma unroll
        for (int j = 0; j < NS; ++j) st.out_maski[((size_t)e * S + s) * NS + j] = j < n ? 1 : 0;
        if (A.raw.sorted_slices) A.raw.sorted_slices[(size_t)e * S + pos] = (int8_t)s;
    }

    // ---- calculate_reward (ib_sched.py:206-221, common.py:381-439) ----------------------------
    double a_s = 0, p_s = 0;
    if (mine && (fl & IRS_F_ACTIVE)) { // common.py:400-408; inactive metrics are already 0 (quirk Q4)
        p_s = prio_raw;
        double mn = dsl[0];
        mn = dsl[1] < mn ? dsl[1] : mn;
        mn = dsl[2] < mn ? dsl[2] : mn;
        a_s = mn;
    }
    bool any = false, viol = false;
    double r_intra = 0, vmn = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (act[k] != 0) { r_intra = any ? (dsl[k] < r_intra ? dsl[k] : r_intra) : dsl[k]; any = true; } // common.py:428-437
        if (act[k] != 0 && dsl[k] < 0) { vmn = viol ? (dsl[k] < vmn ? dsl[k] : vmn) : dsl[k]; viol = true; }
    }
    if (!any) r_intra = 0;
    double Aa[S], Pp[S];
    int nneg = 0, npneg = 0;
#pragma unroll
    for (int j = 0; j < S; ++j) {
        Aa[j] = __shfl_sync(FULL, a_s, j, 8);
        Pp[j] = __shfl_sync(FULL, p_s, j, 8);
        nneg += Aa[j] < 0;
        npneg += (Pp[j] * Aa[j]) < 0;
    }
    double r0;
    {
        double res = 0.;
        int m = 0;
#pragma unroll
        for (int j = 0; j < S; ++j) {
            const bool sel = nneg == 0 ? true : (npneg ? (Aa[j] * Pp[j] < 0) : (Aa[j] < 0));
            if (sel) { res += Aa[j]; ++m; }
        }
        r0 = div_small(0. + res, m) - ((nneg != 0 && npneg) ? 1 : 0); // common.py:410-426
    }
    if (live && g == 0) st.out_reward[(size_t)e * 6] = r0;
    if (live && mine) st.out_reward[(size_t)e * 6 + 1 + s] = r_intra;

    // ---- episode metrics (metrics.py:62-71 as running sums; results/gen_results.py:791-809, 874-1022): lane 0 walks
    //      the slices in sorted order, like the observation vector -----------------------------------------------
    const double mbit = DIVC((double)snt * msg, 1e6);
    const bool act_s = mine && (fl & IRS_F_ACTIVE);
    double* acc = st.acc + (size_t)e * IRS_NUM_METRICS;
    double a9 = acc[9], a12 = acc[12], a13 = acc[13], add10 = 0, add11 = 0;
    double ri[S];
#pragma unroll
    for (int j = 0; j < S; ++j) ri[j] = __shfl_sync(FULL, r_intra, j, 8);
#pragma unroll
    for (int p = 0; p < S; ++p) {
        const uint32_t b = (__ballot_sync(FULL, mine && pos == p) >> (lane_id() & 24)) & 0xffu;
        const int src = b ? __ffs(b) - 1 : 0;
        const int a_j = __shfl_sync(FULL, (int)act_s, src, 8);
        const double mb_j = __shfl_sync(FULL, mbit, src, 8);
        const int vi_j = __shfl_sync(FULL, (int)viol, src, 8);
        const double vm_j = __shfl_sync(FULL, vmn, src, 8);
        const double pr_j = __shfl_sync(FULL, prio_raw, src, 8);
        if (!b || !a_j) continue;
        a9 += mb_j;
        if (vi_j) {
            add10 += 1; a12 += vm_j;
            if (pr_j != 0) { add11 += 1; a13 += vm_j; }
        }
    }
    const bool done = step == c.max_steps; // comm_env.py:370
    if (live && g == 0) {
        acc[0] += r0;
#pragma unroll
        for (int j = 0; j < S; ++j) acc[1 + j] += ri[j];
        acc[6] += (double)t_sent; acc[7] += (double)t_drop; acc[8] += (double)t_inc;
        acc[9] = a9; acc[10] += add10; acc[11] += add11; acc[12] = a12; acc[13] = a13;
        acc[14] += 1;
        if (done) acc[15] += 1;
        st.out_done[e] = done ? 1 : 0;
    }
}
