/*
 * irs_oracle.c - CPU restatement (plain C) of the reference hot path
 * `sixg_radio_mgmt.CommunicationEnv.step` + the IBSched/MARR/MAPF agent callbacks.
 *
 * THIS FILE IS TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may build, load or call it.  The
 * product (intent_radio_sched_multi_slice_b200/) never does.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle_golden.py)
 * against (a) the only known-answer vector the reference holds for this path,
 * `Buffer.main()` (sixg_radio_mgmt/sixg_radio_mgmt/buffer.py:133-149), and (b) traces
 * recorded from the unmodified reference run in the build container by
 * oracle/refharness/gen_golden.py (committed under tests/golden/).
 *
 * Every function cites the reference file:line (relative to /root/reference) it follows.
 * One env is stepped at a time; orc_step_batch loops (optionally with OpenMP) over envs.
 *
 * Build: gcc -O2 -fPIC -shared -ffp-contract=off -fopenmp -mfma -o _build/libirs_oracle.so irs_oracle.c -lm
 * -ffp-contract=off matters: NumPy evaluates a*b+c with two roundings.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define U 25   /* env_config/mult_slice.yml: max_number_ues */
#define S 5    /* max_number_slices */
#define R 135  /* num_available_rbs */
#define W 10   /* agents/ib_sched.py:49 max_obs_memory */
#define NS 5   /* agents/ib_sched.py:50 max_number_ues_slice */
#define RBG 5  /* agents/ib_sched.py:56 rbs_per_rbg */
#define POISSON_K 512

/* ---- scenario bank row (mirrors include/irs_b200.h; restated, not included) ---- */
enum { F_ACTIVE = 1, F_REQ = 2, F_THR = 4, F_REL = 8, F_LAT = 16, F_EMBB = 32, F_URLLC = 64 };
typedef struct {
    double priority, traffic, buffer_size, buffer_latency, message_size;
    double thr_value, rel_value, lat_value;
    int32_t flags, n_ues, lam_idx, pad;
} orc_slice;
typedef struct {
    orc_slice sl[S];
    int8_t ue_slice[32];
} orc_scenario;

typedef struct {
    int32_t num_envs, depth_cap;      /* buffer row stride (ints) */
    int32_t channel_mode;             /* 0 injected SE trace, 1 fixed_se (2.0), 2 native mimic */
    int32_t traffic_mode;             /* 0 injected bits trace, 1 native Poisson */
    int32_t sort_slices;              /* IBSched enable_sort_slices (ib_sched.py:57) */
    int32_t fixed_intra;              /* -1 per-slice action, 0 rr, 1 pf, 2 mt */
    int32_t policy;                   /* 0 external, 1 MARR, 2 MAPF, 3 random (native RNG) */
    int32_t assoc_mode;               /* 0 scen = episode % scen_mod, 1 scen = episode / scen_div,
                                         2 generated per (env, episode): bank rows are slice-type templates */
    int32_t scen_mod, scen_div, n_scen;
    int32_t max_steps, max_episodes, initial_episode;
    int32_t auto_reset, env_id_base;
    int32_t surface;                  /* 0 IBSched family, 1 SchedTWC, 2 SchedColORAN (double window push) */
    int32_t random_episodes;          /* comm_env.py:219-226, 434-444 enable_random_episodes */
    uint64_t seed;
    double bandwidth;
} orc_config;

typedef struct {
    double occ[U], lat[U], sebar[U];
    int32_t sent[U], dropped[U], rb_count[U];
    double thr[U]; /* pkt_throughputs (SchedTWC / SchedColORAN observations) */
} orc_frame;

typedef struct {
    int32_t step, episode, scen, win_len;
    int32_t visit;                    /* resets seen: part of the native generators' key */
    int32_t ep_pos, ep_gen, ep_base;  /* random_episodes: position, permutations drawn, first episode of the range */
    orc_frame win[W]; /* win[0] newest: collections.deque(maxlen=10).appendleft */
    int32_t sorted_slices[S];
    /* last formatted obs, kept for calculate_reward (ib_sched.py:202) */
    double obs0[S * 10], obs_i[S][19];
    double acc[16]; /* episode-metric accumulators */
} orc_env;

typedef struct {
    orc_config cfg;
    orc_env *envs;
    int32_t *buf; /* [E][U][depth_cap] age histograms (buffer.py:47) */
    const orc_scenario *bank;
    const double *poisson_cdf; /* [n_lambda][POISSON_K] */
    orc_scenario *own_bank;    /* assoc_mode 2: templates + two generated rows per env */
    const double *se_bank;     /* channel_mode 3: [bank_eps][bank_steps][U][R] (channels/quadriga.py) */
    int32_t bank_eps, bank_steps, bank_first, bank_wrap;
    const int32_t *ep_order;   /* random_episodes with a given order: [E][ep_order_n] */
    int32_t ep_order_n;
} orc_handle;

/* ------------------------------------------------------------------------------------
 * Counter-based RNG (native mode).  Not in the reference (it uses one sequential PCG64
 * stream per env, comm_env.py:631); this is the CPU restatement of the generators the
 * CUDA path defines in csrc/irs_rng.cuh so that native mode is parity-checkable too.
 * Philox4x32-7 (Salmon et al. 2011), key = env_key(seed, reset count),
 * counter=(q | kind<<16 | ue<<24, frame, episode, global env id).
 * ------------------------------------------------------------------------------------ */
static void philox4x32_7(uint32_t c[4], uint32_t k0, uint32_t k1)
{
    for (int r = 0; r < 7; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
enum { K_SE = 0, K_TRAFFIC = 1, K_MU = 2, K_POLICY = 3, K_SCEN = 4, K_PERM = 5 };
/* key = the seed with the env's reset count in the upper word (csrc/irs_common.cuh env_key): an episode number that
 * is played again draws fresh noise, like the reference's np_random stream that runs on across episodes */
static uint64_t env_key(const orc_config *cfg, int visit) { return cfg->seed + ((uint64_t)(uint32_t)visit << 32); }
static void rng_quad(uint64_t key, int kind, int ue, int q, int frame, int episode, int gid, uint32_t out[4])
{
    out[0] = (uint32_t)q | ((uint32_t)kind << 16) | ((uint32_t)ue << 24);
    out[1] = (uint32_t)frame; out[2] = (uint32_t)episode; out[3] = (uint32_t)gid;
    philox4x32_7(out, (uint32_t)key, (uint32_t)(key >> 32));
}
/* enable_random_episodes (comm_env.py:219-226, 434-444) without a table: element `pos` of a pseudo-random permutation
 * of [0, n) - 8-round balanced Feistel network over the next even power of two, cycle-walked into [0, n)
 * (csrc/irs_rng.cuh feistel_perm) */
static uint32_t mix32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
static uint32_t feistel_perm(uint32_t i, uint32_t n, const uint32_t k[4])
{
    if (n <= 1) return 0;
    int bits = 1;
    while ((1u << bits) < n) ++bits;
    bits += bits & 1;
    int half = bits >> 1;
    uint32_t mask = (1u << half) - 1, x = i;
    do {
        uint32_t l = x >> half, r = x & mask, t;
        for (int j = 0; j < 8; ++j) { t = l ^ (mix32(r ^ k[j & 3] ^ ((uint32_t)j * 0x9E3779B9u)) & mask); l = r; r = t; }
        x = (l << half) | r;
    } while (x >= n);
    return x;
}
static int shuffled_episode(const orc_config *cfg, int gid, int gen, int base, int n, int pos)
{
    uint32_t k[4];
    rng_quad(cfg->seed, K_PERM, 0, 0, gen, 0, gid, k);
    return base + (int)feistel_perm((uint32_t)pos, (uint32_t)n, k);
}
/* One standard normal per 16-bit half-word by table-driven inverse CDF (csrc/irs_rng.cuh normal_from; table and
 * its derivation: scripts/gen_normal_table.py -> csrc/irs_normal_tab.h, shared with the CUDA kernels) */
#include "../intent_radio_sched_multi_slice_b200/csrc/irs_normal_tab.h"
static const uint32_t normal_tab[2 * IRS_NORMAL_TAB_N] = { IRS_NORMAL_TAB_WORDS };
static float bits_to_float(uint32_t b) { float f; memcpy(&f, &b, 4); return f; }
static uint32_t float_to_bits(float f) { uint32_t b; memcpy(&b, &f, 4); return b; }
static float normal16(uint32_t h)
{
    uint32_t m = h & 0x7FFFu, signbit = (h & 0x8000u) << 16;
    uint32_t bits = float_to_bits((float)(int)(32768u - m));
    uint32_t idx = (bits >> (23 - IRS_NORMAL_SUB_BITS)) - (127u << IRS_NORMAL_SUB_BITS);
    float c0 = bits_to_float(normal_tab[2 * idx]), c1 = bits_to_float(normal_tab[2 * idx + 1]);
    float g = bits_to_float(((bits & ((1u << (23 - IRS_NORMAL_SUB_BITS)) - 1u)) << IRS_NORMAL_SUB_BITS) | 0x3F800000u);
    return bits_to_float(float_to_bits(fmaf(c1, g, c0)) ^ signbit);
}
static float normal_half(const uint32_t w[4], int k) { return normal16((w[k >> 1] >> (16 * (k & 1))) & 0xFFFFu); }
/* channels/mimic_quadriga.py:37-41: ues_mean_se = abs(N(10, 7.5)) per UE per episode */
static double native_mu(uint64_t key, int ue, int episode, int gid)
{
    uint32_t w[4];
    rng_quad(key, K_MU, ue, 0, 0, episode, gid, w);
    return fabs(fma(7.5, (double)normal_half(w, 0), 10.0));
}
/* where half-word k of SE quad q lands in the row: quad q < 16 holds the 8 elements that accumulator q & 7 of half
 * q >> 3 of NumPy's pairwise sum visits, in order; quad 16 is the tail 128..134 */
static int se_index(int q, int k)
{
    return q < 16 ? ((q >> 3) << 6) + (q & 7) + 8 * k : 128 + k;
}
/* channels/mimic_quadriga.py:49-56: SE[u,r] = abs(N(mu_u, 1.5)) */
static void native_se_row(uint64_t key, int ue, int frame, int episode, int gid, double *se)
{
    double mu = native_mu(key, ue, episode, gid);
    for (int q = 0; q <= 16; ++q) {
        uint32_t w[4];
        rng_quad(key, K_SE, ue, q, frame, episode, gid, w);
        for (int k = 0; k < 8; ++k)
            if (se_index(q, k) < R) se[se_index(q, k)] = fabs(fma(1.5, (double)normal_half(w, k), mu));
    }
}
/* traffics/mult_slice.py:28-31: Poisson(lambda) * 1e6 bits, by table inversion */
static double native_traffic(const orc_handle *h, uint64_t key, int ue, int lam_idx, int frame, int episode, int gid)
{
    uint32_t w[4];
    rng_quad(key, K_TRAFFIC, ue, 0, frame, episode, gid, w);
    double u = ((double)(w[0] >> 5) * 67108864.0 + (double)(w[1] >> 6)) * (1.0 / 9007199254740992.0);
    const double *cdf = h->poisson_cdf + (size_t)lam_idx * POISSON_K;
    int k = 0;
    for (int j = 0; j < POISSON_K; ++j) k += (cdf[j] <= u);
    return (double)k * 1e6;
}

/* ------------------------------------------------------------------------------------
 * NumPy float64 add.reduce order (numpy/_core/src/umath/loops_utils.h.src,
 * DOUBLE_pairwise_sum): n<8 serial from 0.; n<=128 eight interleaved accumulators, then
 * ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), then the n%8 tail serially; n>128 split at
 * n/2 rounded down to a multiple of 8.  The reduction then adds the result to the
 * identity 0.  ues.py:108-116 and ib_sched.py:111,147 depend on this order.
 * ------------------------------------------------------------------------------------ */
static double np_pairwise(const double *a, int n)
{
    if (n < 8) {
        double res = 0.;
        for (int i = 0; i < n; ++i) res += a[i];
        return res;
    } else if (n <= 128) {
        double r[8];
        int i;
        for (i = 0; i < 8; ++i) r[i] = a[i];
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    } else {
        int n2 = n / 2;
        n2 -= n2 % 8;
        return np_pairwise(a, n2) + np_pairwise(a + n2, n - n2);
    }
}
static double np_sum(const double *a, int n) { return 0. + np_pairwise(a, n); }
static int np_isclose0(double x) { return fabs(x) <= 1e-8; } /* np.isclose(x, 0): atol 1e-8 */

/* buffer.py:53-80 */
static int64_t buffer_receive(int32_t *buf, int depth, int64_t max_pkts, double arrived)
{
    int64_t dropped = buf[depth - 1];
    for (int i = depth - 1; i > 0; --i) buf[i] = buf[i - 1]; /* np.roll(buffer, 1) */
    buf[0] = 0;
    int64_t total = 0;
    for (int i = 0; i < depth; ++i) total += buf[i];
    if ((double)total + arrived <= (double)max_pkts) {
        buf[0] = (int32_t)arrived;
    } else {
        dropped += (int64_t)(arrived - (double)(max_pkts - total));
        buf[0] = (int32_t)(max_pkts - total);
    }
    return dropped;
}
/* buffer.py:82-104; returns sent_packets */
static int64_t buffer_send(int32_t *buf, int depth, double avail)
{
    int64_t before = 0, after = 0;
    for (int i = 0; i < depth; ++i) before += buf[i];
    if (before != 0 || avail != 0) {
        for (int i = depth - 1; i >= 0; --i) {
            if (avail >= (double)buf[i]) {
                avail -= (double)buf[i];
                buf[i] = 0;
            } else {
                buf[i] -= (int32_t)avail;
                break;
            }
        }
    }
    for (int i = 0; i < depth; ++i) after += buf[i];
    return before - after;
}
/* buffer.py:106-130 */
static void buffer_stats(const int32_t *buf, int depth, int64_t max_pkts, double *occ, double *lat)
{
    int64_t tot = 0, wsum = 0;
    for (int i = 0; i < depth; ++i) { tot += buf[i]; wsum += (int64_t)buf[i] * i; }
    *occ = (double)tot / (double)max_pkts;
    *lat = tot != 0 ? (double)wsum / (double)tot : 0.0;
}

/* Known-answer entry point for buffer.py:133-149 (Buffer.main). out: [n][2][depth] rows
 * (after receive, after send), dropped[n], sent[n]; returns final average delay. */
double orc_buffer_main(int max_pkts, int max_age, const int *arrivals, int n, int send_per_step,
                       int32_t *rows, int64_t *dropped, int64_t *sent)
{
    int depth = max_age + 1;
    int32_t *buf = (int32_t *)calloc(depth, sizeof(int32_t));
    for (int k = 0; k < n; ++k) {
        dropped[k] = buffer_receive(buf, depth, max_pkts, (double)arrivals[k]);
        memcpy(rows + (size_t)(2 * k) * depth, buf, depth * sizeof(int32_t));
        sent[k] = buffer_send(buf, depth, (double)send_per_step);
        memcpy(rows + (size_t)(2 * k + 1) * depth, buf, depth * sizeof(int32_t));
    }
    double occ, lat;
    buffer_stats(buf, depth, max_pkts, &occ, &lat);
    free(buf);
    return lat;
}

/* per-UE params come from the UE's slice (associations/mult_slice.py:468-488 update_ues);
 * unassociated UEs keep the env defaults (comm_env.py:184-198). */
static void ue_params(const orc_scenario *sc, int u, int *depth, int64_t *max_pkts, double *pkt_bits)
{
    int s = sc->ue_slice[u];
    if (s < 0) { *depth = 101; *max_pkts = 10240; *pkt_bits = 65536.0; return; }
    *depth = (int)sc->sl[s].buffer_latency + 1;
    *max_pkts = (int64_t)sc->sl[s].buffer_size;
    *pkt_bits = sc->sl[s].message_size;
}

typedef struct {
    double pkt_incoming[U], pkt_throughputs[U];
} orc_ue_aux;

/* ues.py:149-243 UEs.step for one env.  se: [U][R]; traffic: [U] bits; run (start,count).
 * Fills frame->{sent,dropped,occ,lat,sebar}. */
static void ues_step(const orc_handle *h, int e, const orc_scenario *sc, const double *se,
                     const double *traffic, const int16_t *rb_start, const int16_t *rb_count,
                     const uint32_t *sched_bits /* [U][5] or NULL: a scheduling matrix given directly */,
                     orc_frame *fr, orc_ue_aux *aux)
{
    const double bw_rb = h->cfg.bandwidth / (double)R; /* ues.py:110 */
    for (int u = 0; u < U; ++u) {
        int depth; int64_t max_pkts; double pkt_bits;
        ue_params(sc, u, &depth, &max_pkts, &pkt_bits);
        double terms[R];
        for (int r = 0; r < R; ++r) {
            double sched = sched_bits ? (double)((sched_bits[u * 5 + (r >> 5)] >> (r & 31)) & 1u)
                                      : ((r >= rb_start[u] && r < rb_start[u] + rb_count[u]) ? 1.0 : 0.0);
            terms[r] = (bw_rb * sched) * se[u * R + r];
        }
        double thr = floor(np_sum(terms, R) / pkt_bits); /* ues.py:108-116 */
        double inc = floor(traffic[u] / pkt_bits);       /* ues.py:199 */
        int32_t *buf = h->buf + ((size_t)e * U + u) * h->cfg.depth_cap;
        int64_t dropped = buffer_receive(buf, depth, max_pkts, inc);
        int64_t sent = buffer_send(buf, depth, thr);
        buffer_stats(buf, depth, max_pkts, &fr->occ[u], &fr->lat[u]);
        fr->sent[u] = (int32_t)sent;
        fr->dropped[u] = (int32_t)dropped;
        fr->sebar[u] = np_sum(se + u * R, R) / (double)R; /* np.mean(SE[0,u,:]) */
        fr->rb_count[u] = rb_count[u];
        aux->pkt_incoming[u] = inc;
        aux->pkt_throughputs[u] = thr;
        fr->thr[u] = thr;
    }
}

/* ------------------------------------------------------------------------------------
 * Action -> RB allocation (agents/ib_sched.py:223-349, agents/common.py:442-701)
 * ------------------------------------------------------------------------------------ */
/* common.py:481-505.  np.argsort pinned to kind="stable" (SURVEY 7.1): descending value,
 * ties -> higher index first. */
static void round_int_equal_sum(const double *x, int n, int64_t target, int64_t *out)
{
    int idx[8], m = 0;
    double v[8];
    for (int i = 0; i < n; ++i) { out[i] = 0; if (x[i] != 0) { idx[m] = i; v[m] = x[i]; ++m; } }
    if (m == 0) return; /* the reference would raise (i % 0) */
    double sum = np_sum(v, m);
    int64_t p[8], tot = 0;
    for (int i = 0; i < m; ++i) { p[i] = (int64_t)floor(((double)target * v[i]) / sum); tot += p[i]; }
    int64_t adj = target - tot;
    /* stable ascending argsort, then reversed */
    int ord[8];
    for (int i = 0; i < m; ++i) ord[i] = i;
    for (int i = 1; i < m; ++i) { /* insertion sort = stable */
        int k = ord[i], j = i - 1;
        while (j >= 0 && v[ord[j]] > v[k]) { ord[j + 1] = ord[j]; --j; }
        ord[j + 1] = k;
    }
    for (int64_t i = 0; i < adj; ++i) p[ord[m - 1 - (int)(i % m)]] += 1;
    for (int i = 0; i < m; ++i) out[idx[i]] = p[i];
}
/* common.py:442-461 */
static void scores_to_rbs(const double *action, int64_t total, const double *assoc, int64_t *out)
{
    double a1[S], x[S];
    for (int i = 0; i < S; ++i) a1[i] = action[i] + 1;
    double sum = np_sum(a1, S);
    if (sum != 0) {
        for (int i = 0; i < S; ++i) x[i] = ((double)total * a1[i]) / sum;
    } else {
        double sa = np_sum(assoc, S);
        for (int i = 0; i < S; ++i) x[i] = ((double)total / sa) * assoc[i];
    }
    round_int_equal_sum(x, S, total, out);
}
/* common.py:508-555 (returns rbs per eligible UE; eligible indices in `el`) */
static int round_robin(const orc_frame *last, const int *ues, int n, int64_t rbs_s, int account_buffer,
                       int *el, int64_t *rbs)
{
    int m = 0;
    if (account_buffer)
        for (int i = 0; i < n; ++i) if (!np_isclose0(last->occ[ues[i]])) el[m++] = ues[i];
    if (m == 0) { for (int i = 0; i < n; ++i) el[i] = ues[i]; m = n; }
    int64_t base = (int64_t)floor((double)rbs_s / (double)m);
    int64_t rem = rbs_s % m;
    for (int i = 0; i < m; ++i) rbs[i] = base + (i < rem ? 1 : 0);
    return m;
}
/* shared by common.py:567-583 and 648-664 */
static void throughput_available(const orc_handle *h, const orc_scenario *sc, const orc_frame *last,
                                 const int *ues, int n, int64_t rbs_s, double *ta)
{
    for (int i = 0; i < n; ++i) {
        int u = ues[i];
        int depth; int64_t max_pkts; double pkt_bits;
        ue_params(sc, u, &depth, &max_pkts, &pkt_bits);
        double a = (last->sebar[u] * (((double)rbs_s * h->cfg.bandwidth) / (double)R)) / (double)n;
        double b = (last->occ[u] * (double)max_pkts) * pkt_bits;
        ta[i] = a < b ? a : b; /* np.minimum */
    }
}
static void rr_fallback(const orc_frame *last, const int *ues, int n, int64_t rbs_s, int64_t *rbs)
{
    int el[NS];
    round_robin(last, ues, n, rbs_s, 0, el, rbs);
}
/* common.py:558-636 */
static void proportional_fairness(const orc_handle *h, const orc_env *env, const orc_scenario *sc,
                                  const int *ues, int n, int64_t rbs_s, int64_t *rbs)
{
    const orc_frame *last = &env->win[0];
    double ta[NS], snt[NS], w[NS], x[NS];
    throughput_available(h, sc, last, ues, n, rbs_s, ta);
    double tamax = ta[0];
    for (int i = 1; i < n; ++i) if (ta[i] > tamax) tamax = ta[i];
    for (int i = 0; i < n; ++i) {
        int u = ues[i];
        int depth; int64_t max_pkts; double pkt_bits;
        ue_params(sc, u, &depth, &max_pkts, &pkt_bits);
        int64_t tot = 0;
        for (int k = 0; k < env->win_len; ++k) tot += env->win[k].sent[u];
        snt[i] = ((double)tot / (double)env->win_len) * pkt_bits; /* np.mean(list, axis=0) */
        if (np_isclose0(ta[i])) snt[i] = 1;
        w[i] = !np_isclose0(snt[i]) ? ta[i] / snt[i] : 2 * tamax;
    }
    double sw = np_sum(w, n);
    if (sw != 0) {
        for (int i = 0; i < n; ++i) x[i] = ((double)rbs_s * w[i]) / sw;
        round_int_equal_sum(x, n, rbs_s, rbs);
    } else {
        rr_fallback(last, ues, n, rbs_s, rbs);
    }
}
/* common.py:639-701 */
static void max_throughput(const orc_handle *h, const orc_env *env, const orc_scenario *sc,
                           const int *ues, int n, int64_t rbs_s, int64_t *rbs)
{
    const orc_frame *last = &env->win[0];
    double ta[NS], x[NS];
    throughput_available(h, sc, last, ues, n, rbs_s, ta);
    double st = np_sum(ta, n);
    if (st != 0) {
        for (int i = 0; i < n; ++i) x[i] = ((double)rbs_s * ta[i]) / st;
        round_int_equal_sum(x, n, rbs_s, rbs);
    } else {
        rr_fallback(last, ues, n, rbs_s, rbs);
    }
}
static int slice_ues(const orc_scenario *sc, int s, int *ues)
{
    int n = 0;
    for (int u = 0; u < U; ++u) if (sc->ue_slice[u] == s) ues[n++] = u;
    return n;
}
/* ib_sched.py:223-349.  Returns 0, or -1 for an invalid intra action (ValueError :341). */
static int action_format(const orc_handle *h, const orc_env *env, const orc_scenario *sc,
                         const double *a0_in, const int32_t *intra, int16_t *rb_start, int16_t *rb_count,
                         int64_t *rbs_per_slice)
{
    for (int u = 0; u < U; ++u) rb_start[u] = rb_count[u] = 0;
    for (int s = 0; s < S; ++s) rbs_per_slice[s] = 0;
    double assoc[S], nact = 0;
    for (int s = 0; s < S; ++s) { assoc[s] = (sc->sl[s].flags & F_ACTIVE) ? 1.0 : 0.0; nact += assoc[s]; }
    if (nact == 0) return 0; /* ib_sched.py:240-245 */
    double a0[S];
    for (int s = 0; s < S; ++s) a0[s] = a0_in[env->sorted_slices[s]]; /* :247 (quirk Q3) */
    for (int s = 0; s < S; ++s) if (assoc[s] == 0) a0[s] = -1;        /* :248-255 */
    int64_t rbg[S];
    scores_to_rbs(a0, R / RBG, assoc, rbg);
    for (int s = 0; s < S; ++s) rbs_per_slice[s] = rbg[s] * RBG;
    const orc_frame *last = &env->win[0];
    int64_t base = 0;
    for (int s = 0; s < S; ++s) {
        int ues[U], n = slice_ues(sc, s, ues);
        if (n > 0) {
            int algo = h->cfg.fixed_intra >= 0 ? h->cfg.fixed_intra : intra[s];
            int el[U], m = n;
            int64_t rbs[U];
            if (algo == 0) m = round_robin(last, ues, n, rbs_per_slice[s], 1, el, rbs);
            else if (algo == 1) { proportional_fairness(h, env, sc, ues, n, rbs_per_slice[s], rbs); memcpy(el, ues, sizeof(int) * n); }
            else if (algo == 2) { max_throughput(h, env, sc, ues, n, rbs_per_slice[s], rbs); memcpy(el, ues, sizeof(int) * n); }
            else return -1;
            int64_t rb = base; /* common.py:464-478 distribute_rbs_ues */
            for (int i = 0; i < m; ++i) {
                rb_count[el[i]] = (int16_t)rbs[i];
                rb_start[el[i]] = rbs[i] > 0 ? (int16_t)rb : 0;
                rb += rbs[i];
            }
        }
        base += rbs_per_slice[s];
    }
    return 0;
}
/* agents/marr.py:40-47 */
static void policy_marr(const orc_scenario *sc, double *a0)
{
    for (int s = 0; s < S; ++s) a0[s] = sc->sl[s].n_ues > 0 ? 1.0 : -1.0;
}
/* agents/mapf.py:41-111 */
static void policy_mapf(const orc_env *env, const orc_scenario *sc, double *a0)
{
    double occ_mbps[S] = {0}, sent_mbps[S] = {0}, w[S];
    const orc_frame *last = &env->win[0];
    for (int s = 0; s < S; ++s) {
        if (!(sc->sl[s].flags & F_ACTIVE)) continue;
        int ues[U], n = slice_ues(sc, s, ues);
        double occ[NS], snt[NS];
        for (int i = 0; i < n; ++i) {
            occ[i] = last->occ[ues[i]];
            int64_t tot = 0;
            for (int k = 0; k < env->win_len; ++k) tot += env->win[k].sent[ues[i]];
            snt[i] = (double)tot / (double)env->win_len;
        }
        occ_mbps[s] = (((np_sum(occ, n) / n) * sc->sl[s].buffer_size) * sc->sl[s].message_size) / 1e6;
        sent_mbps[s] = ((np_sum(snt, n) / n) * sc->sl[s].message_size) / 1e6;
    }
    double mx = occ_mbps[0];
    for (int s = 1; s < S; ++s) if (occ_mbps[s] > mx) mx = occ_mbps[s];
    for (int s = 0; s < S; ++s) {
        w[s] = !np_isclose0(sent_mbps[s]) ? occ_mbps[s] / sent_mbps[s] : 2 * mx;
        if (!(sc->sl[s].flags & F_ACTIVE)) w[s] = 0;
    }
    double sw = np_sum(w, S);
    for (int s = 0; s < S; ++s) a0[s] = (sw > 0 ? w[s] / sw : 2.0) - 1;
}
/* random policy from the counter-based RNG (BASELINE config 5: actions generated on device) */
static void policy_random(const orc_handle *h, const orc_env *env, int gid, double *a0, int32_t *intra)
{
    uint32_t w[12];
    for (int q = 0; q < 3; ++q) rng_quad(env_key(&h->cfg, env->visit), K_POLICY, 0, q, env->step + 1, env->episode, gid, w + 4 * q);
    for (int i = 0; i < S; ++i) a0[i] = (double)w[i] * (1.0 / 2147483648.0) - 1.0;
    for (int i = 0; i < S; ++i) intra[i] = (int32_t)(((uint64_t)w[5 + i] * 3u) >> 32);
}

/* ------------------------------------------------------------------------------------
 * Observation, masks, reward (agents/common.py:9-439, agents/ib_sched.py:63-221,351-392)
 * ------------------------------------------------------------------------------------ */
static void window_push(orc_env *env, const orc_frame *fr)
{
    for (int k = W - 1; k > 0; --k) env->win[k] = env->win[k - 1]; /* deque.appendleft, maxlen 10 */
    env->win[0] = *fr;
    if (env->win_len < W) env->win_len++;
}
/* common.py:9-65 + 68-340 for one UE of slice `sl`; r = intent_overfulfillment_rate 0.2 */
static void intent_drift_ue(const orc_env *env, const orc_slice *sl, int u, double *drift /*[3]*/)
{
    const double r = 0.2;
    const orc_frame *f0 = &env->win[0];
    drift[0] = drift[1] = drift[2] = 0;
    if (sl->flags & F_THR) {
        double v = sl->thr_value;
        double m = (double)((int64_t)f0->sent[u] * (int64_t)sl->message_size) / 1e6; /* :26-31 */
        int zero = np_isclose0(f0->occ[u]);                                          /* :104-107 */
        if (env->win_len > 1) zero = zero || np_isclose0(env->win[1].occ[u]);         /* :108-116 */
        if (zero) m = v * (1.1 + r);                                                  /* :117-119 */
        if (m >= v) {
            if (m > v * (1 + r)) drift[0] += 1;                 /* :164-168 */
            else drift[0] += (m - v) / (v * r);                  /* :154-162 */
        } else {
            drift[0] -= (v - m) / v;                             /* :172-181 */
        }
    }
    if (sl->flags & F_REL) {
        double v = sl->rel_value;
        int64_t snt = 0, drp = 0;
        for (int k = 0; k < env->win_len; ++k) { snt += env->win[k].sent[u]; drp += env->win[k].dropped[u]; }
        double den = (f0->occ[u] * sl->buffer_size + (double)drp) + (double)snt; /* :40-47 */
        double m = den != 0 ? (double)drp / den : 0.0;                           /* :48-53 */
        double t = (100 - v) / 100;
        if (100 * (1 - m) >= v) {                                                /* :124-126 */
            if (m < t * (1 - r)) drift[1] += 1;                                  /* :188-220 */
            else drift[1] += (t - m) / (t * r);
        } else {
            drift[1] -= (m - t) / (v / 100);                                     /* :223-233 */
        }
    }
    if (sl->flags & F_LAT) {
        double v = sl->lat_value;
        double m = f0->lat[u];
        if (m <= v) {
            if (m < v * (1 - r)) drift[2] += 1;                                  /* :290-319 */
            else drift[2] += (v - m) / (v * r);
        } else {
            drift[2] -= (m - v) / (sl->buffer_latency - v);                      /* :322-335 */
        }
    }
}
/* SchedTWC / SchedColORAN (agents/sched_twc.py:166-413, agents/sched_colran.py:166-417): 50-value observation
 * and scalar reward, evaluated on the window AFTER the agents' second push of the frame (sched_twc.py:176-179):
 * env->win then holds every frame twice, exactly like the reference's deque. */
static double np_mean_small(const double *a, int n) { return np_sum(a, n) / (double)n; }
static void alt_surface(const orc_handle *h, const orc_env *env, const orc_scenario *sc, double *obs /*[50]*/,
                        double *reward /* NULL at reset */)
{
    const orc_frame *f0 = &env->win[0];
    double valid[15], weights[15];
    int nv = 0;
    double r_col = 0;
    for (int s = 0; s < S; ++s) {
        const orc_slice *sl = &sc->sl[s];
        int ues[U], n = slice_ues(sc, s, ues);
        double req[3] = {0, 0, 0};
        if (n != 0) {                                           /* sched_twc.py:217-229 */
            if (sl->flags & F_REL) req[0] = sl->rel_value;
            if (sl->flags & F_LAT) req[1] = sl->lat_value;
            if (sl->flags & F_THR) req[2] = sl->thr_value;
        }
        for (int k = 0; k < 3; ++k) obs[3 * s + k] = req[k];
        double pkt = n != 0 ? sl->message_size : 0;             /* :234-240 */
        double se[NS], thr[NS], eff[NS], occ[NS], lat[NS], loss[NS], thr_pk[NS];
        for (int i = 0; i < n && i < NS; ++i) {
            int u = ues[i];
            se[i] = f0->sebar[u];
            thr_pk[i] = f0->thr[u];
            thr[i] = (f0->thr[u] * pkt) / 1e6;                  /* pkts_to_mbps :430-431 */
            eff[i] = (double)((int64_t)f0->sent[u] * (int64_t)pkt) / 1e6;
            occ[i] = f0->occ[u];
            lat[i] = f0->lat[u];
            int64_t snt = 0, drp = 0;                           /* get_metric_value "reliability", common.py:32-53 */
            for (int k = 0; k < env->win_len; ++k) { snt += env->win[k].sent[u]; drp += env->win[k].dropped[u]; }
            double den = (f0->occ[u] * sl->buffer_size + (double)drp) + (double)snt;
            loss[i] = den != 0 ? (double)drp / den : 0.0;
        }
        obs[15 + s] = n ? np_mean_small(se, n) : 0.0;
        obs[20 + s] = n ? np_mean_small(thr, n) : 0.0;
        obs[25 + s] = n ? np_mean_small(eff, n) : 0.0;
        obs[30 + s] = n ? np_mean_small(occ, n) : 0.0;
        obs[35 + s] = n ? np_mean_small(lat, n) : 0.0;
        obs[40 + s] = n ? np_mean_small(loss, n) : 0.0;
        obs[45 + s] = (sl->flags & F_ACTIVE) ? sl->traffic : 0.0; /* :319-336 */
        if (!reward || n == 0) continue;
        /* TWC: slice drift per declared intent (calculate_slice_ue_obs, common.py:343-378) */
        double dsum[3][NS];
        for (int i = 0; i < n; ++i) {
            double d[3] = {0, 0, 0};
            if (sl->flags & F_REQ) intent_drift_ue(env, sl, ues[i], d);
            for (int k = 0; k < 3; ++k) dsum[k][i] = d[k];
        }
        const int fbit[3] = {F_THR, F_REL, F_LAT};
        for (int k = 0; k < 3; ++k) {
            double d = ((sl->flags & F_REQ) && (sl->flags & fbit[k])) ? np_mean_small(dsum[k], n) : -2.0;
            if (fabs(d - (-2.0)) <= 1e-8 + 1e-5 * 2.0) continue; /* np.isclose(x, -2) */
            valid[nv] = d > 0 ? 0.0 : d;                          /* sched_twc.py:392-394 */
            weights[nv] = sl->priority != 0 ? 2.0 : 1.0;          /* :382-390 */
            ++nv;
        }
        /* ColORAN (sched_colran.py:364-415) */
        if (sl->flags & F_ACTIVE) {
            if (sl->flags & F_EMBB) r_col += ((np_mean_small(thr_pk, n) * sl->message_size) / 1e6) / 200.0;
            if (sl->flags & F_URLLC) r_col -= (((np_mean_small(occ, n) * sl->buffer_size) * sl->message_size) / 1e6) / 2000.0;
        }
    }
    if (!reward) return;
    if (h->cfg.surface == 1) {
        double neg[15], nw[15];
        int m = 0;
        for (int i = 0; i < nv; ++i) if (valid[i] < 0) { neg[m] = valid[i]; nw[m] = weights[i]; ++m; }
        double wsum = np_sum(nw, m);
        if (!np_isclose0(wsum)) {
            double t[15];
            for (int i = 0; i < m; ++i) t[i] = (neg[i] * nw[i]) / wsum;
            *reward = np_sum(t, m);
        } else {
            *reward = 0.0;
        }
    } else {
        *reward = r_col;
    }
}
/* ib_sched.py:351-370: stable ascending argsort of n_ues * traffic */
static void sort_slices(const orc_scenario *sc, int32_t *sorted)
{
    double key[S];
    for (int s = 0; s < S; ++s)
        key[s] = (double)sc->sl[s].n_ues * ((sc->sl[s].flags & F_REQ) ? sc->sl[s].traffic : 0.0);
    for (int s = 0; s < S; ++s) sorted[s] = s;
    for (int i = 1; i < S; ++i) {
        int k = sorted[i], j = i - 1;
        while (j >= 0 && key[sorted[j]] > key[k]) { sorted[j + 1] = sorted[j]; --j; }
        sorted[j + 1] = k;
    }
}
typedef struct {
    double obs0[S * 10];
    int8_t mask0[S];
    double obs_i[S][19];   /* indexed by slice (player_{s+1}) */
    int8_t mask_i[S][NS];
    int8_t sorted_slices[S];
} orc_obs;

/* ib_sched.py:63-204.  The new frame must already be win[0]. */
static void obs_space_format(const orc_handle *h, orc_env *env, const orc_scenario *sc, orc_obs *o)
{
    const orc_frame *f0 = &env->win[0];
    if (h->cfg.sort_slices) sort_slices(sc, env->sorted_slices);
    for (int s = 0; s < S; ++s) o->mask0[s] = (sc->sl[s].flags & F_ACTIVE) ? 1 : 0; /* :76-81 */
    for (int pos = 0; pos < S; ++pos) {
        int s = env->sorted_slices[pos];
        const orc_slice *sl = &sc->sl[s];
        int ues[U], n = slice_ues(sc, s, ues);
        double dsl[3] = {-2, -2, -2}, act[3] = {0, 0, 0};
        double drift[NS][3];
        int has_req = (sl->flags & F_REQ) != 0;
        if (n > 0 && has_req) {
            for (int j = 0; j < n; ++j) intent_drift_ue(env, sl, ues[j], drift[j]);
            const int fl[3] = {F_THR, F_REL, F_LAT};
            for (int k = 0; k < 3; ++k) {
                if (!(sl->flags & fl[k])) continue;
                double col[NS];
                for (int j = 0; j < n; ++j) col[j] = drift[j][k];
                dsl[k] = np_sum(col, n) / (double)n; /* common.py:371-373 np.mean */
            }
        }
        for (int k = 0; k < 3; ++k) { /* ib_sched.py:142-145 */
            int close = fabs(dsl[k] - (-2.0)) <= 1e-8 + 1e-5 * 2.0;
            act[k] = close ? 0 : 1;
            if (close) dsl[k] = 0;
        }
        double sebar[NS] = {0}, occ[NS] = {0};
        for (int j = 0; j < n; ++j) { sebar[j] = f0->sebar[ues[j]]; occ[j] = f0->occ[ues[j]]; }
        double traffic = (sl->flags & F_ACTIVE) ? sl->traffic : 0;       /* :125-134 */
        double prio = n != 0 ? sl->priority : 0;                         /* :135-141 */
        double se_slice = n > 0 ? np_sum(sebar, n) / (double)n : 0;      /* :146-157 */
        double *p0 = o->obs0 + pos * 10;                                 /* :160-173 */
        p0[0] = dsl[0]; p0[1] = dsl[1]; p0[2] = dsl[2];
        p0[3] = act[0]; p0[4] = act[1]; p0[5] = act[2];
        p0[6] = prio; p0[7] = traffic / 120.0; p0[8] = (double)n / 5.0; p0[9] = se_slice / 40.0;
        double rbs_alloc = 0;                                            /* :176-181 */
        for (int j = 0; j < n; ++j) rbs_alloc += (double)f0->rb_count[ues[j]];
        double *pi = o->obs_i[s];                                        /* :186-200 */
        pi[0] = dsl[0]; pi[1] = dsl[1]; pi[2] = dsl[2];
        pi[3] = act[0]; pi[4] = act[1]; pi[5] = act[2];
        pi[6] = rbs_alloc / (double)R; pi[7] = traffic / 120.0; pi[8] = (double)n / 5.0;
        for (int j = 0; j < NS; ++j) { pi[9 + j] = occ[j]; pi[14 + j] = sebar[j] / 40.0; }
        for (int j = 0; j < NS; ++j) o->mask_i[s][j] = j < n ? 1 : 0;
    }
    for (int s = 0; s < S; ++s) o->sorted_slices[s] = (int8_t)env->sorted_slices[s];
    memcpy(env->obs0, o->obs0, sizeof(env->obs0));
    memcpy(env->obs_i, o->obs_i, sizeof(env->obs_i));
}
/* ib_sched.py:206-221 + 372-392 + common.py:381-439.  reward[0] inter, reward[1+s] intra of slice s */
static void calculate_reward(const orc_env *env, const orc_scenario *sc, double *reward)
{
    double un[S * 10];
    for (int pos = 0; pos < S; ++pos) /* unsort_slices */
        memcpy(un + env->sorted_slices[pos] * 10, env->obs0 + pos * 10, 10 * sizeof(double));
    double A[S] = {0}, P[S] = {0};
    for (int e = 0; e < S; ++e) {
        if (!(sc->sl[e].flags & F_ACTIVE)) continue;
        P[e] = sc->sl[e].priority;
        double mn = 0; int cnt = 0;
        for (int k = 0; k < 3; ++k) {
            double x = un[e * 10 + k];
            if (fabs(x + 2.0) <= 1e-8 + 2e-5) continue; /* common.py:406 */
            mn = cnt ? (x < mn ? x : mn) : x; cnt++;
        }
        A[e] = cnt ? mn : 1;
    }
    int nneg = 0, npneg = 0;
    for (int e = 0; e < S; ++e) { nneg += A[e] < 0; npneg += (P[e] * A[e]) < 0; }
    if (nneg == 0) {
        reward[0] = np_sum(A, S) / (double)S;
    } else {
        double sel[S]; int m = 0;
        for (int e = 0; e < S; ++e) if (npneg ? (A[e] * P[e] < 0) : (A[e] < 0)) sel[m++] = A[e];
        reward[0] = np_sum(sel, m) / (double)m - (npneg ? 1 : 0);
    }
    for (int s = 0; s < S; ++s) { /* common.py:428-437 */
        const double *o = env->obs_i[s];
        reward[1 + s] = 0;
        int any = 0; double mn = 0;
        for (int k = 0; k < 3; ++k) if (o[3 + k] != 0) { mn = any ? (o[k] < mn ? o[k] : mn) : o[k]; any = 1; }
        if (o[3] > 0 || o[4] > 0 || o[5] > 0) reward[1 + s] = mn;
    }
}

/* ------------------------------------------------------------------------------------
 * Episode metrics (replaces metrics.py:62-100 history; definitions from
 * results/gen_results.py:791-809 total throughput, 874-970 violations, 973-1022 distance)
 * acc: 0-5 sum reward player_0..5 | 6 sent pkts | 7 dropped pkts | 8 incoming pkts |
 *      9 throughput Mbit | 10 violations | 11 violations (priority slices) |
 *      12 intent distance | 13 intent distance (priority) | 14 env-steps | 15 episodes
 * ------------------------------------------------------------------------------------ */
static void accumulate_metrics(orc_env *env, const orc_scenario *sc, const orc_ue_aux *aux, const double *reward)
{
    const orc_frame *f0 = &env->win[0];
    for (int p = 0; p < 6; ++p) env->acc[p] += reward[p];
    for (int u = 0; u < U; ++u) {
        env->acc[6] += (double)f0->sent[u];
        env->acc[7] += (double)f0->dropped[u];
        env->acc[8] += aux->pkt_incoming[u];
    }
    for (int pos = 0; pos < S; ++pos) {
        int s = env->sorted_slices[pos];
        const orc_slice *sl = &sc->sl[s];
        if (!(sl->flags & F_ACTIVE)) continue;
        int64_t snt = 0;
        for (int u = 0; u < U; ++u) if (sc->ue_slice[u] == s) snt += f0->sent[u];
        env->acc[9] += ((double)snt * sl->message_size) / 1e6;
        double mn = 0; int any = 0;
        for (int k = 0; k < 3; ++k)
            if (env->obs0[pos * 10 + 3 + k] != 0) {
                double d = env->obs0[pos * 10 + k];
                if (d < 0) { mn = any ? (d < mn ? d : mn) : d; any = 1; }
            }
        if (any) {
            env->acc[10] += 1; env->acc[12] += mn;
            if (sl->priority != 0) { env->acc[11] += 1; env->acc[13] += mn; }
        }
    }
    env->acc[14] += 1;
}

/* ------------------------------------------------------------------------------------
 * Env orchestration (comm_env.py:271-482, associations/mult_slice.py:444-452,
 * associations/mult_slice_seq.py:38-46)
 * ------------------------------------------------------------------------------------ */
typedef struct {
    double *pkt_incoming, *pkt_throughputs; /* [E][U] */
    int32_t *sent, *dropped;                /* [E][U] */
    double *occ, *lat;                      /* [E][U] */
    int16_t *rb_start, *rb_count;           /* [E][U] */
    double *obs0;                           /* [E][50] */
    int8_t *mask0;                          /* [E][5] */
    double *obs_i;                          /* [E][5][19] */
    int8_t *mask_i;                         /* [E][5][5] */
    int8_t *sorted_slices;                  /* [E][5] */
    double *reward;                         /* [E][6] */
    uint8_t *done;                          /* [E] */
    double *a0_used;                        /* [E][5] */
    int32_t *intra_used;                    /* [E][5] */
    double *se_used;                        /* [E][U][R] */
    double *traffic_used;                   /* [E][U] */
} orc_io;

static int scen_of_episode(const orc_config *c, int episode)
{
    int s = c->assoc_mode == 0 ? episode % c->scen_mod : episode / c->scen_div;
    return s % c->n_scen;
}
static void write_obs(const orc_io *io, int e, const orc_obs *o)
{
    if (io->obs0) memcpy(io->obs0 + (size_t)e * 50, o->obs0, sizeof(o->obs0));
    if (io->mask0) memcpy(io->mask0 + (size_t)e * S, o->mask0, S);
    if (io->obs_i) memcpy(io->obs_i + (size_t)e * S * 19, o->obs_i, sizeof(o->obs_i));
    if (io->mask_i) memcpy(io->mask_i + (size_t)e * S * NS, o->mask_i, S * NS);
    if (io->sorted_slices) memcpy(io->sorted_slices + (size_t)e * S, o->sorted_slices, S);
}
static void channel_frame(const orc_handle *h, int e, const orc_env *env, int frame, const double *se_in, double *se)
{
    const orc_scenario *sc = &h->bank[env->scen];
    if (h->cfg.channel_mode == 0) {
        memcpy(se, se_in + (size_t)e * U * R, sizeof(double) * U * R);
    } else if (h->cfg.channel_mode == 3) {
        /* channels/quadriga.py:38-77: the file of this episode, frame = the env's step counter
         * (0 at reset AND at the first step: comm_env.py:303-305 runs before step_number += 1) */
        int k = frame > 0 ? frame - 1 : 0;
        k %= h->bank_steps; /* a trace shorter than the episode wraps around */
        int slot = env->episode - h->bank_first;   /* the reference opens this episode's own file (quadriga.py:46-55, 78-87) */
        if (h->bank_wrap) { slot %= h->bank_eps; if (slot < 0) slot += h->bank_eps; }
        else if (slot < 0 || slot >= h->bank_eps) slot = 0; /* the CUDA path flags this (ERR_TRACE_EPISODE) */
        memcpy(se, h->se_bank + ((size_t)slot * h->bank_steps + k) * U * R, sizeof(double) * U * R);
    } else if (h->cfg.channel_mode == 1) {
        for (int i = 0; i < U * R; ++i) se[i] = 2.0; /* channels/fixed_se.py:26 */
    } else {
        for (int u = 0; u < U; ++u) {
            if (sc->ue_slice[u] < 0) { for (int r = 0; r < R; ++r) se[u * R + r] = 0.0; continue; }
            native_se_row(env_key(&h->cfg, env->visit), u, frame, env->episode, h->cfg.env_id_base + e, se + u * R);
        }
    }
}
/* assoc_mode 2: MultSliceAssociation generator mode (associations/mult_slice.py:359-416, 454-466) with the
 * counter-based generator: 3..5 slices, slots and types without replacement (partial Fisher-Yates), UEs per slice
 * uniform in the type's [min, max] (template sl[0].pad .. sl[0].n_ues), random UE ids. */
typedef struct { uint64_t key; int episode, gid, i; uint32_t cur[4]; } scen_words;
static uint32_t sw_next(scen_words *g)
{
    if ((g->i & 3) == 0) rng_quad(g->key, K_SCEN, 0, g->i >> 2, 0, g->episode, g->gid, g->cur);
    return g->cur[g->i++ & 3];
}
static int sw_below(scen_words *g, int n) { return (int)(((uint64_t)sw_next(g) * (uint32_t)n) >> 32); }
static void generate_scenario_row(const orc_handle *h, int e, int ep, orc_scenario *row)
{
    const orc_scenario *types = h->bank;
    const int T = h->cfg.n_scen;
    scen_words g = {env_key(&h->cfg, h->envs[e].visit), ep, h->cfg.env_id_base + e, 0, {0, 0, 0, 0}};
    int slots[S], tys[32], counts[S], ues[U];
    int k = 3 + sw_below(&g, S - 3 + 1);
    for (int j = 0; j < S; ++j) slots[j] = j;
    for (int j = 0; j < k; ++j) { int r = j + sw_below(&g, S - j), t = slots[j]; slots[j] = slots[r]; slots[r] = t; }
    for (int j = 0; j < T && j < 32; ++j) tys[j] = j;
    for (int j = 0; j < k; ++j) { int r = j + sw_below(&g, T - j), t = tys[j]; tys[j] = tys[r]; tys[r] = t; }
    int total = 0;
    for (int j = 0; j < k; ++j) {
        int mn = types[tys[j]].sl[0].pad, mx = types[tys[j]].sl[0].n_ues;
        counts[j] = mn + sw_below(&g, mx - mn + 1);
        total += counts[j];
    }
    for (int j = 0; j < U; ++j) ues[j] = j;
    for (int j = 0; j < total; ++j) { int r = j + sw_below(&g, U - j), t = ues[j]; ues[j] = ues[r]; ues[r] = t; }
    memset(row, 0, sizeof(*row));
    memset(row->ue_slice, -1, sizeof(row->ue_slice));
    int off = 0;
    for (int j = 0; j < k; ++j) {
        orc_slice sl = types[tys[j]].sl[0];
        sl.n_ues = counts[j];
        sl.pad = 0;
        row->sl[slots[j]] = sl;
        for (int c = 0; c < counts[j]; ++c) row->ue_slice[ues[off++]] = (int8_t)slots[j];
    }
}
/* comm_env.py:375-482 for env e.  initial_episode = -1: options {"initial_episode": -1} */
static int reset_one(orc_handle *h, int e, int initial_episode, const double *se_in, const orc_io *io)
{
    orc_env *env = &h->envs[e];
    const orc_config *c = &h->cfg;
    int init = initial_episode != -1 ? initial_episode : c->initial_episode;
    if (c->random_episodes) {                                  /* :434-444 */
        int moved = 0;
        if (initial_episode != -1) { env->ep_base = initial_episode; env->ep_gen += 1; env->ep_pos = 0; moved = 1; }
        if (env->step == c->max_steps) { env->ep_pos += 1; moved = 1; }  /* next(self.random_episodes_gen) */
        int n = h->ep_order ? h->ep_order_n : c->max_episodes - env->ep_base;
        if (n <= 0) return -2;
        env->ep_pos %= n;                                      /* itertools.cycle */
        if (moved)
            env->episode = h->ep_order ? h->ep_order[(size_t)e * n + env->ep_pos]
                                       : shuffled_episode(c, c->env_id_base + e, env->ep_gen, env->ep_base, n, env->ep_pos);
    }
    else if ((env->step == 0 && env->episode == 0) ||
        (env->step == c->max_steps && env->episode == c->max_episodes - 1) || initial_episode != -1)
        env->episode = init;                                   /* :416-421 */
    else if (env->step == 0 && env->episode <= c->max_episodes - 1) { /* :422-425 */ }
    else if (env->episode < c->max_episodes - 1) env->episode += 1;   /* :426-427 */
    else return -2;                                            /* :428-433 */
    env->step = 0;
    env->visit += 1;
    if (c->assoc_mode == 2) {
        int row0 = c->n_scen + 2 * e;                          /* the env's other private row */
        env->scen = env->scen == row0 ? row0 + 1 : row0;
        generate_scenario_row(h, e, env->episode, &h->own_bank[env->scen]);
    } else {
        env->scen = scen_of_episode(c, env->episode);
    }
    memset(h->buf + (size_t)e * U * c->depth_cap, 0, sizeof(int32_t) * U * c->depth_cap); /* :620-625 new UEs */
    double se[U * R];
    channel_frame(h, e, env, 0, se_in, se);
    if (io->se_used) memcpy(io->se_used + (size_t)e * U * R, se, sizeof(se));
    orc_frame fr;
    memset(&fr, 0, sizeof(fr));                                /* :474-478 zero metrics */
    for (int u = 0; u < U; ++u) fr.sebar[u] = np_sum(se + u * R, R) / (double)R;
    window_push(env, &fr);                                     /* quirk Q1: window survives reset */
    if (c->surface) window_push(env, &fr);
    orc_obs o;
    obs_space_format(h, env, &h->bank[env->scen], &o);
    if (c->surface) alt_surface(h, env, &h->bank[env->scen], o.obs0, NULL);
    write_obs(io, e, &o);
    return 0;
}
static int step_one(orc_handle *h, int e, const double *a0_in, const int32_t *intra_in, const double *se_in,
                    const double *traffic_in, const uint32_t *sched_in, const orc_io *io)
{
    orc_env *env = &h->envs[e];
    const orc_config *c = &h->cfg;
    const orc_scenario *sc = &h->bank[env->scen];
    int gid = c->env_id_base + e;
    double a0[S] = {0, 0, 0, 0, 0}; int32_t intra[S] = {0, 0, 0, 0, 0};
    const uint32_t *bits = sched_in ? sched_in + (size_t)e * U * 5 : NULL;
    if (bits) { /* action_format ran outside (comm_env.py:300 with the caller's own callback) */ }
    else if (c->policy == 0) {
        memcpy(a0, a0_in + (size_t)e * S, sizeof(a0));
        if (intra_in) memcpy(intra, intra_in + (size_t)e * S, sizeof(intra));
    } else if (c->policy == 1) policy_marr(sc, a0);
    else if (c->policy == 2) policy_mapf(env, sc, a0);
    else policy_random(h, env, gid, a0, intra);
    if (io->a0_used) memcpy(io->a0_used + (size_t)e * S, a0, sizeof(a0));
    if (io->intra_used) memcpy(io->intra_used + (size_t)e * S, intra, sizeof(intra));
    int16_t rb_start[U], rb_count[U]; int64_t rbs_per_slice[S];
    if (bits) { /* the matrix is given: RBs per UE = its row sum (ib_sched.py:188-190 reads sched_decision) */
        uint32_t seen[5] = {0, 0, 0, 0, 0};
        for (int u = 0; u < U; ++u) {
            int cnt = 0, first = 0;
            for (int r = R - 1; r >= 0; --r)
                if ((bits[u * 5 + (r >> 5)] >> (r & 31)) & 1u) { ++cnt; first = r; }
            rb_start[u] = (int16_t)first; rb_count[u] = (int16_t)cnt;
            for (int w = 0; w < 5; ++w) { if (seen[w] & bits[u * 5 + w]) return -3; seen[w] |= bits[u * 5 + w]; } /* comm_env.py:593-596 */
        }
    } else
    if (action_format(h, env, sc, a0, intra, rb_start, rb_count, rbs_per_slice)) return -1; /* comm_env.py:300 */
    int frame = env->step + 1;
    double se[U * R], traffic[U];
    channel_frame(h, e, env, frame, se_in, se);                /* :303-305 */
    for (int u = 0; u < U; ++u) {                              /* :306-311 */
        int s = sc->ue_slice[u];
        if (c->traffic_mode == 0) traffic[u] = traffic_in[(size_t)e * U + u];
        else traffic[u] = (s >= 0 && (sc->sl[s].flags & F_REQ))
                 ? native_traffic(h, env_key(c, env->visit), u, sc->sl[s].lam_idx, frame, env->episode, gid) : 0.0;
    }
    orc_frame fr; orc_ue_aux aux;
    ues_step(h, e, sc, se, traffic, rb_start, rb_count, bits, &fr, &aux); /* :320-326 */
    env->step += 1;                                            /* :339 */
    window_push(env, &fr);
    if (c->surface) window_push(env, &fr);                     /* sched_twc.py:176-179: the frame is pushed twice */
    orc_obs o;
    obs_space_format(h, env, sc, &o);                          /* :340 */
    double reward[6];
    calculate_reward(env, sc, reward);                         /* :341 */
    if (c->surface) alt_surface(h, env, sc, o.obs0, &reward[0]);
    accumulate_metrics(env, sc, &aux, reward);
    int done = env->step == c->max_steps;                      /* :370 */
    if (done) env->acc[15] += 1;
    size_t eu = (size_t)e * U;
    for (int u = 0; u < U; ++u) {
        if (io->pkt_incoming) io->pkt_incoming[eu + u] = aux.pkt_incoming[u];
        if (io->pkt_throughputs) io->pkt_throughputs[eu + u] = aux.pkt_throughputs[u];
        if (io->sent) io->sent[eu + u] = fr.sent[u];
        if (io->dropped) io->dropped[eu + u] = fr.dropped[u];
        if (io->occ) io->occ[eu + u] = fr.occ[u];
        if (io->lat) io->lat[eu + u] = fr.lat[u];
        if (io->rb_start) io->rb_start[eu + u] = rb_start[u];
        if (io->rb_count) io->rb_count[eu + u] = rb_count[u];
    }
    if (io->se_used) memcpy(io->se_used + eu * R, se, sizeof(se));
    if (io->traffic_used) memcpy(io->traffic_used + eu, traffic, sizeof(traffic));
    write_obs(io, e, &o);
    if (io->reward) memcpy(io->reward + (size_t)e * 6, reward, sizeof(reward));
    if (io->done) io->done[e] = (uint8_t)done;
    if (done && c->auto_reset) return reset_one(h, e, -1, se_in, io);
    return 0;
}

/* ---- batch API (ctypes) ---- */
orc_handle *orc_create(const orc_config *cfg, const orc_scenario *bank, const double *poisson_cdf)
{
    orc_handle *h = (orc_handle *)calloc(1, sizeof(orc_handle));
    h->cfg = *cfg;
    h->envs = (orc_env *)calloc(cfg->num_envs, sizeof(orc_env));
    h->buf = (int32_t *)calloc((size_t)cfg->num_envs * U * cfg->depth_cap, sizeof(int32_t));
    h->bank = bank; h->poisson_cdf = poisson_cdf;
    if (cfg->assoc_mode == 2) {
        h->own_bank = (orc_scenario *)calloc((size_t)cfg->n_scen + 2 * (size_t)cfg->num_envs, sizeof(orc_scenario));
        memcpy(h->own_bank, bank, sizeof(orc_scenario) * cfg->n_scen);
        h->bank = h->own_bank;
    }
    for (int e = 0; e < cfg->num_envs; ++e) {
        h->envs[e].episode = cfg->initial_episode;
        if (cfg->random_episodes) {                            /* comm_env.py:219-226 */
            h->envs[e].ep_base = cfg->initial_episode;
            int n = cfg->max_episodes - cfg->initial_episode;
            if (n > 0) h->envs[e].episode = shuffled_episode(cfg, cfg->env_id_base + e, 0, cfg->initial_episode, n, 0);
        }
        for (int s = 0; s < S; ++s) h->envs[e].sorted_slices[s] = s; /* ib_sched.py:58 */
    }
    return h;
}
#ifdef _OPENMP
#include <omp.h>
void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_max_threads(void) { return omp_get_max_threads(); }
#else
void orc_set_threads(int n) { (void)n; }
int orc_max_threads(void) { return 1; }
#endif
/* random_episodes with a given order [E][n]; the envs move to the table's entry at their position (irs_set_episode_order) */
void orc_set_episode_order(orc_handle *h, const int32_t *order, int32_t n)
{
    h->ep_order = order; h->ep_order_n = order ? n : 0;
    if (order)
        for (int e = 0; e < h->cfg.num_envs; ++e)
            h->envs[e].episode = order[(size_t)e * n + h->envs[e].ep_pos % n];
}
void orc_set_se_trace(orc_handle *h, const double *bank, int32_t n_eps, int32_t n_steps, int32_t first, int32_t wrap)
{
    h->se_bank = bank; h->bank_eps = n_eps; h->bank_steps = n_steps; h->bank_first = first; h->bank_wrap = wrap;
}
/* channels/quadriga.py:56-74 on one frame in file order [R][U] -> SE [U][R] */
void orc_power_to_se(const double *power, double *se, int64_t n_frames, int layout, double tx_power, double noise)
{
    const double c = tx_power / (double)R, den = 0.0 + noise;
    for (int64_t f = 0; f < n_frames; ++f)
        for (int i = 0; i < U * R; ++i) {
            double v = log2(1.0 + (c * power[f * U * R + i]) / den);
            if (layout == 0) { int r = i / U, u = i - r * U; se[f * U * R + u * R + r] = v; }
            else se[f * U * R + i] = v;
        }
}
void orc_destroy(orc_handle *h) { if (h) { free(h->envs); free(h->buf); free(h->own_bank); free(h); } }
/* the scenario row env e is playing (tests: generated scenarios) */
void orc_get_scenario(const orc_handle *h, int e, orc_scenario *out) { *out = h->bank[h->envs[e].scen]; }
int orc_reset(orc_handle *h, const uint8_t *mask, int initial_episode, const double *se, const orc_io *io)
{
    int rc = 0;
#pragma omp parallel for schedule(static) reduction(min : rc)
    for (int e = 0; e < h->cfg.num_envs; ++e) {
        if (mask && !mask[e]) continue;
        int r = reset_one(h, e, initial_episode, se, io);
        if (r < rc) rc = r;
    }
    return rc;
}
int orc_step(orc_handle *h, const double *a0, const int32_t *intra, const double *se, const double *traffic,
             const uint32_t *sched_bits /* [E][U][5] or NULL */, const orc_io *io)
{
    int rc = 0;
#pragma omp parallel for schedule(static) reduction(min : rc)
    for (int e = 0; e < h->cfg.num_envs; ++e) {
        int r = step_one(h, e, a0, intra, se, traffic, sched_bits, io);
        if (r < rc) rc = r;
    }
    return rc;
}
int32_t *orc_buffers(orc_handle *h) { return h->buf; }
void orc_get_state(const orc_handle *h, int32_t *step, int32_t *episode, int32_t *scen)
{
    for (int e = 0; e < h->cfg.num_envs; ++e) {
        step[e] = h->envs[e].step; episode[e] = h->envs[e].episode; scen[e] = h->envs[e].scen;
    }
}
/* per-env accumulators [E][16] and their fixed-order sum over envs [16] */
void orc_episode_metrics(const orc_handle *h, double *per_env, double *total)
{
    for (int k = 0; k < 16; ++k) total[k] = 0;
    for (int e = 0; e < h->cfg.num_envs; ++e)
        for (int k = 0; k < 16; ++k) {
            if (per_env) per_env[(size_t)e * 16 + k] = h->envs[e].acc[k];
            total[k] += h->envs[e].acc[k];
        }
}
int orc_sizeof_config(void) { return (int)sizeof(orc_config); }
int orc_sizeof_scenario(void) { return (int)sizeof(orc_scenario); }
/* exposed for unit tests of the NumPy summation model and the RNG restatement */
double orc_np_sum(const double *a, int n) { return np_sum(a, n); }
void orc_round_int_equal_sum(const double *x, int n, int64_t target, int64_t *out) { round_int_equal_sum(x, n, target, out); }
float orc_normal16(uint32_t h) { return normal16(h); }
uint32_t orc_feistel_perm(uint32_t i, uint32_t n, const uint32_t *k) { return feistel_perm(i, n, k); }
void orc_philox(uint32_t *c, uint32_t k0, uint32_t k1) { philox4x32_7(c, k0, k1); }
