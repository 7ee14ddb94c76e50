// irs_common.cuh - shared device-side definitions of the batched env-step kernels.
//
// Everything on this path that feeds a floor(), a comparison or a tie-break is IEEE binary64
// evaluated in the reference's (NumPy's) order; the translation unit is compiled with
// -fmad=false so that no multiply-add is contracted (NumPy rounds twice).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/irs_b200.h"

#define U IRS_MAX_UES
#define S IRS_MAX_SLICES
#define R IRS_NUM_RBS
#define W IRS_WINDOW
#define NS IRS_MAX_UES_SLICE
#define FULL 0xffffffffu
#define UE_LANES 0x01ffffffu /* lanes 0..24 */

enum { ERR_INTRA_ACTION = 1, ERR_EPISODE_RANGE = 2, ERR_RB_SUM = 4, ERR_TRACE_EPISODE = 8, ERR_RB_SHARED = 16 };

// Per-UE state is kept as small records so that one warp-wide access moves a whole record:
//   ue_i[E][25][8]  int32: the fields below (UI_*)
//   ue_d[E][25][4]  f64  : newest window frame (UD_*)
//   ring[E][25][10] int2 : (sent, dropped) of the last 10 frames (agents/ib_sched.py:49-51)
//   rb[E][25]       short2: (rb_start, rb_count) contiguous RB run of the UE
enum { UI_TOT = 0, UI_SUM_SENT, UI_SUM_DROP, UI_SENT, UI_DROP, UI_INC, UI_OCC1Z, UI_PATH, UI_N };
// UI_PATH word: row path of the step (2 bits) | live extent << 2 (14 bits) | 4-byte slot accesses of
// the step << 16 (bench.py's byte accounting reads the last).  Invariant: every AGE of a buffer row at
// or beyond its live extent is empty, so the row code never reads or writes past it.
#define ROW_LIVE(word) (((word) >> 2) & 0x3fff)
#define ROW_WORD(path, live, slots) ((path) | ((live) << 2) | ((slots) << 16))
enum { UD_OCC = 0, UD_LAT, UD_SEBAR, UD_THR, UD_N };

// Pointers into the caller-owned state blob (see irs_layout).
struct DevState {
    int32_t* buf;
    int32_t* ue_i;
    double* ue_d;
    int2* ring;
    short2* rb;
    int32_t *step, *episode, *scen, *win_len, *win_head, *err;
    int32_t *visit, *ep_pos, *ep_gen, *ep_base;
    long long* wsum;
    int8_t* sorted;
    double* acc;
    const irs_scenario* bank;
    const double* cdf;
    double* in_a0;
    int32_t* in_intra;
    double *out_obs0, *out_obsi, *out_reward;
    int8_t *out_mask0, *out_maski;
    uint8_t* out_done;
    double* metrics_total;
    int32_t* scratch;
    int32_t row_stride;
};

struct StepArgs {
    DevState st;
    irs_config cfg;
    double bw_rb;          // cfg.bandwidth / 135 (ues.py:110), rounded once on the host
    const double* a0;      // actions (device) or NULL
    const int32_t* intra;  // or NULL
    const uint32_t* sched_bits; // [E][25][5] scheduling matrix given directly (irs_actions.sched_bits) or NULL
    const double* se;      // injected SE or NULL
    const double* traffic; // injected traffic or NULL
    const double* se_bank; // channel_mode 3: [bank_eps][bank_steps][U][R] SE trace (irs_set_se_trace)
    int32_t bank_eps, bank_steps;
    int32_t bank_first, bank_wrap; // slot of an episode = episode - bank_first (modulo bank_eps when bank_wrap)
    const int32_t* ep_order; // random_episodes with a given order: [E][ep_order_n] (irs_set_episode_order), else NULL
    int32_t ep_order_n;
    irs_raw_outputs raw;   // optional extra outputs (all may be NULL)
    const uint8_t* reset_mask;
    const int32_t* initial_episode_dev;
    int32_t initial_episode;
    int32_t win_cap;  // distinct frames the observation window keeps: 10 (ib_sched.py:49-51), or 5 when the surface
                      // pushes every frame twice (sched_twc.py:176-179, sched_colran.py:176-179)
    int32_t win_mul;  // how often each kept frame counts in window sums: 1, or 2 with the double push
    int32_t defer_rollover; // k_env_tail leaves finished envs to k_rollover (the alternative surfaces read the
                            // last step's records in between)
    int32_t env_begin, env_end; // env range of this launch (irs_step_host pipelines chunks of the batch)
    int32_t ue_split_log2;      // k_ue_env: 2^this warps share an env (UE u belongs to warp u mod 2^this).  Small batches
                                // cannot fill the GPU with one warp per env, and a warp's serial walk over its UEs is
                                // the step's critical path; nothing in the row kernel couples the UEs of an env
};

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
// Key of env state for the counter-based generators: the seed with the env's reset count in the upper word, so an
// episode number that is visited again does not replay its noise (the reference's np_random stream runs on).
__host__ __device__ __forceinline__ uint64_t env_key(uint64_t seed, int visit) { return seed + ((uint64_t)(uint32_t)visit << 32); }
__device__ __forceinline__ bool isclose0(double x) { return fabs(x) <= 1e-8; } // np.isclose(x, 0)

// x / c for a COMPILE-TIME constant c, correctly rounded (= the IEEE quotient the reference computes) in five float64
// operations instead of the ~30 instructions of a general division: y = RN(1 / c) is folded by the compiler;
// q0 = x y, two residual corrections q <- q + (x - q c) y with the residual exact in an FMA (Markstein's theorem: the
// second correction of a faithful quotient by the correctly rounded reciprocal is the correctly rounded quotient).
// For finite x of moderate magnitude (no overflow / underflow of q c), which is every use below; checked against the
// hardware quotient on 2.4e9 values per constant incl. neighbours of multiples and of rounding midpoints.
__device__ __forceinline__ double div_const(double x, double c, double y)
{
    const double q0 = x * y;
    const double q1 = __fma_rn(__fma_rn(-q0, c, x), y, q0);
    return __fma_rn(__fma_rn(-q1, c, x), y, q1);
}
#define DIVC(x, c) div_const((x), (double)(c), 1.0 / (double)(c))
// x / n for a small count n = 1..10 (UEs of a slice, slices of a reward term, frames of the observation window) the same
// way, the reciprocal read from constant memory.  A float64 division is ~500 cycles of dependent latency on the per-env
// critical paths, this is ~50.  n = 0 (np.mean of an empty selection: 0 / 0) gives nan.
__constant__ double c_recip[W + 1] = {0.0, 1.0, 1.0 / 2, 1.0 / 3, 1.0 / 4, 1.0 / 5, 1.0 / 6, 1.0 / 7, 1.0 / 8, 1.0 / 9, 1.0 / 10};
__device__ __forceinline__ double div_small(double x, int n)
{
    const int i = n < 0 ? 0 : (n > W ? 0 : n);
    const double q = div_const(x, (double)n, c_recip[i]);
    return i >= 1 ? q : x / (double)n;
}

// Several numerators over ONE run-time divisor (a sum of weights, an intent's threshold): y = 1 / c is the only IEEE
// division, every quotient then takes the same five-operation sequence.  With y = RN(1 / c) the first correction leaves a
// faithful quotient and the second rounds it correctly (Markstein's theorem); tests/divc_check.c "rand:N" checks random
// divisors, among them the significands next to 1 and next to 2, with adversarial numerators.  c = 0 gives nan where the
// division gives inf: only used where a zero divisor's quotients are discarded.
struct Recip { double c, y; };
__device__ __forceinline__ Recip recip_of(double c) { Recip r; r.c = c; r.y = 1.0 / c; return r; }
__device__ __forceinline__ double div_by(double x, const Recip& r) { return div_const(x, r.c, r.y); }
// x / m for 0 <= x < 13,000 and a count m = 1..7 (UEs of a slice, non-zero shares): multiply by ceil(2^16 / m)
__constant__ unsigned c_imul[8] = {0u, 65536u, 32768u, 21846u, 16384u, 13108u, 10923u, 9363u};
__device__ __forceinline__ int idiv_small(int x, int m) { return (int)(((unsigned)x * c_imul[m & 7]) >> 16); }

#ifdef IRS_PHASE_CLOCK /* debug builds only (scripts/phase_clock.py): SM clock at the phase boundaries of k_env_step */
__device__ long long g_phase_clock[4096][4][8]; // [block][warp 0 | warp 1 | action core | tail core][stamp]
#define PHASE_STAMP(row, i) do { if (blockDim.x == 64 && (threadIdx.x & 31) == 0 && blockIdx.x < 4096) { long long t_; \
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_) :: "memory"); g_phase_clock[blockIdx.x][row][i] = t_; } } while (0)
#define PHASE_CLOCK(i) do { if (FULL_STEP) PHASE_STAMP(wib, i); } while (0)
#else
#define PHASE_STAMP(row, i) do { } while (0)
#define PHASE_CLOCK(i) do { } while (0)
#endif

// ---- ordered reductions over a group of lanes --------------------------------------------
// A group = the UE lanes of one slice (at most NS = 5, agents/ib_sched.py:50).  Its members are listed ONCE per core
// (GroupIdx: src[k] = lane of the k-th member in ascending lane order, the calling lane itself past the end; all-own
// for idle lanes, whose mask is 0), so that every reduction below is five shuffles plus its arithmetic instead of a
// find-first-set chain per call; all 32 lanes must call.  Members are visited in ascending lane order, which is the
// ascending UE / slice index order NumPy iterates in.  For n < 8 np.add.reduce is a serial sum starting from 0.
// (numpy/_core/src/umath/loops_utils.h.src, DOUBLE_pairwise_sum), then added to the reduction identity: 0. + res.
struct GroupIdx {
    int src[NS];
    int n, rank; // members; how many of them sit on lower lanes than the caller
    uint32_t mask;
};
__device__ __forceinline__ GroupIdx group_index(uint32_t mask)
{
    GroupIdx G;
    const int lane = lane_id();
    G.mask = mask;
    G.n = __popc(mask);
    G.rank = __popc(mask & ((1u << lane) - 1u));
    uint32_t m = mask;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        G.src[k] = m ? (__ffs(m) - 1) : lane;
        m &= m - 1;
    }
    return G;
}
__device__ __forceinline__ double group_sum_ordered(double x, const GroupIdx& G)
{
    double res = 0.;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        const double v = __shfl_sync(FULL, x, G.src[k]);
        if (k < G.n) res += v;
    }
    return 0. + res;
}
__device__ __forceinline__ int group_sum_int(int x, const GroupIdx& G)
{
    int res = 0;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        const int v = __shfl_sync(FULL, x, G.src[k]);
        if (k < G.n) res += v;
    }
    return res;
}
__device__ __forceinline__ double group_max(double x, const GroupIdx& G)
{
    double res = x;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        const double v = __shfl_sync(FULL, x, G.src[k]);
        if (k < G.n && v > res) res = v;
    }
    return res;
}

// agents/common.py:481-505 round_int_equal_sum over the members of G (one value per
// lane).  np.argsort pinned to kind="stable" (SURVEY 7.1): the +1s go to entries in
// descending value order, ties -> higher index first, cyclically.  Returns this lane's
// integer share (0 for non-members and for x == 0).
__device__ __forceinline__ int group_round_int_equal_sum(double x, const GroupIdx& G, int target)
{
    const int lane = lane_id();
    const bool member = (G.mask >> lane) & 1u;
    const bool nz = member && x != 0;
    const uint32_t nzb = __ballot_sync(FULL, nz) & G.mask; // the nonzero members of my group
    double v[NS];
    bool use[NS];
    double res = 0.;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        v[k] = __shfl_sync(FULL, x, G.src[k]);
        use[k] = k < G.n && ((nzb >> G.src[k]) & 1u);
        if (use[k]) res += v[k];
    }
    const double sum = 0. + res;
    const int cnt = __popc(nzb);
    int p = 0;
    if (nz) p = (int)floor(((double)target * x) / sum);
    const int adj = target - group_sum_int(p, G); // p is 0 on the zero members
    int rank = 0; // rank in descending order, ties -> higher index first
#pragma unroll
    for (int k = 0; k < NS; ++k)
        if (use[k] && (v[k] > x || (v[k] == x && G.src[k] > lane))) ++rank;
    if (nz && adj > rank) p += idiv_small(adj - rank - 1, cnt) + 1;
    return p;
}

// Exclusive prefix (ascending lane order) of an int over the members of G.
__device__ __forceinline__ int group_prefix_excl(int x, const GroupIdx& G)
{
    int res = 0;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        const int v = __shfl_sync(FULL, x, G.src[k]);
        if (k < G.rank) res += v;
    }
    return res;
}

// ---- NumPy pairwise float64 sum of a 135-element row held in shared memory ----------------
// n = 135 > 128: split at 64 | 71; each half: 8 interleaved accumulators combined as
// ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), then the 7-element tail of the second half added
// serially, halves added, then "0. +".  Lanes [g*8, g*8+8) reduce half `g&1` of row `g>>1`,
// so one warp produces two row sums at once: result 0 valid in lane 0, result 1 in lane 16.
// ues.py:108-116 (throughput) and np.mean(SE[0,u,:]) (ib_sched.py:111,147) depend on this order.
__device__ __noinline__ double np_sum135_pair(const double* rowA, const double* rowB)
{
    const int lane = lane_id();
    const int j = lane & 7;
    const int half = (lane >> 3) & 1;
    const double* row = (lane < 16) ? rowA : rowB;
    const double* p = row + half * 64 + j;
    double r = p[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) r += p[8 * k];
    r = r + __shfl_down_sync(FULL, r, 1);
    r = r + __shfl_down_sync(FULL, r, 2);
    r = r + __shfl_down_sync(FULL, r, 4);
    if (j == 0 && half == 1) {
#pragma unroll
        for (int k = 0; k < 7; ++k) r += row[128 + k];
    }
    const double other = __shfl_down_sync(FULL, r, 8);
    return 0. + (r + other); // valid in lanes 0 and 16
}

// Same reduction order, both sums of one SE row at once: lanes 0..15 sum the throughput
// terms ((BW/R) * sched[r]) * SE[r] (ues.py:108-116; sched is 1 inside the UE's RB run, else
// 0), lanes 16..31 sum SE[r] itself.  Results valid in lanes 0 and 16.
// sched[u, r] of the step: inside the UE's contiguous RB run, or bit r of a scheduling matrix given directly
__device__ __forceinline__ bool sched_at(const uint32_t* bits, int rb_s, int rb_c, int r)
{
    return bits ? ((bits[r >> 5] >> (r & 31)) & 1u) != 0 : (unsigned)(r - rb_s) < (unsigned)rb_c;
}
__device__ __noinline__ double np_sum135_thr_se(const double* se, int rb_s, int rb_e, double bw_rb, const uint32_t* bits = nullptr)
{
    const int lane = lane_id();
    const int j = lane & 7;
    const int base = ((lane >> 3) & 1) * 64 + j;
    const bool terms = lane < 16;
    auto elem = [&](int r) -> double {
        const double x = se[r];
        const double sched = sched_at(bits, rb_s, rb_e - rb_s, r) ? 1.0 : 0.0;
        return terms ? (bw_rb * sched) * x : x;
    };
    double r = elem(base);
#pragma unroll
    for (int k = 1; k < 8; ++k) r += elem(base + 8 * k);
    r = r + __shfl_down_sync(FULL, r, 1);
    r = r + __shfl_down_sync(FULL, r, 2);
    r = r + __shfl_down_sync(FULL, r, 4);
    if ((lane & 15) == 8) {
#pragma unroll
        for (int k = 0; k < 7; ++k) r += elem(128 + k);
    }
    const double other = __shfl_down_sync(FULL, r, 8);
    return 0. + (r + other);
}

__device__ __forceinline__ int scen_of_episode(const irs_config& c, int episode)
{
    const int s = c.assoc_mode == 0 ? episode % c.scen_mod : episode / c.scen_div;
    return s % c.n_scen;
}

// channel_mode 3: the SE frame env state (episode, step counter k) reads (channels/quadriga.py:38-87)
__device__ __forceinline__ const double* trace_frame(const StepArgs& A, int e, int episode, int k)
{
    const int f = (k < 0 ? 0 : k) % A.bank_steps; // a trace shorter than the episode wraps around
    int slot = episode - A.bank_first;             // the reference opens this episode's own file (quadriga.py:46-55, 78-87)
    if (A.bank_wrap) {
        slot %= A.bank_eps;
        if (slot < 0) slot += A.bank_eps;
    } else if (slot < 0 || slot >= A.bank_eps) {   // an episode the bank does not hold: flagged, never replaced silently
        if (lane_id() == 0) atomicOr(&A.st.err[e], ERR_TRACE_EPISODE);
        slot = 0;
    }
    return A.se_bank + ((size_t)slot * A.bank_steps + (size_t)f) * (U * R);
}

// per-UE buffer parameters: the UE's slice (associations/mult_slice.py:468-488 update_ues);
// unassociated UEs keep the env defaults (comm_env.py:184-198).
struct UeParams {
    int depth;
    int max_pkts;
    double pkt_bits;
};
__device__ __forceinline__ UeParams ue_params(const irs_scenario* sc, int s)
{
    UeParams p;
    if (s < 0) {
        p.depth = 101; p.max_pkts = 10240; p.pkt_bits = 65536.0;
    } else {
        p.depth = (int)__ldg(&sc->sl[s].buffer_latency) + 1;
        p.max_pkts = (int)__ldg(&sc->sl[s].buffer_size);
        p.pkt_bits = __ldg(&sc->sl[s].message_size);
    }
    return p;
}
