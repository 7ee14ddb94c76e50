// irs_ue.cuh - K12: one warp per (env, UE) row.
//
// Replaces, fused in one pass over the UE's buffer row:
//   channel / traffic draws      channels/mimic_quadriga.py:30-58, channels/fixed_se.py:28-41,
//                                traffics/mult_slice.py:15-34          (native mode; else injected)
//   UEs.get_pkt_throughputs      sixg_radio_mgmt/sixg_radio_mgmt/ues.py:71-116
//   pkt_incomings                ues.py:199
//   Buffer.receive_packets       buffer.py:53-80   (age shift, expiry drop, overflow clip)
//   Buffer.send_packets          buffer.py:82-104  (oldest-first drain)
//   get_buffer_occupancy / get_avg_delay   buffer.py:106-130
// and the per-UE part of the 10-frame observation window (agents/ib_sched.py:49-51,64).
//
// Buffer row = age histogram int32[depth]; index = age in steps.  Because the row's packet
// count is kept as state, everything but the mean delay is known before the row is touched:
//   * empty row            -> only slot 0 is written (4 bytes),
//   * everything older than the new arrivals is sent -> the row is zero-filled (write only),
//   * otherwise the row is streamed once: 128-bit loads, the one-slot age shift done in
//     registers (shuffle from the neighbouring lane), the oldest-first drain as a suffix scan
//     from the old end (skipped as soon as the scan has passed `thr` packets), 128-bit stores
//     to the same place.  Rows are walked from the old end in groups of 4 x 128 slots so that
//     the in-place shift never reads a slot it has already overwritten.
#pragma once
#include "irs_common.cuh"
#include "irs_rng.cuh"


#define UE_WARPS 8 /* warps (rows) per block */
#ifndef GC
#define GC 4       /* 128-slot chunks per group */
#endif

enum { ROW_EMPTY = 0, ROW_ZERO_FILL = 1, ROW_STREAM = 2 };

// General case: stream the row once.  `total` = packets in slots >= 1 after the shift,
// `thr_i` = packets that may leave (thr_i < total here), `a_left` = what stays in slot 0.
// `depth` is the row's live extent after the shift (min(live + 1, configured depth)): slots at
// and beyond it are zero and are neither read nor written.
// Returns sum(age * packets left) in the low 48 bits (age < 2^12, packets < 2^31) and the new
// live extent (index of the oldest non-empty slot + 1) above them, both reduced over the warp.
#define ROW_WSUM_BITS 48
#ifndef ROW_PREFETCH
#define ROW_PREFETCH 1 /* L2 prefetch distance in groups within a row (0 = off); the next row's first group too */
#endif
// The 16 128-byte lines of group `g` (GC x 128 slots) of a row, one per lane 0..15, into L2.
__device__ __forceinline__ void prefetch_row_group(const int* row, int depth, int g)
{
    const int i0 = g * (GC << 7) + 32 * lane_id();
    if (ROW_PREFETCH && lane_id() < 4 * GC && i0 < depth) asm volatile("prefetch.global.L2 [%0];" ::"l"(row + i0));
}
__device__ __forceinline__ int row_oldest_group(int depth) { return (((depth + 127) >> 7) + GC - 1) / GC - 1; }

__device__ __noinline__ long long buffer_row_stream(int* __restrict__ row, int depth, int thr_i, int a_left)
{
    const int lane = lane_id();
    const int nch = (depth + 127) >> 7;
    int sfx = 0;        // packets in slots older than the chunk being processed (warp-uniform)
    int hi = 0;         // this lane's oldest non-empty slot + 1 (the walk starts at the old end)
    long long wsum = 0; // sum age * packets left (this lane's slots)
    const int top = (nch + GC - 1) / GC - 1;
    for (int g = top; g >= 0; --g) {
        const int cb = g * GC;
        if (g == top) {
#pragma unroll
            for (int d = 1; d < ROW_PREFETCH; ++d)
                if (g >= d) prefetch_row_group(row, depth, g - d);
        }
        if (g >= ROW_PREFETCH) prefetch_row_group(row, depth, g - ROW_PREFETCH);
        int4 v[GC];
#pragma unroll
        for (int k = 0; k < GC; ++k) {
            const int i0 = ((cb + k) << 7) + 4 * lane;
            v[k] = (i0 < depth) ? *reinterpret_cast<const int4*>(row + i0) : make_int4(0, 0, 0, 0);
        }
        const int carry_in = cb > 0 ? row[(cb << 7) - 1] : 0;
#pragma unroll
        for (int k = GC - 1; k >= 0; --k) {
            const int c0 = (cb + k) << 7;
            if (c0 >= depth) continue; // warp-uniform
            const int i0 = c0 + 4 * lane;
            int n0 = __shfl_up_sync(FULL, v[k].w, 1);
            const int prev_last = (k > 0) ? __shfl_sync(FULL, v[k > 0 ? k - 1 : 0].w, 31) : carry_in;
            if (lane == 0) n0 = prev_last; // new[i] = old[i-1]
            int n1 = v[k].x, n2 = v[k].y, n3 = v[k].z;
            if (c0 + 128 > depth) { // tail chunk: nothing lives at index >= depth
                n0 = (i0 < depth) ? n0 : 0;
                n1 = (i0 + 1 < depth) ? n1 : 0;
                n2 = (i0 + 2 < depth) ? n2 : 0;
                n3 = (i0 + 3 < depth) ? n3 : 0;
            }
            if (sfx < thr_i) { // buffer.py:97-103: the oldest `thr` packets leave
                const int t = n0 + n1 + n2 + n3;
                int incl = t; // inclusive suffix sum over lanes (lane 31 holds the oldest slots)
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int y = __shfl_down_sync(FULL, incl, d);
                    if (lane + d < 32) incl += y;
                }
                const int s3 = sfx + incl - t; // packets strictly older than slot i0+3
                const int s2 = s3 + n3, s1 = s2 + n2, s0 = s1 + n1;
                sfx += __shfl_sync(FULL, incl, 0);
                n3 -= min(n3, max(0, thr_i - s3));
                n2 -= min(n2, max(0, thr_i - s2));
                n1 -= min(n1, max(0, thr_i - s1));
                n0 -= min(n0, max(0, thr_i - s0));
            }
            if (i0 == 0) n0 = a_left; // slot 0: the new arrivals (already net of what was sent)
            wsum += (long long)i0 * (n0 + n1 + n2 + n3) + (n1 + 2 * n2 + 3 * n3);
            if (hi == 0 && (n0 | n1 | n2 | n3) != 0) hi = i0 + (n3 ? 4 : (n2 ? 3 : (n1 ? 2 : 1)));
            if (i0 < depth) *reinterpret_cast<int4*>(row + i0) = make_int4(n0, n1, n2, n3);
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) wsum += __shfl_xor_sync(FULL, wsum, d);
    return wsum | ((long long)__reduce_max_sync(FULL, hi) << ROW_WSUM_BITS);
}


// Everything older than the new arrivals was sent: the row becomes [a_left, 0, 0, ...]
// (`depth` = the live extent before the shift; nothing beyond it needs clearing).
__device__ __noinline__ void buffer_row_zero_fill(int* __restrict__ row, int depth, int a_left)
{
    const int lane = lane_id();
    for (int i0 = 4 * lane; i0 < depth; i0 += 128)
        *reinterpret_cast<int4*>(row + i0) = make_int4(i0 == 0 ? a_left : 0, 0, 0, 0);
}

// SE row of (e, u) for `frame` into shared memory, by the calling warp (reset path and the
// injected / fixed channel modes of the step).
__device__ __noinline__ void channel_row(const StepArgs& A, int e, int u, bool associated, int frame,
                                            int episode, double* se)
{
    const int lane = lane_id();
    if (A.cfg.channel_mode == 0) {
        const double* src = A.se + ((size_t)e * U + u) * R;
        for (int r = lane; r < R; r += 32) se[r] = src[r];
    } else if (A.cfg.channel_mode == 3) { // frame 0 at reset, frame k at the step whose counter is k
        const double* src = trace_frame(A, episode, frame > 0 ? frame - 1 : 0) + (size_t)u * R;
        for (int r = lane; r < R; r += 32) se[r] = src[r];
    } else if (A.cfg.channel_mode == 1) {
        for (int r = lane; r < R; r += 32) se[r] = 2.0; // channels/fixed_se.py:26
    } else if (!associated) {
        for (int r = lane; r < R; r += 32) se[r] = 0.0;
    } else {
        native_se_row(A.cfg.seed, u, frame, episode, A.cfg.env_id_base + e, se);
    }
}

__global__ void __launch_bounds__(UE_WARPS * 32, 4) k_ue_step(const __grid_constant__ StepArgs A)
{
    __shared__ __align__(16) double sm[UE_WARPS][136];
    const DevState& st = A.st;
    const irs_config& cfg = A.cfg;
    const int lane = lane_id();
    const int wib = threadIdx.x >> 5;
    const unsigned gw = (unsigned)A.env_begin * U + blockIdx.x * UE_WARPS + wib; // row index e * 25 + u (host checks E * 25 < 2^31)
    if (gw >= (unsigned)A.env_end * U) return;
    const int e = gw / U, u = gw - e * U;
    const size_t eu = gw;

    // ---- prologue: every scalar the row needs, one lane each, issued before the generators --
    const int win_len = st.win_len[e];
    const int p = (st.win_head[e] + 1) % A.win_cap; // window slot this frame goes to (deque.appendleft)
    const int pre_i = lane < UI_N ? st.ue_i[eu * UI_N + lane] : 0;
    const double occ_old_l = lane == 0 ? st.ue_d[eu * UD_N + UD_OCC] : 0.0;
    int2 evict = make_int2(0, 0);
    if (lane == 0 && win_len == A.win_cap) evict = st.ring[(eu * W) + p];
    const short2 run = st.rb[eu];
    const irs_scenario* sc = st.bank + st.scen[e];
    const int s = sc->ue_slice[u];
    UeParams up = ue_params(sc, s);
    if (up.depth > st.row_stride) up.depth = st.row_stride;
    const int frame = st.step[e] + 1, episode = st.episode[e];
    const int tot_prev = __shfl_sync(FULL, pre_i, UI_TOT);

    // ---- channel + traffic of this frame --------------------------------------------------
    double* se = sm[wib];
    double traffic = 0.0, sum_t = 0.0, sum_se = 0.0;
    const bool native = cfg.channel_mode == 2 && cfg.traffic_mode == 1;
    const bool has_traffic = s >= 0 && (__ldg(&sc->sl[s >= 0 ? s : 0].flags) & IRS_F_REQ);
    if (!(native && s < 0)) { // an unassociated UE has SE 0 and no traffic in native mode
        if (native) {
            traffic = native_row(cfg.seed, st.cdf, u, has_traffic ? __ldg(&sc->sl[s].lam_idx) : 0, has_traffic,
                                 frame, episode, cfg.env_id_base + e, se);
        } else {
            channel_row(A, e, u, s >= 0, frame, episode, se);
            if (cfg.traffic_mode == 0) traffic = A.traffic[eu];
            else if (has_traffic)
                traffic = native_traffic(cfg.seed, st.cdf, u, sc->sl[s].lam_idx, frame, episode, cfg.env_id_base + e);
        }
        __syncwarp();
        const double sums = np_sum135_thr_se(se, run.x, run.x + run.y, A.bw_rb);
        sum_t = __shfl_sync(FULL, sums, 0);
        sum_se = __shfl_sync(FULL, sums, 16);
        if (A.raw.se_used)
            for (int r = lane; r < R; r += 32) A.raw.se_used[eu * R + r] = se[r];
    } else if (A.raw.se_used) {
        for (int r = lane; r < R; r += 32) A.raw.se_used[eu * R + r] = 0.0;
    }
    // three float64 divisions in one instruction stream (lanes 0, 1, 2)
    const double num = lane == 0 ? sum_t : (lane == 1 ? traffic : sum_se);
    const double den = lane == 2 ? (double)R : up.pkt_bits;
    const double quo = num / den;
    const double thr = floor(__shfl_sync(FULL, quo, 0)); // ues.py:108-116
    const double inc = floor(__shfl_sync(FULL, quo, 1)); // ues.py:199
    const double sebar = __shfl_sync(FULL, quo, 2);      // np.mean(SE[0,u,:])
    const int thr_i = (int)fmin(fmax(thr, 0.0), 2147483647.0);

    // ---- Buffer.receive_packets + send_packets (buffer.py:53-104) ---------------------------
    int* row = st.buf + eu * (size_t)st.row_stride;
    int expired = 0, path = ROW_EMPTY;
    const int live_prev = ROW_LIVE(__shfl_sync(FULL, pre_i, UI_PATH)); // slots [0, live_prev) may hold packets
    if (live_prev == up.depth) expired = row[up.depth - 1]; // buffer.py:60 dropped = buffer[-1]
    const int total = tot_prev - expired;             // occupancy after the age shift
    int dropped = expired, kept;
    if ((double)total + inc <= (double)up.max_pkts) { // buffer.py:62-80
        kept = (int)inc;
    } else {
        dropped += (int)(long long)(inc - (double)(up.max_pkts - total));
        kept = up.max_pkts - total;
    }
    const int sent_old = min(thr_i, total);           // oldest first (buffer.py:97-103)
    const int a_left = kept - min(kept, thr_i - sent_old);
    const int sent = sent_old + (kept - a_left);
    const int rem = total + kept - sent;
    long long wsum = 0;
    int live = a_left != 0, slots = live;
    if (tot_prev == 0) {                              // row is all zero: only slot 0 changes
        if (lane == 0 && a_left != 0) row[0] = a_left;
    } else if (sent_old == total) {                   // all older packets left: zero-fill
        slots = live_prev;
        buffer_row_zero_fill(row, slots, a_left);
        path = ROW_ZERO_FILL;
    } else {
        slots = min(live_prev + 1, up.depth);
        wsum = buffer_row_stream(row, slots, thr_i, a_left);
        live = (int)(wsum >> ROW_WSUM_BITS);
        wsum &= (1LL << ROW_WSUM_BITS) - 1;
        path = ROW_STREAM;
    }
    // occupancy and mean delay: two divisions in one stream (buffer.py:106-130)
    const double num2 = lane == 0 ? (double)rem : (double)wsum;
    const double den2 = lane == 0 ? (double)up.max_pkts : (double)rem;
    const double quo2 = num2 / den2;
    const double occ = __shfl_sync(FULL, quo2, 0);
    const double lat = rem != 0 ? __shfl_sync(FULL, quo2, 1) : 0.0;

    // ---- epilogue: window push + newest frame, one lane per field ---------------------------
    const double occ_old = __shfl_sync(FULL, occ_old_l, 0);
    const int ev_s = __shfl_sync(FULL, evict.x, 0), ev_d = __shfl_sync(FULL, evict.y, 0);
    const int sum_sent = __shfl_sync(FULL, pre_i, UI_SUM_SENT) + sent - ev_s;
    const int sum_drop = __shfl_sync(FULL, pre_i, UI_SUM_DROP) + dropped - ev_d;
    int out_i = rem;                                  // UI_TOT
    out_i = lane == UI_SUM_SENT ? sum_sent : out_i;
    out_i = lane == UI_SUM_DROP ? sum_drop : out_i;
    out_i = lane == UI_SENT ? sent : out_i;
    out_i = lane == UI_DROP ? dropped : out_i;
    out_i = lane == UI_INC ? (int)inc : out_i;
    out_i = lane == UI_OCC1Z ? (isclose0(occ_old) ? 1 : 0) : out_i; // win[1].occ of the new window
    out_i = lane == UI_PATH ? ROW_WORD(path, live, slots) : out_i;
    if (lane < UI_N) st.ue_i[eu * UI_N + lane] = out_i;
    double out_d = occ;                               // UD_OCC
    out_d = lane == UD_LAT ? lat : out_d;
    out_d = lane == UD_SEBAR ? sebar : out_d;
    out_d = lane == UD_THR ? thr : out_d;
    if (lane < UD_N) st.ue_d[eu * UD_N + lane] = out_d;
    if (lane == 0) {
        st.ring[eu * W + p] = make_int2(sent, dropped);
        if (A.raw.pkt_incoming) A.raw.pkt_incoming[eu] = inc;
        if (A.raw.pkt_throughputs) A.raw.pkt_throughputs[eu] = thr;
        if (A.raw.sent) A.raw.sent[eu] = sent;
        if (A.raw.dropped) A.raw.dropped[eu] = dropped;
        if (A.raw.occ) A.raw.occ[eu] = occ;
        if (A.raw.lat) A.raw.lat[eu] = lat;
        if (A.raw.traffic_used) A.raw.traffic_used[eu] = traffic;
    }
}
