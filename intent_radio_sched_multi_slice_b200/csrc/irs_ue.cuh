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
// Buffer row = age histogram int32[depth]; index = age in steps.  The row is streamed once:
// 128-bit loads, the one-slot age shift is done in registers (shuffle from the neighbouring
// lane), the drain is a suffix scan from the oldest end, 128-bit stores to the same place.
// Rows are walked from the old end in groups of 4 x 128 slots so that the in-place shift
// never reads a slot it has already overwritten.
#pragma once
#include "irs_common.cuh"
#include "irs_rng.cuh"

#define UE_WARPS 8 /* warps (rows) per block */
#define GC 4       /* 128-slot chunks per group */

struct RowResult {
    int sent, dropped, rem;
    long long wsum;
};

// The whole warp works on one row.  thr_i: packets that can be sent; inc: packets arriving.
__device__ __forceinline__ RowResult buffer_row_step(int* __restrict__ row, int depth, int max_pkts, int thr_i,
                                                     double inc)
{
    const int lane = lane_id();
    const int expired = row[depth - 1]; // buffer.py:60 dropped = buffer[-1]
    const int nch = (depth + 127) >> 7;
    int sfx = 0;        // packets in slots older than the ones being processed (after the shift)
    int rem = 0;        // packets left in this lane's slots
    long long wsum = 0; // sum age * packets left
    int arrived_kept = 0;
    int4 first = make_int4(0, 0, 0, 0);
    for (int g = (nch + GC - 1) / GC - 1; g >= 0; --g) {
        const int cb = g * GC;
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
            int pw = __shfl_up_sync(FULL, v[k].w, 1);
            const int prev_last = (k > 0) ? __shfl_sync(FULL, v[k > 0 ? k - 1 : 0].w, 31) : carry_in;
            if (lane == 0) pw = prev_last;
            // shifted row: new[i] = old[i-1]; new[0] is filled in after the totals are known
            int n0 = (i0 < depth && i0 != 0) ? pw : 0;
            int n1 = (i0 + 1 < depth) ? v[k].x : 0;
            int n2 = (i0 + 2 < depth) ? v[k].y : 0;
            int n3 = (i0 + 3 < depth) ? v[k].z : 0;
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
            // buffer.py:97-103: the oldest `thr` packets leave
            n3 -= min(n3, max(0, thr_i - s3));
            n2 -= min(n2, max(0, thr_i - s2));
            n1 -= min(n1, max(0, thr_i - s1));
            n0 -= min(n0, max(0, thr_i - s0));
            rem += n0 + n1 + n2 + n3;
            wsum += (long long)n0 * i0 + (long long)n1 * (i0 + 1) + (long long)n2 * (i0 + 2) + (long long)n3 * (i0 + 3);
            const int4 out = make_int4(n0, n1, n2, n3);
            if (c0 == 0) first = out; // slot 0 still open
            else if (i0 < depth) *reinterpret_cast<int4*>(row + i0) = out;
        }
    }
    // buffer.py:62-80: sfx is now the occupancy after the shift
    const int total = sfx;
    int dropped = expired;
    if ((double)total + inc <= (double)max_pkts) {
        arrived_kept = (int)inc;
    } else {
        dropped += (int)(long long)(inc - (double)(max_pkts - total));
        arrived_kept = max_pkts - total;
    }
    const int a_left = arrived_kept - min(arrived_kept, max(0, thr_i - total));
    if (lane == 0) { first.x = a_left; rem += a_left; }
    if (4 * lane < depth) *reinterpret_cast<int4*>(row + 4 * lane) = first;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        rem += __shfl_xor_sync(FULL, rem, d);
        wsum += __shfl_xor_sync(FULL, wsum, d);
    }
    RowResult r;
    r.rem = rem;
    r.wsum = wsum;
    r.sent = total + arrived_kept - rem; // buffer.py:85,104 sum before - sum after
    r.dropped = dropped;
    return r;
}

// SE row of (e, u) for `frame` into shared memory, by the calling warp.
__device__ __forceinline__ void channel_row(const StepArgs& A, int e, int u, bool associated, int frame,
                                            int episode, double* se)
{
    const int lane = lane_id();
    if (A.cfg.channel_mode == 0) {
        const double* src = A.se + ((size_t)e * U + u) * R;
        for (int r = lane; r < R; r += 32) se[r] = src[r];
    } else if (A.cfg.channel_mode == 1) {
        for (int r = lane; r < R; r += 32) se[r] = 2.0; // channels/fixed_se.py:26
    } else if (!associated) {
        for (int r = lane; r < R; r += 32) se[r] = 0.0;
    } else {
        native_se_row(A.cfg.seed, u, frame, episode, A.cfg.env_id_base + e, se);
    }
}

__global__ void __launch_bounds__(UE_WARPS * 32) k_ue_step(const __grid_constant__ StepArgs A)
{
    __shared__ double sm[UE_WARPS][2][136];
    const DevState& st = A.st;
    const irs_config& cfg = A.cfg;
    const int lane = lane_id();
    const int wib = threadIdx.x >> 5;
    const long long gw = (long long)blockIdx.x * UE_WARPS + wib;
    if (gw >= (long long)cfg.num_envs * U) return;
    const int e = (int)(gw / U), u = (int)(gw % U);
    const size_t eu = (size_t)gw;
    const irs_scenario* sc = st.bank + st.scen[e];
    const int s = sc->ue_slice[u];
    UeParams up = ue_params(sc, s);
    if (up.depth > st.row_stride) up.depth = st.row_stride;
    const int rb_s = st.rb_start[eu], rb_c = st.rb_count[eu];
    const int frame = st.step[e] + 1, episode = st.episode[e];

    double* se = sm[wib][0];
    double* tm = sm[wib][1];
    channel_row(A, e, u, s >= 0, frame, episode, se);
    __syncwarp();
    const double bw_rb = cfg.bandwidth / (double)R; // ues.py:110
    for (int r = lane; r < R; r += 32) {
        const double sched = (r >= rb_s && r < rb_s + rb_c) ? 1.0 : 0.0;
        tm[r] = (bw_rb * sched) * se[r];
        if (A.raw.se_used) A.raw.se_used[eu * R + r] = se[r];
    }
    __syncwarp();
    const double sums = np_sum135_pair(tm, se);
    const double sum_t = __shfl_sync(FULL, sums, 0), sum_se = __shfl_sync(FULL, sums, 16);
    const double thr = floor(sum_t / up.pkt_bits); // ues.py:108-116
    const double sebar = sum_se / (double)R;       // np.mean(SE[0,u,:])
    double traffic;
    if (cfg.traffic_mode == 0) traffic = A.traffic[eu];
    else traffic = (s >= 0 && (sc->sl[s].flags & IRS_F_REQ))
                       ? native_traffic(cfg.seed, st.cdf, u, sc->sl[s].lam_idx, frame, episode, cfg.env_id_base + e)
                       : 0.0;
    const double inc = floor(traffic / up.pkt_bits); // ues.py:199
    const int thr_i = (int)fmin(fmax(thr, 0.0), 2147483647.0);

    RowResult rr;
    const int tot_prev = st.tot[eu];
    if (tot_prev == 0 && inc == 0.0) { // empty row stays empty: nothing to stream
        rr.sent = 0; rr.dropped = 0; rr.rem = 0; rr.wsum = 0;
    } else {
        rr = buffer_row_step(st.buf + eu * (size_t)st.row_stride, up.depth, up.max_pkts, thr_i, inc);
    }
    if (lane == 0) {
        const double occ = (double)rr.rem / (double)up.max_pkts;               // buffer.py:106-114
        const double lat = rr.rem != 0 ? (double)rr.wsum / (double)rr.rem : 0.0; // buffer.py:116-130
        // window push (deque.appendleft, maxlen 10): slot after the env's current head
        const int win_len = st.win_len[e];
        const int p = (st.win_head[e] + 1) % W;
        const size_t ri = ((size_t)e * W + p) * U + u;
        int ss = st.sum_sent[eu] + rr.sent, sd = st.sum_drop[eu] + rr.dropped;
        if (win_len == W) { ss -= st.ring_sent[ri]; sd -= st.ring_drop[ri]; }
        st.ring_sent[ri] = rr.sent;
        st.ring_drop[ri] = rr.dropped;
        st.sum_sent[eu] = ss;
        st.sum_drop[eu] = sd;
        st.occ1z[eu] = isclose0(st.occ[eu]) ? 1 : 0; // win[1].occ of the new window
        st.occ[eu] = occ;
        st.lat[eu] = lat;
        st.sebar[eu] = sebar;
        st.thr[eu] = thr;
        st.tot[eu] = rr.rem;
        st.last_sent[eu] = rr.sent;
        st.last_drop[eu] = rr.dropped;
        st.last_inc[eu] = (int)inc;
        if (A.raw.pkt_incoming) A.raw.pkt_incoming[eu] = inc;
        if (A.raw.pkt_throughputs) A.raw.pkt_throughputs[eu] = thr;
        if (A.raw.sent) A.raw.sent[eu] = rr.sent;
        if (A.raw.dropped) A.raw.dropped[eu] = rr.dropped;
        if (A.raw.occ) A.raw.occ[eu] = occ;
        if (A.raw.lat) A.raw.lat[eu] = lat;
        if (A.raw.traffic_used) A.raw.traffic_used[eu] = traffic;
    }
}
