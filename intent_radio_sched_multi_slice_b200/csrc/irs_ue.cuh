// irs_ue.cuh - buffer rows as rings, and the SE row of one UE for the reset / slow paths.
//
// Replaces, per UE (the callers are in irs_fused.cuh and irs_env.cuh):
//   Buffer.receive_packets       buffer.py:53-80   (age shift, expiry drop, overflow clip)
//   Buffer.send_packets          buffer.py:82-104  (oldest-first drain)
//   get_buffer_occupancy / get_avg_delay   buffer.py:106-130
//
// The reference keeps an age histogram int[depth] per UE and rolls it by one slot every step
// (np.roll, buffer.py:61).  Here the row is a RING: the packets of age a live in slot
// (k - a) mod depth, k = the env's step counter, so ageing is the counter moving on and a step
// touches only
//   * slot k mod depth: it holds the packets that reach age `depth` now (expired, buffer.py:60)
//     and receives the new arrivals,
//   * the oldest slots, as far as the drain of `thr` packets reaches (buffer.py:97-103).
// The packet count, sum(age * packets) and the live extent (every age at or beyond it is empty)
// are kept per UE, so occupancy and mean delay (buffer.py:106-130) need no pass over the row:
//   shift:  wsum += total - depth * expired        (every packet one step older, the expired ones gone)
//   drain:  wsum -= age * taken                    for every slot the drain takes packets from.
#pragma once
#include "irs_common.cuh"
#include "irs_rng.cuh"

enum { ROW_NONE = 0, ROW_CLEAR = 1, ROW_LANE = 2, ROW_WARP = 3 };
#define RING_LOOK 4 /* oldest slots a UE lane loads ahead of the closed-form arithmetic */

__device__ __forceinline__ int ring_slot(int head, int age, int depth) // 0 <= age <= depth
{
    const int p = head - age;
    return p < 0 ? p + depth : p;
}

// Every packet older than the new arrivals left: ages [1, a_old] become empty.  By the whole warp.
__device__ __noinline__ void ring_clear_warp(int* __restrict__ row, int depth, int head, int a_old)
{
    for (int a = 1 + lane_id(); a <= a_old; a += 32) row[ring_slot(head, a, depth)] = 0;
}

// Oldest-first drain of `want` packets from ages a_start, a_start - 1, ... 1 by the whole warp, 32 slots per
// pass (buffer.py:97-103).  Returns sum(age * packets taken); `live` = the new live extent, `acc` += slot accesses.
__device__ __noinline__ long long ring_drain_warp(int* __restrict__ row, int depth, int head, int a_start, int want,
                                                 int& live, int& acc)
{
    const int lane = lane_id();
    long long wdec = 0;
    live = 1;
    for (int a0 = a_start; a0 >= 1; a0 -= 32) {
        const int a = a0 - lane;
        const bool valid = a >= 1;
        int* slot = row + ring_slot(head, valid ? a : 0, depth);
        const int n = valid ? *slot : 0;
        int incl = n; // packets in this and older slots of the pass (lane 0 holds the oldest)
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(FULL, incl, d);
            if (lane >= d) incl += y;
        }
        const int take = max(0, min(n, want - (incl - n)));
        if (take) { *slot = n - take; wdec += (long long)a * take; }
        acc += __popc(__ballot_sync(FULL, valid)) + __popc(__ballot_sync(FULL, take != 0));
        const uint32_t surv = __ballot_sync(FULL, n - take > 0);
        if (surv) { live = a0 - (__ffs(surv) - 1) + 1; break; }
        want -= min(want, __shfl_sync(FULL, incl, 31));
        live = max(a0 - 32, 0) + 1; // everything at or above age a0 - 31 is empty now
        if (want == 0) break;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) wdec += __shfl_xor_sync(FULL, wdec, d);
    return wdec;
}

// SE row of (e, u) for `frame` into shared memory, by the calling warp (reset path and the
// injected / fixed channel modes of the step).
__device__ __noinline__ void channel_row(const StepArgs& A, uint64_t key, int e, int u, bool associated, int frame,
                                            int episode, double* se)
{
    const int lane = lane_id();
    if (A.cfg.channel_mode == 0) {
        const double* src = A.se + ((size_t)e * U + u) * R;
        for (int r = lane; r < R; r += 32) se[r] = src[r];
    } else if (A.cfg.channel_mode == 3) { // frame 0 at reset, frame k at the step whose counter is k
        const double* src = trace_frame(A, e, episode, frame > 0 ? frame - 1 : 0) + (size_t)u * R;
        for (int r = lane; r < R; r += 32) se[r] = src[r];
    } else if (A.cfg.channel_mode == 1) {
        for (int r = lane; r < R; r += 32) se[r] = 2.0; // channels/fixed_se.py:26
    } else if (!associated) {
        for (int r = lane; r < R; r += 32) se[r] = 0.0;
    } else {
        native_se_row(key, u, frame, episode, A.cfg.env_id_base + e, se);
    }
}
