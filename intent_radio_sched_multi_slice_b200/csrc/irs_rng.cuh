// irs_rng.cuh - counter-based generators of the native (no injected trace) mode.
//
// The reference draws everything from ONE sequential PCG64 stream per env
// (sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:631,666,673,679), which no parallel generator
// can reproduce.  Native mode therefore defines its own stateless generators,
// keyed so that a draw depends only on (seed, reset count, global env id, episode, frame, UE, index):
//   Philox4x32-7 (Salmon et al., SC'11; 7 rounds pass BigCrush), key = env_key(seed, reset count),
//   counter = (index | kind << 16 | ue << 24, frame, episode, global env id).
// The distributions are the reference's: SE ~ |N(mu_u, 1.5)| with mu_u ~ |N(10, 7.5)| per
// episode (channels/mimic_quadriga.py:37-56), traffic ~ Poisson(lambda_slice) * 1e6 bits
// (traffics/mult_slice.py:28-31).
// Normals: ONE per 16-bit half-word (8 per Philox quad) by table-driven inverse CDF - sign bit,
// 15-bit tail probability, 512-entry (exact quantile, slope) table indexed by the exponent and top
// mantissa bits of the tail count, one fused multiply-add (scripts/gen_normal_table.py: max
// quantile error 5e-5, variance 1.000007, |z| <= 4.32).  Every step is an IEEE binary32/64
// operation on shared constants (irs_normal_tab.h), so the plain-C oracle reproduces the bits.
#pragma once
#include "irs_common.cuh"
#include "irs_normal_tab.h"

enum { K_SE = 0, K_TRAFFIC = 1, K_MU = 2, K_POLICY = 3, K_SCEN = 4, K_PERM = 5 };

__device__ __forceinline__ uint4 philox4x32_7(uint4 c, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 7; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c.x, p1 = (uint64_t)0xCD9E8D57u * c.z;
        c = make_uint4((uint32_t)(p1 >> 32) ^ c.y ^ k0, (uint32_t)p1, (uint32_t)(p0 >> 32) ^ c.w ^ k1, (uint32_t)p0);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ uint4 rng_quad(uint64_t seed, int kind, int ue, int q, int frame, int episode, int gid)
{
    const uint4 c = make_uint4((uint32_t)q | ((uint32_t)kind << 16) | ((uint32_t)ue << 24), (uint32_t)frame,
                               (uint32_t)episode, (uint32_t)gid);
    return philox4x32_7(c, (uint32_t)seed, (uint32_t)(seed >> 32));
}

// enable_random_episodes (comm_env.py:219-226, 434-444): element `pos` of a pseudo-random permutation of [0, n),
// without a table: an 8-round balanced Feistel network over the next even power of two >= n, cycle-walked back into
// [0, n) (a bijection, so every episode of the range is played once per cycle, like np_random.permutation + cycle).
// Round keys: one Philox quad of (seed, global env id, permutation number).
__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__host__ __device__ __forceinline__ uint32_t feistel_perm(uint32_t i, uint32_t n, uint32_t k0, uint32_t k1, uint32_t k2, uint32_t k3)
{
    if (n <= 1) return 0;
    int bits = 1;
    while ((1u << bits) < n) ++bits;
    bits += bits & 1;
    const int half = bits >> 1;
    const uint32_t mask = (1u << half) - 1;
    uint32_t x = i;
    do {
        uint32_t l = x >> half, r = x & mask, t;
        const uint32_t k[4] = {k0, k1, k2, k3};
#pragma unroll
        for (int j = 0; j < 8; ++j) { t = l ^ (mix32(r ^ k[j & 3] ^ (j * 0x9E3779B9u)) & mask); l = r; r = t; }
        x = (l << half) | r;
    } while (x >= n);
    return x;
}
__device__ __noinline__ int shuffled_episode(uint64_t seed, int gid, int gen, int base, int n, int pos)
{
    const uint4 k = rng_quad(seed, K_PERM, 0, 0, gen, 0, gid);
    return base + (int)feistel_perm((uint32_t)pos, (uint32_t)n, k.x, k.y, k.z, k.w);
}

// ---- 16-bit inverse-CDF normal -------------------------------------------------------------------
__device__ __align__(16) const uint32_t g_normal_tab[2 * IRS_NORMAL_TAB_N] = { IRS_NORMAL_TAB_WORDS };

// m = 15-bit magnitude word, signbit = 0 or 0x80000000; tab = the (c0, c1) table (shared or global memory)
__device__ __forceinline__ float normal_from(uint32_t m, uint32_t signbit, const float2* __restrict__ tab)
{
    const uint32_t bits = __float_as_uint((float)(int)(32768u - m)); // tail count t in [1, 32768], exact in binary32
    const float2 c = tab[(bits >> (23 - IRS_NORMAL_SUB_BITS)) - (127u << IRS_NORMAL_SUB_BITS)]; // octave, top mantissa bits
    const float g = __uint_as_float(((bits & ((1u << (23 - IRS_NORMAL_SUB_BITS)) - 1u)) << IRS_NORMAL_SUB_BITS) | 0x3F800000u); // position inside the segment, [1, 2)
    return __uint_as_float(__float_as_uint(__fmaf_rn(c.y, g, c.x)) ^ signbit);
}
// half-word k (0 = low) of word w
__device__ __forceinline__ float normal_lo(uint32_t w, const float2* __restrict__ tab)
{
    return normal_from(w & 0x7FFFu, (w << 16) & 0x80000000u, tab);
}
__device__ __forceinline__ float normal_hi(uint32_t w, const float2* __restrict__ tab)
{
    return normal_from((w >> 16) & 0x7FFFu, w & 0x80000000u, tab);
}
__device__ __forceinline__ const float2* normal_tab_global() { return reinterpret_cast<const float2*>(g_normal_tab); }

// Where half-word k (0..7) of SE quad q lands in the 135-RB row.  Quad q < 16 holds exactly the 8 elements
// one accumulator of NumPy's pairwise sum visits (row half q >> 3, accumulator q & 7: elements
// half * 64 + (q & 7) + 8k, k = 0..7 in order), so a lane that generates the quad can add them up in
// registers in NumPy's order; quad 16 is the serial tail 128..134 (its last half-word is unused).
__host__ __device__ __forceinline__ int se_index(int q, int k)
{
    return q < 16 ? ((q >> 3) << 6) + (q & 7) + 8 * k : 128 + k;
}

// channels/mimic_quadriga.py:37-41
__device__ __forceinline__ double native_mu(uint64_t seed, int ue, int episode, int gid)
{
    const uint4 w = rng_quad(seed, K_MU, ue, 0, 0, episode, gid);
    return fabs(__fma_rn(7.5, (double)normal_lo(w.x, normal_tab_global()), 10.0));
}

// channels/mimic_quadriga.py:49-56: one UE row of SE into `se` (shared memory, 135 doubles),
// produced cooperatively by the calling warp (reset path; the step generates in registers, irs_fused.cuh)
__device__ __noinline__ void native_se_row(uint64_t seed, int ue, int frame, int episode, int gid, double* se)
{
    const double mu = native_mu(seed, ue, episode, gid);
    const float2* tab = normal_tab_global();
    const int q = lane_id();
    if (q <= 16) {
        const uint4 w = rng_quad(seed, K_SE, ue, q, frame, episode, gid);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float z = (k & 1) ? normal_hi(ww[k >> 1], tab) : normal_lo(ww[k >> 1], tab);
            if (se_index(q, k) < R) se[se_index(q, k)] = fabs(__fma_rn(1.5, (double)z, mu));
        }
    }
}

// traffics/mult_slice.py:28-31 by table inversion: k = #{j : cdf[j] <= u}, u a 53-bit uniform.
// The table is non-decreasing, so the count is an upper bound search.
__device__ __noinline__ double native_traffic(uint64_t seed, const double* cdf_all, int ue, int lam_idx,
                                                 int frame, int episode, int gid)
{
    const uint4 w = rng_quad(seed, K_TRAFFIC, ue, 0, frame, episode, gid);
    const double u = ((double)(w.x >> 5) * 67108864.0 + (double)(w.y >> 6)) * (1.0 / 9007199254740992.0);
    const double* cdf = cdf_all + (size_t)lam_idx * IRS_POISSON_K;
    int lo = 0, hi = IRS_POISSON_K; // first index with cdf[j] > u
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(&cdf[mid]) <= u) lo = mid + 1; else hi = mid;
    }
    return (double)lo * 1e6;
}
