// irs_rng.cuh - counter-based generators of the native (no injected trace) mode.
//
// The reference draws everything from ONE sequential PCG64 stream per env
// (sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:631,666,673,679), which no parallel generator
// can reproduce.  Native mode therefore defines its own stateless generators,
// keyed so that a draw depends only on (seed, global env id, episode, frame, UE, index):
//   Philox4x32-10 (Salmon et al., SC'11), key = seed, counter =
//   (index | kind << 16 | ue << 24, frame, episode, global env id).
// The distributions are the reference's: SE ~ |N(mu_u, 1.5)| with mu_u ~ |N(10, 7.5)| per
// episode (channels/mimic_quadriga.py:37-56), traffic ~ Poisson(lambda_slice) * 1e6 bits
// (traffics/mult_slice.py:28-31).  Every arithmetic step below is written with explicit
// roundings so that a plain-C restatement reproduces the same bits (the test oracle does).
#pragma once
#include "irs_common.cuh"

enum { K_SE = 0, K_TRAFFIC = 1, K_MU = 2, K_POLICY = 3, K_SCEN = 4, K_PERM = 5 };

__device__ __noinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
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
    return philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
}

// enable_random_episodes (comm_env.py:219-226, 434-444): element `pos` of a pseudo-random permutation of [0, n),
// without a table: a 4-round balanced Feistel network over the next even power of two >= n, cycle-walked back into
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
        t = l ^ (mix32(r ^ k0) & mask); l = r; r = t;
        t = l ^ (mix32(r ^ k1) & mask); l = r; r = t;
        t = l ^ (mix32(r ^ k2) & mask); l = r; r = t;
        t = l ^ (mix32(r ^ k3) & mask); l = r; r = t;
        x = (l << half) | r;
    } while (x >= n);
    return x;
}
__device__ __noinline__ int shuffled_episode(uint64_t seed, int gid, int gen, int base, int n, int pos)
{
    const uint4 k = rng_quad(seed, K_PERM, 0, 0, gen, 0, gid);
    return base + (int)feistel_perm((uint32_t)pos, (uint32_t)n, k.x, k.y, k.z, k.w);
}

// Where word w of SE quad q lands in the 135-RB row.  Quads 0..31 are laid out so that the two
// quads 2m, 2m+1 hold exactly the 8 elements one accumulator of NumPy's pairwise sum visits
// (row half m >> 3, accumulator m & 7: elements half * 64 + (m & 7) + 8k, k = 0..7 in order), so a
// lane that generates both quads can add them up in registers in NumPy's order; quads 32, 33
// are the serial tail 128..134 (word 3 of quad 33 is unused).
__host__ __device__ __forceinline__ int se_index(int q, int w)
{
    return q < 32 ? ((q >> 4) << 6) + ((q >> 1) & 7) + 8 * (4 * (q & 1) + w) : 4 * q + w;
}

// Two standard normals from two 32-bit words: Box-Muller in binary32 with a fixed polynomial
// log (atanh series on the mantissa) and octant-reduced sin/cos polynomials; no libm, no
// fast-math intrinsics, so the result is a pure function of IEEE add/mul/fma/div/sqrt.
__device__ __noinline__ float2 normal_pair2(uint32_t wa, uint32_t wb)
{
    const float u1 = __fmul_rn((float)((wa >> 8) + 1u), 5.9604644775390625e-08f); // (0,1]
    uint32_t bits = __float_as_uint(u1);
    int e = (int)(bits >> 23) - 127;
    float m = __uint_as_float((bits & 0x007FFFFFu) | 0x3F800000u);
    if (m > 1.41421356f) { m = __fmul_rn(m, 0.5f); e += 1; }
    const float s = __fdiv_rn(__fsub_rn(m, 1.0f), __fadd_rn(m, 1.0f));
    const float s2 = __fmul_rn(s, s);
    float p = __fmaf_rn(s2, 0.111111111f, 0.142857143f);
    p = __fmaf_rn(s2, p, 0.2f);
    p = __fmaf_rn(s2, p, 0.333333333f);
    p = __fmaf_rn(s2, p, 1.0f);
    const float lnm = __fmul_rn(__fmul_rn(2.0f, s), p);
    const float lnu = __fmaf_rn((float)e, 0.693147181f, lnm);
    const float r = __fsqrt_rn(__fmul_rn(-2.0f, lnu));
    const uint32_t oct = wb >> 29;
    const uint32_t t = (wb >> 6) & 0x7FFFFFu;
    const float f = __fmul_rn((float)(2u * t + 1u), 5.9604644775390625e-08f);
    const float phi = __fmul_rn(f, 0.785398163f);
    const float x2 = __fmul_rn(phi, phi);
    float sp = __fmaf_rn(x2, -1.98412698e-4f, 8.33333333e-3f);
    sp = __fmaf_rn(x2, sp, -0.166666667f);
    sp = __fmaf_rn(x2, sp, 1.0f);
    const float sn = __fmul_rn(phi, sp);
    float cp = __fmaf_rn(x2, 2.48015873e-5f, -1.38888889e-3f);
    cp = __fmaf_rn(x2, cp, 4.16666667e-2f);
    cp = __fmaf_rn(x2, cp, -0.5f);
    const float cs = __fmaf_rn(x2, cp, 1.0f);
    const bool swap = (oct == 0 || oct == 3 || oct == 4 || oct == 7);
    float a = swap ? cs : sn;
    float b = swap ? sn : cs;
    if (oct >= 2 && oct <= 5) a = -a;
    if (oct >= 4) b = -b;
    return make_float2(__fmul_rn(r, a), __fmul_rn(r, b));
}
__device__ __forceinline__ void normal_pair(uint32_t wa, uint32_t wb, float& z0, float& z1)
{
    const float2 z = normal_pair2(wa, wb);
    z0 = z.x;
    z1 = z.y;
}

// channels/mimic_quadriga.py:37-41
__device__ __forceinline__ double native_mu(uint64_t seed, int ue, int episode, int gid)
{
    const uint4 w = rng_quad(seed, K_MU, ue, 0, 0, episode, gid);
    float z0, z1;
    normal_pair(w.x, w.y, z0, z1);
    return fabs(__fma_rn(7.5, (double)z0, 10.0));
}

// channels/mimic_quadriga.py:49-56: one UE row of SE into `se` (shared memory, 135 doubles),
// produced cooperatively by the calling warp (quad q -> RBs 4q..4q+3).
__device__ __noinline__ void native_se_row(uint64_t seed, int ue, int frame, int episode, int gid, double* se)
{
    const double mu = native_mu(seed, ue, episode, gid);
    for (int q = lane_id(); q < (R + 3) / 4; q += 32) {
        const uint4 w = rng_quad(seed, K_SE, ue, q, frame, episode, gid);
        float z[4];
        normal_pair(w.x, w.y, z[0], z[1]);
        normal_pair(w.z, w.w, z[2], z[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (se_index(q, j) < R) se[se_index(q, j)] = fabs(__fma_rn(1.5, (double)z[j], mu));
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

// One UE row of a step in native mode, by the whole warp, in two generator passes:
//   pass A  lane 0,1: SE quads 32,33 (RBs 128..134); lane 2: the UE's per-episode mean;
//           lane 3: the Poisson arrival          (same counters and values as the functions above)
//   pass B  lane l: SE quad l (RBs 4l..4l+3)
// Returns the offered traffic in bits (0 when the UE's slice has no requirements).
__device__ __noinline__ double native_row(uint64_t seed, const double* cdf_all, int ue, int lam_idx,
                                             bool has_traffic, int frame, int episode, int gid, double* se)
{
    const int lane = lane_id();
    const int kindA = lane == 2 ? K_MU : (lane == 3 ? K_TRAFFIC : K_SE);
    const int qA = lane == 0 ? 32 : (lane == 1 ? 33 : 0);
    const uint4 wa = rng_quad(seed, kindA, ue, qA, lane == 2 ? 0 : frame, episode, gid);
    float z[4];
    normal_pair(wa.x, wa.y, z[0], z[1]);
    normal_pair(wa.z, wa.w, z[2], z[3]);
    const double mu = __shfl_sync(FULL, fabs(__fma_rn(7.5, (double)z[0], 10.0)), 2);
    if (lane < 2) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (4 * qA + j < R) se[4 * qA + j] = fabs(__fma_rn(1.5, (double)z[j], mu));
    }
    // Poisson by table inversion (lane 3's words; the other lanes run along)
    const double u = ((double)(wa.x >> 5) * 67108864.0 + (double)(wa.y >> 6)) * (1.0 / 9007199254740992.0);
    const double* cdf = cdf_all + (size_t)lam_idx * IRS_POISSON_K;
    int lo = 0;
    if (has_traffic) {
#pragma unroll
        for (int step = IRS_POISSON_K / 2; step > 0; step >>= 1) // branch-free upper bound
            if (__ldg(&cdf[lo + step - 1]) <= u) lo += step;
        if (__ldg(&cdf[lo]) <= u) lo += 1; // lo in [0, 256]
    }
    const double traffic = __shfl_sync(FULL, (double)lo * 1e6, 3);
    const uint4 wb = rng_quad(seed, K_SE, ue, lane, frame, episode, gid);
    normal_pair(wb.x, wb.y, z[0], z[1]);
    normal_pair(wb.z, wb.w, z[2], z[3]);
#pragma unroll
    for (int j = 0; j < 4; ++j) se[se_index(lane, j)] = fabs(__fma_rn(1.5, (double)z[j], mu));
    return traffic;
}
