// sb_rng.cuh -- K6: counter-based Philox4x32-10 and the samplers built on it (uniform, normal, Gamma,
// Beta via Gammas).  Replaces the reference's two interleaved global generators (stdlib `random` and
// `np.random`, quirk Q13) -- trajectories cannot match bit for bit, distributions do.
//
// Every random number is a pure function of (seed, GLOBAL chain id, step, purpose, index, attempt), so
//   * all threads of a CTA can draw the same block-uniform decision without communication,
//   * a chain's trajectory does not depend on how chains are sharded over GPUs or on thread scheduling.
#pragma once
#include <stdint.h>

struct RngKey { uint32_t k0, k1; };

__host__ __device__ __forceinline__ void philox_round(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3,
                                                       uint32_t k0, uint32_t k1) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
}

// Philox4x32-10 (Salmon et al., SC'11).
__host__ __device__ __forceinline__ void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                     uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        philox_round(c0, c1, c2, c3, k0, k1);
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// purposes (c2 high byte); the low 24 bits of c2 are the rejection-loop attempt / sub-draw counter
enum {
    RP_OP = 1, RP_ZONE = 2, RP_CONNECTED = 3, RP_SITE_NEW = 4, RP_SITE_REMOVED = 5, RP_ACCEPT = 6,
    RP_SOURCE = 7, RP_SUBSET = 8, RP_GAMMA = 9, RP_FIXED = 10, RP_FEATURE = 11, RP_PAIR = 12, RP_FAMILY = 13,
    RP_BETA_A = 14, RP_BETA_B = 15, RP_PROPOSAL = 16
};

struct Draw {      // two 53-bit uniforms in (0,1) from one Philox block
    double u0, u1;
};

__device__ __forceinline__ double u53(uint32_t hi, uint32_t lo) {
    const uint64_t v = ((uint64_t)hi << 21) ^ (uint64_t)(lo >> 11);     // 53 random bits
    return ((double)v + 0.5) * (1.0 / 9007199254740992.0);
}

__device__ __forceinline__ Draw rng_draw(RngKey key, uint64_t step, uint32_t purpose, uint32_t sub, uint32_t index) {
    uint32_t o[4];
    philox4x32((uint32_t)step, (uint32_t)(step >> 32), (purpose << 24) | (sub & 0xffffffu), index, key.k0, key.k1, o);
    Draw d;
    d.u0 = u53(o[0], o[1]);
    d.u1 = u53(o[2], o[3]);
    return d;
}

__device__ __forceinline__ double rng_uniform(RngKey key, uint64_t step, uint32_t purpose, uint32_t index) {
    return rng_draw(key, step, purpose, 0, index).u0;
}

// uniform integer in [0, n)
__device__ __forceinline__ int rng_int(RngKey key, uint64_t step, uint32_t purpose, uint32_t index, int n) {
    const int v = (int)(rng_uniform(key, step, purpose, index) * (double)n);
    return v < n ? v : n - 1;
}

// Gamma(alpha, 1), Marsaglia & Tsang (2000).  alpha < 1 uses Gamma(alpha+1) * U^(1/alpha).
// Each attempt consumes one Philox block (normal by Box-Muller from u0,u1 of block 2a; accept test from
// block 2a+1).
__device__ double rng_gamma(RngKey key, uint64_t step, uint32_t purpose, uint32_t index, double alpha) {
    double boost = 1.0;
    uint32_t sub = 0;
    if (alpha < 1.0) {
        const Draw b = rng_draw(key, step, purpose, 0xfffffeu, index);
        boost = pow(b.u0, 1.0 / alpha);
        alpha += 1.0;
    }
    const double d = alpha - 1.0 / 3.0;
    const double c = 1.0 / sqrt(9.0 * d);
    for (int attempt = 0; attempt < 64; ++attempt) {
        const Draw n = rng_draw(key, step, purpose, sub++, index);
        const Draw a = rng_draw(key, step, purpose, sub++, index);
        double sn, cs;
        sincospi(2.0 * n.u1, &sn, &cs);
        const double r = sqrt(-2.0 * log(n.u0));
        // both Box-Muller normals are tried before a new block is drawn
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const double x = r * (h == 0 ? cs : sn);
            const double u = h == 0 ? a.u0 : a.u1;
            double v = 1.0 + c * x;
            if (v <= 0.0) continue;
            v = v * v * v;
            const double x2 = x * x;
            if (u < 1.0 - 0.0331 * x2 * x2) return boost * d * v;
            if (log(u) < 0.5 * x2 + d * (1.0 - v + log(v))) return boost * d * v;
        }
    }
    return boost * d;   // unreachable in practice (acceptance > 95 % per attempt)
}
