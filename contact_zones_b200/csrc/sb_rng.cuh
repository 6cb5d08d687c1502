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
// One Philox block per attempt: words 0-1 -> 53-bit radius uniform, word 2 -> Box-Muller angle, word 3 -> accept
// uniform.  The candidate normal x is evaluated with fp32 transcendentals (logf / sqrtf / cospif: relative error
// ~1e-7 on x, i.e. a 1e-7 perturbation of the proposal density -- five orders below the Monte-Carlo error of any run)
// and then treated as an exact fp64 number, so everything downstream (v, the accept test, the variate) is fp64.
// The log-form accept test is decided in fp32 when its margin is clear and in fp64 otherwise.
__device__ double rng_gamma(RngKey key, uint64_t step, uint32_t purpose, uint32_t index, double alpha) {
    double boost = 1.0;
    if (alpha < 1.0) {
        const Draw b = rng_draw(key, step, purpose, 0xfffffeu, index);
        boost = pow(b.u0, 1.0 / alpha);
        alpha += 1.0;
    }
    const double d = alpha - 1.0 / 3.0;
    const double c = rsqrt(9.0 * d);
    for (uint32_t attempt = 0; attempt < 128u; ++attempt) {
        uint32_t o[4];
        philox4x32((uint32_t)step, (uint32_t)(step >> 32), (purpose << 24) | attempt, index, key.k0, key.k1, o);
        const float u0 = (float)u53(o[0], o[1]);                               // in (0, 1]
        const float r = sqrtf(-2.0f * logf(u0));
        const double x = (double)(r * cospif(((float)(o[2] >> 8) + 0.5f) * (2.0f / 16777216.0f)));
        const double u = ((double)o[3] + 0.5) * (1.0 / 4294967296.0);
        double v = 1.0 + c * x;
        if (v <= 0.0) continue;
        v = v * v * v;
        const double x2 = x * x;
        if (u < 1.0 - 0.0331 * x2 * x2) return boost * d * v;                   // squeeze
        const float lu = logf((float)u), lv = logf((float)v);
        const float rhs = (float)(0.5 * x2) + (float)d * ((float)(1.0 - v) + lv);
        const float margin = 2e-6f * (fabsf(lu) + (float)(0.5 * x2) + (float)d * (fabsf((float)(1.0 - v)) + fabsf(lv))) + 1e-6f;
        bool acc;
        if (fabsf(lu - rhs) > margin) acc = lu < rhs;
        else acc = log(u) < 0.5 * x2 + d * (1.0 - v + log(v));
        if (acc) return boost * d * v;
    }
    return boost * d;   // unreachable in practice (acceptance > 95 % per attempt)
}

// debug / parity surface of K6 (sb_rng_sample): n independent draws, kind 0 = uniform(0,1), 1 = Gamma(alpha, 1)
__global__ void rng_sample_kernel(int kind, double alpha, uint32_t seed_lo, uint32_t seed_hi, int n, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RngKey key = {seed_lo, seed_hi};
    out[i] = kind == 0 ? rng_uniform(key, 0, RP_GAMMA, (uint32_t)i) : rng_gamma(key, 0, RP_GAMMA, (uint32_t)i, alpha);
}
