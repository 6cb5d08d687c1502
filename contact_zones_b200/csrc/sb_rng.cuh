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

// the same integer from a uniform that was drawn ahead of time
__device__ __forceinline__ int ahead_int(double u, int n) {
    const int v = (int)(u * (double)n);
    return v < n ? v : n - 1;
}

// Gamma(alpha, 1), Marsaglia & Tsang (2000).  alpha < 1 uses Gamma(alpha+1) * U^(1/alpha).
// One Philox block per attempt: word 0 -> Box-Muller radius, word 1 -> angle, words 2-3 -> the accept uniforms of the
// two normals, which are both tried before a new block is drawn (so a whole warp rarely needs a second attempt:
// 1 - (1 - 0.05^2)^32 ~ 8 %).  The candidate normals are evaluated with the fp32 hardware approximations (MUFU
// lg2 / sqrt / sin / cos, ~12 instructions instead of ~55 for logf / sqrtf / sincospif): absolute error ~5e-7 on x,
// i.e. a ~1e-6 perturbation of the proposal density -- orders of magnitude below the Monte-Carlo error of any run
// (tests/test_rng_parity.py: KS tests against the exact CDFs); the radius uniform has 32 bits, so |x| <= 6.8.  The
// candidates are then treated as exact fp64 numbers: v, the accept test and the variate are fp64.  The log-form accept test is decided by a fast fp32 estimate
// when its margin is clear and in fp64 otherwise.
// the fp64 form of the log test: rare (the fp32 estimate decides unless its margin is unclear), so it is kept out of
// line -- the step kernel is sensitive to the size of its hot loop (instruction cache)
__device__ __noinline__ bool mt_accept_exact(double u, double x2, double d, double v) {
    return log(u) < 0.5 * x2 + d * (1.0 - v + log(v));
}
__device__ __forceinline__ bool mt_accept(double x, double u, double c, double d, double& out) {
    // written without early exits so that two candidates can be interleaved by the compiler
    const double v1 = 1.0 + c * x;
    const double v = v1 * v1 * v1;
    const double x2 = x * x;
    out = d * v;
    const bool pos = v1 > 0.0;
    const bool squeeze = u < 1.0 - 0.0331 * x2 * x2;
    const float hx = (float)(0.5 * x2), df = (float)d, omv = (float)(1.0 - v);
    const float lu = __logf((float)u), lv = __logf(fabsf((float)v));
    const float rhs = hx + df * (omv + lv);
    // fp32 error budget: __logf is good to 2^-21 absolute near 1 and a few ulp elsewhere
    const float margin = 4e-6f * (fabsf(lu) + hx + df * (fabsf(omv) + fabsf(lv))) + 2e-6f * df + 4e-6f;
    const bool clear = fabsf(lu - rhs) > margin;
    bool acc = squeeze || (clear && lu < rhs);
    if (pos && !squeeze && !clear) acc = mt_accept_exact(u, x2, d, v);
    return pos && acc;
}

// Two independent Gamma draws in lock-step (same streams as two rng_gamma calls with alpha >= 1): the two dependency
// chains are laid side by side in one instruction stream, which is what hides their latency -- the step kernel runs
// one CTA of 16 warps per SM, far too few to hide a ~350-instruction dependent chain by warp switching alone.
// valid_x = false skips that draw (result 0).
struct MtCand { double x0, x1, u0, u1; };
__device__ __forceinline__ MtCand mt_candidates(RngKey key, uint64_t step, uint32_t purpose, uint32_t attempt, uint32_t index) {
    uint32_t o[4];
    philox4x32((uint32_t)step, (uint32_t)(step >> 32), (purpose << 24) | attempt, index, key.k0, key.k1, o);
    // hardware approximations (MUFU lg2 / sqrt / sin / cos): absolute error ~5e-7 on the candidate, see above
    float r, sn, cs;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r)
        : "f"(-1.3862943611198906f * __log2f(((float)o[0] + 0.5f) * 2.3283064365386963e-10f)));     // sqrt(-2 ln u)
    // angle uniform on [-pi, pi): the range where sin.approx / cos.approx are most accurate
    __sincosf(((float)(o[1] >> 8) - 8388607.5f) * (6.283185307179586f / 16777216.0f), &sn, &cs);
    MtCand m;
    m.x0 = (double)(r * cs); m.x1 = (double)(r * sn);
    m.u0 = ((double)o[2] + 0.5) * (1.0 / 4294967296.0);
    m.u1 = ((double)o[3] + 0.5) * (1.0 / 4294967296.0);
    return m;
}
__device__ __forceinline__ double mt_scale(double alpha, double& d) {
    d = alpha - 1.0 / 3.0;
    const double nd = 9.0 * d;
    const double c = (double)rsqrtf((float)nd);
    return c * fma(-0.5 * nd, c * c, 1.5);
}
__device__ __forceinline__ void rng_gamma2(RngKey key, uint64_t step, uint32_t purpose_a, uint32_t index_a, double alpha_a,
                                           bool valid_a, uint32_t purpose_b, uint32_t index_b, double alpha_b,
                                           bool valid_b, double& ga, double& gb) {
    double da, db;
    const double ca = mt_scale(valid_a ? alpha_a : 1.0, da), cb = mt_scale(valid_b ? alpha_b : 1.0, db);
    bool need_a = valid_a, need_b = valid_b;
    ga = 0.0; gb = 0.0;
    for (uint32_t attempt = 0; attempt < 64u && (need_a || need_b); ++attempt) {
        const MtCand ma = mt_candidates(key, step, purpose_a, attempt, index_a);
        const MtCand mb = mt_candidates(key, step, purpose_b, attempt, index_b);
        double a0, a1, b0, b1;
        const bool ka0 = mt_accept(ma.x0, ma.u0, ca, da, a0), kb0 = mt_accept(mb.x0, mb.u0, cb, db, b0);
        const bool ka1 = mt_accept(ma.x1, ma.u1, ca, da, a1), kb1 = mt_accept(mb.x1, mb.u1, cb, db, b1);
        if (need_a && (ka0 || ka1)) { ga = ka0 ? a0 : a1; need_a = false; }
        if (need_b && (kb0 || kb1)) { gb = kb0 ? b0 : b1; need_b = false; }
    }
    // 64 rejected attempts (probability < 1e-80 per draw): poison the draw -- the NaN reaches the running
    // log-likelihood and is reported as SB_ERR_NONFINITE instead of passing for a sample
    if (need_a) ga = __longlong_as_double(0x7ff8000000000000LL);
    if (need_b) gb = __longlong_as_double(0x7ff8000000000000LL);
}

// Single draw (same stream as rng_gamma2 for that (purpose, index)); alpha < 1 uses Gamma(alpha+1) * U^(1/alpha).
__device__ double rng_gamma(RngKey key, uint64_t step, uint32_t purpose, uint32_t index, double alpha) {
    double boost = 1.0;
    if (alpha < 1.0) {
        const Draw b = rng_draw(key, step, purpose, 0xfffffeu, index);
        boost = pow(b.u0, 1.0 / alpha);
        alpha += 1.0;
    }
    double d;
    const double c = mt_scale(alpha, d);
    for (uint32_t attempt = 0; attempt < 64u; ++attempt) {
        const MtCand m = mt_candidates(key, step, purpose, attempt, index);
        double g;
        if (mt_accept(m.x0, m.u0, c, d, g)) return boost * g;
        if (mt_accept(m.x1, m.u1, c, d, g)) return boost * g;
    }
    return __longlong_as_double(0x7ff8000000000000LL);   // unreachable in practice (acceptance > 95 % per candidate); see rng_gamma2
}

// debug / parity surface of K6 (sb_rng_sample): n independent draws, kind 0 = uniform(0,1), 1 = Gamma(alpha, 1),
// 3 = the N(0,1) candidates of the Gamma sampler,
// 2 = flog((1 + u_i) * 2^(i % 121 - 60)) with u_i the kind-0 draw (accuracy check of the table-driven logarithm)
__global__ void rng_sample_kernel(int kind, double alpha, uint32_t seed_lo, uint32_t seed_hi, int n, double* out) {
    __shared__ double2 tab[SB_FLOG_N];
    flog_init(tab);
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RngKey key = {seed_lo, seed_hi};
    if (kind == 1) { out[i] = rng_gamma(key, 0, RP_GAMMA, (uint32_t)i, alpha); return; }
    if (kind == 3) {
        const MtCand m = mt_candidates(key, 0, RP_GAMMA, 0, (uint32_t)(i >> 1));
        out[i] = (i & 1) ? m.x1 : m.x0;
        return;
    }
    const double u = rng_uniform(key, 0, RP_GAMMA, (uint32_t)i);
    out[i] = kind == 0 ? u : flog(ldexp(1.0 + u, i % 121 - 60), tab);
}
