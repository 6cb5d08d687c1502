// sb_common.cuh -- device-side data model shared by all kernels of libsbayes_b200.so (sm_100a).
//
// HBM layout (DESIGN.md "Data layout"):
//   immutable, one copy per GPU, shared by all chains (L2-resident at every shape we run):
//     x    u8 [N][ldx]   state index per (site, feature), row = site; 255 = NA, 254 = row padding
//     xT   u8 [F][ldn]   the same, row = feature (column operators: alter_*, gibbs_p_global)
//     fam  i16[N]        family id or -1;  fam_ptr/fam_sites: CSR member lists of the families
//     adj  CSR           Delaunay adjacency (neighbourhood of a zone)
//     cost f64[N][N]     dense cost matrix of the geo prior (optional)
//   per chain (struct-of-arrays over chains, chain-major):
//     zone_id i8[N], zone_members i32[Z][N], zone_size i32[Z], weights f64[F][3], wnorm f64[F][4][3],
//     p_global f64[F][S], p_zones f64[Z][F][S], p_fam f64[Fam][F][S], source u8[N][ldx],
//     cnt_glob i32[F][S], cnt_zone i32[Z][F][S], cnt_fam i32[Fam][F][S], cnt_src i32[F][4][3]
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define SB_NA 255
#define SB_PAD 254
#define SB_MAX_S 32
#define SB_MAX_Z 64
#define SB_LN2 0.693147180559945309417232121458

struct DevData {
    int N, F, S, Z, Fam, C;
    int ldx;   // padded row length of x / source (multiple of 16)
    int ldn;   // padded row length of xT (multiple of 16)
    int inheritance, sample_source, warmup;
    const uint8_t* x;
    const uint8_t* xT;
    const int16_t* fam_id;
    const int* fam_ptr;
    const int* fam_sites;
    const uint32_t* app_mask;   // [F] bit s set <=> state s applicable (data.states)
    const int* n_app;           // [F]
    const int* adj_ptr;
    const int* adj_idx;
    const double* cost;         // may be null
    const double* loc;          // [N][2] site coordinates: implicit Euclidean cost when `cost` is null (may be null)
    // priors
    int size_type, geo_type, pg_type, pf_type;
    double geo_scale;
    const double* gibbs_g;      // [F][S]     prior_p_global.counts
    const double* alpha_g;      // [F][S]     Dirichlet density alpha (counts prior)
    const double* lnorm_g;      // [F]        lgamma(sum a) - sum lgamma(a)
    const double* gibbs_f;      // [Fam][F][S]
    const double* alpha_f;      // [Fam][F][S]
    const double* lnorm_f;      // [Fam][F]
};

// Pointers to chain 0 of every per-chain array plus the per-chain strides (in elements).
struct ChainArrays {
    int8_t* zone_id;      size_t s_zone_id;
    int* zone_members;    size_t s_zone_members;
    int* zone_size;       size_t s_zone_size;
    double* weights;      size_t s_weights;      // [F][3]
    double* wnorm;        size_t s_wnorm;        // [F][4][3]   cfg = has_zone + 2*has_family
    double* p_global;     size_t s_p_global;
    double* p_zones;      size_t s_p_zones;
    double* p_fam;        size_t s_p_fam;
    uint8_t* source;      size_t s_source;
    int* cnt_glob;        size_t s_cnt_glob;
    int* cnt_zone;        size_t s_cnt_zone;
    int* cnt_fam;         size_t s_cnt_fam;
    int* cnt_src;         size_t s_cnt_src;      // [F][4][3]
    double* ll;           // [chains] running log-likelihood
    double* prior;        // [chains] running log-prior
    double* geo;          // [chains] geo-prior term
    double* prior_pg;     size_t s_prior_pg;     // [F]      per-feature Dirichlet prior terms (cache)
    double* prior_pf;     size_t s_prior_pf;     // [Fam][F]
    unsigned long long* acc;   // [chains][SB_N_OPS]
    unsigned long long* rej;   // [chains][SB_N_OPS]
    unsigned long long* step;  // [chains] steps done so far (RNG counter)
    int* err;                  // [chains] sticky error flags
};

// ---------------------------------------------------------------------------------------------
// Sum of logs as the log of a running product: one fp64 multiply per factor instead of one fp64 log.
// The mantissa is renormalised (exponent moved into an integer) every SB_LP_PERIOD factors; factors
// below 2^-120 (incl. 0 and subnormals) take the exact slow path through log().  Relative error of the
// result is ~ n_factors * 2^-53 on the product, i.e. ~1e-16 absolute per factor on the log -- far inside
// the 1e-9 relative parity bound (tests/test_gpu_parity.py).
// ---------------------------------------------------------------------------------------------
#define SB_LP_PERIOD 8
// libm's log() expands to ~100 instructions at every call site; the cold paths share one out-of-line copy (the step
// kernel is sensitive to the size of its hot loop)
__device__ __noinline__ double sb_log(double x) { return log(x); }
struct LogProd {
    double m;      // mantissa product, in [2^-960, 2)
    double side;   // exact logs of tiny factors (may be -inf)
    int e;         // accumulated binary exponent
    int k;         // factors since the last renormalisation
};
__device__ __forceinline__ void lp_init(LogProd& a) { a.m = 1.0; a.side = 0.0; a.e = 0; a.k = 0; }
__device__ __forceinline__ void lp_renorm(LogProd& a) {
    int hi = __double2hiint(a.m);
    a.e += ((hi >> 20) & 0x7ff) - 1023;
    a.m = __hiloint2double((hi & 0x800fffff) | 0x3ff00000, __double2loint(a.m));
    a.k = 0;
}
__device__ __forceinline__ void lp_mul(LogProd& a, double v) {
    if (__double2hiint(v) < 0x38700000) {   // v < 2^-120, zero, subnormal (or negative/NaN): exact path
        a.side += sb_log(v);
    } else {
        a.m *= v;
        if (++a.k == SB_LP_PERIOD) lp_renorm(a);
    }
}
__device__ __forceinline__ double lp_value(LogProd& a) {
    lp_renorm(a);
    return log(a.m) + (double)a.e * SB_LN2 + a.side;
}

// ---------------------------------------------------------------------------------------------
// Rotated column strips (K4, K1s).  A thread holds the 16-byte strip (16 features) of one site; at unrolled step j lane
// l works on feature (l + j) & 15 of ITS site and on table column (l + j) & 31, so the 32 lanes of a warp always touch
// 32 different columns.  Tables laid out [row][32 columns] of 32-bit counters (column c and c + 16 both belong to
// feature c & 15 and are added up at the end) or [row][16 features] of 64-bit values are then updated / read without a
// bank conflict whatever the rows are.  The strip is rotated once per site (12 instructions) so that step j extracts
// the STATIC byte j.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 rot_strip(const uint4 v, const int rot) {
    uint32_t a0 = v.x, a1 = v.y, a2 = v.z, a3 = v.w, t;
    if (rot & 4) { t = a0; a0 = a1; a1 = a2; a2 = a3; a3 = t; }
    if (rot & 8) { t = a0; a0 = a2; a2 = t; t = a1; a1 = a3; a3 = t; }
    const int br = (rot & 3) * 8;
    return make_uint4(__funnelshift_r(a0, a1, br), __funnelshift_r(a1, a2, br), __funnelshift_r(a2, a3, br),
                      __funnelshift_r(a3, a0, br));
}
__device__ __forceinline__ uint32_t strip_byte(const uint4& v, const int j) {      // j is a compile-time constant
    const uint32_t w = (j >> 2) == 0 ? v.x : ((j >> 2) == 1 ? v.y : ((j >> 2) == 2 ? v.z : v.w));
    return (w >> ((j & 3) * 8)) & 0xffu;
}
// state index (6 bits; NA / padding -> 63 / 62) and source component (2 bits) of the 16 cells in one strip
__device__ __forceinline__ uint4 strip_merge(const uint4 x, const uint4 s) {
    return make_uint4((x.x & 0x3f3f3f3fu) | ((s.x << 6) & 0xc0c0c0c0u), (x.y & 0x3f3f3f3fu) | ((s.y << 6) & 0xc0c0c0c0u),
                      (x.z & 0x3f3f3f3fu) | ((s.z << 6) & 0xc0c0c0c0u), (x.w & 0x3f3f3f3fu) | ((s.w << 6) & 0xc0c0c0c0u));
}
// first row of the table a cell's source selects, as a 16-bit field picked by PRMT: [0] global (row 0), [1] the site's
// zone, [2] the site's family, [3] the scrap table (a source the site has no table for)
struct RowSel { uint32_t lo, hi; };
__device__ __forceinline__ RowSel row_sel(int zid, int fam, int Z, int n_tab, int rows_per_tab) {
    RowSel r;
    r.lo = (uint32_t)((zid >= 0 ? 1 + zid : n_tab) * rows_per_tab) << 16;
    r.hi = (uint32_t)((fam >= 0 ? 1 + Z + fam : n_tab) * rows_per_tab) | ((uint32_t)(n_tab * rows_per_tab) << 16);
    return r;
}
__device__ __forceinline__ uint32_t row_of(const RowSel& r, uint32_t sc) {         // sc in 0..3
    return __byte_perm(r.lo, r.hi, sc * 0x22u + 0x10u);
}

// ---------------------------------------------------------------------------------------------
// deterministic block reductions (fixed shuffle tree + fixed warp order)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// log of the product of the 32 lanes' running products (mantissa in [1, 4), binary exponent, exact side logs): the
// mantissas are multiplied across the warp (< 4^32) so that ONE log per warp is taken; lane 0 holds the value, the
// other lanes 0, i.e. the result is ready for block_sum.
__device__ __forceinline__ double lp_warp_value(double m, int e, double side) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m *= __shfl_xor_sync(0xffffffffu, m, o);
        e += __shfl_xor_sync(0xffffffffu, e, o);
        side += __shfl_xor_sync(0xffffffffu, side, o);
    }
    return (threadIdx.x & 31) == 0 ? log(m) + (double)e * SB_LN2 + side : 0.0;
}
// Result valid in every thread. `red` = shared scratch of >= 33 doubles.
__device__ __forceinline__ double block_sum(double v, double* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    if (w == 0) {
        double t = lane < nw ? red[lane] : 0.0;
        t = warp_sum(t);
        if (lane == 0) red[32] = t;
    }
    __syncthreads();
    return red[32];
}
__device__ __forceinline__ int block_sum_i(int v, int* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum_i(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    if (w == 0) {
        int t = lane < nw ? red[lane] : 0;
        t = warp_sum_i(t);
        if (lane == 0) red[32] = t;
    }
    __syncthreads();
    return red[32];
}

// ---------------------------------------------------------------------------------------------
// Table-driven natural logarithm for the step kernel.  fp64 throughput per SM is a quarter of the issue rate, and
// libm's log() costs ~60 fp64 instructions; this one costs ~10.  x = 2^e * m, m in [1,2); the top 7 mantissa bits
// pick c_k = 1 + (k + 1/2)/128 from a 128-entry table of (1/c_k rounded, -log of that rounded value), so
// log x = e ln2 + log(c_k) + log1p(r),  r = m * (1/c_k) - 1,  |r| <= 2^-8, degree-7 Taylor (remainder < 1e-20).
// Absolute error <= ~3e-16 + 1e-16 |log x| -- the sampler only sums these into log-likelihoods of magnitude >= 1,
// where the 1e-9 relative parity bound leaves > 6 orders of head-room.  Non-positive, subnormal, inf and NaN inputs
// go to log().
// ---------------------------------------------------------------------------------------------
#define SB_FLOG_N 128
__device__ __forceinline__ void flog_init(double2* tab) {      // call by the whole CTA, then __syncthreads()
    for (int k = threadIdx.x; k < SB_FLOG_N; k += blockDim.x) {
        const double inv = 1.0 / (1.0 + ((double)k + 0.5) * (1.0 / SB_FLOG_N));
        tab[k] = make_double2(inv, -log(inv));
    }
}
__device__ __forceinline__ double flog(double x, const double2* __restrict__ tab) {
    const int hi = __double2hiint(x);
    if ((unsigned)(hi - 0x00100000) >= 0x7fe00000u) return sb_log(x);     // non-positive / subnormal / inf / NaN inputs
    const int e = (hi >> 20) - 1023;
    const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(x));
    const double2 t = tab[(hi >> 13) & (SB_FLOG_N - 1)];
    const double r = fma(m, t.x, -1.0);
    double p = fma(r, 1.0 / 7.0, -1.0 / 6.0);
    p = fma(p, r, 0.2);
    p = fma(p, r, -0.25);
    p = fma(p, r, 1.0 / 3.0);
    p = fma(p, r, -0.5);
    const double l = fma(p * r, r, r);
    const double ef = (double)e;
    return fma(ef, 6.93147180369123816490e-01, fma(ef, 1.90821492927058770002e-10, t.y + l));   // ln2 = hi + lo
}

// xlogy semantics of scipy (dirichlet._logpdf): 0 * log(0) = 0
__device__ __forceinline__ double xlogy_d(double a, double p) { return a == 0.0 ? 0.0 : a * sb_log(p); }
