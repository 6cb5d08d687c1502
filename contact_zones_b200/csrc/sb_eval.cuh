// sb_eval.cuh -- stateless full-evaluation kernels:
//   K1  lh_full_kernel      batched mixture log-likelihood, chains x sites x features   (model.py:145-184)
//   K2  prior_kernel        size prior + Dirichlet parameter priors                      (model.py:932-972,1142-1219)
//   K3  (inside K2)         geo prior: Prim MST of the last zone over the dense cost     (model.py:1096-1139)
//   K4  count_kernel        integer count tables of the Gibbs operators                  (zone_sampling.py:240-400)
#pragma once
#include "sb_common.cuh"

// ---------------------------------------------------------------------------------------------
// wnorm table: the only distinct normalised weight vectors of update_weights / normalize_weights
// (model.py:257-294, 436-452):  w~[f][cfg][c] = w[f][c]*has_c / sum_c' w[f][c']*has_c',
// cfg = has_zone + 2*has_family, has = (1, has_zone, has_family).  Summation order = numpy's (c = 0,1,2).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void wnorm_row(const double* __restrict__ w, double* __restrict__ out, int C) {
    // w: [3], out: [4][3]
#pragma unroll
    for (int cfg = 0; cfg < 4; ++cfg) {
        const double h1 = (cfg & 1) ? 1.0 : 0.0;
        const double h2 = ((cfg & 2) && C == 3) ? 1.0 : 0.0;
        const double a0 = w[0], a1 = w[1] * h1, a2 = (C == 3 ? w[2] : 0.0) * h2;
        const double s = (a0 + a1) + a2;
        out[cfg * 3 + 0] = a0 / s;
        out[cfg * 3 + 1] = a1 / s;
        out[cfg * 3 + 2] = a2 / s;
    }
}

__global__ void wnorm_kernel(DevData d, ChainArrays st, int chain_begin) {
    const int chain = chain_begin + blockIdx.y;
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= d.F) return;
    wnorm_row(st.weights + chain * st.s_weights + (size_t)f * 3, st.wnorm + chain * st.s_wnorm + (size_t)f * 12, d.C);
}

// ---------------------------------------------------------------------------------------------
// K1: one CTA = one (chain, tile of sites).  Each thread streams 16-byte vectors of x (16 cells of one
// site row), looks the mixture terms up in the chain's tables and folds them into a running product
// (LogProd).  partial[chain][tile] is reduced in fixed order by lh_reduce_kernel.
//   SRC = 0:  sum_{n,f: x!=NA} log sum_c w~[f][cfg(n)][c] * p_c[f][x]                      (model.py:174-176)
//   SRC = 1:  sum_{n,f} log( w~[f][cfg(n)][s_nf] * p_{s_nf}[f][x] ),  p = 1 for NA cells,
//             -inf if a chosen weight is 0                                                  (model.py:177-184)
// ---------------------------------------------------------------------------------------------
#define LH_THREADS 256

template <int SRC>
__global__ void __launch_bounds__(LH_THREADS)
lh_full_kernel(DevData d, ChainArrays st, int chain_begin, int sites_per_tile, double* __restrict__ partial,
               int n_tiles) {
    __shared__ double red[33];
    const int chain = chain_begin + blockIdx.y;
    const int tile = blockIdx.x;
    const int n0 = tile * sites_per_tile;
    const int n1 = min(d.N, n0 + sites_per_tile);
    const int upr = d.ldx >> 4;                       // 16-byte units per site row
    const int units = (n1 - n0) * upr;

    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const double* __restrict__ wn = st.wnorm + chain * st.s_wnorm;
    const double* __restrict__ pg = st.p_global + chain * st.s_p_global;
    const double* __restrict__ pz = st.p_zones + chain * st.s_p_zones;
    const double* __restrict__ pf = st.p_fam + chain * st.s_p_fam;
    const uint8_t* __restrict__ src = SRC ? st.source + chain * st.s_source : nullptr;
    const int S = d.S, F = d.F;
    const size_t FS = (size_t)F * S;

    LogProd acc;
    lp_init(acc);
    bool zero_weight = false;

    for (int u = threadIdx.x; u < units; u += LH_THREADS) {
        const int r = u / upr;
        const int n = n0 + r;
        const int f0 = (u - r * upr) << 4;
        const uint4 xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
        uint4 sv = make_uint4(0, 0, 0, 0);
        if (SRC) sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
        const int zid = zone_id[n];
        const int fam = d.inheritance ? d.fam_id[n] : -1;
        const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
        const double* __restrict__ pzz = pz + (size_t)max(zid, 0) * FS;
        const double* __restrict__ pff = pf + (size_t)max(fam, 0) * FS;
        const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
        const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int xs = (xw[j >> 2] >> ((j & 3) * 8)) & 0xff;
            if (xs == SB_PAD) continue;
            const int f = f0 + j;
            const double* __restrict__ w = wn + (size_t)f * 12 + cfg * 3;
            if (SRC) {
                const int s = (sw[j >> 2] >> ((j & 3) * 8)) & 0xff;
                const double ws = w[s];
                if (ws == 0.0) { zero_weight = true; continue; }
                double p = 1.0;
                if (xs != SB_NA) {
                    const size_t o = (size_t)f * S + xs;
                    p = s == 0 ? pg[o] : (s == 1 ? pzz[o] : pff[o]);
                }
                lp_mul(acc, ws * p);
            } else {
                if (xs == SB_NA) continue;
                const size_t o = (size_t)f * S + xs;
                double mix = w[0] * pg[o];
                if (cfg & 1) mix += w[1] * pzz[o];
                if (cfg & 2) mix += w[2] * pff[o];
                lp_mul(acc, mix);
            }
        }
    }
    double v = lp_value(acc);
    if (zero_weight) v = -INFINITY;
    v = block_sum(v, red);
    if (threadIdx.x == 0) partial[(size_t)blockIdx.y * n_tiles + tile] = v;
}

// K1 with source (SRC = 1), blocked form: one CTA = one (16-feature block, chain); the block's parameter tables
// (w~ [16][12], p_* [1 + Z + Fam][16][S + 1], the extra column = 1 for NA cells) are staged in shared memory, so the
// cell body is two shared-memory look-ups and one fp64 multiply into the running product.  Same result as
// lh_full_kernel<1> (which remains the fallback when the tables do not fit); partial[chain][block] is reduced in
// fixed order by lh_reduce_kernel.
__host__ __device__ __forceinline__ size_t lh_source_block_smem(int S, int Z, int Fam) {
    return ((size_t)16 * (S + 1) * (1 + Z + Fam) + 192) * sizeof(double);
}
__global__ void __launch_bounds__(LH_THREADS)
lh_source_block_kernel(DevData d, ChainArrays st, int chain_begin, double* __restrict__ partial, int n_blocks) {
    extern __shared__ double lb_tab[];
    __shared__ double red[33];
    const int chain = chain_begin + blockIdx.y;
    const int f0 = blockIdx.x << 4;
    const int S = d.S, S1 = S + 1, F = d.F, N = d.N;
    const int nf = min(16, F - f0);
    const int bP = 16 * S1, n_rows = 16 * (1 + d.Z + d.Fam);
    const size_t FS = (size_t)F * S;
    double* W = lb_tab + n_rows * S1;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const uint8_t* __restrict__ src = st.source + chain * st.s_source;
    const double* __restrict__ wn = st.wnorm + chain * st.s_wnorm;
    for (int i = threadIdx.x; i < 192; i += blockDim.x) W[i] = i < nf * 12 ? wn[(size_t)f0 * 12 + i] : 1.0;
    for (int r = threadIdx.x; r < n_rows; r += blockDim.x) {
        const int t = r >> 4, j = r & 15;
        const double* srow = (t == 0 ? st.p_global + chain * st.s_p_global
                                     : (t <= d.Z ? st.p_zones + chain * st.s_p_zones + (size_t)(t - 1) * FS
                                                 : st.p_fam + chain * st.s_p_fam + (size_t)(t - 1 - d.Z) * FS))
                             + (size_t)(f0 + j) * S;
        double* dst = lb_tab + r * S1;
        for (int sx = 0; sx < S; ++sx) dst[sx] = j < nf ? srow[sx] : 1.0;
        dst[S] = 1.0;
    }
    __syncthreads();
    LogProd acc;
    lp_init(acc);
    bool zero_weight = false;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const uint4 xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
        const uint4 sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
        const int zid = zone_id[n];
        const int fam = d.inheritance ? d.fam_id[n] : -1;
        const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
        const int bz = (1 + max(zid, 0)) * bP, bf = (1 + d.Z + max(fam, 0)) * bP;
        const double* wr = W + cfg * 3;
        const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
        const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (j >= nf) break;                                             // CTA-uniform (padding columns)
            const int xs = (xw[j >> 2] >> ((j & 3) * 8)) & 0xff;
            const int sc = (sw[j >> 2] >> ((j & 3) * 8)) & 0xff;
            const double ws = wr[j * 12 + sc];
            if (ws == 0.0) { zero_weight = true; continue; }                // model.py:177-184: log 0
            const int e = j * S1 + min(xs, S);                              // NA -> the column of ones
            lp_mul(acc, ws * lb_tab[(sc == 0 ? 0 : (sc == 1 ? bz : bf)) + e]);
        }
    }
    double v = lp_value(acc);
    if (zero_weight) v = -INFINITY;
    v = block_sum(v, red);
    if (threadIdx.x == 0) partial[(size_t)blockIdx.y * n_blocks + blockIdx.x] = v;
}

__global__ void lh_reduce_kernel(const double* __restrict__ partial, int n_tiles, double* __restrict__ ll,
                                 int chain_begin) {
    // one warp per chain, fixed order
    const int c = blockIdx.x;
    double v = 0.0;
    for (int t = threadIdx.x; t < n_tiles; t += 32) v += partial[(size_t)c * n_tiles + t];
    v = warp_sum(v);
    if (threadIdx.x == 0) ll[chain_begin + c] = v;
}

// ---------------------------------------------------------------------------------------------
// K2 + K3: priors.  One CTA per chain.
// ---------------------------------------------------------------------------------------------
// log C(n,k) as the reference computes it (util.py:1202-1217): -betaln(1+n-k, 1+k) - log(n+1)
__device__ __noinline__ double log_binom_d(double n, double k) {
    const double a = 1.0 + n - k, b = 1.0 + k;
    return -(lgamma(a) + lgamma(b) - lgamma(a + b)) - log(n + 1.0);
}

// Dirichlet log-density term of one (table row): lnorm + sum_{s applicable} xlogy(alpha-1, p)
__device__ __forceinline__ double dirichlet_row(const double* __restrict__ p, const double* __restrict__ alpha,
                                                double lnorm, uint32_t mask, int S) {
    double t = lnorm;
    for (int s = 0; s < S; ++s)
        if (mask >> s & 1u) t += xlogy_d(alpha[s] - 1.0, p[s]);
    return t;
}

// Prim's algorithm over the complete graph on the members of one zone (dense cost sub-matrix), run by one
// CTA.  Returns (sum of non-zero MST edge costs, their number) -- scipy's sparse MST result drops zero-weight
// edges (model.py:1125-1131), infinite entries are non-edges (csgraph_from_dense(null_value=inf)).
// key/intree: shared scratch for M entries.
// Cost of the edge (a, b).  Dense form: the symmetric minimum of the two matrix entries.  Implicit form (no N x N
// matrix, f3): the Euclidean distance of the site coordinates -- what the reference's `dist_mat` holds when the cost
// matrix is the network's distance matrix (simulation.py:155-156, preprocessing.py:101-174 without a CRS); the
// multiplications and the sum are rounded separately (no FMA contraction), like scipy's cdist.
__device__ __forceinline__ double geo_cost(const double* __restrict__ cost, const double* __restrict__ loc, int N, int a, int b) {
    if (cost) return fmin(cost[(size_t)a * N + b], cost[(size_t)b * N + a]);
    const double dx = loc[2 * a] - loc[2 * b], dy = loc[2 * a + 1] - loc[2 * b + 1];
    return sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
}
__device__ void prim_mst(const double* __restrict__ cost, const double* __restrict__ loc, int N, const int* __restrict__ members, int M,
                         double* key, unsigned char* intree, double* red, int* redi,
                         double* out_sum, int* out_cnt) {
    for (int j = threadIdx.x; j < M; j += blockDim.x) { key[j] = INFINITY; intree[j] = 0; }
    __syncthreads();
    double sum = 0.0;
    int cnt = 0;
    int cur = 0;   // first member starts the tree
    for (int it = 0; it < M; ++it) {
        if (threadIdx.x == 0) intree[cur] = 1;
        __syncthreads();
        const int mc = members[cur];
        // relax keys and find the arg-min among the outside vertices
        double best = INFINITY;
        int besti = -1;
        for (int j = threadIdx.x; j < M; j += blockDim.x) {
            if (intree[j]) continue;
            const double cst = geo_cost(cost, loc, N, mc, members[j]);
            double k = key[j];
            if (cst < k) { k = cst; key[j] = k; }
            if (besti < 0 || k < best) { best = k; besti = j; }
        }
        // block arg-min (ties -> lowest index; the multiset of MST weights does not depend on ties)
        for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (oi >= 0 && (besti < 0 || ob < best || (ob == best && oi < besti))) { best = ob; besti = oi; }
        }
        const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
        __syncthreads();
        if (lane == 0) { red[w] = best; redi[w] = besti; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double b = INFINITY; int bi = -1;
            for (int k = 0; k < nw; ++k)
                if (redi[k] >= 0 && (bi < 0 || red[k] < b || (red[k] == b && redi[k] < bi))) { b = red[k]; bi = redi[k]; }
            red[32] = b; redi[32] = bi;
        }
        __syncthreads();
        best = red[32]; besti = redi[32];
        if (besti < 0) break;                       // all vertices in the tree
        if (best != INFINITY && best != 0.0) { sum += best; ++cnt; }
        cur = besti;
    }
    *out_sum = sum;
    *out_cnt = cnt;
}

__global__ void __launch_bounds__(256)
prior_kernel(DevData d, ChainArrays st, int chain_begin, int geo_cap) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ double red[33];
    __shared__ int redi[33];
    const int chain = chain_begin + blockIdx.x;
    const int* zsize = st.zone_size + chain * st.s_zone_size;
    const double* pg = st.p_global + chain * st.s_p_global;
    const double* pf = st.p_fam + chain * st.s_p_fam;
    double* ppg = st.prior_pg + chain * st.s_prior_pg;
    double* ppf = st.prior_pf + chain * st.s_prior_pf;

    // --- Dirichlet parameter priors (model.py:596-623, 698-755 types uniform / counts) ---
    double t = 0.0;
    if (d.pg_type == 1) {
        for (int f = threadIdx.x; f < d.F; f += blockDim.x) {
            const double v = dirichlet_row(pg + (size_t)f * d.S, d.alpha_g + (size_t)f * d.S, d.lnorm_g[f],
                                           d.app_mask[f], d.S);
            ppg[f] = v;
            t += v;
        }
    }
    if (d.inheritance && d.pf_type == 1) {
        for (int i = threadIdx.x; i < d.Fam * d.F; i += blockDim.x) {
            const int f = i % d.F;
            const double v = dirichlet_row(pf + (size_t)i * d.S, d.alpha_f + (size_t)i * d.S, d.lnorm_f[i],
                                           d.app_mask[f], d.S);
            ppf[i] = v;
            t += v;
        }
    }
    const double dir_total = block_sum(t, red);

    // --- size prior (model.py:932-972) ---
    double size_p = 0.0;
    if (threadIdx.x == 0) {
        for (int z = 0; z < d.Z; ++z) {
            const double k = (double)zsize[z];
            if (d.size_type == 1) size_p -= log_binom_d((double)d.N, k);
            else if (d.size_type == 2) size_p -= log(k * k);
        }
    }

    // --- geo prior (model.py:1096-1139), last zone only (quirk Q1) ---
    double geo = 0.0;
    if (d.geo_type == 1 && d.Z > 0) {
        const int z = d.Z - 1;
        const int M = min(zsize[z], geo_cap);
        double* key = reinterpret_cast<double*>(smem_raw);
        unsigned char* intree = smem_raw + (size_t)geo_cap * sizeof(double);
        double sum; int cnt;
        prim_mst(d.cost, d.loc, d.N, st.zone_members + chain * st.s_zone_members + (size_t)z * d.N, M,
                 key, intree, red, redi, &sum, &cnt);
        // mean_e expon.logpdf(cost_e, scale) = -log(scale) - mean(cost_e)/scale ; none -> -log(scale)
        geo = cnt > 0 ? (-(double)cnt * log(d.geo_scale) - sum / d.geo_scale) / (double)cnt : -log(d.geo_scale);
    }
    if (threadIdx.x == 0) {
        st.geo[chain] = geo;
        st.prior[chain] = ((size_p + geo) + 0.0) + dir_total;
    }
}

// ---------------------------------------------------------------------------------------------
// K4 (full recount): integer count tables from the resident source state.  Integer atomics only, so the
// result is exact and independent of scheduling.
//   cnt_src [f][cfg][s]   all cells (NA included: the source of an NA cell is a normal draw and the
//                         reference sums sample.source unmasked, zone_sampling.py:240,257,284)
//   cnt_glob[f][x]        s == 0                       (:339-350)
//   cnt_zone[z][f][x]     s == 1 and site in zone z    (:366-372)
//   cnt_fam [k][f][x]     s == 2 and site in family k  (:389-400)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
count_kernel(DevData d, ChainArrays st, int chain_begin, int sites_per_tile) {
    const int chain = chain_begin + blockIdx.y;
    const int n0 = blockIdx.x * sites_per_tile;
    const int n1 = min(d.N, n0 + sites_per_tile);
    const int upr = d.ldx >> 4;
    const int units = (n1 - n0) * upr;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const uint8_t* __restrict__ src = st.source + chain * st.s_source;
    int* cg = st.cnt_glob + chain * st.s_cnt_glob;
    int* cz = st.cnt_zone + chain * st.s_cnt_zone;
    int* cf = st.cnt_fam + chain * st.s_cnt_fam;
    int* cs = st.cnt_src + chain * st.s_cnt_src;
    const int S = d.S, F = d.F;
    for (int u = threadIdx.x; u < units; u += blockDim.x) {
        const int r = u / upr;
        const int n = n0 + r;
        const int f0 = (u - r * upr) << 4;
        const uint4 xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
        const uint4 sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
        const int zid = zone_id[n];
        const int fam = d.inheritance ? d.fam_id[n] : -1;
        const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
        const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
        const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int xs = (xw[j >> 2] >> ((j & 3) * 8)) & 0xff;
            if (xs == SB_PAD) continue;
            const int s = (sw[j >> 2] >> ((j & 3) * 8)) & 0xff;
            const int f = f0 + j;
            atomicAdd(cs + (size_t)f * 12 + cfg * 3 + s, 1);
            if (xs == SB_NA) continue;
            if (s == 0) atomicAdd(cg + (size_t)f * S + xs, 1);
            else if (s == 1) { if (zid >= 0) atomicAdd(cz + ((size_t)zid * F + f) * S + xs, 1); }
            else if (fam >= 0) atomicAdd(cf + ((size_t)fam * F + f) * S + xs, 1);
        }
    }
}

// K4 (full recount), blocked form: one CTA = one (16-feature block, chain).  The block's counters -- [1 + Z + Fam]
// tables of [16][S] and the source table [16][12] -- live in shared memory, are filled with shared-memory atomics
// while the CTA streams its 16-byte column strip of x and source over all sites, and are flushed with plain stores:
// no global atomics and no zeroing pass (every entry of the block's rows is written).  Same semantics as
// count_kernel, which remains the fallback when the tables of a block do not fit in shared memory.
__host__ __device__ __forceinline__ size_t count_block_smem(int S, int Z, int Fam) {
    return ((size_t)16 * S * (1 + Z + Fam) + 192) * sizeof(int);
}
__global__ void __launch_bounds__(256)
count_block_kernel(DevData d, ChainArrays st, int chain_begin) {
    extern __shared__ int cb_tab[];
    const int chain = chain_begin + blockIdx.y;
    const int f0 = blockIdx.x << 4;
    const int S = d.S, F = d.F, N = d.N;
    const int nf = min(16, F - f0);
    const int bS = 16 * S, n_rows = 16 * (1 + d.Z + d.Fam);
    int* tcs = cb_tab + n_rows * S;
    const int T = n_rows * S + 192;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const uint8_t* __restrict__ src = st.source + chain * st.s_source;
    for (int i = threadIdx.x; i < T; i += blockDim.x) cb_tab[i] = 0;
    __syncthreads();
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const uint4 xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
        const uint4 sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
        const int zid = zone_id[n];
        const int fam = d.inheritance ? d.fam_id[n] : -1;
        const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
        const int bz = (1 + max(zid, 0)) * bS, bf = (1 + d.Z + max(fam, 0)) * bS;
        const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
        const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (j >= nf) break;                                             // CTA-uniform (padding columns)
            const int xs = (xw[j >> 2] >> ((j & 3) * 8)) & 0xff;
            const int s = (sw[j >> 2] >> ((j & 3) * 8)) & 0xff;
            atomicAdd(tcs + j * 12 + cfg * 3 + s, 1);
            if (xs == SB_NA) continue;
            if (s == 0) atomicAdd(cb_tab + j * S + xs, 1);
            else if (s == 1) { if (zid >= 0) atomicAdd(cb_tab + bz + j * S + xs, 1); }
            else if (fam >= 0) atomicAdd(cb_tab + bf + j * S + xs, 1);
        }
    }
    __syncthreads();
    const size_t FS = (size_t)F * S;
    for (int r = threadIdx.x; r < n_rows; r += blockDim.x) {
        const int t = r >> 4, j = r & 15;
        if (j >= nf) continue;
        int* dst = (t == 0 ? st.cnt_glob + chain * st.s_cnt_glob
                           : (t <= d.Z ? st.cnt_zone + chain * st.s_cnt_zone + (size_t)(t - 1) * FS
                                       : st.cnt_fam + chain * st.s_cnt_fam + (size_t)(t - 1 - d.Z) * FS))
                   + (size_t)(f0 + j) * S;
        for (int sx = 0; sx < S; ++sx) dst[sx] = cb_tab[r * S + sx];
    }
    int* cs = st.cnt_src + chain * st.s_cnt_src + (size_t)f0 * 12;
    for (int i = threadIdx.x; i < nf * 12; i += blockDim.x) cs[i] = tcs[i];
}

// ---------------------------------------------------------------------------------------------
// debug / parity kernels: per-cell component likelihoods + normalised weights (model.py:230-294) and the
// source posterior (zone_sampling.py:196-202).  out arrays are (N,F,C) dense f64 on the device.
// Rows of absent components are written as 0 (the reference leaves 0 / stale values there, quirk Q6).
// ---------------------------------------------------------------------------------------------
__global__ void components_kernel(DevData d, ChainArrays st, int chain, double* __restrict__ all_lh,
                                  double* __restrict__ wnorm_out, double* __restrict__ post) {
    const size_t cell = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= (size_t)d.N * d.F) return;
    const int n = (int)(cell / d.F), f = (int)(cell % d.F);
    const int xs = d.x[(size_t)n * d.ldx + f];
    const int zid = (st.zone_id + chain * st.s_zone_id)[n];
    const int fam = d.inheritance ? d.fam_id[n] : -1;
    const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
    const double* w = st.wnorm + chain * st.s_wnorm + (size_t)f * 12 + cfg * 3;
    double l[3] = {0.0, 0.0, 0.0};
    if (xs == SB_NA) {
        l[0] = l[1] = l[2] = 1.0;
    } else {
        const size_t o = (size_t)f * d.S + xs;
        l[0] = (st.p_global + chain * st.s_p_global)[o];
        if (zid >= 0) l[1] = (st.p_zones + chain * st.s_p_zones)[(size_t)zid * d.F * d.S + o];
        if (fam >= 0) l[2] = (st.p_fam + chain * st.s_p_fam)[(size_t)fam * d.F * d.S + o];
    }
    double t[3];
    double tot = 0.0;
    for (int c = 0; c < d.C; ++c) { t[c] = l[c] * w[c]; tot += t[c]; }
    for (int c = 0; c < d.C; ++c) {
        if (all_lh) all_lh[cell * d.C + c] = l[c];
        if (wnorm_out) wnorm_out[cell * d.C + c] = w[c];
        if (post) post[cell * d.C + c] = t[c] / tot;
    }
}
