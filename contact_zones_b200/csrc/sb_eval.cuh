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

// K1 with source (SRC = 1), blocked form: one CTA = one (16-feature block, chain); the block's parameter tables are
// staged in shared memory feature-fastest -- P[table][state column][16] (tables: global, Z zones, Fam families, a scrap
// table of ones; the extra state column = 1 for NA / padding cells) and W[cfg][4][16] (w~ of the chosen source) -- and
// read through rotated strips (see above): two conflict-free LDS.64 and two fp64 multiplies per cell, 8 factors per
// partial product, one exponent fold per site.  A site with a factor below 2^-120 (incl. a zero weight:
// model.py:177-184, log 0) is redone cell by cell on the exact path.  Same result as lh_full_kernel<1> (which remains
// the fallback when the tables do not fit); partial[chain][block] is reduced in fixed order by the chain's last CTA.
__host__ __device__ __forceinline__ size_t lh_source_block_smem(int S, int Z, int Fam) {
    return ((size_t)(2 + Z + Fam) * (S + 1) * 16 + 256) * sizeof(double);
}
__global__ void __launch_bounds__(LH_THREADS)
lh_source_block_kernel(DevData d, ChainArrays st, int chain_begin, double* __restrict__ partial, int n_blocks,
                       unsigned* __restrict__ ticket) {
    extern __shared__ __align__(128) double lb_tab[];
    __shared__ double red[33];
    __shared__ int is_last;
    const int chain = chain_begin + blockIdx.y;
    const int f0 = blockIdx.x << 4;
    const int S = d.S, S1 = S + 1, F = d.F, N = d.N;
    const int nf = min(16, F - f0), nfS = nf * S;
    const int n_tab = 1 + d.Z + d.Fam, n_val = (n_tab + 1) * S1 * 16;
    const size_t FS = (size_t)F * S;
    double* W = lb_tab + n_val;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const uint8_t* __restrict__ src = st.source + chain * st.s_source;
    for (int i = threadIdx.x; i < ((n_val + 256) >> 1); i += blockDim.x)
        reinterpret_cast<double2*>(lb_tab)[i] = make_double2(1.0, 1.0);
    __syncthreads();
    if (threadIdx.x < nf) {                                       // w~ rows of the block (also refreshes the chain's table)
        double row[12];
        wnorm_row(st.weights + chain * st.s_weights + (size_t)(f0 + threadIdx.x) * 3, row, d.C);
        double* g = st.wnorm + chain * st.s_wnorm + (size_t)(f0 + threadIdx.x) * 12;
#pragma unroll
        for (int i = 0; i < 12; ++i) {
            W[((i / 3) * 4 + i % 3) * 16 + threadIdx.x] = row[i];
            g[i] = row[i];
        }
    }
    {   // thread -> (group, entry i = j * S + sx of the block's contiguous [nf][S] slice); groups split the tables
        const int groups = max(1, (int)blockDim.x / nfS);
        for (int i0 = threadIdx.x; i0 < groups * nfS; i0 += blockDim.x) {
            const int g = i0 / nfS, i = i0 - g * nfS;
            const int j = i / S, sx = i - j * S;
            double* dst = lb_tab + sx * 16 + j;
            const size_t o = (size_t)f0 * S + i;
            // eight table entries in flight per thread (the loads are L2 round trips)
            for (int t0 = g; t0 < n_tab; t0 += 8 * groups) {
                double v4[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int t = t0 + u * groups;
                    const double* srow = t == 0 ? st.p_global + chain * st.s_p_global
                                                : (t <= d.Z ? st.p_zones + chain * st.s_p_zones + (size_t)(t - 1) * FS
                                                            : st.p_fam + chain * st.s_p_fam + (size_t)(t - 1 - d.Z) * FS);
                    v4[u] = t < n_tab ? srow[o] : 1.0;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int t = t0 + u * groups;
                    if (t < n_tab) dst[(size_t)t * S1 * 16] = v4[u];
                }
            }
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t p_s = (uint32_t)__cvta_generic_to_shared(lb_tab), w_s = (uint32_t)__cvta_generic_to_shared(W);
    double am = 1.0, side = 0.0;                                   // running product: mantissa in [1, 4), exact logs
    int ae = 0;
    bool zero_weight = false;
    int n = threadIdx.x;
    uint4 xv = make_uint4(0, 0, 0, 0), sv = xv;
    int zid = -1, fam = -1;
    if (n < N) {
        xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
        sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
        zid = zone_id[n];
        fam = d.inheritance ? d.fam_id[n] : -1;
    }
    while (n < N) {
        const uint4 cr = rot_strip(strip_merge(xv, sv), lane & 15);
        const RowSel rs = row_sel(zid, fam, d.Z, n_tab, S1);
        const uint32_t wrow = w_s + (uint32_t)((zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0)) * 512u;        // W[cfg][..][..]
        n += blockDim.x;
        if (n < N) {                                                                   // next site's loads in flight
            xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
            sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
            zid = zone_id[n];
            fam = d.inheritance ? d.fam_id[n] : -1;
        }
        double m0 = am, m1 = 1.0;
        int lo_hi = 0x7fffffff;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const uint32_t col = (uint32_t)((lane + j) & 15) * 8u;                     // loop-invariant per (thread, j)
            const uint32_t c = strip_byte(cr, j);
            const uint32_t xs = min(c & 63u, (uint32_t)S), sc = c >> 6;                // NA / padding -> the column of ones
            double ws, pv;
            asm("ld.shared.f64 %0, [%1];" : "=d"(ws) : "r"((wrow + sc * 128u) | col));
            asm("ld.shared.f64 %0, [%1];" : "=d"(pv) : "r"((p_s + (row_of(rs, sc) + xs) * 128u) | col));
            const double v = ws * pv;
            lo_hi = min(lo_hi, __double2hiint(v));
            if (j & 1) m1 *= v; else m0 *= v;
        }
        if (lo_hi >= 0x38700000) {
            // 8 factors >= 2^-120 per partial product on top of a mantissa in [1, 4): no underflow before the fold
            const int h0 = __double2hiint(m0), h1 = __double2hiint(m1);
            ae += ((h0 >> 20) & 0x7ff) + ((h1 >> 20) & 0x7ff) - 2046;
            am = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(m0))
                 * __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(m1));
            continue;
        }
        // exact path: a tiny / zero factor somewhere in this site's strip
#pragma unroll 1
        for (int j = 0; j < 16; ++j) {
            const uint32_t col = (uint32_t)((lane + j) & 15) * 8u;
            const uint32_t w4 = j < 4 ? cr.x : (j < 8 ? cr.y : (j < 12 ? cr.z : cr.w));
            const uint32_t c = (w4 >> ((j & 3) * 8)) & 0xffu;
            const uint32_t xs = min(c & 63u, (uint32_t)S), sc = c >> 6;
            double ws, pv;
            asm("ld.shared.f64 %0, [%1];" : "=d"(ws) : "r"((wrow + sc * 128u) | col));
            asm("ld.shared.f64 %0, [%1];" : "=d"(pv) : "r"((p_s + (row_of(rs, sc) + xs) * 128u) | col));
            if (ws == 0.0) { zero_weight = true; continue; }
            const double v = ws * pv;
            if (__double2hiint(v) < 0x38700000) { side += sb_log(v); continue; }
            am *= v;
            const int h = __double2hiint(am);
            ae += ((h >> 20) & 0x7ff) - 1023;
            am = __hiloint2double((h & 0x800fffff) | 0x3ff00000, __double2loint(am));
        }
    }
    double v = lp_warp_value(am, ae, side);
    if (zero_weight) v = -INFINITY;
    v = block_sum(v, red);
    if (threadIdx.x == 0) {
        partial[(size_t)blockIdx.y * n_blocks + blockIdx.x] = v;
        __threadfence();
        is_last = atomicAdd(&ticket[blockIdx.y], 1u) == (unsigned)(n_blocks - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double t = 0.0;
        for (int i = threadIdx.x; i < n_blocks; i += blockDim.x) t += __ldcg(&partial[(size_t)blockIdx.y * n_blocks + i]);
        t = block_sum(t, red);
        if (threadIdx.x == 0) {
            st.ll[chain] = t;
            ticket[blockIdx.y] = 0u;
        }
    }
}

// K1 source-free, blocked form (shapes whose zone-free table does not pay: one site per code byte, or few sites): one
// CTA = one (16-feature block, chain).  For a site in no zone the mixture value of a cell depends only on
// (f, x, family), so the CTA builds V[class][state column][16] = w~[f][cfg][0] p_global + w~[f][cfg][2] p_family
// (classes: Fam families, "no family", and a class of ones that zone members read; the extra state column = 1 for
// NA / padding cells) straight from the chain's global tables and streams its 16-byte column strip of x over all sites
// through rotated strips (sb_common.cuh): ONE conflict-free LDS.64 and one fp64 multiply per cell, 8 factors per
// partial product, one exponent fold per site.  The <= Z * MAX_M zone members then get their exact value in a second,
// dense pass over the member lists:  mix = rho_f V[class][x] + w~[f][cfg][1] p_zone[z][x],  rho_f = the weight mass
// of the zone-free components (1 - w~[f][cfg][1] up to rounding; for an NA cell V = p_zone = 1).  A site with a
// factor below 2^-120 is redone cell by cell on the exact path (log 0 = -inf as in model.py:174-176).
// partial[chain][block] is reduced in fixed order by the chain's last CTA.
__host__ __device__ __forceinline__ size_t lh_free_block_smem(int S, int Z, int Fam, int blocks_per_cta = 1) {
    return (size_t)blocks_per_cta * ((size_t)(2 + Fam + Z) * (S + 1) * 16 + 64) * sizeof(double);
}
// the 16 cells of one site folded into (am, ae, side): v_j = rho_j * V[vrow + x_j] + om_j * PZ[zrow + x_j] with ZONE,
// else V[vrow + x_j]
template <bool ZONE>
__device__ __forceinline__ void lh_free_site(const uint4 xr, const int lane, const int S, const uint32_t vrow,
                                             const uint32_t zrow, const uint32_t coef, double& am, int& ae, double& side) {
    double m0 = am, m1 = 1.0;
    int lo_hi = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const uint32_t col = (uint32_t)((lane + j) & 15) * 8u;                         // loop-invariant per (thread, j)
        const uint32_t xs = min(strip_byte(xr, j), (uint32_t)S);                       // NA / padding -> the column of ones
        double v;
        asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"((vrow + xs * 128u) | col));
        if (ZONE) {
            double pz, rho, om;
            asm("ld.shared.f64 %0, [%1];" : "=d"(pz) : "r"((zrow + xs * 128u) | col));
            asm("ld.shared.f64 %0, [%1];" : "=d"(rho) : "r"(coef | col));
            asm("ld.shared.f64 %0, [%1];" : "=d"(om) : "r"((coef + 128u) | col));
            v = rho * v + om * pz;
        }
        lo_hi = min(lo_hi, __double2hiint(v));
        if (j & 1) m1 *= v; else m0 *= v;
    }
    if (lo_hi >= 0x38700000) {
        // 8 factors >= 2^-120 per partial product on top of a mantissa in [1, 4): no underflow before the fold
        const int h0 = __double2hiint(m0), h1 = __double2hiint(m1);
        ae += ((h0 >> 20) & 0x7ff) + ((h1 >> 20) & 0x7ff) - 2046;
        am = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(m0))
             * __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(m1));
        return;
    }
    // exact path: a tiny / zero factor somewhere in this site's strip
#pragma unroll 1
    for (int j = 0; j < 16; ++j) {
        const uint32_t col = (uint32_t)((lane + j) & 15) * 8u;
        const uint32_t w4 = j < 4 ? xr.x : (j < 8 ? xr.y : (j < 12 ? xr.z : xr.w));
        const uint32_t xs = min((w4 >> ((j & 3) * 8)) & 0xffu, (uint32_t)S);
        double v;
        asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"((vrow + xs * 128u) | col));
        if (ZONE) {
            double pz, rho, om;
            asm("ld.shared.f64 %0, [%1];" : "=d"(pz) : "r"((zrow + xs * 128u) | col));
            asm("ld.shared.f64 %0, [%1];" : "=d"(rho) : "r"(coef | col));
            asm("ld.shared.f64 %0, [%1];" : "=d"(om) : "r"((coef + 128u) | col));
            v = rho * v + om * pz;
        }
        if (__double2hiint(v) < 0x38700000) { side += sb_log(v); continue; }
        am *= v;
        const int h = __double2hiint(am);
        ae += ((h >> 20) & 0x7ff) - 1023;
        am = __hiloint2double((h & 0x800fffff) | 0x3ff00000, __double2loint(am));
    }
}
// grid = (parts per chain, chains): CTA `part` owns the feature blocks [part nb / parts, (part + 1) nb / parts) of its
// chain (one block per CTA on wide tables; several on small ones, where a launch would otherwise be waves of CTAs that
// are all prologue).
__global__ void __launch_bounds__(LH_THREADS)
lh_free_block_kernel(DevData d, ChainArrays st, int chain_begin, double* __restrict__ partial, unsigned* __restrict__ ticket) {
    extern __shared__ __align__(128) double fb_tab[];
    __shared__ double red[33];
    __shared__ int is_last;
    __shared__ int zoff[SB_MAX_Z + 1];
    const int chain = chain_begin + blockIdx.y;
    const int parts = gridDim.x, nb = d.ldx >> 4;
    const int b0 = (int)((long)blockIdx.x * nb / parts), b1 = (int)((long)(blockIdx.x + 1) * nb / parts), nbl = b1 - b0;
    const int S = d.S, S1 = S + 1, F = d.F, N = d.N, Fam = d.Fam;
    const int f_begin = b0 << 4, f_end = min(F, b1 << 4), nFS = (f_end - f_begin) * S;
    const size_t FS = (size_t)F * S;
    // per local block: V (classes 0 .. Fam-1 families, Fam "no family", Fam+1 ones) | PZ (Z zones) | coef [cfg 1 | cfg 3][rho, omega][16]
    const int n_v = (Fam + 2) * S1 * 16, n_pz = d.Z * S1 * 16, blk = n_v + n_pz + 64;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    // the first strip of every thread is requested before the tables are built; a thread walks the strips
    // threadIdx.x + k blockDim.x of the flattened (block, site) range.  (Two and four strips in flight per thread were
    // measured: no faster on few sites, slower at shape D -- 94 / 104 registers.)
    int n = threadIdx.x, lb = 0;
    while (n >= N && lb < nbl) { n -= N; ++lb; }
    uint4 xv = make_uint4(0, 0, 0, 0);
    int cls = Fam + 1;
    if (lb < nbl) {
        xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + ((b0 + lb) << 4)));
        const int fam = d.inheritance ? d.fam_id[n] : -1;
        cls = zone_id[n] >= 0 ? Fam + 1 : (fam >= 0 ? fam : Fam);
    }
    for (int i = threadIdx.x; i < (nbl * blk) >> 1; i += blockDim.x)
        reinterpret_cast<double2*>(fb_tab)[i] = make_double2(1.0, 1.0);
    if (threadIdx.x == 0) {
        int o = 0;
        for (int z = 0; z < d.Z; ++z) { zoff[z] = o; o += st.zone_size[chain * st.s_zone_size + z]; }
        zoff[d.Z] = o;
    }
    __syncthreads();
    if (b1 == nb && threadIdx.x < 16 && threadIdx.x >= F - ((nb - 1) << 4)) {       // padding columns: 1 V + 0 pz
        double* co = fb_tab + (size_t)(nbl - 1) * blk + n_v + n_pz;
        co[16 + threadIdx.x] = co[48 + threadIdx.x] = 0.0;
    }
    {   // thread -> (group, entry i of the CTA's contiguous [features][S] slice); the groups split the classes
        const int groups = max(1, (int)blockDim.x / nFS);
        const double* __restrict__ w = st.weights + chain * st.s_weights;
        for (int i0 = threadIdx.x; i0 < groups * nFS; i0 += blockDim.x) {
            const int g = i0 / nFS, i = i0 - g * nFS;
            const int fi = i / S, sx = i - fi * S, f = f_begin + fi, j = f & 15;
            const size_t o = (size_t)f_begin * S + i;
            double* tab = fb_tab + (size_t)((f >> 4) - b0) * blk;
            // the entries of w~ this thread needs, with wnorm_row's arithmetic: a_c / ((a_0 + a_1) + a_2)
            const double a0 = w[(size_t)f * 3], a1 = w[(size_t)f * 3 + 1];
            const double a2 = d.C == 3 ? w[(size_t)f * 3 + 2] : 0.0;
            const double pg = st.p_global[chain * st.s_p_global + o];
            const double s2 = (a0 + 0.0) + a2;
            const double row6 = a0 / s2, row8 = a2 / s2;                               // cfg 2: family, no zone
            double* dst = tab + sx * 16 + j;
            if (g == 0) {
                dst[(size_t)Fam * S1 * 16] = (a0 / ((a0 + 0.0) + 0.0)) * pg;           // no family: cfg 0
                if (sx == 0) {
                    // zone members: cfg 1 -> w~[1][0] V + w~[1][1] pz;  cfg 3 -> (1 - w~[3][1]) V + w~[3][1] pz with
                    // V = (w~[3][0] pg + w~[3][2] pf) / (w~[3][0] + w~[3][2]) = the cfg-2 value
                    const double s1 = (a0 + a1) + 0.0, s3 = (a0 + a1) + a2;
                    double* co = tab + n_v + n_pz;
                    co[j] = a0 / s1;
                    co[16 + j] = a1 / s1;
                    co[32 + j] = a0 / s3 + a2 / s3;
                    co[48 + j] = a1 / s3;
                }
            }
            // the zone and family tables of this entry, eight loads in flight (they are L2 round trips)
            for (int e0 = g; e0 < d.Z + Fam; e0 += 8 * groups) {
                double p8[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int e = e0 + u * groups;
                    const double* src = e < d.Z ? st.p_zones + chain * st.s_p_zones + (size_t)e * FS
                                                : st.p_fam + chain * st.s_p_fam + (size_t)(e - d.Z) * FS;
                    p8[u] = e < d.Z + Fam ? src[o] : 0.0;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int e = e0 + u * groups;
                    if (e < d.Z) tab[n_v + (size_t)e * S1 * 16 + sx * 16 + j] = p8[u];
                    else if (e < d.Z + Fam) dst[(size_t)(e - d.Z) * S1 * 16] = row6 * pg + row8 * p8[u];
                }
            }
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t tab_s = (uint32_t)__cvta_generic_to_shared(fb_tab), blk_b = (uint32_t)blk * 8u;
    const uint32_t pz_o = (uint32_t)n_v * 8u, co_o = (uint32_t)(n_v + n_pz) * 8u;
    double am = 1.0, side = 0.0;                                   // running product: mantissa in [1, 4), exact logs
    int ae = 0;
    // ---- every (site, block) strip as if the site were in no zone (zone members read the class of ones) ----
    while (lb < nbl) {
        const uint4 xr = rot_strip(xv, lane & 15);
        const uint32_t vrow = tab_s + (uint32_t)lb * blk_b + (uint32_t)(cls * S1) * 128u;
        n += blockDim.x;
        while (n >= N && lb < nbl) { n -= N; ++lb; }
        if (lb < nbl) {                                                                // next strip's loads in flight
            xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + ((b0 + lb) << 4)));
            const int fam = d.inheritance ? d.fam_id[n] : -1;
            cls = zone_id[n] >= 0 ? Fam + 1 : (fam >= 0 ? fam : Fam);
        }
        lh_free_site<false>(xr, lane, S, vrow, 0u, 0u, am, ae, side);
    }
    // ---- zone members, exact ----
    {
        const int* __restrict__ members = st.zone_members + chain * st.s_zone_members;
        const int n_mem = zoff[d.Z];
        for (int idx = threadIdx.x; idx < n_mem * nbl; idx += blockDim.x) {
            const int l = idx / n_mem, m = idx - l * n_mem;
            int z = 0;
            while (m >= zoff[z + 1]) ++z;
            const int ns = members[(size_t)z * N + (m - zoff[z])];
            const int fam = d.inheritance ? d.fam_id[ns] : -1;
            const uint4 xr = rot_strip(__ldg(reinterpret_cast<const uint4*>(d.x + (size_t)ns * d.ldx + ((b0 + l) << 4))), lane & 15);
            const uint32_t base = tab_s + (uint32_t)l * blk_b;
            lh_free_site<true>(xr, lane, S, base + (uint32_t)((fam >= 0 ? fam : Fam) * S1) * 128u,
                               base + pz_o + (uint32_t)(z * S1) * 128u, base + co_o + (fam >= 0 ? 256u : 0u), am, ae, side);
        }
    }
    double v = lp_warp_value(am, ae, side);
    v = block_sum(v, red);
    if (parts == 1) {
        if (threadIdx.x == 0) st.ll[chain] = v;
        return;
    }
    if (threadIdx.x == 0) {
        partial[(size_t)blockIdx.y * parts + blockIdx.x] = v;
        __threadfence();
        is_last = atomicAdd(&ticket[blockIdx.y], 1u) == (unsigned)(parts - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double t = 0.0;
        for (int i = threadIdx.x; i < parts; i += blockDim.x) t += __ldcg(&partial[(size_t)blockIdx.y * parts + i]);
        t = block_sum(t, red);
        if (threadIdx.x == 0) {
            st.ll[chain] = t;
            ticket[blockIdx.y] = 0u;
        }
    }
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

// K4 (full recount), blocked form: one CTA = one (16-feature block, chain).  The block's counters live in shared
// memory as cnt[table][state column][32] (32-bit; tables: the global table once per site configuration cfg = has_zone +
// 2 has_family, Z zones, Fam families, one scrap table for cells whose source has no table; the extra state column
// takes NA / padding cells) and are filled with ONE conflict-free shared-memory atomic per cell (rotated strips,
// sb_common.cuh) while the CTA streams its 16-byte column strip of x and source over all sites.  Shared-memory atomics
// retire ~4 lanes per clock whatever their addresses (ncu: mio_throttle, 10 clocks per warp-wide ATOMS), which is
// what bounds this kernel, so the source table cnt_src[f][cfg][s] is not counted but DERIVED: a site of configuration
// cfg only has the sources its configuration allows, hence  src[cfg][0] = sum_x global_cfg[x],
// src[1][1] = n_sites(cfg 1) - src[1][0],  src[2][2] = n_sites(cfg 2) - src[2][0],  src[3][1] = sum_z,x zone_z[x] - src[1][1],
// src[3][2] = sum_k,x family_k[x] - src[2][2].  The flush writes every entry of the block's rows with plain coalesced
// stores: no global atomics and no zeroing pass.  Same result as count_kernel on every state whose sources are valid
// (count_kernel remains the fallback when the tables of a block do not fit in shared memory).
__host__ __device__ __forceinline__ size_t count_block_smem(int S, int Z, int Fam) {
    return ((size_t)(5 + Z + Fam) * (S + 1) * 32 + 128) * sizeof(int);
}
__global__ void __launch_bounds__(256)
count_block_kernel(DevData d, ChainArrays st, int chain_begin) {
    extern __shared__ __align__(128) int cb_tab[];
    const int chain = chain_begin + blockIdx.y;
    const int f0 = blockIdx.x << 4;
    const int S = d.S, S1 = S + 1, F = d.F, N = d.N;
    const int nf = min(16, F - f0);
    const int n_tab = 4 + d.Z + d.Fam, TS = S1 * 32;                    // tables (without the scrap table), ints per table
    int* aux = cb_tab + (n_tab + 1) * TS;                               // [0..3] sites per cfg, [16 + 6 j ..] sums of feature j
    const int T = (n_tab + 1) * TS + 128;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const uint8_t* __restrict__ src = st.source + chain * st.s_source;
    for (int i = threadIdx.x; i < (T >> 2); i += blockDim.x) reinterpret_cast<int4*>(cb_tab)[i] = make_int4(0, 0, 0, 0);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t cnt_s = (uint32_t)__cvta_generic_to_shared(cb_tab);
    int n = threadIdx.x;
    uint4 xv = make_uint4(0, 0, 0, 0), sv = xv;
    int zid = -1, fam = -1;
    if (n < N) {
        xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
        sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
        zid = zone_id[n];
        fam = d.inheritance ? d.fam_id[n] : -1;
    }
    uint32_t n_cfg = 0;                                                 // four 8-bit tallies (<= N / blockDim sites each)
    while (n < N) {
        const uint4 cr = rot_strip(strip_merge(xv, sv), lane & 15);
        const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
        // first row of the table a cell's source selects: [0] global table of the site's cfg, [1] its zone, [2] its
        // family, [3] scrap
        RowSel rs;
        rs.lo = (uint32_t)(cfg * S1) | ((uint32_t)((zid >= 0 ? 4 + zid : n_tab) * S1) << 16);
        rs.hi = (uint32_t)((fam >= 0 ? 4 + d.Z + fam : n_tab) * S1) | ((uint32_t)(n_tab * S1) << 16);
        n_cfg += 1u << (8 * cfg);
        if ((n_cfg & 0x80808080u) != 0u) {                              // (only with > 32768 sites per chain)
            for (int c = 0; c < 4; ++c) atomicAdd(aux + c, (int)((n_cfg >> (8 * c)) & 0xffu));
            n_cfg = 0;
        }
        n += blockDim.x;
        if (n < N) {                                                                   // next site's loads in flight
            xv = __ldg(reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0));
            sv = *reinterpret_cast<const uint4*>(src + (size_t)n * d.ldx + f0);
            zid = zone_id[n];
            fam = d.inheritance ? d.fam_id[n] : -1;
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const uint32_t col = (uint32_t)((lane + j) & 31) * 4u;                     // loop-invariant per (thread, j)
            const uint32_t c = strip_byte(cr, j);
            const uint32_t xs = min(c & 63u, (uint32_t)S), sc = c >> 6;                // NA / padding -> the extra column
            const uint32_t a_cnt = (cnt_s + ((row_of(rs, sc) & 0xffffu) + xs) * 128u) | col;
            asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a_cnt), "r"(1u) : "memory");
        }
    }
    for (int c = 0; c < 4; ++c)
        if ((n_cfg >> (8 * c)) & 0xffu) atomicAdd(aux + c, (int)((n_cfg >> (8 * c)) & 0xffu));
    __syncthreads();
    const size_t FS = (size_t)F * S;
    // flush: thread -> (group, entry i = j * S + sx of the block's contiguous [nf][S] slice); group 0 also adds the four
    // global tables up, the groups split the zone / family tables.  On the way every thread collects its share of the
    // per-feature sums of the derived source table: aux[16 + 6 j + q], q = 0..3 global table of cfg q, 4 zones, 5 families
    {
        const int nfS = nf * S, groups = max(1, (int)blockDim.x / nfS);
        for (int i0 = threadIdx.x; i0 < groups * nfS; i0 += blockDim.x) {
            const int g = i0 / nfS, i = i0 - g * nfS;
            const int j = i / S, sx = i - j * S;
            const int* c = cb_tab + sx * 32 + j;
            const size_t o = (size_t)f0 * S + i;
            int* sums = aux + 16 + 6 * j;
            if (g == 0) {
                int tot = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int v = c[q * TS] + c[q * TS + 16];
                    if (v) atomicAdd(sums + q, v);
                    tot += v;
                }
                st.cnt_glob[chain * st.s_cnt_glob + o] = tot;
            }
            int sz = 0, sf = 0;
            for (int z = g; z < d.Z; z += groups) {
                const int v = c[(4 + z) * TS] + c[(4 + z) * TS + 16];
                st.cnt_zone[chain * st.s_cnt_zone + (size_t)z * FS + o] = v;
                sz += v;
            }
            for (int k = g; k < d.Fam; k += groups) {
                const int v = c[(4 + d.Z + k) * TS] + c[(4 + d.Z + k) * TS + 16];
                st.cnt_fam[chain * st.s_cnt_fam + (size_t)k * FS + o] = v;
                sf += v;
            }
            if (sz) atomicAdd(sums + 4, sz);
            if (sf) atomicAdd(sums + 5, sf);
        }
        // the NA column of every table also counts towards the source sums: (feature, table slice) per thread
        for (int i = threadIdx.x; i < 16 * 8; i += blockDim.x) {
            const int j = i & 15;
            if (j >= nf) continue;
            const int* c = cb_tab + S * 32 + j;
            int acc[6] = {0, 0, 0, 0, 0, 0};
            for (int t = i >> 4; t < n_tab; t += 8) acc[t < 4 ? t : (t < 4 + d.Z ? 4 : 5)] += c[t * TS] + c[t * TS + 16];
#pragma unroll
            for (int q = 0; q < 6; ++q)
                if (acc[q]) atomicAdd(aux + 16 + 6 * j + q, acc[q]);
        }
    }
    __syncthreads();
    int* cs = st.cnt_src + chain * st.s_cnt_src + (size_t)f0 * 12;
    if (threadIdx.x < nf) {
        const int* a = aux + 16 + 6 * threadIdx.x;
        int* o = cs + threadIdx.x * 12;
        const int s11 = aux[1] - a[1], s22 = aux[2] - a[2];
        o[0] = a[0]; o[1] = 0;           o[2] = 0;
        o[3] = a[1]; o[4] = s11;         o[5] = 0;
        o[6] = a[2]; o[7] = 0;           o[8] = s22;
        o[9] = a[3]; o[10] = a[4] - s11; o[11] = a[5] - s22;
    }
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
