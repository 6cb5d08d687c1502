// sb_lhtab.cuh -- K1 (source-free form, model.py:174-176): table-driven batched mixture log-likelihood.
//
//   LL = sum_{n,f: x != NA} log sum_c w~[n,f,c] * lh[n,f,c]
//
// For a site that is in NO zone the mixture value of cell (n,f) depends only on (f, x_nf, family(n)):
//   v[class][f][s] = w~[f][cfg][0] * p_global[f][s] + w~[f][cfg][2] * p_families[fam][f][s]     (cfg = 2*has_family)
// so the kernel streams the packed state codes once per chain and folds  prod v[class][f][x]  into a running
// product (one table look-up + one fp64 multiply per look-up, one log per thread).  The few sites that ARE in a zone
// (<= Z*MAX_M) are then corrected exactly:  + log prod mix_zone(n,f) - log prod v(n,f).
//
// Static layout (built once at upload, shared by all chains):
//   sites are ordered by class (= family id, "no family" last), each class padded to a multiple of G = 16*CPB sites;
//   a GROUP is G consecutive sites of one class.  xcode[f][group][16]: one byte per CPB sites,
//   code = sum_i digit_i * (S+1)^i, digit = state index, S for NA / padding (table value 1); the byte stores 4*code.
//   CPB = the largest of {1,2,3} with (S+1)^CPB <= 32, so a table row (hi and lo words in separate 32-entry arrays)
//   maps one entry per shared-memory bank: every LDS.32 of the inner loop is conflict-free.
// One thread owns one group and walks the features of its CTA's chunk: 16 B of codes per (group, feature).
#pragma once
#include "sb_common.cuh"
#include "sb_eval.cuh"

#define LT_THREADS 128
#define LT_ROW 32                      // padded entries per table row (hi words, then lo words: 256 B per row)
#define LT_CORR_FB 256                 // features per pass of a correction CTA (its w~ rows live in shared memory)
#ifndef LT_CORR_CTAS
#define LT_CORR_CTAS 2                 // correction CTAs per (chain, split): every one of them costs a fixed ~10 us of
                                       // dependent loads; measured 1 / 2 / 4 / 8 / 16 -> 46 / 39 / 44 / 49 / 68 us at N = 500
#endif
#define LT_TINY_HI 0x3C300000          // hi word of 2^-60: smaller factors (and 0) take the exact log path

struct LhTab {
    int cpb, radix, row_entries, G;    // sites per code byte, S+1, radix^cpb, sites per group (16*cpb)
    int n_groups, n_classes;
    // chosen per launch: features per CTA, groups per CTA (multiple of LT_THREADS), class capacity of the tables
    int fc, n_chunks, gsplit, n_splits, max_cls;
    const uint8_t* xcode;              // [F][n_groups][16]
    const int* group_class;            // [n_groups] ascending
    const int* class_fam;              // [n_classes] family id or -1
    const int* perm;                   // [n_groups*G] site id or -1 (padding)
    const uint32_t* chk_mask;          // [F] applicable states | states that occur in the data
    size_t smem;
};

// shared-memory footprint of one CTA (host + device agree through these helpers)
__host__ __device__ __forceinline__ size_t lt_off_val(int fc) { return ((size_t)fc * 12 * 8 + 255) & ~(size_t)255; }
__host__ __device__ __forceinline__ size_t lt_off_tab(int fc, int max_cls, int S) {
    return (lt_off_val(fc) + (size_t)max_cls * fc * (S + 1) * 8 + 255) & ~(size_t)255;
}
// One site per code byte (S + 1 > 5, `compact`): the value array IS the table -- rows of S + 1 doubles read with one
// LDS.64 -- so the 32-entry hi / lo rows (256 B for S + 1 used entries) are not allocated: 3.5x less shared memory per
// row at S = 8, i.e. 3.5x more features per CTA and 3.5x fewer CTAs that rebuild the rows of every family class.
__host__ __device__ __forceinline__ size_t lt_off_slow(int fc, int max_cls, int S, bool compact = false) {
    return lt_off_tab(fc, max_cls, S) + (compact ? 0 : (size_t)max_cls * fc * 2 * LT_ROW * 4);
}
__host__ __device__ __forceinline__ size_t lt_smem_bytes(int fc, int max_cls, int S, bool compact = false) {
    return lt_off_slow(fc, max_cls, S, compact) + (((size_t)max_cls * fc + 15) & ~(size_t)15);
}

__device__ __forceinline__ double lt_free_value(const DevData& d, const double* __restrict__ wn12,
                                                const double* __restrict__ pg, const double* __restrict__ pf,
                                                int fam, int f, int s) {
    // wn12: w~ rows of feature f, [4][3]
    const size_t o = (size_t)f * d.S + s;
    if (fam >= 0) return wn12[6] * pg[o] + wn12[8] * pf[(size_t)fam * d.F * d.S + o];
    return wn12[0] * pg[o];
}

// 16 look-ups of one (group, feature) unit folded into the running product (m, e)
template <int CPB>
__device__ __forceinline__ void lt_unit(const uint4 xv, uint32_t row, double& m, int& e) {
    constexpr int PERIOD = CPB == 1 ? 16 : (CPB == 2 ? 8 : 5);      // look-ups per partial product
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
    double m0 = m, m1 = 1.0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        // rows are 256-byte aligned and the code bytes hold 4*code, so one PRMT forms the LDS address
        const uint32_t a = __byte_perm(row, xw[j >> 2], 0x3214 + (j & 3));
        uint32_t hi, lo;
        asm("ld.shared.u32 %0, [%1];" : "=r"(hi) : "r"(a));
        asm("ld.shared.u32 %0, [%1+128];" : "=r"(lo) : "r"(a));
        const double tv = __hiloint2double((int)hi, (int)lo);
        if (j & 1) m1 *= tv; else m0 *= tv;
        if ((j + 1) % (2 * PERIOD) == 0 || j == 15) {
            // each partial product holds <= PERIOD factors >= 2^-60*CPB: no underflow before the merge
            const int h0 = __double2hiint(m0), h1 = __double2hiint(m1);
            e += ((h0 >> 20) & 0x7ff) + ((h1 >> 20) & 0x7ff) - 2046;
            m0 = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(m0));
            m1 = __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(m1));
            m0 *= m1;                       // in [1, 4)
            m1 = 1.0;
        }
    }
    m = m0;
}

// the same for one site per code byte and rows of S + 1 doubles (the code byte holds 4 * digit)
__device__ __forceinline__ void lt_unit_compact(const uint4 xv, uint32_t row, double& m, int& e) {
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
    double m0 = m, m1 = 1.0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const uint32_t byte = __byte_perm(xw[j >> 2], 0u, 0x4440 + (j & 3));
        double tv;
        asm("ld.shared.f64 %0, [%1];" : "=d"(tv) : "r"(row + 2u * byte));
        if (j & 1) m1 *= tv; else m0 *= tv;
    }
    // 8 factors >= 2^-60 per partial product on top of a mantissa in [1, 4): no underflow before the merge
    const int h0 = __double2hiint(m0), h1 = __double2hiint(m1);
    e += ((h0 >> 20) & 0x7ff) + ((h1 >> 20) & 0x7ff) - 2046;
    m0 = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(m0));
    m1 = __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(m1));
    m = m0 * m1;
}

// Correction CTA: zone members, + log prod mix_zone - log prod v_free over the member's non-NA cells.  CTA `idx` of
// `nidx` takes members idx, idx + nidx, ... of every zone; one warp per member, lanes over the features (coalesced row
// of x); the member ids of a warp are fetched by its lanes up front so that only one dependent load remains per cell.
__device__ double lt_correction(const DevData& d, const ChainArrays& st, int chain, int idx, int nidx, double* wn /* smem */) {
    LogProd num, den;
    lp_init(num);
    lp_init(den);
    const int S = d.S;
    const size_t FS = (size_t)d.F * S;
    const double* __restrict__ w = st.weights + chain * st.s_weights;
    const double* __restrict__ pg = st.p_global + chain * st.s_p_global;
    const double* __restrict__ pz = st.p_zones + chain * st.s_p_zones;
    const double* __restrict__ pf = st.p_fam + chain * st.s_p_fam;
    const int* zsize = st.zone_size + chain * st.s_zone_size;
    const int* members = st.zone_members + chain * st.s_zone_members;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = LT_THREADS >> 5;
    for (int fb = 0; fb < d.F; fb += LT_CORR_FB) {
        const int nfb = min(LT_CORR_FB, d.F - fb);
        __syncthreads();
        for (int i = threadIdx.x; i < nfb; i += LT_THREADS) wnorm_row(w + (size_t)(fb + i) * 3, wn + (size_t)i * 12, d.C);
        __syncthreads();
        int fill = 0;
        double nm = num.m, dm = den.m;
        for (int z = 0; z < d.Z; ++z) {
            const int zs = zsize[z];
            // this warp's members of zone z: mi = idx + (wid + j*nw) * nidx, j = 0, 1, ...
            for (int j0 = 0; idx + (wid + j0 * nw) * nidx < zs; j0 += 32) {
                const int mi_l = idx + (wid + (j0 + lane) * nw) * nidx;
                int n_l = -1, fam_l = -1;
                if (mi_l < zs) {
                    n_l = members[(size_t)z * d.N + mi_l];
                    if (d.inheritance) fam_l = d.fam_id[n_l];
                }
                const int cnt = __popc(__ballot_sync(0xffffffffu, n_l >= 0));
                for (int j = 0; j < cnt; ++j) {
                    const int n = __shfl_sync(0xffffffffu, n_l, j), fam = __shfl_sync(0xffffffffu, fam_l, j);
                    const uint8_t* __restrict__ xr = d.x + (size_t)n * d.ldx + fb;
                    const double* __restrict__ pzz = pz + (size_t)z * FS + (size_t)fb * S;
                    const double* __restrict__ pff = pf + (size_t)max(fam, 0) * FS + (size_t)fb * S;
                    const double* __restrict__ pgg = pg + (size_t)fb * S;
                    // four feature columns per lane in flight: the state codes first, then the table entries they
                    // select, then the arithmetic -- two dependent global loads per four cells instead of per cell
                    for (int fi0 = lane; fi0 < nfb; fi0 += 128) {
                        int xs4[4];
                        double pg4[4], pz4[4], pf4[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int fi = fi0 + 32 * u;
                            xs4[u] = fi < nfb ? (int)xr[fi] : SB_NA;
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int o = (fi0 + 32 * u) * S + xs4[u];
                            const bool ok = xs4[u] != SB_NA;
                            pg4[u] = ok ? pgg[o] : 1.0;
                            pz4[u] = ok ? pzz[o] : 1.0;
                            pf4[u] = (ok && fam >= 0) ? pff[o] : 0.0;
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (xs4[u] == SB_NA) continue;
                            const double* wr = wn + (size_t)(fi0 + 32 * u) * 12;
                            const double pgv = pg4[u], pzv = pz4[u];
                            double mix, vf;
                            if (fam >= 0) {
                                const double pfv = pf4[u];
                                mix = wr[9] * pgv + wr[10] * pzv + wr[11] * pfv;       // cfg = 3
                                vf = wr[6] * pgv + wr[8] * pfv;                          // cfg = 2
                            } else {
                                mix = wr[3] * pgv + wr[4] * pzv;                         // cfg = 1
                                vf = wr[0] * pgv;                                        // cfg = 0
                            }
                            const bool vf_ok = __double2hiint(vf) >= LT_TINY_HI;   // else: the main pass skipped the cell
                            if (__double2hiint(mix) < LT_TINY_HI) {
                                num.side += log(mix);                                    // exact path (may be -inf)
                                if (!vf_ok) continue;
                            } else {
                                nm *= mix;
                            }
                            if (vf_ok) dm *= vf;
                            if (++fill == 8) {                 // factors in [2^-60, 1]: renormalise every 8
                                fill = 0;
                                const int h0 = __double2hiint(nm), h1 = __double2hiint(dm);
                                num.e += ((h0 >> 20) & 0x7ff) - 1023;
                                den.e += ((h1 >> 20) & 0x7ff) - 1023;
                                nm = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(nm));
                                dm = __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(dm));
                            }
                        }
                    }
                }
            }
        }
        num.m = nm;
        den.m = dm;
        lp_renorm(num);
        lp_renorm(den);
    }
    return lp_value(num) - lp_value(den);
}

template <int CPB>
__global__ void __launch_bounds__(LT_THREADS)
lh_table_kernel(DevData d, ChainArrays st, LhTab t, int chain_begin, double* __restrict__ partial,
                unsigned* __restrict__ ticket) {
    extern __shared__ __align__(256) unsigned char lt_smem[];
    __shared__ double red[33];
    __shared__ int is_last;
    // grid = (group splits, chains, n_corr correction parts + n_chunks feature chunks).  The correction CTAs have the
    // lowest blockIdx.z, so the block scheduler starts ALL of them first: they carry the longest dependent-load chain,
    // and placed at the end of every chain's range they were the tail of the launch (108 -> ~95 us at N = 20000)
    // The parts alternate correction / streaming while both kinds last (z = 0, 2, 4, ... correction), so that the
    // latency-bound correction CTAs share their SMs with streaming CTAs from the start.
    const int n_corr = gridDim.z - t.n_chunks;
    const int chain = chain_begin + blockIdx.y;
    const int split = blockIdx.x;
    const int zi = (int)blockIdx.z, m2 = 2 * min(n_corr, t.n_chunks);
    int corr_idx = -1, chunk = -1;
    if (zi < m2) { if (zi & 1) chunk = zi >> 1; else corr_idx = zi >> 1; }
    else if (n_corr > t.n_chunks) corr_idx = (m2 >> 1) + (zi - m2);
    else chunk = (m2 >> 1) + (zi - m2);
    double v;
    if (chunk < 0) {                // correction CTA
#ifdef LT_NO_CORR
        v = 0.0;
#else
        v = lt_correction(d, st, chain, corr_idx * gridDim.x + split, n_corr * gridDim.x,
                          reinterpret_cast<double*>(lt_smem));
#endif
    } else {
    const int f0 = chunk * t.fc, nf = min(d.F, f0 + t.fc) - f0;
    const int g0 = split * t.gsplit, g1 = min(t.n_groups, g0 + t.gsplit);
    const int c_lo = t.group_class[g0], ncls = t.group_class[g1 - 1] - c_lo + 1;
    const int S = d.S, S1 = S + 1;

    // shared layout: wn [fc][12] f64 | val [max_cls*fc][S+1] f64 | tab [max_cls*fc][2][LT_ROW] u32 | slow [..] u8
    double* wn = reinterpret_cast<double*>(lt_smem);
    double* val = reinterpret_cast<double*>(lt_smem + lt_off_val(t.fc));
    uint32_t* tab = reinterpret_cast<uint32_t*>(lt_smem + lt_off_tab(t.fc, t.max_cls, S));
    constexpr bool COMPACT = CPB == 1;
    unsigned char* slow = lt_smem + lt_off_slow(t.fc, t.max_cls, S, COMPACT);
    // compact: rows of S + 1 doubles inside `val`, laid out [class][nf]; else 256-byte hi / lo rows laid out [class][fc]
    const uint32_t tab_s = (uint32_t)__cvta_generic_to_shared(COMPACT ? (void*)val : (void*)tab);
    const uint32_t row_b = COMPACT ? (uint32_t)S1 * 8u : 2u * LT_ROW * 4u;
    const uint32_t cls_rows = COMPACT ? (uint32_t)nf : (uint32_t)t.fc;

    const double* __restrict__ w = st.weights + chain * st.s_weights;
    const double* __restrict__ pg = st.p_global + chain * st.s_p_global;
    const double* __restrict__ pf = st.p_fam + chain * st.s_p_fam;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;

    // ---- tables of this CTA: rows (class, feature) -------------------------------------------------
    for (int i = threadIdx.x; i < nf; i += LT_THREADS) wnorm_row(w + (size_t)(f0 + i) * 3, wn + (size_t)i * 12, d.C);
    for (int i = threadIdx.x; i < ncls * nf; i += LT_THREADS) slow[i] = 0;
    __syncthreads();
    int tiny = 0;
    for (int i = threadIdx.x; i < ncls * nf * S1; i += LT_THREADS) {
        const int r = i / S1, s = i - r * S1, ci = r / nf, fi = r - ci * nf;
        double v = 1.0;                                                      // NA / padding digit
        if (s < S) {
            v = lt_free_value(d, wn + (size_t)fi * 12, pg, pf, t.class_fam[c_lo + ci], f0 + fi, s);
            // a tiny / zero value at a state that can occur sends the row to the exact cell-by-cell path
            if ((t.chk_mask[f0 + fi] >> s & 1u) && __double2hiint(v) < LT_TINY_HI) { slow[r] = 1; tiny = 1; }
        }
        val[i] = v;
    }
    const int any_slow = __syncthreads_or(tiny);
    if (!COMPACT)
    for (int i = threadIdx.x; i < ncls * nf * LT_ROW; i += LT_THREADS) {
        const int r = i / LT_ROW, code = i - r * LT_ROW, ci = r / nf, fi = r - ci * nf;
        if (code >= t.row_entries) continue;
        const double* vr = val + (size_t)r * S1;
        double v = 1.0;
        int c = code;
#pragma unroll
        for (int k = 0; k < CPB; ++k) {
            const int q = c / t.radix;
            v *= vr[c - q * t.radix];
            c = q;
        }
        uint32_t* row = tab + (size_t)(ci * t.fc + fi) * 2 * LT_ROW;
        row[code] = (uint32_t)__double2hiint(v);
        row[LT_ROW + code] = (uint32_t)__double2loint(v);
    }
    __syncthreads();

    // ---- main pass: every site as if it were in no zone -----------------------------------------
    LogProd acc;
    lp_init(acc);
    const size_t xstride = (size_t)t.n_groups * 16;
    constexpr int PF = 4;                                                // code vectors in flight per thread
    for (int g = g0 + threadIdx.x; g < g1; g += LT_THREADS) {
        const int ci = t.group_class[g] - c_lo;
        const uint8_t* __restrict__ xp = t.xcode + (size_t)g * 16 + (size_t)f0 * xstride;
        uint32_t row = tab_s + (uint32_t)ci * cls_rows * row_b;
        if (!any_slow) {
            uint4 buf[PF];
#pragma unroll
            for (int k = 0; k < PF; ++k) {
                if (k < nf) buf[k] = __ldg(reinterpret_cast<const uint4*>(xp));
                xp += xstride;
            }
            double m = acc.m;
            int e = acc.e;
            for (int fb = 0; fb < nf; fb += PF) {
#pragma unroll
                for (int k = 0; k < PF; ++k) {
                    if (fb + k < nf) {
                        const uint4 xv = buf[k];
                        if (fb + k + PF < nf) buf[k] = __ldg(reinterpret_cast<const uint4*>(xp));
                        xp += xstride;
                        if (COMPACT) lt_unit_compact(xv, row, m, e);
                        else lt_unit<CPB>(xv, row, m, e);
                        row += row_b;
                    }
                }
            }
            acc.m = m;
            acc.e = e;
            continue;
        }
        // general path (some row of this CTA holds a tiny / zero value)
        for (int fi = 0; fi < nf; ++fi, row += row_b) {
            if (!slow[ci * nf + fi]) {
                const uint4 xv = __ldg(reinterpret_cast<const uint4*>(xp + (size_t)fi * xstride));
                if (COMPACT) lt_unit_compact(xv, row, acc.m, acc.e);
                else lt_unit<CPB>(xv, row, acc.m, acc.e);
                continue;
            }
            for (int q = 0; q < t.G; ++q) {
                const int n = t.perm[(size_t)g * t.G + q];
                if (n < 0) continue;
                const int xs = d.x[(size_t)n * d.ldx + f0 + fi];
                if (xs == SB_NA) continue;
                const double v = val[(size_t)(ci * nf + fi) * S1 + xs];
                // a zone member with a tiny free value is left to the correction pass (no -inf - -inf)
                if (zone_id[n] >= 0 && __double2hiint(v) < LT_TINY_HI) continue;
                lp_mul(acc, v);
            }
            lp_renorm(acc);
        }
    }
    v = lp_value(acc);
    }

    v = block_sum(v, red);

    // ---- deterministic per-chain reduction: the last CTA of the chain sums the partials in fixed order ---
    const int n_parts = gridDim.x * gridDim.z;
    const int part = blockIdx.z * gridDim.x + blockIdx.x;
    if (threadIdx.x == 0) {
        partial[(size_t)blockIdx.y * n_parts + part] = v;
        __threadfence();
        is_last = atomicAdd(&ticket[blockIdx.y], 1u) == (unsigned)(n_parts - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double s = 0.0;
        for (int i = threadIdx.x; i < n_parts; i += LT_THREADS) s += __ldcg(&partial[(size_t)blockIdx.y * n_parts + i]);
        s = block_sum(s, red);
        if (threadIdx.x == 0) {
            st.ll[chain] = s;
            ticket[blockIdx.y] = 0u;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// K1c: the table-driven likelihood with ONE CTA PER CHAIN, for launches of many chains over few sites.  Same tables,
// same packed code stream and the same exact treatment of tiny values as lh_table_kernel, but a chain's fixed work is
// done once instead of once per (feature chunk, correction part) CTA: at 256 chains x 2000 x 100 x 4 lh_table_kernel
// executes 70 k warp instructions per chain of which 7 k are table values and member cells (ncu: most lanes of its
// 11 CTAs per chain idle through prologue, reductions and the ticket).  Here
//   * w~ rows of all features, p_global and p_zones (with a column of ones for NA) are staged in shared memory once;
//   * the CTA walks the features in chunks of `fc`: builds the chunk's rows, then streams the chunk's (feature, group)
//     units THREAD-STRIDED -- every thread has work whatever the number of groups -- with the next unit's 16 code bytes
//     in flight;
//   * the zone members are corrected by the same CTA, one warp per member, from the staged tables (only the state code
//     is a global load);
//   * one block reduction, one store: no partial sums, no ticket.
// ---------------------------------------------------------------------------------------------------------------
#define LC_THREADS 512
struct LcLayout { size_t wn, pzg, tab, val, slow, gcls, total; };
__host__ __device__ __forceinline__ LcLayout lc_layout(int F, int S, int Z, int fc, int ncls, int n_groups, bool compact) {
    LcLayout L;
    size_t o = 0;
    L.wn = o;   o += (size_t)F * 12 * 8;
    L.pzg = o;  o += (size_t)(1 + Z) * F * (S + 1) * 8;
    o = (o + 255) & ~(size_t)255;
    L.tab = o;  o += compact ? 0 : (size_t)ncls * fc * 2 * LT_ROW * 4;
    L.val = o;  o += (size_t)ncls * fc * (S + 1) * 8;
    L.slow = o; o += ((size_t)ncls * fc + 15) & ~(size_t)15;
    L.gcls = o; o += ((size_t)n_groups * 2 + 15) & ~(size_t)15;
    L.total = o;
    return L;
}

template <int CPB>
__global__ void __launch_bounds__(LC_THREADS)
lh_chain_kernel(DevData d, ChainArrays st, LhTab t, int chain_begin, int fc) {
    extern __shared__ __align__(256) unsigned char lc_smem[];
    __shared__ double red[33];
    __shared__ int zoff[SB_MAX_Z + 1];
    constexpr bool COMPACT = CPB == 1;
    const int chain = chain_begin + blockIdx.x;
    const int S = d.S, S1 = S + 1, F = d.F, N = d.N, Z = d.Z, ncls = t.n_classes, ng = t.n_groups;
    const size_t FS = (size_t)F * S;
    const LcLayout L = lc_layout(F, S, Z, fc, ncls, ng, COMPACT);
    double* wn = reinterpret_cast<double*>(lc_smem + L.wn);
    double* pzg = reinterpret_cast<double*>(lc_smem + L.pzg);           // [1 + Z][F][S1]: p_global, then the zones
    uint32_t* tab = reinterpret_cast<uint32_t*>(lc_smem + L.tab);
    double* val = reinterpret_cast<double*>(lc_smem + L.val);
    unsigned char* slow = lc_smem + L.slow;
    int16_t* gcls = reinterpret_cast<int16_t*>(lc_smem + L.gcls);
    const double* __restrict__ w = st.weights + chain * st.s_weights;
    const double* __restrict__ pg = st.p_global + chain * st.s_p_global;
    const double* __restrict__ pz = st.p_zones + chain * st.s_p_zones;
    const double* __restrict__ pf = st.p_fam + chain * st.s_p_fam;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const int T = blockDim.x, tid = threadIdx.x;

    // ---- once per chain ----------------------------------------------------------------------------------------
    {
        // the first round of every thread's global loads is issued before anything is computed or stored (the loops
        // below would otherwise pay one memory round trip each)
        const int f_w = tid, i_s = tid;
        double w3[3] = {1.0, 0.0, 0.0}, row0[4] = {1.0, 1.0, 1.0, 1.0};
        if (f_w < F) { w3[0] = w[(size_t)f_w * 3]; w3[1] = w[(size_t)f_w * 3 + 1]; w3[2] = w[(size_t)f_w * 3 + 2]; }
        const bool first_row = i_s < (1 + Z) * F && S <= 4;
        if (first_row) {
            const int tt = i_s / F, f = i_s - tt * F;
            const double* src = (tt == 0 ? pg : pz + (size_t)(tt - 1) * FS) + (size_t)f * S;
            for (int s = 0; s < S; ++s) row0[s] = src[s];
        }
        const int gc0 = tid < ng ? t.group_class[tid] : 0;
        int zs = 0;
        if (tid < Z) zs = st.zone_size[chain * st.s_zone_size + tid];
        if (f_w < F) wnorm_row(w3, wn + (size_t)f_w * 12, d.C);
        for (int f = tid + T; f < F; f += T) wnorm_row(w + (size_t)f * 3, wn + (size_t)f * 12, d.C);
        for (int i = tid; i < (1 + Z) * F; i += T) {                    // one (table, feature) row per thread
            double* dst = pzg + (size_t)i * S1;
            if (i == i_s && first_row) {
                for (int s = 0; s < S; ++s) dst[s] = row0[s];
            } else {
                const int tt = i / F, f = i - tt * F;
                const double* src = (tt == 0 ? pg : pz + (size_t)(tt - 1) * FS) + (size_t)f * S;
                for (int s = 0; s < S; ++s) dst[s] = src[s];
            }
            dst[S] = 1.0;
        }
        if (tid < ng) gcls[tid] = (int16_t)gc0;
        for (int g = tid + T; g < ng; g += T) gcls[g] = (int16_t)t.group_class[g];
        if (tid < Z) zoff[tid + 1] = zs;
    }
    __syncthreads();
    if (tid == 0) {                                                     // sizes -> offsets
        int o = 0;
        for (int z = 0; z < Z; ++z) { const int sz = zoff[z + 1]; zoff[z] = o; o += sz; }
        zoff[Z] = o;
    }
    // Without families a member's correction factor depends on (zone, feature, state) only: the staged zone tables are
    // turned in place into RATIO tables  R[z][f][s] = mix_zone / v_free  (= mix_zone where the main pass left the cell
    // out; -1 where mix_zone itself is tiny: those cells take the exact path), so that a member cell costs one look-up
    // and one multiply.
    const bool ratio = !d.inheritance;
    if (ratio) {
        __syncthreads();
        for (int zf = tid; zf < Z * F; zf += T) {                               // one (zone, feature) row per thread
            const int f = zf % F;
            const double* wr = wn + (size_t)f * 12;
            const double w0 = wr[0], w3 = wr[3], w4 = wr[4];                    // cfg 0 / cfg 1
            double* rz = pzg + (size_t)(F + zf) * S1;
            const double* rg = pzg + (size_t)f * S1;
            for (int s = 0; s < S; ++s) {                                       // (the NA column stays 1)
                const double mix = w3 * rg[s] + w4 * rz[s], vf = w0 * rg[s];
                rz[s] = __double2hiint(mix) < LT_TINY_HI ? -1.0 : (__double2hiint(vf) >= LT_TINY_HI ? mix / vf : mix);
            }
        }
    }
    LogProd acc;
    lp_init(acc);
    const uint32_t tab_s = (uint32_t)__cvta_generic_to_shared(COMPACT ? (void*)val : (void*)tab);
    const uint32_t row_b = COMPACT ? (uint32_t)S1 * 8u : 2u * LT_ROW * 4u;
    const size_t xstride = (size_t)ng * 16;

    for (int f0 = 0; f0 < F; f0 += fc) {
        const int nf = min(fc, F - f0);
        const uint32_t cls_rows = COMPACT ? (uint32_t)nf : (uint32_t)fc;
        __syncthreads();                                                // staged tables ready / previous chunk streamed
        for (int i = tid; i < ncls * nf; i += T) slow[i] = 0;
        __syncthreads();
        // ---- rows (class, feature) of the chunk ----
        int tiny = 0;
        for (int i = tid; i < ncls * nf * S1; i += T) {
            const int r = i / S1, s = i - r * S1, ci = r / nf, fi = r - ci * nf;
            double v = 1.0;                                                      // NA / padding digit
            if (s < S) {
                const int f = f0 + fi, fam = t.class_fam[ci];
                const double pgv = pzg[(size_t)f * S1 + s];
                v = fam >= 0 ? wn[(size_t)f * 12 + 6] * pgv + wn[(size_t)f * 12 + 8] * pf[(size_t)fam * FS + (size_t)f * S + s]
                             : wn[(size_t)f * 12] * pgv;
                // a tiny / zero value at a state that can occur sends the row to the exact cell-by-cell path
                if ((t.chk_mask[f] >> s & 1u) && __double2hiint(v) < LT_TINY_HI) { slow[r] = 1; tiny = 1; }
            }
            val[i] = v;
        }
        const int any_slow = __syncthreads_or(tiny);
        if (!COMPACT) {
            // rows of 32 codes: lane = code (its digits are fixed for the thread), one row per warp and iteration
            const int code = tid & 31;
            int dig[CPB], c = code;
#pragma unroll
            for (int k = 0; k < CPB; ++k) { dig[k] = c % t.radix; c /= t.radix; }
            if (code < t.row_entries) {
                int ci = 0, fi2 = tid >> 5;
                while (fi2 >= nf) { fi2 -= nf; ++ci; }
                while (ci < ncls) {
                    const double* vr = val + (size_t)(ci * nf + fi2) * S1;
                    double v = 1.0;
#pragma unroll
                    for (int k = 0; k < CPB; ++k) v *= vr[dig[k]];
                    uint32_t* row = tab + (size_t)(ci * fc + fi2) * 2 * LT_ROW;
                    row[code] = (uint32_t)__double2hiint(v);
                    row[LT_ROW + code] = (uint32_t)__double2loint(v);
                    fi2 += T >> 5;
                    while (fi2 >= nf && ci < ncls) { fi2 -= nf; ++ci; }
                }
            }
            __syncthreads();
        }
        // ---- stream: (feature, group) units, thread-strided; consecutive threads read consecutive groups ----
        int fi = tid / ng, g = tid - fi * ng;
        const int step_f = T / ng, step_g = T - step_f * ng;            // a thread's next unit is T units further
        // (unit u = fi * ng + g is 16 bytes at xcode + (f0 * ng + u) * 16: the thread's pointer just advances by T units)
        const uint4* xp = reinterpret_cast<const uint4*>(t.xcode + (size_t)f0 * xstride) + tid;
        uint4 xv = make_uint4(0, 0, 0, 0);
        if (fi < nf) xv = __ldg(xp);
        double m = acc.m;
        int e = acc.e;
        while (fi < nf) {
            const int ci = gcls[g];
            const uint32_t row = tab_s + ((uint32_t)ci * cls_rows + (uint32_t)fi) * row_b;
            const uint4 xc = xv;
            const int fi_c = fi, g_c = g;
            fi += step_f;
            g += step_g;
            xp += T;
            if (g >= ng) { g -= ng; ++fi; }
            if (fi < nf) xv = __ldg(xp);
            if (!any_slow || !slow[ci * nf + fi_c]) {
                if (COMPACT) lt_unit_compact(xc, row, m, e);
                else lt_unit<CPB>(xc, row, m, e);
                continue;
            }
            // exact path (the row holds a tiny / zero value)
            acc.m = m; acc.e = e;
            for (int q = 0; q < t.G; ++q) {
                const int n = t.perm[(size_t)g_c * t.G + q];
                if (n < 0) continue;
                const int xs = d.x[(size_t)n * d.ldx + f0 + fi_c];
                if (xs == SB_NA) continue;
                const double v = val[(size_t)(ci * nf + fi_c) * S1 + xs];
                // a zone member with a tiny free value is left to the correction pass (no -inf - -inf)
                if (zone_id[n] >= 0 && __double2hiint(v) < LT_TINY_HI) continue;
                lp_mul(acc, v);
            }
            lp_renorm(acc);
            m = acc.m; e = acc.e;
        }
        acc.m = m;
        acc.e = e;
    }

    // ---- zone members: + log prod mix_zone - log prod v_free, one warp per member, lanes over the features ----
    LogProd num, den;
    lp_init(num);
    lp_init(den);
    {
        const int* __restrict__ members = st.zone_members + chain * st.s_zone_members;
        const int lane = tid & 31, wid = tid >> 5, nw = T >> 5, n_mem = zoff[Z];
        int fill = 0;
        double nm = num.m, dm = den.m;
        // the warp's members: wid, wid + nw, ...; their ids are fetched by the lanes up front, and the state codes of
        // the next member are in flight while the current one is processed
        const int my_cnt = wid < n_mem ? (n_mem - wid + nw - 1) / nw : 0;
        for (int j0 = 0; j0 < my_cnt; j0 += 32) {
        int n_l = 0, z_l = 0, fam_l = -1;
        {
            const int mi_l = wid + (j0 + lane) * nw;
            if (mi_l < n_mem) {
                while (mi_l >= zoff[z_l + 1]) ++z_l;
                n_l = members[(size_t)z_l * N + (mi_l - zoff[z_l])];
                if (d.inheritance) fam_l = d.fam_id[n_l];
            }
        }
        const int cnt = min(32, my_cnt - j0);
        int xs_nxt[4];
        {
            const uint8_t* __restrict__ x0 = d.x + (size_t)__shfl_sync(0xffffffffu, n_l, 0) * d.ldx;
#pragma unroll
            for (int u = 0; u < 4; ++u) xs_nxt[u] = lane + 32 * u < F ? (int)x0[lane + 32 * u] : SB_NA;
        }
        for (int j = 0; j < cnt; ++j) {
            const int n = __shfl_sync(0xffffffffu, n_l, j), z = __shfl_sync(0xffffffffu, z_l, j);
            const int fam = __shfl_sync(0xffffffffu, fam_l, j);
            int xs_cur[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) xs_cur[u] = xs_nxt[u];
            if (j + 1 < cnt) {
                const uint8_t* __restrict__ x1 = d.x + (size_t)__shfl_sync(0xffffffffu, n_l, j + 1) * d.ldx;
#pragma unroll
                for (int u = 0; u < 4; ++u) xs_nxt[u] = lane + 32 * u < F ? (int)x1[lane + 32 * u] : SB_NA;
            }
            const uint8_t* __restrict__ xr = d.x + (size_t)n * d.ldx;
            const double* pzz = pzg + (size_t)(1 + z) * F * S1;
            const double* pff = pf + (size_t)max(fam, 0) * FS;
            for (int f_0 = lane; f_0 < F; f_0 += 128) {
                int xs4[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    xs4[u] = f_0 == lane ? xs_cur[u] : (f_0 + 32 * u < F ? (int)xr[f_0 + 32 * u] : SB_NA);
                if (ratio) {
                    // four branch-free look-ups (NA, padding and columns beyond F read a ratio of 1); the exact code
                    // below only runs if one of the four is flagged
                    double r4[4];
                    int lo_hi = 0x7fffffff;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        r4[u] = pzz[(size_t)min(f_0 + 32 * u, F - 1) * S1 + min(xs4[u], S)];
                        lo_hi = min(lo_hi, __double2hiint(r4[u]));
                    }
                    if (lo_hi >= LT_TINY_HI) {
                        nm *= (r4[0] * r4[1]) * (r4[2] * r4[3]);                // factors in [2^-60, 2^60]
                        const int h0 = __double2hiint(nm);
                        num.e += ((h0 >> 20) & 0x7ff) - 1023;
                        nm = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(nm));
                        continue;
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (xs4[u] == SB_NA) continue;
                    const int f = f_0 + 32 * u;
                    const double* wr = wn + (size_t)f * 12;
                    double pzv = pzz[(size_t)f * S1 + xs4[u]];
                    if (ratio) {
                        if (__double2hiint(pzv) >= LT_TINY_HI) {        // the staged ratio: factors in [2^-60, 2^60]
                            nm *= pzv;
                            if (++fill == 8) {                          // (dm too: exact cells may have used it)
                                fill = 0;
                                const int h0 = __double2hiint(nm), h1 = __double2hiint(dm);
                                num.e += ((h0 >> 20) & 0x7ff) - 1023;
                                den.e += ((h1 >> 20) & 0x7ff) - 1023;
                                nm = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(nm));
                                dm = __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(dm));
                            }
                            continue;
                        }
                        pzv = pz[(size_t)z * FS + (size_t)f * S + xs4[u]];      // exact path below, from the chain's table
                    }
                    const double pgv = pzg[(size_t)f * S1 + xs4[u]];
                    double mix, vf;
                    if (fam >= 0) {
                        const double pfv = pff[(size_t)f * S + xs4[u]];
                        mix = wr[9] * pgv + wr[10] * pzv + wr[11] * pfv;       // cfg = 3
                        vf = wr[6] * pgv + wr[8] * pfv;                          // cfg = 2
                    } else {
                        mix = wr[3] * pgv + wr[4] * pzv;                         // cfg = 1
                        vf = wr[0] * pgv;                                        // cfg = 0
                    }
                    const bool vf_ok = __double2hiint(vf) >= LT_TINY_HI;   // else: the main pass skipped the cell
                    if (__double2hiint(mix) < LT_TINY_HI) {
                        num.side += log(mix);                                    // exact path (may be -inf)
                        if (!vf_ok) continue;
                    } else {
                        nm *= mix;
                    }
                    if (vf_ok) dm *= vf;
                    if (++fill == 8) {                 // factors in [2^-60, 1]: renormalise every 8
                        fill = 0;
                        const int h0 = __double2hiint(nm), h1 = __double2hiint(dm);
                        num.e += ((h0 >> 20) & 0x7ff) - 1023;
                        den.e += ((h1 >> 20) & 0x7ff) - 1023;
                        nm = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(nm));
                        dm = __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(dm));
                    }
                }
            }
        }
        }
        num.m = nm;
        den.m = dm;
    }
    // one log per warp: the lanes' mantissas (stream x member numerators / member denominators, each in [1/2, 8)) are
    // multiplied across the warp first
    lp_renorm(acc);
    lp_renorm(num);
    lp_renorm(den);
    double v = lp_warp_value(acc.m * num.m / den.m, acc.e + num.e - den.e, acc.side + num.side - den.side);
    v = block_sum(v, red);
    if (tid == 0) st.ll[chain] = v;
}
