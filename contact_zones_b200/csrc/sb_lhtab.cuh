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
#define LT_TINY_HI 0x3C300000          // hi word of 2^-60: smaller factors (and 0) take the exact log path

struct LhTab {
    int cpb, radix, row_entries, G;    // sites per code byte, S+1, radix^cpb, sites per group (16*cpb)
    int n_groups, n_classes, n_tiles, fc, n_chunks, max_cls_tile, fc_max;   // fc / n_chunks are chosen per launch
    const uint8_t* xcode;              // [F][n_groups][16]
    const int* group_class;            // [n_groups] ascending
    const int* class_fam;              // [n_classes] family id or -1
    const int* perm;                   // [n_groups*G] site id or -1 (padding)
    const uint32_t* chk_mask;          // [F] applicable states | states that occur in the data
    size_t smem;
};

__device__ __forceinline__ double lt_free_value(const DevData& d, const double* __restrict__ wn12,
                                                const double* __restrict__ pg, const double* __restrict__ pf,
                                                int fam, int f, int s) {
    // wn12: w~ rows of feature f, [4][3]
    const size_t o = (size_t)f * d.S + s;
    if (fam >= 0) return wn12[6] * pg[o] + wn12[8] * pf[(size_t)fam * d.F * d.S + o];
    return wn12[0] * pg[o];
}

template <int CPB>
__global__ void __launch_bounds__(LT_THREADS)
lh_table_kernel(DevData d, ChainArrays st, LhTab t, int chain_begin, double* __restrict__ partial,
                unsigned* __restrict__ ticket) {
    extern __shared__ __align__(256) unsigned char lt_smem[];
    __shared__ double red[33];
    __shared__ int is_last;
    const int chain = chain_begin + blockIdx.z;
    const int tile = blockIdx.x, chunk = blockIdx.y;
    const int f0 = chunk * t.fc, nf = min(d.F, f0 + t.fc) - f0;
    const int g0 = tile * LT_THREADS, g1 = min(t.n_groups, g0 + LT_THREADS);
    const int c_lo = t.group_class[g0], ncls = t.group_class[g1 - 1] - c_lo + 1;

    // shared layout: wn [fc][12] f64 | tab [max_cls_tile][fc][2][LT_ROW] u32 (256-byte aligned rows) | slow [..] u8
    double* wn = reinterpret_cast<double*>(lt_smem);
    uint32_t* tab = reinterpret_cast<uint32_t*>(lt_smem + (((size_t)t.fc * 12 * sizeof(double) + 255) & ~(size_t)255));
    unsigned char* slow = reinterpret_cast<unsigned char*>(tab + (size_t)t.max_cls_tile * t.fc * 2 * LT_ROW);
    const uint32_t tab_s = (uint32_t)__cvta_generic_to_shared(tab);

    const double* __restrict__ w = st.weights + chain * st.s_weights;
    const double* __restrict__ pg = st.p_global + chain * st.s_p_global;
    const double* __restrict__ pz = st.p_zones + chain * st.s_p_zones;
    const double* __restrict__ pf = st.p_fam + chain * st.s_p_fam;
    const int8_t* __restrict__ zone_id = st.zone_id + chain * st.s_zone_id;
    const int S = d.S;

    for (int i = threadIdx.x; i < nf; i += LT_THREADS) wnorm_row(w + (size_t)(f0 + i) * 3, wn + (size_t)i * 12, d.C);
    __syncthreads();
    // one thread builds one table row (class, feature): the S+1 cell values, then their CPB-fold products
    for (int r = threadIdx.x; r < ncls * nf; r += LT_THREADS) {
        const int ci = r / nf, fi = r - ci * nf;
        const int fam = t.class_fam[c_lo + ci];
        const uint32_t chk = t.chk_mask[f0 + fi];
        double val[SB_MAX_S + 1];
        bool tiny = false;
        for (int s = 0; s < S; ++s) {
            val[s] = lt_free_value(d, wn + (size_t)fi * 12, pg, pf, fam, f0 + fi, s);
            // a tiny / zero value at a state that can occur sends the row to the exact cell-by-cell path
            if ((chk >> s & 1u) && __double2hiint(val[s]) < LT_TINY_HI) tiny = true;
        }
        val[S] = 1.0;
        slow[ci * nf + fi] = tiny ? 1 : 0;
        uint32_t* row = tab + (size_t)(ci * t.fc + fi) * 2 * LT_ROW;
        int code = 0;
        const int R = t.radix;
        for (int a2 = 0; a2 < (CPB >= 3 ? R : 1); ++a2)
            for (int a1 = 0; a1 < (CPB >= 2 ? R : 1); ++a1)
                for (int a0 = 0; a0 < R; ++a0, ++code) {
                    double v = val[a0];
                    if (CPB >= 2) v *= val[a1];
                    if (CPB >= 3) v *= val[a2];
                    row[code] = (uint32_t)__double2hiint(v);
                    row[LT_ROW + code] = (uint32_t)__double2loint(v);
                }
    }
    __syncthreads();

    // ---- main pass: every site as if it were in no zone -----------------------------------------
    LogProd acc;
    lp_init(acc);
    const int g = g0 + threadIdx.x;
    if (g < g1) {
        const int ci = t.group_class[g] - c_lo;
        const size_t xstride = (size_t)t.n_groups * 16;
        const uint8_t* __restrict__ xp = t.xcode + (size_t)g * 16 + (size_t)f0 * xstride;
        constexpr int PERIOD = CPB == 1 ? 16 : (CPB == 2 ? 8 : 5);      // look-ups between renormalisations
        constexpr int PF = 4;                                            // code vectors in flight per thread
        uint4 buf[PF];
#pragma unroll
        for (int k = 0; k < PF; ++k)
            if (k < nf) buf[k] = __ldg(reinterpret_cast<const uint4*>(xp + (size_t)k * xstride));
        for (int fb = 0; fb < nf; fb += PF) {
#pragma unroll
            for (int k = 0; k < PF; ++k) {
                const int fi = fb + k;
                if (fi >= nf) break;
                const uint4 xv = buf[k];
                if (fi + PF < nf) buf[k] = __ldg(reinterpret_cast<const uint4*>(xp + (size_t)(fi + PF) * xstride));
                if (slow[ci * nf + fi]) {
                    const int fam = t.class_fam[c_lo + ci];
                    for (int q = 0; q < t.G; ++q) {
                        const int n = t.perm[(size_t)g * t.G + q];
                        if (n < 0) continue;
                        const int xs = d.x[(size_t)n * d.ldx + f0 + fi];
                        if (xs == SB_NA) continue;
                        const double v = lt_free_value(d, wn + (size_t)fi * 12, pg, pf, fam, f0 + fi, xs);
                        // a zone member with a tiny free value is left to the correction pass (no -inf - -inf)
                        if (zone_id[n] >= 0 && __double2hiint(v) < LT_TINY_HI) continue;
                        lp_mul(acc, v);
                    }
                    continue;
                }
                // rows are 256-byte aligned and the code bytes hold 4*code, so one PRMT forms the LDS address
                const uint32_t row = tab_s + (uint32_t)(ci * t.fc + fi) * (2 * LT_ROW * 4);
                const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
                double m0 = acc.m, m1 = 1.0;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const uint32_t a = __byte_perm(row, xw[j >> 2], 0x3214 + (j & 3));
                    uint32_t hi, lo;
                    asm volatile("ld.shared.u32 %0, [%2];\n\tld.shared.u32 %1, [%2+128];" : "=r"(hi), "=r"(lo) : "r"(a));
                    const double tv = __hiloint2double((int)hi, (int)lo);
                    if (j & 1) m1 *= tv; else m0 *= tv;
                    if ((j + 1) % (2 * PERIOD) == 0 || j == 15) {
                        // each partial product holds <= PERIOD factors >= 2^-60*CPB: no underflow before the merge
                        const int h0 = __double2hiint(m0), h1 = __double2hiint(m1);
                        acc.e += ((h0 >> 20) & 0x7ff) + ((h1 >> 20) & 0x7ff) - 2046;
                        m0 = __hiloint2double((h0 & 0x800fffff) | 0x3ff00000, __double2loint(m0));
                        m1 = __hiloint2double((h1 & 0x800fffff) | 0x3ff00000, __double2loint(m1));
                        m0 *= m1;                       // in [1, 4)
                        m1 = 1.0;
                    }
                }
                acc.m = m0;
                acc.k = 0;
            }
        }
    }
    double v = lp_value(acc);

    // ---- correction pass: zone members, + log mix_zone - log v_free.  Member i of a zone goes to tile i % n_tiles;
    //      one warp per member, lanes over the features of the chunk (coalesced x row) ------------------------
    if (d.Z > 0) {
        LogProd num, den;
        lp_init(num);
        lp_init(den);
        const int* zsize = st.zone_size + chain * st.s_zone_size;
        const int* members = st.zone_members + chain * st.s_zone_members;
        const size_t FS = (size_t)d.F * S;
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = LT_THREADS >> 5;
        const int n_tiles = gridDim.x;
        for (int z = 0; z < d.Z; ++z) {
            const int zs = zsize[z];
            for (int mi = tile + wid * n_tiles; mi < zs; mi += nw * n_tiles) {
                const int n = members[(size_t)z * d.N + mi];
                const int fam = d.inheritance ? d.fam_id[n] : -1;
                for (int fi = lane; fi < nf; fi += 32) {
                    const int f = f0 + fi;
                    const int xs = d.x[(size_t)n * d.ldx + f];
                    if (xs == SB_NA) continue;
                    const double* wr = wn + (size_t)fi * 12 + (fam >= 0 ? 9 : 3);      // cfg = 1 + 2*has_family
                    const size_t o = (size_t)f * S + xs;
                    double mix = wr[0] * pg[o] + wr[1] * pz[(size_t)z * FS + o];
                    if (fam >= 0) mix += wr[2] * pf[(size_t)fam * FS + o];
                    lp_mul(num, mix);
                    const double vf = lt_free_value(d, wn + (size_t)fi * 12, pg, pf, fam, f, xs);
                    if (__double2hiint(vf) >= LT_TINY_HI) lp_mul(den, vf);          // else: the main pass skipped it
                }
            }
        }
        v += lp_value(num) - lp_value(den);
    }
    v = block_sum(v, red);

    // ---- deterministic per-chain reduction: the last CTA of the chain sums the partials in fixed order ---
    const int n_parts = gridDim.x * gridDim.y;
    const int part = blockIdx.y * gridDim.x + blockIdx.x;
    if (threadIdx.x == 0) {
        partial[(size_t)blockIdx.z * n_parts + part] = v;
        __threadfence();
        is_last = atomicAdd(&ticket[blockIdx.z], 1u) == (unsigned)(n_parts - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double s = 0.0;
        for (int i = threadIdx.x; i < n_parts; i += LT_THREADS) s += __ldcg(&partial[(size_t)blockIdx.z * n_parts + i]);
        s = block_sum(s, red);
        if (threadIdx.x == 0) {
            st.ll[chain] = s;
            ticket[blockIdx.z] = 0u;
        }
    }
}
