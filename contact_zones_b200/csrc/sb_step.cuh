// sb_step.cuh -- K5/K7/K8: the persistent per-chain MCMC step kernel.
//
// A team of CTAs (one thread-block cluster, see `Team`) owns one chain for the whole launch and runs `n_steps`
// Metropolis-Hastings / Gibbs steps (MCMCGenerative.step, mcmc_generative.py:282-329) without any host round-trip.  All proposals are DELTAS
// on the resident chain state (no Sample.copy(), zone_sampling.py:95-115):
//   zone moves      O(F*C) marginal-likelihood ratio of the moved site(s)       (K5)
//   alter_*         O(N) / O(|zone|) / O(|family|) column delta on feature f    (K5)
//   gibbs_*         O(F*S) draws from the integer count tables                   (K4 tables)
// and only an ACCEPTED zone move (or gibbs_sample_sources) pays the O(N*F) source-resampling pass that the
// reference pays three times per zone proposal (zone_sampling.py:801-809,859-862).
#pragma once
#include "sb_common.cuh"
#include "sb_rng.cuh"
#include "sb_eval.cuh"
#include "../../include/sbayes_b200.h"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#ifndef STEP_THREADS
#define STEP_THREADS 512     // threads per CTA of the step kernel
#endif
#ifndef STEP_MIN_CTAS
#define STEP_MIN_CTAS 1       // CTAs the register budget must let share an SM
#endif

// optional phase timing (build with -DSB_PROF): cycles of CTA 0 / thread 0 accumulated per marker
#ifdef SB_PROF
__device__ unsigned long long sb_prof[16];
#define PROF_T(var) long long var = clock64()
#define PROF_ADD(slot, a, b) do { if (threadIdx.x == 0 && blockIdx.x == 0) sb_prof[slot] += (unsigned long long)((b) - (a)); } while (0)
#else
#define PROF_T(var)
#define PROF_ADD(slot, a, b)
#endif

struct StepParams {
    int n_ops;
    int op_ids[SB_N_OPS];
    double op_cdf[SB_N_OPS];
    double prec_w, prec_g, prec_z, prec_f;
    int min_size;
    double frac;
    uint32_t seed_lo, seed_hi;
    int chain_offset;
    const double* p_grow;     // [chains]
    const int* max_size;      // [chains]
    int nwords;               // words of one site bitmap
    int prim_cap;             // member capacity of the Prim scratch
    int s_shift;              // log2 of the smallest power of two >= S (S <= SB_MAX_S = 32)
    int team;                 // CTAs per chain (thread-block cluster size): 1, 2, 4 or 8
    int work_off;             // byte offset of the shared work area in the dynamic smem (16 B aligned)
    int work_counts;          // != 0: resample_all_sources keeps per-feature-block count tables in the work area;
                              // 2: the block's parameter tables are staged behind them as well (par_off)
    int par_off;              // byte offset of the staged parameter tables (8 B aligned)
    // zone cache of the proposal phase (each CTA of a team keeps its own, identical copy in shared memory):
    int zc_off;               // byte offset of { int zsize[Z]; int members[Z][zc_cap]; int8 zone_id[N] }, or -1
    int zc_cap;               // member capacity per zone of the cache (> every max_size)
    int adj_off;              // byte offset of the adjacency CSR { int ptr[N+1]; int idx[nnz] } in shared memory, or -1
    int* tmp_members;         // [chains][N]     candidate member list for the geo prior
    size_t s_tmp;
    int from_prior;           // sample_from_prior (mcmc_generative.py:115-116): the likelihood is identically 0
    struct ForcedMove* forced;   // non-null: evaluate ONE forced zone-move proposal and return (sb_move_log_q)
    int forced_chain;         // the chain whose team evaluates it
    // host-visible events inside a launch (logged samples, MC3 swap points): after `events[k].x` completed steps of the
    // launch every chain records (ll, prior) in ev_scalars[k][chain]; if events[k].y >= 0 the chain also copies its
    // state into snapshot slot events[k].y of the ring
    const int2* events;
    int n_events;
    double2* ev_scalars;      // [n_events][chains]
    unsigned char* ring;      // [slots][chains][ring_rec] snapshot records
    size_t ring_rec;          // bytes of one record: zone_id | weights | p_global | p_zones | p_fam  (16 B aligned parts)
    int ring_chain;           // local chain whose states are recorded, or -1 = every chain (MC3: the logged label moves)
};
// forced zone-move proposal (parity surface of zone_sampling.py:704-933, 1328-1574): the proposal code of op_zone_move
// runs with the zone / sites given instead of drawn and stores log q, log q_back
struct ForcedMove {
    int kind, zone, site_new, site_removed;
    double rho;
    double out[2];
};

// ---- shared scratch of the step kernel -----------------------------------------------------------
#define STEP_WARPS (STEP_THREADS / 32)
#ifndef SB_AHEAD
#define SB_AHEAD 128             // steps per refill of the look-ahead table (multiple of 32)
#endif
// uniforms that depend on the step number only, drawn SB_AHEAD steps at a time (index into StepShared::u_ahead)
enum { UA_ZONE = 0, UA_FAMILY, UA_FIXED, UA_CONNECTED, UA_SITE_NEW, UA_SITE_REMOVED, UA_ACCEPT, UA_FEATURE, UA_PAIR_A,
       UA_PAIR_B, SB_N_AHEAD };
static_assert(SB_AHEAD % 32 == 0, "look-ahead refill works on slabs of 32 steps");
__host__ __device__ constexpr uint32_t ahead_purpose(int q) {
    return q == UA_ZONE ? RP_ZONE : q == UA_FAMILY ? RP_FAMILY : q == UA_FIXED ? RP_FIXED : q == UA_CONNECTED ? RP_CONNECTED
         : q == UA_SITE_NEW ? RP_SITE_NEW : q == UA_SITE_REMOVED ? RP_SITE_REMOVED : q == UA_ACCEPT ? RP_ACCEPT
         : q == UA_FEATURE ? RP_FEATURE : RP_PAIR;
}
__host__ __device__ constexpr uint32_t ahead_index(int q) { return q == UA_PAIR_B ? 1u : 0u; }   // counter index of the draw
struct StepShared {
    double red[33];
    int redi[33];
    int pick;           // bits_select result
    double dshare[4];
    double wpart[2][2][STEP_WARPS];   // per-warp partial sums of the team reductions (double-buffered pair), read by peers
    double wcta[2][2][STEP_WARPS];    // the same for CTA-scope sums: never read by a peer CTA, own parity (Team::cturn)
    // per-step scalar draws of the next SB_AHEAD steps (operator id and the uniforms behind "which zone / family /
    // fixed component"), filled by four warps once per SB_AHEAD steps instead of by every thread in every step
    int op_ahead[SB_AHEAD];
    double u_ahead[SB_N_AHEAD][SB_AHEAD];
};

// ---- site bitmaps in shared memory -----------------------------------------------------------
__device__ __forceinline__ void bits_clear(unsigned* bits, int nwords) {
    for (int i = threadIdx.x; i < nwords; i += blockDim.x) bits[i] = 0u;
}
__device__ __forceinline__ bool bit_test(const unsigned* bits, int n) { return bits[n >> 5] >> (n & 31) & 1u; }

// number of set bits (all threads get the result)
__device__ __forceinline__ int bits_count(const unsigned* bits, int nwords, int* redi) {
    int c = 0;
    for (int i = threadIdx.x; i < nwords; i += blockDim.x) c += __popc(bits[i]);
    return block_sum_i(c, redi);
}

// index of the k-th (0-based) set bit, ascending site order == order of `candidates.nonzero()[0]`
// (zone_sampling.py:744,836).  Executed by warp 0; result broadcast through *out.  Caller syncs.
// With `total` non-null the number of set bits n is stored there and the bit picked is floor(u * n) (capped at
// n - 1; nothing is picked if n == 0); otherwise k is used as given.
__device__ void bits_select(const unsigned* bits, int nwords, int k, int* out, int* total = nullptr, double u = 0.0) {
    if (threadIdx.x >= 32) return;
    const int lane = threadIdx.x;
    const int per = (nwords + 31) / 32;
    const int b = lane * per, e = min(nwords, b + per);
    int c = 0;
    for (int i = b; i < e; ++i) c += __popc(bits[i]);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (total) {
        const int n = __shfl_sync(0xffffffffu, incl, 31);
        if (lane == 0) *total = n;
        k = ahead_int(u, n);            // -1 if n == 0: no lane matches below
    }
    const int excl = incl - c;
    if (k >= excl && k < incl) {
        int r = k - excl;
        for (int i = b; i < e; ++i) {
            const int pc = __popc(bits[i]);
            if (r < pc) {
                unsigned w = bits[i];
                for (int j = 0; j < r; ++j) w &= w - 1;      // drop the r lowest set bits
                *out = (i << 5) + (__ffs(w) - 1);
                break;
            }
            r -= pc;
        }
    }
}

// free[n] = site n is in no zone  (== ~occupied, zone_sampling.py:714,799)
__device__ void build_free_bits(const DevData& d, const int8_t* zone_id, unsigned* bits, int nwords) {
    for (int w = threadIdx.x; w < nwords; w += blockDim.x) {
        unsigned v = 0u;
        const int n0 = w << 5;
        for (int j = 0; j < 32; ++j) {
            const int n = n0 + j;
            if (n < d.N && zone_id[n] < 0) v |= 1u << j;
        }
        bits[w] = v;
    }
}

// neighbours of zone z: (adj_mat @ zone) & ~occupied (util.py:139-155).  `skip` (-1 = none) is a member that is
// treated as already removed (its adjacency does not count and it is itself free) -- the back step of shrink_zone
// (zone_sampling.py:906-908).
__device__ void build_nbr_bits(const int* __restrict__ adj_ptr, const int* __restrict__ adj_idx, const int8_t* zone_id,
                               const int* members, int size, int skip, unsigned* bits, int nwords) {
    bits_clear(bits, nwords);
    __syncthreads();
    // one warp per member, lanes over its adjacency row
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int i = w; i < size; i += nw) {
        const int m = members[i];
        if (m == skip) continue;
        const int b = adj_ptr[m], e = adj_ptr[m + 1];
        for (int j = b + lane; j < e; j += 32) {
            const int n = adj_idx[j];
            if (zone_id[n] < 0 || n == skip) atomicOr(&bits[n >> 5], 1u << (n & 31));
        }
    }
    __syncthreads();
}

// ---- a chain is owned by a TEAM of CTAs (one thread-block cluster) -----------------------------------------------
// With n_chains <= SMs / 2 a single CTA per chain leaves most of the GPU idle, and the step kernel is bound by the
// instruction issue rate of its one SM.  A team of `size` CTAs splits every data-parallel phase of a step (Gamma draws,
// table updates, the source sweep, column deltas); all CTAs take the same block-uniform decisions (counter-based RNG),
// keep identical running ll / prior in registers, and meet in cluster barriers that also order their global-memory
// traffic: between a write by one CTA and a read of that location by another there is always a team barrier.
struct Team {
    int rank, size;
    unsigned turn;      // parity of the reduction buffer (advances identically in every CTA)
    int shift;          // log2(size)
    unsigned cturn;     // parity of the CTA-scope reduction buffer
};
__device__ __forceinline__ void team_sync(const Team& t) {
    if (t.size > 1) cg::this_cluster().sync(); else __syncthreads();
}
// Execution barrier of the team without memory ordering beyond the CTA: enough between a step that only READ the
// chain state and the next step's writes; unlike cluster.sync() it does not invalidate L1 (no CCTL.IVALL).
__device__ __forceinline__ void team_sync_exec(const Team& t) {
    __syncthreads();
    if (t.size > 1) {
        asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
    }
}
// contiguous share [lo, hi) of n items for this CTA
__device__ __forceinline__ void team_share(const Team& t, int n, int& lo, int& hi) {
    const int per = (n + t.size - 1) >> t.shift;
    lo = min(n, t.rank * per);
    hi = min(n, lo + per);
}
// Sum of a and b over all threads of `scope` CTAs (1 = this CTA only, else the whole team), identical in every
// thread: warp tree, then the per-warp partials of all CTAs in fixed (rank, warp) order.  ONE barrier -- for the team
// scope a cluster barrier, so everything the CTAs wrote before the call is visible to all of them after it.  The
// partials are double-buffered: a buffer is rewritten two reductions later, after the barrier in between.
// Out of line: the reductions sit on the Metropolis paths only (no Gibbs step reduces), and ~30 inlined copies of the
// shuffle trees were a quarter of the kernel's code.
__device__ __noinline__ double2 sum2_impl(double a, double b, double* wpart, unsigned turn, int scope) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    a = warp_sum(a);
    b = warp_sum(b);
    double* slot = wpart + (turn & 1u) * (2 * STEP_WARPS);
    if (lane == 0) { slot[w] = a; slot[STEP_WARPS + w] = b; }
    double sa = 0.0, sb = 0.0;
    if (scope > 1) {
        cg::cluster_group cl = cg::this_cluster();
        PROF_T(tb0);
        cl.sync();
        PROF_T(tb1);
        PROF_ADD(11, tb0, tb1);
        for (int r = 0; r < scope; ++r) {
            const double* peer = cl.map_shared_rank(slot, r);
            sa += warp_sum(lane < STEP_WARPS ? peer[lane] : 0.0);
            sb += warp_sum(lane < STEP_WARPS ? peer[STEP_WARPS + lane] : 0.0);
        }
    } else {
        __syncthreads();
        sa = warp_sum(lane < STEP_WARPS ? slot[lane] : 0.0);
        sb = warp_sum(lane < STEP_WARPS ? slot[STEP_WARPS + lane] : 0.0);
    }
    return make_double2(sa, sb);
}
__device__ __forceinline__ void sum2_scope(double& a, double& b, StepShared& sh, Team& t, int scope) {
    // CTA-scope sums use their own double-buffered scratch and parity: they meet in a __syncthreads only, so they must
    // not rewrite a slot that a slower peer CTA may still be reading through DSMEM
    const bool cta = scope <= 1;
    const double2 r = sum2_impl(a, b, cta ? &sh.wcta[0][0][0] : &sh.wpart[0][0][0], cta ? t.cturn : t.turn, scope);
    if (cta) ++t.cturn; else ++t.turn;
    a = r.x; b = r.y;
}
__device__ __forceinline__ void team_sum2(double& a, double& b, StepShared& sh, Team& t) { sum2_scope(a, b, sh, t, t.size); }
__device__ __forceinline__ double team_sum(double v, StepShared& sh, Team& t) {
    double z = 0.0;
    sum2_scope(v, z, sh, t, t.size);
    return v;
}
__device__ __forceinline__ double cta_sum(double v, StepShared& sh, Team& t) {      // this CTA only
    double z = 0.0;
    sum2_scope(v, z, sh, t, 1);
    return v;
}

// ---- per-chain view ------------------------------------------------------------------------------
struct Chain {
    int8_t* zone_id; int* members; int* zsize;
    double* w; double* wn; double* pg; double* pz; double* pf;
    uint8_t* src;
    int* cg; int* cz; int* cf; int* cs;
    double* ppg; double* ppf;
    int* tmp;
    // what the PROPOSAL phase of a zone move reads: the shared-memory zone cache if there is one, else the global
    // arrays (zq_ld = row stride of the member lists); the adjacency CSR from shared memory if it was staged
    const int8_t* zq; const int* memq; const int* szq; int zq_ld;
    const int* adj_ptr; const int* adj_idx;
    int* cnt_tab;       // shared-memory count tables of resample_all_sources_blocked, or null
    double* par_tab;    // shared-memory parameter tables of resample_all_sources_staged, or null
    const double2* flt; // shared-memory table of flog()
};
__device__ __forceinline__ Chain chain_view(const ChainArrays& st, const StepParams& sp, int c) {
    Chain k;
    k.zone_id = st.zone_id + c * st.s_zone_id; k.members = st.zone_members + c * st.s_zone_members;
    k.zsize = st.zone_size + c * st.s_zone_size;
    k.w = st.weights + c * st.s_weights; k.wn = st.wnorm + c * st.s_wnorm;
    k.pg = st.p_global + c * st.s_p_global; k.pz = st.p_zones + c * st.s_p_zones; k.pf = st.p_fam + c * st.s_p_fam;
    k.src = st.source + c * st.s_source;
    k.cg = st.cnt_glob + c * st.s_cnt_glob; k.cz = st.cnt_zone + c * st.s_cnt_zone;
    k.cf = st.cnt_fam + c * st.s_cnt_fam; k.cs = st.cnt_src + c * st.s_cnt_src;
    k.ppg = st.prior_pg + c * st.s_prior_pg; k.ppf = st.prior_pf + c * st.s_prior_pf;
    k.tmp = sp.tmp_members + c * sp.s_tmp;
    k.cnt_tab = nullptr;
    k.par_tab = nullptr;
    k.zq = k.zone_id; k.memq = k.members; k.szq = k.zsize; k.zq_ld = 0;   // zq_ld / adjacency set by the kernel
    k.adj_ptr = nullptr; k.adj_idx = nullptr;
    k.flt = nullptr;
    return k;
}

// mixture value of one cell: sum_c w~[f][cfg][c] * p_c[f][x]   (model.py:174-175); x must not be NA
__device__ __forceinline__ double cell_mix(const DevData& d, const Chain& k, int f, int xs, int zid, int fam,
                                           const double* wrow /* wn[f] or a candidate row, [4][3] */) {
    const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
    const double* w = wrow + cfg * 3;
    const size_t o = (size_t)f * d.S + xs;
    double mix = w[0] * k.pg[o];
    if (zid >= 0) mix += w[1] * k.pz[(size_t)zid * d.F * d.S + o];
    if (fam >= 0) mix += w[2] * k.pf[(size_t)fam * d.F * d.S + o];
    return mix;
}

// K5a: log marginal-likelihood change of ONE site whose zone assignment changes zid_old -> zid_new
// (either may be -1).  Sum over features of log mix_new - log mix_old; NA cells contribute 0 in both the
// source-free likelihood (model.py:247) and the source-marginalised Hastings ratio.
// Marginal-likelihood change of a zone move: site_new joins zone z and / or site_removed leaves it (-1 = none).
// One pass over the features and ONE reduction for both sites.
__device__ double zone_move_delta(const DevData& d, const Chain& k, int site_new, int site_removed, int z,
                                  StepShared& sh, Team& tm) {
    const int fam_n = (d.inheritance && site_new >= 0) ? d.fam_id[site_new] : -1;
    const int fam_r = (d.inheritance && site_removed >= 0) ? d.fam_id[site_removed] : -1;
    double acc = 0.0;
    for (int f = threadIdx.x; f < d.F; f += blockDim.x) {
        const int xn = site_new >= 0 ? d.x[(size_t)site_new * d.ldx + f] : SB_NA;
        const int xr = site_removed >= 0 ? d.x[(size_t)site_removed * d.ldx + f] : SB_NA;
        const double* wrow = k.wn + (size_t)f * 12;
        if (xn != SB_NA) acc += flog(cell_mix(d, k, f, xn, z, fam_n, wrow) / cell_mix(d, k, f, xn, -1, fam_n, wrow), k.flt);
        if (xr != SB_NA) acc += flog(cell_mix(d, k, f, xr, -1, fam_r, wrow) / cell_mix(d, k, f, xr, z, fam_r, wrow), k.flt);
    }
    return cta_sum(acc, sh, tm);        // every CTA of the team evaluates the whole delta (F terms) itself
}

// ---- K7a: resample every source cell from its posterior, rebuild all count tables, return the new
// complete-data log-likelihood sum log(w~ * lh)[source]  (gibbs_sample_sources, zone_sampling.py:180-220;
// sample_categorical, preprocessing.py:321-348: first c with u < cumsum(post)[c], else 0).
__device__ double resample_all_sources(const DevData& d, const Chain& k, RngKey key, uint64_t step, double* red) {
    const int S = d.S, F = d.F, C = d.C;
    const size_t FS = (size_t)F * S;
    for (size_t i = threadIdx.x; i < FS; i += blockDim.x) k.cg[i] = 0;
    for (size_t i = threadIdx.x; i < (size_t)d.Z * FS; i += blockDim.x) k.cz[i] = 0;
    for (size_t i = threadIdx.x; i < (size_t)d.Fam * FS; i += blockDim.x) k.cf[i] = 0;
    for (size_t i = threadIdx.x; i < (size_t)F * 12; i += blockDim.x) k.cs[i] = 0;
    __syncthreads();
    const int upr = d.ldx >> 4;
    const int units = d.N * upr;
    LogProd acc;
    lp_init(acc);
    for (int u = threadIdx.x; u < units; u += blockDim.x) {
        const int n = u / upr;
        const int f0 = (u - n * upr) << 4;
        const uint4 xv = *reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0);
        const int zid = k.zone_id[n];
        const int fam = d.inheritance ? d.fam_id[n] : -1;
        const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
        const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
        uint32_t sw[4] = {0u, 0u, 0u, 0u};
        for (int j = 0; j < 16; j += 2) {
            const Draw dr = rng_draw(key, step, RP_SOURCE, 0, (uint32_t)(u * 8 + (j >> 1)));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int jj = j + h;
                const int xs = (xw[jj >> 2] >> ((jj & 3) * 8)) & 0xff;
                if (xs == SB_PAD) continue;
                const int f = f0 + jj;
                const double* w = k.wn + (size_t)f * 12 + cfg * 3;
                double t0 = w[0], t1 = w[1], t2 = C == 3 ? w[2] : 0.0;
                if (xs != SB_NA) {
                    const size_t o = (size_t)f * S + xs;
                    t0 *= k.pg[o];
                    t1 = zid >= 0 ? t1 * k.pz[(size_t)zid * FS + o] : 0.0;
                    t2 = fam >= 0 ? t2 * k.pf[(size_t)fam * FS + o] : 0.0;
                }
                const double tot = (t0 + t1) + t2;
                const double p0 = t0 / tot, p1 = t1 / tot, p2 = t2 / tot;
                const double uu = h == 0 ? dr.u0 : dr.u1;
                int s = 0;
                if (uu < p0) s = 0;
                else if (uu < p0 + p1) s = 1;
                else if (C == 3 && uu < (p0 + p1) + p2) s = 2;
                const double ts = s == 0 ? t0 : (s == 1 ? t1 : t2);
                lp_mul(acc, ts);
                sw[jj >> 2] |= (uint32_t)s << ((jj & 3) * 8);
                atomicAdd(k.cs + (size_t)f * 12 + cfg * 3 + s, 1);
                if (xs != SB_NA) {
                    if (s == 0) atomicAdd(k.cg + (size_t)f * S + xs, 1);
                    else if (s == 1) atomicAdd(k.cz + (size_t)zid * FS + (size_t)f * S + xs, 1);
                    else atomicAdd(k.cf + (size_t)fam * FS + (size_t)f * S + xs, 1);
                }
            }
        }
        *reinterpret_cast<uint4*>(k.src + (size_t)n * d.ldx + f0) = make_uint4(sw[0], sw[1], sw[2], sw[3]);
    }
    const double v = block_sum(lp_value(acc), red);
    return v;
}

// Same operator with shared-memory count tables: features are processed in blocks of 16 (one 16-byte vector of x per
// site); the counters of the block's features ([16][S] global, [16][12] source, [Z][16][S] zone, [Fam][16][S] family)
// live in shared memory, are filled with shared atomics and flushed with plain coalesced stores -- no global atomics,
// no zeroing pass.  One Philox block serves four cells (32-bit uniforms).
__device__ double resample_all_sources_blocked(const DevData& d, const Chain& k, RngKey key, uint64_t step, StepShared& sh,
                                               int* tab, Team& tm) {
    const int S = d.S, F = d.F, C = d.C, N = d.N;
    const size_t FS = (size_t)F * S;
    const int bS = 16 * S;
    int* tcg = tab;                     // [16][S]
    int* tcs = tcg + bS;                // [16][12]
    int* tcz = tcs + 192;               // [Z][16][S]
    int* tcf = tcz + d.Z * bS;          // [Fam][16][S]
    const int T = bS * (1 + d.Z + d.Fam) + 192;
    const int nb = d.ldx >> 4;
    int fb_lo, fb_hi;
    team_share(tm, nb, fb_lo, fb_hi);                   // each CTA of the team sweeps its own feature blocks
    for (int i = threadIdx.x; i < T; i += blockDim.x) tab[i] = 0;
    __syncthreads();
    LogProd acc;
    lp_init(acc);
    for (int fb = fb_lo; fb < fb_hi; ++fb) {
        const int f0 = fb << 4;
        for (int n = threadIdx.x; n < N; n += blockDim.x) {
            const uint4 xv = *reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0);
            const int zid = k.zone_id[n];
            const int fam = d.inheritance ? d.fam_id[n] : -1;
            const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
            const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
            uint32_t sw[4] = {0u, 0u, 0u, 0u};
            const double* pzr = k.pz + (size_t)(zid >= 0 ? zid : 0) * FS;
            const double* pfr = k.pf + (size_t)(fam >= 0 ? fam : 0) * FS;
            int* czr = tcz + (zid >= 0 ? zid : 0) * bS;
            int* cfr = tcf + (fam >= 0 ? fam : 0) * bS;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint32_t o[4];
                philox4x32((uint32_t)step, (uint32_t)(step >> 32), ((uint32_t)RP_SOURCE << 24) | (uint32_t)q,
                           (uint32_t)(n * nb + fb), key.k0, key.k1, o);
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                    const int xs = (xw[q] >> (h * 8)) & 0xff;
                    if (xs == SB_PAD) continue;
                    const int j = q * 4 + h;
                    const int f = f0 + j;
                    const double* w = k.wn + (size_t)f * 12 + cfg * 3;
                    double t0 = w[0], t1 = zid >= 0 ? w[1] : 0.0, t2 = (C == 3 && fam >= 0) ? w[2] : 0.0;
                    if (xs != SB_NA) {
                        const size_t oo = (size_t)f * S + xs;
                        t0 *= k.pg[oo];
                        if (zid >= 0) t1 *= pzr[oo];
                        if (C == 3 && fam >= 0) t2 *= pfr[oo];
                    }
                    const double tot = (t0 + t1) + t2;
                    const double ut = (((double)o[h] + 0.5) * (1.0 / 4294967296.0)) * tot;
                    // sample_categorical (preprocessing.py:338-341): first component whose cumulative mass exceeds u
                    int sc = 0;
                    if (ut >= t0) sc = (ut < t0 + t1) ? 1 : ((C == 3 && ut < tot) ? 2 : 0);
                    lp_mul(acc, sc == 0 ? t0 : (sc == 1 ? t1 : t2));
                    sw[q] |= (uint32_t)sc << (h * 8);
                    atomicAdd(tcs + j * 12 + cfg * 3 + sc, 1);
                    if (xs != SB_NA) {
                        int* t = sc == 0 ? tcg : (sc == 1 ? czr : cfr);
                        atomicAdd(t + j * S + xs, 1);
                    }
                }
            }
            *reinterpret_cast<uint4*>(k.src + (size_t)n * d.ldx + f0) = make_uint4(sw[0], sw[1], sw[2], sw[3]);
        }
        __syncthreads();
        // flush + clear: the block's counters are contiguous runs of the global tables
        const int nf = min(16, F - f0);
        for (int i = threadIdx.x; i < T; i += blockDim.x) {
            const int v = tab[i];
            tab[i] = 0;
            if (i < bS) {
                if (i < nf * S) k.cg[(size_t)f0 * S + i] = v;
            } else if (i < bS + 192) {
                const int r = i - bS;
                if (r < nf * 12) k.cs[(size_t)f0 * 12 + r] = v;
            } else {
                const int r = i - bS - 192;
                const int t = r / bS, e = r - t * bS;
                if (e < nf * S) {
                    if (t < d.Z) k.cz[(size_t)t * FS + (size_t)f0 * S + e] = v;
                    else k.cf[(size_t)(t - d.Z) * FS + (size_t)f0 * S + e] = v;
                }
            }
        }
        __syncthreads();
    }
    return team_sum(lp_value(acc), sh, tm);
}

// The same sweep with the PARAMETER tables of the 16-feature block staged in shared memory as well: w~ [16][12],
// p_global [16][S+1], p_zones [Z][16][S+1], p_families [Fam][16][S+1], the extra column holding 1 for NA cells
// (model.py:247); the counters get the same [table][16][S+1] shape, so a cell indexes both with one offset.  The cell body is then branch-free on-chip arithmetic -- three weight * probability products from
// shared memory, one categorical draw, two shared-memory counter increments -- with no dependent global load in the
// inner loop (the global gathers were the top stall of the sweep, profiles/r01_k8_op_sources_lines.txt).  The
// complete-data likelihood is NOT accumulated here: it is a function of the count tables (ll_from_counts) and nobody
// needs its absolute value inside a launch, so the caller just marks it stale.  Same counters, same Philox blocks and
// the same arithmetic per cell as resample_all_sources_blocked: the two produce identical source states.
__device__ void resample_all_sources_staged(const DevData& d, const Chain& k, RngKey key, uint64_t step, int* tab,
                                            double* ptab, Team& tm) {
    const int S = d.S, S1 = S + 1, F = d.F, C = d.C, N = d.N;
    const size_t FS = (size_t)F * S;
    const int bP = 16 * S1;             // entries of one (table, feature block): 16 features x (S states + NA)
    const int n_rows = 16 * (1 + d.Z + d.Fam);
    // counters  tab  = cnt[1 + Z + Fam][16][S1] (global, zones, families; the NA column is discarded) | src[16][12]
    // parameters ptab = p  [1 + Z + Fam][16][S1] (NA column = 1)                                         | w~ [16][12]
    int* tcs = tab + n_rows * S1;
    const double* W = ptab + n_rows * S1;
    const int T = n_rows * S1 + 192;
    const int nb = d.ldx >> 4;
    int fb_lo, fb_hi;
    team_share(tm, nb, fb_lo, fb_hi);                   // each CTA of the team sweeps its own feature blocks
    for (int i = threadIdx.x; i < T; i += blockDim.x) tab[i] = 0;
    for (int fb = fb_lo; fb < fb_hi; ++fb) {
        const int f0 = fb << 4;
        const int nf = min(16, F - f0);
        // stage the block's parameters, one table row (table, feature) per thread (the previous block's cells are
        // done: barrier at the end of the loop body)
        for (int i = threadIdx.x; i < 192; i += blockDim.x)
            ptab[n_rows * S1 + i] = i < nf * 12 ? k.wn[(size_t)f0 * 12 + i] : 0.0;
        for (int r = threadIdx.x; r < n_rows; r += blockDim.x) {
            const int t = r >> 4, j = r & 15;
            const double* src = (t == 0 ? k.pg : (t <= d.Z ? k.pz + (size_t)(t - 1) * FS : k.pf + (size_t)(t - 1 - d.Z) * FS))
                                + (size_t)(f0 + j) * S;
            double* dst = ptab + r * S1;
            for (int sx = 0; sx < S; ++sx) dst[sx] = j < nf ? src[sx] : 1.0;
            dst[S] = 1.0;
        }
        __syncthreads();
        for (int n = threadIdx.x; n < N; n += blockDim.x) {
            const uint4 xv = *reinterpret_cast<const uint4*>(d.x + (size_t)n * d.ldx + f0);
            const int zid = k.zone_id[n];
            const int fam = d.inheritance ? d.fam_id[n] : -1;
            const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
            const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
            uint32_t sw[4] = {0u, 0u, 0u, 0u};
            // table bases of this site (entries): component 0 / 1 / 2 -> global / its zone / its family
            const int bz = (1 + (zid >= 0 ? zid : 0)) * bP;
            const int bf = (1 + d.Z + (fam >= 0 ? fam : 0)) * bP;
            const double* wr = W + cfg * 3;
            int* tcc = tcs + cfg * 3;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (q * 4 >= nf) break;                                     // CTA-uniform
                uint32_t o[4];
                philox4x32((uint32_t)step, (uint32_t)(step >> 32), ((uint32_t)RP_SOURCE << 24) | (uint32_t)q,
                           (uint32_t)(n * nb + fb), key.k0, key.k1, o);
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                    const int j = q * 4 + h;
                    if (j >= nf) break;                                     // CTA-uniform (padding columns)
                    const int xs = (xw[q] >> (h * 8)) & 0xff;
                    const int e = j * S1 + min(xs, S);                      // NA -> the extra column
                    const double* w = wr + j * 12;
                    // a weight is 0 where its component is absent (wnorm_row), so no branch on zid / fam is needed
                    const double t0 = w[0] * ptab[e];
                    const double t1 = w[1] * ptab[bz + e];
                    const double t2 = C == 3 ? w[2] * ptab[bf + e] : 0.0;
                    const double t01 = t0 + t1, tot = t01 + t2;
                    // (o + 0.5) 2^-32 tot; the integer -> double conversion as one exact fp64 subtraction
                    const double uo = __hiloint2double(0x43300000, (int)o[h]) - 4503599627370495.5;
                    const double ut = (uo * (1.0 / 4294967296.0)) * tot;
                    // sample_categorical (preprocessing.py:338-341): first component whose cumulative mass exceeds u.
                    // u <= 1 - 2^-33, so ut < tot after rounding and the count of passed boundaries is the component
                    // (with C == 2, t2 = 0 and t01 = tot is never passed)
                    const int sc = (ut >= t0 ? 1 : 0) + (ut >= t01 ? 1 : 0);
                    sw[q] |= (uint32_t)sc << (h * 8);
                    atomicAdd(tcc + j * 12 + sc, 1);
                    atomicAdd(tab + (sc == 0 ? 0 : (sc == 1 ? bz : bf)) + e, 1);
                }
            }
            *reinterpret_cast<uint4*>(k.src + (size_t)n * d.ldx + f0) = make_uint4(sw[0], sw[1], sw[2], sw[3]);
        }
        __syncthreads();
        // flush + clear, one table row per thread: the rows of the block are contiguous runs of the global tables
        for (int r = threadIdx.x; r < n_rows; r += blockDim.x) {
            const int t = r >> 4, j = r & 15;
            int* dst = (t == 0 ? k.cg : (t <= d.Z ? k.cz + (size_t)(t - 1) * FS : k.cf + (size_t)(t - 1 - d.Z) * FS))
                       + (size_t)(f0 + j) * S;
            int* row = tab + r * S1;
            for (int sx = 0; sx < S; ++sx) {
                if (j < nf) dst[sx] = row[sx];
                row[sx] = 0;
            }
            row[S] = 0;
        }
        for (int i = threadIdx.x; i < 192; i += blockDim.x) {
            if (i < nf * 12) k.cs[(size_t)f0 * 12 + i] = tcs[i];
            tcs[i] = 0;
        }
        __syncthreads();
    }
    team_sync(tm);      // the feature blocks of the sweep do not coincide with the feature ranges of the other operators
}

// (the global-atomics variant is only used by one-CTA teams: sb_set_operators picks team = 1 when the tables do not fit)
// ll / ll_stale: the running log-likelihood and its "must be re-formed from the count tables" flag (ll_from_counts)
__device__ __forceinline__ void resample_sources(const DevData& d, const Chain& k, RngKey key, uint64_t step,
                                                 StepShared& sh, Team& tm, double& ll, bool& ll_stale) {
    if (k.par_tab) {
        resample_all_sources_staged(d, k, key, step, k.cnt_tab, k.par_tab, tm);
        ll_stale = true;
        return;
    }
    ll = k.cnt_tab ? resample_all_sources_blocked(d, k, key, step, sh, k.cnt_tab, tm)
                   : resample_all_sources(d, k, key, step, sh.red);
    ll_stale = false;
}

// prior terms that depend on the zones: size prior (model.py:932-972) and geo prior (:1096-1139, last zone only)
__device__ __forceinline__ double size_prior_of(const DevData& d, const int* zsize, int z_changed, int new_size) {
    if (d.size_type == 0) return 0.0;
    double p = 0.0;
    for (int z = 0; z < d.Z; ++z) {
        const double k = (double)(z == z_changed ? new_size : zsize[z]);
        p -= d.size_type == 1 ? log_binom_d((double)d.N, k) : log(k * k);
    }
    return p;
}

__device__ double geo_prior_of(const DevData& d, const int* members, int M, double* key, unsigned char* intree,
                               StepShared& sh) {
    double sum; int cnt;
    prim_mst(d.cost, d.loc, d.N, members, M, key, intree, sh.red, sh.redi, &sum, &cnt);
    return cnt > 0 ? (-(double)cnt * log(d.geo_scale) - sum / d.geo_scale) / (double)cnt : -log(d.geo_scale);
}

// q of adding `site` to a zone whose neighbour bitmap is `nbr`: (1-rho)/n_free + [site in nbr] rho/n_nbr
// (zone_sampling.py:839-845, 757-770, 910-917)
__device__ __forceinline__ double grow_q(bool in_nbr, int n_free, int n_nbr, double rho) {
    double q = (1.0 - rho) / (double)n_free;
    if (in_nbr) q += rho / (double)n_nbr;
    return q;
}

// K7b: grow / shrink / swap (zone_sampling.py:704-933 and the warm-up variants :1328-1574).
// Returns true if accepted.  ll / prior / geo are the chain's running values (updated on accept).
__device__ bool op_zone_move(const DevData& d, const StepParams& sp, const Chain& k, int kind, int chain,
                             RngKey key, uint64_t step, double& ll, double& prior, double& geo,
                             unsigned* nbr, unsigned* freeb, double* prim_key, unsigned char* prim_tree,
                             StepShared& sh, Team& tm, const double* ua, bool& synced, bool& wrote_nothing,
                             bool& need_sweep, int* err) {
    wrote_nothing = true;       // until the proposal is accepted: a rejected move only read the chain state
    // ua: this step's column of the look-ahead table (ua[q * SB_AHEAD] = uniform q of the step)
    PROF_T(tzp);
    // fm != null (sb_move_log_q): the zone and the sites are given, the proposal probabilities are stored and the
    // move is dropped -- the SAME code computes q / q_back for the sampler and for the parity surface
    ForcedMove* fm = sp.forced;
    const double rho = fm ? fm->rho : sp.p_grow[chain];
    const int z = fm ? fm->zone : ahead_int(ua[UA_ZONE * SB_AHEAD], d.Z);
    const int size = k.szq[z];
    int* mem = k.members + (size_t)z * d.N;                 // global list (written on accept)
    const int* memq = k.memq + (size_t)z * k.zq_ld;         // the list the proposal reads (shared-memory cache if any)
    if (!fm) {
        if (kind == SB_OP_GROW_ZONE && size >= sp.max_size[chain]) return false;      // :818-820
        if (kind == SB_OP_SHRINK_ZONE && size <= sp.min_size) return false;           // :894-896
    }

    int n_occ = 0;
    for (int i = 0; i < d.Z; ++i) n_occ += k.szq[i];
    const int n_free = d.N - n_occ;
    int site_new = -1, site_removed = -1;
    double q, q_back;
    PROF_T(tz0);
    PROF_ADD(12, tzp, tz0);

    if (kind != SB_OP_SHRINK_ZONE) {
        build_nbr_bits(k.adj_ptr, k.adj_idx, k.zq, memq, size, -1, nbr, sp.nwords);
        PROF_T(tz1);
        PROF_ADD(8, tz0, tz1);
        const bool connected = fm ? true : ua[UA_CONNECTED * SB_AHEAD] < rho;     // :731,823
        // warp 0 counts the neighbourhood and, for a connected step, picks the candidate in the same scan; the
        // other warps meanwhile build the bitmap of free sites if the step is not connected
        bits_select(nbr, sp.nwords, 0, connected ? &sh.pick : &sh.redi[1], &sh.redi[0], ua[UA_SITE_NEW * SB_AHEAD]);
        if (!connected) build_free_bits(d, k.zq, freeb, sp.nwords);
        __syncthreads();
        const int n_nbr = sh.redi[0];
        PROF_T(tz2);
        PROF_ADD(9, tz1, tz2);
        const int n_cand = connected ? n_nbr : n_free;
        if (n_cand == 0 && !fm) return false;                                     // :740-741, 832-833
        if (!connected) {
            bits_select(freeb, sp.nwords, ahead_int(ua[UA_SITE_NEW * SB_AHEAD], n_cand), &sh.pick);
            __syncthreads();
        }
        site_new = fm ? fm->site_new : sh.pick;
        q = grow_q(bit_test(nbr, site_new), n_free, n_nbr, rho);
        if (kind == SB_OP_GROW_ZONE) {
            q_back = 1.0 / (double)(size + 1);                                    // :848
        } else {
            site_removed = fm ? fm->site_removed : memq[ahead_int(ua[UA_SITE_REMOVED * SB_AHEAD], size)];  // :748-750
            q_back = grow_q(false, n_free, n_nbr, rho);    // old zone's neighbourhood never holds a member (:753,768)
        }
    } else {
        site_removed = fm ? fm->site_removed : memq[ahead_int(ua[UA_SITE_REMOVED * SB_AHEAD], size)];      // :899-901
        q = 1.0 / (double)size;                                                   // :904
        build_nbr_bits(k.adj_ptr, k.adj_idx, k.zq, memq, size, site_removed, nbr, sp.nwords);    // :906-908
        bits_select(nbr, sp.nwords, 0, &sh.redi[1], &sh.redi[0], 0.0);            // warp 0: the count only
        __syncthreads();
        const int n_nbr_back = sh.redi[0];
        q_back = grow_q(bit_test(nbr, site_removed), n_free + 1, n_nbr_back, rho);   // :911-917
        if (d.warmup) q_back = 1.0 / (double)(size + 1);                          // :1552 (quirk Q3)
    }
    __syncthreads();
    PROF_T(tz3);
    PROF_ADD(10, tz0, tz3);
    if (fm) {
        if (threadIdx.x == 0 && tm.rank == 0) { fm->out[0] = log(q); fm->out[1] = log(q_back); }
        return false;
    }

    // likelihood ratio of the moved site(s); sample_from_prior: the likelihood is identically 0 (mcmc_generative.py:115-116)
    const double dll = sp.from_prior ? 0.0 : zone_move_delta(d, k, site_new, site_removed, z, sh, tm);
    PROF_T(tz4);
    PROF_ADD(11, tz3, tz4);

    // prior ratio: only the size and geo terms depend on the zones
    const int new_size = size + (site_new >= 0 ? 1 : 0) - (site_removed >= 0 ? 1 : 0);
    double dprior = 0.0;
    if (d.size_type != 0) dprior += size_prior_of(d, k.szq, z, new_size) - size_prior_of(d, k.szq, -1, 0);
    double geo_new = geo;
    if (d.geo_type == 1 && z == d.Z - 1) {
        int m = 0;
        if (threadIdx.x == 0) {
            for (int i = 0; i < size; ++i)
                if (memq[i] != site_removed) k.tmp[m++] = memq[i];
            if (site_new >= 0) k.tmp[m++] = site_new;
        }
        __syncthreads();
        if (new_size > sp.prim_cap) {      // cannot happen through the operators (prim_cap > every max_size): a state
            if (threadIdx.x == 0 && tm.rank == 0) atomicOr(err, SB_ERR_GEO_CAP);      // uploaded later with a larger zone
            return false;
        }
        geo_new = geo_prior_of(d, k.tmp, new_size, prim_key, prim_tree, sh);
        dprior += geo_new - geo;
    }

    bool accept;
    if (d.warmup && d.sample_source) {
        accept = true;      // warm-up forces as_gibbs=True in gibbs_sample_sources => log_q = -inf => accept (:1293-1296)
    } else {
        const double mh = dll - flog(q / q_back, k.flt) + dprior;                 // mcmc_generative.py:345-349
        accept = flog(ua[UA_ACCEPT * SB_AHEAD], k.flt) < mh;                      // :318
    }
    PROF_T(tz5);
    PROF_ADD(13, tz4, tz5);
    if (!accept) return false;

    // Proposal and decision above are computed redundantly (and identically) by every CTA of the team from the old
    // state.  Barrier 1: nobody reads the old zone state any more; rank 0 writes; barrier 2: everybody sees the new one.
    wrote_nothing = false;
    team_sync(tm);
    // the same update on the global state (rank 0) and on this CTA's shared-memory cache (every CTA, so the copies of
    // a team stay identical without any traffic between them)
    const bool cached = k.zq != k.zone_id;
    if (threadIdx.x == 0 && (tm.rank == 0 || cached)) {
        for (int pass = (tm.rank == 0 ? 0 : 1); pass < (cached ? 2 : 1); ++pass) {
            int* m_ = pass == 0 ? mem : const_cast<int*>(memq);
            int8_t* zid_ = pass == 0 ? k.zone_id : const_cast<int8_t*>(k.zq);
            int* zs_ = pass == 0 ? k.zsize : const_cast<int*>(k.szq);
            if (site_removed >= 0) {
                int i = 0;
                while (m_[i] != site_removed) ++i;
                m_[i] = m_[size - 1];
                zid_[site_removed] = -1;
            }
            if (site_new >= 0) {
                m_[new_size - 1] = site_new;
                zid_[site_new] = (int8_t)z;
            }
            zs_[z] = new_size;
        }
    }
    team_sync(tm);
    synced = true;
    prior += dprior;
    geo = geo_new;
    if (d.sample_source) need_sweep = true;     // gibbs_sample_sources on accept (:859-862): run by the caller
    else ll += dll;
    return true;
}

// With SAMPLE_SOURCE the likelihood (model.py:177-184) is a function of the integer count tables (K4) and the
// parameter tables alone:
//   ll = sum_f [ sum_{cfg,c} cnt_src[f][cfg][c] log w~[f][cfg][c] + sum_s ( cnt_g[f][s] log p_global[f][s]
//              + sum_z cnt_z[z][f][s] log p_zones[z][f][s] + sum_k cnt_f[k][f][s] log p_families[k][f][s] ) ]
// (cells with count 0 contribute nothing; a positive count on a zero probability gives -inf, as the cell-by-cell
// form does).  The Gibbs operators therefore do not track their likelihood change; the running value is re-formed
// from the tables when it is needed -- once per launch -- at O(F * (12 + S * (1 + Z + Fam))) logarithms.
__device__ __forceinline__ double count_term(int c, double p) { return c > 0 ? (double)c * sb_log(p) : 0.0; }
__device__ double ll_from_counts(const DevData& d, const Chain& k, StepShared& sh, Team& tm) {
    int f_lo, f_hi;
    team_share(tm, d.F, f_lo, f_hi);
    const int S = d.S, FS = d.F * S;
    double acc = 0.0;
    for (int i = 12 * f_lo + threadIdx.x; i < 12 * f_hi; i += blockDim.x) acc += count_term(k.cs[i], k.wn[i]);
    for (int i = f_lo * S + threadIdx.x; i < f_hi * S; i += blockDim.x) {
        double t = count_term(k.cg[i], k.pg[i]);
        for (int z = 0; z < d.Z; ++z) t += count_term(k.cz[(size_t)z * FS + i], k.pz[(size_t)z * FS + i]);
        if (d.C == 3)
            for (int m = 0; m < d.Fam; ++m) t += count_term(k.cf[(size_t)m * FS + i], k.pf[(size_t)m * FS + i]);
        acc += t;
    }
    return team_sum(acc, sh, tm);
}

// ---- K7c: Gibbs operators on the count tables ----------------------------------------------------
// gibbs_sample_weights (zone_sampling.py:222-331).  Always accepted (incl. quirk Q2: with inheritance the
// per-feature MH decision is overwritten by `sample_new.weights = w_new`, :325-327).
__device__ void op_gibbs_weights(const DevData& d, const Chain& k, RngKey key, uint64_t step, bool fixed_inheritance,
                                 const Team& tm) {
    // fixed = choice(['inheritance', 'contact'])  (:249): fixed_inheritance = (uniform(step, RP_FIXED) < 0.5)
    // One thread per feature: the two Gammas of the feature's Beta / Dirichlet draw in lock-step, the new weight row
    // and its normalised table (wnorm_row: the same arithmetic as every other producer of w~).  The likelihood change
    // is not formed here: with SAMPLE_SOURCE the likelihood is a function of the count tables, see ll_from_counts.
    const int a = (d.C == 2) ? 1 : (fixed_inheritance ? 1 : 2);   // cfg of the "only this component" sites; also column x
    int f_lo, f_hi;
    team_share(tm, d.F, f_lo, f_hi);                    // each CTA of the team owns a contiguous range of features
    PROF_T(tw0);
    // The features go to the LAST warps of the CTA: in the table operators the first warps own two items each and the
    // last ones one, and with barrier-free Gibbs steps (see mcmc_step_kernel) it is the per-warp total that counts.
    const int nf = f_hi - f_lo, bd = blockDim.x;
    const int first = bd - min(bd, (nf + 31) & ~31);
    const int t0 = (int)threadIdx.x >= first ? (int)threadIdx.x - first : (int)threadIdx.x + bd - first;
    for (int f = f_lo + t0; f < f_hi; f += bd) {
        const int4* cs4 = reinterpret_cast<const int4*>(k.cs + (size_t)f * 12);
        const int4 ca = cs4[0], cb = cs4[1], cc = cs4[2];
        double* w = k.w + (size_t)f * 3;
        double v[3] = {w[0], w[1], d.C == 3 ? w[2] : 0.0};
        int c0, c1;
        if (d.C == 2) { c0 = ca.w; c1 = cb.x; }                            // counts over sites with a zone (:239-243)
        else if (a == 1) { c0 = cb.x + cc.z; c1 = ca.w + cc.y; }           // c_x, c_univ over has_zone sites (:256-259)
        else { c0 = cc.x + cc.w; c1 = cb.z + cc.y; }                       // ... over has_family sites     (:283-286)
        double g0, g1;
        rng_gamma2(key, step, RP_BETA_A, (uint32_t)f, 1.0 + (double)c0, true, RP_BETA_B, (uint32_t)f, 1.0 + (double)c1,
                   true, g0, g1);
        if (d.C == 2) {
            v[0] = g0; v[1] = g1;                           // Dir(1 + counts) = normalised Gammas
        } else {
            // a_x ~ Beta(1 + c_x, 1 + c_univ) = g0 / (g0 + g1)  (:262-265, 289-292);  w_x = a_x w_univ / (1 - a_x)
            // (:270, 297), and a_x / (1 - a_x) = g0 / g1
            const double nv = v[0] * (g0 / g1);
            if (a == 1) v[1] = nv; else v[2] = nv;
        }
        const double rs = 1.0 / ((v[0] + v[1]) + v[2]);     // normalize(w_new) :271, 298
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            v[c] *= rs;
            w[c] = v[c];
        }
        // the entries of w~ that can change (wnorm_row's arithmetic: a_c / ((a_0 + a_1) + a_2)); the others are the
        // constants 1 and 0 written at upload
        double* wn = k.wn + (size_t)f * 12;
        const double s1 = v[0] + v[1];
        const double q3 = v[0] / s1, q4 = v[1] / s1;
        wn[3] = q3; wn[4] = q4;
        if (d.C == 3) {
            const double s2 = v[0] + v[2], s3 = s1 + v[2];
            wn[6] = v[0] / s2; wn[8] = v[2] / s2;
            wn[9] = v[0] / s3; wn[10] = v[1] / s3; wn[11] = v[2] / s3;
        } else {
            wn[9] = q3; wn[10] = q4;                        // cfg 3 = cfg 1 without a family component
        }
    }
    PROF_T(tw1);
    PROF_ADD(7, tw0, tw1);
}

// Shared body of gibbs_sample_p_global / p_zones / p_families (zone_sampling.py:334-406):
// p[f][applicable] ~ Dirichlet(prior_counts[f] + cnt[f]) for the selected features.
//   p, cnt: tables of the row being resampled ([F][S]);  pc: prior pseudo-counts [F][S] or null (= 1)
//   use_subset: features chosen with probability sp.frac (:335,385), else all (:370)
//   alpha/lnorm/pcache: Dirichlet prior density tables for the prior delta (null if that prior is uniform)
__device__ void gibbs_table_row(const DevData& d, const StepParams& sp, const Chain& k, RngKey key, uint64_t step,
                                double* p, const int* cnt, const double* pc, bool use_subset,
                                const double* alpha, const double* lnorm, double* pcache,
                                double& prior_acc, const Team& tm) {
    const int S = d.S;
    int f_lo, f_hi;
    team_share(tm, d.F, f_lo, f_hi);                    // whole rows per CTA: the row totals stay CTA-local
    {
        // Warp-autonomous form (S <= SB_MAX_S = 32): a row (feature) occupies 2^s_shift consecutive lanes of one warp, one lane
        // per state.  Every lane draws its Gamma, the row total is a shuffle tree inside the lane group, and the lane
        // writes its own probability: no shared memory, no CTA barrier, and a (feature, state) item belongs to the
        // same thread in every step.  The subset draw of a feature (:335, 385) is recomputed by the lanes of its row
        // (same counter, same value; one instruction stream per warp either way).
        const int Sp = 1 << sp.s_shift;
        const int n_slots = (f_hi - f_lo) << sp.s_shift;
        double dpr = 0.0;
        PROF_T(t0);
        for (int base = 0; base < n_slots; base += blockDim.x) {           // CTA-uniform trip count
            const int slot = base + (int)threadIdx.x;
            const int f = f_lo + (slot >> sp.s_shift), s = slot & (Sp - 1);
            const bool row = slot < n_slots;
            const int i = f * S + s;
            const bool app = row && s < S && (d.app_mask[f] >> s & 1u);
            bool fsel = row;
            if (use_subset) fsel = fsel && rng_uniform(key, step, RP_SUBSET, f) < sp.frac;
            const bool sel = fsel && app;
            double g = 0.0;
            if (sel) g = rng_gamma(key, step, RP_GAMMA, i, (pc ? pc[i] : 1.0) + (double)cnt[i]);
            double tot = g;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)                                // levels above the row width are skipped (uniform)
                if (o < Sp) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (sel) p[i] = g / tot;
            if (alpha) {                                                    // prior change of the rows just written
                __syncwarp();
                if (fsel && s == 0) {
                    const double v = dirichlet_row(p + (size_t)f * S, alpha + (size_t)f * S, lnorm[f], d.app_mask[f], S);
                    dpr += v - pcache[f];
                    pcache[f] = v;
                }
                __syncwarp();
            }
        }
        PROF_T(t1);
        PROF_ADD(0, t0, t1);
        prior_acc += dpr;           // thread-private: summed over the team at the end of the launch
    }
}


// ---- K7d: alter_* operators (source-free mode), zone_sampling.py:408-612 --------------------------
// dirichlet_proposal on a renormalised pair (zone_sampling.py:538-569): v' ~ Dir(1 + kappa v); returns the new pair
// (scaled back by the pair's mass) and log_q - log_q_back, the difference of two 2-component Dirichlet log-densities
//   [lgamma(a0+a1) - lgamma(a0) - lgamma(a1) + (a0-1) ln v0 + (a1-1) ln v1]        a = 1 + kappa t  (forward)
// - [lgamma(b0+b1) - lgamma(b0) - lgamma(b1) + (b0-1) ln t0 + (b1-1) ln t1]        b = 1 + kappa v  (back).
// Evaluated ONCE per CTA, by warp 0, and broadcast through shared memory: lanes 0 / 1 draw the two Gammas in lock-step,
// then lane l takes one of the six lgamma terms and one of the four x log y terms.  (Every thread used to evaluate all of
// it -- ~3000 instructions per warp and step, which is what an alter_* step cost whatever the number of sites.)
// sh.dshare[0..1] are rewritten in the next alter_* step (after this step's reduction barrier), sh.dshare[2..3] behind that
// step's first CTA barrier, i.e. after every thread has read this step's values.
// log q(t -> v) - log q(v -> t) of the pair proposal; called by (at least) 8 consecutive lanes of a warp with
// l = lane & 7, result valid in all of them
__device__ __forceinline__ double pair_log_q_ratio(int l, double t0, double t1, double v0, double v1, double kappa) {
    const double a0 = 1.0 + kappa * t0, a1 = 1.0 + kappa * t1;
    const double b0 = 1.0 + kappa * v0, b1 = 1.0 + kappa * v1;
    const double X = l == 0 ? a0 : l == 1 ? a1 : l == 2 ? b0 : l == 3 ? b1 : l == 4 ? a0 + a1 : b0 + b1;
    const double sgn = (l == 0 || l == 1 || l == 5) ? -1.0 : 1.0;
    double term = l < 6 ? sgn * lgamma(X) : 0.0;
    const double coef = l == 0 ? a0 - 1.0 : l == 1 ? a1 - 1.0 : l == 2 ? -(b0 - 1.0) : -(b1 - 1.0);
    const double Y = l == 0 ? v0 : l == 1 ? v1 : l == 2 ? t0 : t1;
    if (l < 4) term += xlogy_d(coef, Y);
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) term += __shfl_xor_sync(0xffffffffu, term, o);
    return term;
}
// parity surface of dirichlet_proposal (zone_sampling.py:538-569) for a FORCED new pair: the log-density code of
// pair_proposal with v' given instead of drawn.  out[0] = log_q - log_q_back
__global__ void pair_log_q_kernel(double c0, double c1, double n0, double n1, double kappa, double* out) {
    const double mass = c0 + c1;
    const double term = pair_log_q_ratio(threadIdx.x & 7, c0 / mass, c1 / mass, n0 / mass, n1 / mass, kappa);
    if (threadIdx.x == 0) out[0] = term;
}
// Two phases: warp 0 draws the new pair and publishes it (CTA barrier); then warp 0 goes on with log q - log q_back (six
// lgamma, four x log y: the longest dependent chain of an alter_* step) and leaves it in sh.dshare[2] WITHOUT a barrier,
// while the other warps already walk the column (column_delta keeps warp 0 out of the site loop).  The caller reads
// sh.dshare[2] after the barrier of its reduction.
__device__ void pair_proposal(RngKey key, uint64_t step, double c0, double c1, double kappa, double& n0, double& n1,
                              StepShared& sh) {
    double t0 = 0.0, t1 = 0.0, v0 = 0.0, v1 = 0.0;
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        const double mass = c0 + c1;
        t0 = c0 / mass; t1 = c1 / mass;
        const double a0 = 1.0 + kappa * t0, a1 = 1.0 + kappa * t1;
        const double g = rng_gamma(key, step, RP_PROPOSAL, (uint32_t)(lane & 1), (lane & 1) ? a1 : a0);
        const double g0 = __shfl_sync(0xffffffffu, g, 0), g1 = __shfl_sync(0xffffffffu, g, 1);
        v0 = g0 / (g0 + g1); v1 = g1 / (g0 + g1);
        if (lane == 0) { sh.dshare[0] = v0 * mass; sh.dshare[1] = v1 * mass; }
    }
    __syncthreads();
    n0 = sh.dshare[0]; n1 = sh.dshare[1];
    if (threadIdx.x < 32) {
        const double term = pair_log_q_ratio(threadIdx.x & 7, t0, t1, v0, v1, kappa);
        if (threadIdx.x == 0) sh.dshare[2] = term;
    }
}

// ordered pair of distinct indices out of n (== random.sample(list, 2))
__device__ __forceinline__ void pick_pair(const double* ua, int n, int& a, int& b) {
    a = ahead_int(ua[UA_PAIR_A * SB_AHEAD], n);
    b = ahead_int(ua[UA_PAIR_B * SB_AHEAD], n - 1);
    if (b >= a) ++b;
}
// index of the j-th applicable state of a feature
__device__ __forceinline__ int nth_state(uint32_t mask, int j) {
    for (int i = 0; i < j; ++i) mask &= mask - 1;
    return __ffs(mask) - 1;
}

// Column delta of feature f: sum over the given sites of log mix_new - log mix_old, where "new" replaces either the
// weight row (wrow_new) or one probability row (which = 0 global / 1 zone z / 2 family fam; states s0,s1 -> v0,v1).
__device__ double column_delta(const DevData& d, const Chain& k, int f, const int* sites, int n_sites,
                               const double* wrow_new, int which, int tab, int s0, int s1, double v0, double v1,
                               StepShared& sh, Team& tm) {
    const double* wrow = k.wn + (size_t)f * 12;
    const size_t FS = (size_t)d.F * d.S;
    double acc = 0.0;
    int i_lo, i_hi;
    team_share(tm, n_sites, i_lo, i_hi);
    // warp 0 is still busy with the proposal's log-density ratio (pair_proposal): the other warps take the sites
    const int nthr = (int)blockDim.x - 32;
    for (int i = i_lo + (int)threadIdx.x - 32; threadIdx.x >= 32 && i < i_hi; i += nthr) {
        const int n = sites ? sites[i] : i;
        const int xs = sites ? d.x[(size_t)n * d.ldx + f] : d.xT[(size_t)f * d.ldn + n];
        if (xs == SB_NA) continue;
        if (which >= 0 && xs != s0 && xs != s1) continue;      // cell does not see the changed entries
        const int zid = k.zone_id[n];
        const int fam = d.inheritance ? d.fam_id[n] : -1;
        const double old = cell_mix(d, k, f, xs, zid, fam, wrow);
        double nw;
        if (which < 0) {
            nw = cell_mix(d, k, f, xs, zid, fam, wrow_new);
        } else {
            const int cfg = (zid >= 0 ? 1 : 0) + (fam >= 0 ? 2 : 0);
            const double* w = wrow + cfg * 3;
            const size_t o = (size_t)f * d.S + xs;
            const double pv = xs == s0 ? v0 : v1;
            double g = k.pg[o], zt = zid >= 0 ? k.pz[(size_t)zid * FS + o] : 0.0, ft = fam >= 0 ? k.pf[(size_t)fam * FS + o] : 0.0;
            if (which == 0) g = pv;
            else if (which == 1) { if (zid == tab) zt = pv; }
            else if (fam == tab) ft = pv;
            nw = w[0] * g;
            if (zid >= 0) nw += w[1] * zt;
            if (fam >= 0) nw += w[2] * ft;
        }
        acc += flog(nw / old, k.flt);
    }
    return team_sum(acc, sh, tm);       // barrier: every read of the old state (the callers read before calling) is done
}

// alter_weights / alter_p_global / alter_p_zones / alter_p_families (zone_sampling.py:408-612). Returns accepted.
__device__ bool op_alter(const DevData& d, const StepParams& sp, const Chain& k, int op, RngKey key, uint64_t step,
                         double& ll, double& prior, StepShared& sh, Team& tm, const double* ua, bool& synced, int* err) {
    // ua: this step's column of the look-ahead table (ua[q * SB_AHEAD] = uniform q of the step)
    const int f = ahead_int(ua[UA_FEATURE * SB_AHEAD], d.F);          // np.random.choice(range(n_features))
    double lqr, n0, n1, dll, dprior = 0.0;
    if (op == SB_OP_ALTER_WEIGHTS) {
        int a = 0, b = 1;
        if (d.C == 3) pick_pair(ua, 3, a, b);                         // _random.sample([0,1,2], 2)  :423
        double* w = k.w + (size_t)f * 3;
        pair_proposal(key, step, w[a], w[b], sp.prec_w, n0, n1, sh);
        __shared__ double wrow_new[12];
        if (threadIdx.x == 32) {                                      // (warp 0 is busy, see pair_proposal)
            double v[3] = {w[0], w[1], w[2]};
            v[a] = n0; v[b] = n1;
            wnorm_row(v, wrow_new, d.C);
        }
        // wrow_new is read by warps 1.. only (column_delta): a named barrier over them keeps warp 0 out of the wait
        if (threadIdx.x >= 32) asm volatile("bar.sync 1, %0;" ::"r"((int)blockDim.x - 32) : "memory");
        // sample_from_prior: no likelihood term; the team barrier of the reduction is still needed
        if (sp.from_prior) { team_sync(tm); dll = 0.0; }
        else dll = column_delta(d, k, f, nullptr, d.N, wrow_new, -1, 0, 0, 0, 0.0, 0.0, sh, tm);
        lqr = sh.dshare[2];
        const bool accept = flog(ua[UA_ACCEPT * SB_AHEAD], k.flt) < dll - lqr;
        synced = !accept;       // a rejected proposal wrote nothing after the reduction's barrier
        if (accept) {
            __syncthreads();
            if (threadIdx.x == 0 && tm.rank == 0) {
                w[a] = n0; w[b] = n1;
                for (int i = 0; i < 12; ++i) k.wn[(size_t)f * 12 + i] = wrow_new[i];
            }
            __syncthreads();
            ll += dll;
        }
        return accept;
    }
    const uint32_t mask = d.app_mask[f];
    const int napp = d.n_app[f];
    if (napp < 2) {                   // the reference's random.sample(list, 2) raises ValueError here: the step is
        if (threadIdx.x == 0 && tm.rank == 0) atomicOr(err, SB_ERR_NAPP);      // rejected and the chain flagged
        return false;
    }
    int ia, ib;
    pick_pair(ua, napp, ia, ib);                                      // _random.sample(list(f_states), 2)
    const int s0 = nth_state(mask, ia), s1 = nth_state(mask, ib);
    double* row;
    int which, tab = 0;
    const int* sites = nullptr;
    int n_sites = d.N;
    double kappa;
    const double* alpha = nullptr; const double* lnorm = nullptr; double* pcache = nullptr;
    if (op == SB_OP_ALTER_P_GLOBAL) {
        which = 0; row = k.pg + (size_t)f * d.S; kappa = sp.prec_g;
        if (d.pg_type == 1) { alpha = d.alpha_g + (size_t)f * d.S; lnorm = d.lnorm_g + f; pcache = k.ppg + f; }
    } else if (op == SB_OP_ALTER_P_ZONES) {
        which = 1; tab = ahead_int(ua[UA_ZONE * SB_AHEAD], d.Z); kappa = sp.prec_z;
        row = k.pz + ((size_t)tab * d.F + f) * d.S;
        sites = k.members + (size_t)tab * d.N; n_sites = k.zsize[tab];
    } else {
        which = 2; tab = ahead_int(ua[UA_FAMILY * SB_AHEAD], d.Fam); kappa = sp.prec_f;
        row = k.pf + ((size_t)tab * d.F + f) * d.S;
        sites = d.fam_sites + d.fam_ptr[tab]; n_sites = d.fam_ptr[tab + 1] - d.fam_ptr[tab];
        if (d.pf_type == 1) {
            alpha = d.alpha_f + ((size_t)tab * d.F + f) * d.S; lnorm = d.lnorm_f + (size_t)tab * d.F + f;
            pcache = k.ppf + (size_t)tab * d.F + f;
        }
    }
    pair_proposal(key, step, row[s0], row[s1], kappa, n0, n1, sh);
    double new_prior_term = 0.0;
    if (alpha && (threadIdx.x >> 5) == 1) {
        // Dirichlet prior of the changed row with the two entries replaced (read before the reduction's barrier):
        // warp 1, one state per lane (warp 0 is busy, see pair_proposal); published by the reduction's barrier
        const int lane = threadIdx.x & 31;
        double t = 0.0;
        for (int s = lane; s < d.S; s += 32)
            if (mask >> s & 1u) t += xlogy_d(alpha[s] - 1.0, s == s0 ? n0 : (s == s1 ? n1 : row[s]));
        t = warp_sum(t);
        if (lane == 0) sh.dshare[3] = *lnorm + t;
    }
    if (sp.from_prior) { team_sync(tm); dll = 0.0; }
    else dll = column_delta(d, k, f, sites, n_sites, nullptr, which, tab, s0, s1, n0, n1, sh, tm);
    lqr = sh.dshare[2];
    if (alpha) {
        new_prior_term = sh.dshare[3];
        dprior = new_prior_term - *pcache;
    }
    const bool accept = flog(ua[UA_ACCEPT * SB_AHEAD], k.flt) < dll - lqr + dprior;
    synced = !accept;
    if (accept) {
        __syncthreads();
        if (threadIdx.x == 0 && tm.rank == 0) {
            row[s0] = n0; row[s1] = n1;
            if (alpha) *pcache = new_prior_term;
        }
        __syncthreads();
        ll += dll;
        prior += dprior;
    }
    return accept;
}

// ---- snapshot of one chain's state into the ring: record = zone_id i8[s_zone_id] | weights
// f64[F][3] | p_global | p_zones | p_fam, every part 16 B aligned.  All CTAs of the team copy, 16 bytes per thread and
// iteration; the caller has published the team's writes (team barrier) before and fences with one after.
__host__ __device__ __forceinline__ size_t ring_align(size_t v) { return (v + 15) & ~(size_t)15; }
__device__ void snapshot_state(const DevData& d, const ChainArrays& st, const StepParams& sp, int chain, int slot,
                               int n_chains, const Team& tm) {
    unsigned char* rec = sp.ring + ((size_t)slot * n_chains + chain) * sp.ring_rec;
    const void* src[5] = {st.zone_id + chain * st.s_zone_id, st.weights + chain * st.s_weights,
                          st.p_global + chain * st.s_p_global, st.p_zones + chain * st.s_p_zones,
                          st.p_fam + chain * st.s_p_fam};
    const size_t len[5] = {st.s_zone_id, st.s_weights * 8, st.s_p_global * 8, st.s_p_zones * 8, st.s_p_fam * 8};
    size_t off = 0;
    const int tid = tm.rank * blockDim.x + threadIdx.x, nthr = tm.size * blockDim.x;
    for (int a = 0; a < 5; ++a) {
        // every per-chain array starts 16 B aligned only if its stride is a multiple of 2 doubles: copy 8 B units
        const size_t n8 = len[a] >> 3;
        const unsigned long long* s8 = reinterpret_cast<const unsigned long long*>(src[a]);
        unsigned long long* d8 = reinterpret_cast<unsigned long long*>(rec + off);
        for (size_t i = tid; i < n8; i += nthr) d8[i] = __ldcg(s8 + i);
        off += ring_align(len[a]);
    }
}

// ---- K8: the persistent kernel ---------------------------------------------------------------------
__global__ void __launch_bounds__(STEP_THREADS, STEP_MIN_CTAS)
mcmc_step_kernel(DevData d, ChainArrays st, StepParams sp, long long n_steps) {
    extern __shared__ unsigned char step_smem[];
    __shared__ StepShared sh;
    const int chain = blockIdx.x / sp.team;             // one cluster of sp.team CTAs per chain (cluster dims (team,1,1))
    Team tm = {(int)(blockIdx.x % sp.team), sp.team, 0u, 31 - __clz(sp.team), 0u};
    unsigned* nbr = reinterpret_cast<unsigned*>(step_smem);
    unsigned* freeb = nbr + sp.nwords;
    double* prim_key = reinterpret_cast<double*>(freeb + sp.nwords);       // 2 * nwords words from the start: 8 B aligned
    unsigned char* prim_tree = reinterpret_cast<unsigned char*>(prim_key + sp.prim_cap);

    Chain kv = chain_view(st, sp, chain);
    if (sp.work_counts) kv.cnt_tab = reinterpret_cast<int*>(step_smem + sp.work_off);
    if (sp.work_counts == 2) kv.par_tab = reinterpret_cast<double*>(step_smem + sp.par_off);
    __shared__ double2 flog_tab[SB_FLOG_N];
    flog_init(flog_tab);
    kv.zq_ld = d.N;
    kv.adj_ptr = d.adj_ptr; kv.adj_idx = d.adj_idx;
    if (sp.zc_off >= 0) {
        // zone cache: sizes, member lists (first zc_cap entries of every row) and zone ids of this chain
        int* zs = reinterpret_cast<int*>(step_smem + sp.zc_off);
        int* zm = zs + d.Z;
        int8_t* zi = reinterpret_cast<int8_t*>(zm + (size_t)d.Z * sp.zc_cap);
        for (int i = threadIdx.x; i < d.Z; i += blockDim.x) zs[i] = kv.zsize[i];
        for (int i = threadIdx.x; i < d.Z * sp.zc_cap; i += blockDim.x) {
            const int z = i / sp.zc_cap, j = i - z * sp.zc_cap;
            zm[i] = j < d.N ? kv.members[(size_t)z * d.N + j] : 0;
        }
        for (int i = threadIdx.x; i < d.N; i += blockDim.x) zi[i] = kv.zone_id[i];
        kv.szq = zs; kv.memq = zm; kv.zq = zi; kv.zq_ld = sp.zc_cap;
    }
    if (sp.adj_off >= 0) {
        int* ap = reinterpret_cast<int*>(step_smem + sp.adj_off);
        int* ai = ap + d.N + 1;
        for (int i = threadIdx.x; i <= d.N; i += blockDim.x) ap[i] = d.adj_ptr[i];
        const int nnz = d.adj_ptr[d.N];
        for (int i = threadIdx.x; i < nnz; i += blockDim.x) ai[i] = d.adj_idx[i];
        kv.adj_ptr = ap; kv.adj_idx = ai;
    }
    __syncthreads();
    kv.flt = flog_tab;
    const Chain& k = kv;
    const unsigned gchain = (unsigned)(sp.chain_offset + chain);
    const RngKey key = {sp.seed_lo ^ (gchain * 0x9E3779B9u), sp.seed_hi ^ gchain};
    double ll = sp.from_prior ? 0.0 : st.ll[chain], prior = st.prior[chain], geo = st.geo[chain];
    const uint64_t step0 = st.step[chain];
    if (sp.forced) {
        // parity surface (sb_move_log_q): one forced proposal through op_zone_move, behind the same set-up (zone cache,
        // staged adjacency) as a sampling launch; nothing of the chain state is written
        if (chain == sp.forced_chain) {
            for (int i = threadIdx.x; i < SB_N_AHEAD; i += blockDim.x) sh.u_ahead[i][0] = 0.5;
            __syncthreads();
            bool f0 = false, f1 = false, f2 = false;
            op_zone_move(d, sp, k, sp.forced->kind, chain, key, step0, ll, prior, geo, nbr, freeb, prim_key, prim_tree, sh,
                         tm, &sh.u_ahead[0][0], f0, f1, f2, st.err + chain);
        }
        if (tm.size > 1) cg::this_cluster().sync();
        return;
    }
    int ev = 0;                                         // cursor into sp.events (team-uniform)

    bool dirty = false;         // a Gibbs operator wrote chain tables since the last full team barrier
    // The Gibbs operators are always accepted, and nobody needs the absolute ll / prior inside a launch (the
    // Metropolis steps work on deltas).  So a Gibbs step has no reduction and no team traffic: its prior change stays
    // in a thread-private accumulator that is summed over the team at the end of the launch, and its likelihood
    // change is not formed at all -- the running ll is re-formed from the count tables (ll_from_counts) if a Gibbs
    // operator ran after the last from-scratch value.
    double prior_acc = 0.0;
    bool ll_stale = false;
    // the steps run in segments that end at the next host-visible event (or at n_steps): the hot loop itself knows
    // nothing about events
    long long it = 0;
    for (;;) {
    const long long seg_end = ev < sp.n_events ? (long long)sp.events[ev].x : n_steps;
    for (; it < seg_end; ++it) {
        const uint64_t step = step0 + (uint64_t)it;
        PROF_T(ts0);
        // operator choice: np.random.choice(fn_operators, p=p_operators)  (mcmc_generative.py:294).  The draws that
        // depend on the step number only are made SB_AHEAD steps at a time: warp w, lane l -> draw w of step + l.
        const int ia = (int)(it & (SB_AHEAD - 1));
        if (ia == 0) {
            __syncthreads();            // Gibbs steps end without a barrier: wait until nobody reads the old table
            // SB_N_AHEAD + 1 tasks per slab of 32 steps (the operator id and the SB_N_AHEAD uniforms), one warp each,
            // dealt round-robin to the warps; lane l -> step + 32 * slab + l
            const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
            for (int item = w; item < (SB_N_AHEAD + 1) * (SB_AHEAD / 32); item += STEP_WARPS) {
                const int slab = item / (SB_N_AHEAD + 1), task = item - slab * (SB_N_AHEAD + 1);
                const int a = slab * 32 + l;
                if (task == 0) {
                    const double uo = rng_uniform(key, step + a, RP_OP, 0);
                    int oi = 0;
                    while (oi < sp.n_ops - 1 && uo >= sp.op_cdf[oi]) ++oi;
                    sh.op_ahead[a] = sp.op_ids[oi];
                } else {
#pragma unroll
                    for (int q = 0; q < SB_N_AHEAD; ++q)
                        if (q == task - 1) sh.u_ahead[q][a] = rng_uniform(key, step + a, ahead_purpose(q), ahead_index(q));
                }
            }
            __syncthreads();
        }
        const int op = sh.op_ahead[ia];
        bool accepted = true;
        bool synced = false;    // the operator ended in a team barrier that followed all of its global writes
        bool read_only = false; // the operator wrote nothing (a rejected zone move)
        bool need_sweep = false;
        PROF_T(ts1);
        // A Gibbs operator leaves its writes (own feature range only) unpublished; every other operator reads the
        // chain's tables across feature ranges, so it starts behind a full team barrier if anything is pending.
        const bool gibbs_table_op = op == SB_OP_GIBBS_WEIGHTS || op == SB_OP_GIBBS_P_GLOBAL ||
                                    op == SB_OP_GIBBS_P_ZONES || op == SB_OP_GIBBS_P_FAMILIES;
        if (gibbs_table_op) {
            dirty = true;
            ll_stale = true;
        } else if (dirty) {
            team_sync(tm);
            dirty = false;
        }
        switch (op) {
            case SB_OP_SHRINK_ZONE: case SB_OP_GROW_ZONE: case SB_OP_SWAP_ZONE:
                accepted = op_zone_move(d, sp, k, op, chain, key, step, ll, prior, geo, nbr, freeb, prim_key,
                                        prim_tree, sh, tm, &sh.u_ahead[0][ia], synced, read_only, need_sweep,
                                        st.err + chain);
                break;
            case SB_OP_GIBBS_SOURCES:
                need_sweep = true;
                break;
            case SB_OP_GIBBS_WEIGHTS:
                op_gibbs_weights(d, k, key, step, sh.u_ahead[UA_FIXED][ia] < 0.5, tm);
                synced = true;
                break;
            case SB_OP_GIBBS_P_GLOBAL: case SB_OP_GIBBS_P_ZONES: case SB_OP_GIBBS_P_FAMILIES: {
                // one call site for the three table operators (one copy of the Gamma sampler in the instruction cache)
                double* p = k.pg; const int* cnt = k.cg; const double* pc = d.gibbs_g;
                const double* alpha = d.pg_type == 1 ? d.alpha_g : nullptr; const double* lnorm = d.lnorm_g;
                double* pcache = k.ppg;
                if (op == SB_OP_GIBBS_P_ZONES) {
                    const size_t o = (size_t)ahead_int(sh.u_ahead[UA_ZONE][ia], d.Z) * d.F * d.S;
                    p = k.pz + o; cnt = k.cz + o; pc = nullptr; alpha = nullptr; lnorm = nullptr; pcache = nullptr;
                } else if (op == SB_OP_GIBBS_P_FAMILIES) {
                    const int fam = ahead_int(sh.u_ahead[UA_FAMILY][ia], d.Fam);
                    const size_t o = (size_t)fam * d.F * d.S;
                    p = k.pf + o; cnt = k.cf + o; pc = d.gibbs_f + o;
                    alpha = d.pf_type == 1 ? d.alpha_f + o : nullptr; lnorm = d.lnorm_f + (size_t)fam * d.F;
                    pcache = k.ppf + (size_t)fam * d.F;
                }
                gibbs_table_row(d, sp, k, key, step, p, cnt, pc, op != SB_OP_GIBBS_P_ZONES, alpha, lnorm, pcache,
                                prior_acc, tm);
                synced = true;
                break;
            }
            default:
                accepted = op_alter(d, sp, k, op, key, step, ll, prior, sh, tm, &sh.u_ahead[0][ia], synced, st.err + chain);
                break;
        }
        // the source sweep has ONE call site (gibbs_sample_sources and every accepted zone move): it is the largest
        // block of code of the kernel
        if (need_sweep) {
            resample_sources(d, k, key, step, sh, tm, ll, ll_stale);
            if (sp.from_prior) ll = 0.0;
            synced = true;
        }
        PROF_T(ts2);
        if (threadIdx.x == 0 && tm.rank == 0)      // fire-and-forget reduction: nobody waits for the L2 round trip
            atomicAdd((accepted ? st.acc : st.rej) + (size_t)chain * SB_N_OPS + op, 1ull);
        // Step boundary.  The Gibbs table operators are warp-autonomous -- a thread only ever reads and writes the
        // items it owns, in every step -- so the warps run on without a barrier until the next operator of another
        // kind (which starts behind the `dirty` team barrier above) or the next refill of the look-ahead table.
        // After any other operator: the next step's writes must not overtake this step's reads in a slower thread
        // or CTA of the team.
        // A step that wrote nothing needs the team barrier only for execution order (no release / acquire, so L1
        // stays valid).
        if (!gibbs_table_op) {
            if (synced) __syncthreads();
            else if (read_only) team_sync_exec(tm);
            else team_sync(tm);
        }
        PROF_T(ts3);
        PROF_ADD(4, ts0, ts1); PROF_ADD(5, ts1, ts2); PROF_ADD(6, ts2, ts3);
    }
    if (ev >= sp.n_events) break;
    // ---- host-visible event after `it` completed steps: (ll, prior) of the chain, optionally a snapshot of its
    // state into the ring -- the launch goes on (generate_samples logs without ending the persistent kernel)
    {
        __syncthreads();                // the warps may arrive from barrier-free Gibbs steps
        if (dirty) { team_sync(tm); dirty = false; }
        if (ll_stale && !sp.from_prior) { ll = ll_from_counts(d, k, sh, tm); ll_stale = false; }
        prior += team_sum(prior_acc, sh, tm);
        prior_acc = 0.0;
        const int n_chains = (int)(gridDim.x / sp.team);
        if (threadIdx.x == 0 && tm.rank == 0) sp.ev_scalars[(size_t)ev * n_chains + chain] = make_double2(ll, prior);
        const int slot = sp.events[ev].y;
        if (slot >= 0 && (sp.ring_chain < 0 || sp.ring_chain == chain))
            snapshot_state(d, st, sp, chain, slot, n_chains, tm);
        ++ev;
        team_sync(tm);                  // the copy read the whole state: done before the next step writes
    }
    }
    __syncthreads();            // the warps may arrive from barrier-free Gibbs steps
    if (ll_stale && !sp.from_prior) ll = ll_from_counts(d, k, sh, tm);       // reads this CTA's own feature range only
    prior += team_sum(prior_acc, sh, tm);
    if (threadIdx.x == 0 && tm.rank == 0) {
        st.ll[chain] = ll; st.prior[chain] = prior; st.geo[chain] = geo;
        st.step[chain] = step0 + (uint64_t)n_steps;
        // NaN (or +inf) in the running values: a Gamma draw that exhausted its attempts returns NaN, a NaN weight row
        // poisons the likelihood -- surfaced by sb_read_chain_scalars / sb_chain_errors (SB_E_NUMERIC)
        if (!(ll == ll) || !(prior == prior) || ll == INFINITY || prior == INFINITY) atomicOr(st.err + chain, SB_ERR_NONFINITE);
    }
    if (tm.size > 1) cg::this_cluster().sync();         // keep every CTA's shared memory alive until the last peer read
}

