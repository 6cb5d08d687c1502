/*
 * sbayes_b200.h -- C ABI of the B200-native sBayes MCMC inner loop (libsbayes_b200.so).
 *
 * The reference (derpetermann/contact_zones) is pure Python and has no FFI of its own; the boundary it
 * offers for this path is its Python object API.  Each entry point below names the reference interface
 * whose arithmetic it replaces (paths relative to the reference root).  The Python mirror of that object
 * API (contact_zones_b200/{model,sampling,...}.py) binds these symbols through ctypes -- see
 * INTEGRATION.md for the stub a reference maintainer would add.
 *
 * Conventions
 *  - every function returns 0 on success, a negative SB_E_* code on failure, never throws;
 *    sb_last_error(ctx) returns a message owned by the context (valid until the next call on it);
 *  - all pointer arguments are HOST pointers borrowed for the duration of the call; device memory is
 *    owned by the context and released by sb_destroy;
 *  - one context per (process, GPU); calls on one context are serialised by the caller and are
 *    stream-ordered on the context's CUDA stream (sb_set_stream); contexts are independent;
 *  - arrays are C-contiguous, layouts as in the reference: zones (Z,N) u8 0/1, weights (F,C) f64,
 *    p_global (F,S), p_zones (Z,F,S), p_families (Fam,F,S) f64, x / source (N,F) u8.
 *    x[n,f] = state index, 255 = NA (all-zero one-hot, model.py:112); source[n,f] in {0,1,2} =
 *    [universal, contact, inheritance] (model.py:243-245).
 */
#ifndef SBAYES_B200_H
#define SBAYES_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sb_ctx sb_ctx;

enum {
    SB_OK = 0,
    SB_E_ARG = -1,        /* invalid argument / shape                      */
    SB_E_CUDA = -2,       /* CUDA runtime error (message in sb_last_error) */
    SB_E_STATE = -3,      /* call order (data / state / operators not set) */
    SB_E_OVERLAP = -4,    /* sb_set_state: overlapping zones (quirk Q7)    */
    SB_E_NOGPU = -5,      /* no CUDA device                                */
    SB_E_NUMERIC = -6     /* a chain carries a sticky SB_ERR_* flag (sb_chain_errors) */
};

/* sticky per-chain error flags raised by the step kernel.  The reference's counterparts: the NaN / Inf warning and
 * `assert 0 < q` of dirichlet_proposal (zone_sampling.py:559-567), the ValueError of random.sample(states, 2) on a
 * feature with one applicable state (:471, 512, 588), the ValueError of geo_prior_distance (model.py:1134-1135). */
enum {
    SB_ERR_NONFINITE = 1, /* NaN or +inf in the running log-likelihood / log-prior (incl. an exhausted Gamma sampler) */
    SB_ERR_GEO_CAP = 2,   /* last zone outgrew the geo prior's member scratch (state uploaded after sb_set_operators) */
    SB_ERR_NAPP = 4       /* alter_p_* picked a feature with fewer than 2 applicable states (step rejected)           */
};

/* sb_create flags */
enum {
    SB_F_INHERITANCE = 1u << 0,    /* model.py:45  config['INHERITANCE']   -> C = 3 */
    SB_F_SAMPLE_SOURCE = 1u << 1,  /* model.py:46  config['SAMPLE_SOURCE'] -> per-cell source state */
    SB_F_WARMUP = 1u << 2          /* ZoneMCMCWarmup semantics (zone_sampling.py:1272-1577) */
};

/* prior types (model.py:567-1056) */
enum { SB_SIZE_NONE = 0, SB_SIZE_UNIFORM = 1, SB_SIZE_QUADRATIC = 2 };      /* ZoneSizePrior :893-976 */
enum { SB_GEO_UNIFORM = 0, SB_GEO_COST_BASED = 1 };                          /* GeoPrior      :979-1056 */
enum { SB_DIR_UNIFORM = 0, SB_DIR_COUNTS = 1 };                              /* PGlobalPrior :567, PFamiliesPrior :633 */

/* operator ids (zone_sampling.py; table built by mcmc_setup.py:70-95) */
enum {
    SB_OP_SHRINK_ZONE = 0, SB_OP_GROW_ZONE = 1, SB_OP_SWAP_ZONE = 2,
    SB_OP_GIBBS_SOURCES = 3, SB_OP_GIBBS_WEIGHTS = 4, SB_OP_GIBBS_P_GLOBAL = 5,
    SB_OP_GIBBS_P_ZONES = 6, SB_OP_GIBBS_P_FAMILIES = 7,
    SB_OP_ALTER_WEIGHTS = 8, SB_OP_ALTER_P_GLOBAL = 9, SB_OP_ALTER_P_ZONES = 10, SB_OP_ALTER_P_FAMILIES = 11,
    SB_N_OPS = 12
};

typedef struct {
    int32_t n_sites;       /* N   */
    int32_t n_features;    /* F   */
    int32_t n_states;      /* S   (max states per feature, <= 32) */
    int32_t n_zones;       /* Z   */
    int32_t n_families;    /* Fam (0 without inheritance) */
    int32_t n_chains;      /* chains resident in this context (this rank's shard) */
    int32_t chain_offset;  /* global id of local chain 0: RNG streams are keyed by the GLOBAL chain id, so
                              trajectories do not depend on how chains are sharded over GPUs */
    uint32_t flags;        /* SB_F_* */
} sb_dims;

typedef struct {
    int32_t size_type;       /* SB_SIZE_*  */
    int32_t geo_type;        /* SB_GEO_*   */
    double geo_scale;        /* GeoPrior.scale, model.py:1010 */
    int32_t p_global_type;   /* SB_DIR_*   */
    int32_t p_families_type; /* SB_DIR_*   */
} sb_prior_cfg;

typedef struct {
    int32_t n_ops;                   /* <= SB_N_OPS */
    int32_t op_ids[SB_N_OPS];        /* SB_OP_*                                              */
    double op_prob[SB_N_OPS];        /* selection probabilities (mcmc_generative.py:294)     */
    double precision_weights;        /* PROPOSAL_PRECISION (zone_sampling.py:165-169)        */
    double precision_p_global;
    double precision_p_zones;
    double precision_p_families;
    int32_t min_size;                /* MIN_M  (model.py:43)                                 */
    double gibbs_feature_fraction;   /* 0.4: zone_sampling.py:334,381                        */
    uint64_t seed;
} sb_operator_cfg;

/* ---- lifetime -------------------------------------------------------------------------------- */
int sb_version(void);
int sb_device_count(void);
int sb_create(sb_ctx** out, int device, const sb_dims* dims);
int sb_destroy(sb_ctx* ctx);
const char* sb_last_error(const sb_ctx* ctx);
int sb_set_stream(sb_ctx* ctx, void* cuda_stream);   /* NULL = the context's own stream */
int sb_sync(sb_ctx* ctx);

/* ---- immutable data: replaces Likelihood.__init__ (model.py:105-131), ZoneMCMC.__init__
 *      (zone_sampling.py:127-151) and the Prior constructors (model.py:470-482) ------------------- */
int sb_upload_data(sb_ctx* ctx,
                   const uint8_t* x,            /* (N,F)  state index, 255 = NA                       */
                   const uint8_t* applicable,   /* (F,S)  data.states                                  */
                   const int16_t* family_id,    /* (N)    -1 = none; NULL without inheritance          */
                   const int32_t* adj_indptr,   /* (N+1)  network['adj_mat'] CSR (util.py:139-155)     */
                   const int32_t* adj_indices,  /* (nnz)                                               */
                   const double* cost);         /* (N,N)  geo_prior['cost_matrix'] or NULL             */
/* Site coordinates (N,2) for the IMPLICIT cost matrix of the cost_based geo prior: cost(i, j) = Euclidean distance of
 * the sites -- exactly the `dist_mat` the reference uses as cost matrix when no cost file is given
 * (simulation.py:155-156, load_data.py:116-132 -> preprocessing.py:101-174), without its O(N^2) memory (3.2 GB at
 * N = 20000).  Used when sb_upload_data got no dense cost matrix. */
int sb_upload_locations(sb_ctx* ctx, const double* xy);
int sb_set_priors(sb_ctx* ctx, const sb_prior_cfg* cfg,
                  const double* gibbs_counts_global,  /* (F,S)     prior_p_global.counts   (model.py:577,585)  or NULL = 1 */
                  const double* alpha_global,         /* (F,S)     Dirichlet density alpha (util.py:589-623)   or NULL     */
                  const double* gibbs_counts_fam,     /* (Fam,F,S) prior_p_families.counts (model.py:647,664)  or NULL = 1 */
                  const double* alpha_fam);           /* (Fam,F,S)                                              or NULL     */

/* ---- per-chain state: the device image of a `Sample` (zone_sampling.py:49-115) ------------------ */
int sb_set_state(sb_ctx* ctx, int chain, const uint8_t* zones, const double* weights, const double* p_global,
                 const double* p_zones, const double* p_families, const uint8_t* source /* or NULL */);
/* n_chains consecutive chains in one call: the same arrays stacked along a leading chain axis (C-contiguous).  One
 * host->device copy per device array whatever n_chains is -- the bulk form of `for c in chains: generate_initial_sample`
 * (mcmc_generative.py:166-171). */
int sb_set_states(sb_ctx* ctx, int chain_begin, int n_chains, const uint8_t* zones, const double* weights,
                  const double* p_global, const double* p_zones, const double* p_families, const uint8_t* source);
int sb_get_state(sb_ctx* ctx, int chain, uint8_t* zones, double* weights, double* p_global,
                 double* p_zones, double* p_families, uint8_t* source /* any may be NULL */);
int sb_copy_state(sb_ctx* ctx, int dst_chain, int src_chain);

/* ---- stateless full evaluation of resident states (kernels K1, K2, K3) --------------------------
 * ll    : Likelihood.__call__  model.py:145-184   (use_source: 0 -> :174-176, 1 -> :177-184)
 * prior : Prior.__call__       model.py:484-505
 * geo   : GeoPrior.__call__    model.py:1018-1043 / geo_prior_distance :1096-1139 (quirk Q1: last zone)
 * Outputs are host arrays of n_chains doubles (any may be NULL).  Also refreshes the running ll/prior of
 * the sampler for those chains. */
int sb_eval(sb_ctx* ctx, int chain_begin, int n_chains, int use_source, double* ll, double* prior, double* geo);
/* launch-only variant for timing the likelihood kernel with device-resident inputs (no D2H, no sync) */
int sb_eval_likelihood_async(sb_ctx* ctx, int chain_begin, int n_chains, int use_source);
int sb_read_ll(sb_ctx* ctx, int chain_begin, int n_chains, double* ll);
/* Allocation-free variant of the source-free likelihood evaluation (kernel K1).  It has NO counterpart in the reference:
 * it was added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  It evaluates Likelihood.__call__ (model.py:174-176) of the
 * resident chains [chain_begin, chain_begin + n_chains) like sb_eval_likelihood_async, but the kernel's scratch (per-CTA
 * partial sums + per-chain tickets) is a file-scope `static __device__` array of 528 KiB inside the library instead of
 * context-owned cudaMalloc memory, and the call itself performs no cudaMalloc / cudaFree / allocating library call.
 *   ll_out : n_chains doubles, DEVICE memory (written stream-ordered on the context's stream, no synchronisation) or
 *            HOST memory (filled and synchronised before the call returns).
 * Limits (SB_E_ARG with a message beyond them): n_chains <= 4096 per call and n_chains * CTAs-per-chain <= 65536
 * partial sums (a launch uses about max(n_chains, 2 * 148 * 14) of them); n_states <= 31; a context whose family-class
 * tables do not fit the kernel's shared memory is refused.  Every call on a GPU shares the one array: calls that
 * overlap on the GPU (different streams or contexts on the same device) race; serialise them.  The first call on a
 * device resolves the array's address (cudaGetSymbolAddress) -- make it once outside a graph capture. */
int sb_eval_likelihood_noalloc(sb_ctx* ctx, int chain_begin, int n_chains, double* ll_out);

/* debug / parity: per-cell arrays of update_component_likelihoods (:230-249) and update_weights (:257-294) */
int sb_eval_components(sb_ctx* ctx, int chain, double* all_lh /* (N,F,C) */, double* wnorm /* (N,F,C) */);
/* source posterior of gibbs_sample_sources (zone_sampling.py:196-202) */
int sb_source_posterior(sb_ctx* ctx, int chain, double* post /* (N,F,C) */);

/* ---- integer count tables of the Gibbs operators (kernel K4), bit-exact parity surface ----------
 * cnt_glob (F,S)      zone_sampling.py:339-350     cnt_zone (Z,F,S)   :366-372
 * cnt_fam  (Fam,F,S)  :389-400                     cnt_src_zone (F,C) :239-240,256-257
 * cnt_src_fam (F,C)   :283-284
 * recount != 0 recomputes from the resident source state; 0 returns the incrementally maintained tables. */
/* launch-only full recount (kernel K4) of chains [chain_begin, chain_begin + n_chains) for timing with device-resident
 * inputs (no D2H, no synchronisation) */
int sb_recount_async(sb_ctx* ctx, int chain_begin, int n_chains);
int sb_counts(sb_ctx* ctx, int chain, int recount, int32_t* cnt_glob, int32_t* cnt_zone, int32_t* cnt_fam,
              int32_t* cnt_src_zone, int32_t* cnt_src_fam);

/* ---- zone-move proposal probabilities for a forced move (parity of zone_sampling.py:704-933, warm-up :1328-1574) ----
 * kind: SB_OP_GROW_ZONE / SB_OP_SHRINK_ZONE / SB_OP_SWAP_ZONE; unused site = -1.  Does not modify state.
 * Runs the step kernel itself (same set-up, same proposal code as a sampling launch) with the zone and the sites
 * given instead of drawn. */
int sb_move_log_q(sb_ctx* ctx, int chain, int kind, int zone, int site_new, int site_removed,
                  double p_grow_connected, double* log_q, double* log_q_back);
/* log_q - log_q_back of dirichlet_proposal (zone_sampling.py:538-569) for a forced new pair: the current pair (c0, c1)
 * of weights / probabilities, the proposed pair (n0, n1) with n0 + n1 = c0 + c1, the proposal precision -- evaluated by
 * the code alter_weights / alter_p_* use inside the step kernel (:408-612). */
int sb_pair_log_q(sb_ctx* ctx, double c0, double c1, double n0, double n1, double precision, double* log_q_ratio);

/* ---- sampler: MCMCGenerative.step / generate_samples (mcmc_generative.py:149-351) ---------------- */
int sb_set_operators(sb_ctx* ctx, const sb_operator_cfg* cfg,
                     const double* p_grow_connected,  /* (n_chains) zone_sampling.py:138,1288-1291 */
                     const int32_t* max_size);        /* (n_chains) MAX_M / warm-up list :1280-1285 */
/* Run n_steps MH steps on every resident chain in one persistent launch (kernel K8). Asynchronous. */
int sb_run(sb_ctx* ctx, int64_t n_steps);
/* sample_from_prior of MCMCGenerative (mcmc_generative.py:115-116): the likelihood is identically 0 in every
 * Metropolis ratio and in the logged values.  Takes effect at the next sb_run. */
int sb_set_sample_from_prior(sb_ctx* ctx, int on);
/* The same launch with host-visible EVENTS inside it (generate_samples' logging and MC3 swap points,
 * mcmc_generative.py:205-231) instead of one launch per event: after event_steps[k] completed steps of this launch
 * (ascending, 1 <= step <= n_steps) every chain records (ll, prior); if event_slots[k] >= 0 the state of chain
 * `ring_chain` (-1 = of every chain: under MC3 the logged label moves between chains) is copied into snapshot slot
 * event_slots[k] of a device ring.  Asynchronous; read with sb_read_event_scalars / sb_read_snapshot.
 * The ring grows to the largest slot ever named and keeps its content, and sb_read_snapshot(slot) waits only for the
 * launch that wrote that slot (it copies on a stream of its own): a caller that alternates between two ranges of slots
 * can enqueue launch b + 1 right after sb_read_event_scalars of launch b and read b's snapshots while b + 1 runs
 * (contact_zones_b200/sampling.py:generate_samples does). */
int sb_run_events(sb_ctx* ctx, int64_t n_steps, int n_events, const int32_t* event_steps, const int32_t* event_slots,
                  int ring_chain);
int sb_read_event_scalars(sb_ctx* ctx, int n_events, double* ll, double* prior);   /* (n_events, n_chains) each */
int sb_read_snapshot(sb_ctx* ctx, int slot, int chain, uint8_t* zones, double* weights, double* p_global,
                     double* p_zones, double* p_families /* any may be NULL */);
/* DEVICE address and layout of the snapshot ring, for callers that move logged samples between GPUs without a host hop
 * (contact_zones_b200/mc3.py:gather_records all-gathers the records with NCCL): record (slot, chain) is record_bytes
 * bytes at index slot * n_chains + chain; field a of a record -- 0 zone ids int8[N] (-1 = no zone), 1 weights f64[F][3],
 * 2 p_global f64[F][S], 3 p_zones f64[Z][F][S], 4 p_families f64[Fam][F][S] -- is lengths[a] bytes at offsets[a].  The
 * address is valid until the next sb_run_events that names a larger slot (ring == NULL before the first one); the
 * caller orders its reads behind the launch that wrote a slot (sb_read_event_scalars of that launch has returned).
 * No reference counterpart (the reference logs Python objects, mcmc_generative.py:360-365). */
int sb_snapshot_ring(sb_ctx* ctx, void** ring, int64_t* record_bytes, int64_t* n_records, int64_t offsets[5], int64_t lengths[5]);
int sb_read_chain_scalars(sb_ctx* ctx, double* ll, double* prior);              /* (n_chains) each; SB_E_NUMERIC if flagged */
/* DEVICE addresses of the running (ll, prior) arrays, n_chains doubles each, valid until sb_destroy: the MC3 exchange
 * (mcmc_generative.py:242-280) all-gathers them with NCCL on the context's stream without a host hop */
int sb_device_scalars(sb_ctx* ctx, void** ll, void** prior);
int sb_chain_errors(sb_ctx* ctx, int32_t* flags);                                /* (n_chains) SB_ERR_* bits, sticky until sb_set_state */
int sb_read_op_stats(sb_ctx* ctx, int64_t* accepted, int64_t* rejected);        /* (n_chains, SB_N_OPS) each */
int sb_step_counter(sb_ctx* ctx, int64_t* steps_done);                          /* per context      */
int sb_kernel_launches(sb_ctx* ctx, int64_t* launches);                         /* launches issued so far */

/* ---- K6 parity surface: n independent draws of the device generator (Philox4x32-10 + Marsaglia-Tsang), which
 *      replaces np.random.dirichlet / scipy.stats.beta.rvs (zone_sampling.py:243,262-265,351,373,401,560).
 *      kind 0 = uniform(0,1), 1 = Gamma(alpha, 1), 2 = the step kernel's table-driven log of
 *      (1 + u_i) * 2^(i % 121 - 60), u_i the kind-0 draw (accuracy surface), 3 = the standard-normal candidates of
 *      the Gamma sampler (draws 2j, 2j+1 = the Box-Muller pair of block j). */
int sb_rng_sample(sb_ctx* ctx, int kind, double alpha, uint64_t seed, int n, double* out);

#ifdef __cplusplus
}
#endif
#endif /* SBAYES_B200_H */
