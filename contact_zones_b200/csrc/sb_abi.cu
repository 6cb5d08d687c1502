// sb_abi.cu -- host side of libsbayes_b200.so: the C ABI declared in include/sbayes_b200.h.
// Pure CUDA runtime (no torch types); built by __graft_entry__.build() with
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include "../../include/sbayes_b200.h"
#include "sb_common.cuh"
#include "sb_eval.cuh"
#include "sb_lhtab.cuh"
#include "sb_step.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

// pinned upload staging buffer of sb_set_state
struct Staging {
    unsigned char* buf = nullptr;
    cudaEvent_t done = nullptr;
    bool pending = false;
};

struct sb_ctx {
    int device = 0;
    sb_dims dims{};
    DevData d{};
    ChainArrays st{};
    cudaStream_t own_stream = nullptr, stream = nullptr;
    // snapshot read-back next to a launch in flight: sb_read_snapshot copies on its own stream, after the event of the
    // launch that wrote the slot (two events, by launch parity: waiting for a later launch of the same parity is safe)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t run_done[2] = {nullptr, nullptr};
    long long n_event_launches = 0;
    std::vector<long long> slot_launch;          // per ring slot: the sb_run_events launch that wrote it last (0 = never)
    std::string err;
    bool has_data = false, has_priors = false, has_ops = false;
    std::vector<void*> allocs;
    double* partial = nullptr;       // [n_chains][n_tiles]
    int adj_nnz = 0;                 // entries of the adjacency CSR
    cudaEvent_t direct_done = nullptr;   // last direct (unstaged) upload from page-locked caller memory
    uint8_t* src_scratch = nullptr;      // device scratch of the unstaged source upload ([chains][N][F], unpadded)
    size_t src_scratch_bytes = 0;
    int n_tiles = 0, sites_per_tile = 0;
    LhTab lt{};                      // static layout of the table-driven K1 (sb_lhtab.cuh)
    double* lt_partial = nullptr;    // [n_chains][lt.n_tiles * lt.n_chunks]
    unsigned* lt_ticket = nullptr;   // [n_chains]
    long lt_want = 148L * 14;        // CTA target of one K1 launch (SB_K1_CTAS overrides, for tuning)
    int k1b_threads = LH_THREADS, k1s_threads = LH_THREADS;   // CTA sizes of the blocked likelihood kernels (SB_K1B_THREADS, SB_K1S_THREADS)
    int n_sms = 148;                 // cudaDevAttrMultiProcessorCount of the context's device
    int k1b_parts = 0;               // CTAs per chain of the blocked source-free kernel (0 = chosen per launch; SB_K1B_PARTS)
    int k1c_max_n = 12288, k1c_min_chains = 96;   // one-CTA-per-chain table kernel: site and chain range (SB_K1C_MAX_N, SB_K1C_MIN_CHAINS)
    int k1_block_max_n = 1024;       // source-free launches over at most this many sites take the blocked kernel (SB_K1_BLOCK_MAX_N)
    long lt_want_compact = 148L * 7; // the same for shapes with one site per code byte (SB_K1_CTAS_COMPACT; 444..2072 measured)
    std::vector<int> h_gclass;       // host copy of lt.group_class (class span of a group range)
    int geo_cap = 0;                 // capacity (members) of the Prim scratch
    int64_t launches = 0;
    int64_t steps_done = 0;
    StepParams sp{};                 // operator configuration of the persistent kernel
    double* d_pgrow = nullptr;
    int* d_maxsize = nullptr;
    size_t step_smem = 0;
    std::vector<int> h_zone_size;    // [chains][Z] host mirror kept by set_state (shape checks only)
    struct Staging stage[2];         // pinned upload staging (sb_set_state)
    size_t stage_bytes = 0;
    int stage_next = 0;
    bool has_layout = false;         // shared-memory layout / team size of the step kernel chosen (configure_step)
    ForcedMove* d_forced = nullptr;  // device record of sb_move_log_q
    double* d_pair = nullptr;        // device result of sb_pair_log_q
    // events of sb_run_events: device copies + the snapshot ring (grown on demand, freed by sb_destroy)
    int2* d_events = nullptr;        int events_cap = 0;
    double2* d_ev_scalars = nullptr; size_t ev_scalars_cap = 0;
    unsigned char* d_ring = nullptr; size_t ring_cap = 0;
    size_t ring_rec = 0;
    int ring_slots = 0;              // slots of the last sb_run_events
};

#define CK(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess) {                                                                  \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                        \
            return SB_E_CUDA;                                                                     \
        }                                                                                         \
    } while (0)
#define FAIL(code, msg)        \
    do {                       \
        ctx->err = (msg);      \
        return (code);         \
    } while (0)

template <typename T>
static int dalloc(sb_ctx* ctx, T** p, size_t n) {
    if (n == 0) n = 1;
    CK(cudaMalloc((void**)p, n * sizeof(T)));
    ctx->allocs.push_back((void*)*p);
    CK(cudaMemsetAsync(*p, 0, n * sizeof(T), ctx->stream));
    return SB_OK;
}
#define DA(p, n)                                       \
    do {                                               \
        int r_ = dalloc(ctx, &(p), (size_t)(n));       \
        if (r_ != SB_OK) return r_;                    \
    } while (0)

template <typename T>
static int h2d(sb_ctx* ctx, T* dst, const T* src, size_t n) {
    if (n == 0) return SB_OK;
    CK(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));   // host buffers are borrowed for the duration of the call only
    return SB_OK;
}
template <typename T>
static int d2h(sb_ctx* ctx, T* dst, const T* src, size_t n) {
    if (n == 0) return SB_OK;
    CK(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SB_OK;
}
#define H2D(dst, src, n)                                 \
    do {                                                 \
        int r_ = h2d(ctx, (dst), (src), (size_t)(n));    \
        if (r_ != SB_OK) return r_;                      \
    } while (0)
#define D2H(dst, src, n)                                 \
    do {                                                 \
        int r_ = d2h(ctx, (dst), (src), (size_t)(n));    \
        if (r_ != SB_OK) return r_;                      \
    } while (0)

static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

extern "C" {

int sb_version(void) { return 100; }

int sb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

const char* sb_last_error(const sb_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int sb_create(sb_ctx** out, int device, const sb_dims* dims) {
    if (!out || !dims) return SB_E_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return SB_E_NOGPU;
    if (device < 0 || device >= ndev) return SB_E_ARG;
    if (dims->n_sites < 1 || dims->n_features < 1 || dims->n_states < 1 || dims->n_states > SB_MAX_S ||
        dims->n_zones < 0 || dims->n_zones > SB_MAX_Z || dims->n_families < 0 || dims->n_families > 32767 ||
        dims->n_chains < 1)
        return SB_E_ARG;
    sb_ctx* ctx = new sb_ctx();
    ctx->device = device;
    ctx->dims = *dims;
    if (cudaDeviceGetAttribute(&ctx->n_sms, cudaDevAttrMultiProcessorCount, ctx->device) != cudaSuccess || ctx->n_sms < 1) ctx->n_sms = 148;
    if (const char* e = getenv("SB_K1_CTAS")) ctx->lt_want = std::max(1L, atol(e));
    if (const char* e = getenv("SB_K1_CTAS_COMPACT")) ctx->lt_want_compact = std::max(1L, atol(e));
    if (const char* e = getenv("SB_K1_BLOCK_MAX_N")) ctx->k1_block_max_n = atoi(e);
    if (const char* e = getenv("SB_K1C_MAX_N")) ctx->k1c_max_n = atoi(e);
    if (const char* e = getenv("SB_K1C_MIN_CHAINS")) ctx->k1c_min_chains = atoi(e);
    if (const char* e = getenv("SB_K1B_PARTS")) ctx->k1b_parts = atoi(e);
    if (const char* e = getenv("SB_K1B_THREADS")) ctx->k1b_threads = std::min(LH_THREADS, std::max(32, atoi(e) & ~31));
    if (const char* e = getenv("SB_K1S_THREADS")) ctx->k1s_threads = std::min(LH_THREADS, std::max(32, atoi(e) & ~31));
    *out = ctx;   // returned even on failure so that the caller can read sb_last_error, then sb_destroy
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    ctx->stream = ctx->own_stream;

    DevData& d = ctx->d;
    d.N = dims->n_sites; d.F = dims->n_features; d.S = dims->n_states; d.Z = dims->n_zones;
    d.inheritance = (dims->flags & SB_F_INHERITANCE) ? 1 : 0;
    d.sample_source = (dims->flags & SB_F_SAMPLE_SOURCE) ? 1 : 0;
    d.warmup = (dims->flags & SB_F_WARMUP) ? 1 : 0;
    d.Fam = d.inheritance ? dims->n_families : 0;
    d.C = d.inheritance ? 3 : 2;
    d.ldx = round_up(d.F, 16);
    d.ldn = round_up(d.N, 16);
    if (d.inheritance && d.Fam < 1) FAIL(SB_E_ARG, "inheritance needs n_families >= 1");

    const size_t nc = (size_t)dims->n_chains;
    ChainArrays& st = ctx->st;
    const int N = d.N, F = d.F, S = d.S, Z = d.Z, Fam = d.Fam;
    st.s_zone_id = (size_t)round_up(N, 16);        DA(st.zone_id, nc * st.s_zone_id);
    st.s_zone_members = (size_t)Z * N;              DA(st.zone_members, nc * st.s_zone_members);
    st.s_zone_size = (size_t)(Z > 0 ? Z : 1);       DA(st.zone_size, nc * st.s_zone_size);
    st.s_weights = (size_t)F * 3;                   DA(st.weights, nc * st.s_weights);
    st.s_wnorm = (size_t)F * 12;                    DA(st.wnorm, nc * st.s_wnorm);
    st.s_p_global = (size_t)F * S;                  DA(st.p_global, nc * st.s_p_global);
    st.s_p_zones = (size_t)Z * F * S;               DA(st.p_zones, nc * st.s_p_zones);
    st.s_p_fam = (size_t)Fam * F * S;               DA(st.p_fam, nc * st.s_p_fam);
    st.s_source = d.sample_source ? (size_t)N * d.ldx : 0;   DA(st.source, nc * st.s_source);
    st.s_cnt_glob = (size_t)F * S;                  DA(st.cnt_glob, nc * st.s_cnt_glob);
    st.s_cnt_zone = (size_t)Z * F * S;              DA(st.cnt_zone, nc * st.s_cnt_zone);
    st.s_cnt_fam = (size_t)Fam * F * S;             DA(st.cnt_fam, nc * st.s_cnt_fam);
    st.s_cnt_src = (size_t)F * 12;                  DA(st.cnt_src, nc * st.s_cnt_src);
    DA(st.ll, nc); DA(st.prior, nc); DA(st.geo, nc);
    st.s_prior_pg = (size_t)F;                      DA(st.prior_pg, nc * st.s_prior_pg);
    st.s_prior_pf = (size_t)Fam * F;                DA(st.prior_pf, nc * st.s_prior_pf);
    DA(st.acc, nc * SB_N_OPS); DA(st.rej, nc * SB_N_OPS); DA(st.step, nc); DA(st.err, nc);

    // K1 tiling: tiles x chains CTAs; aim at >= 4 waves of 148 SMs x 8 resident CTAs when the shape allows
    int spt = 64;
    ctx->sites_per_tile = spt;
    ctx->n_tiles = (N + spt - 1) / spt;
    DA(ctx->partial, nc * (size_t)std::max(ctx->n_tiles, d.ldx >> 4));     // per site tile or per 16-feature block
    ctx->geo_cap = N < 4096 ? N : 4096;
    ctx->h_zone_size.assign(nc * (size_t)(Z > 0 ? Z : 1), 0);
    CK(cudaStreamSynchronize(ctx->stream));
    return SB_OK;
}

#ifdef SB_PROF
extern "C" int sb_prof_read(unsigned long long* out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, sb_prof, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(sb_prof, z, sizeof(z)); }
    return 0;
}
#endif

int sb_destroy(sb_ctx* ctx) {
    if (!ctx) return SB_OK;
    cudaSetDevice(ctx->device);
    if (ctx->own_stream) cudaStreamSynchronize(ctx->own_stream);
    for (void* p : ctx->allocs) cudaFree(p);
    for (Staging& sg : ctx->stage) {
        if (sg.done) { cudaEventSynchronize(sg.done); cudaEventDestroy(sg.done); }
        if (sg.buf) cudaFreeHost(sg.buf);
    }
    if (ctx->direct_done) cudaEventDestroy(ctx->direct_done);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    for (cudaEvent_t e : ctx->run_done)
        if (e) cudaEventDestroy(e);
    if (ctx->src_scratch) cudaFree(ctx->src_scratch);
    if (ctx->d_events) cudaFree(ctx->d_events);
    if (ctx->d_ev_scalars) cudaFree(ctx->d_ev_scalars);
    if (ctx->d_ring) cudaFree(ctx->d_ring);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
    return SB_OK;
}

int sb_set_stream(sb_ctx* ctx, void* cuda_stream) {
    if (!ctx) return SB_E_ARG;
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return SB_OK;
}

int sb_sync(sb_ctx* ctx) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    return SB_OK;
}

int sb_upload_data(sb_ctx* ctx, const uint8_t* x, const uint8_t* applicable, const int16_t* family_id,
                   const int32_t* adj_indptr, const int32_t* adj_indices, const double* cost) {
    if (!ctx || !x || !applicable) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    DevData& d = ctx->d;
    const int N = d.N, F = d.F, S = d.S;
    if (d.inheritance && !family_id) FAIL(SB_E_ARG, "family_id required with inheritance");
    if (d.geo_type == SB_GEO_COST_BASED && !cost) FAIL(SB_E_ARG, "cost matrix required by cost_based geo prior");

    std::vector<uint8_t> hx((size_t)N * d.ldx, SB_PAD), hxT((size_t)F * d.ldn, SB_PAD);
    for (int n = 0; n < N; ++n)
        for (int f = 0; f < F; ++f) {
            const uint8_t v = x[(size_t)n * F + f];
            if (v != SB_NA && v >= S) FAIL(SB_E_ARG, "state index out of range");
            hx[(size_t)n * d.ldx + f] = v;
            hxT[(size_t)f * d.ldn + n] = v;
        }
    std::vector<uint32_t> mask(F, 0);
    std::vector<int> napp(F, 0);
    for (int f = 0; f < F; ++f)
        for (int s = 0; s < S; ++s)
            if (applicable[(size_t)f * S + s]) { mask[f] |= 1u << s; ++napp[f]; }
    for (int f = 0; f < F; ++f)
        if (napp[f] < 1) FAIL(SB_E_ARG, "feature without applicable state");

    uint8_t *dx, *dxT; uint32_t* dmask; int* dnapp;
    DA(dx, hx.size()); DA(dxT, hxT.size()); DA(dmask, F); DA(dnapp, F);
    H2D(dx, hx.data(), hx.size()); H2D(dxT, hxT.data(), hxT.size());
    H2D(dmask, mask.data(), F); H2D(dnapp, napp.data(), F);
    d.x = dx; d.xT = dxT; d.app_mask = dmask; d.n_app = dnapp;

    // families: id per site + CSR member lists (ascending site order)
    std::vector<int16_t> fid(N, -1);
    std::vector<int> fptr(d.Fam + 1, 0), fsites;
    if (d.inheritance) {
        for (int n = 0; n < N; ++n) {
            fid[n] = family_id[n];
            if (fid[n] >= d.Fam || fid[n] < -1) FAIL(SB_E_ARG, "family id out of range");
            if (fid[n] >= 0) ++fptr[fid[n] + 1];
        }
        for (int k = 0; k < d.Fam; ++k) fptr[k + 1] += fptr[k];
        fsites.resize(fptr[d.Fam] > 0 ? fptr[d.Fam] : 1);
        std::vector<int> pos(fptr.begin(), fptr.end() - 1);
        for (int n = 0; n < N; ++n)
            if (fid[n] >= 0) fsites[pos[fid[n]]++] = n;
    } else {
        fsites.resize(1);
    }
    int16_t* dfid; int *dfptr, *dfs;
    DA(dfid, N); DA(dfptr, fptr.size()); DA(dfs, fsites.size());
    H2D(dfid, fid.data(), N); H2D(dfptr, fptr.data(), fptr.size()); H2D(dfs, fsites.data(), fsites.size());
    d.fam_id = dfid; d.fam_ptr = dfptr; d.fam_sites = dfs;

    // adjacency CSR
    int *dptr = nullptr, *didx = nullptr;
    if (adj_indptr && adj_indices) {
        const int nnz = adj_indptr[N];
        ctx->adj_nnz = nnz;
        for (int i = 0; i < nnz; ++i)
            if (adj_indices[i] < 0 || adj_indices[i] >= N) FAIL(SB_E_ARG, "adjacency index out of range");
        DA(dptr, N + 1); DA(didx, nnz);
        H2D(dptr, adj_indptr, N + 1); H2D(didx, adj_indices, nnz);
    }
    d.adj_ptr = dptr; d.adj_idx = didx;

    double* dcost = nullptr;
    if (cost) {
        DA(dcost, (size_t)N * N);
        H2D(dcost, cost, (size_t)N * N);
    }
    d.cost = dcost;

    // ---- static layout of the table-driven K1 (sb_lhtab.cuh): class-ordered groups + packed state codes ----
    {
        LhTab& t = ctx->lt;
        t.radix = S + 1;
        t.cpb = 1;
        while (t.cpb < 3 && std::pow((double)t.radix, t.cpb + 1) <= (double)LT_ROW) ++t.cpb;
        t.row_entries = 1;
        for (int i = 0; i < t.cpb; ++i) t.row_entries *= t.radix;
        t.G = 16 * t.cpb;
        const int ncls_all = d.Fam + 1;                      // class = family id, "no family" last
        std::vector<std::vector<int>> by_class(ncls_all);
        for (int n = 0; n < N; ++n) by_class[fid[n] >= 0 ? fid[n] : d.Fam].push_back(n);
        std::vector<int> perm, gclass, cfam;
        for (int c = 0; c < ncls_all; ++c) {
            if (by_class[c].empty()) continue;
            const int ci = (int)cfam.size();
            cfam.push_back(c < d.Fam ? c : -1);
            for (size_t i = 0; i < by_class[c].size(); ++i) {
                if (i % t.G == 0) gclass.push_back(ci);
                perm.push_back(by_class[c][i]);
            }
            while (perm.size() % t.G) perm.push_back(-1);
        }
        t.n_groups = (int)gclass.size();
        t.n_classes = (int)cfam.size();
        ctx->h_gclass = gclass;
        std::vector<uint8_t> code((size_t)F * t.n_groups * 16);
        for (int f = 0; f < F; ++f)
            for (int g = 0; g < t.n_groups; ++g)
                for (int b = 0; b < 16; ++b) {
                    int c = 0, mul = 1;
                    for (int i = 0; i < t.cpb; ++i) {
                        const int n = perm[(size_t)g * t.G + b * t.cpb + i];
                        int digit = S;
                        if (n >= 0 && x[(size_t)n * F + f] != SB_NA) digit = x[(size_t)n * F + f];
                        c += digit * mul;
                        mul *= t.radix;
                    }
                    code[((size_t)f * t.n_groups + g) * 16 + b] = (uint8_t)(4 * c);   // byte offset of the table entry
                }
        // states that occur in the data count as applicable for the zero / tiny check of the table rows
        for (int n = 0; n < N; ++n)
            for (int f = 0; f < F; ++f)
                if (x[(size_t)n * F + f] != SB_NA) mask[f] |= 1u << x[(size_t)n * F + f];
        uint32_t* dchk;
        DA(dchk, F);
        H2D(dchk, mask.data(), F);
        t.chk_mask = dchk;
        uint8_t* dcode; int *dgc, *dcf, *dperm;
        DA(dcode, code.size()); DA(dgc, gclass.size()); DA(dcf, cfam.size()); DA(dperm, perm.size());
        H2D(dcode, code.data(), code.size()); H2D(dgc, gclass.data(), gclass.size());
        H2D(dcf, cfam.data(), cfam.size()); H2D(dperm, perm.data(), perm.size());
        t.xcode = dcode; t.group_class = dgc; t.class_fam = dcf; t.perm = dperm;
        DA(ctx->lt_partial, (size_t)ctx->dims.n_chains * ((t.n_groups + LT_THREADS - 1) / LT_THREADS) * (F + LT_CORR_CTAS));
        DA(ctx->lt_ticket, ctx->dims.n_chains);
        CK(cudaStreamSynchronize(ctx->stream));
    }
    ctx->has_data = true;
    return SB_OK;
}

int sb_upload_locations(sb_ctx* ctx, const double* xy) {
    if (!ctx || !xy) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (!ctx->has_data) FAIL(SB_E_STATE, "sb_upload_data must precede sb_upload_locations");
    for (int i = 0; i < 2 * ctx->d.N; ++i)
        if (!std::isfinite(xy[i])) FAIL(SB_E_ARG, "non-finite site coordinate");
    double* dloc;
    DA(dloc, (size_t)2 * ctx->d.N);
    H2D(dloc, xy, (size_t)2 * ctx->d.N);
    ctx->d.loc = dloc;
    return SB_OK;
}

int sb_set_priors(sb_ctx* ctx, const sb_prior_cfg* cfg, const double* gibbs_g, const double* alpha_g,
                  const double* gibbs_f, const double* alpha_f) {
    if (!ctx || !cfg) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    DevData& d = ctx->d;
    if (!ctx->has_data) FAIL(SB_E_STATE, "sb_upload_data must precede sb_set_priors");
    if (cfg->size_type < 0 || cfg->size_type > 2 || cfg->geo_type < 0 || cfg->geo_type > 1 ||
        cfg->p_global_type < 0 || cfg->p_global_type > 1 || cfg->p_families_type < 0 || cfg->p_families_type > 1)
        FAIL(SB_E_ARG, "unknown prior type");
    if (cfg->geo_type == SB_GEO_COST_BASED && !d.cost && !d.loc)
        FAIL(SB_E_ARG, "cost_based geo prior needs a cost matrix (sb_upload_data) or site locations (sb_upload_locations)");
    if (cfg->geo_type == SB_GEO_COST_BASED && !(cfg->geo_scale > 0)) FAIL(SB_E_ARG, "geo scale must be > 0");
    if (cfg->p_global_type == SB_DIR_COUNTS && !alpha_g) FAIL(SB_E_ARG, "alpha_global required");
    if (d.inheritance && cfg->p_families_type == SB_DIR_COUNTS && !alpha_f) FAIL(SB_E_ARG, "alpha_fam required");
    d.size_type = cfg->size_type; d.geo_type = cfg->geo_type; d.geo_scale = cfg->geo_scale;
    d.pg_type = cfg->p_global_type; d.pf_type = d.inheritance ? cfg->p_families_type : 0;
    const int F = d.F, S = d.S, Fam = d.Fam;
    const size_t FS = (size_t)F * S;

    auto upload_rows = [&](const double* gibbs, const double* alpha, size_t rows, const double** out_gibbs,
                           const double** out_alpha, const double** out_lnorm) -> int {
        std::vector<double> g(rows * S, 1.0), a(rows * S, 1.0), ln(rows, 0.0);
        if (gibbs) memcpy(g.data(), gibbs, g.size() * sizeof(double));
        if (alpha) {
            memcpy(a.data(), alpha, a.size() * sizeof(double));
            std::vector<uint32_t> mask(F);
            CK(cudaMemcpy(mask.data(), d.app_mask, F * sizeof(uint32_t), cudaMemcpyDeviceToHost));
            for (size_t r = 0; r < rows; ++r) {
                double sa = 0.0, sl = 0.0;
                for (int s = 0; s < S; ++s)
                    if (mask[r % F] >> s & 1u) { sa += a[r * S + s]; sl += std::lgamma(a[r * S + s]); }
                ln[r] = std::lgamma(sa) - sl;
            }
        }
        double *dg, *da, *dl;
        DA(dg, g.size()); DA(da, a.size()); DA(dl, ln.size());
        H2D(dg, g.data(), g.size()); H2D(da, a.data(), a.size()); H2D(dl, ln.data(), ln.size());
        *out_gibbs = dg; *out_alpha = da; *out_lnorm = dl;
        return SB_OK;
    };
    int r = upload_rows(gibbs_g, d.pg_type ? alpha_g : nullptr, F, &d.gibbs_g, &d.alpha_g, &d.lnorm_g);
    if (r != SB_OK) return r;
    if (d.inheritance) {
        r = upload_rows(gibbs_f, d.pf_type ? alpha_f : nullptr, (size_t)Fam * F, &d.gibbs_f, &d.alpha_f, &d.lnorm_f);
        if (r != SB_OK) return r;
    }
    (void)FS;
    ctx->has_priors = true;
    return SB_OK;
}

// ---------------------------------------------------------------------------------------------
// per-chain state
// ---------------------------------------------------------------------------------------------
static int check_chain(sb_ctx* ctx, int chain) {
    if (chain < 0 || chain >= ctx->dims.n_chains) FAIL(SB_E_ARG, "chain index out of range");
    if (!ctx->has_data) FAIL(SB_E_STATE, "sb_upload_data has not been called");
    return SB_OK;
}

// SBAYES_B200_SWEEP=0 (tests) also selects the tile-per-site fallbacks of the evaluation kernels: no shared-memory
// tables anywhere
static bool blocked_kernels_allowed() {
    const char* env = getenv("SBAYES_B200_SWEEP");
    return !(env && atoi(env) < 1);
}

static int recount_chains(sb_ctx* ctx, int chain, int n) {
    const DevData& d = ctx->d;
    ChainArrays& st = ctx->st;
    const size_t nn = (size_t)n;
    const size_t cb_smem = count_block_smem(d.S, d.Z, d.Fam);
    if (d.sample_source && cb_smem <= 200 * 1024 && n <= 65535 && blocked_kernels_allowed()) {
        // blocked recount: shared-memory counters per (16-feature block, chain), every table entry written once
        if (cb_smem > 48 * 1024)
            CK(cudaFuncSetAttribute(count_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cb_smem));
        dim3 grid((unsigned)(d.ldx >> 4), (unsigned)n);
        count_block_kernel<<<grid, 256, cb_smem, ctx->stream>>>(d, st, chain);
        ++ctx->launches;
        CK(cudaGetLastError());
        return SB_OK;
    }
    CK(cudaMemsetAsync(st.cnt_glob + chain * st.s_cnt_glob, 0, nn * st.s_cnt_glob * sizeof(int), ctx->stream));
    CK(cudaMemsetAsync(st.cnt_zone + chain * st.s_cnt_zone, 0, nn * st.s_cnt_zone * sizeof(int), ctx->stream));
    CK(cudaMemsetAsync(st.cnt_fam + chain * st.s_cnt_fam, 0, nn * st.s_cnt_fam * sizeof(int), ctx->stream));
    CK(cudaMemsetAsync(st.cnt_src + chain * st.s_cnt_src, 0, nn * st.s_cnt_src * sizeof(int), ctx->stream));
    if (!d.sample_source) return SB_OK;
    dim3 grid(ctx->n_tiles, n);
    count_kernel<<<grid, 256, 0, ctx->stream>>>(d, st, chain, ctx->sites_per_tile);
    ++ctx->launches;
    CK(cudaGetLastError());
    return SB_OK;
}
static int recount_chain(sb_ctx* ctx, int chain) { return recount_chains(ctx, chain, 1); }

// Upload path: the caller's arrays are packed into a pinned staging buffer (two of them, used alternately, so that
// packing chain c+1 overlaps the DMA of chain c), every device array is filled by one asynchronous copy from there,
// and the call returns without a stream synchronisation -- the borrowed host pointers are not touched after return.
static size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

static int staging_acquire(sb_ctx* ctx, size_t bytes, Staging** out) {
    if (ctx->stage_bytes < bytes) {
        for (Staging& sg : ctx->stage) {
            if (sg.pending) { CK(cudaEventSynchronize(sg.done)); sg.pending = false; }
            if (sg.buf) cudaFreeHost(sg.buf);
            sg.buf = nullptr;
            CK(cudaMallocHost((void**)&sg.buf, bytes));
            if (!sg.done) CK(cudaEventCreateWithFlags(&sg.done, cudaEventDisableTiming));
        }
        ctx->stage_bytes = bytes;
    }
    Staging& sg = ctx->stage[ctx->stage_next];
    ctx->stage_next ^= 1;
    if (sg.pending) { CK(cudaEventSynchronize(sg.done)); sg.pending = false; }
    *out = &sg;
    return SB_OK;
}

// rows of F bytes -> rows of ld bytes (ld = F rounded up to 16), padding zero; one thread per 16-byte unit
__global__ void pad_rows_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, size_t n_rows, int F, int ld) {
    const int upr = ld >> 4;
    const size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_rows * upr) return;
    const size_t r = u / upr;
    const int b0 = (int)(u - r * upr) << 4;
    const uint8_t* s = src + r * (size_t)F + b0;
    uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if (b0 + i < F) w[i >> 2] |= (uint32_t)s[i] << ((i & 3) * 8);
    *reinterpret_cast<uint4*>(dst + r * (size_t)ld + b0) = make_uint4(w[0], w[1], w[2], w[3]);
}

// Page-locked (cudaMallocHost / cudaHostRegister / torch pin_memory) caller memory can be read by the copy engine
// directly; pageable memory has to go through the pinned staging buffer.
static bool host_pinned(const void* p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// n consecutive chains from arrays stacked along a leading chain axis.  Staging layout = device layout (array-major,
// per-chain strides st.s_*), so every device array is filled by ONE copy whatever n is.
static int set_states_impl(sb_ctx* ctx, int chain, int n, const uint8_t* zones, const double* weights,
                           const double* p_global, const double* p_zones, const double* p_families,
                           const uint8_t* source) {
    const DevData& d = ctx->d;
    ChainArrays& st = ctx->st;
    const int N = d.N, F = d.F, S = d.S, Z = d.Z, C = d.C, Fam = d.Fam;
    if (!weights || !p_global || (Z > 0 && (!zones || !p_zones)) || (d.inheritance && !p_families))
        FAIL(SB_E_ARG, "sb_set_state: missing array");
    if (d.sample_source && !source) FAIL(SB_E_ARG, "sb_set_state: SAMPLE_SOURCE context needs a source array");

    const size_t FS = (size_t)F * S;
    const size_t o_zid = 0;
    const size_t o_mem = align16(o_zid + (size_t)n * st.s_zone_id);
    const size_t o_siz = align16(o_mem + (size_t)n * st.s_zone_members * sizeof(int));
    const size_t o_w = align16(o_siz + (size_t)n * st.s_zone_size * sizeof(int));
    const size_t o_pg = align16(o_w + (size_t)n * st.s_weights * sizeof(double));
    const size_t o_pz = o_pg + (size_t)n * st.s_p_global * sizeof(double);
    const size_t o_pf = o_pz + (size_t)n * st.s_p_zones * sizeof(double);
    const size_t o_src = align16(o_pf + (size_t)n * st.s_p_fam * sizeof(double));
    const size_t total = o_src + (size_t)n * st.s_source;
    Staging* sg;
    int r = staging_acquire(ctx, total, &sg);
    if (r != SB_OK) return r;
    unsigned char* b = sg->buf;

    // Large arrays in page-locked caller memory skip the staging copy: their device layout equals the caller's
    // (the source rows via a device scratch and a row-padding kernel), so the copy engine
    // reads them in place while the host threads below pack the zones and validate the source.  The call then waits
    // for these copies before it returns -- the host pointers are borrowed for the duration of the call only.
    cudaStream_t sm = ctx->stream;
    const size_t nn = (size_t)n;
    const bool dir_pg = host_pinned(p_global), dir_pz = Z > 0 && host_pinned(p_zones);
    const bool dir_pf = d.inheritance && host_pinned(p_families), dir_src = d.sample_source && host_pinned(source);
    if (dir_pg) CK(cudaMemcpyAsync(st.p_global + chain * st.s_p_global, p_global, nn * FS * sizeof(double), cudaMemcpyHostToDevice, sm));
    if (dir_pz) CK(cudaMemcpyAsync(st.p_zones + chain * st.s_p_zones, p_zones, nn * Z * FS * sizeof(double), cudaMemcpyHostToDevice, sm));
    if (dir_pf) CK(cudaMemcpyAsync(st.p_fam + chain * st.s_p_fam, p_families, nn * Fam * FS * sizeof(double), cudaMemcpyHostToDevice, sm));
    if (dir_src) {
        // one contiguous copy into a device scratch, then a kernel pads the rows from F to ldx bytes (a pitched copy
        // of 200-byte rows runs at a row per ~0.1 us on the copy engine: 13 ms for 64 chains at shape D)
        const size_t need = nn * N * F;
        if (ctx->src_scratch_bytes < need) {
            if (ctx->src_scratch) { CK(cudaStreamSynchronize(sm)); CK(cudaFree(ctx->src_scratch)); ctx->src_scratch = nullptr; }
            CK(cudaMalloc((void**)&ctx->src_scratch, need));
            ctx->src_scratch_bytes = need;
        }
        CK(cudaMemcpyAsync(ctx->src_scratch, source, need, cudaMemcpyHostToDevice, sm));
        const size_t units = nn * N * (size_t)(d.ldx >> 4);
        pad_rows_kernel<<<(unsigned)((units + 255) / 256), 256, 0, sm>>>(ctx->src_scratch, st.source + chain * st.s_source,
                                                                          nn * N, F, d.ldx);
        ++ctx->launches;
        CK(cudaGetLastError());
    }
    const bool any_direct = dir_pg || dir_pz || dir_pf || dir_src;
    if (any_direct) {
        if (!ctx->direct_done) CK(cudaEventCreateWithFlags(&ctx->direct_done, cudaEventDisableTiming));
        CK(cudaEventRecord(ctx->direct_done, sm));
    }

    // packing is host memory bandwidth: spread the chains over a few threads when there are many
    std::vector<int> h_sizes((size_t)n * st.s_zone_size, 0);
    auto pack = [&](int c) -> int {
        int8_t* zid = reinterpret_cast<int8_t*>(b + o_zid) + (size_t)c * st.s_zone_id;
        int* members = reinterpret_cast<int*>(b + o_mem) + (size_t)c * st.s_zone_members;
        int* sizes = reinterpret_cast<int*>(b + o_siz) + (size_t)c * st.s_zone_size;
        memset(zid, 0xff, st.s_zone_id);                       // -1
        sizes[0] = 0;
        for (int z = 0; z < Z; ++z) {
            int m = 0;
            const uint8_t* row = zones + ((size_t)c * Z + z) * N;
            int* mrow = members + (size_t)z * N;
            for (int i = 0; i < N; ++i)
                if (row[i]) {
                    if (zid[i] >= 0) return SB_E_OVERLAP;
                    zid[i] = (int8_t)z;
                    mrow[m++] = i;
                }
            for (int i = m; i < N; ++i) mrow[i] = 0;
            sizes[z] = m;
            h_sizes[(size_t)c * st.s_zone_size + z] = m;
        }
        double* w3 = reinterpret_cast<double*>(b + o_w) + (size_t)c * st.s_weights;
        const double* wc = weights + (size_t)c * F * C;
        for (int f = 0; f < F; ++f)
            for (int k = 0; k < 3; ++k) w3[(size_t)f * 3 + k] = k < C ? wc[(size_t)f * C + k] : 0.0;
        if (!dir_pg) memcpy(b + o_pg + (size_t)c * FS * sizeof(double), p_global + (size_t)c * FS, FS * sizeof(double));
        if (Z > 0 && !dir_pz)
            memcpy(b + o_pz + (size_t)c * Z * FS * sizeof(double), p_zones + (size_t)c * Z * FS, (size_t)Z * FS * sizeof(double));
        if (d.inheritance && !dir_pf)
            memcpy(b + o_pf + (size_t)c * Fam * FS * sizeof(double), p_families + (size_t)c * Fam * FS,
                   (size_t)Fam * FS * sizeof(double));
        if (d.sample_source) {
            unsigned char* hs = b + o_src + (size_t)c * st.s_source;
            unsigned bad = 0;
            for (int i = 0; i < N; ++i) {
                const uint8_t* srow = source + ((size_t)c * N + i) * F;
                if (!dir_src) {
                    unsigned char* drow = hs + (size_t)i * d.ldx;
                    memcpy(drow, srow, (size_t)F);
                    memset(drow + F, 0, (size_t)(d.ldx - F));
                }
                unsigned char mx = 0;
                for (int f = 0; f < F; ++f) mx = srow[f] > mx ? srow[f] : mx;      // vectorises (pmaxub)
                bad |= (unsigned)(mx >= C);
            }
            if (bad) return SB_E_ARG;
        }
        return SB_OK;
    };

    int status = SB_OK;
    const int n_thr = std::min<int>({n / 4, 8, (int)std::max(1u, std::thread::hardware_concurrency())});
    if (n_thr <= 1) {
        for (int c = 0; c < n && status == SB_OK; ++c) status = pack(c);
    } else {
        std::vector<int> rc(n_thr, SB_OK);
        std::vector<std::thread> pool;
        for (int t = 0; t < n_thr; ++t)
            pool.emplace_back([&, t]() {
                for (int c = t; c < n; c += n_thr) {
                    const int e = pack(c);
                    if (e != SB_OK) { rc[t] = e; return; }
                }
            });
        for (auto& th : pool) th.join();
        for (int e : rc) if (e != SB_OK) status = e;
    }
    if (status != SB_OK && any_direct) CK(cudaEventSynchronize(ctx->direct_done));       // borrowed pointers: drain first
    if (status == SB_E_OVERLAP) FAIL(SB_E_OVERLAP, "sb_set_state: zones overlap (a site belongs to two zones)");
    if (status != SB_OK) FAIL(SB_E_ARG, "sb_set_state: source component out of range");
    for (size_t i = 0; i < h_sizes.size(); ++i) ctx->h_zone_size[(size_t)chain * st.s_zone_size + i] = h_sizes[i];

    CK(cudaMemcpyAsync(st.zone_id + chain * st.s_zone_id, b + o_zid, nn * st.s_zone_id, cudaMemcpyHostToDevice, sm));
    if (Z > 0) {
        CK(cudaMemcpyAsync(st.zone_members + chain * st.s_zone_members, b + o_mem, nn * st.s_zone_members * sizeof(int),
                           cudaMemcpyHostToDevice, sm));
        CK(cudaMemcpyAsync(st.zone_size + chain * st.s_zone_size, b + o_siz, nn * st.s_zone_size * sizeof(int),
                           cudaMemcpyHostToDevice, sm));
    }
    CK(cudaMemcpyAsync(st.weights + chain * st.s_weights, b + o_w, nn * st.s_weights * sizeof(double), cudaMemcpyHostToDevice, sm));
    if (!dir_pg)
        CK(cudaMemcpyAsync(st.p_global + chain * st.s_p_global, b + o_pg, nn * FS * sizeof(double), cudaMemcpyHostToDevice, sm));
    if (Z > 0 && !dir_pz)
        CK(cudaMemcpyAsync(st.p_zones + chain * st.s_p_zones, b + o_pz, nn * Z * FS * sizeof(double), cudaMemcpyHostToDevice, sm));
    if (d.inheritance && !dir_pf)
        CK(cudaMemcpyAsync(st.p_fam + chain * st.s_p_fam, b + o_pf, nn * Fam * FS * sizeof(double), cudaMemcpyHostToDevice, sm));
    if (d.sample_source && !dir_src)
        CK(cudaMemcpyAsync(st.source + chain * st.s_source, b + o_src, nn * st.s_source, cudaMemcpyHostToDevice, sm));
    CK(cudaEventRecord(sg->done, sm));
    sg->pending = true;
    CK(cudaMemsetAsync(st.err + chain, 0, nn * sizeof(int), sm));          // a new state starts without error flags
    dim3 grid((F + 127) / 128, n);
    wnorm_kernel<<<grid, 128, 0, ctx->stream>>>(d, st, chain);
    ++ctx->launches;
    CK(cudaGetLastError());
    r = recount_chains(ctx, chain, n);
    if (any_direct) CK(cudaEventSynchronize(ctx->direct_done));     // the copy engine is done with the caller's memory
    return r;
}

int sb_set_state(sb_ctx* ctx, int chain, const uint8_t* zones, const double* weights, const double* p_global,
                 const double* p_zones, const double* p_families, const uint8_t* source) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, chain);
    if (r != SB_OK) return r;
    return set_states_impl(ctx, chain, 1, zones, weights, p_global, p_zones, p_families, source);
}

int sb_set_states(sb_ctx* ctx, int chain_begin, int n_chains, const uint8_t* zones, const double* weights,
                  const double* p_global, const double* p_zones, const double* p_families, const uint8_t* source) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (chain_begin < 0 || n_chains < 1 || chain_begin + n_chains > ctx->dims.n_chains) FAIL(SB_E_ARG, "chain range out of bounds");
    if (!ctx->has_data) FAIL(SB_E_STATE, "sb_upload_data has not been called");
    return set_states_impl(ctx, chain_begin, n_chains, zones, weights, p_global, p_zones, p_families, source);
}

int sb_get_state(sb_ctx* ctx, int chain, uint8_t* zones, double* weights, double* p_global, double* p_zones,
                 double* p_families, uint8_t* source) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, chain);
    if (r != SB_OK) return r;
    const DevData& d = ctx->d;
    ChainArrays& st = ctx->st;
    const int N = d.N, F = d.F, S = d.S, Z = d.Z, C = d.C, Fam = d.Fam;
    if (source && !d.sample_source) FAIL(SB_E_STATE, "context has no source state");
    // every requested array is copied into the pinned staging buffer asynchronously; ONE stream synchronisation, then
    // the host-side unpacking (zone ids -> indicator rows, [F][3] -> [F][C] weights, padded -> dense source rows)
    const size_t FS = (size_t)F * S;
    const bool want_z = zones && Z > 0, want_pz = p_zones && Z > 0, want_pf = p_families && d.inheritance;
    const size_t o_zid = 0;
    const size_t o_w = align16(o_zid + (want_z ? st.s_zone_id : 0));
    const size_t o_pg = align16(o_w + (weights ? (size_t)F * 3 * sizeof(double) : 0));
    const size_t o_pz = o_pg + (p_global ? FS * sizeof(double) : 0);
    const size_t o_pf = o_pz + (want_pz ? (size_t)Z * FS * sizeof(double) : 0);
    const size_t o_src = align16(o_pf + (want_pf ? (size_t)Fam * FS * sizeof(double) : 0));
    const size_t total = o_src + (source ? (size_t)N * d.ldx : 0);
    Staging* sg;
    r = staging_acquire(ctx, total > 0 ? total : 16, &sg);
    if (r != SB_OK) return r;
    unsigned char* b = sg->buf;
    cudaStream_t sm = ctx->stream;
#define GET(off, arr, bytes) CK(cudaMemcpyAsync(b + (off), (arr), (bytes), cudaMemcpyDeviceToHost, sm))
    if (want_z) GET(o_zid, st.zone_id + chain * st.s_zone_id, st.s_zone_id);
    if (weights) GET(o_w, st.weights + chain * st.s_weights, (size_t)F * 3 * sizeof(double));
    if (p_global) GET(o_pg, st.p_global + chain * st.s_p_global, FS * sizeof(double));
    if (want_pz) GET(o_pz, st.p_zones + chain * st.s_p_zones, (size_t)Z * FS * sizeof(double));
    if (want_pf) GET(o_pf, st.p_fam + chain * st.s_p_fam, (size_t)Fam * FS * sizeof(double));
    if (source) GET(o_src, st.source + chain * st.s_source, (size_t)N * d.ldx);
#undef GET
    CK(cudaStreamSynchronize(sm));
    if (want_z) {
        const int8_t* zid = reinterpret_cast<const int8_t*>(b + o_zid);
        memset(zones, 0, (size_t)Z * N);
        for (int n = 0; n < N; ++n)
            if (zid[n] >= 0) zones[(size_t)zid[n] * N + n] = 1;
    }
    if (weights) {
        const double* w3 = reinterpret_cast<const double*>(b + o_w);
        for (int f = 0; f < F; ++f)
            for (int c = 0; c < C; ++c) weights[(size_t)f * C + c] = w3[(size_t)f * 3 + c];
    }
    if (p_global) memcpy(p_global, b + o_pg, FS * sizeof(double));
    if (want_pz) memcpy(p_zones, b + o_pz, (size_t)Z * FS * sizeof(double));
    if (want_pf) memcpy(p_families, b + o_pf, (size_t)Fam * FS * sizeof(double));
    if (source)
        for (int n = 0; n < N; ++n) memcpy(source + (size_t)n * F, b + o_src + (size_t)n * d.ldx, F);
    return SB_OK;
}

int sb_copy_state(sb_ctx* ctx, int dst, int src) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, dst);
    if (r == SB_OK) r = check_chain(ctx, src);
    if (r != SB_OK) return r;
    if (dst == src) return SB_OK;
    ChainArrays& st = ctx->st;
#define CP(arr, stride, T) \
    CK(cudaMemcpyAsync(st.arr + dst * (stride), st.arr + src * (stride), (stride) * sizeof(T), cudaMemcpyDeviceToDevice, ctx->stream))
    CP(zone_id, st.s_zone_id, int8_t); CP(zone_members, st.s_zone_members, int); CP(zone_size, st.s_zone_size, int);
    CP(weights, st.s_weights, double); CP(wnorm, st.s_wnorm, double); CP(p_global, st.s_p_global, double);
    CP(p_zones, st.s_p_zones, double); CP(p_fam, st.s_p_fam, double); CP(source, st.s_source, uint8_t);
    CP(cnt_glob, st.s_cnt_glob, int); CP(cnt_zone, st.s_cnt_zone, int); CP(cnt_fam, st.s_cnt_fam, int);
    CP(cnt_src, st.s_cnt_src, int); CP(prior_pg, st.s_prior_pg, double); CP(prior_pf, st.s_prior_pf, double);
    CP(ll, (size_t)1, double); CP(prior, (size_t)1, double); CP(geo, (size_t)1, double);
#undef CP
    for (size_t z = 0; z < st.s_zone_size; ++z)
        ctx->h_zone_size[dst * st.s_zone_size + z] = ctx->h_zone_size[src * st.s_zone_size + z];
    return SB_OK;
}

// ---------------------------------------------------------------------------------------------
// stateless evaluation
// ---------------------------------------------------------------------------------------------
static int launch_likelihood(sb_ctx* ctx, int chain_begin, int n, int use_source) {
    const DevData& d = ctx->d;
    // table-driven K1: one launch (weights normalisation and the per-chain reduction are fused).  Its table rows hold
    // 32 codes, i.e. S + 1 <= 32 digits: S = 32 (SB_MAX_S) takes the tile-per-site kernel below
    // one CTA per chain (sb_lhtab.cuh:lh_chain_kernel): launches of many chains over few sites, where the table
    // kernel's per-chain prologue, repeated in each of its ~11 CTAs per chain, is most of the launch
    if (!use_source && d.S + 1 <= LT_ROW && ctx->lt.cpb >= 2 && n >= ctx->k1c_min_chains && d.N <= ctx->k1c_max_n &&
        n <= 65535 && blocked_kernels_allowed()) {
        const LhTab& t = ctx->lt;
        int fc = d.F;
        while (fc > 1 && lc_layout(d.F, d.S, d.Z, fc, t.n_classes, t.n_groups, false).total > 96 * 1024) fc = (fc + 1) / 2;
        const size_t smem = lc_layout(d.F, d.S, d.Z, fc, t.n_classes, t.n_groups, false).total;
        if (smem <= 200 * 1024) {
            auto kern = t.cpb == 2 ? lh_chain_kernel<2> : lh_chain_kernel<3>;
            if (smem > 48 * 1024) CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            kern<<<n, LC_THREADS, smem, ctx->stream>>>(d, ctx->st, t, chain_begin, fc);
            ++ctx->launches;
            CK(cudaGetLastError());
            return SB_OK;
        }
    }
    // blocked form (one CTA per 16-feature block and chain, sb_eval.cuh): shapes with one site per code byte (S + 1 > 5),
    // S = 32, and launches over few sites, where the table kernel's per-chain prologue dominates (SB_K1_BLOCK_MAX_N)
    const size_t fb_one = lh_free_block_smem(d.S, d.Z, d.Fam);
    if (!use_source && fb_one <= 200 * 1024 && n <= 65535 && blocked_kernels_allowed() &&
        (ctx->lt.cpb == 1 || d.S + 1 > LT_ROW || d.N <= ctx->k1_block_max_n)) {
        // CTAs per chain: enough for two CTAs per SM over the launch (a CTA's prologue and epilogue are dependent
        // memory round trips, so few fat CTAs win: 256 chains x 500 x 100 x 4 take 29 / 34 / 36 / 50 us with 1 / 2 /
        // 3 / 7 CTAs per chain), every CTA's tables within 48 KB (wide tables: one block per CTA).  SB_K1B_PARTS overrides.
        const int nb = d.ldx >> 4;
        int parts = std::min(nb, std::max(1, 2 * ctx->n_sms / n));
        while (parts < nb && lh_free_block_smem(d.S, d.Z, d.Fam, (nb + parts - 1) / parts) > 48 * 1024) ++parts;
        if (ctx->k1b_parts > 0) parts = std::min(nb, ctx->k1b_parts);
        const size_t fb_smem = lh_free_block_smem(d.S, d.Z, d.Fam, (nb + parts - 1) / parts);
        if (fb_smem > 200 * 1024) FAIL(SB_E_ARG, "SB_K1B_PARTS: the tables of a CTA exceed shared memory");
        if (fb_smem > 48 * 1024)
            CK(cudaFuncSetAttribute(lh_free_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_smem));
        lh_free_block_kernel<<<dim3(parts, n), ctx->k1b_threads, fb_smem, ctx->stream>>>(
            d, ctx->st, chain_begin, ctx->partial + (size_t)chain_begin * nb, ctx->lt_ticket + chain_begin);
        ++ctx->launches;
        CK(cudaGetLastError());
        return SB_OK;
    }
    if (!use_source && d.S + 1 <= LT_ROW) {
        LhTab t = ctx->lt;
        // grid = (group splits, feature chunks, chains).  Every CTA pays a fixed latency chain (tables, zone-member
        // correction, reduction); measured best on B200 (scripts/gpu_k1_sweep.sh) at about two waves of
        // 148 SMs x 7 resident CTAs, each walking all groups of its split for its feature chunk
        const int n_tiles = (t.n_groups + LT_THREADS - 1) / LT_THREADS;
        // compact rows (one site per code byte): every CTA rebuilds the rows of ALL family classes of its feature chunk,
        // so fewer, fatter chunks win (shape D: 17 chunks of 12 features, 77 us, against 34 chunks of 6 with 256-byte rows, 96 us)
        const long want = t.cpb == 1 ? std::min(ctx->lt_want, ctx->lt_want_compact) : ctx->lt_want;
        const int fc_min = std::min(d.F, 4);
        int chunks = (int)std::min<long>((d.F + fc_min - 1) / fc_min, std::max<long>(1, (want + n - 1) / n));
        t.n_splits = (int)std::min<long>(n_tiles, std::max<long>(1, (want + (long)chunks * n - 1) / ((long)chunks * n)));
        t.gsplit = (n_tiles + t.n_splits - 1) / t.n_splits * LT_THREADS;
        t.n_splits = (t.n_groups + t.gsplit - 1) / t.gsplit;
        t.max_cls = 1;
        for (int sp = 0; sp < t.n_splits; ++sp) {
            const int g0 = sp * t.gsplit, g1 = std::min(t.n_groups, g0 + t.gsplit);
            t.max_cls = std::max(t.max_cls, ctx->h_gclass[g1 - 1] - ctx->h_gclass[g0] + 1);
        }
        t.fc = std::max(1, (d.F + chunks - 1) / chunks);
        while (t.fc > 1 && lt_smem_bytes(t.fc, t.max_cls, d.S, t.cpb == 1) > 46 * 1024) --t.fc;
        // (hundreds of family classes in one split: even one feature's rows do not fit -> tile-per-site kernel below)
        const bool tables_fit = lt_smem_bytes(t.fc, t.max_cls, d.S, t.cpb == 1) <= 46 * 1024;
        if (tables_fit) {
        t.n_chunks = (d.F + t.fc - 1) / t.fc;
        t.smem = std::max(lt_smem_bytes(t.fc, t.max_cls, d.S, t.cpb == 1), (size_t)std::min(d.F, LT_CORR_FB) * 12 * sizeof(double));
        // blockIdx.z < n_corr: correction CTAs (zone members), scheduled first, concurrent with the streaming CTAs
        // (one is enough once the streaming pass dominates: 96.7 vs 99.1 us at N = 20000)
        const int n_corr = d.Z > 0 ? (d.N >= 16384 ? 1 : std::max(1, LT_CORR_CTAS / t.n_splits)) : 0;
        if (n > 65535 || t.n_chunks + n_corr > 65535) FAIL(SB_E_ARG, "K1 grid exceeds the launch limits (chains or feature chunks > 65535)");
        dim3 grid(t.n_splits, n, t.n_chunks + n_corr);
        double* part = ctx->lt_partial + (size_t)chain_begin * n_tiles * (d.F + LT_CORR_CTAS);
        unsigned* tick = ctx->lt_ticket + chain_begin;
        if (t.cpb == 1)
            lh_table_kernel<1><<<grid, LT_THREADS, t.smem, ctx->stream>>>(d, ctx->st, t, chain_begin, part, tick);
        else if (t.cpb == 2)
            lh_table_kernel<2><<<grid, LT_THREADS, t.smem, ctx->stream>>>(d, ctx->st, t, chain_begin, part, tick);
        else
            lh_table_kernel<3><<<grid, LT_THREADS, t.smem, ctx->stream>>>(d, ctx->st, t, chain_begin, part, tick);
        ++ctx->launches;
        CK(cudaGetLastError());
        return SB_OK;
        }
    }
    const size_t lb_smem = lh_source_block_smem(d.S, d.Z, d.Fam);
    if (use_source && lb_smem <= 200 * 1024 && n <= 65535 && blocked_kernels_allowed()) {
        // blocked form, ONE launch: parameter tables of a 16-feature block staged in shared memory (w~ rows included),
        // the chain's last CTA reduces the per-block partial sums
        const int nb = d.ldx >> 4;
        if (lb_smem > 48 * 1024)
            CK(cudaFuncSetAttribute(lh_source_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lb_smem));
        lh_source_block_kernel<<<dim3(nb, n), ctx->k1s_threads, lb_smem, ctx->stream>>>(
            d, ctx->st, chain_begin, ctx->partial + (size_t)chain_begin * nb, nb, ctx->lt_ticket + chain_begin);
        ++ctx->launches;
        CK(cudaGetLastError());
        return SB_OK;
    }
    dim3 gw((d.F + 127) / 128, n);
    wnorm_kernel<<<gw, 128, 0, ctx->stream>>>(d, ctx->st, chain_begin);
    // fallbacks (with source: tables of a feature block exceed shared memory; source-free: S = 32 or too many family
    // classes for K1's tables): one CTA per (site tile, chain), global gathers
    dim3 grid(ctx->n_tiles, n);
    if (use_source)
        lh_full_kernel<1><<<grid, LH_THREADS, 0, ctx->stream>>>(d, ctx->st, chain_begin, ctx->sites_per_tile, ctx->partial,
                                                               ctx->n_tiles);
    else
        lh_full_kernel<0><<<grid, LH_THREADS, 0, ctx->stream>>>(d, ctx->st, chain_begin, ctx->sites_per_tile, ctx->partial,
                                                               ctx->n_tiles);
    lh_reduce_kernel<<<n, 32, 0, ctx->stream>>>(ctx->partial, ctx->n_tiles, ctx->st.ll, chain_begin);
    ctx->launches += 3;
    CK(cudaGetLastError());
    return SB_OK;
}

static int check_range(sb_ctx* ctx, int chain_begin, int n, int use_source) {
    if (!ctx->has_data) FAIL(SB_E_STATE, "sb_upload_data has not been called");
    if (chain_begin < 0 || n < 1 || chain_begin + n > ctx->dims.n_chains) FAIL(SB_E_ARG, "chain range out of bounds");
    if (use_source && !ctx->d.sample_source) FAIL(SB_E_STATE, "use_source on a context created without SB_F_SAMPLE_SOURCE");
    return SB_OK;
}

int sb_eval_likelihood_async(sb_ctx* ctx, int chain_begin, int n, int use_source) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_range(ctx, chain_begin, n, use_source);
    if (r != SB_OK) return r;
    return launch_likelihood(ctx, chain_begin, n, use_source);
}

int sb_read_ll(sb_ctx* ctx, int chain_begin, int n, double* ll) {
    if (!ctx || !ll) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_range(ctx, chain_begin, n, 0);
    if (r != SB_OK) return r;
    D2H(ll, ctx->st.ll + chain_begin, n);
    return SB_OK;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free likelihood evaluation (no counterpart in the reference; see include/sbayes_b200.h).
// The scratch of the table-driven K1 -- the per-CTA partial sums and the per-chain arrival tickets -- lives in ONE
// file-scope device array instead of the context's cudaMalloc'ed buffers.  Every GPU that loads the library holds
// the array (SB_NA_SCRATCH_BYTES = 528 KiB) until the process exits; every call on a device shares it.
//   [0, SB_NA_MAX_PARTIALS * 8)            double  partial[chains][parts]
//   [SB_NA_MAX_PARTIALS * 8, ... + 4 * SB_NA_MAX_CHAINS)   unsigned ticket[chains]   (must start at zero)
// In host code `sb_na_scratch` is only a stand-in for the symbol: its device address is resolved once per device, with
// that device current, through cudaGetSymbolAddress (na_scratch_address) and handed to the kernel as a pointer.
// ---------------------------------------------------------------------------------------------
#define SB_NA_MAX_PARTIALS 65536
#define SB_NA_MAX_CHAINS 4096
#define SB_NA_MAX_DEVICES 64
#define SB_NA_SCRATCH_BYTES (SB_NA_MAX_PARTIALS * 8 + SB_NA_MAX_CHAINS * 4)
static __device__ __align__(256) unsigned char sb_na_scratch[SB_NA_SCRATCH_BYTES];
static std::mutex g_na_mutex;
static unsigned char* g_na_address[SB_NA_MAX_DEVICES] = {};

// device address of sb_na_scratch on ctx->device (which the caller has made current)
static int na_scratch_address(sb_ctx* ctx, unsigned char** out) {
    if (ctx->device < 0 || ctx->device >= SB_NA_MAX_DEVICES) FAIL(SB_E_ARG, "sb_eval_likelihood_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_address[ctx->device]) {
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, sb_na_scratch));
        if (!p) FAIL(SB_E_CUDA, "cudaGetSymbolAddress(sb_na_scratch) returned a null address");
        g_na_address[ctx->device] = static_cast<unsigned char*>(p);
    }
    *out = g_na_address[ctx->device];
    return SB_OK;
}

int sb_eval_likelihood_noalloc(sb_ctx* ctx, int chain_begin, int n, double* ll_out) {
    if (!ctx || !ll_out) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_range(ctx, chain_begin, n, 0);
    if (r != SB_OK) return r;
    const DevData& d = ctx->d;
    if (d.S + 1 > LT_ROW) FAIL(SB_E_ARG, "sb_eval_likelihood_noalloc: n_states = 32 is not served by the table-driven kernel");
    if (n > SB_NA_MAX_CHAINS) FAIL(SB_E_ARG, "sb_eval_likelihood_noalloc: more than 4096 chains per call (static scratch limit)");
    // launch geometry of the table-driven K1, as in launch_likelihood
    LhTab t = ctx->lt;
    const int n_tiles = (t.n_groups + LT_THREADS - 1) / LT_THREADS;
    const long want = t.cpb == 1 ? std::min(ctx->lt_want, ctx->lt_want_compact) : ctx->lt_want;
    const int fc_min = std::min(d.F, 4);
    const int chunks = (int)std::min<long>((d.F + fc_min - 1) / fc_min, std::max<long>(1, (want + n - 1) / n));
    t.n_splits = (int)std::min<long>(n_tiles, std::max<long>(1, (want + (long)chunks * n - 1) / ((long)chunks * n)));
    t.gsplit = (n_tiles + t.n_splits - 1) / t.n_splits * LT_THREADS;
    t.n_splits = (t.n_groups + t.gsplit - 1) / t.gsplit;
    t.max_cls = 1;
    for (int sp = 0; sp < t.n_splits; ++sp) {
        const int g0 = sp * t.gsplit, g1 = std::min(t.n_groups, g0 + t.gsplit);
        t.max_cls = std::max(t.max_cls, ctx->h_gclass[g1 - 1] - ctx->h_gclass[g0] + 1);
    }
    t.fc = std::max(1, (d.F + chunks - 1) / chunks);
    while (t.fc > 1 && lt_smem_bytes(t.fc, t.max_cls, d.S, t.cpb == 1) > 46 * 1024) --t.fc;
    if (lt_smem_bytes(t.fc, t.max_cls, d.S, t.cpb == 1) > 46 * 1024)
        FAIL(SB_E_ARG, "sb_eval_likelihood_noalloc: the family-class tables of one feature exceed shared memory");
    t.n_chunks = (d.F + t.fc - 1) / t.fc;
    t.smem = std::max(lt_smem_bytes(t.fc, t.max_cls, d.S, t.cpb == 1), (size_t)std::min(d.F, LT_CORR_FB) * 12 * sizeof(double));
    const int n_corr = d.Z > 0 ? (d.N >= 16384 ? 1 : std::max(1, LT_CORR_CTAS / t.n_splits)) : 0;
    if (t.n_chunks + n_corr > 65535) FAIL(SB_E_ARG, "sb_eval_likelihood_noalloc: more than 65535 feature chunks");
    const size_t parts = (size_t)t.n_splits * (t.n_chunks + n_corr);
    if ((size_t)n * parts > SB_NA_MAX_PARTIALS) {
        ctx->err = "sb_eval_likelihood_noalloc: " + std::to_string((size_t)n * parts) +
                   " partial sums exceed the static scratch (65536); evaluate fewer chains per call";
        return SB_E_ARG;
    }
    unsigned char* scratch = nullptr;
    r = na_scratch_address(ctx, &scratch);
    if (r != SB_OK) return r;
    double* part = reinterpret_cast<double*>(scratch);
    unsigned* tick = reinterpret_cast<unsigned*>(scratch + (size_t)SB_NA_MAX_PARTIALS * 8);
    // the array keeps the previous call's contents: the tickets must start at zero (the partials are all rewritten)
    CK(cudaMemsetAsync(tick, 0, (size_t)n * sizeof(unsigned), ctx->stream));
    dim3 grid(t.n_splits, n, t.n_chunks + n_corr);
    if (t.cpb == 1)
        lh_table_kernel<1><<<grid, LT_THREADS, t.smem, ctx->stream>>>(d, ctx->st, t, chain_begin, part, tick);
    else if (t.cpb == 2)
        lh_table_kernel<2><<<grid, LT_THREADS, t.smem, ctx->stream>>>(d, ctx->st, t, chain_begin, part, tick);
    else
        lh_table_kernel<3><<<grid, LT_THREADS, t.smem, ctx->stream>>>(d, ctx->st, t, chain_begin, part, tick);
    ++ctx->launches;
    CK(cudaGetLastError());
    // result: device memory is written stream-ordered (no synchronisation), host memory is filled before returning
    cudaPointerAttributes attr;
    bool on_device = false;
    if (cudaPointerGetAttributes(&attr, ll_out) == cudaSuccess) on_device = attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
    else cudaGetLastError();
    CK(cudaMemcpyAsync(ll_out, ctx->st.ll + chain_begin, (size_t)n * sizeof(double),
                       on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, ctx->stream));
    if (!on_device) CK(cudaStreamSynchronize(ctx->stream));
    return SB_OK;
}

int sb_eval(sb_ctx* ctx, int chain_begin, int n, int use_source, double* ll, double* prior, double* geo) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_range(ctx, chain_begin, n, use_source);
    if (r != SB_OK) return r;
    if (!ctx->has_priors) FAIL(SB_E_STATE, "sb_set_priors has not been called");
    const DevData& d = ctx->d;
    if (d.geo_type == SB_GEO_COST_BASED && d.Z > 0 && d.N > ctx->geo_cap) {
        // zone sizes as they are on the DEVICE (sb_run moves zones; the host mirror only follows sb_set_state)
        std::vector<int> zs((size_t)n * ctx->st.s_zone_size);
        D2H(zs.data(), ctx->st.zone_size + chain_begin * ctx->st.s_zone_size, zs.size());
        for (int c = 0; c < n; ++c)
            if (zs[c * ctx->st.s_zone_size + d.Z - 1] > ctx->geo_cap)
                FAIL(SB_E_ARG, "geo prior: last zone larger than the Prim scratch (4096 members)");
    }
    r = launch_likelihood(ctx, chain_begin, n, use_source);
    if (r != SB_OK) return r;
    const size_t smem = d.geo_type == SB_GEO_COST_BASED ? (size_t)ctx->geo_cap * 9 + 16 : 16;
    if (smem > 48 * 1024)
        CK(cudaFuncSetAttribute(prior_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    prior_kernel<<<n, 256, smem, ctx->stream>>>(d, ctx->st, chain_begin, ctx->geo_cap);
    ++ctx->launches;
    CK(cudaGetLastError());
    if (ll) D2H(ll, ctx->st.ll + chain_begin, n);
    if (prior) D2H(prior, ctx->st.prior + chain_begin, n);
    if (geo) D2H(geo, ctx->st.geo + chain_begin, n);
    CK(cudaStreamSynchronize(ctx->stream));
    return SB_OK;
}

static int run_components(sb_ctx* ctx, int chain, double* all_lh, double* wnorm, double* post) {
    const DevData& d = ctx->d;
    const size_t cells = (size_t)d.N * d.F, n = cells * d.C;
    double *dl = nullptr, *dw = nullptr, *dp = nullptr;
    if (all_lh) CK(cudaMalloc((void**)&dl, n * sizeof(double)));
    if (wnorm) CK(cudaMalloc((void**)&dw, n * sizeof(double)));
    if (post) CK(cudaMalloc((void**)&dp, n * sizeof(double)));
    dim3 gw((d.F + 127) / 128, 1);
    wnorm_kernel<<<gw, 128, 0, ctx->stream>>>(d, ctx->st, chain);
    components_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, ctx->stream>>>(d, ctx->st, chain, dl, dw, dp);
    ctx->launches += 2;
    int r = SB_OK;
    if (cudaGetLastError() != cudaSuccess) { ctx->err = "components_kernel launch failed"; r = SB_E_CUDA; }
    if (r == SB_OK && all_lh) r = d2h(ctx, all_lh, dl, n);
    if (r == SB_OK && wnorm) r = d2h(ctx, wnorm, dw, n);
    if (r == SB_OK && post) r = d2h(ctx, post, dp, n);
    cudaStreamSynchronize(ctx->stream);
    if (dl) cudaFree(dl);
    if (dw) cudaFree(dw);
    if (dp) cudaFree(dp);
    return r;
}

int sb_eval_components(sb_ctx* ctx, int chain, double* all_lh, double* wnorm) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, chain);
    if (r != SB_OK) return r;
    return run_components(ctx, chain, all_lh, wnorm, nullptr);
}

int sb_source_posterior(sb_ctx* ctx, int chain, double* post) {
    if (!ctx || !post) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, chain);
    if (r != SB_OK) return r;
    return run_components(ctx, chain, nullptr, nullptr, post);
}

int sb_recount_async(sb_ctx* ctx, int chain_begin, int n) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_range(ctx, chain_begin, n, 1);
    if (r != SB_OK) return r;
    return recount_chains(ctx, chain_begin, n);
}

int sb_device_scalars(sb_ctx* ctx, void** ll, void** prior) {
    if (!ctx) return SB_E_ARG;
    if (ll) *ll = ctx->st.ll;
    if (prior) *prior = ctx->st.prior;
    return SB_OK;
}

int sb_counts(sb_ctx* ctx, int chain, int recount, int32_t* cnt_glob, int32_t* cnt_zone, int32_t* cnt_fam,
              int32_t* cnt_src_zone, int32_t* cnt_src_fam) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, chain);
    if (r != SB_OK) return r;
    const DevData& d = ctx->d;
    ChainArrays& st = ctx->st;
    if (!d.sample_source) FAIL(SB_E_STATE, "count tables need a SAMPLE_SOURCE context");
    if (recount) {
        r = recount_chain(ctx, chain);
        if (r != SB_OK) return r;
    }
    if (cnt_glob) D2H(cnt_glob, st.cnt_glob + chain * st.s_cnt_glob, st.s_cnt_glob);
    if (cnt_zone) D2H(cnt_zone, st.cnt_zone + chain * st.s_cnt_zone, st.s_cnt_zone);
    if (cnt_fam && d.inheritance) D2H(cnt_fam, st.cnt_fam + chain * st.s_cnt_fam, st.s_cnt_fam);
    if (cnt_src_zone || cnt_src_fam) {
        std::vector<int> cs(st.s_cnt_src);
        D2H(cs.data(), st.cnt_src + chain * st.s_cnt_src, cs.size());
        for (int f = 0; f < d.F; ++f)
            for (int c = 0; c < d.C; ++c) {
                const int* row = cs.data() + (size_t)f * 12;
                if (cnt_src_zone) cnt_src_zone[(size_t)f * d.C + c] = row[1 * 3 + c] + row[3 * 3 + c];
                if (cnt_src_fam) cnt_src_fam[(size_t)f * d.C + c] = row[2 * 3 + c] + row[3 * 3 + c];
            }
    }
    return SB_OK;
}

// ---------------------------------------------------------------------------------------------
// sampler
// ---------------------------------------------------------------------------------------------
static size_t step_smem_base(const StepParams& sp) {
    const size_t words = 2 * (size_t)sp.nwords + (sp.nwords & 1);
    return (words * sizeof(unsigned) + (size_t)sp.prim_cap * 9 + 16 + 15) & ~(size_t)15;
}
// Shared work area of the step kernel (one CTA per SM, so up to ~200 KB): the count tables of one 16-feature block
// of the source sweep; if they do not fit, the sweep falls back to global atomics.
static size_t step_smem_bytes(const DevData& d, StepParams& sp, int zone_cap, int adj_nnz) {
    const size_t base = step_smem_base(sp);
    const size_t counts = ((size_t)16 * (d.S + 1) * (1 + d.Z + d.Fam) + 192) * sizeof(int);   // S+1: staged variant
    const size_t room = 160 * 1024 > base ? 160 * 1024 - base : 0;
    sp.work_off = (int)base;
    sp.work_counts = d.sample_source && counts <= room;
    // SBAYES_B200_SWEEP caps the variant of the source sweep (tests): 0 global atomics, 1 shared counters, 2 staged
    int cap = 2;
    if (const char* env = getenv("SBAYES_B200_SWEEP")) cap = atoi(env);
    if (cap < 1) sp.work_counts = 0;
    size_t end = base;
    if (sp.work_counts) {
        end = base + counts;
        // the sweep's parameter tables (w~ [16][12], p_* [1+Z+Fam][16][S+1], f64) behind the counters, if they fit too
        const size_t par_off = (base + counts + 7) & ~(size_t)7;
        const size_t params = ((size_t)192 + (size_t)16 * (d.S + 1) * (1 + d.Z + d.Fam)) * sizeof(double);
        sp.par_off = (int)par_off;
        if (cap >= 2 && par_off + params <= base + room) {
            sp.work_counts = 2;
            end = par_off + params;
        }
    }
    // Proposal phase of the zone moves on chip: this chain's zone sizes, member lists and zone ids, then the adjacency
    // CSR ("adjacency rows staged in smem"), each only if it still fits next to the sweep's tables.
    // SBAYES_B200_ZONE_CACHE=0 disables both (tests).
    const size_t limit = 200 * 1024;
    bool allow = true;
    if (const char* env = getenv("SBAYES_B200_ZONE_CACHE")) allow = atoi(env) != 0;
    sp.zc_off = -1; sp.adj_off = -1; sp.zc_cap = zone_cap;
    if (allow && d.Z > 0) {
        const size_t zc_off = (end + 15) & ~(size_t)15;
        const size_t zc = (((size_t)d.Z * (1 + zone_cap)) * sizeof(int) + (size_t)d.N + 15) & ~(size_t)15;
        if (zc_off + zc <= limit) {
            sp.zc_off = (int)zc_off;
            end = zc_off + zc;
            const size_t adj = ((size_t)d.N + 1 + adj_nnz) * sizeof(int);
            if (d.adj_ptr && end + adj <= limit) {
                sp.adj_off = (int)end;
                end += (adj + 15) & ~(size_t)15;
            }
        }
    }
    return end;
}

static int ensure_step_scratch(sb_ctx* ctx) {
    StepParams& sp = ctx->sp;
    const DevData& d = ctx->d;
    if (!sp.tmp_members) {
        sp.s_tmp = (size_t)d.N;
        DA(sp.tmp_members, (size_t)ctx->dims.n_chains * sp.s_tmp);
    }
    sp.nwords = (d.N + 31) / 32;
    sp.chain_offset = ctx->dims.chain_offset;
    return SB_OK;
}

// Shared-memory layout and team (thread-block cluster) size of the step kernel for zones of up to `cap` members.
static int configure_step(sb_ctx* ctx, int cap) {
    const DevData& d = ctx->d;
    StepParams& sp = ctx->sp;
    const int nc = ctx->dims.n_chains;
    int r = ensure_step_scratch(ctx);
    if (r != SB_OK) return r;
    sp.s_shift = 0;
    while ((1 << sp.s_shift) < d.S) ++sp.s_shift;         // S <= SB_MAX_S = 32
    sp.prim_cap = d.geo_type == SB_GEO_COST_BASED ? cap : 0;
    ctx->step_smem = step_smem_bytes(d, sp, cap, ctx->adj_nnz);
    if (getenv("SBAYES_B200_DEBUG"))
        fprintf(stderr, "[sbayes_b200] step kernel: smem %zu B (work_off %d counts %d par_off %d zc_off %d zc_cap %d adj_off %d), nwords %d prim_cap %d\n",
                ctx->step_smem, sp.work_off, sp.work_counts, sp.par_off, sp.zc_off, sp.zc_cap, sp.adj_off, sp.nwords, sp.prim_cap);
    if (ctx->step_smem > 200 * 1024) FAIL(SB_E_ARG, "step kernel scratch exceeds shared memory (N or MAX_M too large)");
    // always opt in: the 48 KB default limit applies to static + dynamic shared memory together, and the kernel's
    // static part (look-ahead tables, reduction scratch, flog table) is ~14 KB
    CK(cudaFuncSetAttribute(mcmc_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->step_smem));
    // team (thread-block cluster) size: as many CTAs per chain as the SM count allows, each with at least one
    // 16-feature block of work; SBAYES_B200_TEAM overrides (1, 2, 4, 8 or 16 -- 16 is a non-portable cluster size)
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
    int team = 1;
    for (int t = 8; t >= 2; t >>= 1)
        if (nc * t <= sms && d.F >= 16 * t) { team = t; break; }
    if (const char* env = getenv("SBAYES_B200_TEAM")) {
        const int t = atoi(env);
        if (t == 1 || t == 2 || t == 4 || t == 8 || t == 16) team = t;
    }
    if (d.sample_source && !sp.work_counts) team = 1;     // the global-atomics source sweep is single-CTA
    if (team > 8) CK(cudaFuncSetAttribute(mcmc_step_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    // every chain's cluster must be resident at once: clusters live inside one GPC (~18 SMs on B200), so e.g. 16 teams of
    // 8 do NOT fit although 128 <= 148 SMs -- a second wave doubles the launch time (measured: 16 chains, team 8 -> 3.2 us
    // per step against 1.8 us for team 4).  An explicit SBAYES_B200_TEAM keeps its value as long as ONE cluster fits.
    const int need_clusters = getenv("SBAYES_B200_TEAM") ? 1 : nc;
    while (team > 1) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(nc * team), 1, 1);
        cfg.blockDim = dim3(STEP_THREADS, 1, 1);
        cfg.dynamicSmemBytes = ctx->step_smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)team; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        int n_clusters = 0;
        if (cudaOccupancyMaxActiveClusters(&n_clusters, mcmc_step_kernel, &cfg) == cudaSuccess &&
            n_clusters >= std::min(need_clusters, sms / team)) break;
        cudaGetLastError();
        team >>= 1;
    }
    sp.team = team;
    ctx->has_layout = true;
    return SB_OK;
}

// one launch of the step kernel: a thread-block cluster (team) of sp.team CTAs per chain
static int launch_steps(sb_ctx* ctx, const StepParams& sp_in, long long n_steps) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(ctx->dims.n_chains * sp_in.team), 1, 1);
    cfg.blockDim = dim3(STEP_THREADS, 1, 1);
    cfg.dynamicSmemBytes = ctx->step_smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)sp_in.team;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // a state uploaded after sb_set_operators may hold a zone that outgrows the member capacity of the shared
    // zone cache: such a launch reads the global lists instead
    StepParams sp = sp_in;
    if (sp.zc_off >= 0 && sp.zc_cap < ctx->d.N)
        for (int v : ctx->h_zone_size)
            if (v + 1 > sp.zc_cap) { sp.zc_off = -1; break; }
    CK(cudaLaunchKernelEx(&cfg, mcmc_step_kernel, ctx->d, ctx->st, sp, n_steps));
    ++ctx->launches;
    CK(cudaGetLastError());
    return SB_OK;
}

int sb_set_operators(sb_ctx* ctx, const sb_operator_cfg* cfg, const double* p_grow, const int32_t* max_size) {
    if (!ctx || !cfg || !p_grow || !max_size) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (!ctx->has_data || !ctx->has_priors) FAIL(SB_E_STATE, "sb_upload_data / sb_set_priors must precede sb_set_operators");
    const DevData& d = ctx->d;
    if (cfg->n_ops < 1 || cfg->n_ops > SB_N_OPS) FAIL(SB_E_ARG, "n_ops out of range");
    StepParams& sp = ctx->sp;
    double total = 0.0;
    for (int i = 0; i < cfg->n_ops; ++i) {
        const int op = cfg->op_ids[i];
        if (op < 0 || op >= SB_N_OPS) FAIL(SB_E_ARG, "unknown operator id");
        if (!(cfg->op_prob[i] >= 0.0)) FAIL(SB_E_ARG, "negative operator probability");
        if (cfg->op_prob[i] > 0.0) {
            const bool gibbs = op >= SB_OP_GIBBS_SOURCES && op <= SB_OP_GIBBS_P_FAMILIES;
            const bool alter = op >= SB_OP_ALTER_WEIGHTS;
            if (gibbs && !d.sample_source) FAIL(SB_E_ARG, "gibbs_* operators need SAMPLE_SOURCE");
            if (alter && d.sample_source) FAIL(SB_E_ARG, "alter_* operators are not defined with SAMPLE_SOURCE (mcmc_setup.py:80-94)");
            if ((op == SB_OP_GIBBS_P_FAMILIES || op == SB_OP_ALTER_P_FAMILIES) && !d.inheritance)
                FAIL(SB_E_ARG, "family operator without inheritance");
            if ((op <= SB_OP_SWAP_ZONE || op == SB_OP_GIBBS_P_ZONES || op == SB_OP_ALTER_P_ZONES) && d.Z < 1)
                FAIL(SB_E_ARG, "zone operator with zero zones");
            if (op <= SB_OP_SWAP_ZONE && !d.adj_ptr) FAIL(SB_E_ARG, "zone moves need the adjacency CSR");
        }
        total += cfg->op_prob[i];
    }
    if (!(total > 0.0)) FAIL(SB_E_ARG, "operator probabilities sum to zero");
    // keep operators with positive probability only; CDF normalised like np.random.choice(p=...)
    int m = 0;
    double run = 0.0;
    for (int i = 0; i < cfg->n_ops; ++i) {
        if (cfg->op_prob[i] <= 0.0) continue;
        run += cfg->op_prob[i] / total;
        sp.op_ids[m] = cfg->op_ids[i];
        sp.op_cdf[m] = run;
        ++m;
    }
    sp.n_ops = m;
    sp.prec_w = cfg->precision_weights; sp.prec_g = cfg->precision_p_global;
    sp.prec_z = cfg->precision_p_zones; sp.prec_f = cfg->precision_p_families;
    sp.min_size = cfg->min_size;
    sp.frac = cfg->gibbs_feature_fraction;
    if ((uint64_t)d.F * d.S >= ((uint64_t)1 << 27)) FAIL(SB_E_ARG, "F*S too large for the step kernel");
    sp.seed_lo = (uint32_t)cfg->seed; sp.seed_hi = (uint32_t)(cfg->seed >> 32);
    int r = ensure_step_scratch(ctx);
    if (r != SB_OK) return r;
    const int nc = ctx->dims.n_chains;
    if (!ctx->d_pgrow) { DA(ctx->d_pgrow, nc); DA(ctx->d_maxsize, nc); }
    int cap = 2;
    for (int c = 0; c < nc; ++c) {
        if (!(p_grow[c] >= 0.0 && p_grow[c] <= 1.0)) FAIL(SB_E_ARG, "p_grow_connected outside [0,1]");
        if (max_size[c] < 1) FAIL(SB_E_ARG, "max_size < 1");
        if (max_size[c] + 1 > cap) cap = max_size[c] + 1;
    }
    for (int v : ctx->h_zone_size) if (v + 1 > cap) cap = v + 1;
    if (cap > d.N) cap = d.N;
    H2D(ctx->d_pgrow, p_grow, nc);
    H2D(ctx->d_maxsize, max_size, nc);
    sp.p_grow = ctx->d_pgrow; sp.max_size = ctx->d_maxsize;
    r = configure_step(ctx, cap);
    if (r != SB_OK) return r;
    ctx->has_ops = true;
    return SB_OK;
}

int sb_run(sb_ctx* ctx, int64_t n_steps) {
    if (!ctx || n_steps < 0) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (!ctx->has_ops) FAIL(SB_E_STATE, "sb_set_operators has not been called");
    if (n_steps == 0) return SB_OK;
    StepParams sp = ctx->sp;
    sp.forced = nullptr; sp.n_events = 0; sp.events = nullptr;
    int r = launch_steps(ctx, sp, (long long)n_steps);
    if (r != SB_OK) return r;
    ctx->steps_done += n_steps;
    return SB_OK;
}

int sb_set_sample_from_prior(sb_ctx* ctx, int on) {
    if (!ctx) return SB_E_ARG;
    ctx->sp.from_prior = on ? 1 : 0;
    return SB_OK;
}

static size_t ring_record_bytes(const ChainArrays& st) {
    return ring_align(st.s_zone_id) + ring_align(st.s_weights * 8) + ring_align(st.s_p_global * 8) +
           ring_align(st.s_p_zones * 8) + ring_align(st.s_p_fam * 8);
}

int sb_run_events(sb_ctx* ctx, int64_t n_steps, int n_events, const int32_t* event_steps, const int32_t* event_slots,
                  int ring_chain) {
    if (!ctx || n_steps < 0 || n_events < 0 || (n_events > 0 && (!event_steps || !event_slots))) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (!ctx->has_ops) FAIL(SB_E_STATE, "sb_set_operators has not been called");
    if (ring_chain < -1 || ring_chain >= ctx->dims.n_chains) FAIL(SB_E_ARG, "ring_chain out of range");
    if (n_steps == 0) return SB_OK;
    const size_t nc = (size_t)ctx->dims.n_chains;
    int n_slots = 0;
    std::vector<int2> ev((size_t)n_events);
    for (int k = 0; k < n_events; ++k) {
        if (event_steps[k] < 1 || event_steps[k] > n_steps || (k > 0 && event_steps[k] <= event_steps[k - 1]))
            FAIL(SB_E_ARG, "event steps must be ascending in [1, n_steps]");
        if (event_slots[k] < -1) FAIL(SB_E_ARG, "event slot < -1");
        ev[k] = make_int2(event_steps[k], event_slots[k]);
        n_slots = std::max(n_slots, event_slots[k] + 1);
    }
    StepParams sp = ctx->sp;
    sp.forced = nullptr;
    if (n_events > 0) {
        // device buffers grow on demand (the stream is drained first: an earlier launch may still use the old ones)
        if (n_events > ctx->events_cap) {
            CK(cudaStreamSynchronize(ctx->stream));
            if (ctx->d_events) CK(cudaFree(ctx->d_events));
            ctx->d_events = nullptr;
            CK(cudaMalloc((void**)&ctx->d_events, (size_t)n_events * sizeof(int2)));
            ctx->events_cap = n_events;
        }
        if ((size_t)n_events * nc > ctx->ev_scalars_cap) {
            CK(cudaStreamSynchronize(ctx->stream));
            if (ctx->d_ev_scalars) CK(cudaFree(ctx->d_ev_scalars));
            ctx->d_ev_scalars = nullptr;
            CK(cudaMalloc((void**)&ctx->d_ev_scalars, (size_t)n_events * nc * sizeof(double2)));
            ctx->ev_scalars_cap = (size_t)n_events * nc;
        }
        ctx->ring_rec = ring_record_bytes(ctx->st);
        const size_t need = (size_t)n_slots * nc * ctx->ring_rec;
        if (need > ctx->ring_cap) {
            // (the slots of an earlier launch may not have been read yet: the ring keeps its content when it grows)
            CK(cudaStreamSynchronize(ctx->stream));
            if (ctx->copy_stream) CK(cudaStreamSynchronize(ctx->copy_stream));
            unsigned char* grown = nullptr;
            CK(cudaMalloc((void**)&grown, need));
            if (ctx->d_ring) {
                CK(cudaMemcpy(grown, ctx->d_ring, ctx->ring_cap, cudaMemcpyDeviceToDevice));
                CK(cudaFree(ctx->d_ring));
            }
            ctx->d_ring = grown;
            ctx->ring_cap = need;
        }
        // the events travel through the pinned staging buffer (the call returns without a synchronisation)
        Staging* sg;
        int r = staging_acquire(ctx, std::max<size_t>(ctx->stage_bytes, (size_t)n_events * sizeof(int2)), &sg);
        if (r != SB_OK) return r;
        memcpy(sg->buf, ev.data(), (size_t)n_events * sizeof(int2));
        CK(cudaMemcpyAsync(ctx->d_events, sg->buf, (size_t)n_events * sizeof(int2), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaEventRecord(sg->done, ctx->stream));
        sg->pending = true;
    }
    sp.events = ctx->d_events; sp.n_events = n_events; sp.ev_scalars = ctx->d_ev_scalars;
    sp.ring = ctx->d_ring; sp.ring_rec = ctx->ring_rec; sp.ring_chain = ring_chain;
    ctx->ring_slots = std::max(ctx->ring_slots, n_slots);
    int r = launch_steps(ctx, sp, (long long)n_steps);
    if (r != SB_OK) return r;
    ctx->steps_done += n_steps;
    const long long id = ++ctx->n_event_launches;
    if (!ctx->run_done[id & 1]) CK(cudaEventCreateWithFlags(&ctx->run_done[id & 1], cudaEventDisableTiming));
    CK(cudaEventRecord(ctx->run_done[id & 1], ctx->stream));
    if ((int)ctx->slot_launch.size() < ctx->ring_slots) ctx->slot_launch.resize((size_t)ctx->ring_slots, 0);
    for (int k = 0; k < n_events; ++k)
        if (event_slots[k] >= 0) ctx->slot_launch[(size_t)event_slots[k]] = id;
    return SB_OK;
}

static int check_chain_errors(sb_ctx* ctx) {
    std::vector<int> e((size_t)ctx->dims.n_chains);
    D2H(e.data(), ctx->st.err, e.size());
    for (size_t c = 0; c < e.size(); ++c)
        if (e[c]) {
            ctx->err = "chain " + std::to_string(c + ctx->dims.chain_offset) + " carries error flags " + std::to_string(e[c]) +
                       " (1 = non-finite log-likelihood / prior, 2 = zone beyond the geo-prior scratch, "
                       "4 = alter_p_* on a feature with < 2 applicable states)";
            return SB_E_NUMERIC;
        }
    return SB_OK;
}

int sb_chain_errors(sb_ctx* ctx, int32_t* flags) {
    if (!ctx || !flags) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    D2H(flags, ctx->st.err, (size_t)ctx->dims.n_chains);
    return SB_OK;
}

int sb_read_event_scalars(sb_ctx* ctx, int n_events, double* ll, double* prior) {
    if (!ctx || n_events < 0) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    const size_t n = (size_t)n_events * ctx->dims.n_chains;
    if (n > ctx->ev_scalars_cap) FAIL(SB_E_ARG, "more events than the last sb_run_events recorded");
    std::vector<double2> h(n);
    if (n) D2H(h.data(), ctx->d_ev_scalars, n);
    for (size_t i = 0; i < n; ++i) {
        if (ll) ll[i] = h[i].x;
        if (prior) prior[i] = h[i].y;
    }
    return check_chain_errors(ctx);
}

int sb_read_snapshot(sb_ctx* ctx, int slot, int chain, uint8_t* zones, double* weights, double* p_global,
                     double* p_zones, double* p_families) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, chain);
    if (r != SB_OK) return r;
    if (slot < 0 || slot >= (int)ctx->slot_launch.size() || ctx->slot_launch[(size_t)slot] == 0)
        FAIL(SB_E_ARG, "snapshot slot out of range (no launch has written it)");
    const DevData& d = ctx->d;
    const ChainArrays& st = ctx->st;
    const int N = d.N, F = d.F, S = d.S, Z = d.Z, C = d.C, Fam = d.Fam;
    const size_t FS = (size_t)F * S;
    Staging* sg;
    r = staging_acquire(ctx, std::max(ctx->stage_bytes, ctx->ring_rec), &sg);
    if (r != SB_OK) return r;
    unsigned char* b = sg->buf;
    const unsigned char* rec = ctx->d_ring + ((size_t)slot * ctx->dims.n_chains + chain) * ctx->ring_rec;
    // on the copy stream, behind the launch that wrote the slot: a later launch in flight (which writes other slots) is
    // not waited for
    if (!ctx->copy_stream) CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->run_done[ctx->slot_launch[(size_t)slot] & 1], 0));
    CK(cudaMemcpyAsync(b, rec, ctx->ring_rec, cudaMemcpyDeviceToHost, ctx->copy_stream));
    CK(cudaStreamSynchronize(ctx->copy_stream));
    size_t off = 0;
    if (zones && Z > 0) {
        const int8_t* zid = reinterpret_cast<const int8_t*>(b);
        memset(zones, 0, (size_t)Z * N);
        for (int n = 0; n < N; ++n)
            if (zid[n] >= 0) zones[(size_t)zid[n] * N + n] = 1;
    }
    off += ring_align(st.s_zone_id);
    if (weights) {
        const double* w3 = reinterpret_cast<const double*>(b + off);
        for (int f = 0; f < F; ++f)
            for (int c = 0; c < C; ++c) weights[(size_t)f * C + c] = w3[(size_t)f * 3 + c];
    }
    off += ring_align(st.s_weights * 8);
    if (p_global) memcpy(p_global, b + off, FS * sizeof(double));
    off += ring_align(st.s_p_global * 8);
    if (p_zones && Z > 0) memcpy(p_zones, b + off, (size_t)Z * FS * sizeof(double));
    off += ring_align(st.s_p_zones * 8);
    if (p_families && d.inheritance) memcpy(p_families, b + off, (size_t)Fam * FS * sizeof(double));
    return SB_OK;
}

int sb_snapshot_ring(sb_ctx* ctx, void** ring, int64_t* record_bytes, int64_t* n_records, int64_t* offsets, int64_t* lengths) {
    if (!ctx || !ring || !record_bytes || !n_records) return SB_E_ARG;
    const ChainArrays& st = ctx->st;
    const size_t len[5] = {st.s_zone_id, st.s_weights * 8, st.s_p_global * 8, st.s_p_zones * 8, st.s_p_fam * 8};
    size_t off = 0;
    for (int a = 0; a < 5; ++a) {
        if (offsets) offsets[a] = (int64_t)off;
        if (lengths) lengths[a] = (int64_t)len[a];
        off += ring_align(len[a]);
    }
    *ring = ctx->d_ring;
    *record_bytes = (int64_t)off;
    *n_records = ctx->d_ring ? (int64_t)(ctx->ring_cap / off) : 0;
    return SB_OK;
}

int sb_read_chain_scalars(sb_ctx* ctx, double* ll, double* prior) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    if (ll) D2H(ll, ctx->st.ll, ctx->dims.n_chains);
    if (prior) D2H(prior, ctx->st.prior, ctx->dims.n_chains);
    return check_chain_errors(ctx);
}

int sb_read_op_stats(sb_ctx* ctx, int64_t* accepted, int64_t* rejected) {
    if (!ctx) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    const size_t n = (size_t)ctx->dims.n_chains * SB_N_OPS;
    if (accepted) D2H((unsigned long long*)accepted, ctx->st.acc, n);
    if (rejected) D2H((unsigned long long*)rejected, ctx->st.rej, n);
    CK(cudaStreamSynchronize(ctx->stream));
    return SB_OK;
}

int sb_step_counter(sb_ctx* ctx, int64_t* steps_done) {
    if (!ctx || !steps_done) return SB_E_ARG;
    *steps_done = ctx->steps_done;
    return SB_OK;
}

int sb_kernel_launches(sb_ctx* ctx, int64_t* launches) {
    if (!ctx || !launches) return SB_E_ARG;
    *launches = ctx->launches;
    return SB_OK;
}

int sb_move_log_q(sb_ctx* ctx, int chain, int kind, int zone, int site_new, int site_removed, double rho,
                  double* log_q, double* log_q_back) {
    if (!ctx || !log_q || !log_q_back) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    int r = check_chain(ctx, chain);
    if (r != SB_OK) return r;
    const DevData& d = ctx->d;
    if (!d.adj_ptr) FAIL(SB_E_STATE, "no adjacency uploaded");
    if (zone < 0 || zone >= d.Z) FAIL(SB_E_ARG, "zone out of range");
    if (kind != SB_OP_GROW_ZONE && kind != SB_OP_SHRINK_ZONE && kind != SB_OP_SWAP_ZONE) FAIL(SB_E_ARG, "kind");
    if (kind != SB_OP_SHRINK_ZONE && (site_new < 0 || site_new >= d.N)) FAIL(SB_E_ARG, "site_new out of range");
    if (kind != SB_OP_GROW_ZONE && (site_removed < 0 || site_removed >= d.N)) FAIL(SB_E_ARG, "site_removed out of range");
    // the step kernel itself evaluates the proposal (its own set-up: zone cache, staged adjacency).  Without
    // sb_set_operators the layout is chosen for the zones as they are now
    if (!ctx->has_layout) {
        int cap = 2;
        for (int v : ctx->h_zone_size) cap = std::max(cap, v + 2);
        r = configure_step(ctx, std::min(cap, d.N));
        if (r != SB_OK) return r;
    }
    if (!ctx->d_forced) DA(ctx->d_forced, 1);
    ForcedMove fm = {kind, zone, site_new, site_removed, rho, {NAN, NAN}};
    CK(cudaMemcpyAsync(ctx->d_forced, &fm, sizeof(fm), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    StepParams sp = ctx->sp;
    sp.forced = ctx->d_forced; sp.n_events = 0; sp.events = nullptr;
    // the kernel runs on the whole grid (one team per chain); only the team of `chain` stores its result
    sp.forced_chain = chain;
    r = launch_steps(ctx, sp, 0);
    if (r != SB_OK) return r;
    CK(cudaMemcpyAsync(&fm, ctx->d_forced, sizeof(fm), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *log_q = fm.out[0]; *log_q_back = fm.out[1];
    return SB_OK;
}

int sb_pair_log_q(sb_ctx* ctx, double c0, double c1, double n0, double n1, double precision, double* log_q_ratio) {
    if (!ctx || !log_q_ratio) return SB_E_ARG;
    if (!(c0 > 0.0 && c1 > 0.0 && n0 > 0.0 && n1 > 0.0 && precision >= 0.0)) FAIL(SB_E_ARG, "sb_pair_log_q: pairs must be positive");
    CK(cudaSetDevice(ctx->device));
    if (!ctx->d_pair) DA(ctx->d_pair, 1);
    pair_log_q_kernel<<<1, 32, 0, ctx->stream>>>(c0, c1, n0, n1, precision, ctx->d_pair);
    ++ctx->launches;
    CK(cudaGetLastError());
    D2H(log_q_ratio, ctx->d_pair, 1);
    return SB_OK;
}

int sb_rng_sample(sb_ctx* ctx, int kind, double alpha, uint64_t seed, int n, double* out) {
    if (!ctx || !out || n < 1 || kind < 0 || kind > 3 || (kind == 1 && !(alpha > 0.0))) return SB_E_ARG;
    CK(cudaSetDevice(ctx->device));
    double* dout;
    CK(cudaMalloc((void**)&dout, (size_t)n * sizeof(double)));
    rng_sample_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(kind, alpha, (uint32_t)seed, (uint32_t)(seed >> 32), n, dout);
    ++ctx->launches;
    const int r = d2h(ctx, out, dout, (size_t)n);
    cudaFree(dout);
    return r;
}

}  // extern "C"
