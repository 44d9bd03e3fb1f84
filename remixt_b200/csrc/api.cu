// C-ABI of libremixt_b200 (see include/remixt_b200.h): batches of fits, their device buffers,
// generation tracking of cached quantities, and the host-side sequencing of kernel launches.
//
// A BATCH is G fits that share the data (x, l, y), the state tables and the breakpoint structure and
// differ in h, the dispersion parameters, divergence_weight, the likelihood masks and the variational
// state -- exactly the fits of the reference's initialisation grid (analysis/pipeline.py:42-103).
// Every per-fit array has the fit as its leading dimension and every kernel launch covers a range of
// fits.  A HANDLE addresses a contiguous range [g0, g0+gn) of a batch; rmx_create makes a batch of one.
//
// State semantics kept from the reference (remixt/bpmodel.pyx):
//   framelogprob / log_transmat / joint_posterior_marginals are "as of the last update_p_cn" (:921-962);
//   cached_log_transmat is "as of the constructor or the last update_p_breakpoint" (:604, :985);
//   the likelihood terms always use the CURRENT h / dispersion parameters (:751-853).
#include <cuda_runtime.h>
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cstdlib>
#include "../../include/remixt_b200.h"
#include "kernels.h"

static thread_local std::string g_create_error;

// ---- NCCL, bound at run time (no link-time dependency): the five entry points the sharded fits need
#include <dlfcn.h>
typedef struct ncclComm* rmx_ncclComm_t;
struct rmx_ncclUniqueId { char internal[128]; };
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(rmx_ncclUniqueId*) = nullptr;
    int (*CommInitRank)(rmx_ncclComm_t*, int, rmx_ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(rmx_ncclComm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, rmx_ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string error;
    bool load(const char* path) {
        if (lib) return true;
        const char* names[] = {path, "libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            if (!nm) continue;
            lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) { error = std::string("cannot load NCCL: ") + (dlerror() ? dlerror() : "?"); return false; }
        GetUniqueId = (decltype(GetUniqueId))dlsym(lib, "ncclGetUniqueId");
        CommInitRank = (decltype(CommInitRank))dlsym(lib, "ncclCommInitRank");
        CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
        AllReduce = (decltype(AllReduce))dlsym(lib, "ncclAllReduce");
        GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !CommDestroy || !AllReduce) { error = "NCCL symbols missing"; lib = nullptr; return false; }
        return true;
    }
};
static NcclApi g_nccl;
static const int RMX_NCCL_DOUBLE = 8, RMX_NCCL_SUM = 0;      // ncclFloat64, ncclSum (nccl.h)

// Device-memory pool.  A fit of one sample is followed by the fit of the next with the same shapes (init grids,
// restarts, batches of genomes), and cudaMalloc / cudaFree of a 12 GB working set costs more than the E-step it
// serves.  Freed blocks are kept per (device, size) and handed back to the next batch; rmx_release_cached_memory
// returns them to the driver, and the pool gives way when an allocation fails.
struct DevPool {
    std::mutex mu;
    std::multimap<std::pair<int, size_t>, void*> free_blocks;
    size_t cached = 0;
    cudaError_t get(int device, size_t bytes, void** p, cudaStream_t stream) {
        *p = nullptr;
        {
            std::lock_guard<std::mutex> lk(mu);
            auto it = free_blocks.find({device, bytes});
            if (it != free_blocks.end()) {
                *p = it->second;
                free_blocks.erase(it);
                cached -= bytes;
            }
        }
        // a recycled block starts zeroed, as nothing may depend on the previous owner's contents: cleared on the new
        // owner's stream, ahead of everything the owner does with it (no host synchronisation)
        if (*p) return cudaMemsetAsync(*p, 0, bytes, stream);
        cudaError_t e = cudaMalloc(p, bytes);
        if (e != cudaSuccess) {
            cudaGetLastError();
            release(device);
            e = cudaMalloc(p, bytes);
        }
        return e;
    }
    void put(int device, size_t bytes, void* p) {
        std::lock_guard<std::mutex> lk(mu);
        free_blocks.insert({{device, bytes}, p});
        cached += bytes;
    }
    void release(int device) {
        std::lock_guard<std::mutex> lk(mu);
        for (auto it = free_blocks.begin(); it != free_blocks.end();) {
            if (device < 0 || it->first.first == device) {
                cudaFree(it->second);
                cached -= it->first.second;
                it = free_blocks.erase(it);
            } else {
                ++it;
            }
        }
    }
};
static DevPool g_pool;

// the same for the batch's pinned staging buffers (page-locking a few MB costs milliseconds and is erratic)
struct PinnedPool {
    std::mutex mu;
    std::multimap<size_t, void*> free_blocks;
    cudaError_t get(size_t bytes, void** p, bool zero = true) {
        {
            std::lock_guard<std::mutex> lk(mu);
            auto it = free_blocks.find(bytes);
            if (it != free_blocks.end()) {
                *p = it->second;
                free_blocks.erase(it);
                if (zero) memset(*p, 0, bytes);
                return cudaSuccess;
            }
        }
        return cudaMallocHost(p, bytes);
    }
    void put(size_t bytes, void* p) {
        std::lock_guard<std::mutex> lk(mu);
        free_blocks.insert({bytes, p});
    }
    void release() {
        std::lock_guard<std::mutex> lk(mu);
        for (auto& kv : free_blocks) cudaFreeHost(kv.second);
        free_blocks.clear();
    }
};
static PinnedPool g_pinned;
static std::atomic<int64_t> g_launches{0};

struct ClassMats {
    double *Tmat = nullptr, *Emat = nullptr, *flipc = nullptr, *flipE = nullptr;
    bool built = false;
};

struct Batch {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    std::recursive_mutex mu;
    RmxDims d{};
    int G = 1;
    double pen = 0., prior_ot = 0.01, prior_oa = 0.01;
    std::vector<std::pair<void*, size_t>> allocs;

    // per-fit host state
    std::vector<RmxParams> P;
    std::vector<int> tm, tm_used, tm_cached;
    std::vector<double> logZ;
    std::vector<char> ll_dirty, cn_updated, hist_valid, asum_dirty, path_valid;   // ll_dirty / asum_dirty: term-group bits
    bool params_dirty = true;

    // host copies of the small integer tables
    std::vector<int8_t> h_tab, h_nrm, h_nsub, h_brk;
    std::vector<int64_t> h_tmask, h_amask;   // (G,N)
    std::vector<int> h_path;                 // (G,N)

    // device: shared tables
    int8_t *tab = nullptr, *nrm = nullptr, *nsub = nullptr, *brk = nullptr, *d0 = nullptr, *class_nrm = nullptr;
    double *l = nullptr, *x = nullptr, *y0 = nullptr, *y1 = nullptr;
    int *bidx = nullptr, *borient = nullptr, *tclass = nullptr, *brow = nullptr, *brk_rows = nullptr, *rinfo = nullptr;
    int *krow_ptr = nullptr, *krow_idx = nullptr, *ch_start = nullptr, *ch_end = nullptr, *ch_order = nullptr;
    int *ch_bin_ptr = nullptr, *ch_bin_idx = nullptr, NB = 0;
    int NC = 0, C = 0, R = 0;
    // likelihood-table fast path: normal-copy kinds, wide log table, per-(fit, kind, state) table
    uint8_t *nkind = nullptr, *sflag = nullptr;
    int8_t* kind_nrm = nullptr;
    double2* logw = nullptr;
    double* stab = nullptr;
    double* segf = nullptr;
    double* segk = nullptr;      // (N,3) data-only part of the hoisted constants
    double* logq = nullptr;      // (G,K,B) log q of update_p_breakpoint before normalisation (sharded fits)
    rmx_ncclComm_t comm = nullptr;   // sharded fit over several GPUs: in-stream reductions (rmx_comm_init)
    int comm_rank = 0, comm_size = 1;
    bool is_shard = false;
    uint8_t* bc_tab = nullptr;   // states grouped by tumour-clone totals: [perm (4 KBB)][kbcls (KBB)][scls (S)][didx (BC,BC)][ddel (ND,M)]
    int BC = 0, KBB = 0, ND = 0;
    uint16_t* bc_didx16 = nullptr;   // the same (BC,BC) table with 16-bit entries (more than 255 difference tuples: chain_mma3.cu)
    uint8_t* hord = nullptr;     // (M,S) states ordered by each clone's total (k_hist2)
    int* hgstart = nullptr;
    uint16_t* hord16 = nullptr;  // the same order for any S (k_hist4), with the group starts hgs16 (M, TVN + 1)
    int* hgs16 = nullptr;
    int TVN = 0;                 // number of distinct totals of a clone (0 .. max total)
    int NGS = 0;                 // entries of hgs16: [number of classes, class starts ..., S]
    int Q = 0;
    bool fast_ll = false;
    int64_t rmd = 1;          // R*M*D, fit stride of the breakend-row tables
    ClassMats mats[2];

    // device: per-fit (leading dimension G)
    RmxParams* dP = nullptr;
    uint8_t *tmask = nullptr, *amask = nullptr;
    double *LL = nullptr, *segc = nullptr, *F = nullptr, *fexp = nullptr, *mF = nullptr;
    double *ahat = nullptr, *bhat = nullptr, *post = nullptr, *Asum = nullptr;
    double *q_sw = nullptr, *q_ot = nullptr, *q_oa = nullptr, *p_brk = nullptr;
    double *pd_used = nullptr, *wexp_used = nullptr, *pd_cached = nullptr, *wexp_cached = nullptr;
    double *hist = nullptr, *chain_logz = nullptr, *rowv = nullptr, *rowsum = nullptr;
    double *red_out = nullptr, *part = nullptr, *lattice = nullptr, *segc_s = nullptr, *evalbuf = nullptr;
    int *path = nullptr, *d_rows = nullptr, *d_Rn = nullptr, *dflags = nullptr;
    // batched M-step: the staged sample (row lists in ms_rows, per-fit counts on the host) and the per-call
    // effective counts (0 for inactive fits)
    int *ms_rows = nullptr, *ms_Rn = nullptr, *ms_Rn_all = nullptr;
    double* ms_rowc = nullptr;      // (G, 256, 3) parameter-independent constants of the staged rows
    double* ms_out = nullptr;       // (G, 33, chunks) partial sums of the one-launch sampled objective (slot 32: constants)
    double* hms = nullptr;          // its pinned mirror
    size_t hms_bytes = 0;
    // E[LL] of one dispersion parameter moves only the likelihood terms that parameter enters: the sum of the other
    // terms over the staged sample is computed once per (sample, parameter) and added on the host
    std::vector<double> ms_const;
    int ms_const_id = -1;
    bool ms_const_valid = false;
    std::vector<int> ms_count;      // (G) staged rows per fit; -1 = nothing staged
    bool ms_all = false;            // staged sample = every segment of every fit
    int ms_rmax = 0;
    int64_t part_stride = 0;
    int lattice_fits = 0;
    static const int RED = 64;   // doubles of reduction output per fit

    // pinned host staging
    double* hp = nullptr;       // G*RED doubles
    int* hflags = nullptr;      // 1 int
    int* hrows = nullptr;       // G*N ints
    int* hRn = nullptr;         // G ints
    size_t hp_bytes = 0, hrows_bytes = 0, hRn_bytes = 0;

    // per-kernel-class timing (CUDA events on the stream around each launch, when enabled)
    bool timing = false;
    struct Tev { int cls; cudaEvent_t e0, e1; };
    std::vector<Tev> tev;
    double cls_ms[RMX_K_COUNT_] = {};
    int64_t cls_n[RMX_K_COUNT_] = {};
    int64_t launches = 0;
    cudaEvent_t tbegin(int cls) {
        if (!timing) return nullptr;
        Tev t;
        t.cls = cls;
        cudaEventCreate(&t.e0);
        cudaEventCreate(&t.e1);
        cudaEventRecord(t.e0, stream);
        tev.push_back(t);
        return t.e1;
    }
    void tend(cudaEvent_t e1) {
        if (e1) cudaEventRecord(e1, stream);
    }

    ~Batch() {
        cudaSetDevice(device);
        if (stream) cudaStreamSynchronize(stream);
        if (comm && g_nccl.CommDestroy) g_nccl.CommDestroy(comm);
        for (auto& t : tev) { cudaEventDestroy(t.e0); cudaEventDestroy(t.e1); }
        for (auto& p : allocs) g_pool.put(device, p.second, p.first);
        if (hp) g_pinned.put(hp_bytes, hp);
        if (hflags) g_pinned.put(sizeof(int), hflags);
        if (hrows) g_pinned.put(hrows_bytes, hrows);
        if (hRn) g_pinned.put(hRn_bytes, hRn);
        if (hms) g_pinned.put(hms_bytes, hms);
        if (stream) cudaStreamDestroy(stream);
    }
    void count(int k = 1) {
        launches += k;
        g_launches += k;
    }
};

struct rmx_model_s {
    std::shared_ptr<Batch> b;
    int g0 = 0, gn = 1;
    std::string err;
    cudaEvent_t ev[16] = {};
    int fail(int code, const std::string& m) {
        err = m;
        return code;
    }
};
typedef rmx_model_s Model;
typedef std::lock_guard<std::recursive_mutex> Lock;

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) {                                                                    \
            char buf__[512];                                                                         \
            snprintf(buf__, sizeof buf__, "CUDA error %s at %s:%d: %s", cudaGetErrorName(e__),       \
                     __FILE__, __LINE__, cudaGetErrorString(e__));                                   \
            return m->fail(RMX_ERR_CUDA, buf__);                                                     \
        }                                                                                            \
    } while (0)

template <typename T>
static cudaError_t dalloc(Batch* b, T** p, size_t n) {
    if (n == 0) n = 1;
    cudaError_t e = g_pool.get(b->device, n * sizeof(T), (void**)p, b->stream);
    if (e == cudaSuccess) b->allocs.push_back({(void*)*p, n * sizeof(T)});
    return e;
}
template <typename T>
static cudaError_t upload(Batch* b, T** dptr, const std::vector<T>& v) {
    cudaError_t e = dalloc(b, dptr, v.size());
    if (e != cudaSuccess) return e;
    if (!v.empty()) {   // on the batch stream (after the block's clearing); the source may be a temporary
        e = cudaMemcpyAsync(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, b->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
    }
    return e;
}

static RmxTables tables(const Batch* b) {
    RmxTables T;
    T.tab = b->tab; T.nrm = b->nrm; T.nsub = b->nsub; T.l = b->l; T.x = b->x; T.y0 = b->y0; T.y1 = b->y1;
    T.tmask = b->tmask; T.amask = b->amask;
    return T;
}
static RmxRows rowsof(const Batch* b) {
    RmxRows r;
    r.tclass = b->tclass; r.brow = b->brow; r.d0 = b->d0; r.brk_rows = b->brk_rows;
    return r;
}
static RmxChains chainsof(const Batch* b) {
    RmxChains c;
    c.start = b->ch_start; c.end = b->ch_end; c.order = b->ch_order; c.NC = b->NC;
    c.bin_ptr = b->ch_bin_ptr; c.bin_idx = b->ch_bin_idx; c.NB = b->NB;
    return c;
}

// [g0, g0+gn) split into maximal runs over which key[] is constant and flag[] (if given) is nonzero
template <typename K, typename F>
static void for_runs(int g0, int gn, const std::vector<K>& key, const std::vector<char>* flag, F fn) {
    int g = g0;
    const int ge = g0 + gn;
    while (g < ge) {
        if (flag && !(*flag)[g]) { ++g; continue; }
        int e = g + 1;
        while (e < ge && key[e] == key[g] && (!flag || (*flag)[e])) ++e;
        fn(g, e - g);
        g = e;
    }
}

// ---------------------------------------------------------------------------------------------
// cached-quantity maintenance (all launch-only; callers synchronise)
// ---------------------------------------------------------------------------------------------
// likelihood term groups a parameter enters (bit u: total-read term u; bit 2 + v: the allele terms of v); 15 = all
static int planes_of(int scalar_id) {
    switch (scalar_id) {
    case RMX_S_NEGBIN_R_0: case RMX_S_NEGBIN_HDEL_R_0: return 1;
    case RMX_S_NEGBIN_R_1: case RMX_S_NEGBIN_HDEL_R_1: return 2;
    case RMX_S_NEGBIN_HDEL_MU: return 3;
    case RMX_S_BETABIN_M_0: case RMX_S_BETABIN_LOH_M_0: return 4;
    case RMX_S_BETABIN_M_1: case RMX_S_BETABIN_LOH_M_1: return 8;
    case RMX_S_BETABIN_LOH_P: return 12;
    default: return 15;
    }
}

static int push_params(Model* m) {
    Batch* b = m->b.get();
    if (!b->params_dirty) return RMX_OK;
    // pageable source: the runtime stages small copies before returning, so P may be modified right after
    CK(cudaMemcpyAsync(b->dP, b->P.data(), sizeof(RmxParams) * b->G, cudaMemcpyHostToDevice, b->stream));
    b->params_dirty = false;
    return RMX_OK;
}

// class matrices for transition model tm (penalty is fixed per batch), built on demand
static int ensure_mats(Model* m, int tm) {
    Batch* b = m->b.get();
    ClassMats& cm = b->mats[tm];
    size_t n = (size_t)b->C * b->d.S * b->d.S;
    if (!cm.Tmat) {
        if (dalloc(b, &cm.Tmat, n) != cudaSuccess || dalloc(b, &cm.Emat, n + 64) != cudaSuccess ||
            dalloc(b, &cm.flipc, n) != cudaSuccess || dalloc(b, &cm.flipE, n) != cudaSuccess)
            return m->fail(RMX_ERR_CUDA, "out of device memory for transition class matrices");
    }
    if (!cm.built) {
        rmx_launch_build_class(b->stream, tables(b), b->d, b->C, b->class_nrm, b->pen, tm, cm.Tmat, cm.Emat, cm.flipc,
                               cm.flipE);
        b->count();
        cm.built = true;
    }
    return RMX_OK;
}

static RmxTransSpec spec_of(Batch* b, bool cached, int tm) {
    RmxTransSpec s;
    ClassMats& cm = b->mats[tm];
    s.pd = cached ? b->pd_cached : b->pd_used;
    s.wexp = cached ? b->wexp_cached : b->wexp_used;
    s.Tmat = cm.Tmat; s.Emat = cm.Emat; s.flipc = cm.flipc; s.flipE = cm.flipE;
    s.pen = b->pen;
    s.tm = tm;
    return s;
}

// snapshot the current (p_breakpoint, transition_model) of the fits as their `used` or `cached` spec
static int snapshot_spec(Model* m, int g0, int gn, bool cached) {
    Batch* b = m->b.get();
    int rc = RMX_OK;
    for_runs(g0, gn, b->tm, nullptr, [&](int g, int n) {
        if (rc) return;
        rc = ensure_mats(m, b->tm[g]);
        if (rc) return;
        rmx_launch_prep_brk(b->stream, b->d, g, n, b->R, b->brk_rows, b->bidx, b->borient, b->brk, b->p_brk, b->pen,
                            b->tm[g], b->rmd, cached ? b->pd_cached : b->pd_used, cached ? b->wexp_cached : b->wexp_used);
        if (b->R > 0) b->count();
        for (int k = g; k < g + n; ++k) (cached ? b->tm_cached : b->tm_used)[k] = b->tm[g];
    });
    return rc;
}

// likelihood table of the fits whose h / dispersions / masks changed since it was last built (`only`, indexed by the
// absolute fit, restricts the work to a subset; the others keep their stale planes until something needs them)
static int ensure_ll(Model* m, int g0, int gn, const uint8_t* only = nullptr) {
    Batch* b = m->b.get();
    bool any = false;
    for (int g = g0; g < g0 + gn; ++g) any |= b->ll_dirty[g] != 0 && (!only || only[g]);
    if (!any) return RMX_OK;
    int rc = push_params(m);
    if (rc) return rc;
    RmxTables T = tables(b);
    RmxFast Fp;
    Fp.nkind = b->nkind; Fp.kind_nrm = b->kind_nrm; Fp.Q = b->Q; Fp.logw = b->logw; Fp.stab = b->stab; Fp.sflag = b->sflag; Fp.segf = b->segf;
    // ll_dirty[g] holds the term groups whose parameters changed (planes_of); only those planes are rebuilt (fast
    // path; the general kernel rebuilds all).  Up to RMX_SEL_MAX fits: ONE launch sequence over the list of stale
    // fits, each with its own groups.  Larger batches: one launch sequence per run of fits with the same groups.
    if (b->fast_ll && b->G <= RMX_SEL_MAX) {
        RmxFitSel sel;
        sel.n = 0;
        for (int g = g0; g < g0 + gn; ++g)
            if (b->ll_dirty[g] && (!only || only[g])) {
                sel.fit[sel.n] = (unsigned char)g;
                sel.planes[sel.n] = (unsigned char)b->ll_dirty[g];
                ++sel.n;
            }
        rmx_launch_segconst(b->stream, T, b->d, b->dP, 0, sel.n, nullptr, nullptr, b->d.N, b->segc, b->segk, 0, &sel);
        cudaEvent_t te_ = b->tbegin(RMX_K_LLTABLE);
        rmx_launch_lltable2(b->stream, b->sm_count, T, b->d, b->dP, 0, sel.n, b->segc, Fp, b->LL, b->dflags, 0, &sel);
        b->tend(te_);
        b->count(4);
        for (int k = 0; k < sel.n; ++k) {
            b->asum_dirty[sel.fit[k]] |= (char)sel.planes[k];
            b->ll_dirty[sel.fit[k]] = 0;
        }
        return RMX_OK;
    }
    std::vector<int> key(b->G, 0);
    std::vector<char> need(b->G, 0);
    for (int g = g0; g < g0 + gn; ++g) {
        need[g] = b->ll_dirty[g] != 0 && (!only || only[g]);
        key[g] = b->fast_ll ? (int)b->ll_dirty[g] : (b->ll_dirty[g] ? 15 : 0);
    }
    for_runs(g0, gn, key, &need, [&](int g, int n) {
        const int pm = key[g];
        rmx_launch_segconst(b->stream, T, b->d, b->dP, g, n, nullptr, nullptr, b->d.N, b->segc, b->segk, pm);
        cudaEvent_t te_ = b->tbegin(RMX_K_LLTABLE);
        if (b->fast_ll) {
            rmx_launch_lltable2(b->stream, b->sm_count, T, b->d, b->dP, g, n, b->segc, Fp, b->LL, b->dflags, pm);
            b->count(2);
        } else {
            rmx_launch_lltable(b->stream, b->sm_count, T, b->d, b->dP, g, n, b->segc, b->LL, b->dflags);
        }
        b->tend(te_);
        b->count(2);
        for (int k = g; k < g + n; ++k) { b->ll_dirty[k] = 0; b->asum_dirty[k] |= (char)pm; }
    });
    return RMX_OK;
}

// posterior-weighted likelihood sums of the fits whose posterior or likelihood table changed
static int ensure_asum(Model* m, int g0, int gn, const uint8_t* only = nullptr) {
    Batch* b = m->b.get();
    int rc = ensure_ll(m, g0, gn, only);
    if (rc) return rc;
    if (b->G <= RMX_SEL_MAX) {
        RmxFitSel sel;
        sel.n = 0;
        for (int g = g0; g < g0 + gn; ++g)
            if (b->asum_dirty[g] && (!only || only[g])) {
                sel.fit[sel.n] = (unsigned char)g;
                sel.planes[sel.n] = (unsigned char)b->asum_dirty[g];
                ++sel.n;
            }
        if (sel.n == 0) return RMX_OK;
        { cudaEvent_t te_ = b->tbegin(RMX_K_QSUMS); rmx_launch_qsums(b->stream, b->d, 0, sel.n, b->LL, b->post, b->Asum, 0, &sel); b->tend(te_); }
        b->count();
        for (int k = 0; k < sel.n; ++k) b->asum_dirty[sel.fit[k]] = 0;
        return RMX_OK;
    }
    std::vector<int> key(b->G, 0);
    std::vector<char> need(b->G, 0);
    for (int g = g0; g < g0 + gn; ++g) { need[g] = b->asum_dirty[g] != 0 && (!only || only[g]); key[g] = b->asum_dirty[g]; }
    for_runs(g0, gn, key, &need, [&](int g, int n) {
        { cudaEvent_t te_ = b->tbegin(RMX_K_QSUMS); rmx_launch_qsums(b->stream, b->d, g, n, b->LL, b->post, b->Asum, key[g]); b->tend(te_); }
        b->count();
        for (int k = g; k < g + n; ++k) b->asum_dirty[k] = 0;
    });
    return RMX_OK;
}

static const char* flag_message(int f, int* code) {
    *code = RMX_ERR_VALUE;
    if (f & RMX_F_TOTAL_DEPTH) return "total_depth <= 0 for a copy number state";
    if (f & RMX_F_LOH_P) return "unexpected expected allele ratio p for loh state";
    if (f & RMX_F_BB_P_RANGE) return "p <= 0 or (1 - p) <= 0. for expected allele ratio p";
    if (f & RMX_F_NB_NAN) return "ll is nan in negbin_log_likelihood";
    if (f & RMX_F_BB_NAN) return "ll is nan in betabin_log_likelihood";
    if (f & RMX_F_NB_PARTIAL_NAN) return "partial_mu is nan in negbin_log_likelihood_partial_mu";
    if (f & RMX_F_BB_PARTIAL_NAN) return "partial_p is nan in betabin_log_likelihood_partial_p";
    if (f & RMX_F_DIGAMMA_X) return "x <= 0.0 in digamma";
    *code = RMX_ERR_ASSERT;
    if (f & RMX_F_FRAME_NAN) return "assert not np.any(np.isnan(self.framelogprob))";
    if (f & RMX_F_MSG_NAN) return "assert not np.any(np.isnan(alphas)) / betas (chain message underflow or NaN)";
    if (f & RMX_F_POST_NAN) return "assert not np.any(np.isnan(self.posterior_marginals))";
    if (f & RMX_F_JOINT_NAN) return "assert not np.any(np.isnan(self.joint_posterior_marginals))";
    return "unknown device error flag";
}

// synchronise the stream, collect timing events and device error flags
static int sync_check(Model* m) {
    Batch* b = m->b.get();
    cudaError_t e = cudaMemcpyAsync(b->hflags, b->dflags, sizeof(int), cudaMemcpyDeviceToHost, b->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
    // a launch that could not be configured (the launchers return nothing) leaves its error here
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
        char buf[512];
        snprintf(buf, sizeof buf, "CUDA error %s: %s", cudaGetErrorName(e), cudaGetErrorString(e));
        return m->fail(RMX_ERR_CUDA, buf);
    }
    for (auto& t : b->tev) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.e0, t.e1) == cudaSuccess) {
            b->cls_ms[t.cls] += ms;
            b->cls_n[t.cls] += 1;
        }
        cudaEventDestroy(t.e0);
        cudaEventDestroy(t.e1);
    }
    b->tev.clear();
    int f = *b->hflags;
    if (f) {
        cudaMemsetAsync(b->dflags, 0, sizeof(int), b->stream);
        int code;
        const char* msg = flag_message(f, &code);
        return m->fail(code, msg);
    }
    return RMX_OK;
}

// sharded fit: the reduction outputs red_out[g][0..RED) of the handle's fits summed over the ranks, in stream.
// Every caller has just produced the slots it will read; slots it does not read may hold anything.
static int shard_allreduce(Model* m, double* buf, size_t count) {
    Batch* b = m->b.get();
    if (!b->comm) return RMX_OK;
    const int r = g_nccl.AllReduce(buf, buf, count, RMX_NCCL_DOUBLE, RMX_NCCL_SUM, b->comm, b->stream);
    if (r != 0) return m->fail(RMX_ERR_CUDA, std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error"));
    return RMX_OK;
}
static int shard_reduce_red(Model* m, int slot0, int nslots) {
    Batch* b = m->b.get();
    if (!b->comm) return RMX_OK;
    for (int g = m->g0; g < m->g0 + m->gn; ++g) {
        int rc = shard_allreduce(m, b->red_out + (size_t)g * Batch::RED + slot0, (size_t)nslots);
        if (rc) return rc;
    }
    return RMX_OK;
}
// sharded fit: a shard without rows in a sample still takes part in the reduction, with zeros
static int shard_zero_red(Model* m, int slot0, int nslots) {
    Batch* b = m->b.get();
    if (!b->comm) return RMX_OK;
    for (int g = m->g0; g < m->g0 + m->gn; ++g)
        CK(cudaMemsetAsync(b->red_out + (size_t)g * Batch::RED + slot0, 0, sizeof(double) * nslots, b->stream));
    return RMX_OK;
}

// ---------------------------------------------------------------------------------------------
// coordinate updates over the handle's fits (launch only; callers synchronise)
// ---------------------------------------------------------------------------------------------
static int do_update_framelogprob(Model* m) {
    Batch* b = m->b.get();
    int rc = ensure_ll(m, m->g0, m->gn);
    if (rc) return rc;
    { cudaEvent_t te_ = b->tbegin(RMX_K_FRAME); rmx_launch_frame(b->stream, tables(b), b->d, b->dP, m->g0, m->gn, b->LL, b->q_ot, b->q_oa, b->q_sw, b->F, b->fexp,
                     b->mF, b->dflags); b->tend(te_); }
    b->count();
    return RMX_OK;
}

// have_frame: F / fexp / mF of this update were already produced (by k_round at the end of the previous round);
// defer_post: the posterior marginals are left to the k_round that follows (rmx_variational_update's fused loop)
static int do_update_p_cn(Model* m, bool have_frame = false, bool defer_post = false) {
    Batch* b = m->b.get();
    const int g0 = m->g0, gn = m->gn;
    int rc = have_frame ? ensure_ll(m, g0, gn) : do_update_framelogprob(m);
    if (rc) return rc;
    rc = snapshot_spec(m, g0, gn, false);
    if (rc) return rc;
    int lc = 0;
    for_runs(g0, gn, b->tm_used, nullptr, [&](int g, int n) {
        if (lc) return;
        RmxTransSpec sp = spec_of(b, false, b->tm_used[g]);
        cudaEvent_t te = b->tbegin(RMX_K_CHAIN);
        static const bool force_streamed = getenv("RMX_CHAIN_STREAMED") != nullptr;   // A/B measurements only
        static const bool force_v1 = getenv("RMX_CHAIN_V1") != nullptr;               // A/B measurements only
        lc = (force_streamed || force_v1)
                 ? 2
                 : rmx_launch_chain_mma2(b->stream, tables(b), b->d, chainsof(b), b->rinfo, sp, b->rmd, g, n, b->fexp, b->mF,
                                         b->ahat, b->bhat, b->chain_logz, b->dflags, b->bc_tab, b->BC, b->KBB, b->ND);
        if (lc == 2 && !force_streamed && !force_v1)      // 129 .. 168 states: matrix split between registers and shared memory
            lc = rmx_launch_chain_mma3(b->stream, tables(b), b->d, chainsof(b), b->rinfo, sp, b->rmd, g, n, b->fexp, b->mF,
                                       b->ahat, b->bhat, b->chain_logz, b->dflags, b->bc_tab, b->bc_didx16, b->BC, b->KBB, b->ND);
        if (lc == 2 && !force_streamed && !force_v1)      // 169 .. 360 states: B fragments streamed from L2 under the DMMAs
            lc = rmx_launch_chain_mma_s(b->stream, tables(b), b->d, chainsof(b), b->rinfo, sp, b->rmd, g, n, b->fexp, b->mF,
                                        b->ahat, b->bhat, b->chain_logz, b->dflags, b->hord16, b->hgs16);
        if (lc == 2 && !force_streamed)
            lc = rmx_launch_chain_mma(b->stream, tables(b), b->d, chainsof(b), b->rinfo, sp, b->rmd, g, n, b->fexp,
                                      b->mF, b->ahat, b->bhat, b->chain_logz, b->dflags);
        if (lc == 2)    // state spaces above 128: transition matrix streamed from L2 (chain.cu)
            lc = rmx_launch_chain(b->stream, tables(b), b->d, chainsof(b), rowsof(b), sp, b->rmd, g, n, b->fexp, b->mF,
                                  b->ahat, b->bhat, b->chain_logz, b->dflags);
        b->tend(te);
        b->count();
    });
    if (lc == 2) return m->fail(RMX_ERR_UNSUPPORTED, "this number of copy number states is not supported by the chain kernel");
    if (lc) return m->fail(RMX_ERR_CUDA, "could not configure or launch the chain kernel");
    rmx_launch_reduce_cols(b->stream, b->chain_logz, b->NC, 1, g0, gn, b->NC, b->red_out + 15, Batch::RED);
    if (!defer_post) {
        cudaEvent_t te_ = b->tbegin(RMX_K_POST);
        rmx_launch_post(b->stream, b->d, g0, gn, b->ahat, b->bhat, b->post, b->dflags);
        b->tend(te_);
        b->count();
    }
    b->count();
    for_runs(g0, gn, b->tm_used, nullptr, [&](int g, int n) {
        RmxTransSpec sp = spec_of(b, false, b->tm_used[g]);
        { cudaEvent_t te_ = b->tbegin(RMX_K_HIST); rmx_launch_hist(b->stream, tables(b), b->d, g, n, b->R, rowsof(b), sp, b->rmd, 0, b->ahat, b->bhat, b->fexp, b->hist,
                        b->dflags, b->hord, b->hgstart, b->hord16, b->hgs16, b->TVN, b->NGS); b->tend(te_); }
        if (b->R > 0) b->count();
    });
    for (int g = g0; g < g0 + gn; ++g) {
        b->cn_updated[g] = 1;
        b->hist_valid[g] = 1;
        b->path_valid[g] = 0;
        b->asum_dirty[g] = 15;
    }
    {
        int rc2 = shard_reduce_red(m, 15, 1);      // hmm_log_norm_const = the chains of every shard
        if (rc2) return rc2;
    }
    CK(cudaMemcpyAsync(b->hp + (size_t)g0 * Batch::RED, b->red_out + (size_t)g0 * Batch::RED,
                       sizeof(double) * Batch::RED * gn, cudaMemcpyDeviceToHost, b->stream));
    return RMX_OK;
}

// mode 0: update_p_breakpoint; 1: partial log q of this model's breakend rows into b->logq; 2: p_breakpoint from
// the log q in b->logq (sharded fits: modes 1 and 2 around the allreduce of the shards' partial sums)
static int do_update_p_breakpoint(Model* m, int mode = 0) {
    Batch* b = m->b.get();
    const int g0 = m->g0, gn = m->gn;
    if (mode == 0 && b->comm && b->is_shard && b->d.K > 0) {
        // this shard's partial sums, the sum over the ranks on the stream, the same normalisation everywhere
        const size_t kb = (size_t)b->d.K * b->d.B;
        CK(cudaMemsetAsync(b->logq + (size_t)g0 * kb, 0, sizeof(double) * kb * gn, b->stream));
        int rc = do_update_p_breakpoint(m, 1);
        if (rc) return rc;
        rc = shard_allreduce(m, b->logq + (size_t)g0 * kb, kb * gn);
        if (rc) return rc;
        return do_update_p_breakpoint(m, 2);
    }
    if (mode != 2 && b->d.K > 0 && b->R > 0) {
        // fits still in the constructor state take the histogram of the uniform joint (bpmodel.pyx:566-567)
        std::vector<char> need(b->G, 0);
        for (int g = g0; g < g0 + gn; ++g) need[g] = !b->hist_valid[g];
        for_runs(g0, gn, b->tm_cached, &need, [&](int g, int n) {
            RmxTransSpec sp = spec_of(b, true, b->tm_cached[g]);
            { cudaEvent_t te_ = b->tbegin(RMX_K_HIST); rmx_launch_hist(b->stream, tables(b), b->d, g, n, b->R, rowsof(b), sp, b->rmd, 1, b->ahat, b->bhat, b->fexp,
                            b->hist, b->dflags, b->hord, b->hgstart); b->tend(te_); }
            b->count();
            for (int k = g; k < g + n; ++k) b->hist_valid[k] = 1;
        });
    }
    if (b->d.K > 0) {
        for_runs(g0, gn, b->tm, nullptr, [&](int g, int n) {
            { cudaEvent_t te_ = b->tbegin(RMX_K_BRK_UPDATE); rmx_launch_brk_update(b->stream, b->d, g, n, b->krow_ptr, b->krow_idx, b->brk_rows, b->borient, b->brk, b->hist,
                                  b->rmd, b->pen, b->tm[g], b->p_brk, mode, b->logq); b->tend(te_); }
            b->count();
        });
    }
    if (mode == 1) return RMX_OK;
    return snapshot_spec(m, g0, gn, true);
}

static int do_update_q(Model* m, int which) {
    Batch* b = m->b.get();
    int rc = ensure_asum(m, m->g0, m->gn);
    if (rc) return rc;
    double* out = which == 0 ? b->q_sw : (which == 1 ? b->q_ot : b->q_oa);
    double prior = which == 1 ? b->prior_ot : b->prior_oa;
    rmx_launch_qupdate(b->stream, b->d, m->g0, m->gn, which, b->Asum, b->q_oa, b->q_sw, prior, out);
    b->count();
    return RMX_OK;
}

static int do_update(Model* m, int id) {
    switch (id) {
    case RMX_U_ALLELE_SWAP: return do_update_q(m, 0);
    case RMX_U_CN: return do_update_p_cn(m);
    case RMX_U_BREAKPOINT: return do_update_p_breakpoint(m);
    case RMX_U_OUTLIER_TOTAL: return do_update_q(m, 1);
    case RMX_U_OUTLIER_ALLELE: return do_update_q(m, 2);
    case RMX_U_FRAMELOGPROB: return do_update_framelogprob(m);
    default: return m->fail(RMX_ERR_ARG, "unknown update id");
    }
}

// ---------------------------------------------------------------------------------------------
extern "C" {

const char* rmx_last_error(rmx_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int rmx_device_count(int* count) {
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    *count = (e == cudaSuccess) ? c : 0;
    if (e != cudaSuccess) {
        g_create_error = std::string("cudaGetDeviceCount failed: ") + cudaGetErrorString(e);
        return RMX_ERR_CUDA;
    }
    return RMX_OK;
}

int rmx_release_cached_memory(int device) {
    g_pool.release(device);
    if (device < 0) g_pinned.release();
    return RMX_OK;
}

int rmx_destroy(rmx_handle m) {
    if (!m) return RMX_OK;
    if (m->b) cudaSetDevice(m->b->device);
    for (int i = 0; i < 16; ++i)
        if (m->ev[i]) cudaEventDestroy(m->ev[i]);
    delete m;    // the batch goes with its last handle
    return RMX_OK;
}

#define CREATE_FAIL(code, msg)          \
    do {                                \
        g_create_error = (msg);         \
        delete m;                       \
        return (code);                  \
    } while (0)
#define CCK(call)                                                                               \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            char buf__[512];                                                                    \
            snprintf(buf__, sizeof buf__, "CUDA error %s at %s:%d: %s", cudaGetErrorName(e__),  \
                     __FILE__, __LINE__, cudaGetErrorString(e__));                              \
            CREATE_FAIL(RMX_ERR_CUDA, buf__);                                                   \
        }                                                                                       \
    } while (0)

int rmx_batch_create(const rmx_desc* desc, int num_fits, const double* h_init, const double* divergence_weight,
                     rmx_handle* out) {
    *out = nullptr;
    Model* m = new Model();
    m->b = std::make_shared<Batch>();
    Batch* b = m->b.get();
    const int G = num_fits;
    const int N = desc->num_segments, S = desc->num_cn_states, M = desc->num_clones, B = desc->num_brk_states,
              K = desc->num_breakpoints;
    if (G < 1) CREATE_FAIL(RMX_ERR_VALUE, "num_fits must be positive");
    if (N < 1 || S < 1 || B < 1) CREATE_FAIL(RMX_ERR_VALUE, "num_segments, num_cn_states and num_brk_states must be positive");
    if (M < 2 || M > RMX_MAXM) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "num_clones must be between 2 and 8");
    if (S > 1024) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "num_cn_states > 1024 is not supported");
    if ((int64_t)N * S >= (int64_t)1 << 31) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "num_segments * num_cn_states must be below 2^31");
    if (B > 64) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "num_brk_states > 64 is not supported");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
        CREATE_FAIL(RMX_ERR_CUDA, "no CUDA device available: remixt_b200 has no CPU fallback");
    if (desc->device < 0 || desc->device >= ndev) CREATE_FAIL(RMX_ERR_ARG, "device ordinal out of range");
    b->device = desc->device;
    b->G = G;
    m->g0 = 0;
    m->gn = G;
    CCK(cudaSetDevice(b->device));
    cudaDeviceGetAttribute(&b->sm_count, cudaDevAttrMultiProcessorCount, b->device);
    CCK(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));

    // ---- state table + per-segment normal copies (cn_model.py:359-364 builds cn_states this way)
    b->h_tab.assign((size_t)S * M * 2, 0);
    b->h_nrm.assign((size_t)N * 2, 0);
    int64_t vmax = 0;
    auto fits = [](int64_t v) { return v >= 0 && v <= 100; };
    if (desc->cn_states) {
        const int64_t* cs = desc->cn_states;
        const size_t row = (size_t)S * M * 2;
        for (int s = 0; s < S; ++s)
            for (int mm = 1; mm < M; ++mm)
                for (int a = 0; a < 2; ++a) {
                    int64_t v = cs[((size_t)s * M + mm) * 2 + a];
                    if (!fits(v)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "cn_states entries must be in [0, 100]");
                    b->h_tab[((size_t)s * M + mm) * 2 + a] = (int8_t)v;
                    if (v > vmax) vmax = v;
                }
        for (int n = 0; n < N; ++n) {
            const int64_t* r = cs + (size_t)n * row;
            int64_t n0 = r[0], n1 = r[1];
            if (!fits(n0) || !fits(n1)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "cn_states entries must be in [0, 100]");
            b->h_nrm[2 * n] = (int8_t)n0;
            b->h_nrm[2 * n + 1] = (int8_t)n1;
            if (n0 > vmax) vmax = n0;
            if (n1 > vmax) vmax = n1;
            for (int s = 0; s < S; ++s) {
                const int64_t* rs = r + (size_t)s * M * 2;
                if (rs[0] != n0 || rs[1] != n1)
                    CREATE_FAIL(RMX_ERR_UNSUPPORTED,
                                "cn_states[n,s,0,:] (normal copies) must not depend on the state s");
                for (int e = 2; e < M * 2; ++e)
                    if (rs[e] != (int64_t)b->h_tab[(size_t)s * M * 2 + e])
                        CREATE_FAIL(RMX_ERR_UNSUPPORTED,
                                    "cn_states[n,s,1:,:] (tumour clone copies) must be the same state table for every segment");
            }
        }
    } else if (desc->cn_table && desc->normal_copies) {
        for (int s = 0; s < S; ++s)
            for (int mm = 1; mm < M; ++mm)
                for (int a = 0; a < 2; ++a) {
                    int64_t v = desc->cn_table[((size_t)s * M + mm) * 2 + a];
                    if (!fits(v)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "cn_states entries must be in [0, 100]");
                    b->h_tab[((size_t)s * M + mm) * 2 + a] = (int8_t)v;
                    if (v > vmax) vmax = v;
                }
        for (int n = 0; n < 2 * N; ++n) {
            int64_t v = desc->normal_copies[n];
            if (!fits(v)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "cn_states entries must be in [0, 100]");
            b->h_nrm[n] = (int8_t)v;
            if (v > vmax) vmax = v;
        }
    } else {
        CREATE_FAIL(RMX_ERR_ARG, "either cn_states or (cn_table, normal_copies) must be given");
    }
    b->h_brk.assign((size_t)B * M, 0);
    for (int e = 0; e < B * M; ++e) {
        int64_t v = desc->brk_states[e];
        if (!fits(v)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "brk_states entries must be in [0, 100]");
        b->h_brk[e] = (int8_t)v;
        if (v > vmax) vmax = v;
    }
    // num_alleles_subclonal (bpmodel.pyx:505): alleles whose copy number differs between tumour clones
    b->h_nsub.assign(S, 0);
    int maxtot = 0;
    for (int s = 0; s < S; ++s) {
        int cnt = 0;
        for (int a = 0; a < 2; ++a) {
            int lo = 127, hi = -1;
            for (int mm = 1; mm < M; ++mm) {
                int v = b->h_tab[((size_t)s * M + mm) * 2 + a];
                lo = v < lo ? v : lo;
                hi = v > hi ? v : hi;
            }
            cnt += (hi != lo);
        }
        b->h_nsub[s] = (int8_t)cnt;
        for (int mm = 1; mm < M; ++mm) {
            int t = b->h_tab[((size_t)s * M + mm) * 2] + b->h_tab[((size_t)s * M + mm) * 2 + 1];
            if (t > maxtot) maxtot = t;
        }
    }
    for (int n = 0; n < N; ++n) {
        int t = b->h_nrm[2 * n] + b->h_nrm[2 * n + 1];
        if (t > maxtot) maxtot = t;
    }
    RmxDims& d = b->d;
    d.N = N; d.S = S; d.M = M; d.B = B; d.K = K;
    d.cn_max = (int)vmax;                      // bpmodel.pyx:489
    if (maxtot > d.cn_max + 1)
        CREATE_FAIL(RMX_ERR_UNSUPPORTED,
                    "total copy number of a clone exceeds cn_states.max()+1: the reference's _p_d table would alias");
    d.D = 2 * d.cn_max + 3;
    d.doff = d.cn_max + 1;
    if (d.D > 64) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "maximum copy number above 30 is not supported");

    // ---- rows: telomeres, breakends, transition classes, chains
    std::vector<int> bidx(N), borient(N), tclass(N > 1 ? N - 1 : 0), brow(N > 1 ? N - 1 : 0), brk_rows;
    std::vector<int8_t> d0(N > 1 ? N - 1 : 0), class_nrm;
    std::vector<int> ch_start, ch_end;
    int64_t max_bidx = -1;
    for (int n = 0; n < N; ++n) {
        bidx[n] = (int)desc->breakpoint_idx[n];
        borient[n] = (int)desc->breakpoint_orient[n];
        if (desc->breakpoint_idx[n] > max_bidx) max_bidx = desc->breakpoint_idx[n];
    }
    // a shard of a chromosome-sharded fit holds the breakends of its own chromosomes only (RMX_DESC_SHARD)
    b->is_shard = (desc->flags & RMX_DESC_SHARD) != 0;
    if ((desc->flags & RMX_DESC_SHARD) ? (max_bidx + 1 > K) : (max_bidx + 1 != K))
        CREATE_FAIL(RMX_ERR_VALUE, "breakpoint_idx must have maximum of num_breakpoints positive indices");
    std::map<uint32_t, int> class_of;
    ch_start.push_back(0);
    for (int n = 0; n + 1 < N; ++n) {
        bool tel = desc->is_telomere[n] > 0;
        brow[n] = -1;
        if (bidx[n] >= 0) {
            brow[n] = (int)brk_rows.size();
            brk_rows.push_back(n);
        }
        d0[n] = (int8_t)((b->h_nrm[2 * n] + b->h_nrm[2 * n + 1]) - (b->h_nrm[2 * n + 2] + b->h_nrm[2 * n + 3]));
        if (tel) {
            tclass[n] = -1;
            ch_end.push_back(n);
            ch_start.push_back(n + 1);
        } else {
            uint32_t key = ((uint32_t)(uint8_t)b->h_nrm[2 * n] << 24) | ((uint32_t)(uint8_t)b->h_nrm[2 * n + 1] << 16) |
                           ((uint32_t)(uint8_t)b->h_nrm[2 * n + 2] << 8) | (uint32_t)(uint8_t)b->h_nrm[2 * n + 3];
            auto it = class_of.find(key);
            if (it == class_of.end()) {
                it = class_of.emplace(key, (int)class_of.size()).first;
                for (int e = 0; e < 4; ++e) class_nrm.push_back(b->h_nrm[2 * n + e]);
            }
            tclass[n] = it->second;
        }
    }
    ch_end.push_back(N - 1);
    b->NC = (int)ch_start.size();
    b->C = (int)class_of.size();
    if (b->C == 0) { b->C = 1; class_nrm.assign(4, 1); }
    b->R = (int)brk_rows.size();
    b->rmd = (int64_t)b->R * M * d.D;
    if (b->rmd < 1) b->rmd = 1;
    // chains longest first: the launch order of the chain kernel (longest sequential jobs start first)
    std::vector<int> ch_order(b->NC);
    for (int c = 0; c < b->NC; ++c) ch_order[c] = c;
    std::stable_sort(ch_order.begin(), ch_order.end(), [&](int a2, int b2) {
        return (ch_end[a2] - ch_start[a2]) > (ch_end[b2] - ch_start[b2]);
    });
    // bins of about equal total length for the Viterbi kernel: as many as the longest chain allows, filled longest
    // chain first into the currently shortest bin
    std::vector<int> bin_ptr, bin_idx;
    {
        int64_t total = 0, longest = 1;
        for (int c = 0; c < b->NC; ++c) {
            const int64_t len = ch_end[c] - ch_start[c] + 1;
            total += len;
            if (len > longest) longest = len;
        }
        int nb = (int)(total / longest);
        if (nb < 1) nb = 1;
        if (nb > b->NC) nb = b->NC;
        std::vector<int64_t> load(nb, 0);
        std::vector<std::vector<int>> members(nb);
        for (int c : ch_order) {
            int best = 0;
            for (int q = 1; q < nb; ++q)
                if (load[q] < load[best]) best = q;
            load[best] += ch_end[c] - ch_start[c] + 1;
            members[best].push_back(c);
        }
        bin_ptr.push_back(0);
        for (int q = 0; q < nb; ++q) {
            if (members[q].empty()) continue;
            bin_idx.insert(bin_idx.end(), members[q].begin(), members[q].end());
            bin_ptr.push_back((int)bin_idx.size());
        }
        b->NB = (int)bin_ptr.size() - 1;
    }
    // breakpoint -> its breakend rows, ascending n (the order bpmodel.pyx:971-980 visits them)
    std::vector<int> krow_ptr(K + 1, 0), krow_idx(b->R);
    for (int r = 0; r < b->R; ++r) krow_ptr[bidx[brk_rows[r]] + 1]++;
    for (int k = 0; k < K; ++k) krow_ptr[k + 1] += krow_ptr[k];
    {
        std::vector<int> fill(krow_ptr.begin(), krow_ptr.end() - 1);
        for (int r = 0; r < b->R; ++r) krow_idx[fill[bidx[brk_rows[r]]]++] = r;
    }

    // ---- uploads (shared)
    std::vector<double> y0(N), y1(N);
    for (int n = 0; n < N; ++n) { y0[n] = desc->y[2 * n]; y1[n] = desc->y[2 * n + 1]; }
    CCK(upload(b, &b->tab, b->h_tab));
    CCK(upload(b, &b->nrm, b->h_nrm));
    CCK(upload(b, &b->nsub, b->h_nsub));
    CCK(upload(b, &b->brk, b->h_brk));
    CCK(upload(b, &b->d0, d0));
    CCK(upload(b, &b->class_nrm, class_nrm));
    CCK(upload(b, &b->l, std::vector<double>(desc->l, desc->l + N)));
    CCK(upload(b, &b->x, std::vector<double>(desc->x, desc->x + N)));
    CCK(upload(b, &b->y0, y0));
    CCK(upload(b, &b->y1, y1));
    CCK(upload(b, &b->bidx, bidx));
    CCK(upload(b, &b->borient, borient));
    CCK(upload(b, &b->tclass, tclass));
    CCK(upload(b, &b->brow, brow));
    bool d0_fits = true;        // the packed row info holds d0 + 8 in four bits
    for (size_t n = 0; n < d0.size(); ++n) d0_fits = d0_fits && d0[n] >= -8 && d0[n] <= 7;
    if (S <= 250) {
        // states grouped by the totals of their tumour clones (the breakend rows of chain_mma2.cu): classes in order of
        // first appearance, each padded to whole row blocks of four states (255 = padding)
        std::vector<std::vector<int>> keys;
        std::vector<uint8_t> scls(S);
        for (int s2 = 0; s2 < S; ++s2) {
            std::vector<int> key(M, 0);
            for (int mm = 1; mm < M; ++mm)
                key[mm] = b->h_tab[((size_t)s2 * M + mm) * 2] + b->h_tab[((size_t)s2 * M + mm) * 2 + 1];
            size_t c = 0;
            while (c < keys.size() && keys[c] != key) ++c;
            if (c == keys.size()) keys.push_back(key);
            scls[s2] = (uint8_t)(c < 255 ? c : 255);
        }
        if (keys.size() <= 64) {
            std::vector<uint8_t> perm, kbcls;
            for (size_t c = 0; c < keys.size(); ++c) {
                for (int s2 = 0; s2 < S; ++s2)
                    if (scls[s2] == c) perm.push_back((uint8_t)s2);
                while (perm.size() % 4) perm.push_back(255);
                while (kbcls.size() < perm.size() / 4) kbcls.push_back((uint8_t)c);
            }
            // the distinct tuples of per-clone total differences between two classes, and which one each pair is
            std::vector<std::vector<int>> dels;
            std::vector<uint8_t> didx(keys.size() * keys.size());
            std::vector<uint16_t> didx16(keys.size() * keys.size());
            for (size_t ci = 0; ci < keys.size(); ++ci)
                for (size_t cj = 0; cj < keys.size(); ++cj) {
                    std::vector<int> del(M, 0);
                    for (int mm = 1; mm < M; ++mm) del[mm] = keys[ci][mm] - keys[cj][mm];
                    size_t q = 0;
                    while (q < dels.size() && dels[q] != del) ++q;
                    if (q == dels.size()) dels.push_back(del);
                    didx[ci * keys.size() + cj] = (uint8_t)(q < 255 ? q : 255);
                    didx16[ci * keys.size() + cj] = (uint16_t)(q < 65535 ? q : 65535);
                }
            std::vector<uint8_t> bc(perm);
            bc.insert(bc.end(), kbcls.begin(), kbcls.end());
            bc.insert(bc.end(), scls.begin(), scls.end());
            bc.insert(bc.end(), didx.begin(), didx.end());
            for (size_t q = 0; q < dels.size(); ++q)
                for (int mm = 0; mm < M; ++mm) bc.push_back((uint8_t)(dels[q][mm] + 128));
            b->ND = (int)dels.size();
            b->BC = (int)keys.size();
            b->KBB = (int)kbcls.size();
            CCK(upload(b, &b->bc_tab, bc));
            CCK(upload(b, &b->bc_didx16, didx16));
        }
    }
    if (b->C <= 254 && b->R < (1 << 19) && d0_fits) {     // packed row info of the one-barrier chain kernels
        std::vector<int> rinfo(tclass.size());
        for (size_t n = 0; n < tclass.size(); ++n)
            rinfo[n] = (tclass[n] + 1) | ((d0[n] + 8) << 8) | ((brow[n] + 1) << 12);
        CCK(upload(b, &b->rinfo, rinfo));
    }
    CCK(upload(b, &b->brk_rows, brk_rows));
    CCK(upload(b, &b->krow_ptr, krow_ptr));
    CCK(upload(b, &b->krow_idx, krow_idx));
    CCK(upload(b, &b->ch_start, ch_start));
    CCK(upload(b, &b->ch_end, ch_end));
    CCK(upload(b, &b->ch_order, ch_order));
    CCK(upload(b, &b->ch_bin_ptr, bin_ptr));
    CCK(upload(b, &b->ch_bin_idx, bin_idx));

    // ---- per-fit parameters (bpmodel.pyx:573-597)
    b->P.resize(G);
    for (int g = 0; g < G; ++g) {
        RmxParams& P = b->P[g];
        memset(&P, 0, sizeof P);
        for (int mm = 0; mm < M; ++mm) P.h[mm] = h_init[(size_t)g * M + mm];
        P.negbin_r[0] = 500.; P.negbin_r[1] = 10.;
        P.negbin_hdel_mu = 1e-5; P.negbin_hdel_r[0] = 10.; P.negbin_hdel_r[1] = 1.;
        P.betabin_M[0] = 500.; P.betabin_M[1] = 10.;
        P.betabin_loh_p = 1e-3; P.betabin_loh_M[0] = 10.; P.betabin_loh_M[1] = 1.;
        P.divergence_weight = fabs(divergence_weight[g]);
        P.normal_contamination = desc->normal_contamination ? 1 : 0;
    }
    b->pen = fabs(desc->transition_penalty);
    b->tm.assign(G, 0);
    b->tm_used.assign(G, 0);
    b->tm_cached.assign(G, 0);
    b->logZ.assign(G, 0.);
    b->ll_dirty.assign(G, 15);
    b->cn_updated.assign(G, 0);
    b->hist_valid.assign(G, 0);
    b->asum_dirty.assign(G, 15);
    b->path_valid.assign(G, 0);
    b->h_tmask.assign((size_t)G * N, 1);
    b->h_amask.assign((size_t)G * N, 1);
    b->h_path.assign((size_t)G * N, 0);

    // ---- per-fit device state
    const size_t NS = (size_t)N * S, GNS = (size_t)G * NS, GN = (size_t)G * N;
    CCK(dalloc(b, &b->dP, (size_t)G));
    CCK(upload(b, &b->tmask, std::vector<uint8_t>(GN, 1)));
    CCK(upload(b, &b->amask, std::vector<uint8_t>(GN, 1)));
    CCK(dalloc(b, &b->LL, GNS * RMX_NTERM));
    CCK(dalloc(b, &b->segc, GN * 12));
    CCK(dalloc(b, &b->segk, (size_t)N * 3));
    rmx_launch_segk(b->stream, tables(b), d, b->segk);
    b->count();
    {
        // distinct normal-copy pairs (kinds) and the wide log table of the fast path (common.cuh)
        std::vector<uint8_t> nkind(N);
        std::vector<int8_t> kind_nrm;
        std::map<int, int> kind_of;
        bool ok = true;
        for (int n = 0; n < N; ++n) {
            int key = ((int)(uint8_t)b->h_nrm[2 * n] << 8) | (int)(uint8_t)b->h_nrm[2 * n + 1];
            auto it = kind_of.find(key);
            if (it == kind_of.end()) {
                if (kind_of.size() >= 255) { ok = false; break; }
                it = kind_of.emplace(key, (int)kind_of.size()).first;
                kind_nrm.push_back(b->h_nrm[2 * n]);
                kind_nrm.push_back(b->h_nrm[2 * n + 1]);
            }
            nkind[n] = (uint8_t)it->second;
        }
        b->Q = (int)kind_of.size();
        b->fast_ll = ok && b->Q > 0 && rmx_lltable2_ok(d, b->Q);
        if (b->fast_ll) {
            std::vector<double2> lw(RMX_LW_SIZE);
            for (int o = 0; o < RMX_LW_NOCT; ++o)
                for (int j = 0; j < 128; ++j) {
                    const long double c = ldexpl(1.0L + ((long double)j + 0.5L) / 128.0L, RMX_LW_E0 + o);
                    const double inv = (double)(1.0L / c);
                    lw[o * 128 + j] = make_double2(inv, (double)(-logl((long double)inv)));
                }
            CCK(upload(b, &b->nkind, nkind));
            CCK(upload(b, &b->kind_nrm, kind_nrm));
            CCK(upload(b, &b->logw, lw));
            CCK(dalloc(b, &b->stab, (size_t)G * 8 * b->Q * S));
            CCK(dalloc(b, &b->sflag, (size_t)G * b->Q * S));
            CCK(dalloc(b, &b->segf, GN * 12));
        }
    }
    if (M >= 2 && S <= 65535) {
        // states ordered by each clone's total copy number, and where every total's group starts (k_hist4)
        int maxtot = 0;
        for (int s2 = 0; s2 < S; ++s2)
            for (int mm = 1; mm < M; ++mm)
                maxtot = std::max(maxtot, (int)b->h_tab[((size_t)s2 * M + mm) * 2] + (int)b->h_tab[((size_t)s2 * M + mm) * 2 + 1]);
        const int TVN = maxtot + 1;
        // class order: states sorted (stably) by the tuple of their tumour-clone totals; gs = [number of classes,
        // class starts ..., S]
        auto key_of = [&](int s2) {
            std::vector<int> key(M, 0);
            for (int mm = 1; mm < M; ++mm)
                key[mm] = (int)b->h_tab[((size_t)s2 * M + mm) * 2] + (int)b->h_tab[((size_t)s2 * M + mm) * 2 + 1];
            return key;
        };
        std::vector<int> order(S);
        for (int s2 = 0; s2 < S; ++s2) order[s2] = s2;
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return key_of(x) < key_of(y); });
        std::vector<uint16_t> ord16((size_t)M * S, 0);
        std::vector<int> gs(1, 0);
        for (int k = 0; k < S; ++k) {
            ord16[k] = (uint16_t)order[k];
            if (k == 0 || key_of(order[k]) != key_of(order[k - 1])) gs.push_back(k);
        }
        gs[0] = (int)gs.size() - 1;
        gs.push_back(S);
        b->NGS = (int)gs.size();
        CCK(upload(b, &b->hord16, ord16));
        CCK(upload(b, &b->hgs16, gs));
        b->TVN = TVN;
    }
    if (rmx_hist2_ok(d)) {
        CCK(dalloc(b, &b->hord, (size_t)M * S));
        CCK(dalloc(b, &b->hgstart, (size_t)rmx_hist2_gstart_count(d)));
        rmx_launch_hist_prep(b->stream, tables(b), d, b->hord, b->hgstart);
    }
    CCK(dalloc(b, &b->F, GNS));
    CCK(dalloc(b, &b->fexp, GNS));
    CCK(dalloc(b, &b->mF, GN));
    CCK(dalloc(b, &b->ahat, GNS));
    CCK(dalloc(b, &b->bhat, GNS));
    CCK(dalloc(b, &b->post, GNS));
    CCK(dalloc(b, &b->Asum, GN * RMX_NTERM));
    CCK(dalloc(b, &b->q_sw, GN * 2));
    CCK(dalloc(b, &b->q_ot, GN * 2));
    CCK(dalloc(b, &b->q_oa, GN * 2));
    CCK(dalloc(b, &b->p_brk, (size_t)G * K * B));
    CCK(dalloc(b, &b->logq, (size_t)G * K * B));
    const size_t grmd = (size_t)G * b->rmd;
    CCK(dalloc(b, &b->pd_used, grmd));
    CCK(dalloc(b, &b->wexp_used, grmd));
    CCK(dalloc(b, &b->pd_cached, grmd));
    CCK(dalloc(b, &b->wexp_cached, grmd));
    CCK(dalloc(b, &b->hist, grmd));
    CCK(dalloc(b, &b->chain_logz, (size_t)G * b->NC));
    CCK(dalloc(b, &b->rowv, GN * 8 + (size_t)G * K * B));
    CCK(dalloc(b, &b->rowsum, GN));
    CCK(dalloc(b, &b->red_out, (size_t)G * Batch::RED));
    b->part_stride = (int64_t)rmx_ell_blocks(d, N) * RMX_MAXM + 64;
    CCK(dalloc(b, &b->part, (size_t)G * b->part_stride));
    CCK(dalloc(b, &b->segc_s, GN * 12));
    CCK(dalloc(b, &b->evalbuf, 16));
    CCK(dalloc(b, &b->path, GN));
    CCK(dalloc(b, &b->d_rows, GN));
    CCK(dalloc(b, &b->d_Rn, (size_t)G));
    CCK(dalloc(b, &b->ms_rows, GN));
    CCK(dalloc(b, &b->ms_Rn, (size_t)G));
    CCK(dalloc(b, &b->ms_Rn_all, (size_t)G));
    CCK(dalloc(b, &b->ms_rowc, (size_t)G * 256 * 3));
    CCK(dalloc(b, &b->ms_out, (size_t)G * 33 * rmx_ell_sample_chunks()));
    b->hms_bytes = (size_t)G * 33 * rmx_ell_sample_chunks() * sizeof(double);
    CCK(g_pinned.get(b->hms_bytes, (void**)&b->hms));
    b->ms_count.assign(G, -1);
    b->ms_const.assign(G, 0.);
    CCK(dalloc(b, &b->dflags, 1));
    CCK(cudaMemsetAsync(b->dflags, 0, sizeof(int), b->stream));
    b->hp_bytes = (size_t)G * Batch::RED * sizeof(double);
    b->hrows_bytes = GN * sizeof(int);
    b->hRn_bytes = (size_t)G * sizeof(int);
    CCK(g_pinned.get(b->hp_bytes, (void**)&b->hp));
    CCK(g_pinned.get(sizeof(int), (void**)&b->hflags));
    CCK(g_pinned.get(b->hrows_bytes, (void**)&b->hrows));
    CCK(g_pinned.get(b->hRn_bytes, (void**)&b->hRn));
    *b->hflags = 0;

    {   // initial factors (bpmodel.pyx:546-584), the same for every fit, filled on the device
        double sum = 0.;    // posterior_marginals[:] = 1; /= sum  (bpmodel.pyx:564-565)
        for (int s = 0; s < S; ++s) sum += 1.0;
        rmx_launch_fill2(b->stream, b->F, (int64_t)GNS, 1.0, 1.0);
        rmx_launch_fill2(b->stream, b->post, (int64_t)GNS, 1.0 / sum, 1.0 / sum);
        rmx_launch_fill2(b->stream, b->q_sw, (int64_t)(2 * GN), 0.5, 0.5);
        rmx_launch_fill2(b->stream, b->q_ot, (int64_t)(2 * GN), 1. - 0.01, 0.01);
        rmx_launch_fill2(b->stream, b->q_oa, (int64_t)(2 * GN), 1. - 0.01, 0.01);
        b->count(5);
        // p_breakpoint favours states with at most one copy in every clone (bpmodel.pyx:546-554)
        std::vector<double> pb((size_t)G * K * B, 0.);
        double cnt = 0.;
        std::vector<double> rowp(B, 0.);
        for (int bb = 0; bb < B; ++bb) {
            int mx = 0;
            for (int mm = 0; mm < M; ++mm) mx = b->h_brk[(size_t)bb * M + mm] > mx ? b->h_brk[(size_t)bb * M + mm] : mx;
            if (mx <= 1) { rowp[bb] = 1.; cnt += 1.; }
        }
        for (size_t k = 0; k < (size_t)G * K; ++k)
            for (int bb = 0; bb < B; ++bb) pb[k * B + bb] = rowp[bb] / cnt;
        // on the batch stream: a recycled block's clearing is queued there
        if (K > 0) {
            CCK(cudaMemcpyAsync(b->p_brk, pb.data(), pb.size() * sizeof(double), cudaMemcpyHostToDevice, b->stream));
            CCK(cudaStreamSynchronize(b->stream));
        }
    }
    // first cached transition expectation (bpmodel.pyx:604), transition_model still 0 here
    if (snapshot_spec(m, 0, G, true) != RMX_OK) { std::string e = m->err; CREATE_FAIL(RMX_ERR_CUDA, e); }
    if (snapshot_spec(m, 0, G, false) != RMX_OK) { std::string e = m->err; CREATE_FAIL(RMX_ERR_CUDA, e); }
    CCK(cudaStreamSynchronize(b->stream));
    CCK(cudaGetLastError());
    *out = m;
    return RMX_OK;
}

int rmx_create(const rmx_desc* desc, rmx_handle* out) {
    return rmx_batch_create(desc, 1, desc->h_init, &desc->divergence_weight, out);
}

int rmx_batch_view(rmx_handle m, int first_fit, int num_fits, rmx_handle* out) {
    *out = nullptr;
    if (!m) return RMX_ERR_ARG;
    if (first_fit < 0 || num_fits < 1 || first_fit + num_fits > m->gn)
        return m->fail(RMX_ERR_ARG, "fit range outside the handle's fits");
    Model* v = new Model();
    v->b = m->b;
    v->g0 = m->g0 + first_fit;
    v->gn = num_fits;
    *out = v;
    return RMX_OK;
}

int rmx_num_fits(rmx_handle m, int* num_fits) {
    if (!m) return RMX_ERR_ARG;
    *num_fits = m->gn;
    return RMX_OK;
}

// per-fit element count of an array attribute
static int64_t array_count(Batch* b, int id, bool* is_i64) {
    const RmxDims& d = b->d;
    *is_i64 = false;
    switch (id) {
    case RMX_A_H: return d.M;
    case RMX_A_P_BREAKPOINT: return (int64_t)d.K * d.B;
    case RMX_A_P_ALLELE_SWAP:
    case RMX_A_P_OUTLIER_TOTAL:
    case RMX_A_P_OUTLIER_ALLELE: return (int64_t)d.N * 2;
    case RMX_A_POSTERIOR_MARGINALS:
    case RMX_A_FRAMELOGPROB: return (int64_t)d.N * d.S;
    case RMX_A_TOTAL_LIKELIHOOD_MASK:
    case RMX_A_ALLELE_LIKELIHOOD_MASK: *is_i64 = true; return d.N;
    case RMX_A_LOG_TRANSMAT:
    case RMX_A_CACHED_LOG_TRANSMAT:
    case RMX_A_JOINT_POSTERIOR_MARGINALS: return (int64_t)(d.N - 1) * d.S * d.S;
    case RMX_A_L:
    case RMX_A_X: return d.N;
    case RMX_A_Y: return (int64_t)d.N * 2;
    case RMX_A_STATE_PATH: *is_i64 = true; return d.N;
    case RMX_A_LL_TABLE: return (int64_t)RMX_NTERM * d.N * d.S;
    default: return -1;
    }
}

int rmx_set_array(rmx_handle m, int id, const void* src, int64_t count) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    bool i64;
    const int64_t per = array_count(b, id, &i64);
    if (per < 0) return m->fail(RMX_ERR_ARG, "unknown array id");
    const bool bcast = m->gn > 1 && count == per && id != RMX_A_H;      // one value for every fit of the handle
    if (per * m->gn != count && !bcast)
        return m->fail(RMX_ERR_VALUE, "array has the wrong number of elements for this attribute");
    const int g0 = m->g0, gn = m->gn;
    double* dst = nullptr;
    switch (id) {
    case RMX_A_H: {
        const double* s = (const double*)src;
        for (int g = 0; g < gn; ++g) {
            for (int k = 0; k < b->d.M; ++k) b->P[g0 + g].h[k] = s[(size_t)g * b->d.M + k];
            b->ll_dirty[g0 + g] = 15;
        }
        b->params_dirty = true;
        b->ms_const_valid = false;
        return RMX_OK;
    }
    case RMX_A_P_BREAKPOINT: dst = b->p_brk; break;
    case RMX_A_P_ALLELE_SWAP: dst = b->q_sw; break;
    case RMX_A_P_OUTLIER_TOTAL: dst = b->q_ot; break;
    case RMX_A_P_OUTLIER_ALLELE: dst = b->q_oa; break;
    case RMX_A_POSTERIOR_MARGINALS:
        dst = b->post;
        for (int g = g0; g < g0 + gn; ++g) b->asum_dirty[g] = 15;
        break;
    case RMX_A_FRAMELOGPROB: dst = b->F; break;
    case RMX_A_TOTAL_LIKELIHOOD_MASK:
    case RMX_A_ALLELE_LIKELIHOOD_MASK: {
        const int64_t* s = (const int64_t*)src;
        const size_t n = (size_t)gn * b->d.N, off = (size_t)g0 * b->d.N, N1 = (size_t)b->d.N;
        std::vector<uint8_t> v(n);
        std::vector<int64_t>& keep = id == RMX_A_TOTAL_LIKELIHOOD_MASK ? b->h_tmask : b->h_amask;
        if (bcast) {
            for (size_t e = 0; e < N1; ++e) { keep[off + e] = s[e]; v[e] = s[e] != 0; }
            for (int g = 1; g < gn; ++g) {
                memcpy(&keep[off + (size_t)g * N1], &keep[off], N1 * sizeof(int64_t));
                memcpy(&v[(size_t)g * N1], &v[0], N1);
            }
        } else {
            for (size_t e = 0; e < n; ++e) { keep[off + e] = s[e]; v[e] = s[e] != 0; }
        }
        CK(cudaMemcpyAsync((id == RMX_A_TOTAL_LIKELIHOOD_MASK ? b->tmask : b->amask) + off, v.data(), n,
                           cudaMemcpyHostToDevice, b->stream));
        CK(cudaStreamSynchronize(b->stream));
        for (int g = g0; g < g0 + gn; ++g) b->ll_dirty[g] = 15;
        return RMX_OK;
    }
    default: return m->fail(RMX_ERR_ARG, "this attribute cannot be assigned");
    }
    b->ms_const_valid = false;
    if (count > 0) {
        if (bcast) {
            for (int g = 0; g < gn; ++g)
                CK(cudaMemcpyAsync(dst + (size_t)(g0 + g) * per, src, (size_t)per * sizeof(double), cudaMemcpyHostToDevice,
                                   b->stream));
        } else {
            CK(cudaMemcpyAsync(dst + (size_t)g0 * per, src, (size_t)count * sizeof(double), cudaMemcpyHostToDevice,
                               b->stream));
        }
        CK(cudaStreamSynchronize(b->stream));
    }
    return RMX_OK;
}

// the (N-1,S,S) attributes of ONE fit, materialised in chunks (inspection / tests only)
static int materialize(Model* m, int g, int id, double* dst) {
    Batch* b = m->b.get();
    const RmxDims& d = b->d;
    const int64_t SS = (int64_t)d.S * d.S;
    const int64_t nrows = d.N - 1;
    if (nrows <= 0) return RMX_OK;
    if (id == RMX_A_LOG_TRANSMAT && !b->cn_updated[g]) {   // constructor state: zeros (bpmodel.pyx:558)
        memset(dst, 0, (size_t)(nrows * SS) * sizeof(double));
        return RMX_OK;
    }
    const bool cached = id == RMX_A_CACHED_LOG_TRANSMAT;
    const int tm = cached ? b->tm_cached[g] : b->tm_used[g];
    int rc = ensure_mats(m, tm);
    if (rc) return rc;
    RmxTransSpec sp = spec_of(b, cached, tm);
    sp.pd += (int64_t)g * b->rmd;
    sp.wexp += (int64_t)g * b->rmd;
    const int64_t NSo = (int64_t)g * d.N * d.S;
    int64_t chunk = (int64_t)(64 << 20) / (SS * 8);
    if (chunk < 1) chunk = 1;
    if (chunk > nrows) chunk = nrows;
    double* buf = nullptr;
    CK(cudaMalloc((void**)&buf, (size_t)(chunk * SS) * sizeof(double)));
    for (int64_t r0 = 0; r0 < nrows && rc == RMX_OK; r0 += chunk) {
        int64_t nr = nrows - r0 < chunk ? nrows - r0 : chunk;
        if (id == RMX_A_JOINT_POSTERIOR_MARGINALS)
            rmx_launch_materialize_joint(b->stream, tables(b), d, rowsof(b), sp, b->cn_updated[g] ? 0 : 1, b->ahat + NSo,
                                         b->bhat + NSo, b->fexp + NSo, r0, nr, buf);
        else
            rmx_launch_materialize_T(b->stream, tables(b), d, rowsof(b), sp, r0, nr, buf);
        b->count();
        cudaError_t e = cudaMemcpyAsync(dst + r0 * SS, buf, (size_t)(nr * SS) * sizeof(double), cudaMemcpyDeviceToHost,
                                        b->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
        if (e != cudaSuccess) rc = m->fail(RMX_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaFree(buf);
    return rc;
}

int rmx_get_array(rmx_handle m, int id, void* dstv, int64_t count) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    bool i64;
    const int64_t per = array_count(b, id, &i64);
    if (per < 0) return m->fail(RMX_ERR_ARG, "unknown array id");
    if (per * m->gn != count) return m->fail(RMX_ERR_VALUE, "array has the wrong number of elements for this attribute");
    const int g0 = m->g0, gn = m->gn, N = b->d.N;
    const double* src = nullptr;
    bool shared = false;
    switch (id) {
    case RMX_A_H:
        for (int g = 0; g < gn; ++g) memcpy((double*)dstv + (size_t)g * b->d.M, b->P[g0 + g].h, sizeof(double) * b->d.M);
        return RMX_OK;
    case RMX_A_P_BREAKPOINT: src = b->p_brk; break;
    case RMX_A_P_ALLELE_SWAP: src = b->q_sw; break;
    case RMX_A_P_OUTLIER_TOTAL: src = b->q_ot; break;
    case RMX_A_P_OUTLIER_ALLELE: src = b->q_oa; break;
    case RMX_A_POSTERIOR_MARGINALS: src = b->post; break;
    case RMX_A_FRAMELOGPROB: src = b->F; break;
    case RMX_A_LL_TABLE: {
        int rc = ensure_ll(m, g0, gn);
        if (rc) return rc;
        rc = sync_check(m);
        if (rc) return rc;
        src = b->LL;
        break;
    }
    case RMX_A_TOTAL_LIKELIHOOD_MASK: memcpy(dstv, b->h_tmask.data() + (size_t)g0 * N, sizeof(int64_t) * N * gn); return RMX_OK;
    case RMX_A_ALLELE_LIKELIHOOD_MASK: memcpy(dstv, b->h_amask.data() + (size_t)g0 * N, sizeof(int64_t) * N * gn); return RMX_OK;
    case RMX_A_LOG_TRANSMAT:
    case RMX_A_CACHED_LOG_TRANSMAT:
    case RMX_A_JOINT_POSTERIOR_MARGINALS:
        for (int g = 0; g < gn; ++g) {
            int rc = materialize(m, g0 + g, id, (double*)dstv + (size_t)g * per);
            if (rc) return rc;
        }
        return RMX_OK;
    case RMX_A_L: src = b->l; shared = true; break;
    case RMX_A_X: src = b->x; shared = true; break;
    case RMX_A_Y: {
        std::vector<double> a(N), c(N);
        CK(cudaMemcpy(a.data(), b->y0, sizeof(double) * N, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(c.data(), b->y1, sizeof(double) * N, cudaMemcpyDeviceToHost));
        double* o = (double*)dstv;
        for (int g = 0; g < gn; ++g)
            for (int n = 0; n < N; ++n) { o[((size_t)g * N + n) * 2] = a[n]; o[((size_t)g * N + n) * 2 + 1] = c[n]; }
        return RMX_OK;
    }
    case RMX_A_STATE_PATH: {
        int64_t* o = (int64_t*)dstv;
        for (int g = 0; g < gn; ++g) {
            if (!b->path_valid[g0 + g]) return m->fail(RMX_ERR_VALUE, "no state path: call rmx_infer_cn first");
            for (int n = 0; n < N; ++n) o[(size_t)g * N + n] = b->h_path[(size_t)(g0 + g) * N + n];
        }
        return RMX_OK;
    }
    default: return m->fail(RMX_ERR_ARG, "unknown array id");
    }
    if (count > 0) {
        if (shared) {
            for (int g = 0; g < gn; ++g)
                CK(cudaMemcpyAsync((double*)dstv + (size_t)g * per, src, (size_t)per * sizeof(double), cudaMemcpyDeviceToHost,
                                   b->stream));
        } else if ((size_t)count * sizeof(double) >= ((size_t)4 << 20) && (size_t)count * sizeof(double) <= ((size_t)256 << 20)) {
            // a large (but not huge: the block stays in the pool) read into the caller's pageable, usually fresh array:
            // through a pooled page-locked block at
            // link speed, then copied out by a few host threads (the page faults of the destination run in parallel)
            const size_t bytes = (size_t)count * sizeof(double);
            void* stage = nullptr;
            if (g_pinned.get(bytes, &stage, false) != cudaSuccess) return m->fail(RMX_ERR_CUDA, "out of page-locked memory");
            cudaError_t e = cudaMemcpyAsync(stage, src + (size_t)g0 * per, bytes, cudaMemcpyDeviceToHost, b->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
            if (e != cudaSuccess) {
                g_pinned.put(bytes, stage);
                return m->fail(RMX_ERR_CUDA, cudaGetErrorString(e));
            }
            int nthr = (int)std::thread::hardware_concurrency();
            if (nthr > 8) nthr = 8;
            if (nthr < 1) nthr = 1;
            std::vector<std::thread> pool;
            for (int k = 0; k < nthr; ++k) {
                const size_t lo = bytes * k / nthr, hi = bytes * (k + 1) / nthr;
                pool.emplace_back([=] { memcpy((char*)dstv + lo, (const char*)stage + lo, hi - lo); });
            }
            for (auto& th : pool) th.join();
            g_pinned.put(bytes, stage);
            return RMX_OK;
        } else {
            CK(cudaMemcpyAsync(dstv, src + (size_t)g0 * per, (size_t)count * sizeof(double), cudaMemcpyDeviceToHost,
                               b->stream));
        }
        CK(cudaStreamSynchronize(b->stream));
    }
    return RMX_OK;
}

static double* scalar_slot(RmxParams& P, int id) {
    switch (id) {
    case RMX_S_NEGBIN_R_0: return &P.negbin_r[0];
    case RMX_S_NEGBIN_R_1: return &P.negbin_r[1];
    case RMX_S_NEGBIN_HDEL_MU: return &P.negbin_hdel_mu;
    case RMX_S_NEGBIN_HDEL_R_0: return &P.negbin_hdel_r[0];
    case RMX_S_NEGBIN_HDEL_R_1: return &P.negbin_hdel_r[1];
    case RMX_S_BETABIN_M_0: return &P.betabin_M[0];
    case RMX_S_BETABIN_M_1: return &P.betabin_M[1];
    case RMX_S_BETABIN_LOH_P: return &P.betabin_loh_p;
    case RMX_S_BETABIN_LOH_M_0: return &P.betabin_loh_M[0];
    case RMX_S_BETABIN_LOH_M_1: return &P.betabin_loh_M[1];
    case RMX_S_DIVERGENCE_WEIGHT: return &P.divergence_weight;
    default: return nullptr;
    }
}

// sets the scalar on every fit of the handle
int rmx_set_scalar(rmx_handle m, int id, double v) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    const int g0 = m->g0, gn = m->gn;
    switch (id) {
    case RMX_S_PRIOR_OUTLIER_TOTAL:
    case RMX_S_PRIOR_OUTLIER_ALLELE:
        if (b->G != gn && v != (id == RMX_S_PRIOR_OUTLIER_TOTAL ? b->prior_ot : b->prior_oa))
            return m->fail(RMX_ERR_UNSUPPORTED, "the outlier priors are shared by the fits of a batch: set them on the batch handle");
        (id == RMX_S_PRIOR_OUTLIER_TOTAL ? b->prior_ot : b->prior_oa) = v;
        return RMX_OK;
    case RMX_S_TRANSITION_PENALTY:
        if (v != b->pen) return m->fail(RMX_ERR_UNSUPPORTED, "transition_penalty is fixed when the model is created");
        return RMX_OK;
    case RMX_S_TRANSITION_MODEL:
        if (v != 0. && v != 1.) return m->fail(RMX_ERR_VALUE, "transition_model must be 0 or 1");
        for (int g = g0; g < g0 + gn; ++g) b->tm[g] = (int)v;
        return RMX_OK;
    case RMX_S_HMM_LOG_NORM_CONST:
        for (int g = g0; g < g0 + gn; ++g) b->logZ[g] = v;
        return RMX_OK;
    default: break;
    }
    if (!scalar_slot(b->P[g0], id)) return m->fail(RMX_ERR_ARG, "unknown scalar id");
    for (int g = g0; g < g0 + gn; ++g) {
        *scalar_slot(b->P[g], id) = v;
        if (id != RMX_S_DIVERGENCE_WEIGHT) b->ll_dirty[g] |= (char)planes_of(id);
    }
    b->params_dirty = true;
    b->ms_const_valid = false;
    return RMX_OK;
}

// one value per fit of the handle
int rmx_get_scalar(rmx_handle m, int id, double* v) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    for (int g = m->g0; g < m->g0 + m->gn; ++g, ++v) {
        switch (id) {
        case RMX_S_PRIOR_OUTLIER_TOTAL: *v = b->prior_ot; break;
        case RMX_S_PRIOR_OUTLIER_ALLELE: *v = b->prior_oa; break;
        case RMX_S_TRANSITION_PENALTY: *v = b->pen; break;
        case RMX_S_TRANSITION_MODEL: *v = (double)b->tm[g]; break;
        case RMX_S_HMM_LOG_NORM_CONST: *v = b->logZ[g]; break;
        default: {
            double* s = scalar_slot(b->P[g], id);
            if (!s) return m->fail(RMX_ERR_ARG, "unknown scalar id");
            *v = *s;
        }
        }
    }
    return RMX_OK;
}

// identical for every fit of a batch: filled once (not per fit)
int rmx_get_derived(rmx_handle m, int which, int64_t* dst, int64_t count) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    const RmxDims& d = b->d;
    int64_t want = which == 0 ? (int64_t)d.N * d.S * d.M : (int64_t)d.N * d.S;
    if (which < 0 || which > 3) return m->fail(RMX_ERR_ARG, "unknown derived table");
    if (want != count) return m->fail(RMX_ERR_VALUE, "array has the wrong number of elements");
    for (int n = 0; n < d.N; ++n) {
        int n0 = b->h_nrm[2 * n], n1 = b->h_nrm[2 * n + 1];
        for (int s = 0; s < d.S; ++s) {
            int a0 = n0, a1 = n1;
            for (int mm = 1; mm < d.M; ++mm) {
                a0 += b->h_tab[((size_t)s * d.M + mm) * 2];
                a1 += b->h_tab[((size_t)s * d.M + mm) * 2 + 1];
            }
            int64_t o = (int64_t)n * d.S + s;
            if (which == 0) {
                dst[o * d.M] = n0 + n1;
                for (int mm = 1; mm < d.M; ++mm)
                    dst[o * d.M + mm] = b->h_tab[((size_t)s * d.M + mm) * 2] + b->h_tab[((size_t)s * d.M + mm) * 2 + 1];
            } else if (which == 1) dst[o] = b->h_nsub[s];
            else if (which == 2) dst[o] = (a0 + a1) == 0;
            else dst[o] = (a0 == 0) || (a1 == 0);
        }
    }
    return RMX_OK;
}

static void take_logz(Model* m) {
    Batch* b = m->b.get();
    for (int g = m->g0; g < m->g0 + m->gn; ++g) b->logZ[g] = b->hp[(size_t)g * Batch::RED + 15];
}

int rmx_update(rmx_handle m, int id) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    b->ms_const_valid = false;      // the posteriors the M-step sums are weighted with change
    int rc = do_update(m, id);
    if (rc) return rc;
    rc = sync_check(m);
    if (rc) return rc;
    if (id == RMX_U_CN) take_logz(m);
    return RMX_OK;
}

int rmx_breakpoint_logq(rmx_handle m, double* out) {
    if (!m || !out) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const size_t kb = (size_t)b->d.K * b->d.B;
    if (kb == 0) return RMX_OK;
    CK(cudaMemsetAsync(b->logq + (size_t)m->g0 * kb, 0, sizeof(double) * kb * m->gn, b->stream));
    int rc = do_update_p_breakpoint(m, 1);
    if (rc) return rc;
    CK(cudaMemcpyAsync(out, b->logq + (size_t)m->g0 * kb, sizeof(double) * kb * m->gn, cudaMemcpyDeviceToHost, b->stream));
    return sync_check(m);
}

int rmx_breakpoint_apply_logq(rmx_handle m, const double* logq) {
    if (!m || !logq) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const size_t kb = (size_t)b->d.K * b->d.B;
    if (kb > 0) {
        CK(cudaMemcpyAsync(b->logq + (size_t)m->g0 * kb, logq, sizeof(double) * kb * m->gn, cudaMemcpyHostToDevice, b->stream));
        CK(cudaStreamSynchronize(b->stream));   // the caller's buffer is free when this returns
    }
    int rc = do_update_p_breakpoint(m, 2);
    if (rc) return rc;
    return sync_check(m);
}

// ---- communicator of a chromosome-sharded fit
int rmx_comm_load(const char* nccl_library) {
    if (!g_nccl.load(nccl_library)) { g_create_error = g_nccl.error; return RMX_ERR_CUDA; }
    return RMX_OK;
}
int rmx_comm_unique_id(void* unique_id_128) {
    if (!unique_id_128) return RMX_ERR_ARG;
    if (!g_nccl.load(nullptr)) { g_create_error = g_nccl.error; return RMX_ERR_CUDA; }
    rmx_ncclUniqueId id;
    const int r = g_nccl.GetUniqueId(&id);
    if (r != 0) { g_create_error = "ncclGetUniqueId failed"; return RMX_ERR_CUDA; }
    memcpy(unique_id_128, &id, sizeof id);
    return RMX_OK;
}
int rmx_comm_init(rmx_handle m, const void* unique_id_128, int rank, int nranks) {
    if (!m || !unique_id_128 || nranks < 1 || rank < 0 || rank >= nranks) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    if (!g_nccl.load(nullptr)) return m->fail(RMX_ERR_CUDA, g_nccl.error);
    if (b->comm) return m->fail(RMX_ERR_ARG, "a communicator is already attached to this model");
    if (!b->is_shard) return m->fail(RMX_ERR_ARG, "rmx_comm_init needs a model created with RMX_DESC_SHARD");
    rmx_ncclUniqueId id;
    memcpy(&id, unique_id_128, sizeof id);
    const int r = g_nccl.CommInitRank(&b->comm, nranks, id, rank);
    if (r != 0) {
        b->comm = nullptr;
        return m->fail(RMX_ERR_CUDA, std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error"));
    }
    b->comm_rank = rank;
    b->comm_size = nranks;
    return RMX_OK;
}
int rmx_comm_destroy(rmx_handle m) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    if (b->comm) {
        cudaStreamSynchronize(b->stream);
        g_nccl.CommDestroy(b->comm);
        b->comm = nullptr;
    }
    return RMX_OK;
}

int rmx_variational_update(rmx_handle m, int n_iter) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    b->ms_const_valid = false;      // the posteriors the M-step sums are weighted with change
    static const int order[5] = {RMX_U_ALLELE_SWAP, RMX_U_CN, RMX_U_BREAKPOINT, RMX_U_OUTLIER_TOTAL,
                                 RMX_U_OUTLIER_ALLELE};
    static const bool no_fuse = getenv("RMX_NO_FUSED_ROUND") != nullptr;     // A/B measurements only
    bool fuse = !no_fuse && rmx_round_ok(b->d);
    if (fuse) {
        // swap -> cn -> breakpoint -> outlier_total -> outlier_allele per round (cn_model.py:444-460); the posterior
        // marginals, both outlier updates and the next round's swap update and frame log-probability are one
        // kernel (k_round)
        for (int it = 0; it < n_iter; ++it) {
            int rc = RMX_OK;
            if (it == 0) {
                // update_p_allele_swap and the frame log-probability of the first round: the same kernel in its
                // head mode (likelihood sums of the stored posterior, swap update, frame) instead of three passes
                rc = ensure_ll(m, m->g0, m->gn);
                if (rc) return rc;
                cudaEvent_t th_ = b->tbegin(RMX_K_ROUND);
                rmx_launch_round(b->stream, tables(b), b->d, b->dP, m->g0, m->gn, b->LL, b->ahat, b->bhat, b->post, b->Asum,
                                 b->q_ot, b->q_oa, b->q_sw, b->prior_ot, b->prior_oa, 1, 1, b->F, b->fexp, b->mF, b->dflags);
                b->tend(th_);
                b->count();
            }
            rc = do_update_p_cn(m, true, true);
            if (rc) return rc;
            rc = do_update_p_breakpoint(m);
            if (rc) return rc;
            cudaEvent_t te_ = b->tbegin(RMX_K_ROUND);
            rmx_launch_round(b->stream, tables(b), b->d, b->dP, m->g0, m->gn, b->LL, b->ahat, b->bhat, b->post, b->Asum,
                             b->q_ot, b->q_oa, b->q_sw, b->prior_ot, b->prior_oa, it + 1 < n_iter ? 1 : 0, 0, b->F,
                             b->fexp, b->mF, b->dflags);
            b->tend(te_);
            b->count();
            for (int g = m->g0; g < m->g0 + m->gn; ++g) b->asum_dirty[g] = 0;
        }
    } else {
        for (int it = 0; it < n_iter; ++it)
            for (int k = 0; k < 5; ++k) {
                int rc = do_update(m, order[k]);
                if (rc) return rc;
            }
    }
    int rc = sync_check(m);
    if (rc) return rc;
    if (n_iter > 0) take_logz(m);
    return RMX_OK;
}

// launches the ELBO reductions of the handle's fits; results land in hp[g][0..9] after sync.
// full = false: only the ELBO is wanted, so of sum joint*cached_T - sum joint*T only the breakend rows are
// evaluated (everywhere else the two transition matrices are identical and the terms cancel exactly;
// this needs both snapshots to have been taken under the same transition_model).
static int elbo_terms(Model* m, bool full) {
    Batch* b = m->b.get();
    const int g0 = m->g0, gn = m->gn;
    int rc = ensure_asum(m, g0, gn);
    if (rc) return rc;
    rc = push_params(m);
    if (rc) return rc;
    const RmxDims& d = b->d;
    RmxTables T = tables(b);
    const int RED = Batch::RED;
    { cudaEvent_t te_ = b->tbegin(RMX_K_ELBO_ROWS); rmx_launch_elbo_rows(b->stream, T, d, b->dP, g0, gn, b->Asum, b->post, b->F, b->q_ot, b->q_oa, b->q_sw, b->prior_ot,
                         b->prior_oa, b->rowv); b->tend(te_); }
    rmx_launch_reduce_cols(b->stream, b->rowv, d.N, 7, g0, gn, (int64_t)8 * d.N, b->red_out, RED);
    b->count(2);
    // entropy of p_breakpoint
    double* pl = b->rowv + (size_t)b->G * 8 * d.N;
    const int64_t kb = (int64_t)d.K * d.B;
    if (kb > 0) {
        rmx_launch_plogp(b->stream, b->p_brk, kb, g0, gn, pl, kb);
        rmx_launch_reduce_cols(b->stream, pl, kb, 1, g0, gn, kb, b->red_out + 7, RED);
        b->count(2);
    }
    // transition terms -> red_out[g][8] (energy side) and [9] (entropy side)
    for (int g = g0; g < g0 + gn; ++g)
        CK(cudaMemsetAsync(b->red_out + (size_t)g * RED + (kb > 0 ? 8 : 7), 0, (kb > 0 ? 2 : 3) * sizeof(double), b->stream));
    if (d.N > 1) {
        std::vector<int> key(b->G, 0);
        for (int g = g0; g < g0 + gn; ++g) key[g] = (b->cn_updated[g] ? 4 : 0) | (b->tm_used[g] << 1) | b->tm_cached[g];
        for_runs(g0, gn, key, nullptr, [&](int g, int n) {
            if (rc) return;
            rc = ensure_mats(m, b->tm_cached[g]);
            if (!rc) rc = ensure_mats(m, b->tm_used[g]);
            if (rc) return;
            RmxTransSpec used = spec_of(b, false, b->tm_used[g]), cached = spec_of(b, true, b->tm_cached[g]);
            const bool upd = b->cn_updated[g] != 0;
            if (upd && !full && b->tm_used[g] == b->tm_cached[g]) {
                if (b->R > 0) {
                    bool hv = true;
                    for (int k = g; k < g + n; ++k) hv = hv && b->hist_valid[k];
                    cudaEvent_t te_ = b->tbegin(RMX_K_PAIRSUM);
                    if (hv)     // hist = histograms of the current joint (update_p_cn made them for update_p_breakpoint)
                        rmx_launch_pairsum_hist(b->stream, d, g, n, b->R, rowsof(b), b->hist, cached, used, b->rmd, b->rowsum, d.N);
                    else
                        rmx_launch_pairsum(b->stream, T, d, g, n, rowsof(b), used, cached, used, 1, b->rmd, b->brk_rows, b->R, 0,
                                           b->ahat, b->bhat, b->fexp, b->rowsum, d.N);
                    b->tend(te_);
                    rmx_launch_reduce_cols(b->stream, b->rowsum, b->R, 1, g, n, d.N, b->red_out + 8, RED);
                    b->count(2);
                }
            } else {
                { cudaEvent_t te_ = b->tbegin(RMX_K_PAIRSUM); rmx_launch_pairsum(b->stream, T, d, g, n, rowsof(b), used, cached, cached, 0, b->rmd, nullptr, d.N - 1,
                                   upd ? 0 : 1, b->ahat, b->bhat, b->fexp, b->rowsum, d.N); b->tend(te_); }
                rmx_launch_reduce_cols(b->stream, b->rowsum, d.N - 1, 1, g, n, d.N, b->red_out + 8, RED);
                b->count(2);
                if (upd) {
                    { cudaEvent_t te_ = b->tbegin(RMX_K_PAIRSUM); rmx_launch_pairsum(b->stream, T, d, g, n, rowsof(b), used, used, used, 0, b->rmd, nullptr, d.N - 1, 0,
                                       b->ahat, b->bhat, b->fexp, b->rowsum, d.N); b->tend(te_); }
                    rmx_launch_reduce_cols(b->stream, b->rowsum, d.N - 1, 1, g, n, d.N, b->red_out + 9, RED);
                    b->count(2);
                }
            }
        });
        if (rc) return rc;
    }
    if (b->comm) {
        // the entropy of p_breakpoint is the one term every shard holds in full: counted on rank 0 only
        if (b->comm_rank != 0)
            for (int g = g0; g < g0 + gn; ++g) CK(cudaMemsetAsync(b->red_out + (size_t)g * RED + 7, 0, sizeof(double), b->stream));
        rc = shard_reduce_red(m, 0, 10);
        if (rc) return rc;
    }
    CK(cudaMemcpyAsync(b->hp + (size_t)g0 * RED, b->red_out + (size_t)g0 * RED, sizeof(double) * RED * gn,
                       cudaMemcpyDeviceToHost, b->stream));
    return sync_check(m);
}

static double energy_of(const Batch* b, int g) {
    const double* r = b->hp + (size_t)g * Batch::RED;
    // prior, total (+ outlier prior), allele (+ outlier prior), transitions  (bpmodel.pyx:1068-1115)
    return r[0] + (r[1] + r[3]) + (r[2] + r[4]) + r[8];
}
static double entropy_of(const Batch* b, int g) {
    const double* r = b->hp + (size_t)g * Batch::RED;
    // bpmodel.pyx:1050-1056
    return -b->logZ[g] + r[5] + r[9] + r[7] + r[6];
}

static int elbo_like(rmx_handle m, double* out, int what) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    // the logZ copy of an update may still be in hp: elbo_terms overwrites hp[15] with the same device value
    int rc = elbo_terms(m, what != 2);
    if (rc) return rc;
    for (int g = m->g0; g < m->g0 + m->gn; ++g)
        out[g - m->g0] = what == 0 ? energy_of(b, g) : (what == 1 ? entropy_of(b, g) : energy_of(b, g) - entropy_of(b, g));
    return RMX_OK;
}
int rmx_variational_energy(rmx_handle m, double* out) { return elbo_like(m, out, 0); }
int rmx_variational_entropy(rmx_handle m, double* out) { return elbo_like(m, out, 1); }
int rmx_elbo(rmx_handle m, double* out) { return elbo_like(m, out, 2); }

// sample masks (gn, N) -> per-fit row lists on the device; *rows = nullptr when every fit takes all segments
static int stage_rows(Model* m, const int64_t* sample, int* Rmax, const int** rows) {
    Batch* b = m->b.get();
    const int N = b->d.N, g0 = m->g0, gn = m->gn;
    int rmax = 0;
    bool all = true;
    for (int g = 0; g < gn; ++g) {
        int cnt = 0;
        int* hr = b->hrows + (size_t)(g0 + g) * N;
        const int64_t* s = sample + (size_t)g * N;
        for (int n = 0; n < N; ++n)
            if (s[n] != 0) hr[cnt++] = n;
        b->hRn[g0 + g] = cnt;
        all &= cnt == N;
        rmax = cnt > rmax ? cnt : rmax;
    }
    *Rmax = rmax;
    if (all) {
        *rows = nullptr;
        return RMX_OK;
    }
    CK(cudaMemcpyAsync(b->d_rows + (size_t)g0 * N, b->hrows + (size_t)g0 * N, (size_t)gn * N * sizeof(int),
                       cudaMemcpyHostToDevice, b->stream));
    CK(cudaMemcpyAsync(b->d_Rn + g0, b->hRn + g0, (size_t)gn * sizeof(int), cudaMemcpyHostToDevice, b->stream));
    *rows = b->d_rows;
    return RMX_OK;
}

// E[LL] of the handle's fits on the staged rows -> red_out[g][slot]
static int launch_ell(Model* m, const int* rows, int Rmax, int slot) {
    Batch* b = m->b.get();
    int rc = push_params(m);
    if (rc) return rc;
    RmxTables T = tables(b);
    const int g0 = m->g0, gn = m->gn;
    rmx_launch_segconst(b->stream, T, b->d, b->dP, g0, gn, rows, b->d_Rn, Rmax, b->segc_s);
    const int nb = rmx_ell_blocks(b->d, Rmax);
    { cudaEvent_t te_ = b->tbegin(RMX_K_ELL); rmx_launch_ell(b->stream, T, b->d, b->dP, g0, gn, rows, b->d_Rn, Rmax, b->segc_s, b->post, b->q_ot, b->q_oa, b->q_sw,
                   b->part, b->part_stride, b->dflags); b->tend(te_); }
    rmx_launch_reduce_cols(b->stream, b->part, nb, 1, g0, gn, b->part_stride, b->red_out + slot, Batch::RED);
    b->count(3);
    return RMX_OK;
}

static int fetch_red(Model* m) {
    Batch* b = m->b.get();
    {
        int rc = shard_reduce_red(m, 12, Batch::RED - 12);      // constants, E[LL] values, gradient
        if (rc) return rc;
    }
    CK(cudaMemcpyAsync(b->hp + (size_t)m->g0 * Batch::RED, b->red_out + (size_t)m->g0 * Batch::RED,
                       sizeof(double) * Batch::RED * m->gn, cudaMemcpyDeviceToHost, b->stream));
    return sync_check(m);
}

int rmx_expected_log_likelihood(rmx_handle m, const int64_t* sample, double* out) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    int Rmax;
    const int* rows;
    int rc = stage_rows(m, sample, &Rmax, &rows);
    if (rc) return rc;
    for (int g = 0; g < m->gn; ++g) out[g] = 0.;
    if (Rmax == 0 && !b->comm) return RMX_OK;
    rc = shard_zero_red(m, 16, 1);
    if (rc) return rc;
    if (Rmax > 0) {
        rc = launch_ell(m, rows, Rmax, 16);
        if (rc) return rc;
    }
    rc = fetch_red(m);
    if (rc) return rc;
    for (int g = 0; g < m->gn; ++g)
        out[g] = (b->hRn[m->g0 + g] > 0 || b->comm) ? b->hp[(size_t)(m->g0 + g) * Batch::RED + 16] : 0.;
    return RMX_OK;
}

int rmx_expected_log_likelihood_grid(rmx_handle m, const int64_t* sample, int scalar_id, const double* values,
                                     int n_values, double* out) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    if (n_values < 0 || n_values > 32) return m->fail(RMX_ERR_ARG, "n_values must be in [0, 32]");
    if (scalar_id < RMX_S_NEGBIN_R_0 || scalar_id > RMX_S_BETABIN_LOH_M_1)
        return m->fail(RMX_ERR_ARG, "scalar_id must name a likelihood parameter");
    int Rmax;
    const int* rows;
    int rc = stage_rows(m, sample, &Rmax, &rows);
    if (rc) return rc;
    // values is (gn, n_values): fit g evaluates values[g][v]; launches are stream ordered, so the per-launch
    // scratch (segc_s, part) is reused safely
    for (int v = 0; v < n_values; ++v) {
        for (int g = 0; g < m->gn; ++g) {
            *scalar_slot(b->P[m->g0 + g], scalar_id) = values[(size_t)g * n_values + v];
            b->ll_dirty[m->g0 + g] |= (char)planes_of(scalar_id);
        }
        b->params_dirty = true;
        if (Rmax == 0) continue;
        rc = launch_ell(m, rows, Rmax, 16 + v);
        if (rc) return rc;
    }
    for (int e = 0; e < m->gn * n_values; ++e) out[e] = 0.;
    if (Rmax == 0 || n_values == 0) return RMX_OK;
    rc = fetch_red(m);
    if (rc) return rc;
    for (int g = 0; g < m->gn; ++g)
        for (int v = 0; v < n_values; ++v)
            out[(size_t)g * n_values + v] = b->hRn[m->g0 + g] > 0 ? b->hp[(size_t)(m->g0 + g) * Batch::RED + 16 + v] : 0.;
    return RMX_OK;
}

int rmx_expected_log_likelihood_partial_h(rmx_handle m, const int64_t* sample, double* out) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const int M = b->d.M;
    int Rmax;
    const int* rows;
    int rc = stage_rows(m, sample, &Rmax, &rows);
    if (rc) return rc;
    for (int e = 0; e < m->gn * M; ++e) out[e] = 0.;
    if (Rmax == 0 && !b->comm) return RMX_OK;
    rc = push_params(m);
    if (rc) return rc;
    rc = shard_zero_red(m, 48, RMX_MAXM);
    if (rc) return rc;
    if (Rmax > 0) {
        const int nb = rmx_ell_blocks(b->d, Rmax);
        rmx_launch_ell_grad(b->stream, tables(b), b->d, b->dP, m->g0, m->gn, rows, b->d_Rn, Rmax, b->post, b->q_ot, b->q_oa,
                            b->q_sw, b->part, b->part_stride, b->dflags);
        rmx_launch_reduce_cols(b->stream, b->part, nb, RMX_MAXM, m->g0, m->gn, b->part_stride, b->red_out + 48, Batch::RED);
        b->count(2);
    }
    rc = fetch_red(m);
    if (rc) return rc;
    for (int g = 0; g < m->gn; ++g)
        for (int k = 0; k < M; ++k)
            out[(size_t)g * M + k] = (b->hRn[m->g0 + g] > 0 || b->comm) ? b->hp[(size_t)(m->g0 + g) * Batch::RED + 48 + k] : 0.;
    return RMX_OK;
}

// ---------------------------------------------------------------------------------------------
// batched M-step
// ---------------------------------------------------------------------------------------------
int rmx_mstep_stage_sample(rmx_handle m, const int64_t* sample) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const int N = b->d.N, g0 = m->g0, gn = m->gn;
    if (g0 != 0 || gn != b->G) return m->fail(RMX_ERR_ARG, "the batched M-step works on the handle of the whole batch");
    b->ms_const_valid = false;
    if (!sample) {
        b->ms_all = true;
        b->ms_rmax = N;
        for (int g = 0; g < gn; ++g) b->ms_count[g] = N;
        return RMX_OK;
    }
    int rmax = 0;
    for (int g = 0; g < gn; ++g) {
        int cnt = 0;
        int* hr = b->hrows + (size_t)g * N;
        const int64_t* s = sample + (size_t)g * N;
        for (int n = 0; n < N; ++n)
            if (s[n] != 0) hr[cnt++] = n;
        b->ms_count[g] = cnt;
        rmax = cnt > rmax ? cnt : rmax;
    }
    b->ms_all = false;
    b->ms_rmax = rmax;
    // only the filled head of every fit's list travels
    for (int g = 0; g < gn; ++g)
        if (b->ms_count[g] > 0)
            CK(cudaMemcpyAsync(b->ms_rows + (size_t)g * N, b->hrows + (size_t)g * N, (size_t)b->ms_count[g] * sizeof(int),
                               cudaMemcpyHostToDevice, b->stream));
    for (int g = 0; g < gn; ++g) b->hRn[g] = b->ms_count[g];
    CK(cudaMemcpyAsync(b->ms_Rn_all, b->hRn, (size_t)gn * sizeof(int), cudaMemcpyHostToDevice, b->stream));
    if (rmx_ell_sample_ok(rmax)) {      // the parameter-independent constants of these rows, once
        rmx_launch_ell_rowcache(b->stream, tables(b), b->d, 0, gn, b->ms_rows, b->ms_Rn_all, rmax, b->ms_rowc);
        b->count();
    }
    CK(cudaStreamSynchronize(b->stream));   // hrows is reused by the next staging / decode
    return RMX_OK;
}

int rmx_mstep_stage_rows(rmx_handle m, const int32_t* rows, const int32_t* counts, int max_rows) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const int N = b->d.N, G = b->G;
    if (m->g0 != 0 || m->gn != G) return m->fail(RMX_ERR_ARG, "the batched M-step works on the handle of the whole batch");
    if (!rows || !counts || max_rows < 0 || max_rows > N) return m->fail(RMX_ERR_ARG, "bad row lists");
    b->ms_const_valid = false;
    int rmax = 0;
    for (int g = 0; g < G; ++g) {
        const int cnt = counts[g];
        if (cnt < 0 || cnt > max_rows) return m->fail(RMX_ERR_ARG, "counts must lie in [0, max_rows]");
        for (int r = 0; r < cnt; ++r) {
            const int n = rows[(size_t)g * max_rows + r];
            if (n < 0 || n >= N) return m->fail(RMX_ERR_VALUE, "segment index out of range in the sample");
            b->hrows[(size_t)g * N + r] = n;
        }
        b->ms_count[g] = cnt;
        rmax = cnt > rmax ? cnt : rmax;
    }
    b->ms_all = false;
    b->ms_rmax = rmax;
    // one strided copy for all fits (the caller's (G, max_rows) block into the (G, N) row lists; entries beyond a fit's
    // count are not read); the stream is synchronised before this call returns
    if (rmax > 0)
        CK(cudaMemcpy2DAsync(b->ms_rows, (size_t)N * sizeof(int), rows, (size_t)max_rows * sizeof(int32_t),
                             (size_t)rmax * sizeof(int), (size_t)G, cudaMemcpyHostToDevice, b->stream));
    for (int g = 0; g < G; ++g) b->hRn[g] = b->ms_count[g];
    CK(cudaMemcpyAsync(b->ms_Rn_all, b->hRn, (size_t)G * sizeof(int), cudaMemcpyHostToDevice, b->stream));
    if (rmx_ell_sample_ok(rmax)) {
        rmx_launch_ell_rowcache(b->stream, tables(b), b->d, 0, G, b->ms_rows, b->ms_Rn_all, rmax, b->ms_rowc);
        b->count();
    }
    CK(cudaStreamSynchronize(b->stream));
    return RMX_OK;
}

int rmx_mstep_evaluate(rmx_handle m, const uint8_t* active, const double* h_values, int scalar_id,
                       const double* values, int n_values, double* ell, double* grad) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const int M = b->d.M, G = b->G;
    if (m->g0 != 0 || m->gn != G) return m->fail(RMX_ERR_ARG, "the batched M-step works on the handle of the whole batch");
    if (b->ms_count[0] < 0) return m->fail(RMX_ERR_ARG, "rmx_mstep_stage_sample has not been called");
    if (n_values < 1 || n_values > 32) return m->fail(RMX_ERR_ARG, "n_values must be in [1, 32]");
    if (scalar_id >= 0 && (scalar_id > RMX_S_BETABIN_LOH_M_1 || !values))
        return m->fail(RMX_ERR_ARG, "scalar_id must name a likelihood parameter and come with values");
    if (scalar_id < 0 && n_values != 1) return m->fail(RMX_ERR_ARG, "n_values must be 1 without a parameter");
    // likelihood terms the parameter enters (bit u: total-read term u; bit 2 + v: the allele terms of v)
    int planes = planes_of(scalar_id);
    static const bool no_planes = getenv("RMX_MSTEP_ALL_TERMS") != nullptr;      // A/B measurements only
    if (b->ms_all || grad || h_values || no_planes) planes = 15;
    if (h_values || scalar_id != b->ms_const_id) b->ms_const_valid = false;
    bool any = false;
    for (int g = 0; g < G; ++g) {
        const bool on = !active || active[g];
        b->hRn[g] = on ? b->ms_count[g] : 0;
        any |= on && b->ms_count[g] > 0;
        if (on && h_values) {
            for (int k = 0; k < M; ++k) b->P[g].h[k] = h_values[(size_t)g * M + k];
            b->ll_dirty[g] = 15;
            b->params_dirty = true;
        }
    }
    for (int e = 0; e < G * n_values; ++e) ell[e] = 0.;
    if (grad) for (int e = 0; e < G * M; ++e) grad[e] = 0.;
    const int Rmax = b->ms_rmax;
    const int* rows = b->ms_all ? nullptr : b->ms_rows;
    // a shard of a multi-GPU fit takes part in every reduction, rows or not
    const bool launch = any && Rmax > 0;
    const bool work = launch || b->comm != nullptr;
    if (b->comm) {
        int rcz = shard_zero_red(m, 12, Batch::RED - 12);
        if (rcz) return rcz;
    }
    if (launch)   // pageable-style small copy from the pinned block: staged before the call returns to the stream order
        CK(cudaMemcpyAsync(b->ms_Rn, b->hRn, (size_t)G * sizeof(int), cudaMemcpyHostToDevice, b->stream));
    RmxTables T = tables(b);
    const int nb = launch ? rmx_ell_blocks(b->d, Rmax) : 0;
    bool fetch_const = false;
    if (work && planes != 15 && !b->ms_const_valid) fetch_const = true;
    // samples (<= 256 rows per fit, one GPU): the whole objective is one launch per value and the partial sums travel
    // to the host in one copy; otherwise constants + objective + column reduction as three launches per value
    const int NCH = rmx_ell_sample_chunks();
    const bool fused = launch && !b->ms_all && !b->comm && rmx_ell_sample_ok(Rmax);
    if (fused) {
        if (fetch_const) {
            int rc = push_params(m);
            if (rc) return rc;
            rmx_launch_ell_sample(b->stream, T, b->d, b->dP, 0, G, rows, b->ms_Rn_all, b->post, b->q_ot, b->q_oa, b->q_sw,
                                  b->ms_rowc, b->ms_out + (size_t)32 * G * NCH, NCH, b->dflags, 15 ^ planes);
            b->count();
        }
        for (int v = 0; v < n_values; ++v) {
            if (scalar_id >= 0) {
                for (int g = 0; g < G; ++g)
                    if (!active || active[g]) {
                        *scalar_slot(b->P[g], scalar_id) = values[(size_t)g * n_values + v];
                        b->ll_dirty[g] |= (char)planes_of(scalar_id);
                    }
                b->params_dirty = true;
            }
            int rc = push_params(m);
            if (rc) return rc;
            cudaEvent_t te_ = b->tbegin(RMX_K_ELL);
            rmx_launch_ell_sample(b->stream, T, b->d, b->dP, 0, G, rows, b->ms_Rn, b->post, b->q_ot, b->q_oa, b->q_sw,
                                  b->ms_rowc, b->ms_out + (size_t)v * G * NCH, NCH, b->dflags, planes);
            b->tend(te_);
            b->count();
        }
        // only what this call produced travels: the first n_values slots, and the constants' slot when it is new
        CK(cudaMemcpyAsync(b->hms, b->ms_out, (size_t)n_values * G * NCH * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
        if (fetch_const)
            CK(cudaMemcpyAsync(b->hms + (size_t)32 * G * NCH, b->ms_out + (size_t)32 * G * NCH, (size_t)G * NCH * sizeof(double),
                               cudaMemcpyDeviceToHost, b->stream));
    } else {
    if (launch && planes != 15 && !b->ms_const_valid) {
        // the terms the parameter does not enter, for EVERY fit with staged rows (fits join and leave the active set
        // from call to call), with the parameters as they stand
        int rc = push_params(m);
        if (rc) return rc;
        rmx_launch_segconst(b->stream, T, b->d, b->dP, 0, G, rows, b->ms_Rn_all, Rmax, b->segc_s);
        rmx_launch_ell(b->stream, T, b->d, b->dP, 0, G, rows, b->ms_Rn_all, Rmax, b->segc_s, b->post, b->q_ot, b->q_oa, b->q_sw,
                       b->part, b->part_stride, b->dflags, 15 ^ planes);
        rmx_launch_reduce_cols(b->stream, b->part, nb, 1, 0, G, b->part_stride, b->red_out + 12, Batch::RED);
        b->count(3);
    }
    for (int v = 0; v < n_values; ++v) {
        if (scalar_id >= 0) {
            for (int g = 0; g < G; ++g)
                if (!active || active[g]) {
                    *scalar_slot(b->P[g], scalar_id) = values[(size_t)g * n_values + v];
                    b->ll_dirty[g] |= (char)planes_of(scalar_id);
                }
            b->params_dirty = true;
        }
        if (!launch) continue;
        if (b->ms_all) {
            // every segment: the likelihood table of the fits whose parameters changed (the kernel of the E-step,
            // which the next E-step then finds ready), its posterior-weighted sums, and one value per segment
            // (fits outside the active set keep their stale table until something needs it)
            int rc = ensure_asum(m, 0, G, active);
            if (rc) return rc;
            double* rowe = b->rowv + 7 * (size_t)b->d.N;
            rmx_launch_ell_rows_from_sums(b->stream, b->d, 0, G, b->Asum, b->q_ot, b->q_oa, b->q_sw, rowe, (int64_t)8 * b->d.N);
            rmx_launch_reduce_cols(b->stream, rowe, b->d.N, 1, 0, G, (int64_t)8 * b->d.N, b->red_out + 16 + v, Batch::RED);
            b->count(2);
            continue;
        }
        int rc = push_params(m);
        if (rc) return rc;
        rmx_launch_segconst(b->stream, T, b->d, b->dP, 0, G, rows, b->ms_Rn, Rmax, b->segc_s);
        { cudaEvent_t te_ = b->tbegin(RMX_K_ELL);
          rmx_launch_ell(b->stream, T, b->d, b->dP, 0, G, rows, b->ms_Rn, Rmax, b->segc_s, b->post, b->q_ot, b->q_oa, b->q_sw,
                         b->part, b->part_stride, b->dflags, planes);
          b->tend(te_); }
        rmx_launch_reduce_cols(b->stream, b->part, nb, 1, 0, G, b->part_stride, b->red_out + 16 + v, Batch::RED);
        b->count(3);
    }
    }
    if (launch && grad) {
        rmx_launch_ell_grad(b->stream, T, b->d, b->dP, 0, G, rows, b->ms_Rn, Rmax, b->post, b->q_ot, b->q_oa, b->q_sw, b->part,
                            b->part_stride, b->dflags);
        rmx_launch_reduce_cols(b->stream, b->part, nb, RMX_MAXM, 0, G, b->part_stride, b->red_out + 48, Batch::RED);
        b->count(2);
    }
    if (!work) return RMX_OK;
    int rc = (fused && !grad) ? sync_check(m) : fetch_red(m);
    if (rc) return rc;
    auto chunk_sum = [&](int g, int slot) {      // the partial sums of a fit in chunk order
        const double* q = b->hms + ((size_t)slot * G + g) * NCH;
        double acc = 0.;
        for (int c = 0; c < NCH; ++c) acc += q[c];
        return acc;
    };
    if (fetch_const) {
        for (int g = 0; g < G; ++g)
            b->ms_const[g] = fused ? (b->ms_count[g] > 0 ? chunk_sum(g, 32) : 0.)
                                   : ((b->ms_count[g] > 0 || b->comm) ? b->hp[(size_t)g * Batch::RED + 12] : 0.);
        b->ms_const_id = scalar_id;
        b->ms_const_valid = true;
    }
    for (int g = 0; g < G; ++g) {
        if (b->hRn[g] <= 0 && !(b->comm && (!active || active[g]))) continue;
        const double* r = b->hp + (size_t)g * Batch::RED;
        const double c0 = planes != 15 ? b->ms_const[g] : 0.;
        for (int v = 0; v < n_values; ++v) ell[(size_t)g * n_values + v] = (fused ? chunk_sum(g, v) : r[16 + v]) + c0;
        if (grad) for (int k = 0; k < M; ++k) grad[(size_t)g * M + k] = r[48 + k];
    }
    return RMX_OK;
}

// ---- update_param's optimiser for all fits of a grid, driven from inside the library --------------------------------
// scipy.optimize.fmin (Nelder-Mead, one variable, default options) as a state machine: next() is the point to
// evaluate, feed() takes its value.  Statement for statement remixt_b200/mstep.py:neldermead_1d, which
// tests/test_mstep_cpu.py pins to scipy evaluation by evaluation.
namespace {
struct NelderMead1 {
    enum Phase { INIT0, INIT1, REFLECT, EXPAND, CONTRACT_OUT, CONTRACT_IN, SHRINK, DONE };
    static constexpr int MAXFUN = 200, MAXITER = 200;
    static constexpr double XATOL = 1e-4, FATOL = 1e-4;
    double s0 = 0., s1 = 0., f0 = INFINITY, f1 = INFINITY, x = 0., xr = 0., fxr = 0.;
    int calls = 0, iterations = 0;
    Phase phase = DONE;

    static bool second_first(double a, double b) {      // np.argsort([a, b]) == [1, 0]
        if (b != b) return false;
        if (a != a) return true;
        return b < a;
    }
    void sort() { if (second_first(f0, f1)) { std::swap(s0, s1); std::swap(f0, f1); } }
    void start(double x0) {
        s0 = x0;
        s1 = x0 != 0. ? 1.05 * x0 : 0.00025;
        f0 = f1 = INFINITY;
        calls = 0;
        iterations = 0;
        request(INIT0, s0);
    }
    bool done() const { return phase == DONE; }
    double next() const { return x; }
    // an evaluation is due: out of budget ends the search the way the `aborted` path of the generator does
    void request(Phase p, double at) {
        if (calls >= MAXFUN) {
            if (p != INIT0 && p != INIT1) sort();
            phase = DONE;
            return;
        }
        ++calls;
        phase = p;
        x = at;
    }
    void begin_iteration() {
        if (!(calls < MAXFUN && iterations < MAXITER)) { phase = DONE; return; }
        if (fabs(s1 - s0) <= XATOL && fabs(f0 - f1) <= FATOL) { phase = DONE; return; }
        xr = 2 * s0 - 1 * s1;
        request(REFLECT, xr);
    }
    void end_iteration() {
        ++iterations;
        sort();
        begin_iteration();
    }
    void feed(double f) {
        switch (phase) {
        case INIT0: f0 = f; request(INIT1, s1); break;
        case INIT1:
            f1 = f;
            sort();
            iterations = 1;
            begin_iteration();
            break;
        case REFLECT:
            fxr = f;
            if (fxr < f0) request(EXPAND, 3 * s0 - 2 * s1);
            else if (fxr < f1) request(CONTRACT_OUT, 1.5 * s0 - 0.5 * s1);
            else request(CONTRACT_IN, 0.5 * s0 + 0.5 * s1);
            break;
        case EXPAND:
            if (f < fxr) { s1 = x; f1 = f; } else { s1 = xr; f1 = fxr; }
            end_iteration();
            break;
        case CONTRACT_OUT:
            if (f <= fxr) { s1 = x; f1 = f; end_iteration(); }
            else { s1 = s0 + 0.5 * (s1 - s0); request(SHRINK, s1); }
            break;
        case CONTRACT_IN:
            if (f < f1) { s1 = x; f1 = f; end_iteration(); }
            else { s1 = s0 + 0.5 * (s1 - s0); request(SHRINK, s1); }
            break;
        case SHRINK: f1 = f; end_iteration(); break;
        case DONE: break;
        }
    }
};
}  // namespace

int rmx_mstep_param_search(rmx_handle m, const uint8_t* active, int scalar_id, double lo, double hi, const double* grid,
                           int n_grid, double* result, int32_t* evals, int64_t* rounds, double* trace, int trace_cap) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    const int G = b->G;
    if (m->g0 != 0 || m->gn != G) return m->fail(RMX_ERR_ARG, "the batched M-step works on the handle of the whole batch");
    if (n_grid < 1 || n_grid > 32 || !grid || !result) return m->fail(RMX_ERR_ARG, "the grid holds 1 to 32 values");
    if (scalar_id < 0 || scalar_id > RMX_S_BETABIN_LOH_M_1) return m->fail(RMX_ERR_ARG, "scalar_id must name a likelihood parameter");
    std::vector<uint8_t> act(G), alive(G);
    std::vector<double> vals((size_t)G * n_grid), ell((size_t)G * n_grid), J((size_t)G * n_grid, INFINITY);
    std::vector<NelderMead1> nm(G);
    int64_t n_rounds = 0, n_evals = 0;
    int n_alive = 0;
    for (int g = 0; g < G; ++g) { alive[g] = !active || active[g]; n_alive += alive[g]; if (evals) evals[g] = 0; }
    // scipy.optimize.brute: the grid in ascending order; values outside the bounds are +inf and never assigned
    std::vector<int> inside;
    for (int i = 0; i < n_grid; ++i) if (!(grid[i] < lo || grid[i] > hi)) inside.push_back(i);
    const int ni = (int)inside.size();
    if (ni > 0 && n_alive > 0) {
        for (int g = 0; g < G; ++g) for (int k = 0; k < ni; ++k) vals[(size_t)g * ni + k] = grid[inside[k]];
        int rc = rmx_mstep_evaluate(m, alive.data(), nullptr, scalar_id, vals.data(), ni, ell.data(), nullptr);
        if (rc) return rc;
        for (int g = 0; g < G; ++g) if (alive[g]) for (int k = 0; k < ni; ++k) J[(size_t)g * n_grid + inside[k]] = -ell[(size_t)g * ni + k];
        n_rounds += ni;
        n_evals += (int64_t)ni * n_alive;
    }
    // fmin from the grid minimum (np.argmin: the first minimum, the first NaN if there is one)
    int running = 0;
    for (int g = 0; g < G; ++g) {
        if (!alive[g]) continue;
        const double* Jg = J.data() + (size_t)g * n_grid;
        int best = 0;
        for (int i = 0; i < n_grid; ++i) {
            if (Jg[i] != Jg[i]) { best = i; break; }
            if (Jg[i] < Jg[best]) best = i;
        }
        nm[g].start(grid[best]);
        running += !nm[g].done();
    }
    std::vector<double> v1(G, 0.), e1(G, 0.);
    while (running > 0) {
        int n_act = 0;
        for (int g = 0; g < G; ++g) {
            act[g] = 0;
            if (!alive[g] || nm[g].done()) continue;
            const double x = nm[g].next();
            if (!(x < lo || x > hi)) { act[g] = 1; v1[g] = x; ++n_act; }
        }
        if (n_act > 0) {
            int rc = rmx_mstep_evaluate(m, act.data(), nullptr, scalar_id, v1.data(), 1, e1.data(), nullptr);
            if (rc) return rc;
            ++n_rounds;
            n_evals += n_act;
        }
        running = 0;
        for (int g = 0; g < G; ++g) {
            if (!alive[g] || nm[g].done()) continue;
            const double x = nm[g].next();
            const double f = act[g] ? -e1[g] : INFINITY;
            if (trace && evals && evals[g] < trace_cap) {
                trace[((size_t)g * trace_cap + evals[g]) * 2] = x;
                trace[((size_t)g * trace_cap + evals[g]) * 2 + 1] = f;
            }
            if (evals) ++evals[g];
            nm[g].feed(f);
            running += !nm[g].done();
        }
    }
    for (int g = 0; g < G; ++g) result[g] = alive[g] ? nm[g].s0 : 0.;
    if (rounds) { rounds[0] = n_rounds; rounds[1] = n_evals; }
    return RMX_OK;
}

int rmx_set_scalar_array(rmx_handle m, int id, const uint8_t* active, const double* values) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    if (!scalar_slot(b->P[m->g0], id)) return m->fail(RMX_ERR_ARG, "scalar_id must name a per-fit parameter");
    for (int g = 0; g < m->gn; ++g) {
        if (active && !active[g]) continue;
        *scalar_slot(b->P[m->g0 + g], id) = values[g];
        if (id != RMX_S_DIVERGENCE_WEIGHT) b->ll_dirty[m->g0 + g] |= (char)planes_of(id);
    }
    b->params_dirty = true;
    b->ms_const_valid = false;
    return RMX_OK;
}

int rmx_infer_cn(rmx_handle m, int64_t* cn) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const RmxDims& d = b->d;
    const int g0 = m->g0, gn = m->gn, N = d.N;
    // max_product(self.framelogprob, self.log_transmat) (bpmodel.pyx:1203): both as of the last update_p_cn;
    // constructor state (framelogprob == 1, log_transmat == 0) gives the all-zero path (first max)
    bool any = false;
    for (int g = g0; g < g0 + gn; ++g) any |= b->cn_updated[g] != 0;
    if (any) {
        if (b->lattice_fits < gn) {
            // scratch sized for the largest range decoded so far (old one is freed with the batch)
            cudaError_t e = dalloc(b, &b->lattice, (size_t)gn * N * d.S);
            if (e != cudaSuccess) return m->fail(RMX_ERR_CUDA, "out of device memory for the Viterbi lattice");
            b->lattice_fits = gn;
        }
        int rc = RMX_OK;
        for_runs(g0, gn, b->tm_used, &b->cn_updated, [&](int g, int n) {
            if (rc) return;
            rc = ensure_mats(m, b->tm_used[g]);
            if (rc) return;
            rmx_launch_viterbi(b->stream, tables(b), d, g, n, chainsof(b), rowsof(b), spec_of(b, false, b->tm_used[g]),
                               b->rmd, b->F, b->lattice, b->path, b->C);
            b->count();
        });
        if (rc) return rc;
        CK(cudaMemcpyAsync(b->hrows + (size_t)g0 * N, b->path + (size_t)g0 * N, sizeof(int) * (size_t)gn * N,
                           cudaMemcpyDeviceToHost, b->stream));
        rc = sync_check(m);
        if (rc) return rc;
    }
    // cn[n, m, ell] = cn_states[n, path[n], m, ell]; the allele-swap branch of bpmodel.pyx:1208-1209
    // indexes both sides with the swapped ell, so it does not change the output.  One pass per fit (path kept for
    // RMX_A_STATE_PATH, states expanded into the caller's array), the fits spread over a few host threads: the
    // output of a whole grid is tens of MB of fresh pages.
    auto expand = [&](int gbeg, int gend) {
        for (int g = gbeg; g < gend; ++g) {
            const bool dec = any && b->cn_updated[g0 + g];
            b->path_valid[g0 + g] = 1;
            for (int n = 0; n < N; ++n) {
                const int s = dec ? b->hrows[(size_t)(g0 + g) * N + n] : 0;
                b->h_path[(size_t)(g0 + g) * N + n] = s;
                int64_t* o = cn + ((size_t)g * N + n) * d.M * 2;
                o[0] = b->h_nrm[2 * n];
                o[1] = b->h_nrm[2 * n + 1];
                for (int e = 2; e < d.M * 2; ++e) o[e] = b->h_tab[(size_t)s * d.M * 2 + e];
            }
        }
    };
    int nthr = 1;
    if ((int64_t)gn * N >= (1 << 18)) {
        nthr = (int)std::thread::hardware_concurrency();
        if (nthr > 8) nthr = 8;
        if (nthr > gn) nthr = gn;
        if (nthr < 1) nthr = 1;
    }
    if (nthr == 1) {
        expand(0, gn);
    } else {
        std::vector<std::thread> pool;
        for (int k = 0; k < nthr; ++k)
            pool.emplace_back(expand, (int)((int64_t)gn * k / nthr), (int)((int64_t)gn * (k + 1) / nthr));
        for (auto& th : pool) th.join();
    }
    return RMX_OK;
}

int rmx_calculate_log_transmat(rmx_handle m, double* out) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    if (m->gn != 1) return m->fail(RMX_ERR_ARG, "calculate_log_transmat works on a single-fit handle");
    // calculate_log_transmat(out) evaluates with the CURRENT p_breakpoint / transition_model (bpmodel.pyx:640),
    // without disturbing the `used` / `cached` snapshots
    const RmxDims& d = b->d;
    const int g = m->g0, tm = b->tm[g];
    if (d.N <= 1) return RMX_OK;
    double *pd = nullptr, *we = nullptr;
    CK(cudaMalloc((void**)&pd, (size_t)b->rmd * sizeof(double)));
    CK(cudaMalloc((void**)&we, (size_t)b->rmd * sizeof(double)));
    int rc = ensure_mats(m, tm);
    if (rc == RMX_OK) {
        // fit_stride 0 with the fit's own p_breakpoint slice: kernel indexes p_brk by absolute fit, tables by stride
        rmx_launch_prep_brk(b->stream, d, g, 1, b->R, b->brk_rows, b->bidx, b->borient, b->brk, b->p_brk, b->pen, tm, 0,
                            pd - 0, we - 0);
        RmxTransSpec s = spec_of(b, false, tm);
        s.pd = pd;
        s.wexp = we;
        const int64_t SS = (int64_t)d.S * d.S, nrows = d.N - 1;
        int64_t chunk = (int64_t)(64 << 20) / (SS * 8);
        if (chunk < 1) chunk = 1;
        if (chunk > nrows) chunk = nrows;
        double* buf = nullptr;
        cudaError_t e = cudaMalloc((void**)&buf, (size_t)(chunk * SS) * sizeof(double));
        if (e != cudaSuccess) rc = m->fail(RMX_ERR_CUDA, cudaGetErrorString(e));
        for (int64_t r0 = 0; r0 < nrows && rc == RMX_OK; r0 += chunk) {
            int64_t nr = nrows - r0 < chunk ? nrows - r0 : chunk;
            rmx_launch_materialize_T(b->stream, tables(b), d, rowsof(b), s, r0, nr, buf);
            b->count();
            e = cudaMemcpyAsync(out + r0 * SS, buf, (size_t)(nr * SS) * sizeof(double), cudaMemcpyDeviceToHost, b->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
            if (e != cudaSuccess) rc = m->fail(RMX_ERR_CUDA, cudaGetErrorString(e));
        }
        if (buf) cudaFree(buf);
    }
    cudaStreamSynchronize(b->stream);
    cudaFree(pd);
    cudaFree(we);
    return rc;
}

int rmx_eval(rmx_handle m, int what, int n, int s, int u, int v, int w, double* out) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    cudaSetDevice(b->device);
    const RmxDims& d = b->d;
    if (m->gn != 1) return m->fail(RMX_ERR_ARG, "the per-(n,s) accessors work on a single-fit handle");
    if (n < 0 || n >= d.N || s < 0 || s >= d.S || u < 0 || u > 1 || v < 0 || v > 1 || w < 0 || w > 1)
        return m->fail(RMX_ERR_ARG, "index out of range");
    if (what < 0 || what > 8) return m->fail(RMX_ERR_ARG, "unknown eval id");
    int rc = push_params(m);
    if (rc) return rc;
    rmx_launch_eval(b->stream, tables(b), d, b->dP, m->g0, what, n, s, u, v, w, b->evalbuf, b->dflags);
    b->count();
    double* hp = b->hp + (size_t)m->g0 * Batch::RED + 32;
    CK(cudaMemcpyAsync(hp, b->evalbuf, RMX_MAXM * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
    rc = sync_check(m);
    if (rc) return rc;
    for (int k = 0; k < RMX_MAXM; ++k) out[k] = hp[k];
    return RMX_OK;
}

// ---- stateless module functions
static int stateless_fail(const char* msg) {
    g_create_error = msg;
    return RMX_ERR_CUDA;
}
#define SCK(call)                                                         \
    do {                                                                  \
        cudaError_t e__ = (call);                                         \
        if (e__ != cudaSuccess) {                                         \
            for (void* p__ : bufs) cudaFree(p__);                         \
            return stateless_fail(cudaGetErrorString(e__));               \
        }                                                                 \
    } while (0)

int rmx_sum_product(int device, int64_t N, int64_t S, const double* F, const double* Tm, double* alphas,
                    double* betas) {
    std::vector<void*> bufs;
    if (N < 1 || S < 1 || S > 4096) return stateless_fail("bad shape");
    SCK(cudaSetDevice(device));
    double *dF = nullptr, *dT = nullptr, *dA = nullptr, *dB = nullptr;
    size_t ns = (size_t)N * S * sizeof(double), nt = (size_t)(N - 1) * S * S * sizeof(double);
    SCK(cudaMalloc((void**)&dF, ns)); bufs.push_back(dF);
    SCK(cudaMalloc((void**)&dT, nt ? nt : 8)); bufs.push_back(dT);
    SCK(cudaMalloc((void**)&dA, ns)); bufs.push_back(dA);
    SCK(cudaMalloc((void**)&dB, ns)); bufs.push_back(dB);
    SCK(cudaMemcpy(dF, F, ns, cudaMemcpyHostToDevice));
    if (nt) SCK(cudaMemcpy(dT, Tm, nt, cudaMemcpyHostToDevice));
    rmx_launch_sum_product_log(0, N, (int)S, dF, dT, dA, dB);
    g_launches += 1;
    SCK(cudaGetLastError());
    SCK(cudaMemcpy(alphas, dA, ns, cudaMemcpyDeviceToHost));
    SCK(cudaMemcpy(betas, dB, ns, cudaMemcpyDeviceToHost));
    for (void* p : bufs) cudaFree(p);
    return RMX_OK;
}

int rmx_sum_product_2paramtrans(int device, int64_t N, int64_t S, const double* F, double* alphas, double* betas,
                                double e0, double e1) {
    std::vector<void*> bufs;
    if (N < 1 || S < 1 || S > 4096) return stateless_fail("bad shape");
    SCK(cudaSetDevice(device));
    double *dF = nullptr, *dA = nullptr, *dB = nullptr;
    size_t ns = (size_t)N * S * sizeof(double);
    SCK(cudaMalloc((void**)&dF, ns)); bufs.push_back(dF);
    SCK(cudaMalloc((void**)&dA, ns)); bufs.push_back(dA);
    SCK(cudaMalloc((void**)&dB, ns)); bufs.push_back(dB);
    SCK(cudaMemcpy(dF, F, ns, cudaMemcpyHostToDevice));
    rmx_launch_sum_product_2p(0, N, (int)S, dF, dA, dB, e0, e1);
    g_launches += 1;
    SCK(cudaGetLastError());
    SCK(cudaMemcpy(alphas, dA, ns, cudaMemcpyDeviceToHost));
    SCK(cudaMemcpy(betas, dB, ns, cudaMemcpyDeviceToHost));
    for (void* p : bufs) cudaFree(p);
    return RMX_OK;
}

int rmx_max_product(int device, int64_t N, int64_t S, const double* F, const double* Tm, int64_t* path,
                    double* logprob) {
    std::vector<void*> bufs;
    if (N < 1 || S < 1 || S > 4096) return stateless_fail("bad shape");
    SCK(cudaSetDevice(device));
    double *dF = nullptr, *dT = nullptr, *dL = nullptr, *dP = nullptr;
    int64_t* dpath = nullptr;
    size_t ns = (size_t)N * S * sizeof(double), nt = (size_t)(N - 1) * S * S * sizeof(double);
    SCK(cudaMalloc((void**)&dF, ns)); bufs.push_back(dF);
    SCK(cudaMalloc((void**)&dT, nt ? nt : 8)); bufs.push_back(dT);
    SCK(cudaMalloc((void**)&dL, ns)); bufs.push_back(dL);
    SCK(cudaMalloc((void**)&dP, 8)); bufs.push_back(dP);
    SCK(cudaMalloc((void**)&dpath, (size_t)N * sizeof(int64_t))); bufs.push_back(dpath);
    SCK(cudaMemcpy(dF, F, ns, cudaMemcpyHostToDevice));
    if (nt) SCK(cudaMemcpy(dT, Tm, nt, cudaMemcpyHostToDevice));
    rmx_launch_max_product(0, N, (int)S, dF, dT, dL, dpath, dP);
    g_launches += 1;
    SCK(cudaGetLastError());
    SCK(cudaMemcpy(path, dpath, (size_t)N * sizeof(int64_t), cudaMemcpyDeviceToHost));
    SCK(cudaMemcpy(logprob, dP, sizeof(double), cudaMemcpyDeviceToHost));
    for (void* p : bufs) cudaFree(p);
    return RMX_OK;
}

// Viterbi decode without any allocation: device pointers in and out, caller's stream, lattice in the library's
// file-scope device array (chain.cu).  No reference counterpart; see include/remixt_b200.h.
int rmx_max_product_noalloc(int device, void* stream, int64_t N, int64_t S, const double* F, const double* Tm,
                            int64_t* path, double* logprob) {
    if (N < 1 || S < 1 || S > 4096) { g_create_error = "bad shape"; return RMX_ERR_ARG; }
    if (N > (int64_t)RMX_NOALLOC_MAX_CELLS / S) {
        g_create_error = "rmx_max_product_noalloc: n_obs * n_comp = " + std::to_string(N * S) +
                         " exceeds the static lattice of RMX_NOALLOC_MAX_CELLS = " +
                         std::to_string((long long)RMX_NOALLOC_MAX_CELLS) + " cells; use rmx_max_product";
        return RMX_ERR_ARG;
    }
    if (!F || !path || !logprob || (N > 1 && !Tm)) { g_create_error = "null pointer"; return RMX_ERR_ARG; }
    const char* err = rmx_launch_max_product_noalloc((cudaStream_t)stream, device, N, (int)S, F, Tm, path, logprob);
    if (err) { g_create_error = err; return RMX_ERR_CUDA; }
    g_launches += 1;
    return RMX_OK;
}

// ---- instrumentation
int rmx_launch_count(rmx_handle m, int64_t* count) {
    if (!m) return RMX_ERR_ARG;
    *count = m->b->launches;
    return RMX_OK;
}
int rmx_global_launch_count(int64_t* count) {
    *count = g_launches.load();
    return RMX_OK;
}
int rmx_kernel_timing(rmx_handle m, int enable, double* total_ms, int64_t* launches) {
    if (!m) return RMX_ERR_ARG;
    Batch* b = m->b.get();
    Lock lk(b->mu);
    for (int c = 0; c < RMX_K_COUNT_; ++c) {
        if (total_ms) total_ms[c] = b->cls_ms[c];
        if (launches) launches[c] = b->cls_n[c];
        b->cls_ms[c] = 0.;
        b->cls_n[c] = 0;
    }
    b->timing = enable != 0;
    return RMX_OK;
}
int rmx_chain_timing(rmx_handle m, int enable, double* total_ms, int64_t* launches) {
    double ms[RMX_K_COUNT_];
    int64_t n[RMX_K_COUNT_];
    int rc = rmx_kernel_timing(m, enable, ms, n);
    if (rc) return rc;
    if (total_ms) *total_ms = ms[RMX_K_CHAIN];
    if (launches) *launches = n[RMX_K_CHAIN];
    return RMX_OK;
}
int rmx_synchronize(rmx_handle m) {
    if (!m) return RMX_ERR_ARG;
    Lock lk(m->b->mu);
    cudaSetDevice(m->b->device);
    return sync_check(m);
}
int rmx_stream(rmx_handle m, void** stream) {
    if (!m) return RMX_ERR_ARG;
    *stream = (void*)m->b->stream;
    return RMX_OK;
}
int rmx_event_record(rmx_handle m, int slot) {
    if (!m || slot < 0 || slot >= 16) return RMX_ERR_ARG;
    Lock lk(m->b->mu);
    cudaSetDevice(m->b->device);
    if (!m->ev[slot]) CK(cudaEventCreate(&m->ev[slot]));
    CK(cudaEventRecord(m->ev[slot], m->b->stream));
    return RMX_OK;
}
int rmx_event_elapsed(rmx_handle m, int slot0, rmx_handle m1, int slot1, double* ms) {
    if (!m || !m1 || slot0 < 0 || slot0 >= 16 || slot1 < 0 || slot1 >= 16) return RMX_ERR_ARG;
    if (!m->ev[slot0] || !m1->ev[slot1]) return m->fail(RMX_ERR_ARG, "event was never recorded");
    cudaSetDevice(m->b->device);
    CK(cudaEventSynchronize(m->ev[slot0]));
    CK(cudaEventSynchronize(m1->ev[slot1]));
    float f = 0.f;
    CK(cudaEventElapsedTime(&f, m->ev[slot0], m1->ev[slot1]));
    *ms = (double)f;
    return RMX_OK;
}
int rmx_fp64_peak(int device, double* tflops) {
    double v = 0.;
    const char* e = rmx_measure_fp64_peak(device, &v);
    if (e) { g_create_error = e; return RMX_ERR_CUDA; }
    *tflops = v;
    g_launches += 6;
    return RMX_OK;
}

}  // extern "C"
