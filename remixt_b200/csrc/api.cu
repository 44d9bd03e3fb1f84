// C-ABI of libremixt_b200 (see include/remixt_b200.h): the model handle, its device buffers,
// generation tracking of cached quantities, and the host-side sequencing of kernel launches.
//
// State semantics kept from the reference (remixt/bpmodel.pyx):
//   framelogprob / log_transmat / joint_posterior_marginals are "as of the last update_p_cn" (:921-962);
//   cached_log_transmat is "as of the constructor or the last update_p_breakpoint" (:604, :985);
//   the likelihood terms always use the CURRENT h / dispersion parameters (:751-853).
#include <cuda_runtime.h>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/remixt_b200.h"
#include "kernels.h"

static thread_local std::string g_create_error;
static std::atomic<int64_t> g_launches{0};

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) {                                                                    \
            char buf__[512];                                                                         \
            snprintf(buf__, sizeof buf__, "CUDA error %s at %s:%d: %s", cudaGetErrorName(e__),       \
                     __FILE__, __LINE__, cudaGetErrorString(e__));                                   \
            return m->fail(RMX_ERR_CUDA, buf__);                                                        \
        }                                                                                            \
    } while (0)

struct ClassMats {
    double *Tmat = nullptr, *Emat = nullptr, *flipc = nullptr, *flipE = nullptr;
    double pen = NAN;
    bool built = false;
};

struct rmx_model_s {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    RmxDims d{};
    RmxParams P{};
    double pen = 0., prior_ot = 0.01, prior_oa = 0.01;
    int tm = 0;
    double logZ = 0.;
    std::string err;
    int64_t launches = 0;
    std::vector<void*> allocs;

    // host copies of the small integer tables
    std::vector<int8_t> h_tab, h_nrm, h_nsub, h_brk;
    std::vector<int64_t> h_tmask, h_amask;
    std::vector<int> h_path;

    // device: static tables
    int8_t *tab = nullptr, *nrm = nullptr, *nsub = nullptr, *brk = nullptr, *d0 = nullptr, *class_nrm = nullptr;
    double *l = nullptr, *x = nullptr, *y0 = nullptr, *y1 = nullptr;
    uint8_t *tmask = nullptr, *amask = nullptr;
    int *bidx = nullptr, *borient = nullptr, *tclass = nullptr, *brow = nullptr, *brk_rows = nullptr;
    int *krow_ptr = nullptr, *krow_idx = nullptr, *ch_start = nullptr, *ch_end = nullptr;
    int NC = 0, C = 0, R = 0;

    // device: variational state and scratch
    double *LL = nullptr, *segc = nullptr, *F = nullptr, *fexp = nullptr, *mF = nullptr;
    double *ahat = nullptr, *bhat = nullptr, *cnorm = nullptr, *zp = nullptr, *post = nullptr;
    double *q_sw = nullptr, *q_ot = nullptr, *q_oa = nullptr, *p_brk = nullptr;
    double *pd_used = nullptr, *wexp_used = nullptr, *pd_cached = nullptr, *wexp_cached = nullptr;
    double *hist = nullptr, *logq = nullptr, *chain_logz = nullptr, *rowv = nullptr, *rowsum = nullptr;
    double *red_out = nullptr, *part = nullptr, *lattice = nullptr, *segc_s = nullptr, *evalbuf = nullptr;
    int *path = nullptr, *d_rows = nullptr, *dflags = nullptr;
    int64_t part_cap = 0;
    ClassMats mats[2];
    int tm_used = 0, tm_cached = 0;
    double pen_used = 0., pen_cached = 0.;

    // pinned host staging
    double* hp = nullptr;   // 64 doubles
    int* hflags = nullptr;  // 1 int
    int* hrows = nullptr;   // N ints

    bool ll_dirty = true;
    bool cn_updated = false;   // false: constructor state (uniform posteriors, F == 1, log_transmat == 0)
    bool hist_valid = false;
    bool path_valid = false;

    cudaEvent_t ev[16] = {};
    // chain timing
    bool timing = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> tev;
    double chain_ms = 0.;
    int64_t chain_launches = 0;

    int fail(int code, const std::string& m) {
        err = m;
        return code;
    }
    void count(int k = 1) {
        launches += k;
        g_launches += k;
    }
};

typedef rmx_model_s Model;

template <typename T>
static cudaError_t dalloc(Model* m, T** p, size_t n) {
    if (n == 0) n = 1;
    cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
    if (e == cudaSuccess) m->allocs.push_back((void*)*p);
    return e;
}

template <typename T>
static cudaError_t upload(Model* m, T** dptr, const std::vector<T>& v) {
    cudaError_t e = dalloc(m, dptr, v.size());
    if (e != cudaSuccess) return e;
    if (!v.empty()) e = cudaMemcpy(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
    return e;
}

static RmxTables tables(const Model* m) {
    RmxTables T;
    T.tab = m->tab; T.nrm = m->nrm; T.nsub = m->nsub; T.l = m->l; T.x = m->x; T.y0 = m->y0; T.y1 = m->y1;
    T.tmask = m->tmask; T.amask = m->amask;
    return T;
}
static RmxRows rowsof(const Model* m) {
    RmxRows r;
    r.tclass = m->tclass; r.brow = m->brow; r.d0 = m->d0; r.brk_rows = m->brk_rows;
    return r;
}
static RmxChains chainsof(const Model* m) {
    RmxChains c;
    c.start = m->ch_start; c.end = m->ch_end; c.NC = m->NC;
    return c;
}

// class matrices for (tm, pen), built on demand
static int ensure_mats(Model* m, int tm, double pen) {
    ClassMats& cm = m->mats[tm];
    size_t n = (size_t)m->C * m->d.S * m->d.S;
    if (!cm.Tmat) {
        if (dalloc(m, &cm.Tmat, n) != cudaSuccess || dalloc(m, &cm.Emat, n) != cudaSuccess ||
            dalloc(m, &cm.flipc, n) != cudaSuccess || dalloc(m, &cm.flipE, n) != cudaSuccess)
            return m->fail(RMX_ERR_CUDA, "out of device memory for transition class matrices");
    }
    if (!cm.built || cm.pen != pen) {
        rmx_launch_build_class(m->stream, tables(m), m->d, m->C, m->class_nrm, pen, tm, cm.Tmat, cm.Emat, cm.flipc,
                               cm.flipE);
        m->count();
        cm.built = true;
        cm.pen = pen;
    }
    return RMX_OK;
}

static RmxTransSpec spec_of(Model* m, bool cached) {
    RmxTransSpec s;
    int tm = cached ? m->tm_cached : m->tm_used;
    ClassMats& cm = m->mats[tm];
    s.pd = cached ? m->pd_cached : m->pd_used;
    s.wexp = cached ? m->wexp_cached : m->wexp_used;
    s.Tmat = cm.Tmat; s.Emat = cm.Emat; s.flipc = cm.flipc; s.flipE = cm.flipE;
    s.pen = cached ? m->pen_cached : m->pen_used;
    s.tm = tm;
    return s;
}

// snapshot the current (p_breakpoint, transition_model, penalty) as the `used` or `cached` spec
static int snapshot_spec(Model* m, bool cached) {
    int rc = ensure_mats(m, m->tm, m->pen);
    if (rc) return rc;
    rmx_launch_prep_brk(m->stream, m->d, m->R, m->brk_rows, m->bidx, m->borient, m->brk, m->p_brk, m->pen, m->tm,
                        cached ? m->pd_cached : m->pd_used, cached ? m->wexp_cached : m->wexp_used);
    if (m->R > 0) m->count();
    if (cached) { m->tm_cached = m->tm; m->pen_cached = m->pen; }
    else { m->tm_used = m->tm; m->pen_used = m->pen; }
    return RMX_OK;
}

static int ensure_ll(Model* m) {
    if (!m->ll_dirty) return RMX_OK;
    RmxTables T = tables(m);
    rmx_launch_segconst(m->stream, T, m->P, nullptr, m->d.N, m->segc);
    rmx_launch_lltable(m->stream, m->sm_count, T, m->d, m->P, m->segc, m->LL, m->dflags);
    m->count(2);
    m->ll_dirty = false;
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
    cudaError_t e = cudaMemcpyAsync(m->hflags, m->dflags, sizeof(int), cudaMemcpyDeviceToHost, m->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
    if (e != cudaSuccess) {
        char buf[512];
        snprintf(buf, sizeof buf, "CUDA error %s: %s", cudaGetErrorName(e), cudaGetErrorString(e));
        return m->fail(RMX_ERR_CUDA, buf);
    }
    for (auto& pr : m->tev) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) {
            m->chain_ms += ms;
            m->chain_launches += 1;
        }
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    m->tev.clear();
    int f = *m->hflags;
    if (f) {
        cudaMemsetAsync(m->dflags, 0, sizeof(int), m->stream);
        int code;
        const char* msg = flag_message(f, &code);
        return m->fail(code, msg);
    }
    return RMX_OK;
}

// ---------------------------------------------------------------------------------------------
// coordinate updates (launch only; callers synchronise)
// ---------------------------------------------------------------------------------------------
static int do_update_framelogprob(Model* m) {
    int rc = ensure_ll(m);
    if (rc) return rc;
    rmx_launch_frame(m->stream, tables(m), m->d, m->P, m->LL, m->q_ot, m->q_oa, m->q_sw, m->F, m->fexp, m->mF,
                     m->dflags);
    m->count();
    return RMX_OK;
}

static int do_update_p_cn(Model* m) {
    int rc = do_update_framelogprob(m);
    if (rc) return rc;
    rc = snapshot_spec(m, false);
    if (rc) return rc;
    RmxTransSpec sp = spec_of(m, false);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (m->timing) {
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, m->stream);
    }
    int lc = rmx_launch_chain(m->stream, tables(m), m->d, chainsof(m), rowsof(m), sp, m->fexp, m->mF, m->ahat, m->bhat,
                              m->cnorm, m->chain_logz, m->dflags);
    if (m->timing) {
        cudaEventRecord(e1, m->stream);
        m->tev.push_back({e0, e1});
    }
    if (lc == 2) return m->fail(RMX_ERR_UNSUPPORTED, "num_cn_states > 1024 is not supported by the chain kernel");
    if (lc) return m->fail(RMX_ERR_CUDA, "could not configure shared memory for the chain kernel");
    m->count();
    rmx_launch_reduce_cols(m->stream, m->chain_logz, m->NC, 1, m->red_out + 15);
    rmx_launch_post(m->stream, m->d, m->ahat, m->bhat, m->post, m->zp, m->dflags);
    m->count(2);
    m->cn_updated = true;
    rmx_launch_hist(m->stream, tables(m), m->d, m->R, rowsof(m), sp, 0, m->ahat, m->bhat, m->fexp, m->hist, m->dflags);
    if (m->R > 0) m->count();
    m->hist_valid = true;
    m->path_valid = false;
    CK(cudaMemcpyAsync(m->hp + 15, m->red_out + 15, sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    return RMX_OK;
}

static int do_update_p_breakpoint(Model* m) {
    if (m->d.K > 0 && m->R > 0) {
        if (!m->hist_valid) {
            RmxTransSpec sp = spec_of(m, true);
            rmx_launch_hist(m->stream, tables(m), m->d, m->R, rowsof(m), sp, 1, m->ahat, m->bhat, m->fexp, m->hist,
                            m->dflags);
            m->count();
            m->hist_valid = true;
        }
    }
    if (m->d.K > 0) {
        rmx_launch_brk_update(m->stream, m->d, m->krow_ptr, m->krow_idx, m->brk_rows, m->borient, m->brk, m->hist,
                              m->pen, m->tm, m->p_brk, m->logq);
        m->count();
    }
    return snapshot_spec(m, true);
}

static int do_update_q(Model* m, int which) {
    int rc = ensure_ll(m);
    if (rc) return rc;
    double* out = which == 0 ? m->q_sw : (which == 1 ? m->q_ot : m->q_oa);
    double prior = which == 1 ? m->prior_ot : m->prior_oa;
    rmx_launch_qfactor(m->stream, m->d, which, m->LL, m->post, m->q_oa, m->q_sw, prior, out);
    m->count();
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

int rmx_destroy(rmx_handle m) {
    if (!m) return RMX_OK;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    for (auto& pr : m->tev) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
    for (int i = 0; i < 16; ++i) if (m->ev[i]) cudaEventDestroy(m->ev[i]);
    for (void* p : m->allocs) cudaFree(p);
    if (m->hp) cudaFreeHost(m->hp);
    if (m->hflags) cudaFreeHost(m->hflags);
    if (m->hrows) cudaFreeHost(m->hrows);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
    return RMX_OK;
}

#define CREATE_FAIL(code, msg)          \
    do {                                \
        g_create_error = (msg);         \
        rmx_destroy(m);                 \
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

int rmx_create(const rmx_desc* desc, rmx_handle* out) {
    *out = nullptr;
    Model* m = new Model();
    const int N = desc->num_segments, S = desc->num_cn_states, M = desc->num_clones, B = desc->num_brk_states,
              K = desc->num_breakpoints;
    if (N < 1 || S < 1 || B < 1) CREATE_FAIL(RMX_ERR_VALUE, "num_segments, num_cn_states and num_brk_states must be positive");
    if (M < 2 || M > RMX_MAXM) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "num_clones must be between 2 and 8");
    if (S > 1024) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "num_cn_states > 1024 is not supported");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
        CREATE_FAIL(RMX_ERR_CUDA, "no CUDA device available: remixt_b200 has no CPU fallback");
    if (desc->device < 0 || desc->device >= ndev) CREATE_FAIL(RMX_ERR_ARG, "device ordinal out of range");
    m->device = desc->device;
    CCK(cudaSetDevice(m->device));
    cudaDeviceGetAttribute(&m->sm_count, cudaDevAttrMultiProcessorCount, m->device);
    CCK(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));

    // ---- state table + per-segment normal copies (cn_model.py:359-364 builds cn_states this way)
    m->h_tab.assign((size_t)S * M * 2, 0);
    m->h_nrm.assign((size_t)N * 2, 0);
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
                    m->h_tab[((size_t)s * M + mm) * 2 + a] = (int8_t)v;
                    if (v > vmax) vmax = v;
                }
        for (int n = 0; n < N; ++n) {
            const int64_t* r = cs + (size_t)n * row;
            int64_t n0 = r[0], n1 = r[1];
            if (!fits(n0) || !fits(n1)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "cn_states entries must be in [0, 100]");
            m->h_nrm[2 * n] = (int8_t)n0;
            m->h_nrm[2 * n + 1] = (int8_t)n1;
            if (n0 > vmax) vmax = n0;
            if (n1 > vmax) vmax = n1;
            for (int s = 0; s < S; ++s) {
                const int64_t* rs = r + (size_t)s * M * 2;
                if (rs[0] != n0 || rs[1] != n1)
                    CREATE_FAIL(RMX_ERR_UNSUPPORTED,
                                "cn_states[n,s,0,:] (normal copies) must not depend on the state s");
                for (int e = 2; e < M * 2; ++e)
                    if (rs[e] != (int64_t)m->h_tab[(size_t)s * M * 2 + e])
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
                    m->h_tab[((size_t)s * M + mm) * 2 + a] = (int8_t)v;
                    if (v > vmax) vmax = v;
                }
        for (int n = 0; n < 2 * N; ++n) {
            int64_t v = desc->normal_copies[n];
            if (!fits(v)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "cn_states entries must be in [0, 100]");
            m->h_nrm[n] = (int8_t)v;
            if (v > vmax) vmax = v;
        }
    } else {
        CREATE_FAIL(RMX_ERR_ARG, "either cn_states or (cn_table, normal_copies) must be given");
    }
    m->h_brk.assign((size_t)B * M, 0);
    for (int e = 0; e < B * M; ++e) {
        int64_t v = desc->brk_states[e];
        if (!fits(v)) CREATE_FAIL(RMX_ERR_UNSUPPORTED, "brk_states entries must be in [0, 100]");
        m->h_brk[e] = (int8_t)v;
        if (v > vmax) vmax = v;
    }
    // num_alleles_subclonal (bpmodel.pyx:505): alleles whose copy number differs between tumour clones
    m->h_nsub.assign(S, 0);
    int maxtot = 0;
    for (int s = 0; s < S; ++s) {
        int cnt = 0;
        for (int a = 0; a < 2; ++a) {
            int lo = 127, hi = -1;
            for (int mm = 1; mm < M; ++mm) {
                int v = m->h_tab[((size_t)s * M + mm) * 2 + a];
                lo = v < lo ? v : lo;
                hi = v > hi ? v : hi;
            }
            cnt += (hi != lo);
        }
        m->h_nsub[s] = (int8_t)cnt;
        for (int mm = 1; mm < M; ++mm) {
            int t = m->h_tab[((size_t)s * M + mm) * 2] + m->h_tab[((size_t)s * M + mm) * 2 + 1];
            if (t > maxtot) maxtot = t;
        }
    }
    for (int n = 0; n < N; ++n) {
        int t = m->h_nrm[2 * n] + m->h_nrm[2 * n + 1];
        if (t > maxtot) maxtot = t;
    }
    RmxDims& d = m->d;
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
    if (max_bidx + 1 != K)
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
        d0[n] = (int8_t)((m->h_nrm[2 * n] + m->h_nrm[2 * n + 1]) - (m->h_nrm[2 * n + 2] + m->h_nrm[2 * n + 3]));
        if (tel) {
            tclass[n] = -1;
            ch_end.push_back(n);
            ch_start.push_back(n + 1);
        } else {
            uint32_t key = ((uint32_t)(uint8_t)m->h_nrm[2 * n] << 24) | ((uint32_t)(uint8_t)m->h_nrm[2 * n + 1] << 16) |
                           ((uint32_t)(uint8_t)m->h_nrm[2 * n + 2] << 8) | (uint32_t)(uint8_t)m->h_nrm[2 * n + 3];
            auto it = class_of.find(key);
            if (it == class_of.end()) {
                it = class_of.emplace(key, (int)class_of.size()).first;
                for (int e = 0; e < 4; ++e) class_nrm.push_back(m->h_nrm[2 * n + e]);
            }
            tclass[n] = it->second;
        }
    }
    ch_end.push_back(N - 1);
    m->NC = (int)ch_start.size();
    m->C = (int)class_of.size();
    if (m->C == 0) { m->C = 1; class_nrm.assign(4, 1); }
    m->R = (int)brk_rows.size();
    // breakpoint -> its breakend rows, ascending n (the order bpmodel.pyx:971-980 visits them)
    std::vector<int> krow_ptr(K + 1, 0), krow_idx(m->R);
    for (int r = 0; r < m->R; ++r) krow_ptr[bidx[brk_rows[r]] + 1]++;
    for (int k = 0; k < K; ++k) krow_ptr[k + 1] += krow_ptr[k];
    {
        std::vector<int> fill(krow_ptr.begin(), krow_ptr.end() - 1);
        for (int r = 0; r < m->R; ++r) krow_idx[fill[bidx[brk_rows[r]]]++] = r;
    }

    // ---- uploads
    std::vector<double> y0(N), y1(N);
    for (int n = 0; n < N; ++n) { y0[n] = desc->y[2 * n]; y1[n] = desc->y[2 * n + 1]; }
    m->h_tmask.assign(N, 1);
    m->h_amask.assign(N, 1);
    std::vector<uint8_t> ones(N, 1);
    CCK(upload(m, &m->tab, m->h_tab));
    CCK(upload(m, &m->nrm, m->h_nrm));
    CCK(upload(m, &m->nsub, m->h_nsub));
    CCK(upload(m, &m->brk, m->h_brk));
    CCK(upload(m, &m->d0, d0));
    CCK(upload(m, &m->class_nrm, class_nrm));
    CCK(upload(m, &m->l, std::vector<double>(desc->l, desc->l + N)));
    CCK(upload(m, &m->x, std::vector<double>(desc->x, desc->x + N)));
    CCK(upload(m, &m->y0, y0));
    CCK(upload(m, &m->y1, y1));
    CCK(upload(m, &m->tmask, ones));
    CCK(upload(m, &m->amask, ones));
    CCK(upload(m, &m->bidx, bidx));
    CCK(upload(m, &m->borient, borient));
    CCK(upload(m, &m->tclass, tclass));
    CCK(upload(m, &m->brow, brow));
    CCK(upload(m, &m->brk_rows, brk_rows));
    CCK(upload(m, &m->krow_ptr, krow_ptr));
    CCK(upload(m, &m->krow_idx, krow_idx));
    CCK(upload(m, &m->ch_start, ch_start));
    CCK(upload(m, &m->ch_end, ch_end));

    // ---- parameters (bpmodel.pyx:573-597)
    RmxParams& P = m->P;
    memset(&P, 0, sizeof P);
    for (int mm = 0; mm < M; ++mm) P.h[mm] = desc->h_init[mm];
    P.negbin_r[0] = 500.; P.negbin_r[1] = 10.;
    P.negbin_hdel_mu = 1e-5; P.negbin_hdel_r[0] = 10.; P.negbin_hdel_r[1] = 1.;
    P.betabin_M[0] = 500.; P.betabin_M[1] = 10.;
    P.betabin_loh_p = 1e-3; P.betabin_loh_M[0] = 10.; P.betabin_loh_M[1] = 1.;
    P.divergence_weight = fabs(desc->divergence_weight);
    P.normal_contamination = desc->normal_contamination ? 1 : 0;
    m->pen = fabs(desc->transition_penalty);
    m->tm = 0;

    // ---- variational state
    const size_t NS = (size_t)N * S;
    CCK(dalloc(m, &m->LL, NS * RMX_NTERM));
    CCK(dalloc(m, &m->segc, (size_t)N * 12));
    CCK(dalloc(m, &m->F, NS));
    CCK(dalloc(m, &m->fexp, NS));
    CCK(dalloc(m, &m->mF, (size_t)N));
    CCK(dalloc(m, &m->ahat, NS));
    CCK(dalloc(m, &m->bhat, NS));
    CCK(dalloc(m, &m->cnorm, (size_t)N));
    CCK(dalloc(m, &m->zp, (size_t)N));
    CCK(dalloc(m, &m->post, NS));
    CCK(dalloc(m, &m->q_sw, (size_t)N * 2));
    CCK(dalloc(m, &m->q_ot, (size_t)N * 2));
    CCK(dalloc(m, &m->q_oa, (size_t)N * 2));
    CCK(dalloc(m, &m->p_brk, (size_t)K * B));
    size_t rmd = (size_t)m->R * M * d.D;
    CCK(dalloc(m, &m->pd_used, rmd));
    CCK(dalloc(m, &m->wexp_used, rmd));
    CCK(dalloc(m, &m->pd_cached, rmd));
    CCK(dalloc(m, &m->wexp_cached, rmd));
    CCK(dalloc(m, &m->hist, rmd));
    CCK(dalloc(m, &m->logq, (size_t)K * B));
    CCK(dalloc(m, &m->chain_logz, (size_t)m->NC));
    CCK(dalloc(m, &m->rowv, (size_t)N * 8 + (size_t)K * B));
    CCK(dalloc(m, &m->rowsum, (size_t)N));
    CCK(dalloc(m, &m->red_out, 64));
    m->part_cap = (int64_t)rmx_ell_blocks(d, N) * RMX_MAXM + 64;
    CCK(dalloc(m, &m->part, (size_t)m->part_cap));
    CCK(dalloc(m, &m->segc_s, (size_t)N * 12));
    CCK(dalloc(m, &m->evalbuf, 16));
    CCK(dalloc(m, &m->path, (size_t)N));
    CCK(dalloc(m, &m->d_rows, (size_t)N));
    CCK(dalloc(m, &m->dflags, 1));
    CCK(cudaMemset(m->dflags, 0, sizeof(int)));
    CCK(cudaMallocHost((void**)&m->hp, 64 * sizeof(double)));
    CCK(cudaMallocHost((void**)&m->hflags, sizeof(int)));
    CCK(cudaMallocHost((void**)&m->hrows, (size_t)N * sizeof(int)));
    *m->hflags = 0;

    {   // initial factors (bpmodel.pyx:546-584)
        std::vector<double> v(NS, 1.0);
        CCK(cudaMemcpy(m->F, v.data(), NS * sizeof(double), cudaMemcpyHostToDevice));
        std::fill(v.begin(), v.end(), 1.0 / (double)S);
        {   // posterior_marginals[:] = 1; /= sum  (bpmodel.pyx:564-565)
            double sum = 0.;
            for (int s = 0; s < S; ++s) sum += 1.0;
            std::fill(v.begin(), v.end(), 1.0 / sum);
        }
        CCK(cudaMemcpy(m->post, v.data(), NS * sizeof(double), cudaMemcpyHostToDevice));
        std::vector<double> q(2 * (size_t)N);
        for (int n = 0; n < N; ++n) { q[2 * n] = 0.5; q[2 * n + 1] = 0.5; }
        CCK(cudaMemcpy(m->q_sw, q.data(), q.size() * sizeof(double), cudaMemcpyHostToDevice));
        for (int n = 0; n < N; ++n) { q[2 * n] = 1. - 0.01; q[2 * n + 1] = 0.01; }
        CCK(cudaMemcpy(m->q_ot, q.data(), q.size() * sizeof(double), cudaMemcpyHostToDevice));
        CCK(cudaMemcpy(m->q_oa, q.data(), q.size() * sizeof(double), cudaMemcpyHostToDevice));
        // p_breakpoint favours states with at most one copy in every clone (bpmodel.pyx:546-554)
        std::vector<double> pb((size_t)K * B, 0.);
        double cnt = 0.;
        std::vector<double> rowp(B, 0.);
        for (int b = 0; b < B; ++b) {
            int mx = 0;
            for (int mm = 0; mm < M; ++mm) mx = m->h_brk[(size_t)b * M + mm] > mx ? m->h_brk[(size_t)b * M + mm] : mx;
            if (mx <= 1) { rowp[b] = 1.; cnt += 1.; }
        }
        for (int k = 0; k < K; ++k)
            for (int b = 0; b < B; ++b) pb[(size_t)k * B + b] = rowp[b] / cnt;
        if (K > 0) CCK(cudaMemcpy(m->p_brk, pb.data(), pb.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    CCK(cudaDeviceSynchronize());   // pageable uploads above went through the legacy stream
    // first cached transition expectation (bpmodel.pyx:604), transition_model still 0 here
    if (snapshot_spec(m, true) != RMX_OK) { std::string e = m->err; CREATE_FAIL(RMX_ERR_CUDA, e); }
    if (snapshot_spec(m, false) != RMX_OK) { std::string e = m->err; CREATE_FAIL(RMX_ERR_CUDA, e); }
    CCK(cudaStreamSynchronize(m->stream));
    CCK(cudaGetLastError());
    *out = m;
    return RMX_OK;
}

static int64_t array_count(Model* m, int id, bool* is_i64) {
    const RmxDims& d = m->d;
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
    default: return -1;
    }
}

int rmx_set_array(rmx_handle m, int id, const void* src, int64_t count) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    bool i64;
    int64_t want = array_count(m, id, &i64);
    if (want < 0) return m->fail(RMX_ERR_ARG, "unknown array id");
    if (want != count) return m->fail(RMX_ERR_VALUE, "array has the wrong number of elements for this attribute");
    double* dst = nullptr;
    switch (id) {
    case RMX_A_H: {
        const double* s = (const double*)src;
        for (int k = 0; k < m->d.M; ++k) m->P.h[k] = s[k];
        m->ll_dirty = true;
        return RMX_OK;
    }
    case RMX_A_P_BREAKPOINT: dst = m->p_brk; break;
    case RMX_A_P_ALLELE_SWAP: dst = m->q_sw; break;
    case RMX_A_P_OUTLIER_TOTAL: dst = m->q_ot; break;
    case RMX_A_P_OUTLIER_ALLELE: dst = m->q_oa; break;
    case RMX_A_POSTERIOR_MARGINALS: dst = m->post; break;
    case RMX_A_FRAMELOGPROB: dst = m->F; break;
    case RMX_A_TOTAL_LIKELIHOOD_MASK:
    case RMX_A_ALLELE_LIKELIHOOD_MASK: {
        const int64_t* s = (const int64_t*)src;
        std::vector<uint8_t> v(m->d.N);
        std::vector<int64_t>& keep = id == RMX_A_TOTAL_LIKELIHOOD_MASK ? m->h_tmask : m->h_amask;
        for (int n = 0; n < m->d.N; ++n) { keep[n] = s[n]; v[n] = s[n] != 0; }
        CK(cudaMemcpyAsync(id == RMX_A_TOTAL_LIKELIHOOD_MASK ? m->tmask : m->amask, v.data(), v.size(),
                           cudaMemcpyHostToDevice, m->stream));
        CK(cudaStreamSynchronize(m->stream));
        m->ll_dirty = true;
        return RMX_OK;
    }
    default: return m->fail(RMX_ERR_ARG, "this attribute cannot be assigned");
    }
    if (count > 0) {
        CK(cudaMemcpyAsync(dst, src, (size_t)count * sizeof(double), cudaMemcpyHostToDevice, m->stream));
        CK(cudaStreamSynchronize(m->stream));
    }
    return RMX_OK;
}

static int materialize(Model* m, int id, double* dst) {
    const RmxDims& d = m->d;
    const int64_t SS = (int64_t)d.S * d.S;
    const int64_t nrows = d.N - 1;
    if (nrows <= 0) return RMX_OK;
    if (id == RMX_A_LOG_TRANSMAT && !m->cn_updated) {   // constructor state: zeros (bpmodel.pyx:558)
        memset(dst, 0, (size_t)(nrows * SS) * sizeof(double));
        return RMX_OK;
    }
    int64_t chunk = (int64_t)(64 << 20) / (SS * 8);
    if (chunk < 1) chunk = 1;
    if (chunk > nrows) chunk = nrows;
    double* buf = nullptr;
    CK(cudaMalloc((void**)&buf, (size_t)(chunk * SS) * sizeof(double)));
    int rc = RMX_OK;
    for (int64_t r0 = 0; r0 < nrows && rc == RMX_OK; r0 += chunk) {
        int64_t nr = nrows - r0 < chunk ? nrows - r0 : chunk;
        if (id == RMX_A_JOINT_POSTERIOR_MARGINALS)
            rmx_launch_materialize_joint(m->stream, tables(m), d, rowsof(m), spec_of(m, false), m->cn_updated ? 0 : 1,
                                         m->ahat, m->bhat, m->fexp, r0, nr, buf);
        else
            rmx_launch_materialize_T(m->stream, tables(m), d, rowsof(m), spec_of(m, id == RMX_A_CACHED_LOG_TRANSMAT),
                                     r0, nr, buf);
        m->count();
        cudaError_t e = cudaMemcpyAsync(dst + r0 * SS, buf, (size_t)(nr * SS) * sizeof(double), cudaMemcpyDeviceToHost,
                                        m->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
        if (e != cudaSuccess) rc = m->fail(RMX_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaFree(buf);
    return rc;
}

int rmx_get_array(rmx_handle m, int id, void* dstv, int64_t count) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    bool i64;
    int64_t want = array_count(m, id, &i64);
    if (want < 0) return m->fail(RMX_ERR_ARG, "unknown array id");
    if (want != count) return m->fail(RMX_ERR_VALUE, "array has the wrong number of elements for this attribute");
    const double* src = nullptr;
    switch (id) {
    case RMX_A_H: memcpy(dstv, m->P.h, sizeof(double) * m->d.M); return RMX_OK;
    case RMX_A_P_BREAKPOINT: src = m->p_brk; break;
    case RMX_A_P_ALLELE_SWAP: src = m->q_sw; break;
    case RMX_A_P_OUTLIER_TOTAL: src = m->q_ot; break;
    case RMX_A_P_OUTLIER_ALLELE: src = m->q_oa; break;
    case RMX_A_POSTERIOR_MARGINALS: src = m->post; break;
    case RMX_A_FRAMELOGPROB: src = m->F; break;
    case RMX_A_TOTAL_LIKELIHOOD_MASK: memcpy(dstv, m->h_tmask.data(), sizeof(int64_t) * m->d.N); return RMX_OK;
    case RMX_A_ALLELE_LIKELIHOOD_MASK: memcpy(dstv, m->h_amask.data(), sizeof(int64_t) * m->d.N); return RMX_OK;
    case RMX_A_LOG_TRANSMAT:
    case RMX_A_CACHED_LOG_TRANSMAT:
    case RMX_A_JOINT_POSTERIOR_MARGINALS: return materialize(m, id, (double*)dstv);
    case RMX_A_L: src = m->l; break;
    case RMX_A_X: src = m->x; break;
    case RMX_A_Y: {
        std::vector<double> a(m->d.N), b(m->d.N);
        CK(cudaMemcpy(a.data(), m->y0, sizeof(double) * m->d.N, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(b.data(), m->y1, sizeof(double) * m->d.N, cudaMemcpyDeviceToHost));
        double* o = (double*)dstv;
        for (int n = 0; n < m->d.N; ++n) { o[2 * n] = a[n]; o[2 * n + 1] = b[n]; }
        return RMX_OK;
    }
    case RMX_A_STATE_PATH: {
        int64_t* o = (int64_t*)dstv;
        if (!m->path_valid) return m->fail(RMX_ERR_VALUE, "no state path: call rmx_infer_cn first");
        for (int n = 0; n < m->d.N; ++n) o[n] = m->h_path[n];
        return RMX_OK;
    }
    default: return m->fail(RMX_ERR_ARG, "unknown array id");
    }
    if (count > 0) {
        CK(cudaMemcpyAsync(dstv, src, (size_t)count * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
        CK(cudaStreamSynchronize(m->stream));
    }
    return RMX_OK;
}

int rmx_set_scalar(rmx_handle m, int id, double v) {
    if (!m) return RMX_ERR_ARG;
    RmxParams& P = m->P;
    switch (id) {
    case RMX_S_NEGBIN_R_0: P.negbin_r[0] = v; break;
    case RMX_S_NEGBIN_R_1: P.negbin_r[1] = v; break;
    case RMX_S_NEGBIN_HDEL_MU: P.negbin_hdel_mu = v; break;
    case RMX_S_NEGBIN_HDEL_R_0: P.negbin_hdel_r[0] = v; break;
    case RMX_S_NEGBIN_HDEL_R_1: P.negbin_hdel_r[1] = v; break;
    case RMX_S_BETABIN_M_0: P.betabin_M[0] = v; break;
    case RMX_S_BETABIN_M_1: P.betabin_M[1] = v; break;
    case RMX_S_BETABIN_LOH_P: P.betabin_loh_p = v; break;
    case RMX_S_BETABIN_LOH_M_0: P.betabin_loh_M[0] = v; break;
    case RMX_S_BETABIN_LOH_M_1: P.betabin_loh_M[1] = v; break;
    case RMX_S_PRIOR_OUTLIER_TOTAL: m->prior_ot = v; return RMX_OK;
    case RMX_S_PRIOR_OUTLIER_ALLELE: m->prior_oa = v; return RMX_OK;
    case RMX_S_TRANSITION_PENALTY: m->pen = v; return RMX_OK;
    case RMX_S_DIVERGENCE_WEIGHT: P.divergence_weight = v; return RMX_OK;
    case RMX_S_TRANSITION_MODEL:
        if (v != 0. && v != 1.) return m->fail(RMX_ERR_VALUE, "transition_model must be 0 or 1");
        m->tm = (int)v;
        return RMX_OK;
    case RMX_S_HMM_LOG_NORM_CONST: m->logZ = v; return RMX_OK;
    default: return m->fail(RMX_ERR_ARG, "unknown scalar id");
    }
    m->ll_dirty = true;
    return RMX_OK;
}

int rmx_get_scalar(rmx_handle m, int id, double* v) {
    if (!m) return RMX_ERR_ARG;
    const RmxParams& P = m->P;
    switch (id) {
    case RMX_S_NEGBIN_R_0: *v = P.negbin_r[0]; break;
    case RMX_S_NEGBIN_R_1: *v = P.negbin_r[1]; break;
    case RMX_S_NEGBIN_HDEL_MU: *v = P.negbin_hdel_mu; break;
    case RMX_S_NEGBIN_HDEL_R_0: *v = P.negbin_hdel_r[0]; break;
    case RMX_S_NEGBIN_HDEL_R_1: *v = P.negbin_hdel_r[1]; break;
    case RMX_S_BETABIN_M_0: *v = P.betabin_M[0]; break;
    case RMX_S_BETABIN_M_1: *v = P.betabin_M[1]; break;
    case RMX_S_BETABIN_LOH_P: *v = P.betabin_loh_p; break;
    case RMX_S_BETABIN_LOH_M_0: *v = P.betabin_loh_M[0]; break;
    case RMX_S_BETABIN_LOH_M_1: *v = P.betabin_loh_M[1]; break;
    case RMX_S_PRIOR_OUTLIER_TOTAL: *v = m->prior_ot; break;
    case RMX_S_PRIOR_OUTLIER_ALLELE: *v = m->prior_oa; break;
    case RMX_S_TRANSITION_PENALTY: *v = m->pen; break;
    case RMX_S_DIVERGENCE_WEIGHT: *v = P.divergence_weight; break;
    case RMX_S_TRANSITION_MODEL: *v = (double)m->tm; break;
    case RMX_S_HMM_LOG_NORM_CONST: *v = m->logZ; break;
    default: return m->fail(RMX_ERR_ARG, "unknown scalar id");
    }
    return RMX_OK;
}

int rmx_get_derived(rmx_handle m, int which, int64_t* dst, int64_t count) {
    if (!m) return RMX_ERR_ARG;
    const RmxDims& d = m->d;
    int64_t want = which == 0 ? (int64_t)d.N * d.S * d.M : (int64_t)d.N * d.S;
    if (which < 0 || which > 3) return m->fail(RMX_ERR_ARG, "unknown derived table");
    if (want != count) return m->fail(RMX_ERR_VALUE, "array has the wrong number of elements");
    for (int n = 0; n < d.N; ++n) {
        int n0 = m->h_nrm[2 * n], n1 = m->h_nrm[2 * n + 1];
        for (int s = 0; s < d.S; ++s) {
            int a0 = n0, a1 = n1;
            for (int mm = 1; mm < d.M; ++mm) {
                a0 += m->h_tab[((size_t)s * d.M + mm) * 2];
                a1 += m->h_tab[((size_t)s * d.M + mm) * 2 + 1];
            }
            int64_t o = (int64_t)n * d.S + s;
            if (which == 0) {
                dst[o * d.M] = n0 + n1;
                for (int mm = 1; mm < d.M; ++mm)
                    dst[o * d.M + mm] = m->h_tab[((size_t)s * d.M + mm) * 2] + m->h_tab[((size_t)s * d.M + mm) * 2 + 1];
            } else if (which == 1) dst[o] = m->h_nsub[s];
            else if (which == 2) dst[o] = (a0 + a1) == 0;
            else dst[o] = (a0 == 0) || (a1 == 0);
        }
    }
    return RMX_OK;
}

int rmx_update(rmx_handle m, int id) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    int rc = do_update(m, id);
    if (rc) return rc;
    rc = sync_check(m);
    if (rc) return rc;
    if (id == RMX_U_CN) m->logZ = m->hp[15];
    return RMX_OK;
}

int rmx_variational_update(rmx_handle m, int n_iter) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    static const int order[5] = {RMX_U_ALLELE_SWAP, RMX_U_CN, RMX_U_BREAKPOINT, RMX_U_OUTLIER_TOTAL,
                                 RMX_U_OUTLIER_ALLELE};
    for (int it = 0; it < n_iter; ++it)
        for (int k = 0; k < 5; ++k) {
            int rc = do_update(m, order[k]);
            if (rc) return rc;
        }
    int rc = sync_check(m);
    if (rc) return rc;
    if (n_iter > 0) m->logZ = m->hp[15];
    return RMX_OK;
}

// launches the ELBO reductions; results land in hp[0..9] after sync
static int elbo_terms(Model* m) {
    int rc = ensure_ll(m);
    if (rc) return rc;
    const RmxDims& d = m->d;
    RmxTables T = tables(m);
    rmx_launch_elbo_rows(m->stream, T, d, m->P, m->LL, m->post, m->F, m->q_ot, m->q_oa, m->q_sw, m->prior_ot,
                         m->prior_oa, m->rowv);
    rmx_launch_reduce_cols(m->stream, m->rowv, d.N, 7, m->red_out);
    m->count(2);
    // entropy of p_breakpoint
    double* pl = m->rowv + (size_t)d.N * 8;
    int64_t kb = (int64_t)d.K * d.B;
    if (kb > 0) {
        rmx_launch_plogp(m->stream, m->p_brk, kb, pl);
        rmx_launch_reduce_cols(m->stream, pl, kb, 1, m->red_out + 7);
        m->count(2);
    } else {
        CK(cudaMemsetAsync(m->red_out + 7, 0, sizeof(double), m->stream));
    }
    // transition terms
    if (d.N > 1) {
        rc = ensure_mats(m, m->tm_cached, m->pen_cached);
        if (rc) return rc;
        rc = ensure_mats(m, m->tm_used, m->pen_used);
        if (rc) return rc;
        RmxTransSpec used = spec_of(m, false), cached = spec_of(m, true);
        int uniform = m->cn_updated ? 0 : 1;
        rmx_launch_pairsum(m->stream, T, d, rowsof(m), used, cached, uniform, m->ahat, m->bhat, m->fexp, m->rowsum);
        rmx_launch_reduce_cols(m->stream, m->rowsum, d.N - 1, 1, m->red_out + 8);
        m->count(2);
        if (m->cn_updated) {
            rmx_launch_pairsum(m->stream, T, d, rowsof(m), used, used, 0, m->ahat, m->bhat, m->fexp, m->rowsum);
            rmx_launch_reduce_cols(m->stream, m->rowsum, d.N - 1, 1, m->red_out + 9);
            m->count(2);
        } else {
            CK(cudaMemsetAsync(m->red_out + 9, 0, sizeof(double), m->stream));
        }
    } else {
        CK(cudaMemsetAsync(m->red_out + 8, 0, 2 * sizeof(double), m->stream));
    }
    CK(cudaMemcpyAsync(m->hp, m->red_out, 10 * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    return sync_check(m);
}

static double energy_of(const Model* m) {
    const double* r = m->hp;
    // prior, total (+ outlier prior), allele (+ outlier prior), transitions  (bpmodel.pyx:1068-1115)
    return r[0] + (r[1] + r[3]) + (r[2] + r[4]) + r[8];
}
static double entropy_of(const Model* m) {
    const double* r = m->hp;
    // bpmodel.pyx:1050-1056
    return -m->logZ + r[5] + r[9] + r[7] + r[6];
}

int rmx_variational_energy(rmx_handle m, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    int rc = elbo_terms(m);
    if (rc) return rc;
    *out = energy_of(m);
    return RMX_OK;
}
int rmx_variational_entropy(rmx_handle m, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    int rc = elbo_terms(m);
    if (rc) return rc;
    *out = entropy_of(m);
    return RMX_OK;
}
int rmx_elbo(rmx_handle m, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    int rc = elbo_terms(m);
    if (rc) return rc;
    *out = energy_of(m) - entropy_of(m);
    return RMX_OK;
}

// sample mask -> row list on the device; returns R and sets *rows (nullptr = identity)
static int stage_rows(Model* m, const int64_t* sample, int* R, const int** rows) {
    const int N = m->d.N;
    int cnt = 0;
    for (int n = 0; n < N; ++n)
        if (sample[n] != 0) m->hrows[cnt++] = n;
    *R = cnt;
    if (cnt == N) {
        *rows = nullptr;
        return RMX_OK;
    }
    if (cnt > 0) CK(cudaMemcpyAsync(m->d_rows, m->hrows, (size_t)cnt * sizeof(int), cudaMemcpyHostToDevice, m->stream));
    *rows = m->d_rows;
    return RMX_OK;
}

static int launch_ell(Model* m, const int* rows, int R, double* out_dev) {
    RmxTables T = tables(m);
    rmx_launch_segconst(m->stream, T, m->P, rows, R, m->segc_s);
    int nb = rmx_ell_blocks(m->d, R);
    rmx_launch_ell(m->stream, T, m->d, m->P, rows, R, m->segc_s, m->post, m->q_ot, m->q_oa, m->q_sw, m->part,
                   m->dflags);
    rmx_launch_reduce_cols(m->stream, m->part, nb, 1, out_dev);
    m->count(3);
    return RMX_OK;
}

int rmx_expected_log_likelihood(rmx_handle m, const int64_t* sample, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    int R;
    const int* rows;
    int rc = stage_rows(m, sample, &R, &rows);
    if (rc) return rc;
    if (R == 0) { *out = 0.; return RMX_OK; }
    launch_ell(m, rows, R, m->red_out + 16);
    CK(cudaMemcpyAsync(m->hp + 16, m->red_out + 16, sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    rc = sync_check(m);
    if (rc) return rc;
    *out = m->hp[16];
    return RMX_OK;
}

int rmx_expected_log_likelihood_grid(rmx_handle m, const int64_t* sample, int scalar_id, const double* values,
                                     int n_values, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    if (n_values < 0 || n_values > 32) return m->fail(RMX_ERR_ARG, "n_values must be in [0, 32]");
    if (scalar_id < RMX_S_NEGBIN_R_0 || scalar_id > RMX_S_BETABIN_LOH_M_1)
        return m->fail(RMX_ERR_ARG, "scalar_id must name a likelihood parameter");
    int R;
    const int* rows;
    int rc = stage_rows(m, sample, &R, &rows);
    if (rc) return rc;
    // the per-launch partial buffer is reused, which is safe because launches are stream ordered
    for (int g = 0; g < n_values; ++g) {
        rc = rmx_set_scalar(m, scalar_id, values[g]);
        if (rc) return rc;
        if (R == 0) { out[g] = 0.; continue; }
        launch_ell(m, rows, R, m->red_out + 16 + g);
    }
    if (R == 0 || n_values == 0) return RMX_OK;
    CK(cudaMemcpyAsync(m->hp + 16, m->red_out + 16, n_values * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    rc = sync_check(m);
    if (rc) return rc;
    for (int g = 0; g < n_values; ++g) out[g] = m->hp[16 + g];
    return RMX_OK;
}

int rmx_expected_log_likelihood_partial_h(rmx_handle m, const int64_t* sample, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    int R;
    const int* rows;
    int rc = stage_rows(m, sample, &R, &rows);
    if (rc) return rc;
    for (int k = 0; k < m->d.M; ++k) out[k] = 0.;
    if (R == 0) return RMX_OK;
    int nb = rmx_ell_blocks(m->d, R);
    rmx_launch_ell_grad(m->stream, tables(m), m->d, m->P, rows, R, m->post, m->q_ot, m->q_oa, m->q_sw, m->part,
                        m->dflags);
    rmx_launch_reduce_cols(m->stream, m->part, nb, RMX_MAXM, m->red_out + 48);
    m->count(2);
    CK(cudaMemcpyAsync(m->hp + 48, m->red_out + 48, RMX_MAXM * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    rc = sync_check(m);
    if (rc) return rc;
    for (int k = 0; k < m->d.M; ++k) out[k] = m->hp[48 + k];
    return RMX_OK;
}

int rmx_infer_cn(rmx_handle m, int64_t* cn) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    const RmxDims& d = m->d;
    m->h_path.assign(d.N, 0);
    if (m->cn_updated) {
        // max_product(self.framelogprob, self.log_transmat) (bpmodel.pyx:1203): both as of the last update_p_cn
        if (!m->lattice) {
            cudaError_t e = dalloc(m, &m->lattice, (size_t)d.N * d.S);
            if (e != cudaSuccess) return m->fail(RMX_ERR_CUDA, "out of device memory for the Viterbi lattice");
        }
        int rc = ensure_mats(m, m->tm_used, m->pen_used);
        if (rc) return rc;
        rmx_launch_viterbi(m->stream, tables(m), d, chainsof(m), rowsof(m), spec_of(m, false), m->F, m->lattice,
                           m->path);
        m->count();
        CK(cudaMemcpyAsync(m->h_path.data(), m->path, sizeof(int) * d.N, cudaMemcpyDeviceToHost, m->stream));
        rc = sync_check(m);
        if (rc) return rc;
    }
    // constructor state: framelogprob == 1 and log_transmat == 0 give the all-zero path (first max)
    m->path_valid = true;
    // cn[n, m, ell] = cn_states[n, path[n], m, ell]; the allele-swap branch of bpmodel.pyx:1208-1209
    // indexes both sides with the swapped ell, so it does not change the output
    for (int n = 0; n < d.N; ++n) {
        int s = m->h_path[n];
        int64_t* o = cn + (size_t)n * d.M * 2;
        o[0] = m->h_nrm[2 * n];
        o[1] = m->h_nrm[2 * n + 1];
        for (int e = 2; e < d.M * 2; ++e) o[e] = m->h_tab[(size_t)s * d.M * 2 + e];
    }
    return RMX_OK;
}

int rmx_calculate_log_transmat(rmx_handle m, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    // calculate_log_transmat(out) evaluates with the CURRENT p_breakpoint / transition_model
    // (bpmodel.pyx:640); borrow the `used` slot's scratch tables without disturbing its snapshot.
    const RmxDims& d = m->d;
    if (d.N <= 1) return RMX_OK;
    size_t rmd = (size_t)m->R * d.M * d.D;
    double *pd = nullptr, *we = nullptr;
    CK(cudaMalloc((void**)&pd, (rmd ? rmd : 1) * sizeof(double)));
    CK(cudaMalloc((void**)&we, (rmd ? rmd : 1) * sizeof(double)));
    int rc = ensure_mats(m, m->tm, m->pen);
    if (rc == RMX_OK) {
        rmx_launch_prep_brk(m->stream, d, m->R, m->brk_rows, m->bidx, m->borient, m->brk, m->p_brk, m->pen, m->tm, pd, we);
        RmxTransSpec s;
        ClassMats& cm = m->mats[m->tm];
        s.pd = pd; s.wexp = we; s.Tmat = cm.Tmat; s.Emat = cm.Emat; s.flipc = cm.flipc; s.flipE = cm.flipE;
        s.pen = m->pen; s.tm = m->tm;
        const int64_t SS = (int64_t)d.S * d.S, nrows = d.N - 1;
        int64_t chunk = (int64_t)(64 << 20) / (SS * 8);
        if (chunk < 1) chunk = 1;
        if (chunk > nrows) chunk = nrows;
        double* buf = nullptr;
        cudaError_t e = cudaMalloc((void**)&buf, (size_t)(chunk * SS) * sizeof(double));
        if (e != cudaSuccess) rc = m->fail(RMX_ERR_CUDA, cudaGetErrorString(e));
        for (int64_t r0 = 0; r0 < nrows && rc == RMX_OK; r0 += chunk) {
            int64_t nr = nrows - r0 < chunk ? nrows - r0 : chunk;
            rmx_launch_materialize_T(m->stream, tables(m), d, rowsof(m), s, r0, nr, buf);
            m->count();
            e = cudaMemcpyAsync(out + r0 * SS, buf, (size_t)(nr * SS) * sizeof(double), cudaMemcpyDeviceToHost, m->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
            if (e != cudaSuccess) rc = m->fail(RMX_ERR_CUDA, cudaGetErrorString(e));
        }
        if (buf) cudaFree(buf);
    }
    cudaStreamSynchronize(m->stream);
    cudaFree(pd);
    cudaFree(we);
    return rc;
}

int rmx_eval(rmx_handle m, int what, int n, int s, int u, int v, int w, double* out) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    const RmxDims& d = m->d;
    if (n < 0 || n >= d.N || s < 0 || s >= d.S || u < 0 || u > 1 || v < 0 || v > 1 || w < 0 || w > 1)
        return m->fail(RMX_ERR_ARG, "index out of range");
    if (what < 0 || what > 8) return m->fail(RMX_ERR_ARG, "unknown eval id");
    rmx_launch_eval(m->stream, tables(m), d, m->P, what, n, s, u, v, w, m->evalbuf, m->dflags);
    m->count();
    CK(cudaMemcpyAsync(m->hp + 32, m->evalbuf, RMX_MAXM * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    int rc = sync_check(m);
    if (rc) return rc;
    for (int k = 0; k < RMX_MAXM; ++k) out[k] = m->hp[32 + k];
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

// ---- instrumentation
int rmx_launch_count(rmx_handle m, int64_t* count) {
    if (!m) return RMX_ERR_ARG;
    *count = m->launches;
    return RMX_OK;
}
int rmx_global_launch_count(int64_t* count) {
    *count = g_launches.load();
    return RMX_OK;
}
int rmx_chain_timing(rmx_handle m, int enable, double* total_ms, int64_t* launches) {
    if (!m) return RMX_ERR_ARG;
    if (total_ms) *total_ms = m->chain_ms;
    if (launches) *launches = m->chain_launches;
    m->chain_ms = 0.;
    m->chain_launches = 0;
    m->timing = enable != 0;
    return RMX_OK;
}
int rmx_synchronize(rmx_handle m) {
    if (!m) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    return sync_check(m);
}
int rmx_stream(rmx_handle m, void** stream) {
    if (!m) return RMX_ERR_ARG;
    *stream = (void*)m->stream;
    return RMX_OK;
}

int rmx_event_record(rmx_handle m, int slot) {
    if (!m || slot < 0 || slot >= 16) return RMX_ERR_ARG;
    cudaSetDevice(m->device);
    if (!m->ev[slot]) CK(cudaEventCreate(&m->ev[slot]));
    CK(cudaEventRecord(m->ev[slot], m->stream));
    return RMX_OK;
}
int rmx_event_elapsed(rmx_handle m, int slot0, rmx_handle m1, int slot1, double* ms) {
    if (!m || !m1 || slot0 < 0 || slot0 >= 16 || slot1 < 0 || slot1 >= 16) return RMX_ERR_ARG;
    if (!m->ev[slot0] || !m1->ev[slot1]) return m->fail(RMX_ERR_ARG, "event was never recorded");
    cudaSetDevice(m->device);
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
