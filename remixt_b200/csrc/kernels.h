// Host-side launcher declarations shared by api.cu, likelihood.cu and chain.cu.
#pragma once
#include "common.cuh"

// A transition specification: everything that defines log_transmat[n,:,:] for every row n.
// Two of these are alive per model: `used` (as of the last update_p_cn, bpmodel.pyx:939) and
// `cached` (as of the constructor / last update_p_breakpoint, bpmodel.pyx:604,985).
// The (N-1,S,S) arrays of the reference are never materialised; a row is either
//   telomere  : all zeros                                   (bpmodel.pyx:649-650)
//   plain     : one of C class matrices Tmat[c]             (bpmodel.pyx:652-656, 670-684)
//   breakend  : -pen*(sum_m pd[r][m][dtot_m] + flipc[c])    (bpmodel.pyx:658-668, 670-684)
struct RmxTransSpec {
    const double* pd;     // (G)(R,M,D) expected |d - orient*brk| under p_breakpoint, per fit and breakend row
    const double* wexp;   // (G)(R,M,D) exp(-pen*pd)
    const double* Tmat;   // (C,S,S) exact reference-order log transition of plain rows
    const double* Emat;   // (C,S,S) exp(Tmat); symmetric in (i,j) because the cost is even in the difference
    const double* flipc;  // (C,S,S) min-over-allele-flip excess cost (integer valued)
    const double* flipE;  // (C,S,S) exp(-pen*flipc)
    double pen;           // transition penalty
    int tm;               // transition model (0: |d|, 1: d != 0)
};

struct RmxRows {
    const int* tclass;    // (N-1) -1 = telomere row, else class id
    const int* brow;      // (N-1) breakend-row index r, or -1
    const int8_t* d0;     // (N-1) normal total copies[n] - normal total copies[n+1]
    const int* brk_rows;  // (R) segment index n of breakend row r
};

struct RmxChains {
    const int* start;     // (NC) first segment of chain c
    const int* end;       // (NC) last segment of chain c (inclusive)
    const int* order;     // (NC) chain ids, longest chain first (launch order of the chain kernel)
    int NC;
    // the chains packed into NB bins of about equal total length (longest-first greedy), the unit of work of the
    // Viterbi kernel: bin b holds the chains bin_idx[bin_ptr[b] .. bin_ptr[b+1])
    const int* bin_ptr;
    const int* bin_idx;
    int NB;
};

// ---- likelihood.cu   (every launcher covers the fits [g0, g0+gn) of a batch; P is the per-fit parameter array)
void rmx_launch_segconst(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                         const int* rows_all, const int* Rn, int Rmax, double* segc, const double* segk = nullptr,
                         int planes = 15, const RmxFitSel* sel = nullptr);
// per-segment data-only constants (N,3) the full-table form of the above reuses
void rmx_launch_segk(cudaStream_t st, const RmxTables& T, const RmxDims& d, double* segk);
void rmx_launch_lltable(cudaStream_t st, int sm_count, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0,
                        int gn, const double* segc, double* LL, int* dflags);
// fast path (per-state tables + wide log table); rmx_lltable2_ok says whether the shape fits its shared memory
int rmx_lltable2_ok(const RmxDims& d, int Q);
void rmx_launch_lltable2(cudaStream_t st, int sm_count, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0,
                         int gn, const double* segc, const RmxFast& Fp, double* LL, int* dflags, int planes = 15,
                         const RmxFitSel* sel = nullptr);
void rmx_launch_frame(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                      const double* LL, const double* q_ot, const double* q_oa, const double* q_sw, double* F,
                      double* fexp, double* mF, int* dflags);
void rmx_launch_qsums(cudaStream_t st, const RmxDims& d, int g0, int gn, const double* LL, const double* post,
                      double* A, int planes = 15, const RmxFitSel* sel = nullptr);
// post + qsums + outlier updates (+ next round's swap update and frame) in one pass; S <= 128
int rmx_round_ok(const RmxDims& d);
void rmx_launch_round(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                      const double* LL, const double* ahat, const double* bhat, double* post, double* A, double* q_ot,
                      double* q_oa, double* q_sw, double prior_ot, double prior_oa, int next, int head, double* F,
                      double* fexp, double* mF, int* dflags);
void rmx_launch_qupdate(cudaStream_t st, const RmxDims& d, int g0, int gn, int which, const double* A,
                        const double* q_oa, const double* q_sw, double prior, double* out);
void rmx_launch_reduce_cols(cudaStream_t st, const double* in, int64_t n, int ncol, int g0, int gn,
                            int64_t in_fit_stride, double* out, int out_fit_stride);
void rmx_launch_elbo_rows(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                          const double* A, const double* post, const double* F, const double* q_ot, const double* q_oa,
                          const double* q_sw, double prior_ot, double prior_oa, double* rowv);
void rmx_launch_ell_rows_from_sums(cudaStream_t st, const RmxDims& d, int g0, int gn, const double* A, const double* q_ot,
                                   const double* q_oa, const double* q_sw, double* out, int64_t out_fit_stride);
void rmx_launch_plogp(cudaStream_t st, const double* p, int64_t n, int g0, int gn, double* out, int64_t out_fit_stride);
void rmx_launch_fill2(cudaStream_t st, double* out, int64_t n, double v0, double v1);
int rmx_ell_blocks(const RmxDims& d, int R);
// the sampled objective in one launch: out[g * out_fit_stride + chunk], chunk < rmx_ell_sample_chunks()
int rmx_ell_sample_ok(int Rmax);
int rmx_ell_sample_chunks();
void rmx_launch_ell_sample(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                           const int* rows_all, const int* Rn, const double* post, const double* q_ot, const double* q_oa,
                           const double* q_sw, const double* rowc, double* out, int out_fit_stride, int* dflags,
                           int planes);
// rowc[(g * 256 + r) * 3 ...]: the parameter-independent constants of the staged rows (once per staged sample)
void rmx_launch_ell_rowcache(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, const int* rows_all,
                             const int* Rn, int Rmax, double* rowc);
void rmx_launch_ell(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                    const int* rows_all, const int* Rn, int Rmax, const double* segc, const double* post,
                    const double* q_ot, const double* q_oa, const double* q_sw, double* part, int64_t part_fit_stride,
                    int* dflags, int planes = 15);
void rmx_launch_ell_grad(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                         const int* rows_all, const int* Rn, int Rmax, const double* post, const double* q_ot,
                         const double* q_oa, const double* q_sw, double* part, int64_t part_fit_stride, int* dflags);
void rmx_launch_eval(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g, int what, int n,
                     int s, int u, int v, int w, double* out, int* dflags);
// DFMA micro-benchmark; returns nullptr on success, else an error string
const char* rmx_measure_fp64_peak(int device, double* tflops);

// ---- chain.cu
void rmx_launch_build_class(cudaStream_t st, const RmxTables& T, const RmxDims& d, int C, const int8_t* class_nrm,
                            double pen, int tm, double* Tmat, double* Emat, double* flipc, double* flipE);
void rmx_launch_prep_brk(cudaStream_t st, const RmxDims& d, int g0, int gn, int R, const int* brk_rows, const int* bidx,
                         const int* borient, const int8_t* brk, const double* p_brk, double pen, int tm,
                         int64_t fit_stride, double* pd, double* wexp);
// returns 0 on success, 1 if the shared-memory configuration failed, 2 if the state space is unsupported
int rmx_launch_chain(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const RmxRows& rows,
                     const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp, const double* mF,
                     double* ahat, double* bhat, double* chain_logz, int* dflags);
// S <= 128: eight fits per CTA on the FP64 tensor path, one barrier per segment (chain_mma.cu); returns 2 if the
// shape is not covered
int rmx_launch_chain_mma(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                         const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                         const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags);
// second generation of the same (chain_mma2.cu): four DMMA warps with balanced column blocks, the last state of
// S = 97 on the vector pipe, per-segment bookkeeping inside the DMMA block; returns 2 for shapes it does not cover
int rmx_launch_chain_mma2(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                          const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                          const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags,
                          const uint8_t* bc, int BC, int KBB, int ND);
// 129 .. 168 states (chain_mma3.cu): seven DMMA warps, the class matrix split between registers and shared memory
int rmx_launch_chain_mma3(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                          const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                          const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags,
                          const uint8_t* bc, const uint16_t* didx, int BC, int KBB, int ND);
// 169 .. 360 states (chain_mma3.cu: k_chain_mma_s): eight fits per read of the L2-resident class matrix, B fragments
// loaded under the DMMA stream
int rmx_launch_chain_mma_s(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                           const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                           const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags,
                           const uint16_t* cord, const int* cstart);
void rmx_launch_post(cudaStream_t st, const RmxDims& d, int g0, int gn, const double* ahat, const double* bhat,
                     double* post, int* dflags);
void rmx_launch_hist(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, int R, const RmxRows& rows,
                     const RmxTransSpec& sp, int64_t sp_fit_stride, int uniform, const double* ahat, const double* bhat,
                     const double* fexp, double* hist, int* dflags, const uint8_t* hord = nullptr,
                     const int* hgstart = nullptr, const uint16_t* hord16 = nullptr, const int* hgs16 = nullptr, int TVN = 0,
                     int NGS = 0);
// atomic-free histogram kernel (S <= 128): per-batch state order tables
int rmx_hist2_ok(const RmxDims& d);
int rmx_hist2_gstart_count(const RmxDims& d);
void rmx_launch_hist_prep(cudaStream_t st, const RmxTables& T, const RmxDims& d, uint8_t* hord, int* hgstart);
void rmx_launch_brk_update(cudaStream_t st, const RmxDims& d, int g0, int gn, const int* krow_ptr, const int* krow_idx,
                           const int* brk_rows, const int* borient, const int8_t* brk, const double* hist,
                           int64_t hist_fit_stride, double pen, int tm, double* p_brk, int mode = 0,
                           double* logq = nullptr);
// rowsum[g][r] = sum_ij joint[n,i,j] * (T_tsp - (diff ? T_t2sp : 0))[n,i,j]; rowlist == nullptr: n = r over N-1 rows
void rmx_launch_pairsum(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, const RmxRows& rows,
                        const RmxTransSpec& joint_sp, const RmxTransSpec& t_sp, const RmxTransSpec& t2_sp, int diff,
                        int64_t sp_fit_stride, const int* rowlist, int nrows, int uniform, const double* ahat,
                        const double* bhat, const double* fexp, double* rowsum, int64_t rowsum_fit_stride);
// the breakend rows' sum joint * (T_cached - T_used) from the histograms of the last update_p_cn (same transition model)
void rmx_launch_pairsum_hist(cudaStream_t st, const RmxDims& d, int g0, int gn, int R, const RmxRows& rows,
                             const double* hist, const RmxTransSpec& cached, const RmxTransSpec& used,
                             int64_t sp_fit_stride, double* rowsum, int64_t rowsum_fit_stride);
void rmx_launch_viterbi(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, const RmxChains& ch,
                        const RmxRows& rows, const RmxTransSpec& sp, int64_t sp_fit_stride, const double* F,
                        double* lattice, int* path, int nclass);
// single fit: pointers (and sp.pd / sp.wexp) already offset to that fit
void rmx_launch_materialize_T(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxRows& rows,
                              const RmxTransSpec& sp, int64_t row0, int64_t nrows, double* out);
void rmx_launch_materialize_joint(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxRows& rows,
                                  const RmxTransSpec& sp, int uniform, const double* ahat, const double* bhat,
                                  const double* fexp, int64_t row0, int64_t nrows, double* out);
void rmx_launch_sum_product_log(cudaStream_t st, int64_t N, int S, const double* F, const double* Tm, double* alphas,
                                double* betas);
void rmx_launch_sum_product_2p(cudaStream_t st, int64_t N, int S, const double* F, double* alphas, double* betas,
                               double e0, double e1);
void rmx_launch_max_product(cudaStream_t st, int64_t N, int S, const double* F, const double* Tm, double* lattice,
                            int64_t* path, double* logprob);
// Viterbi decode with the lattice in the library's file-scope device array (no allocation); DEVICE pointers.  Returns
// nullptr or an error message.
const char* rmx_launch_max_product_noalloc(cudaStream_t st, int device, int64_t N, int S, const double* F,
                                           const double* Tm, int64_t* path, double* logprob);
