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
    const double* pd;     // (R,M,D) expected |d - orient*brk| under p_breakpoint, per breakend row
    const double* wexp;   // (R,M,D) exp(-pen*pd)
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
    int NC;
};

// ---- likelihood.cu
void rmx_launch_segconst(cudaStream_t st, const RmxTables& T, const RmxParams& P, const int* rows, int R, double* segc);
void rmx_launch_lltable(cudaStream_t st, int sm_count, const RmxTables& T, const RmxDims& d, const RmxParams& P,
                        const double* segc, double* LL, int* dflags);
void rmx_launch_frame(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams& P, const double* LL,
                      const double* q_ot, const double* q_oa, const double* q_sw, double* F, double* fexp, double* mF,
                      int* dflags);
void rmx_launch_qfactor(cudaStream_t st, const RmxDims& d, int which, const double* LL, const double* post,
                        const double* q_oa, const double* q_sw, double prior, double* out);
void rmx_launch_reduce_cols(cudaStream_t st, const double* in, int64_t n, int ncol, double* out);
void rmx_launch_elbo_rows(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams& P, const double* LL,
                          const double* post, const double* F, const double* q_ot, const double* q_oa,
                          const double* q_sw, double prior_ot, double prior_oa, double* rowv);
void rmx_launch_plogp(cudaStream_t st, const double* p, int64_t n, double* out);
int rmx_ell_blocks(const RmxDims& d, int R);
void rmx_launch_ell(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams& P, const int* rows, int R,
                    const double* segc, const double* post, const double* q_ot, const double* q_oa,
                    const double* q_sw, double* part, int* dflags);
void rmx_launch_ell_grad(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams& P, const int* rows,
                         int R, const double* post, const double* q_ot, const double* q_oa, const double* q_sw,
                         double* part, int* dflags);
void rmx_launch_eval(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams& P, int what, int n, int s,
                     int u, int v, int w, double* out, int* dflags);

// ---- chain.cu
void rmx_launch_build_class(cudaStream_t st, const RmxTables& T, const RmxDims& d, int C, const int8_t* class_nrm,
                            double pen, int tm, double* Tmat, double* Emat, double* flipc, double* flipE);
void rmx_launch_prep_brk(cudaStream_t st, const RmxDims& d, int R, const int* brk_rows, const int* bidx,
                         const int* borient, const int8_t* brk, const double* p_brk, double pen, int tm, double* pd,
                         double* wexp);
// returns 0 on success, nonzero if the shared-memory configuration failed
int rmx_launch_chain(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const RmxRows& rows,
                     const RmxTransSpec& sp, const double* fexp, const double* mF, double* ahat, double* bhat,
                     double* cnorm, double* chain_logz, int* dflags);
void rmx_launch_post(cudaStream_t st, const RmxDims& d, const double* ahat, const double* bhat, double* post,
                     double* zp, int* dflags);
void rmx_launch_hist(cudaStream_t st, const RmxTables& T, const RmxDims& d, int R, const RmxRows& rows,
                     const RmxTransSpec& sp, int uniform, const double* ahat, const double* bhat, const double* fexp,
                     double* hist, int* dflags);
void rmx_launch_brk_update(cudaStream_t st, const RmxDims& d, const int* krow_ptr, const int* krow_idx,
                           const int* brk_rows, const int* borient, const int8_t* brk, const double* hist, double pen,
                           int tm, double* p_brk, double* logq_scratch);
void rmx_launch_pairsum(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxRows& rows,
                        const RmxTransSpec& joint_sp, const RmxTransSpec& t_sp, int uniform, const double* ahat,
                        const double* bhat, const double* fexp, double* rowsum);
void rmx_launch_viterbi(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch,
                        const RmxRows& rows, const RmxTransSpec& sp, const double* F, double* lattice, int* path);
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

// DFMA micro-benchmark; returns nullptr on success, else an error string
const char* rmx_measure_fp64_peak(int device, double* tflops);
