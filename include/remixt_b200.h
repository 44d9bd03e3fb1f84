/*
 * remixt_b200 -- C-ABI of the B200-native ReMixT variational-inference core.
 *
 * This is the drop-in boundary for the reference's Cython extension `remixt/bpmodel.pyx`
 * (class RemixtModel and the module functions sum_product / max_product).  Every entry
 * point below names the reference interface it replaces (paths relative to the reference
 * tree).  Plain pointers and sizes only; all pointers are HOST pointers unless stated;
 * all arrays are C-contiguous, float64 / int64 exactly as the reference's typed buffers.
 *
 * Every function returns an rmx_status.  On failure rmx_last_error() returns the message;
 * RMX_ERR_VALUE maps to Python ValueError, RMX_ERR_ASSERT to AssertionError (the two
 * exception classes the reference raises on this path, bpmodel.pyx:208,270,336,345,
 * 384,392,722,739,829 and :934-962).
 *
 * There is no CPU fallback behind this interface: every compute entry point launches
 * sm_100a kernels on the handle's device and fails with RMX_ERR_CUDA if it cannot.
 *
 * Batches.  The reference fits one RemixtModel per initialisation of its (h, ploidy) grid, each in its own
 * process (workflow.py:329-340, analysis/pipeline.py:42-103).  Those fits share the data, the state tables
 * and the breakpoint structure, so here they live in one BATCH whose kernels each serve all fits in one
 * launch.  A handle addresses a contiguous range of `gn` fits of a batch: rmx_create makes a batch of one
 * (gn = 1, the reference's semantics exactly), rmx_batch_create a batch of G, rmx_batch_view a sub-range.
 * Every entry point acts on all fits of its handle: array attributes gain a leading dimension gn, scalar
 * outputs become arrays of gn doubles, scalar setters assign the value to every fit of the handle.
 * Calls on handles of one batch are serialised by a per-batch mutex (one CUDA stream per batch).
 */
#ifndef REMIXT_B200_H
#define REMIXT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rmx_model_s* rmx_handle;

typedef enum {
    RMX_OK = 0,
    RMX_ERR_CUDA = 1,        /* CUDA runtime failure (no device, launch error, OOM)      */
    RMX_ERR_VALUE = 2,       /* reference raises ValueError                              */
    RMX_ERR_ASSERT = 3,      /* reference `assert` fails (NaN checks, bpmodel.pyx:934-962) */
    RMX_ERR_UNSUPPORTED = 4, /* input outside what the device layout supports            */
    RMX_ERR_ARG = 5          /* bad id / size passed through this C-ABI                  */
} rmx_status;

/* rmx_desc.flags: the model is one shard of a chromosome-sharded fit: it holds a subset of the chromosomes and
 * all K breakpoints, breakpoint_idx marks the breakends on its own chromosomes only. */
#define RMX_DESC_SHARD 1

/* Constructor arguments: RemixtModel.__cinit__ (bpmodel.pyx:461-476), same order/meaning. */
typedef struct {
    int32_t num_clones;            /* M  (bpmodel.pyx:462) */
    int32_t num_segments;          /* N1 (bpmodel.pyx:463) */
    int32_t num_breakpoints;       /* K  (bpmodel.pyx:464) */
    int32_t normal_contamination;  /* bint (bpmodel.pyx:465) */
    int32_t num_cn_states;         /* S = cn_states.shape[1] */
    int32_t num_brk_states;        /* B = brk_states.shape[0] */
    int32_t device;                /* CUDA device ordinal the model lives on */
    int32_t flags;                 /* RMX_DESC_* bits; 0 = the reference's constructor exactly */
    const int64_t* cn_states;      /* (N1,S,M,2) (bpmodel.pyx:466); may be NULL if the compact pair below is given */
    const int64_t* cn_table;       /* optional compact form: (S,M,2), clone 0 row ignored */
    const int64_t* normal_copies;  /* optional compact form: (N1,2) = cn_states[n,:,0,:]  */
    const int64_t* brk_states;     /* (B,M)      (bpmodel.pyx:467) */
    const double* h_init;          /* (M)        (bpmodel.pyx:468) */
    const double* l;               /* (N1)       (bpmodel.pyx:469) */
    const double* x;               /* (N1)       (bpmodel.pyx:470) */
    const double* y;               /* (N1,2)     (bpmodel.pyx:471) */
    const int64_t* is_telomere;        /* (N1)   (bpmodel.pyx:472) */
    const int64_t* breakpoint_idx;     /* (N1)   (bpmodel.pyx:473) */
    const int64_t* breakpoint_orient;  /* (N1)   (bpmodel.pyx:474) */
    double transition_penalty;     /* (bpmodel.pyx:475), stored as fabs() (:534) */
    double divergence_weight;      /* (bpmodel.pyx:476), stored as fabs() (:535) */
} rmx_desc;

/* Array attributes (`cdef public` memoryviews, bpmodel.pyx:405-442). */
typedef enum {
    RMX_A_H = 0,                    /* f64 (M)        :442 */
    RMX_A_P_BREAKPOINT = 1,         /* f64 (K,B)      :420 */
    RMX_A_P_ALLELE_SWAP = 2,        /* f64 (N1,2)     :429 */
    RMX_A_P_OUTLIER_TOTAL = 3,      /* f64 (N1,2)     :430 */
    RMX_A_P_OUTLIER_ALLELE = 4,     /* f64 (N1,2)     :431 */
    RMX_A_POSTERIOR_MARGINALS = 5,  /* f64 (N1,S)     :426 */
    RMX_A_FRAMELOGPROB = 6,         /* f64 (N1,S)     :423 */
    RMX_A_TOTAL_LIKELIHOOD_MASK = 7,  /* i64 (N1)     :439 */
    RMX_A_ALLELE_LIKELIHOOD_MASK = 8, /* i64 (N1)     :440 */
    RMX_A_LOG_TRANSMAT = 9,           /* f64 (N1-1,S,S) :424, get only, materialised on demand */
    RMX_A_CACHED_LOG_TRANSMAT = 10,   /* f64 (N1-1,S,S) :425, get only, materialised on demand */
    RMX_A_JOINT_POSTERIOR_MARGINALS = 11, /* f64 (N1-1,S,S) :427, get only, materialised on demand */
    RMX_A_L = 12, RMX_A_X = 13, RMX_A_Y = 14,   /* f64, get only (:436-438) */
    RMX_A_STATE_PATH = 15,            /* i64 (N1): Viterbi path of the last rmx_infer_cn, get only */
    RMX_A_LL_TABLE = 16               /* f64 (6,N1,S): the cached likelihood terms calculate_log_likelihood_total(n,s,u)
                                         (:751, planes 0-1) and calculate_log_likelihood_allele(n,s,v,w) (:809, planes
                                         2+2v+w) for the current h / dispersions / masks; get only */
} rmx_array_id;

/* Scalar attributes (`cdef public` doubles / ints, bpmodel.pyx:417-456). */
typedef enum {
    RMX_S_NEGBIN_R_0 = 0, RMX_S_NEGBIN_R_1 = 1, RMX_S_NEGBIN_HDEL_MU = 2,
    RMX_S_NEGBIN_HDEL_R_0 = 3, RMX_S_NEGBIN_HDEL_R_1 = 4,
    RMX_S_BETABIN_M_0 = 5, RMX_S_BETABIN_M_1 = 6, RMX_S_BETABIN_LOH_P = 7,
    RMX_S_BETABIN_LOH_M_0 = 8, RMX_S_BETABIN_LOH_M_1 = 9,
    RMX_S_PRIOR_OUTLIER_TOTAL = 10, RMX_S_PRIOR_OUTLIER_ALLELE = 11,
    RMX_S_TRANSITION_PENALTY = 12, RMX_S_DIVERGENCE_WEIGHT = 13,
    RMX_S_TRANSITION_MODEL = 14, RMX_S_HMM_LOG_NORM_CONST = 15,
    RMX_S_COUNT_ = 16
} rmx_scalar_id;

/* Coordinate updates (cpdef methods, bpmodel.pyx:898-1042). */
typedef enum {
    RMX_U_ALLELE_SWAP = 0,    /* update_p_allele_swap   :1025 */
    RMX_U_CN = 1,             /* update_p_cn            :921  */
    RMX_U_BREAKPOINT = 2,     /* update_p_breakpoint    :964  */
    RMX_U_OUTLIER_TOTAL = 3,  /* update_p_outlier_total :987  */
    RMX_U_OUTLIER_ALLELE = 4, /* update_p_outlier_allele:1005 */
    RMX_U_FRAMELOGPROB = 5    /* update_framelogprob    :898  */
} rmx_update_id;

/* Per-(n,s) scalar accessors (cpdef, bpmodel.pyx:686-896). */
typedef enum {
    RMX_E_EXPECTED_TOTAL_READS = 0,           /* :686  -> out[0]      */
    RMX_E_EXPECTED_TOTAL_READS_PARTIAL_H = 1, /* :700  -> out[0..M)   */
    RMX_E_EXPECTED_ALLELE_RATIO = 2,          /* :709  -> out[0]      */
    RMX_E_EXPECTED_ALLELE_RATIO_PARTIAL_H = 3,/* :726  -> out[0..M)   */
    RMX_E_LOG_PRIOR_CN = 4,                   /* :746  -> out[0]      */
    RMX_E_LOG_LIKELIHOOD_TOTAL = 5,           /* :751  (n,s,u)        */
    RMX_E_LOG_LIKELIHOOD_TOTAL_PARTIAL_H = 6, /* :778  (n,s,u)        */
    RMX_E_LOG_LIKELIHOOD_ALLELE = 7,          /* :809  (n,s,v,w)      */
    RMX_E_LOG_LIKELIHOOD_ALLELE_PARTIAL_H = 8 /* :855  (n,s,v,w)      */
} rmx_eval_id;

/* ---- lifetime ---------------------------------------------------------------------- */
/* RemixtModel.__cinit__ (bpmodel.pyx:461-604): validates, uploads, builds derived tables,
 * initial q-factors and the first cached transition expectation. */
int rmx_create(const rmx_desc* desc, rmx_handle* out);
/* G fits sharing desc's data and tables; h_init is (G,M) and divergence_weight (G) (desc->h_init and
 * desc->divergence_weight are ignored).  The returned handle covers all G fits. */
int rmx_batch_create(const rmx_desc* desc, int num_fits, const double* h_init, const double* divergence_weight,
                     rmx_handle* out);
/* A handle on fits [first_fit, first_fit + num_fits) of h's range.  The batch lives until its last handle
 * is destroyed. */
int rmx_batch_view(rmx_handle h, int first_fit, int num_fits, rmx_handle* out);
int rmx_num_fits(rmx_handle h, int* num_fits);
int rmx_destroy(rmx_handle h);
/* Device blocks of destroyed batches are kept for the next batch of the same shapes (cudaMalloc / cudaFree of a
 * whole-genome init grid cost more than its E-step); this returns them to the driver (device < 0: every device). */
int rmx_release_cached_memory(int device);
/* Message of the last failure on this handle (or of rmx_create when h == NULL). */
const char* rmx_last_error(rmx_handle h);
int rmx_device_count(int* count);

/* ---- attribute access (cdef public attributes, bpmodel.pyx:398-456) ------------------ */
/* `count` is the number of ELEMENTS of the host buffer: gn times the attribute's size, fit-major.
 * rmx_set_array also accepts ONE attribute's worth of elements on a multi-fit handle: every fit gets that value.
 * rmx_get_scalar writes gn doubles.  The derived tables are the same for every fit (no leading gn).
 * transition_penalty is fixed at creation; the outlier priors are per batch. */
int rmx_set_array(rmx_handle h, int array_id, const void* src, int64_t count);
int rmx_get_array(rmx_handle h, int array_id, void* dst, int64_t count);
int rmx_set_scalar(rmx_handle h, int scalar_id, double value);
int rmx_get_scalar(rmx_handle h, int scalar_id, double* value);
/* Derived int64 tables of the constructor (bpmodel.pyx:497-507): which = 0 cn_states_total
 * (N1,S,M); 1 num_alleles_subclonal (N1,S); 2 is_hdel (N1,S); 3 is_loh (N1,S). */
int rmx_get_derived(rmx_handle h, int which, int64_t* dst, int64_t count);

/* ---- coordinate updates -------------------------------------------------------------- */
int rmx_update(rmx_handle h, int update_id);
/* n_iter rounds of swap -> cn -> breakpoint -> outlier_total -> outlier_allele, the order of
 * cn_model.py:444-460, on the handle's stream with ONE host synchronisation at the end. */
int rmx_variational_update(rmx_handle h, int n_iter);
/* update_p_breakpoint (bpmodel.pyx:964-985) split around its one reduction, for a fit whose chromosomes are
 * sharded over several models / GPUs: rmx_breakpoint_logq returns this model's partial
 * log q (gn,K,B), the sums of bpmodel.pyx:971-980 over ITS breakend rows; the caller adds the shards' partials
 * (an allreduce of K*B doubles) and hands the total to rmx_breakpoint_apply_logq on every shard, which does the
 * exp-normalise (:982) and refreshes the cached transition expectation (:985). */
int rmx_breakpoint_logq(rmx_handle h, double* logq);
int rmx_breakpoint_apply_logq(rmx_handle h, const double* logq);

/* ---- chromosome-sharded fits over several GPUs ---------------------------------------------
 * A model created with RMX_DESC_SHARD on every rank (its own chromosomes, all K breakpoints) plus an NCCL communicator
 * attached to it behaves as ONE model: the (K,B) log q of update_p_breakpoint (bpmodel.pyx:971-980) is summed over the
 * ranks with ncclAllReduce ON THE MODEL'S STREAM between the shard's partial sums and the normalisation, so
 * rmx_update / rmx_variational_update run their rounds without a host round trip; hmm_log_norm_const, the ELBO terms,
 * calculate_expected_log_likelihood[_partial_h] and rmx_mstep_evaluate return the totals over all shards (the entropy
 * of p_breakpoint, identical on every rank, is counted once).  Every rank must make the same calls in the same order.
 * NCCL is loaded at run time (dlopen of `nccl_library`, or of libnccl.so.2 as the process already has it when NULL):
 * the library itself does not link against it.  rmx_comm_unique_id fills the 128-byte ncclUniqueId that rank 0
 * hands to the others (by any means; the host mirror broadcasts it over torch.distributed). */
int rmx_comm_load(const char* nccl_library);
int rmx_comm_unique_id(void* unique_id_128);
int rmx_comm_init(rmx_handle h, const void* unique_id_128, int rank, int nranks);
int rmx_comm_destroy(rmx_handle h);

/* ---- objective / M-step reductions ---------------------------------------------------- */
int rmx_variational_energy(rmx_handle h, double* energy);    /* bpmodel.pyx:1060 */
int rmx_variational_entropy(rmx_handle h, double* entropy);  /* bpmodel.pyx:1044 */
int rmx_elbo(rmx_handle h, double* elbo);                    /* bpmodel.pyx:1119 */
/* All of the following write gn values (energy / entropy / elbo / E[LL]) or (gn,M) (gradient).
 * calculate_expected_log_likelihood(sample) (bpmodel.pyx:1125); sample is the (gn,N1) int64 mask. */
int rmx_expected_log_likelihood(rmx_handle h, const int64_t* sample, double* out);
/* calculate_expected_log_likelihood_partial_h(sample, partial_h) (bpmodel.pyx:1159); out (M). */
int rmx_expected_log_likelihood_partial_h(rmx_handle h, const int64_t* sample, double* out);
/* A whole 1-D parameter grid with one synchronisation (cn_model.py:553-558 evaluates
 * scipy.optimize.brute's grid one point at a time): values and out are (gn, n_values);
 * out[g][v] = E[LL](sample[g]) of fit g with scalar `scalar_id` set to values[g][v].
 * The stored value of the scalar is left at values[g][n_values-1], which is what the
 * reference's point-by-point evaluation leaves behind (stale-parameter quirk). */
int rmx_expected_log_likelihood_grid(rmx_handle h, const int64_t* sample, int scalar_id,
                                     const double* values, int n_values, double* out);

/* ---- batched M-step (the call pattern of cn_model.py:482-569, served for all fits of a grid at once) --------
 * update_h and update_param draw ONE segment sample per call (cn_model.py:475-480) and evaluate the objective on it
 * some 10-70 times (L-BFGS-B; brute grid + Nelder-Mead).  rmx_mstep_stage_sample uploads the row lists of the
 * handle's fits once: `sample` is the (gn,N1) int64 mask of rmx_expected_log_likelihood, NULL = every segment.
 * rmx_mstep_evaluate then evaluates calculate_expected_log_likelihood (bpmodel.pyx:1125) -- and, when grad != NULL,
 * calculate_expected_log_likelihood_partial_h (:1159) -- on the staged rows with ONE host synchronisation:
 *   active   (gn) uint8 or NULL (= all): fits that take part; the others are left untouched (outputs 0).
 *   h_values (gn,M) or NULL: assigned to h of the active fits first (model.h = h, cn_model.py:484).
 *   scalar_id, values (gn,n_values), n_values: if scalar_id >= 0, ell[g][v] is evaluated with the named likelihood
 *            parameter set to values[g][v], v ascending; the parameter is LEFT at values[g][n_values-1] as the
 *            reference's point-by-point evaluation leaves it (cn_model.py:544).  scalar_id < 0: n_values must be 1.
 *   ell (gn,n_values), grad (gn,M) or NULL. */
int rmx_mstep_stage_sample(rmx_handle h, const int64_t* sample);
/* The same from row lists: rows (gn,max_rows) int32 segment indices, fit g uses its first counts[g] entries, in the
 * order given (ascending order = the order of the mask form, hence the same rounding of the sums). */
int rmx_mstep_stage_rows(rmx_handle h, const int32_t* rows, const int32_t* counts, int max_rows);
int rmx_mstep_evaluate(rmx_handle h, const uint8_t* active, const double* h_values, int scalar_id,
                       const double* values, int n_values, double* ell, double* grad);
/* update_param's optimiser (cn_model.py:533-569: scipy.optimize.brute over `grid`, then scipy.optimize.fmin from the
 * grid minimum) for every active fit of the batch, driven from inside the library: the `n_grid` (<= 32) grid values
 * inside [lo, hi] are evaluated on the staged sample in one synchronisation, then the fits' Nelder-Mead simplices
 * advance in lock step, one evaluation per fit and round (one launch + one synchronisation per round, no host
 * interpreter in between).  Values outside [lo, hi] count as +inf and are never assigned.  Per fit the evaluation
 * points, their order and the optimum are those of the per-fit scipy calls; on return the parameter of every fit holds
 * the LAST value evaluated for it (the state the reference's acceptance test then sees, cn_model.py:561-563).
 * result (gn): the optimum; evals (gn, optional): fmin evaluations per fit; rounds (2, optional): device rounds and
 * objective evaluations spent; trace (gn, trace_cap, 2; optional): the fmin evaluations (x, f) of every fit. */
int rmx_mstep_param_search(rmx_handle h, const uint8_t* active, int scalar_id, double lo, double hi,
                           const double* grid, int n_grid, double* result, int32_t* evals, int64_t* rounds,
                           double* trace, int trace_cap);
/* The segment samples of the M-step for all fits at once, on HOST threads (no device work): for draw d and fit g,
 * numpy.random.RandomState.choice(n, size, replace=False, p=weights[d][g]) of cn_model.py:475-480, consuming the fit's
 * own MT19937 stream exactly as numpy's (frozen) legacy algorithm does.  mt_key (num_fits,624) / mt_pos (num_fits) are the
 * streams' states (RandomState.get_state()), updated in place; weights (num_draws,num_fits,n) with weighted
 * (num_draws,num_fits) flags (0 = uniform, p=None); chosen (num_draws,num_fits,size); status (num_draws,num_fits):
 * 0 ok, 1-5 = the ValueError numpy raises (sum != 1, negative, NaN, fewer non-zero entries than size, size > n),
 * -1 = not drawn because an earlier draw of the fit failed. */
int rmx_host_sample_segments(int num_fits, int num_draws, int n, int size, uint32_t* mt_key, int32_t* mt_pos,
                             const double* weights, const uint8_t* weighted, int64_t* chosen, int32_t* status,
                             int num_threads);
/* setattr(model, name, value) with one value per fit of the handle (values (gn)); active as above. */
int rmx_set_scalar_array(rmx_handle h, int scalar_id, const uint8_t* active, const double* values);

/* ---- decode ----------------------------------------------------------------------------- */
/* infer_cn(cn) (bpmodel.pyx:1197): cn is (gn,N1,M,2) int64, filled in place. */
int rmx_infer_cn(rmx_handle h, int64_t* cn);
/* calculate_log_transmat(out) (bpmodel.pyx:640): out is (N1-1,S,S) f64 filled in place; gn must be 1,
 * as for rmx_eval. */
int rmx_calculate_log_transmat(rmx_handle h, double* out);
int rmx_eval(rmx_handle h, int eval_id, int n, int s, int u, int v, int w, double* out);

/* ---- stateless module functions --------------------------------------------------------- */
/* sum_product(framelogprob, log_transmat, alphas, betas) (bpmodel.pyx:1213). */
int rmx_sum_product(int device, int64_t n_obs, int64_t n_comp, const double* framelogprob,
                    const double* log_transmat, double* alphas, double* betas);
/* sum_product_2paramtrans(framelogprob, alphas, betas, e0, e1) (bpmodel.pyx:1249). */
int rmx_sum_product_2paramtrans(int device, int64_t n_obs, int64_t n_comp, const double* framelogprob,
                                double* alphas, double* betas, double e0, double e1);
/* max_product(framelogprob, log_transmat, state_sequence) -> logprob (bpmodel.pyx:1296). */
int rmx_max_product(int device, int64_t n_obs, int64_t n_comp, const double* framelogprob,
                    const double* log_transmat, int64_t* state_sequence, double* logprob);

/* Allocation-free Viterbi decode.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernel and same results as rmx_max_product, but
 *   - framelogprob (n_obs, n_comp), log_transmat (n_obs - 1, n_comp, n_comp), state_sequence (n_obs int64) and
 *     logprob (1 double) are DEVICE pointers on `device`, owned by the caller;
 *   - the kernel is enqueued on `stream` (a cudaStream_t; NULL = the default stream) and the call returns without
 *     synchronising: the outputs are valid once the stream has reached that point;
 *   - the kernel's scratch (the (n_obs, n_comp) lattice of running maxima) lives in a file-scope device array of
 *     RMX_NOALLOC_MAX_CELLS doubles (16 MiB) that every GPU that loads the library holds until the process exits.
 *     n_obs * n_comp > RMX_NOALLOC_MAX_CELLS (or n_comp > 4096) is rejected with RMX_ERR_ARG and a message.
 * The call performs no cudaMalloc / cudaFree and calls no allocating library routine.  All calls on one device share
 * the array (it keeps the previous call's contents; the kernel writes every cell before it reads it, so nothing
 * needs clearing): calls whose kernels can overlap on the GPU (different streams without a dependency) RACE --
 * serialise them on one stream or with events. */
#define RMX_NOALLOC_MAX_CELLS (1 << 21)
int rmx_max_product_noalloc(int device, void* stream, int64_t n_obs, int64_t n_comp, const double* framelogprob,
                            const double* log_transmat, int64_t* state_sequence, double* logprob);

/* ---- instrumentation ---------------------------------------------------------------------- */
/* Number of kernels this handle has launched since creation. */
int rmx_launch_count(rmx_handle h, int64_t* count);
/* Number of kernels launched by the whole library in this process. */
int rmx_global_launch_count(int64_t* count);
/* Kernel classes of rmx_kernel_timing. */
typedef enum {
    RMX_K_CHAIN = 0, RMX_K_LLTABLE = 1, RMX_K_FRAME = 2, RMX_K_QSUMS = 3, RMX_K_POST = 4, RMX_K_HIST = 5,
    RMX_K_BRK_UPDATE = 6, RMX_K_ELBO_ROWS = 7, RMX_K_PAIRSUM = 8, RMX_K_ELL = 9, RMX_K_ROUND = 10, RMX_K_COUNT_ = 11
} rmx_kernel_class;
/* Device-side duration (CUDA events on the batch's stream around every launch) of each kernel class since
 * the last call: total milliseconds and number of launches, arrays of RMX_K_COUNT_; then switches the
 * timing on or off.  Durations are collected at the next synchronising call. */
int rmx_kernel_timing(rmx_handle h, int enable, double* total_ms, int64_t* launches);
/* The same for the chain-message kernel alone (scalars). */
int rmx_chain_timing(rmx_handle h, int enable, double* total_ms, int64_t* launches);
int rmx_synchronize(rmx_handle h);
/* The cudaStream_t the handle launches on (for CUDA-event timing by the caller). */
int rmx_stream(rmx_handle h, void** stream);
/* CUDA-event timing on the handle's stream: record event `slot` (0..15); elapsed ms between two
 * recorded events (possibly of two handles on one device); both must have completed. */
int rmx_event_record(rmx_handle h, int slot);
int rmx_event_elapsed(rmx_handle h0, int slot0, rmx_handle h1, int slot1, double* ms);
/* Measured FP64 FMA peak of the device (dependent-chain-free DFMA micro-benchmark, best of 5),
 * in TFLOP/s: the denominator of the FP64 roofline. */
int rmx_fp64_peak(int device, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* REMIXT_B200_H */
