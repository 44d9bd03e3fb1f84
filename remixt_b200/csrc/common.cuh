// Shared device-side definitions for the remixt_b200 kernels (sm_100a).
//
// Notation follows the reference (remixt/bpmodel.pyx): N segments, S copy-number states,
// M clones (clone 0 = normal), B breakpoint states, K breakpoints, two alleles.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define RMX_MAXM 8            // max clones (incl. normal)
#define RMX_NTERM 6           // LL table planes: total u=0,1 ; allele (v,w) = 00,01,10,11

// device error bits (checked at host sync points)
#define RMX_F_NB_NAN        (1 << 0)   // bpmodel.pyx:269
#define RMX_F_BB_P_RANGE    (1 << 1)   // bpmodel.pyx:335 / :383
#define RMX_F_BB_NAN        (1 << 2)   // bpmodel.pyx:344
#define RMX_F_TOTAL_DEPTH   (1 << 3)   // bpmodel.pyx:721 / :738
#define RMX_F_LOH_P         (1 << 4)   // bpmodel.pyx:829
#define RMX_F_FRAME_NAN     (1 << 5)   // bpmodel.pyx:936
#define RMX_F_MSG_NAN       (1 << 6)   // bpmodel.pyx:943-944
#define RMX_F_POST_NAN      (1 << 7)   // bpmodel.pyx:952
#define RMX_F_JOINT_NAN     (1 << 8)   // bpmodel.pyx:962
#define RMX_F_NB_PARTIAL_NAN (1 << 9)  // bpmodel.pyx:298
#define RMX_F_BB_PARTIAL_NAN (1 << 10) // bpmodel.pyx:391
#define RMX_F_DIGAMMA_X     (1 << 11)  // bpmodel.pyx:207

struct RmxDims {
    int N, S, M, B, K;
    int cn_max;     // max(cn_states.max(), brk_states.max())  (bpmodel.pyx:489)
    int D;          // 2*cn_max+3 copy-number-difference slots, d in [-cn_max-1, cn_max+1]
    int doff;       // cn_max+1 : slot = d + doff
};

// Likelihood parameters, passed to kernels by value (bpmodel.pyx:442-456, :573-597).
struct RmxParams {
    double h[RMX_MAXM];
    double negbin_r[2];
    double negbin_hdel_mu;
    double negbin_hdel_r[2];
    double betabin_M[2];
    double betabin_loh_p;
    double betabin_loh_M[2];
    double divergence_weight;
    int normal_contamination;
};

// A launch covers the fits [g0, g0 + gridDim.y) of a batch; every per-fit array is laid out with the
// fit index as its leading dimension.
#define RMX_FIT() (g0 + (int)blockIdx.y)

// Some launches address an explicit LIST of fits instead (blockIdx.y -> list entry), each fit with its own mask of
// likelihood term groups: the full-data evaluations of the M-step touch whichever fits have stale table planes, and
// one launch per maximal run of equal masks fragments into tens of small launches.  n == 0: the range form above.
#define RMX_SEL_MAX 256
struct RmxFitSel {
    int n;
    unsigned char fit[RMX_SEL_MAX];
    unsigned char planes[RMX_SEL_MAX];
};
#define RMX_FIT_SEL(sel) ((sel).n ? (int)(sel).fit[blockIdx.y] : g0 + (int)blockIdx.y)
#define RMX_PLANES_SEL(sel, all_) ((sel).n ? (int)(sel).planes[blockIdx.y] : (all_))

// Read-only tables in HBM, shared by every fit of a batch (the fits of an initialisation grid see the
// same data and state space; only h, the dispersions, divergence_weight and the variational state differ).
struct RmxTables {
    const int8_t* tab;      // (S,M,2) clone copy number per state; clone 0 row is zero
    const int8_t* nrm;      // (N,2)  normal copies per segment (cn_states[n,:,0,:])
    const int8_t* nsub;     // (S)    num_alleles_subclonal
    const double* l;        // (N)
    const double* x;        // (N)
    const double* y0;       // (N)  y[:,0]
    const double* y1;       // (N)  y[:,1]
    const uint8_t* tmask;   // (G,N)  total_likelihood_mask != 0   (per fit)
    const uint8_t* amask;   // (G,N)  allele_likelihood_mask != 0  (per fit)
};

// Inputs of the likelihood table's fast path (likelihood.cu: k_statetab, k_lltable2)
struct RmxFast {
    const uint8_t* nkind;     // (N) index of the segment's normal-copy pair among the distinct pairs
    const int8_t* kind_nrm;   // (Q,2) the distinct normal-copy pairs
    int Q;
    const double2* logw;      // (RMX_LW_SIZE) wide log table
    double* stab;             // (G)(8,Q,S) per-(fit, kind, state) quantities
    uint8_t* sflag;           // (G)(Q,S)   bit 0: total-read fast path valid, bit 1: allele fast path valid
    double* segf;             // (G)(N,12)  per-(fit, segment) inputs of the fast path (k_segfast)
};

// The opt-in to more than 48 KB of dynamic shared memory is a per-DEVICE attribute of a kernel: a flag per process
// would leave a second device of the same process without it (models take `device=`).  One bit per device ordinal.
struct RmxPerDeviceOnce {
    unsigned long long done = 0;
    bool need() {
        int dev = 0;
        cudaGetDevice(&dev);
        const unsigned long long bit = 1ull << (dev & 63);
        if (done & bit) return false;
        done |= bit;
        return true;
    }
};

// ---------------------------------------------------------------------------------------------
// warp / block helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------------------------------------
// per-(segment,state) quantities
// ---------------------------------------------------------------------------------------------
struct StateCN {
    int tot[RMX_MAXM];   // total copies per clone (clone 0 = normal of this segment)
    int a0[RMX_MAXM];    // allele-0 copies per clone
    bool hdel;           // all copies zero (bpmodel.pyx:506)
    bool loh;            // some allele has zero copies in every clone (bpmodel.pyx:507)
};

__device__ __forceinline__ StateCN load_state2(const int8_t* __restrict__ tab, const RmxDims& d, int n0, int n1, int s);
__device__ __forceinline__ StateCN load_state(const RmxTables& T, const RmxDims& d, int n, int s) {
    return load_state2(T.tab, d, T.nrm[2 * n], T.nrm[2 * n + 1], s);
}
// tab may point to shared memory (a staged copy of the state table)
__device__ __forceinline__ StateCN load_state2(const int8_t* __restrict__ tab, const RmxDims& d, int n0, int n1, int s) {
    StateCN c;
    c.tot[0] = n0 + n1;
    c.a0[0] = n0;
    int s0 = n0, s1 = n1;
#pragma unroll
    for (int m = 1; m < RMX_MAXM; ++m) {
        if (m < d.M) {
            int v0 = tab[(s * d.M + m) * 2], v1 = tab[(s * d.M + m) * 2 + 1];
            c.tot[m] = v0 + v1;
            c.a0[m] = v0;
            s0 += v0;
            s1 += v1;
        } else {
            c.tot[m] = 0;
            c.a0[m] = 0;
        }
    }
    c.hdel = (s0 + s1) == 0;
    c.loh = (s0 == 0) || (s1 == 0);
    return c;
}

// Expected total reads (bpmodel.pyx:686-698), reference summation order.
__device__ __forceinline__ double expected_total_reads(const RmxParams& P, const RmxDims& d,
                                                       const StateCN& c, double l) {
    double mu = 0.;
#pragma unroll
    for (int m = 0; m < RMX_MAXM; ++m)
        if (m < d.M) mu += P.h[m] * (double)c.tot[m];
    return mu * l;
}

// Expected allele ratio (bpmodel.pyx:709-724).  Returns total_depth through *td.
__device__ __forceinline__ double expected_allele_ratio(const RmxParams& P, const RmxDims& d,
                                                        const StateCN& c, double* md, double* td) {
    double minor = 0., total = 0.;
#pragma unroll
    for (int m = 0; m < RMX_MAXM; ++m)
        if (m < d.M) {
            minor += P.h[m] * (double)c.a0[m];
            total += P.h[m] * (double)c.tot[m];
        }
    *md = minor;
    *td = total;
    return minor / total;
}

// ---------------------------------------------------------------------------------------------
// log / lgamma for the likelihood table.  CUDA's double log costs ~54 FP64-pipe slots and lgamma 85-260
// (measured, scripts/micro/lgam_rate.cu); the likelihood kernel is FP64-pipe bound and calls them 16 times per
// (segment, state).  rmx_log: argument reduction to m in [0.75, 1.5), a 96-entry table of (1/c, -log(1/c)) at the
// bucket centres c (shared memory, built per CTA with the library log) and a degree-7 series in r = m/c - 1,
// |r| <= 5.3e-3: ~12 FP64 operations, absolute error < 3e-16 * max(1, |log x|).  rmx_lgamma_pos: Stirling series
// with 7 terms above 8 (truncation < 1e-15), upward recurrence below; agrees with glibc's lgamma to 8e-15 *
// max(1, |lgamma|) over [1e-3, 1.6e5] (scripts/micro/lgam_test.cpp).  Positive finite normal arguments only.
// ---------------------------------------------------------------------------------------------
#define RMX_LOGTAB 96
__device__ __forceinline__ void rmx_logtab_init(double2* tab) {      // all threads call; includes the barrier
    for (int i = threadIdx.x; i < RMX_LOGTAB; i += blockDim.x) {
        const double c = 0.75 + ((double)i + 0.5) * (1.0 / 128.0);
        const double inv = 1.0 / c;
        tab[i] = make_double2(inv, -log(inv));
    }
    __syncthreads();
}
__device__ __forceinline__ double rmx_log(double x, const double2* __restrict__ tab) {
    int hi = __double2hiint(x);
    if (hi < 0x00100000 || hi >= 0x7ff00000) return log(x);   // zero, denormal, negative, inf, NaN: library semantics
    const int lo = __double2loint(x);
    int e = (hi >> 20) - 1023;
    int mh = (hi & 0x000fffff) | 0x3ff00000;                 // m in [1, 2)
    if (mh >= 0x3ff80000) { mh -= 0x00100000; e += 1; }      // m >= 1.5: halve -> m in [0.75, 1.5)
    const int idx = mh < 0x3ff00000 ? ((mh >> 14) & 63) - 32 : 32 + ((mh >> 13) & 127);
    const double m = __hiloint2double(mh, lo);
    const double2 t = tab[idx];
    const double r = fma(m, t.x, -1.0);
    double p = 1.0 / 7.0;
    p = fma(p, r, -1.0 / 6.0);
    p = fma(p, r, 0.2);
    p = fma(p, r, -0.25);
    p = fma(p, r, 1.0 / 3.0);
    p = fma(p, r, -0.5);
    p = fma(p * r, r, r);
    const double ed = (double)e;
    return fma(ed, 6.93147180369123816490e-01, t.y) + fma(ed, 1.90821492927058770002e-10, p);
}
// 1/z for positive normal z well inside the exponent range: hardware seed + two Newton steps (5 instructions)
__device__ __forceinline__ double rmx_rcp(double z) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(z));
    double e = fma(-z, r, 1.0);
    r = fma(r, e, r);
    e = fma(-z, r, 1.0);
    return fma(r, e, r);
}
__device__ __forceinline__ double rmx_lgamma_pos(double z, const double2* __restrict__ tab) {
    double lp = 0.;
    if (z < 8.) {
        double p = z;
        z += 1.;
        while (z < 8.) { p *= z; z += 1.; }
        lp = rmx_log(p, tab);
    }
    // z in [8, 1e300): the seed's exponent handling is safe; larger arguments go through the library
    if (!(z < 1e300)) return lgamma(z);
    const double r = rmx_rcp(z), r2 = r * r;
    double s = 6.4102564102564102564e-3;                 //  1/156
    s = fma(s, r2, -1.9175269175269175269e-3);           // -691/360360
    s = fma(s, r2, 8.4175084175084175084e-4);            //  1/1188
    s = fma(s, r2, -5.9523809523809523810e-4);           // -1/1680
    s = fma(s, r2, 7.9365079365079365079e-4);            //  1/1260
    s = fma(s, r2, -2.7777777777777777778e-3);           // -1/360
    s = fma(s, r2, 8.3333333333333333333e-2);            //  1/12
    return (z - 0.5) * rmx_log(z, tab) - z + 0.91893853320467274178 + s * r - lp;
}

// ---------------------------------------------------------------------------------------------
// Wide log table (fast path of the likelihood table).  One entry per (binary exponent, top 7 mantissa bits) for
// arguments in [2^RMX_LW_E0, 2^(RMX_LW_E0+RMX_LW_NOCT)): (1/c, -log(1/c)) at the bucket centre c, built on the host
// in long double (api.cu) and copied to shared memory by each CTA.  log z = -log(1/c) + log1p(z/c - 1) with
// |z/c - 1| <= 2^-8: no exponent / mantissa surgery, one FMA for the reduced argument, a degree-6 series
// (truncation < 2e-18).  8 FP64 operations per log.  Arguments outside the table take the library.
// ---------------------------------------------------------------------------------------------
#define RMX_LW_E0 4
#define RMX_LW_NOCT 18
#define RMX_LW_SIZE (RMX_LW_NOCT * 128)
#define RMX_LW_BASE ((1023 + RMX_LW_E0) << 7)
__device__ __forceinline__ bool rmx_lw_in(double z, int* idx) {
    *idx = (__double2hiint(z) >> 13) - RMX_LW_BASE;
    return (unsigned)*idx < (unsigned)RMX_LW_SIZE;
}
__device__ __forceinline__ double rmx_lw_log(double z, const double2 t) {
    const double r = fma(z, t.x, -1.0);
    double p = -1.0 / 6.0;
    p = fma(p, r, 0.2);
    p = fma(p, r, -0.25);
    p = fma(p, r, 1.0 / 3.0);
    p = fma(p, r, -0.5);
    p = fma(p * r, r, r);
    return t.y + p;
}
// Stirling core of lgamma for z in the table range (z >= 16): (z - 1/2) log z + series(1/z), i.e. lgamma(z) + z -
// log(2 pi)/2.  Five series terms (next term 1.9e-3 / z^11 < 1.1e-16 at 16); 1/z from the hardware seed and one
// Newton step (relative 1e-12 on a term below 5.3e-3).  18 FP64 operations.
__device__ __forceinline__ double rmx_lw_lgcore(double z, const double2 t) {
    const double L = rmx_lw_log(z, t);
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(z));
    r = fma(r, fma(-z, r, 1.0), r);
    const double r2 = r * r;
    double s = 8.4175084175084175084e-4;                 //  1/1188
    s = fma(s, r2, -5.9523809523809523810e-4);           // -1/1680
    s = fma(s, r2, 7.9365079365079365079e-4);            //  1/1260
    s = fma(s, r2, -2.7777777777777777778e-3);           // -1/360
    s = fma(s, r2, 8.3333333333333333333e-2);            //  1/12
    return fma(z - 0.5, L, s * r);
}
// the same quantity for any positive finite z (library log, upward recurrence below 16): the out-of-table branch
static __device__ __noinline__ double rmx_lgcore_slow(double z) {
    double lp = 0.;
    const double z0 = z;
    if (z < 16.) {
        double p = z;
        z += 1.;
        while (z < 16.) { p *= z; z += 1.; }
        lp = log(p);
    }
    if (!(z < 1e300)) return lgamma(z0) + z0 - 0.91893853320467274178;
    const double r = 1.0 / z, r2 = r * r;
    double s = 8.4175084175084175084e-4;
    s = fma(s, r2, -5.9523809523809523810e-4);
    s = fma(s, r2, 7.9365079365079365079e-4);
    s = fma(s, r2, -2.7777777777777777778e-3);
    s = fma(s, r2, 8.3333333333333333333e-2);
    // lgamma(z0) = lgamma(z) - lp ; core(z0) = lgamma(z0) + z0 - c
    return fma(z - 0.5, log(z), s * r) - z - lp + z0;
}
__device__ __forceinline__ double rmx_lgcore(double z, const double2* __restrict__ lw) {
    int idx;
    if (rmx_lw_in(z, &idx)) return rmx_lw_lgcore(z, lw[idx]);
    return rmx_lgcore_slow(z);
}
// W Stirling cores at once: one range test for all of them, then straight-line code the scheduler can interleave
// (W independent dependency chains per thread; a branch per evaluation serialises them).
template <int W>
__device__ __forceinline__ void rmx_lgcore_n(const double (&z)[W], const double2* __restrict__ lw, double (&out)[W]) {
    int idx[W];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < W; ++i) ok = rmx_lw_in(z[i], &idx[i]) && ok;
    if (ok) {
        double2 t[W];
        double r[W], p[W], q[W], s[W];
#pragma unroll
        for (int i = 0; i < W; ++i) t[i] = lw[idx[i]];
#pragma unroll
        for (int i = 0; i < W; ++i) asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(q[i]) : "d"(z[i]));
#pragma unroll
        for (int i = 0; i < W; ++i) r[i] = fma(z[i], t[i].x, -1.0);
#pragma unroll
        for (int i = 0; i < W; ++i) q[i] = fma(q[i], fma(-z[i], q[i], 1.0), q[i]);
#pragma unroll
        for (int i = 0; i < W; ++i) p[i] = fma(-1.0 / 6.0, r[i], 0.2);
#pragma unroll
        for (int i = 0; i < W; ++i) s[i] = q[i] * q[i];
#pragma unroll
        for (int i = 0; i < W; ++i) p[i] = fma(p[i], r[i], -0.25);
#pragma unroll
        for (int i = 0; i < W; ++i) out[i] = fma(8.4175084175084175084e-4, s[i], -5.9523809523809523810e-4);
#pragma unroll
        for (int i = 0; i < W; ++i) p[i] = fma(p[i], r[i], 1.0 / 3.0);
#pragma unroll
        for (int i = 0; i < W; ++i) out[i] = fma(out[i], s[i], 7.9365079365079365079e-4);
#pragma unroll
        for (int i = 0; i < W; ++i) p[i] = fma(p[i], r[i], -0.5);
#pragma unroll
        for (int i = 0; i < W; ++i) out[i] = fma(out[i], s[i], -2.7777777777777777778e-3);
#pragma unroll
        for (int i = 0; i < W; ++i) p[i] = fma(p[i] * r[i], r[i], r[i]);
#pragma unroll
        for (int i = 0; i < W; ++i) out[i] = fma(out[i], s[i], 8.3333333333333333333e-2);
#pragma unroll
        for (int i = 0; i < W; ++i) out[i] = fma(z[i] - 0.5, t[i].y + p[i], out[i] * q[i]);
    } else {
#pragma unroll
        for (int i = 0; i < W; ++i) out[i] = rmx_lgcore_slow(z[i]);
    }
}
__device__ __forceinline__ double rmx_wlog(double z, const double2* __restrict__ lw) {
    int idx;
    if (rmx_lw_in(z, &idx)) return rmx_lw_log(z, lw[idx]);
    return log(z);
}

// ---------------------------------------------------------------------------------------------
// Arithmetic shared by the separate update kernels and the fused round kernel (likelihood.cu: k_round), written
// with explicit roundings so that both paths give the same bits whatever the compiler would contract.
// ---------------------------------------------------------------------------------------------
// framelogprob term of one (segment, state) (bpmodel.pyx:898-919); l = the six likelihood terms
__device__ __forceinline__ double rmx_frame_value(double ot0, double ot1, double oa0, double oa1, double sw0, double sw1,
                                                  double l0, double l1, double l2, double l3, double l4, double l5,
                                                  double nsub, double ln, double dw) {
    double f = __dmul_rn(ot0, l0);
    f = fma(ot1, l1, f);
    f = fma(__dmul_rn(oa0, sw0), l2, f);
    f = fma(__dmul_rn(oa0, sw1), l3, f);
    f = fma(__dmul_rn(oa1, sw0), l4, f);
    f = fma(__dmul_rn(oa1, sw1), l5, f);
    return __dadd_rn(f, __dmul_rn(__dmul_rn(-nsub, ln), dw));      // bpmodel.pyx:746-749
}
// exp-normalise of a 2-vector (bpmodel.pyx:120-128: exp(x - logsumexp(x)), then divided by its sum).  In exact
// arithmetic that is the 2-way softmax; it is evaluated as 1 / (1 + t), t / (1 + t) with t = exp(-|a - b|): one exp
// and two divisions instead of four exps, a log and two divisions, and within 2 ulp of the reference's sequence.
__device__ __forceinline__ void rmx_exp_normalize2(double a, double b, double* out) {
    const double dlt = __dadd_rn(b, -a);
    if (isnan(dlt)) {                 // NaN or inf - inf: the reference's sequence gives NaN as well
        out[0] = out[1] = dlt;
        return;
    }
    const double t = exp(-fabs(dlt));
    const double z = __dadd_rn(1., t);
    const double big = __ddiv_rn(1., z), small = __ddiv_rn(t, z);
    out[0] = dlt <= 0. ? big : small;
    out[1] = dlt <= 0. ? small : big;
}
// Two independent 2-way exp-normalisations, (a0, a1) -> oa and (b0, b1) -> ob, evaluated by a whole warp whose lanes all
// hold the same arguments (k_round: every lane has the segment's sums) with ONE exp and ONE division instead of two
// and four: lane bit 0 picks the problem, lane bit 1 the quotient (1 / z or t / z), four shuffles hand every lane all
// four quotients.  Each value goes through the operations of rmx_exp_normalize2 on the same inputs: same bits.
__device__ __forceinline__ void rmx_exp_normalize2_x2(double a0, double a1, double b0, double b1, int lane, double* oa,
                                                      double* ob) {
    const double dA = __dadd_rn(a1, -a0), dB = __dadd_rn(b1, -b0);
    const double dl = (lane & 1) ? dB : dA;
    const double t = exp(-fabs(dl));
    const double z = __dadd_rn(1., t);
    const double quo = __ddiv_rn((lane & 2) ? t : 1., z);
    const double bigA = __shfl_sync(0xffffffffu, quo, 0), bigB = __shfl_sync(0xffffffffu, quo, 1);
    const double smallA = __shfl_sync(0xffffffffu, quo, 2), smallB = __shfl_sync(0xffffffffu, quo, 3);
    if (isnan(dA)) { oa[0] = oa[1] = dA; }
    else { oa[0] = dA <= 0. ? bigA : smallA; oa[1] = dA <= 0. ? smallA : bigA; }
    if (isnan(dB)) { ob[0] = ob[1] = dB; }
    else { ob[0] = dB <= 0. ? bigB : smallB; ob[1] = dB <= 0. ? smallB : bigB; }
}
// arguments of the three 2-way factor updates from the sums A[0..6) (bpmodel.pyx:987-1042)
// lp0 = log(1 - prior), lp1 = log(prior), evaluated once on the host
__device__ __forceinline__ void rmx_q_args(int which, const double* A, double q0, double q1, double lp0, double lp1,
                                           double* a) {
    if (which == 1) {            // outlier total
        a[0] = __dadd_rn(A[0], lp0);
        a[1] = __dadd_rn(A[1], lp1);
    } else if (which == 0) {     // allele swap; q = p_outlier_allele
        a[0] = fma(q1, A[4], __dmul_rn(q0, A[2]));
        a[1] = fma(q1, A[5], __dmul_rn(q0, A[3]));
    } else {                     // outlier allele; q = p_allele_swap
        a[0] = __dadd_rn(fma(q1, A[3], __dmul_rn(q0, A[2])), lp0);
        a[1] = __dadd_rn(fma(q1, A[5], __dmul_rn(q0, A[4])), lp1);
    }
}

// AS 103 digamma (J. Bernardo 1976), the algorithm bpmodel.pyx:162-235 uses.
__device__ __forceinline__ double digamma_as103(double x, int* flags) {
    if (x <= 0.0) {
        *flags |= RMX_F_DIGAMMA_X;
        return 0.;
    }
    if (x <= 0.000001) return -0.57721566490153286060 - 1.0 / x + 1.6449340668482264365 * x;
    double value = 0.0;
    while (x < 8.5) {
        value -= 1.0 / x;
        x += 1.0;
    }
    double r = 1.0 / x;
    value += log(x) - 0.5 * r;
    r = r * r;
    value -= r * (1.0 / 12.0 - r * (1.0 / 120.0 - r * (1.0 / 252.0 - r * (1.0 / 240.0 - r * (1.0 / 132.0)))));
    return value;
}
