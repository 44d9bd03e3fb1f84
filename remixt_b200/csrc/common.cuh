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

__device__ __forceinline__ StateCN load_state(const RmxTables& T, const RmxDims& d, int n, int s) {
    StateCN c;
    int n0 = T.nrm[2 * n], n1 = T.nrm[2 * n + 1];
    c.tot[0] = n0 + n1;
    c.a0[0] = n0;
    int s0 = n0, s1 = n1;
#pragma unroll
    for (int m = 1; m < RMX_MAXM; ++m) {
        if (m < d.M) {
            int v0 = T.tab[(s * d.M + m) * 2], v1 = T.tab[(s * d.M + m) * 2 + 1];
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
