// Likelihood kernels: the per-(segment,state) negative-binomial / beta-binomial expected
// log-likelihood table, the frame log-probability, the 2-way q-factor posteriors
// (allele swap, outliers), and the M-step / ELBO reductions.  Every kernel serves a range of
// fits of a batch in one launch (blockIdx.y = fit within the range).
//
// Reference: remixt/bpmodel.pyx
//   negbin_log_likelihood :238-272      betabin_log_likelihood :304-347
//   calculate_log_likelihood_total :751  calculate_log_likelihood_allele :809
//   update_framelogprob :898            update_p_outlier_total :987
//   update_p_outlier_allele :1005       update_p_allele_swap :1025
//   calculate_expected_log_likelihood :1125   ..._partial_h :1159
#include "common.cuh"
#include "kernels.h"

// ---------------------------------------------------------------------------------------------
// Per-segment hoisted constants.  The state-independent lgamma terms of the two pmfs:
//   NB (u):    lgamma(x+r_u) - lgamma(x+1) - lgamma(r_u)
//   BB (v,w):  lgamma(n+1) - lgamma(k_w+1) - lgamma(n-k_w+1) - lgamma(n+M_v) + lgamma(M_v)
// plus the same with the hdel / LOH dispersions (only used when normal_contamination is off).
// Layout per fit: segc[c * R + r], c in [0,12), r = position in the row list; fit stride given.
// ---------------------------------------------------------------------------------------------
// (kernel: k_segconst_rows below)

// Six log-likelihood terms of one (segment, state): ll[0..1] total (u), ll[2..5] allele (v*2+w).
// Per-segment inputs of ll_core: either read straight from HBM (ll_terms) or staged in shared memory (k_lltable).
struct SegIn {
    const double* segc;   // hoisted constants c = 0..11 at segc[c * sstride]
    int sstride;
    double x, l, y0, y1;
    int nrm0, nrm1;
};

__device__ __forceinline__ void ll_core(const int8_t* __restrict__ tab, const RmxDims& d, const RmxParams& P, bool tmask,
                                        bool amask, const SegIn& in, int s, double* ll, int* flags,
                                        const double2* __restrict__ ltab, const double* lnr = nullptr, int planes = 15) {
    StateCN c = load_state2(tab, d, in.nrm0, in.nrm1, s);
    const bool special = !P.normal_contamination;
    // planes: bit u = total-read term u, bit 2 + v = the two allele terms of v; terms left out are returned as 0
    // (the M-step objective of one dispersion parameter moves only the terms that parameter enters)
    // ---- total reads: negative binomial (bpmodel.pyx:751-776, :238-272)
    if (tmask && (planes & 3)) {
        double x = in.x;
        bool hd = special && c.hdel;
        double mu = hd ? P.negbin_hdel_mu : expected_total_reads(P, d, c, in.l);
        // x log p + r log(1-p) with p = mu/(r+mu) is x (log mu - log(r+mu)) + r (log r - log(r+mu)): three logs for
        // the two dispersions instead of four and no division; log r comes hoisted (lnr).  Only for 0 < mu < inf
        // (p strictly inside (0,1)); anything else takes the reference's literal form, clamp included.
        const bool plain = lnr != nullptr && mu > 0. && mu < 1e300;
        const double lmu = plain ? rmx_log(mu, ltab) : 0.;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            if (!((planes >> u) & 1)) { ll[u] = 0.; continue; }
            double rr = hd ? P.negbin_hdel_r[u] : P.negbin_r[u];
            double v;
            if (plain) {
                const double lrm = rmx_log(rr + mu, ltab);
                v = in.segc[((hd ? 2 : 0) + u) * in.sstride] + x * (lmu - lrm) + rr * (lnr[(hd ? 2 : 0) + u] - lrm);
            } else {
                double p = mu / (rr + mu);
                if (p < 0. || p > 1.) p = 0.5;
                v = in.segc[((hd ? 2 : 0) + u) * in.sstride] + x * rmx_log(p, ltab) + rr * rmx_log(1. - p, ltab);
            }
            if (isnan(v)) *flags |= RMX_F_NB_NAN;
            ll[u] = v;
        }
    } else {
        ll[0] = 0.;
        ll[1] = 0.;
    }
    // ---- allele reads: beta binomial (bpmodel.pyx:809-853, :304-347)
    ll[2] = ll[3] = ll[4] = ll[5] = 0.;
    if (amask && (planes & 12)) {
        double p;
        if (c.hdel) {
            p = 0.;
        } else {
            double md, td;
            p = expected_allele_ratio(P, d, c, &md, &td);
            if (td <= 0.) {
                *flags |= RMX_F_TOTAL_DEPTH;
                return;
            }
        }
        bool loh = special && c.loh;
        if (loh) {
            if (p == 0.) p = P.betabin_loh_p;
            else if (p == 1.) p = 1. - P.betabin_loh_p;
            else {
                *flags |= RMX_F_LOH_P;
                return;
            }
        }
        double y0 = in.y0, y1 = in.y1;
        double na = y0 + y1;
        if (na == 0.) return;
        if (p <= 0. || (1. - p) <= 0.) {
            *flags |= RMX_F_BB_P_RANGE;
            return;
        }
#pragma unroll
        for (int v = 0; v < 2; ++v) {
            if (!((planes >> (2 + v)) & 1)) continue;
            double Mv = loh ? P.betabin_loh_M[v] : P.betabin_M[v];
            double a = Mv * p, b = Mv * (1. - p);
            double base = -rmx_lgamma_pos(a, ltab) - rmx_lgamma_pos(b, ltab);
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                double k = w ? y1 : y0;
                double val = in.segc[((loh ? 8 : 4) + v * 2 + w) * in.sstride] + rmx_lgamma_pos(k + a, ltab) + rmx_lgamma_pos(na - k + b, ltab) + base;
                if (isnan(val)) *flags |= RMX_F_BB_NAN;
                ll[2 + v * 2 + w] = val;
            }
        }
    }
}

// the same, per-segment inputs read from HBM (row lists of the M-step)
__device__ __forceinline__ void ll_terms(const RmxTables& T, const RmxDims& d, const RmxParams& P, bool tmask,
                                         bool amask, const double* __restrict__ segc, int R, int r, int n, int s,
                                         double* ll, int* flags, const double2* __restrict__ ltab, int planes = 15) {
    SegIn in;
    in.segc = segc + r;
    in.sstride = R;
    in.x = T.x[n]; in.l = T.l[n]; in.y0 = T.y0[n]; in.y1 = T.y1[n];
    in.nrm0 = T.nrm[2 * n]; in.nrm1 = T.nrm[2 * n + 1];
    ll_core(T.tab, d, P, tmask, amask, in, s, ll, flags, ltab, nullptr, planes);
}

// K1: the likelihood table, LL[g][t][n*S+s], one thread per (n,s), fully coalesced planar writes.  A CTA takes
// 256 consecutive (n,s) at a time -- a handful of segments -- and stages their per-segment inputs (12 hoisted
// constants, x, l, y, masks, normal copies) and the state table in shared memory, so a thread's ~20 per-segment
// reads are one cooperative round trip instead of 20 dependent broadcast loads.
#define LLT_SEGW 16      // doubles per staged segment: 12 constants, x, l, y0, y1
__global__ void __launch_bounds__(256, 3) k_lltable(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                                  const double* __restrict__ segc_all,
                                                  double* __restrict__ LL_all, int* __restrict__ dflags, int segmax) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double2 ltab[RMX_LOGTAB];
    double* segs = reinterpret_cast<double*>(smem_raw);                 // [segmax][LLT_SEGW]
    int* segi = reinterpret_cast<int*>(segs + (size_t)segmax * LLT_SEGW);   // [segmax] masks + normal copies, packed
    int8_t* stab = reinterpret_cast<int8_t*>(segi + segmax);            // [S*M*2]
    const int t = threadIdx.x;
    for (int e = t; e < d.S * d.M * 2; e += blockDim.x) stab[e] = T.tab[e];
    rmx_logtab_init(ltab);
    const int g = RMX_FIT();
    const RmxParams P = Pg[g];
    const int NS = d.N * d.S;                                           // per fit, < 2^31 (checked at creation)
    const double* segc = segc_all + (int64_t)g * 12 * d.N;
    double* LL = LL_all + (int64_t)g * RMX_NTERM * NS;
    const uint8_t* tm = T.tmask + (int64_t)g * d.N;
    const uint8_t* am = T.amask + (int64_t)g * d.N;
    int flags = 0;
    // log of the four negative-binomial dispersions of this fit (plain u = 0, 1; hdel u = 0, 1)
    const double lnr[4] = {log(P.negbin_r[0]), log(P.negbin_r[1]), log(P.negbin_hdel_r[0]), log(P.negbin_hdel_r[1])};
    for (int base = blockIdx.x * blockDim.x; base < NS; base += gridDim.x * blockDim.x) {
        const int nfirst = base / d.S;
        const int last = min(base + (int)blockDim.x, NS) - 1;
        const int nseg = last / d.S - nfirst + 1;
        for (int e = t; e < nseg * LLT_SEGW; e += blockDim.x) {
            const int sg = e / LLT_SEGW, c = e - sg * LLT_SEGW, n = nfirst + sg;
            double v;
            if (c < 12) v = segc[(int64_t)c * d.N + n];
            else if (c == 12) v = T.x[n];
            else if (c == 13) v = T.l[n];
            else if (c == 14) v = T.y0[n];
            else v = T.y1[n];
            segs[e] = v;
        }
        for (int sg = t; sg < nseg; sg += blockDim.x) {
            const int n = nfirst + sg;
            segi[sg] = (tm[n] != 0 ? 1 : 0) | (am[n] != 0 ? 2 : 0) | ((int)T.nrm[2 * n] << 8) | ((int)T.nrm[2 * n + 1] << 16);
        }
        __syncthreads();
        const int idx = base + t;
        if (idx < NS) {
            const int n = idx / d.S, s = idx - n * d.S, sg = n - nfirst;
            const double* sp = segs + sg * LLT_SEGW;
            const int si = segi[sg];
            SegIn in;
            in.segc = sp;
            in.sstride = 1;
            in.x = sp[12]; in.l = sp[13]; in.y0 = sp[14]; in.y1 = sp[15];
            in.nrm0 = (si >> 8) & 0xff; in.nrm1 = (si >> 16) & 0xff;
            double ll[RMX_NTERM];
            ll_core(stab, d, P, (si & 1) != 0, (si & 2) != 0, in, s, ll, &flags, ltab, lnr);
#pragma unroll
            for (int k = 0; k < RMX_NTERM; ++k) LL[(int64_t)k * NS + idx] = ll[k];
        }
        __syncthreads();
    }
    if (flags) atomicOr(dflags, flags);
}

// ---------------------------------------------------------------------------------------------
// K1 fast path.  Everything of a (segment, state) term that depends only on (fit, normal-copy kind, state) is
// tabulated once per likelihood-table build by k_statetab: haploid depth sum_m h_m tot_m and its log, the
// beta-binomial shape parameters a_v = M_v p, b_v = M_v (1-p) and -lgamma(a_v) - lgamma(b_v).  What remains per
// (n,s) is 2 logs (negative binomial) and 8 lgammas of arguments k + a (beta binomial), evaluated with the wide
// log table and the Stirling core above: ~190 FP64 operations instead of ~540.  Entries the fast path does not
// cover (hdel / LOH special parameters, p outside (0,1), zero depth, non-finite or negative data, zero allele
// reads) are flagged and go through ll_core, which also raises the reference's error conditions.
// stab layout per fit: [field f in 0..8)[kind q][state s]; sflag per fit: [q][s], bit 0 total ok, bit 1 allele ok.
// ---------------------------------------------------------------------------------------------
#define ST_NF 8
__global__ void k_statetab(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0, RmxFast Fp,
                           const __grid_constant__ RmxFitSel sel) {
    const int g = RMX_FIT_SEL(sel);
    const int QS = Fp.Q * d.S;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= QS) return;
    const int q = e / d.S, s = e - q * d.S;
    const RmxParams& P = Pg[g];
    const StateCN c = load_state2(T.tab, d, Fp.kind_nrm[2 * q], Fp.kind_nrm[2 * q + 1], s);
    const bool special = !P.normal_contamination;
    double* st = Fp.stab + (int64_t)g * ST_NF * QS;
    double depth = 0.;
#pragma unroll
    for (int m = 0; m < RMX_MAXM; ++m)
        if (m < d.M) depth += P.h[m] * (double)c.tot[m];
    int flag = 0;
    if (!(special && c.hdel) && depth > 0. && depth < 1e300) flag |= 1;
    double v[ST_NF] = {depth, flag & 1 ? log(depth) : 0., 1., 1., 1., 1., 0., 0.};
    if (!c.hdel && !(special && c.loh)) {
        double md, td;
        const double p = expected_allele_ratio(P, d, c, &md, &td);
        if (td > 0. && p > 0. && (1. - p) > 0.) {
            bool ok = true;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const double a = P.betabin_M[k] * p, b = P.betabin_M[k] * (1. - p);
                ok = ok && a > 0. && b > 0. && a < 1e300 && b < 1e300;
                v[2 + 2 * k] = a;
                v[3 + 2 * k] = b;
                v[6 + k] = -lgamma(a) - lgamma(b);
            }
            if (ok) flag |= 2;
            else { v[2] = v[3] = v[4] = v[5] = 1.; v[6] = v[7] = 0.; }
        }
    }
#pragma unroll
    for (int f = 0; f < ST_NF; ++f) st[f * QS + e] = v[f];
    Fp.sflag[(int64_t)g * QS + e] = (uint8_t)flag;
}

// slow branch of k_lltable2: the general evaluation with every reference special case
__device__ __noinline__ void ll_slow(const RmxTables& T, const RmxDims& d, const RmxParams& P, int si,
                                     const double* __restrict__ segc, int n, int s, double* ll, int* flags,
                                     const double2* __restrict__ ltab) {
    ll_terms(T, d, P, (si & 1) != 0, (si & 2) != 0, segc, d.N, n, n, s, ll, flags, ltab);
}

// Per-(fit, segment) inputs of the fast path, 12 doubles each (segf[g][n][12], three 32-byte sectors):
//   0-1 negative-binomial constants (u)     2-5 beta-binomial constants (v,w) with the Stirling cores'
//   -z + log(2 pi)/2 of their two lgammas folded in      6 log l     7 x   8 l   9 y0   10 y1
//   11 (as an integer) bit 0 total mask, bit 1 allele mask, bit 2 / 3 total / allele data valid for the fast
//   path, bits 8.. normal-copy kind
#define LLT2_SEGW 12
__global__ void k_segfast(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                          const double* __restrict__ segc_all, RmxFast Fp, const __grid_constant__ RmxFitSel sel) {
    const int g = RMX_FIT_SEL(sel);
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= d.N) return;
    const RmxParams& P = Pg[g];
    const double* segc = segc_all + (int64_t)g * 12 * d.N;
    double* o = Fp.segf + ((int64_t)g * d.N + n) * LLT2_SEGW;
    const double xv = T.x[n], lv = T.l[n], a0 = T.y0[n], a1 = T.y1[n];
    const double na = a0 + a1;
    const bool okT = xv >= 0. && xv < 1e300 && lv > 0. && lv < 1e300;
    const bool okA = a0 >= 0. && a1 >= 0. && na > 0. && na < 1e300;
    const int si = (T.tmask[(int64_t)g * d.N + n] != 0 ? 1 : 0) | (T.amask[(int64_t)g * d.N + n] != 0 ? 2 : 0) |
                   (okT ? 4 : 0) | (okA ? 8 : 0) | ((int)Fp.nkind[n] << 8);
    o[0] = segc[n];
    o[1] = segc[(int64_t)d.N + n];
#pragma unroll
    for (int c = 0; c < 4; ++c)
        o[2 + c] = segc[(int64_t)(4 + c) * d.N + n] - (na + P.betabin_M[c >> 1]) + 2. * 0.91893853320467274178;
    o[6] = lv > 0. ? log(lv) : 0.;
    o[7] = xv;
    o[8] = lv;
    o[9] = a0;
    o[10] = a1;
    o[11] = __longlong_as_double((long long)si);
}

// No barrier inside the unit loop: every thread reads its segment's 96 bytes itself (three sectors, shared by the
// 97 consecutive units of the segment, so L1 hits after the first touch) and warps proceed independently.
__global__ void __launch_bounds__(256, 3) k_lltable2(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                                   const double* __restrict__ segc_all, RmxFast Fp,
                                                   double* __restrict__ LL_all, int* __restrict__ dflags, int planes_all,
                                                   const __grid_constant__ RmxFitSel sel) {
    // planes: which term groups to (re)build -- bit u: total-read term u; bit 2 + v: the two allele terms of v.  The
    // M-step changes one dispersion at a time, and each enters one group only; the other planes of the table stand.
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double2 ltab[RMX_LOGTAB];
    const int QS = Fp.Q * d.S;
    double2* lw = reinterpret_cast<double2*>(smem_raw);                          // [RMX_LW_SIZE]
    double* st = reinterpret_cast<double*>(lw + RMX_LW_SIZE);                   // [ST_NF][QS]
    uint8_t* sfl = reinterpret_cast<uint8_t*>(st + (size_t)ST_NF * QS);         // [QS]
    const int t = threadIdx.x;
    const int g = RMX_FIT_SEL(sel);
    const int planes = RMX_PLANES_SEL(sel, planes_all);
    for (int e = t; e < RMX_LW_SIZE; e += blockDim.x) lw[e] = Fp.logw[e];
    {
        const double* sg = Fp.stab + (int64_t)g * ST_NF * QS;
        for (int e = t; e < ST_NF * QS; e += blockDim.x) st[e] = sg[e];
        const uint8_t* fg = Fp.sflag + (int64_t)g * QS;
        for (int e = t; e < QS; e += blockDim.x) sfl[e] = fg[e];
    }
    rmx_logtab_init(ltab);
    const RmxParams& P = Pg[g];
    const int NS = d.N * d.S;
    const double* segc = segc_all + (int64_t)g * 12 * d.N;
    const double2* segf = reinterpret_cast<const double2*>(Fp.segf + (int64_t)g * d.N * LLT2_SEGW);
    double* LL = LL_all + (int64_t)g * RMX_NTERM * NS;
    int flags = 0;
    const double r0 = P.negbin_r[0], r1 = P.negbin_r[1];
    const double lnr0 = log(r0), lnr1 = log(r1);
    for (int idx = blockIdx.x * blockDim.x + t; idx < NS; idx += gridDim.x * blockDim.x) {
        const int n = idx / d.S, s = idx - n * d.S;
        const double2* sp = segf + (int64_t)n * (LLT2_SEGW / 2);
        // all six sectors of the segment's record at once: one latency round instead of three dependent ones
        const double2 c01 = __ldg(sp), cb0 = __ldg(sp + 1), cb1 = __ldg(sp + 2), lx = __ldg(sp + 3), ly = __ldg(sp + 4),
                      s45 = __ldg(sp + 5);                       // (c0,c1) (cb00,cb01) (cb10,cb11) (logl,x) (l,y0) (y1,flags)
        const int si = (int)__double_as_longlong(s45.y);
        const int e = (si >> 8) * d.S + s;
        const int sf = sfl[e];
        const bool tmk = (si & 1) != 0, amk = (si & 2) != 0;
        double ll[RMX_NTERM];
        if ((tmk && (planes & 3) && !((si & 4) && (sf & 1))) || (amk && (planes & 12) && !((si & 8) && (sf & 2)))) {
            ll_slow(T, d, P, si, segc, n, s, ll, &flags, ltab);
        } else {
#pragma unroll
            for (int k = 0; k < RMX_NTERM; ++k) ll[k] = 0.;
            if (tmk && (planes & 3)) {
                const double x = lx.y;
                const double mu = st[e] * ly.x;
                const double lmu = st[QS + e] + lx.x;
                double l0, l1;
                {
                    const double z0 = r0 + mu, z1 = r1 + mu;
                    int i0, i1;
                    const bool in0 = rmx_lw_in(z0, &i0), in1 = rmx_lw_in(z1, &i1);
                    if (in0 && in1) {
                        const double2 t0 = lw[i0], t1 = lw[i1];
                        l0 = rmx_lw_log(z0, t0);
                        l1 = rmx_lw_log(z1, t1);
                    } else {
                        l0 = log(z0);
                        l1 = log(z1);
                    }
                }
                ll[0] = c01.x + fma(x, lmu - l0, r0 * (lnr0 - l0));
                ll[1] = c01.y + fma(x, lmu - l1, r1 * (lnr1 - l1));
            }
            if (amk) {
                const double y0 = ly.y, y1 = s45.x;
#pragma unroll
                for (int v = 0; v < 2; ++v) {
                    if (!((planes >> (2 + v)) & 1)) continue;
                    const double a = st[(2 + 2 * v) * QS + e], b = st[(3 + 2 * v) * QS + e];
                    const double bs = st[(6 + v) * QS + e];
                    const double z[4] = {y0 + a, y1 + b, y1 + a, y0 + b};
                    double c[4];
                    rmx_lgcore_n<4>(z, lw, c);
                    ll[2 + 2 * v] = ((v ? cb1.x : cb0.x) + bs) + (c[0] + c[1]);
                    ll[3 + 2 * v] = ((v ? cb1.y : cb0.y) + bs) + (c[2] + c[3]);
                }
            }
            // finite inputs cannot produce NaN here; parameters can (bpmodel.pyx:269, :344)
            if (isnan(ll[0] + ll[1])) flags |= RMX_F_NB_NAN;
            if (isnan((ll[2] + ll[3]) + (ll[4] + ll[5]))) flags |= RMX_F_BB_NAN;
        }
        if (planes == 15) {
#pragma unroll
            for (int k = 0; k < RMX_NTERM; ++k) LL[(int64_t)k * NS + idx] = ll[k];
        } else {
            if (planes & 1) LL[idx] = ll[0];
            if (planes & 2) LL[NS + idx] = ll[1];
            if (planes & 4) { LL[2 * NS + idx] = ll[2]; LL[3 * NS + idx] = ll[3]; }
            if (planes & 8) { LL[4 * NS + idx] = ll[4]; LL[5 * NS + idx] = ll[5]; }
        }
    }
    if (flags) atomicOr(dflags, flags);
}

// Frame log-probability (bpmodel.pyx:898-919) from the LL table, one warp per (fit, segment).
// Also writes the row max mF[n] and fexp[n,s] = exp(F[n,s] - mF[n]) for the scaled chain messages.
__global__ void __launch_bounds__(256) k_frame(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                               const double* __restrict__ LL_all,
                                               const double* __restrict__ q_ot_all, const double* __restrict__ q_oa_all,
                                               const double* __restrict__ q_sw_all,
                                               double* __restrict__ F_all, double* __restrict__ fexp_all,
                                               double* __restrict__ mF_all, int* __restrict__ dflags) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    const int g = RMX_FIT();
    const int n = warp;
    const int64_t NS = (int64_t)d.N * d.S;
    const double* LL = LL_all + (int64_t)g * RMX_NTERM * NS;
    const int64_t q = ((int64_t)g * d.N + n) * 2;
    double ot0 = q_ot_all[q], ot1 = q_ot_all[q + 1];
    double oa0 = q_oa_all[q], oa1 = q_oa_all[q + 1];
    double sw0 = q_sw_all[q], sw1 = q_sw_all[q + 1];
    double ln = T.l[n], dw = Pg[g].divergence_weight;
    double* F = F_all + (int64_t)g * NS;
    double* fexp = fexp_all + (int64_t)g * NS;
    double mx = -INFINITY;
    bool bad = false;
    for (int s = lane; s < d.S; s += 32) {
        int64_t i = (int64_t)n * d.S + s;
        const double f = rmx_frame_value(ot0, ot1, oa0, oa1, sw0, sw1, LL[i], LL[NS + i], LL[2 * NS + i], LL[3 * NS + i],
                                         LL[4 * NS + i], LL[5 * NS + i], (double)T.nsub[s], ln, dw);
        F[i] = f;
        bad |= isnan(f);
        mx = fmax(mx, f);
    }
    mx = warp_max(mx);
    for (int s = lane; s < d.S; s += 32) {
        int64_t i = (int64_t)n * d.S + s;
        fexp[i] = exp(F[i] - mx);   // own store, L1/L2 hit
    }
    if (lane == 0) mF_all[(int64_t)g * d.N + n] = mx;
    if (bad) atomicOr(dflags, RMX_F_FRAME_NAN);
}

// k_frame for state spaces up to 128: the six likelihood rows of the segment in registers (all loads in flight at
// once), one pass.  Same arithmetic as k_frame (rmx_frame_value), so the results are bit-identical.
template <int KS>
__global__ void __launch_bounds__(256) k_frame_reg(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                                   const double* __restrict__ LL_all,
                                                   const double* __restrict__ q_ot_all, const double* __restrict__ q_oa_all,
                                                   const double* __restrict__ q_sw_all,
                                                   double* __restrict__ F_all, double* __restrict__ fexp_all,
                                                   double* __restrict__ mF_all, int* __restrict__ dflags) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    const int g = RMX_FIT();
    const int n = warp;
    const int64_t NS = (int64_t)d.N * d.S;
    const int64_t row = (int64_t)n * d.S;
    const double* LL = LL_all + (int64_t)g * RMX_NTERM * NS + row;
    double ll[RMX_NTERM][KS];
#pragma unroll
    for (int k = 0; k < KS; ++k) {
        const int s = lane + 32 * k;
#pragma unroll
        for (int t = 0; t < RMX_NTERM; ++t) ll[t][k] = s < d.S ? LL[(int64_t)t * NS + s] : 0.;
    }
    const int64_t q = ((int64_t)g * d.N + n) * 2;
    const double ot0 = q_ot_all[q], ot1 = q_ot_all[q + 1];
    const double oa0 = q_oa_all[q], oa1 = q_oa_all[q + 1];
    const double sw0 = q_sw_all[q], sw1 = q_sw_all[q + 1];
    const double ln = T.l[n], dw = Pg[g].divergence_weight;
    double* F = F_all + (int64_t)g * NS + row;
    double* fexp = fexp_all + (int64_t)g * NS + row;
    double f[KS];
    double mx = -INFINITY;
    bool bad = false;
#pragma unroll
    for (int k = 0; k < KS; ++k) {
        const int s = lane + 32 * k;
        if (s < d.S) {
            f[k] = rmx_frame_value(ot0, ot1, oa0, oa1, sw0, sw1, ll[0][k], ll[1][k], ll[2][k], ll[3][k], ll[4][k], ll[5][k],
                                   (double)T.nsub[s], ln, dw);
            F[s] = f[k];
            bad |= isnan(f[k]);
            mx = fmax(mx, f[k]);
        }
    }
    mx = warp_max(mx);
#pragma unroll
    for (int k = 0; k < KS; ++k)
        if (lane + 32 * k < d.S) fexp[lane + 32 * k] = exp(f[k] - mx);
    if (lane == 0) mF_all[(int64_t)g * d.N + n] = mx;
    if (bad) atomicOr(dflags, RMX_F_FRAME_NAN);
}

// exp-normalise of a 2-vector as bpmodel.pyx:120-128 does it.
__device__ __forceinline__ void exp_normalize2(double a, double b, double* out) { rmx_exp_normalize2(a, b, out); }

// Posterior-weighted likelihood sums A[g][n][t] = sum_s post[g,n,s] * LL[g,t,n,s], t in [0,6): the
// sufficient statistic of all three 2-way factor updates and of the ELBO's likelihood terms.
// One warp per (fit, segment); one pass over post + the six LL planes (56 B per (n,s)).
__global__ void __launch_bounds__(256) k_qsums(RmxDims d, int g0, const double* __restrict__ LL_all,
                                               const double* __restrict__ post_all, double* __restrict__ A_all, int planes_all,
                                               const __grid_constant__ RmxFitSel sel) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    const int g = RMX_FIT_SEL(sel);
    const int planes = RMX_PLANES_SEL(sel, planes_all);
    const int n = warp;
    const int64_t NS = (int64_t)d.N * d.S;
    const double* LL = LL_all + (int64_t)g * RMX_NTERM * NS;
    const double* post = post_all + (int64_t)g * NS;
    double a[RMX_NTERM];
#pragma unroll
    for (int t = 0; t < RMX_NTERM; ++t) a[t] = 0.;
    // term t belongs to plane group t (t < 2) or 2 + (t - 2) / 2; only the groups in `planes` are summed and written
    const int tmask6 = (planes & 1) | (planes & 2) | ((planes & 4) ? 12 : 0) | ((planes & 8) ? 48 : 0);
    for (int s = lane; s < d.S; s += 32) {
        int64_t i = (int64_t)n * d.S + s;
        double p = post[i];
#pragma unroll
        for (int t = 0; t < RMX_NTERM; ++t)
            if ((tmask6 >> t) & 1) a[t] = fma(p, LL[(int64_t)t * NS + i], a[t]);
    }
#pragma unroll
    for (int t = 0; t < RMX_NTERM; ++t)
        if ((tmask6 >> t) & 1) a[t] = warp_sum(a[t]);
    if (lane < RMX_NTERM && ((tmask6 >> lane) & 1)) {
        double v = a[0];
#pragma unroll
        for (int t = 1; t < RMX_NTERM; ++t) v = lane == t ? a[t] : v;
        A_all[((int64_t)g * d.N + n) * RMX_NTERM + lane] = v;
    }
}

// The tail of one coordinate-ascent round and the head of the next in ONE pass over the likelihood table, inside
// rmx_variational_update's loop (cn_model.py:444-460 order): after the chain messages of update_p_cn,
//   post = normalise(ahat .* bhat)                       (k_post; bpmodel.pyx:946-952)
//   A[t] = sum_s post * LL[t]                            (k_qsums)
//   update_p_outlier_total, update_p_outlier_allele      (k_qupdate 1, 2; :987-1023)
// and, when another round follows,
//   update_p_allele_swap of that round                   (k_qupdate 0; :1025-1042)
//   framelogprob of that round, its row max and exp      (k_frame; :898-919)
// One warp per (fit, segment), the segment's post and six LL rows in registers: 88 bytes per (n,s) instead of the
// 144 of the four separate passes.  Expressions and summation orders are the separate kernels' (results agree with
// the unfused sequence to the last bits; FMA contraction may differ).  KS = ceil(S / 32).
template <int KS>
__global__ void __launch_bounds__(256, KS <= 4 ? 3 : 2) k_round(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                               const double* __restrict__ LL_all, const double* __restrict__ ahat_all,
                                               const double* __restrict__ bhat_all, double* __restrict__ post_all,
                                               double* __restrict__ A_all, double* __restrict__ q_ot_all,
                                               double* __restrict__ q_oa_all, double* __restrict__ q_sw_all,
                                               double lpt0, double lpt1, double lpa0, double lpa1, int next, int head,
                                               double* __restrict__ F_all,
                                               double* __restrict__ fexp_all, double* __restrict__ mF_all,
                                               int* __restrict__ dflags) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    const int g = RMX_FIT();
    const int n = warp;
    const int64_t NS = (int64_t)d.N * d.S;
    const int64_t row = (int64_t)n * d.S;
    const double* LL = LL_all + (int64_t)g * RMX_NTERM * NS + row;
    const double* ah = ahat_all + (int64_t)g * NS + row;
    const double* bh = bhat_all + (int64_t)g * NS + row;
    double ll[RMX_NTERM][KS], p[KS];
#pragma unroll
    for (int k = 0; k < KS; ++k) {
        const int s = lane + 32 * k;
#pragma unroll
        for (int t = 0; t < RMX_NTERM; ++t) ll[t][k] = s < d.S ? LL[(int64_t)t * NS + s] : 0.;
        p[k] = s < d.S ? (head ? post_all[(int64_t)g * NS + row + s] : __dmul_rn(ah[s], bh[s])) : 0.;
    }
    // head = 1: the first round of a loop.  The posterior marginals are the stored ones (previous EM iteration or
    // the constructor's), only the swap update and the frame log-probability follow (update_p_allele_swap and the
    // head of update_p_cn, cn_model.py:447-449): k_qsums + k_qupdate(0) + k_frame in one pass.
    // ---- posterior marginals
    double z = 0.;
    bool bad = false;
    if (!head) {
#pragma unroll
    for (int k = 0; k < KS; ++k)
        if (lane + 32 * k < d.S) z = __dadd_rn(z, p[k]);
    z = warp_sum(z);
    bad = !(z > 0.);
    double* post = post_all + (int64_t)g * NS + row;
#pragma unroll
    for (int k = 0; k < KS; ++k)
        if (lane + 32 * k < d.S) {
            p[k] = __ddiv_rn(p[k], z);
            post[lane + 32 * k] = p[k];
            bad |= isnan(p[k]);
        }
    if (bad) atomicOr(dflags, RMX_F_POST_NAN);
    }
    // ---- posterior-weighted likelihood sums
    double a[RMX_NTERM];
#pragma unroll
    for (int t = 0; t < RMX_NTERM; ++t) a[t] = 0.;
#pragma unroll
    for (int k = 0; k < KS; ++k)
        if (lane + 32 * k < d.S) {
#pragma unroll
            for (int t = 0; t < RMX_NTERM; ++t) a[t] = fma(p[k], ll[t][k], a[t]);
        }
#pragma unroll
    for (int t = 0; t < RMX_NTERM; ++t) a[t] = warp_sum(a[t]);
    const int64_t gn = (int64_t)g * d.N + n;
    if (lane < RMX_NTERM) {
        double v = a[0];
#pragma unroll
        for (int t = 1; t < RMX_NTERM; ++t) v = lane == t ? a[t] : v;
        A_all[gn * RMX_NTERM + lane] = v;
    }
    // ---- outlier factors (every lane holds the same sums; lane 0 stores)
    double ot[2], oa[2], sw[2], qa[2];
    if (head) {
        ot[0] = q_ot_all[gn * 2]; ot[1] = q_ot_all[gn * 2 + 1];
        oa[0] = q_oa_all[gn * 2]; oa[1] = q_oa_all[gn * 2 + 1];
    } else {
        // the two outlier factors do not depend on each other: one exp and one division for both (common.cuh)
        double qb[2];
        rmx_q_args(1, a, 0., 0., lpt0, lpt1, qa);
        sw[0] = q_sw_all[gn * 2];
        sw[1] = q_sw_all[gn * 2 + 1];
        rmx_q_args(2, a, sw[0], sw[1], lpa0, lpa1, qb);
        rmx_exp_normalize2_x2(qa[0], qa[1], qb[0], qb[1], lane, ot, oa);
        if (lane == 0) {
            q_ot_all[gn * 2] = ot[0]; q_ot_all[gn * 2 + 1] = ot[1];
            q_oa_all[gn * 2] = oa[0]; q_oa_all[gn * 2 + 1] = oa[1];
        }
    }
    if (!next) return;
    // ---- next round: allele swap factor, frame log-probability
    rmx_q_args(0, a, oa[0], oa[1], 0., 0., qa);
    exp_normalize2(qa[0], qa[1], sw);
    if (lane == 0) { q_sw_all[gn * 2] = sw[0]; q_sw_all[gn * 2 + 1] = sw[1]; }
    const double ln = T.l[n], dw = Pg[g].divergence_weight;
    double* F = F_all + (int64_t)g * NS + row;
    double* fexp = fexp_all + (int64_t)g * NS + row;
    double mx = -INFINITY;
    bad = false;
#pragma unroll
    for (int k = 0; k < KS; ++k) {
        const int s = lane + 32 * k;
        if (s < d.S) {
            const double f = rmx_frame_value(ot[0], ot[1], oa[0], oa[1], sw[0], sw[1], ll[0][k], ll[1][k], ll[2][k],
                                             ll[3][k], ll[4][k], ll[5][k], (double)T.nsub[s], ln, dw);
            F[s] = f;
            p[k] = f;
            bad |= isnan(f);
            mx = fmax(mx, f);
        }
    }
    mx = warp_max(mx);
#pragma unroll
    for (int k = 0; k < KS; ++k)
        if (lane + 32 * k < d.S) fexp[lane + 32 * k] = exp(p[k] - mx);
    if (lane == 0) mF_all[gn] = mx;
    if (bad) atomicOr(dflags, RMX_F_FRAME_NAN);
}

// The three 2-way factor updates from the sums A, one thread per (fit, segment).
//   which = 0: update_p_allele_swap   (bpmodel.pyx:1025-1042)
//   which = 1: update_p_outlier_total (bpmodel.pyx:987-1003)
//   which = 2: update_p_outlier_allele(bpmodel.pyx:1005-1023)
__global__ void __launch_bounds__(256) k_qupdate(RmxDims d, int g0, int which, const double* __restrict__ A_all,
                                                 const double* __restrict__ q_oa_all, const double* __restrict__ q_sw_all,
                                                 double lp0, double lp1, double* __restrict__ out_all) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= d.N) return;
    const int g = RMX_FIT();
    const int64_t gn = (int64_t)g * d.N + n;
    const double* A = A_all + gn * RMX_NTERM;
    double a[2];
    if (which == 0) rmx_q_args(0, A, q_oa_all[gn * 2], q_oa_all[gn * 2 + 1], lp0, lp1, a);
    else if (which == 1) rmx_q_args(1, A, 0., 0., lp0, lp1, a);
    else rmx_q_args(2, A, q_sw_all[gn * 2], q_sw_all[gn * 2 + 1], lp0, lp1, a);
    const double a0 = a[0], a1 = a[1];
    exp_normalize2(a0, a1, out_all + gn * 2);
}

// ---------------------------------------------------------------------------------------------
// Reductions.  Every scalar the host reads is produced by the same two-stage scheme: per-row
// values (fixed lane order) then one CTA per (column, fit) summing the row values in a fixed strided
// order, so results are reproducible run to run.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_reduce_cols(const double* __restrict__ in_all, int64_t n, int ncol, int g0,
                                                      int64_t in_fit_stride, double* __restrict__ out_all,
                                                      int out_fit_stride) {
    // out[g][c] = sum_i in[g][c*n + i]
    __shared__ double sm[32];
    const int c = blockIdx.x;
    const int g = RMX_FIT();
    const double* p = in_all + (int64_t)g * in_fit_stride + (int64_t)c * n;
    double acc = 0.;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) acc += p[i];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.;
        v = warp_sum(v);
        if (threadIdx.x == 0) out_all[(int64_t)g * out_fit_stride + c] = v;
    }
}

__device__ __forceinline__ double plogp(double p) { return p > 0. ? p * log(p) : 0.; }

// Per-segment ELBO terms (bpmodel.pyx:1044-1117), one warp per (fit, segment), from post, F and the sums A.
// rowv columns (each length N, fit stride 8*N):
//   0: sum_s post*prior_cn          1: sum_s post * sum_u q_ot*LLtot   2: sum_s post * sum_vw q_oa q_sw LLal
//   3: q_ot . log prior            4: q_oa . log prior                 5: sum_s post*F
//   6: entropy terms of q_ot,q_oa,q_sw  (sum p log p)
__global__ void __launch_bounds__(256) k_elbo_rows(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                                   const double* __restrict__ A_all, const double* __restrict__ post_all,
                                                   const double* __restrict__ F_all, const double* __restrict__ q_ot_all,
                                                   const double* __restrict__ q_oa_all, const double* __restrict__ q_sw_all,
                                                   double prior_ot, double prior_oa, double* __restrict__ rowv_all) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    const int g = RMX_FIT();
    const int n = warp;
    const int64_t NS = (int64_t)d.N * d.S;
    const int64_t gn = (int64_t)g * d.N + n;
    const double* post = post_all + (int64_t)g * NS;
    const double* F = F_all + (int64_t)g * NS;
    double ln = T.l[n], dw = Pg[g].divergence_weight;
    double e0 = 0., e5 = 0.;
    for (int s = lane; s < d.S; s += 32) {
        int64_t i = (int64_t)n * d.S + s;
        double p = post[i];
        e0 += p * (-1.0 * (double)T.nsub[s] * ln * dw);
        e5 += p * F[i];
    }
    e0 = warp_sum(e0);
    e5 = warp_sum(e5);
    // The segment's scalar tail needs ten logarithms (six p log p, four log prior).  On one lane they were ten passes of
    // the log routine per warp and made this kernel FP64-issue bound (1.63 ms per 48 fits for 2.2 GB of reads); here
    // lanes 0..9 take one each in ONE pass and lane 0 combines them in the original order (same values).
    double arg = 1.;
    if (lane < 6) arg = (lane < 2 ? q_ot_all : lane < 4 ? q_oa_all : q_sw_all)[gn * 2 + (lane & 1)];
    else if (lane < 10) arg = lane == 6 ? 1. - prior_ot : lane == 7 ? prior_ot : lane == 8 ? 1. - prior_oa : prior_oa;
    const double q = arg;
    double lg = 0.;
    if (lane < 10 && (lane >= 6 || arg > 0.)) lg = log(arg);
    const double pl = (lane < 6 && q > 0.) ? q * lg : 0.;        // plogp of lanes 0..5
    double qv[6], plv[6], lpr[4];
#pragma unroll
    for (int k = 0; k < 6; ++k) { qv[k] = __shfl_sync(0xffffffffu, q, k); plv[k] = __shfl_sync(0xffffffffu, pl, k); }
#pragma unroll
    for (int k = 0; k < 4; ++k) lpr[k] = __shfl_sync(0xffffffffu, lg, 6 + k);
    if (lane == 0) {
        const double* A = A_all + gn * RMX_NTERM;
        const double ot0 = qv[0], ot1 = qv[1], oa0 = qv[2], oa1 = qv[3], sw0 = qv[4], sw1 = qv[5];
        double* rowv = rowv_all + (int64_t)g * 8 * d.N;
        int64_t N = d.N;
        rowv[0 * N + n] = e0;
        rowv[1 * N + n] = ot0 * A[0] + ot1 * A[1];
        rowv[2 * N + n] = oa0 * sw0 * A[2] + oa0 * sw1 * A[3] + oa1 * sw0 * A[4] + oa1 * sw1 * A[5];
        rowv[3 * N + n] = ot0 * lpr[0] + ot1 * lpr[1];
        rowv[4 * N + n] = oa0 * lpr[2] + oa1 * lpr[3];
        rowv[5 * N + n] = e5;
        rowv[6 * N + n] = plv[0] + plv[1] + plv[2] + plv[3] + plv[4] + plv[5];
    }
}

// Full-data E_q[log lik] (bpmodel.pyx:1125-1157 with an all-ones sample) from the posterior-weighted likelihood sums
// A[n][t] = sum_s post[n,s] LL_t[n,s]: the per-segment value sum_u q_ot A_u + sum_vw q_oa q_sw A_vw, one thread per
// (fit, segment); k_reduce_cols adds the rows.
__global__ void __launch_bounds__(256) k_ell_rows_from_sums(RmxDims d, int g0, const double* __restrict__ A_all,
                                                            const double* __restrict__ q_ot_all,
                                                            const double* __restrict__ q_oa_all,
                                                            const double* __restrict__ q_sw_all,
                                                            double* __restrict__ out_all, int64_t out_fit_stride) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= d.N) return;
    const int g = RMX_FIT();
    const int64_t gn = (int64_t)g * d.N + n;
    const double* A = A_all + gn * RMX_NTERM;
    const double ot0 = q_ot_all[gn * 2], ot1 = q_ot_all[gn * 2 + 1];
    const double oa0 = q_oa_all[gn * 2], oa1 = q_oa_all[gn * 2 + 1];
    const double sw0 = q_sw_all[gn * 2], sw1 = q_sw_all[gn * 2 + 1];
    out_all[(int64_t)g * out_fit_stride + n] = (ot0 * A[0] + ot1 * A[1]) +
        (oa0 * sw0 * A[2] + oa0 * sw1 * A[3] + oa1 * sw0 * A[4] + oa1 * sw1 * A[5]);
}

// p log p over a flat per-fit array (p_breakpoint)
__global__ void k_plogp(const double* __restrict__ p_all, int64_t n, int g0, double* __restrict__ out_all,
                        int64_t out_fit_stride) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int g = RMX_FIT();
    if (i < n) out_all[(int64_t)g * out_fit_stride + i] = plogp(p_all[(int64_t)g * n + i]);
}

// ---------------------------------------------------------------------------------------------
// M-step objective on per-fit row lists: E_q[log lik] (bpmodel.pyx:1125-1157) and its gradient in h
// (bpmodel.pyx:1159-1195), with the LL terms evaluated for the CURRENT parameters (the LL table
// belongs to the E-step generation and is not touched).  One thread per (row, state); the CTA
// sum goes to part[g][blockIdx.x] (ELL) or part[g][m*gridDim.x + blockIdx.x] (gradient).
// rows_all == nullptr: every segment of every fit; else fit g uses rows_all[g*N .. g*N + Rn[g]).
// ---------------------------------------------------------------------------------------------
template <int NV>
__device__ __forceinline__ void block_sum_store(double* v, double* __restrict__ part, int stride) {
    __shared__ double sm[NV][8];
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        double x = warp_sum(v[k]);
        if (lane == 0) sm[k][wid] = x;
    }
    __syncthreads();
    if (threadIdx.x < NV) {
        double a = 0.;
        int nw = blockDim.x >> 5;
        for (int w = 0; w < nw; ++w) a += sm[threadIdx.x][w];
        part[(int64_t)threadIdx.x * stride + blockIdx.x] = a;
    }
}

// the twelve hoisted per-segment constants of one row (see the top of this file), written with stride `st`
__device__ __forceinline__ void segconst_row(const RmxParams& P, double x, double y0, double y1, double* __restrict__ o, int st) {
    double na = y0 + y1;
    double lgx1 = lgamma(x + 1.);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        double rr = P.negbin_r[u];
        o[(0 + u) * st] = lgamma(x + rr) - lgx1 - lgamma(rr);
        double rh = P.negbin_hdel_r[u];
        o[(2 + u) * st] = P.normal_contamination ? 0. : (lgamma(x + rh) - lgx1 - lgamma(rh));
    }
    double lgn1 = lgamma(na + 1.);
    double comb[2];
    comb[0] = lgn1 - lgamma(y0 + 1.) - lgamma(na - y0 + 1.);
    comb[1] = lgn1 - lgamma(y1 + 1.) - lgamma(na - y1 + 1.);
#pragma unroll
    for (int v = 0; v < 2; ++v) {
        double Mv = P.betabin_M[v];
        double t = -lgamma(na + Mv) + lgamma(Mv);
        double Ml = P.betabin_loh_M[v];
        double tl = P.normal_contamination ? 0. : (-lgamma(na + Ml) + lgamma(Ml));
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            o[(4 + v * 2 + w) * st] = comb[w] + t;
            o[(8 + v * 2 + w) * st] = comb[w] + tl;
        }
    }
}

#define RMX_ELL_CHUNKS 16
#define RMX_ELL_PF 7             // posterior entries a thread of k_ell_sample fetches ahead (16 rows x 97 states / 256)
#define RMX_ELL_MAXROWS 16       // rows per chunk held in shared memory (256 sampled rows per fit)
// The parameter-independent part of those constants for the staged sample rows of the batched M-step, computed once
// per sample: rowc[(g * 256 + r) * 3 + {0,1,2}] = lgamma(x+1), and the two log binomial coefficients.
__global__ void k_ell_rowcache(RmxTables T, RmxDims d, int g0, const int* __restrict__ rows_all, const int* __restrict__ Rn,
                               double* __restrict__ rowc) {
    const int g = RMX_FIT();
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= Rn[g]) return;
    const int n = rows_all[(int64_t)g * d.N + r];
    const double x = T.x[n], y0 = T.y0[n], y1 = T.y1[n], na = y0 + y1;
    const double lgn1 = lgamma(na + 1.);
    double* o = rowc + ((int64_t)g * (RMX_ELL_CHUNKS * RMX_ELL_MAXROWS) + r) * 3;
    o[0] = lgamma(x + 1.);
    o[1] = lgn1 - lgamma(y0 + 1.) - lgamma(na - y0 + 1.);
    o[2] = lgn1 - lgamma(y1 + 1.) - lgamma(na - y1 + 1.);
}

// segconst_row with the cached part given and only the constants of the terms in `planes` (the same operations in the
// same order on those, hence the same values)
__device__ __forceinline__ void segconst_row_cached(const RmxParams& P, double x, double na, const double* __restrict__ rc,
                                                    int planes, double* __restrict__ o, int st) {
    const double lgx1 = rc[0];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        if (!((planes >> u) & 1)) continue;
        double rr = P.negbin_r[u];
        o[(0 + u) * st] = lgamma(x + rr) - lgx1 - lgamma(rr);
        double rh = P.negbin_hdel_r[u];
        o[(2 + u) * st] = P.normal_contamination ? 0. : (lgamma(x + rh) - lgx1 - lgamma(rh));
    }
#pragma unroll
    for (int v = 0; v < 2; ++v) {
        if (!((planes >> (2 + v)) & 1)) continue;
        double Mv = P.betabin_M[v];
        double t = -lgamma(na + Mv) + lgamma(Mv);
        double Ml = P.betabin_loh_M[v];
        double tl = P.normal_contamination ? 0. : (-lgamma(na + Ml) + lgamma(Ml));
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            o[(4 + v * 2 + w) * st] = rc[1 + w] + t;
            o[(8 + v * 2 + w) * st] = rc[1 + w] + tl;
        }
    }
}

// segconst for the row lists of the M-step (same contents as k_segconst, per-fit row count).  With `segk` (the
// parameter-independent part per segment, k_segk) and no row list, only the constants of the term groups in `planes`
// are recomputed -- 2 to 8 lgamma per segment instead of 20.
__global__ void k_segconst_rows(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                const int* __restrict__ rows_all, const int* __restrict__ Rn,
                                double* __restrict__ segc_all, const double* __restrict__ segk, int planes_all,
                                const __grid_constant__ RmxFitSel sel) {
    const int g = RMX_FIT_SEL(sel);
    const int planes = RMX_PLANES_SEL(sel, planes_all);
    const int R = rows_all ? Rn[g] : d.N;
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    int n = rows_all ? rows_all[(int64_t)g * d.N + r] : r;
    if (segk && !rows_all)
        segconst_row_cached(Pg[g], T.x[n], T.y0[n] + T.y1[n], segk + (int64_t)n * 3, planes,
                            segc_all + (int64_t)g * 12 * d.N + r, R);
    else
        segconst_row(Pg[g], T.x[n], T.y0[n], T.y1[n], segc_all + (int64_t)g * 12 * d.N + r, R);
}

// segk[n][0..2] = lgamma(x+1) and the two log binomial coefficients of segment n: data only, once per batch
__global__ void k_segk(RmxTables T, RmxDims d, double* __restrict__ segk) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= d.N) return;
    const double x = T.x[n], y0 = T.y0[n], y1 = T.y1[n], na = y0 + y1;
    const double lgn1 = lgamma(na + 1.);
    segk[(int64_t)n * 3] = lgamma(x + 1.);
    segk[(int64_t)n * 3 + 1] = lgn1 - lgamma(y0 + 1.) - lgamma(na - y0 + 1.);
    segk[(int64_t)n * 3 + 2] = lgn1 - lgamma(y1 + 1.) - lgamma(na - y1 + 1.);
}

// The sampled M-step objective in ONE launch (the lock-step M-step evaluates it ~300 times per EM iteration, each a
// host round trip): CTA (chunk c, fit g) takes a contiguous sixteenth of the fit's sample rows, computes their hoisted
// constants into shared memory, sums post * q * LL over (row, state) in a fixed order and writes ONE partial;
// the host adds the RMX_ELL_CHUNKS partials of a fit in order.  `planes` as in ll_core.
__global__ void __launch_bounds__(256, 3) k_ell_sample(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                                    const int* __restrict__ rows_all, const int* __restrict__ Rn,
                                                    const double* __restrict__ post_all,
                                                    const double* __restrict__ q_ot_all, const double* __restrict__ q_oa_all,
                                                    const double* __restrict__ q_sw_all, const double* __restrict__ rowc,
                                                    double* __restrict__ out, int out_fit_stride,
                                                    int* __restrict__ dflags, int planes) {
    __shared__ double2 ltab[RMX_LOGTAB];
    __shared__ double segc[12 * RMX_ELL_MAXROWS];
    // per-row inputs staged once (x, l, y0, y1, the six q-factor entries), masks / normal copies, and the state table:
    // the (row, state) loop then touches global memory for the posterior only
    __shared__ double rowd[10 * RMX_ELL_MAXROWS];
    __shared__ int rown[RMX_ELL_MAXROWS], rowi[RMX_ELL_MAXROWS];
    __shared__ int8_t stab[2 * RMX_MAXM * 128];
    const int g = RMX_FIT();
    const int R = Rn[g];
    const int per = (R + RMX_ELL_CHUNKS - 1) / RMX_ELL_CHUNKS;
    const int r0 = blockIdx.x * per;
    const int nr = max(0, min(per, R - r0));
    const RmxParams& P = Pg[g];
    const bool tab_sm = d.S <= 128;
    // this thread's posterior entries first: two dependent global loads each (row list, then the posterior), in flight
    // under the set-up below instead of inside the unit loop (up to RMX_ELL_PF units; further ones load in the loop).
    // The second load is a cp.async into the thread's own shared-memory slots: nothing waits for it before the unit
    // loop (a register array was spilled to local memory by the compiler, and each spill store waited for its load:
    // 17 % of the kernel's stall samples, ncu source view).
    const int units = nr * d.S;
    __shared__ double pps[RMX_ELL_PF * 256];
    {
        int rn[RMX_ELL_PF];
#pragma unroll
        for (int k = 0; k < RMX_ELL_PF; ++k) {
            const int idx = threadIdx.x + k * 256;
            rn[k] = idx < units ? rows_all[(int64_t)g * d.N + r0 + idx / d.S] : -1;
        }
#pragma unroll
        for (int k = 0; k < RMX_ELL_PF; ++k) {
            const int idx = threadIdx.x + k * 256;
            if (rn[k] >= 0) {
                const int r = idx / d.S, s = idx - r * d.S;
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(pps + k * 256 + threadIdx.x)),
                             "l"(post_all + ((int64_t)g * d.N + rn[k]) * d.S + s)
                             : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (tab_sm)
        for (int e = threadIdx.x; e < d.S * d.M * 2; e += blockDim.x) stab[e] = T.tab[e];
    if ((int)threadIdx.x < nr) {
        const int r = threadIdx.x;
        const int n = rows_all[(int64_t)g * d.N + r0 + r];
        const int64_t gn = (int64_t)g * d.N + n;
        rown[r] = n;
        const double x = T.x[n], y0 = T.y0[n], y1 = T.y1[n];
        rowd[0 * RMX_ELL_MAXROWS + r] = x;
        rowd[1 * RMX_ELL_MAXROWS + r] = T.l[n];
        rowd[2 * RMX_ELL_MAXROWS + r] = y0;
        rowd[3 * RMX_ELL_MAXROWS + r] = y1;
        rowd[4 * RMX_ELL_MAXROWS + r] = q_ot_all[2 * gn];
        rowd[5 * RMX_ELL_MAXROWS + r] = q_ot_all[2 * gn + 1];
        rowd[6 * RMX_ELL_MAXROWS + r] = q_oa_all[2 * gn];
        rowd[7 * RMX_ELL_MAXROWS + r] = q_oa_all[2 * gn + 1];
        rowd[8 * RMX_ELL_MAXROWS + r] = q_sw_all[2 * gn];
        rowd[9 * RMX_ELL_MAXROWS + r] = q_sw_all[2 * gn + 1];
        rowi[r] = (T.tmask[gn] != 0 ? 1 : 0) | (T.amask[gn] != 0 ? 2 : 0) | ((int)T.nrm[2 * n] << 8) | ((int)T.nrm[2 * n + 1] << 16);
    }
    // the hoisted constants of the rows (segconst_row_cached: up to sixteen lgamma per row, four per term group), one
    // lgamma per thread instead of all of a row's on one thread while the others wait at the barrier: thread (row r,
    // group j, q) with the four q of a group on adjacent lanes, combined through shuffles by the lane q = 0 with the
    // operations and the order of segconst_row_cached (same values)
    {
        const int r = threadIdx.x >> 4, j = (threadIdx.x >> 2) & 3, q = threadIdx.x & 3;
        const bool on = r < nr && ((planes >> j) & 1);
        double lg = 0.;
        const double* rc = rowc + ((int64_t)g * (RMX_ELL_CHUNKS * RMX_ELL_MAXROWS) + r0 + (r < nr ? r : 0)) * 3;
        if (on && (q < 2 || !P.normal_contamination)) {
            const int n = rows_all[(int64_t)g * d.N + r0 + r];
            const double base = j < 2 ? T.x[n] : T.y0[n] + T.y1[n];
            const double par = j < 2 ? (q < 2 ? P.negbin_r[j] : P.negbin_hdel_r[j]) : (q < 2 ? P.betabin_M[j - 2] : P.betabin_loh_M[j - 2]);
            lg = lgamma((q & 1) ? par : base + par);
        }
        const double lg1 = __shfl_down_sync(0xffffffffu, lg, 1), lg2 = __shfl_down_sync(0xffffffffu, lg, 2),
                     lg3 = __shfl_down_sync(0xffffffffu, lg, 3);
        if (on && q == 0) {
            double* o = segc + r;
            if (j < 2) {
                const double lgx1 = rc[0];
                o[(0 + j) * RMX_ELL_MAXROWS] = lg - lgx1 - lg1;
                o[(2 + j) * RMX_ELL_MAXROWS] = P.normal_contamination ? 0. : (lg2 - lgx1 - lg3);
            } else {
                const int v = j - 2;
                const double t = -lg + lg1;
                const double tl = P.normal_contamination ? 0. : (-lg2 + lg3);
#pragma unroll
                for (int w = 0; w < 2; ++w) {
                    o[(4 + v * 2 + w) * RMX_ELL_MAXROWS] = rc[1 + w] + t;
                    o[(8 + v * 2 + w) * RMX_ELL_MAXROWS] = rc[1 + w] + tl;
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");       // own slots only: no barrier needed for them
    rmx_logtab_init(ltab);      // includes the barrier
    double acc[1] = {0.};
    int flags = 0;
    const int8_t* tab = tab_sm ? stab : T.tab;
    int kpf = 0;
    for (int idx = threadIdx.x; idx < units; idx += blockDim.x, ++kpf) {
        const int r = idx / d.S, s = idx - r * d.S;
        const int n = rown[r], ri = rowi[r];
        const double p = kpf < RMX_ELL_PF ? pps[kpf * 256 + threadIdx.x] : post_all[((int64_t)g * d.N + n) * d.S + s];
        SegIn in;
        in.segc = segc + r;
        in.sstride = RMX_ELL_MAXROWS;
        in.x = rowd[r]; in.l = rowd[RMX_ELL_MAXROWS + r]; in.y0 = rowd[2 * RMX_ELL_MAXROWS + r]; in.y1 = rowd[3 * RMX_ELL_MAXROWS + r];
        in.nrm0 = (int)(int8_t)((ri >> 8) & 0xff); in.nrm1 = (int)(int8_t)((ri >> 16) & 0xff);
        double ll[RMX_NTERM];
        ll_core(tab, d, P, (ri & 1) != 0, (ri & 2) != 0, in, s, ll, &flags, ltab, nullptr, planes);
        const double ot0 = rowd[4 * RMX_ELL_MAXROWS + r], ot1 = rowd[5 * RMX_ELL_MAXROWS + r];
        const double oa0 = rowd[6 * RMX_ELL_MAXROWS + r], oa1 = rowd[7 * RMX_ELL_MAXROWS + r];
        const double sw0 = rowd[8 * RMX_ELL_MAXROWS + r], sw1 = rowd[9 * RMX_ELL_MAXROWS + r];
        double a = 0.;
        if (planes & 1) a += p * ot0 * ll[0];
        if (planes & 2) a += p * ot1 * ll[1];
        if (planes & 4) a += p * oa0 * sw0 * ll[2] + p * oa0 * sw1 * ll[3];
        if (planes & 8) a += p * oa1 * sw0 * ll[4] + p * oa1 * sw1 * ll[5];
        acc[0] += a;
    }
    // block_sum_store writes part[threadIdx.x * stride + blockIdx.x]: one value per (fit, chunk)
    block_sum_store<1>(acc, out + (int64_t)g * out_fit_stride, gridDim.x);
    if (flags) atomicOr(dflags, flags);
}

__global__ void __launch_bounds__(256) k_ell(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                             const int* __restrict__ rows_all, const int* __restrict__ Rn,
                                             const double* __restrict__ segc_all, const double* __restrict__ post_all,
                                             const double* __restrict__ q_ot_all, const double* __restrict__ q_oa_all,
                                             const double* __restrict__ q_sw_all, double* __restrict__ part_all,
                                             int64_t part_fit_stride, int* __restrict__ dflags, int planes) {
    __shared__ double2 ltab[RMX_LOGTAB];
    rmx_logtab_init(ltab);
    const int g = RMX_FIT();
    const int R = rows_all ? Rn[g] : d.N;
    int64_t RS = (int64_t)R * d.S;
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double acc[1] = {0.};
    int flags = 0;
    if (idx < RS) {
        int r = (int)(idx / d.S), s = (int)(idx - (int64_t)r * d.S);
        int n = rows_all ? rows_all[(int64_t)g * d.N + r] : r;
        const int64_t gn = (int64_t)g * d.N + n;
        double ll[RMX_NTERM];
        ll_terms(T, d, Pg[g], T.tmask[gn] != 0, T.amask[gn] != 0, segc_all + (int64_t)g * 12 * d.N, R, r, n, s, ll,
                 &flags, ltab, planes);
        double p = post_all[gn * d.S + s];
        double ot0 = q_ot_all[2 * gn], ot1 = q_ot_all[2 * gn + 1];
        double oa0 = q_oa_all[2 * gn], oa1 = q_oa_all[2 * gn + 1];
        double sw0 = q_sw_all[2 * gn], sw1 = q_sw_all[2 * gn + 1];
        if (planes == 15) {
            acc[0] = p * ot0 * ll[0] + p * ot1 * ll[1] + p * oa0 * sw0 * ll[2] + p * oa0 * sw1 * ll[3] +
                     p * oa1 * sw0 * ll[4] + p * oa1 * sw1 * ll[5];
        } else {        // the same terms in the same order, the ones left out skipped
            double a = 0.;
            if (planes & 1) a += p * ot0 * ll[0];
            if (planes & 2) a += p * ot1 * ll[1];
            if (planes & 4) a += p * oa0 * sw0 * ll[2] + p * oa0 * sw1 * ll[3];
            if (planes & 8) a += p * oa1 * sw0 * ll[4] + p * oa1 * sw1 * ll[5];
            acc[0] = a;
        }
    }
    block_sum_store<1>(acc, part_all + (int64_t)g * part_fit_stride, gridDim.x);
    if (flags) atomicOr(dflags, flags);
}

__global__ void __launch_bounds__(256) k_ell_grad(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g0,
                                                  const int* __restrict__ rows_all, const int* __restrict__ Rn,
                                                  const double* __restrict__ post_all,
                                                  const double* __restrict__ q_ot_all, const double* __restrict__ q_oa_all,
                                                  const double* __restrict__ q_sw_all, double* __restrict__ part_all,
                                                  int64_t part_fit_stride, int* __restrict__ dflags) {
    const int gf = RMX_FIT();
    const RmxParams& P = Pg[gf];
    const int R = rows_all ? Rn[gf] : d.N;
    const int* rows = rows_all ? rows_all + (int64_t)gf * d.N : nullptr;
    const double* post = post_all + (int64_t)gf * d.N * d.S;
    const double* q_ot = q_ot_all + (int64_t)gf * d.N * 2;
    const double* q_oa = q_oa_all + (int64_t)gf * d.N * 2;
    const double* q_sw = q_sw_all + (int64_t)gf * d.N * 2;
    const uint8_t* tmask = T.tmask + (int64_t)gf * d.N;
    const uint8_t* amask = T.amask + (int64_t)gf * d.N;
    double* part = part_all + (int64_t)gf * part_fit_stride;
    int64_t RS = (int64_t)R * d.S;
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double g[RMX_MAXM];
#pragma unroll
    for (int m = 0; m < RMX_MAXM; ++m) g[m] = 0.;
    int flags = 0;
    if (idx < RS) {
        int r = (int)(idx / d.S), s = (int)(idx - (int64_t)r * d.S);
        int n = rows ? rows[r] : r;
        StateCN c = load_state(T, d, n, s);
        const bool special = !P.normal_contamination;
        double p = post[(int64_t)n * d.S + s];
        // total reads (bpmodel.pyx:778-807, :275-301)
        if (tmask[n] && !(special && c.hdel)) {
            double l = T.l[n], x = T.x[n];
            double mu = expected_total_reads(P, d, c, l);
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                double rr = P.negbin_r[u];
                double dmu = x / mu - (rr + x) / (rr + mu);
                if (isnan(dmu)) flags |= RMX_F_NB_PARTIAL_NAN;
                double wgt = p * q_ot[2 * n + u];
#pragma unroll
                for (int m = 0; m < RMX_MAXM; ++m)
                    if (m < d.M) g[m] += wgt * ((l * (double)c.tot[m]) * dmu);
            }
        }
        // allele reads (bpmodel.pyx:855-896, :350-394, :726-744)
        if (amask[n] && !(special && c.loh)) {
            double md, td;
            double pr = expected_allele_ratio(P, d, c, &md, &td);
            if (td <= 0.) {
                flags |= RMX_F_TOTAL_DEPTH;
            } else {
                double y0 = T.y0[n], y1 = T.y1[n];
                double na = y0 + y1;
                if (na != 0.) {
                    if (pr <= 0. || (1. - pr) <= 0.) {
                        flags |= RMX_F_BB_P_RANGE;
                    } else {
                        double rp[RMX_MAXM];
#pragma unroll
                        for (int m = 0; m < RMX_MAXM; ++m)
                            rp[m] = (m < d.M) ? ((double)c.a0[m] * td - md * (double)c.tot[m]) / (td * td) : 0.;
#pragma unroll
                        for (int v = 0; v < 2; ++v) {
                            double Mv = P.betabin_M[v];
                            double a = Mv * pr, b = Mv * (1. - pr);
                            double dga = digamma_as103(a, &flags), dgb = digamma_as103(b, &flags);
#pragma unroll
                            for (int w = 0; w < 2; ++w) {
                                double k = w ? y1 : y0;
                                double dp = Mv * digamma_as103(k + a, &flags) + (-Mv) * digamma_as103(na - k + b, &flags) -
                                            Mv * dga - (-Mv) * dgb;
                                if (isnan(dp)) flags |= RMX_F_BB_PARTIAL_NAN;
                                double wgt = p * q_oa[2 * n + v] * q_sw[2 * n + w];
#pragma unroll
                                for (int m = 0; m < RMX_MAXM; ++m)
                                    if (m < d.M) g[m] += wgt * (rp[m] * dp);
                            }
                        }
                    }
                }
            }
        }
    }
    block_sum_store<RMX_MAXM>(g, part, gridDim.x);
    if (flags) atomicOr(dflags, flags);
}

// Single-(n,s) scalar accessors (bpmodel.pyx:686-896), evaluated on the device.
__global__ void k_eval(RmxTables T, RmxDims d, const RmxParams* __restrict__ Pg, int g, int what, int n, int s, int u, int v, int w,
                       double* __restrict__ out, int* __restrict__ dflags) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int flags = 0;
    const RmxParams& P = Pg[g];
    const uint8_t* tmask = T.tmask + (int64_t)g * d.N;
    const uint8_t* amask = T.amask + (int64_t)g * d.N;
    for (int m = 0; m < RMX_MAXM; ++m) out[m] = 0.;
    StateCN c = load_state(T, d, n, s);
    const bool special = !P.normal_contamination;
    double l = T.l[n];
    switch (what) {
    case 0: out[0] = expected_total_reads(P, d, c, l); break;
    case 1: for (int m = 0; m < d.M; ++m) out[m] = l * (double)c.tot[m]; break;
    case 2:
    case 3: {
        double md, td;
        double pr = expected_allele_ratio(P, d, c, &md, &td);
        if (td <= 0.) { flags |= RMX_F_TOTAL_DEPTH; break; }
        if (what == 2) out[0] = pr;
        else for (int m = 0; m < d.M; ++m) out[m] = ((double)c.a0[m] * td - md * (double)c.tot[m]) / (td * td);
        break;
    }
    case 4: out[0] = -1.0 * (double)T.nsub[s] * l * P.divergence_weight; break;
    case 5: {
        if (!tmask[n]) break;
        bool hd = special && c.hdel;
        double mu = hd ? P.negbin_hdel_mu : expected_total_reads(P, d, c, l);
        double rr = hd ? P.negbin_hdel_r[u] : P.negbin_r[u];
        double x = T.x[n];
        double p = mu / (rr + mu);
        if (p < 0. || p > 1.) p = 0.5;
        double val = lgamma(x + rr) - lgamma(x + 1.) - lgamma(rr) + x * log(p) + rr * log(1. - p);
        if (isnan(val)) flags |= RMX_F_NB_NAN;
        out[0] = val;
        break;
    }
    case 6: {
        if (!tmask[n] || (special && c.hdel)) break;
        double mu = expected_total_reads(P, d, c, l);
        double rr = P.negbin_r[u], x = T.x[n];
        double dmu = x / mu - (rr + x) / (rr + mu);
        if (isnan(dmu)) flags |= RMX_F_NB_PARTIAL_NAN;
        for (int m = 0; m < d.M; ++m) out[m] = (l * (double)c.tot[m]) * dmu;
        break;
    }
    case 7: {
        if (!amask[n]) break;
        double p = 0.;
        if (!c.hdel) {
            double md, td;
            p = expected_allele_ratio(P, d, c, &md, &td);
            if (td <= 0.) { flags |= RMX_F_TOTAL_DEPTH; break; }
        }
        bool loh = special && c.loh;
        if (loh) {
            if (p == 0.) p = P.betabin_loh_p;
            else if (p == 1.) p = 1. - P.betabin_loh_p;
            else { flags |= RMX_F_LOH_P; break; }
        }
        double Mv = loh ? P.betabin_loh_M[v] : P.betabin_M[v];
        double y0 = T.y0[n], y1 = T.y1[n], na = y0 + y1;
        if (na == 0.) break;
        double k = w ? y1 : y0;
        if (p <= 0. || (1. - p) <= 0.) { flags |= RMX_F_BB_P_RANGE; break; }
        double val = lgamma(na + 1.) - lgamma(k + 1.) - lgamma(na - k + 1.) + lgamma(k + Mv * p) +
                     lgamma(na - k + Mv * (1. - p)) - lgamma(na + Mv) - lgamma(Mv * p) - lgamma(Mv * (1. - p)) +
                     lgamma(Mv);
        if (isnan(val)) flags |= RMX_F_BB_NAN;
        out[0] = val;
        break;
    }
    case 8: {
        if (!amask[n] || (special && c.loh)) break;
        double md, td;
        double pr = expected_allele_ratio(P, d, c, &md, &td);
        if (td <= 0.) { flags |= RMX_F_TOTAL_DEPTH; break; }
        double Mv = P.betabin_M[v];
        double y0 = T.y0[n], y1 = T.y1[n], na = y0 + y1;
        if (na == 0.) break;
        double k = w ? y1 : y0;
        if (pr <= 0. || (1. - pr) <= 0.) { flags |= RMX_F_BB_P_RANGE; break; }
        double dp = Mv * digamma_as103(k + Mv * pr, &flags) + (-Mv) * digamma_as103(na - k + Mv * (1. - pr), &flags) -
                    Mv * digamma_as103(Mv * pr, &flags) - (-Mv) * digamma_as103(Mv * (1. - pr), &flags);
        if (isnan(dp)) flags |= RMX_F_BB_PARTIAL_NAN;
        for (int m = 0; m < d.M; ++m) out[m] = (((double)c.a0[m] * td - md * (double)c.tot[m]) / (td * td)) * dp;
        break;
    }
    default: break;
    }
    if (flags) atomicOr(dflags, flags);
}

// ---------------------------------------------------------------------------------------------
// FP64 peak: 8 independent DFMA chains per thread, 1024 iterations, 148*8 CTAs of 256 threads.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_dfma_peak(double* __restrict__ out, double a, double b, int iters) {
    double x0 = threadIdx.x, x1 = x0 + 1., x2 = x0 + 2., x3 = x0 + 3., x4 = x0 + 4., x5 = x0 + 5., x6 = x0 + 6., x7 = x0 + 7.;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true; keeps the chains live
}

const char* rmx_measure_fp64_peak(int device, double* tflops) {
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cudaGetErrorString(e);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    double* buf = nullptr;
    const int grid = sms * 8, iters = 4096;
    e = cudaMalloc((void**)&buf, (size_t)grid * 256 * sizeof(double));
    if (e != cudaSuccess) return cudaGetErrorString(e);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0, 0);
        k_dfma_peak<<<grid, 256>>>(buf, 0.999999, 1e-9, iters);
        cudaEventRecord(e1, 0);
        e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double tf = 2.0 * 8.0 * iters * (double)grid * 256.0 / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    if (e != cudaSuccess) return cudaGetErrorString(e);
    *tflops = best;
    return nullptr;
}

// ---------------------------------------------------------------------------------------------
// launchers (g0 = first fit, gn = number of fits; blockIdx.y = fit)
// ---------------------------------------------------------------------------------------------
static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// constructor state: out[i] = (i & 1) ? v1 : v0 (v0 == v1 for a plain fill), grid-stride
__global__ void k_fill2(double* __restrict__ out, int64_t n, double v0, double v1) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = (i & 1) ? v1 : v0;
}

void rmx_launch_fill2(cudaStream_t st, double* out, int64_t n, double v0, double v1) {
    if (n <= 0) return;
    int64_t want = (n + 255) / 256;
    k_fill2<<<(int)(want < 148 * 16 ? want : 148 * 16), 256, 0, st>>>(out, n, v0, v1);
}

static const RmxFitSel g_no_sel = {};      // n == 0: the launch addresses the range [g0, g0 + gn)

void rmx_launch_segconst(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                         const int* rows_all, const int* Rn, int Rmax, double* segc, const double* segk, int planes,
                         const RmxFitSel* sel) {
    if (Rmax <= 0 || gn <= 0) return;
    k_segconst_rows<<<dim3(cdiv(Rmax, 128), gn), 128, 0, st>>>(T, d, P, g0, rows_all, Rn, segc, segk, planes,
                                                               sel ? *sel : g_no_sel);
}
void rmx_launch_segk(cudaStream_t st, const RmxTables& T, const RmxDims& d, double* segk) {
    if (d.N <= 0) return;
    k_segk<<<cdiv(d.N, 128), 128, 0, st>>>(T, d, segk);
}

void rmx_launch_lltable(cudaStream_t st, int sm_count, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0,
                        int gn, const double* segc, double* LL, int* dflags) {
    int64_t NS = (int64_t)d.N * d.S;
    if (NS <= 0 || gn <= 0) return;
    int64_t want = (NS + 255) / 256;
    int64_t cap = ((int64_t)sm_count * 8 + gn - 1) / gn;    // whole grid sized in multiples of the SM count
    if (cap < 1) cap = 1;
    int grid = (int)(want < cap ? want : cap);
    const int segmax = 256 / d.S + 2;
    const size_t smem = (size_t)segmax * (LLT_SEGW * sizeof(double) + sizeof(int)) + (size_t)d.S * d.M * 2 + 16;
    k_lltable<<<dim3(grid, gn), 256, smem, st>>>(T, d, P, g0, segc, LL, dflags, segmax);
}

static size_t lltable2_smem(const RmxDims& d, int Q) {
    return (size_t)RMX_LW_SIZE * 16 + (size_t)ST_NF * Q * d.S * 8 + (size_t)Q * d.S + 16;
}
int rmx_lltable2_ok(const RmxDims& d, int Q) { return lltable2_smem(d, Q) <= 72 * 1024; }

void rmx_launch_lltable2(cudaStream_t st, int sm_count, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0,
                         int gn, const double* segc, const RmxFast& Fp, double* LL, int* dflags, int planes,
                         const RmxFitSel* selp) {
    int64_t NS = (int64_t)d.N * d.S;
    if (NS <= 0 || gn <= 0) return;
    const int QS = Fp.Q * d.S;
    const RmxFitSel& sel = selp ? *selp : g_no_sel;
    k_statetab<<<dim3(cdiv(QS, 128), gn), 128, 0, st>>>(T, d, P, g0, Fp, sel);
    k_segfast<<<dim3(cdiv(d.N, 128), gn), 128, 0, st>>>(T, d, P, g0, segc, Fp, sel);
    int64_t want = (NS + 255) / 256;
    int64_t cap = ((int64_t)sm_count * 12 + gn - 1) / gn;   // whole grid sized in multiples of the SM count
    if (cap < 1) cap = 1;
    int grid = (int)(want < cap ? want : cap);
    static RmxPerDeviceOnce attr_once;
    if (attr_once.need()) {
        cudaFuncSetAttribute(k_lltable2, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024);
    }
    k_lltable2<<<dim3(grid, gn), 256, lltable2_smem(d, Fp.Q), st>>>(T, d, P, g0, segc, Fp, LL, dflags, planes, sel);
}

void rmx_launch_frame(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                      const double* LL, const double* q_ot, const double* q_oa, const double* q_sw, double* F,
                      double* fexp, double* mF, int* dflags) {
    if (d.N <= 0 || gn <= 0) return;
    const dim3 grid(cdiv((int64_t)d.N * 32, 256), gn);
    const int ks = (d.S + 31) / 32;
    if (ks == 1) k_frame_reg<1><<<grid, 256, 0, st>>>(T, d, P, g0, LL, q_ot, q_oa, q_sw, F, fexp, mF, dflags);
    else if (ks == 2) k_frame_reg<2><<<grid, 256, 0, st>>>(T, d, P, g0, LL, q_ot, q_oa, q_sw, F, fexp, mF, dflags);
    else if (ks == 3) k_frame_reg<3><<<grid, 256, 0, st>>>(T, d, P, g0, LL, q_ot, q_oa, q_sw, F, fexp, mF, dflags);
    else if (ks == 4) k_frame_reg<4><<<grid, 256, 0, st>>>(T, d, P, g0, LL, q_ot, q_oa, q_sw, F, fexp, mF, dflags);
    else k_frame<<<grid, 256, 0, st>>>(T, d, P, g0, LL, q_ot, q_oa, q_sw, F, fexp, mF, dflags);
}

// k_qsums with the row in registers (S <= 32 KS): every load of the segment's row -- the posterior and the planes that are
// summed -- is issued before the first multiply-add, instead of one (posterior, planes) group per trip of a loop whose
// trips wait for their own loads.  Same products in the same order per lane, same warp reduction: same values.
template <int KS>
__global__ void __launch_bounds__(256) k_qsums_reg(RmxDims d, int g0, const double* __restrict__ LL_all,
                                                   const double* __restrict__ post_all, double* __restrict__ A_all,
                                                   int planes_all, const __grid_constant__ RmxFitSel sel) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    const int g = RMX_FIT_SEL(sel);
    const int planes = RMX_PLANES_SEL(sel, planes_all);
    const int n = warp;
    const int64_t NS = (int64_t)d.N * d.S;
    const double* LL = LL_all + (int64_t)g * RMX_NTERM * NS + (int64_t)n * d.S;
    const double* post = post_all + (int64_t)g * NS + (int64_t)n * d.S;
    const int tmask6 = (planes & 1) | (planes & 2) | ((planes & 4) ? 12 : 0) | ((planes & 8) ? 48 : 0);
    double p[KS], ll[RMX_NTERM][KS];
#pragma unroll
    for (int k = 0; k < KS; ++k) {
        const int s = lane + 32 * k;
        p[k] = s < d.S ? post[s] : 0.;
#pragma unroll
        for (int t = 0; t < RMX_NTERM; ++t) ll[t][k] = (s < d.S && ((tmask6 >> t) & 1)) ? LL[(int64_t)t * NS + s] : 0.;
    }
    double a[RMX_NTERM];
#pragma unroll
    for (int t = 0; t < RMX_NTERM; ++t) a[t] = 0.;
#pragma unroll
    for (int k = 0; k < KS; ++k)
        if (lane + 32 * k < d.S) {
#pragma unroll
            for (int t = 0; t < RMX_NTERM; ++t)
                if ((tmask6 >> t) & 1) a[t] = fma(p[k], ll[t][k], a[t]);
        }
#pragma unroll
    for (int t = 0; t < RMX_NTERM; ++t)
        if ((tmask6 >> t) & 1) a[t] = warp_sum(a[t]);
    if (lane < RMX_NTERM && ((tmask6 >> lane) & 1)) {
        double v = a[0];
#pragma unroll
        for (int t = 1; t < RMX_NTERM; ++t) v = lane == t ? a[t] : v;
        A_all[((int64_t)g * d.N + n) * RMX_NTERM + lane] = v;
    }
}

void rmx_launch_qsums(cudaStream_t st, const RmxDims& d, int g0, int gn, const double* LL, const double* post,
                      double* A, int planes, const RmxFitSel* sel) {
    if (d.N <= 0 || gn <= 0) return;
    const dim3 grid(cdiv((int64_t)d.N * 32, 256), gn);
    const RmxFitSel& fs = sel ? *sel : g_no_sel;
    const int ks = (d.S + 31) / 32;
    if (ks <= 1) k_qsums_reg<1><<<grid, 256, 0, st>>>(d, g0, LL, post, A, planes, fs);
    else if (ks == 2) k_qsums_reg<2><<<grid, 256, 0, st>>>(d, g0, LL, post, A, planes, fs);
    else if (ks == 3) k_qsums_reg<3><<<grid, 256, 0, st>>>(d, g0, LL, post, A, planes, fs);
    else if (ks == 4) k_qsums_reg<4><<<grid, 256, 0, st>>>(d, g0, LL, post, A, planes, fs);
    else if (ks <= 6) k_qsums_reg<6><<<grid, 256, 0, st>>>(d, g0, LL, post, A, planes, fs);
    else k_qsums<<<grid, 256, 0, st>>>(d, g0, LL, post, A, planes, fs);
}

int rmx_round_ok(const RmxDims& d) { return d.S <= 192; }

void rmx_launch_round(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                      const double* LL, const double* ahat, const double* bhat, double* post, double* A, double* q_ot,
                      double* q_oa, double* q_sw, double prior_ot, double prior_oa, int next, int head, double* F,
                      double* fexp, double* mF, int* dflags) {
    if (d.N <= 0 || gn <= 0) return;
    const dim3 grid(cdiv((int64_t)d.N * 32, 256), gn);
    const int ks = (d.S + 31) / 32;
#define RMX_ROUND(KS) k_round<KS><<<grid, 256, 0, st>>>(T, d, P, g0, LL, ahat, bhat, post, A, q_ot, q_oa, q_sw, \
                                                       log(1. - prior_ot), log(prior_ot), log(1. - prior_oa), log(prior_oa), next, head, F, fexp, mF, dflags)
    if (ks <= 1) RMX_ROUND(1);
    else if (ks == 2) RMX_ROUND(2);
    else if (ks == 3) RMX_ROUND(3);
    else if (ks == 4) RMX_ROUND(4);
    else RMX_ROUND(6);       // 129 .. 192 states (max copy number 8: S = 165), two CTAs per SM
#undef RMX_ROUND
}

void rmx_launch_qupdate(cudaStream_t st, const RmxDims& d, int g0, int gn, int which, const double* A,
                        const double* q_oa, const double* q_sw, double prior, double* out) {
    if (d.N <= 0 || gn <= 0) return;
    k_qupdate<<<dim3(cdiv(d.N, 256), gn), 256, 0, st>>>(d, g0, which, A, q_oa, q_sw, log(1. - prior), log(prior), out);
}

void rmx_launch_reduce_cols(cudaStream_t st, const double* in, int64_t n, int ncol, int g0, int gn,
                            int64_t in_fit_stride, double* out, int out_fit_stride) {
    if (ncol <= 0 || gn <= 0) return;
    k_reduce_cols<<<dim3(ncol, gn), 1024, 0, st>>>(in, n, ncol, g0, in_fit_stride, out, out_fit_stride);
}

void rmx_launch_elbo_rows(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                          const double* A, const double* post, const double* F, const double* q_ot, const double* q_oa,
                          const double* q_sw, double prior_ot, double prior_oa, double* rowv) {
    if (d.N <= 0 || gn <= 0) return;
    k_elbo_rows<<<dim3(cdiv((int64_t)d.N * 32, 256), gn), 256, 0, st>>>(T, d, P, g0, A, post, F, q_ot, q_oa, q_sw,
                                                                        prior_ot, prior_oa, rowv);
}

void rmx_launch_ell_rows_from_sums(cudaStream_t st, const RmxDims& d, int g0, int gn, const double* A, const double* q_ot,
                                   const double* q_oa, const double* q_sw, double* out, int64_t out_fit_stride) {
    if (d.N <= 0 || gn <= 0) return;
    k_ell_rows_from_sums<<<dim3(cdiv(d.N, 256), gn), 256, 0, st>>>(d, g0, A, q_ot, q_oa, q_sw, out, out_fit_stride);
}
void rmx_launch_plogp(cudaStream_t st, const double* p, int64_t n, int g0, int gn, double* out, int64_t out_fit_stride) {
    if (n <= 0 || gn <= 0) return;
    k_plogp<<<dim3(cdiv(n, 256), gn), 256, 0, st>>>(p, n, g0, out, out_fit_stride);
}

int rmx_ell_sample_ok(int Rmax) { return Rmax > 0 && Rmax <= RMX_ELL_CHUNKS * RMX_ELL_MAXROWS; }
int rmx_ell_sample_chunks() { return RMX_ELL_CHUNKS; }
void rmx_launch_ell_sample(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                           const int* rows_all, const int* Rn, const double* post, const double* q_ot, const double* q_oa,
                           const double* q_sw, const double* rowc, double* out, int out_fit_stride, int* dflags,
                           int planes) {
    if (gn <= 0) return;
    k_ell_sample<<<dim3(RMX_ELL_CHUNKS, gn), 256, 0, st>>>(T, d, P, g0, rows_all, Rn, post, q_ot, q_oa, q_sw, rowc, out,
                                                          out_fit_stride, dflags, planes);
}
void rmx_launch_ell_rowcache(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, const int* rows_all,
                             const int* Rn, int Rmax, double* rowc) {
    if (gn <= 0 || Rmax <= 0) return;
    k_ell_rowcache<<<dim3(cdiv(Rmax, 128), gn), 128, 0, st>>>(T, d, g0, rows_all, Rn, rowc);
}
int rmx_ell_blocks(const RmxDims& d, int R) { return cdiv((int64_t)R * d.S, 256); }

void rmx_launch_ell(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                    const int* rows_all, const int* Rn, int Rmax, const double* segc, const double* post,
                    const double* q_ot, const double* q_oa, const double* q_sw, double* part, int64_t part_fit_stride,
                    int* dflags, int planes) {
    if (Rmax <= 0 || gn <= 0) return;
    k_ell<<<dim3(rmx_ell_blocks(d, Rmax), gn), 256, 0, st>>>(T, d, P, g0, rows_all, Rn, segc, post, q_ot, q_oa, q_sw,
                                                            part, part_fit_stride, dflags, planes);
}

void rmx_launch_ell_grad(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g0, int gn,
                         const int* rows_all, const int* Rn, int Rmax, const double* post, const double* q_ot,
                         const double* q_oa, const double* q_sw, double* part, int64_t part_fit_stride, int* dflags) {
    if (Rmax <= 0 || gn <= 0) return;
    k_ell_grad<<<dim3(rmx_ell_blocks(d, Rmax), gn), 256, 0, st>>>(T, d, P, g0, rows_all, Rn, post, q_ot, q_oa, q_sw, part,
                                                                 part_fit_stride, dflags);
}

void rmx_launch_eval(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxParams* P, int g, int what, int n,
                     int s, int u, int v, int w, double* out, int* dflags) {
    k_eval<<<1, 32, 0, st>>>(T, d, P, g, what, n, s, u, v, w, out, dflags);
}
