// Chain kernels: the structured part of the variational update, batched over the fits of a grid.
//
//   k_build_class  : the S x S log-transition matrix of "plain" rows, exp'd once     (bpmodel.pyx:640-684)
//   k_prep_brk     : per-breakend-row expected transition tables under p_breakpoint  (bpmodel.pyx:658-664)
//   k_chain        : forward / backward sum-product messages; one CTA per (chromosome chain,
//                    direction, tile of GT fits), in the scaled probability domain    (bpmodel.pyx:1213-1246)
//   k_post         : posterior marginals                                              (bpmodel.pyx:948-950)
//   k_hist         : joint-marginal histograms by total-copy difference at breakend
//                    rows                                                             (bpmodel.pyx:619-633, 954-960)
//   k_brk_update   : update_p_breakpoint                                              (bpmodel.pyx:964-983)
//   k_pairsum      : sum_ij joint[n,i,j] * log_transmat[n,i,j] per row (ELBO terms)   (bpmodel.pyx:1052, 1112-1115)
//   k_viterbi      : max_product + traceback per chain                                (bpmodel.pyx:1296-1333)
//
// The reference materialises three (N-1,S,S) arrays; nothing here does.  A log-domain message
//   alpha[n,j] = LSE_i(alpha[n-1,i] + T[i,j]) + F[n,j]
// is evaluated as a scaled matrix-vector product  u_n = ((u_{n-1} / s_{n-1})^T E) .* f_n  with
// E = exp(T) held in registers across the whole chain, f_n = exp(F[n]-max F[n]), s_n = sum(u_n); the
// chain's log normaliser is sum_n (log s_n + max F[n]).  Every fit of an initialisation grid has the same
// E (it depends on the state table and the penalty only), so one CTA advances GT fits in lock step and
// each E entry loaded into a register is used GT times per step.  Chromosomes are coupled only through
// telomere rows whose log_transmat is identically zero, so chains are exactly independent.
#include <algorithm>
#include <cstdlib>
#include <mutex>
#include "../../include/remixt_b200.h"
#include "common.cuh"
#include "kernels.h"

__device__ __forceinline__ double ftrans(int dd, int tm) {
    // calc_transition (bpmodel.pyx:607-616)
    return tm == 0 ? (double)(dd < 0 ? -dd : dd) : (dd == 0 ? 0. : 1.);
}

__device__ __forceinline__ int tab_cn(const int8_t* tab, int M, int s, int m, int a) { return tab[(s * M + m) * 2 + a]; }

// ---------------------------------------------------------------------------------------------
// class matrices
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_build_class(const int8_t* __restrict__ tab, int S, int M,
                                                     const int8_t* __restrict__ class_nrm, double pen, int tm,
                                                     double* __restrict__ Tmat, double* __restrict__ Emat,
                                                     double* __restrict__ flipc, double* __restrict__ flipE) {
    int c = blockIdx.y;
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)S * S) return;
    int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
    const int8_t* cn = class_nrm + 4 * c;   // normal copies of segment n (2) and n+1 (2)
    double t = 0.;
    double acc[2] = {0., 0.};
    for (int m = 0; m < M; ++m) {
        int ci[2], cj[2];
        if (m == 0) {
            ci[0] = cn[0]; ci[1] = cn[1]; cj[0] = cn[2]; cj[1] = cn[3];
        } else {
            ci[0] = tab_cn(tab, M, i, m, 0); ci[1] = tab_cn(tab, M, i, m, 1);
            cj[0] = tab_cn(tab, M, j, m, 0); cj[1] = tab_cn(tab, M, j, m, 1);
        }
        double ft = ftrans((ci[0] + ci[1]) - (cj[0] + cj[1]), tm);
        t = __dadd_rn(t, __dmul_rn(-pen, ft));
        for (int flip = 0; flip < 2; ++flip) {
            for (int a = 0; a < 2; ++a) acc[flip] += ftrans(ci[a] - cj[flip ? 1 - a : a], tm);
            acc[flip] -= ft;
        }
    }
    double fc = acc[0] <= acc[1] ? acc[0] : acc[1];
    t = __dadd_rn(t, __dmul_rn(-pen, fc));
    int64_t o = (int64_t)c * S * S + idx;
    Tmat[o] = t;
    Emat[o] = exp(t);
    flipc[o] = fc;
    flipE[o] = exp(__dmul_rn(-pen, fc));
}

// pd[r][m][slot] = sum_b p_breakpoint[k,b] * f(d - orient*brk[b,m]), d = slot - doff (bpmodel.pyx:660-664)
__global__ void k_prep_brk(RmxDims d, int g0, const int* __restrict__ brk_rows, const int* __restrict__ bidx,
                           const int* __restrict__ borient, const int8_t* __restrict__ brk,
                           const double* __restrict__ p_brk_all, double pen, int tm, int64_t fit_stride,
                           double* __restrict__ pd_all, double* __restrict__ wexp_all) {
    const int g = RMX_FIT();
    const double* p_brk = p_brk_all + (int64_t)g * d.K * d.B;
    double* pd = pd_all + (int64_t)g * fit_stride;
    double* wexp = wexp_all + (int64_t)g * fit_stride;
    int r = blockIdx.x;
    int n = brk_rows[r];
    int k = bidx[n], o = borient[n];
    for (int e = threadIdx.x; e < d.M * d.D; e += blockDim.x) {
        int m = e / d.D, slot = e - m * d.D;
        int dd = slot - d.doff;
        double acc = 0.;
        for (int b = 0; b < d.B; ++b)
            acc = __dadd_rn(acc, __dmul_rn(p_brk[(int64_t)k * d.B + b], ftrans(dd - o * (int)brk[b * d.M + m], tm)));
        pd[((int64_t)r * d.M + m) * d.D + slot] = acc;
        wexp[((int64_t)r * d.M + m) * d.D + slot] = exp(-pen * acc);
    }
}

// exact reference-order log transition of a breakend row (bpmodel.pyx:658-684)
__device__ __forceinline__ double brk_logT(const double* __restrict__ pdr, const int8_t* __restrict__ tots, int M, int D,
                                           int doff, int d0, int i, int j, double fc, double pen) {
    double t = 0.;
    for (int m = 0; m < M; ++m) {
        int dd = (m == 0) ? d0 : (int)tots[i * M + m] - (int)tots[j * M + m];
        t = __dadd_rn(t, __dmul_rn(-pen, pdr[m * D + dd + doff]));
    }
    return __dadd_rn(t, __dmul_rn(-pen, fc));
}

// exp of the above, factorised: flipE * prod_m wexp[m][d_m]
__device__ __forceinline__ double brk_E(const double* __restrict__ wr, const int8_t* __restrict__ tots, int M, int D,
                                        int doff, int d0, int i, int j, double fe) {
    double e = fe * wr[d0 + doff];
    for (int m = 1; m < M; ++m) e *= wr[m * D + (int)tots[i * M + m] - (int)tots[j * M + m] + doff];
    return e;
}


// ---------------------------------------------------------------------------------------------
// K2: forward / backward messages for GT fits in lock step.
//
// grid.x = ((chain_rank * 2 + dir) * ntiles + tile): chains are ranked longest first so the long
// sequential jobs start first.  Thread t -> (column group jg = t % NJ, row slice q = t / NJ); the
// thread owns the RR x CW tile E[q*RR .. q*RR+RR) x [jg*CW .. jg*CW+CW) of the transition matrix, in
// registers for the whole chain (REG) or streamed from L2 every step (!REG, large state spaces).
//
// Per step, for every fit g of the tile:  acc[g][c] = sum_k u[g][i0+k] * E[i0+k][j0+c]   (main loop),
// partial sums of the Q row slices are combined through shared memory, multiplied by 1/s_prev[g]
// (the LAGGED normaliser: the sum of the previous message, reduced while the main loop runs) and by
// f_n; the new message goes to shared memory (next step's input) and to HBM (ahat / bhat).
// Messages are stored un-normalised by their own sum; every consumer normalises (k_post, pair
// weights), and logZ accumulates log s_n.
//
// shared: u[2][GT][Sp] | part[Q][GT][CW][NJ] | red[2][GT][NCH] | wtab[GT][M*D] | tots[S*M] (int8)
// ---------------------------------------------------------------------------------------------
struct ChainArgs {
    RmxDims d;
    RmxChains ch;
    RmxRows rows;
    RmxTransSpec sp;       // pd / wexp are per fit: fit stride sp_fit_stride
    int64_t sp_fit_stride;
    const int8_t* tab;
    int g0, gn;            // fits of this launch
    int ntiles;
    int Q, NJ, rr;         // thread layout (rr = rows per slice; == RR when REG)
    int fast_brk;          // breakend fast path allowed (flip costs fit the table)
    const double* fexp;    // (G,N,S)
    const double* mF;      // (G,N)
    double* ahat;          // (G,N,S)
    double* bhat;          // (G,N,S)
    double* chain_logz;    // (G,NC)
    int* dflags;
};

// biased exponent field of a non-negative double (0 for zero / denormal, 2047 for inf / NaN)
__device__ __forceinline__ int dexp_field(double v) { return (__double2hiint(v) >> 20) & 0x7ff; }
// 2^(1023 - be): the power of two that brings a value with exponent field be into [1, 2)
__device__ __forceinline__ double pow2_scale(int be) { return __hiloint2double((2046 - be) << 20, 0); }

#define RMX_EPI 3    // epilogue slots per thread (3 * threads >= GT * SJ, checked by the launcher)
#define RMX_FETAB 64 // flip-cost table entries of the breakend fast path

// QT > 0: the number of row slices is the compile-time constant QT (register-resident E); QT == 0: A.Q.
template <int RR, int CW, int GT, int QT, bool REG>
__global__ void __launch_bounds__(REG ? 256 : 512) k_chain(ChainArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const RmxDims& d = A.d;
    const int S = d.S, M = d.M, NJ = A.NJ;
    const int Q = QT > 0 ? QT : A.Q;
    const int rr = REG ? RR : A.rr;
    const int Sp = Q * rr;                 // padded number of rows (>= S)
    const int SJ = (S + 31) & ~31;         // epilogue slots per fit
    const int NCH = SJ >> 5;               // warps (chunks) per fit in the epilogue
    const int MD = M * d.D;
    const int T = blockDim.x;
    // breakend fast path: two tumour clones, combined (d1,d2) table index fits a byte
    const bool fast_brk = REG && M == 3 && d.D * d.D <= 256 && A.fast_brk;
    double* u = reinterpret_cast<double*>(smem_raw);        // [2][GT][Sp]   messages (double buffered)
    double* part = u + 2 * GT * Sp;                          // [Q][GT][CW][NJ] slice partial sums
    double* wtab = part + Q * GT * CW * NJ;                  // [GT][M*D]     breakend tables of the step
    double* redd = wtab + GT * MD;                           // [GT][NCH]     sums of the last message
    double* w2 = redd + GT * NCH;                            // [GT][256]     fast path: w1[d1]*w2[d2]
    double* fetab = w2 + (fast_brk ? GT * 256 : 0);          // [RMX_FETAB]   fast path: exp(-pen*flipcost)
    int* redi = reinterpret_cast<int*>(fetab + (fast_brk ? RMX_FETAB : 0));   // [2][GT][NCH] max exponent fields
    unsigned short* bix = reinterpret_cast<unsigned short*>(redi + 2 * GT * NCH);   // [RR*CW][T] fast path
    int8_t* tots = reinterpret_cast<int8_t*>(bix + (fast_brk ? RR * CW * T : 0));   // [S*M]

    const int tile = blockIdx.x % A.ntiles;
    const int cd = blockIdx.x / A.ntiles;
    const int dir = cd & 1, chain = A.ch.order[cd >> 1];
    const int s0 = A.ch.start[chain], s1 = A.ch.end[chain];
    const int t = threadIdx.x;
    const int jg = t % NJ, q = t / NJ;
    const bool act = q < Q;
    const int i0 = q * rr, j0 = jg * CW;
    const int64_t SS = (int64_t)S * S;
    const int64_t NS = (int64_t)d.N * S;
    const int gbase = A.g0 + tile * GT;                      // first fit of the tile
    const int gcount = min(GT, A.g0 + A.gn - gbase);         // live fits in the tile
    const int QS = GT * CW * NJ;                             // part stride between row slices

    for (int e = t; e < 2 * GT * Sp; e += T) u[e] = 0.;
    for (int e = t; e < 2 * GT * NCH; e += T) redi[e] = 1023;   // first message: max f = 1
    for (int e = t; e < S * M; e += T) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(A.tab[(s * M + m) * 2] + A.tab[(s * M + m) * 2 + 1]);
    }
    if (fast_brk)
        for (int e = t; e < RMX_FETAB; e += T) fetab[e] = exp(__dmul_rn(-A.sp.pen, (double)e));

    // ---- per-thread epilogue slots: slot -> (fit g, column j), consecutive j per lane, one fit per warp
    int sl_u[RMX_EPI], sl_p[RMX_EPI], sl_r[RMX_EPI], sl_rb[RMX_EPI];
    int64_t sl_g[RMX_EPI];
    const double* sl_mf[RMX_EPI];
    bool sl_ok[RMX_EPI], sl_lead[RMX_EPI];
#pragma unroll
    for (int e = 0; e < RMX_EPI; ++e) {
        const int slot = t + e * T;
        const int g = slot / SJ, j = slot - g * SJ;
        sl_ok[e] = slot < GT * SJ && j < S && g < gcount;
        sl_lead[e] = sl_ok[e] && j == 0;
        sl_u[e] = g * Sp + j;
        sl_p[e] = (g * CW + (j % CW)) * NJ + j / CW;
        sl_r[e] = slot < GT * SJ ? g * NCH + (j >> 5) : -1;
        sl_rb[e] = g * NCH;
        sl_g[e] = (int64_t)(gbase + g) * NS + j;
        sl_mf[e] = A.mF + (int64_t)(gbase + min(g, GT - 1)) * d.N;
    }
    __syncthreads();

    // ---- first segment of the sweep: u = f (forward: ahat = f; backward: bhat = 1)
    {
        const int n = dir == 0 ? s0 : s1;
#pragma unroll
        for (int e = 0; e < RMX_EPI; ++e)
            if (sl_ok[e]) {
                const int64_t o = sl_g[e] + (int64_t)n * S;
                const double v = A.fexp[o];
                if (dir == 0) A.ahat[o] = v; else A.bhat[o] = 1.;
                u[sl_u[e]] = v;
            }
        __syncthreads();
    }

    double ereg[REG ? RR : 1][REG ? CW : 1];
    int cur_class = -1, bix_class = -1;
    int cur = 0;
    // the lead thread of fit g (slot with j == 0) accumulates the fit's log normaliser:
    // logZ = log(sum of the last message) + ln2 * esum + sum_n max F[n]
    double mfsum[RMX_EPI];
    int esum[RMX_EPI];
#pragma unroll
    for (int e = 0; e < RMX_EPI; ++e) { mfsum[e] = 0.; esum[e] = 0; }
    int bad = 0;

    const int nsteps = s1 - s0;
    for (int step = 0; step < nsteps; ++step) {
        const int n = dir == 0 ? s0 + 1 + step : s1 - 1 - step;   // segment being produced
        const int row = dir == 0 ? n - 1 : n;                     // transition row n -> n+1
        const int prev = dir == 0 ? n - 1 : n + 1;                // segment whose message is the input
        const int cls = A.rows.tclass[row];
        const int br = A.rows.brow[row];
        const double* uin = u + cur * GT * Sp + i0;

        // this step's f, fetched early (used in the epilogue)
        double fpre[RMX_EPI];
#pragma unroll
        for (int e = 0; e < RMX_EPI; ++e) fpre[e] = sl_ok[e] ? A.fexp[sl_g[e] + (int64_t)n * S] : 0.;

        double acc[GT][CW];
#pragma unroll
        for (int g = 0; g < GT; ++g)
#pragma unroll
            for (int c = 0; c < CW; ++c) acc[g][c] = 0.;

        if (br < 0) {
            const double* E = A.sp.Emat + (int64_t)cls * SS;   // symmetric: E[i][j] == E[j][i]
            if (REG) {
                if (cls != cur_class) {
#pragma unroll
                    for (int k = 0; k < RR; ++k)
#pragma unroll
                        for (int c = 0; c < CW; ++c) {
                            const int i = i0 + k, j = j0 + c;
                            ereg[REG ? k : 0][REG ? c : 0] = (act && i < S && j < S) ? E[(int64_t)i * S + j] : 0.;
                        }
                    cur_class = cls;
                }
#pragma unroll
                for (int k = 0; k < RR; ++k) {
                    double v[GT];
#pragma unroll
                    for (int g = 0; g < GT; ++g) v[g] = uin[g * Sp + k];
#pragma unroll
                    for (int g = 0; g < GT; ++g)
#pragma unroll
                        for (int c = 0; c < CW; ++c) acc[g][c] = fma(v[g], ereg[REG ? k : 0][REG ? c : 0], acc[g][c]);
                }
            } else if (act) {
                const int kmax = min(rr, S - i0);
#pragma unroll 1
                for (int k = 0; k < kmax; ++k) {
                    double e[CW];
#pragma unroll
                    for (int c = 0; c < CW; ++c) e[c] = (j0 + c < S) ? E[(int64_t)(i0 + k) * S + j0 + c] : 0.;
#pragma unroll
                    for (int g = 0; g < GT; ++g) {
                        const double v = uin[g * Sp + k];
#pragma unroll
                        for (int c = 0; c < CW; ++c) acc[g][c] = fma(v, e[c], acc[g][c]);
                    }
                }
            }
        } else {
            // breakend row (rare): E_n[from][to] = flipE[cls][from][to] * prod_m wexp[g][r][m][tot_from - tot_to].
            // The m = 0 factor (normal copies) does not depend on the states: applied once at the end.
            const int sgn = dir == 0 ? 1 : -1;    // forward: from = row index; backward: from = column index
            for (int e = t; e < GT * MD; e += T) {
                const int g = e / MD, o = e - g * MD;
                const int gg = gbase + min(g, gcount - 1);
                wtab[e] = A.sp.wexp[(int64_t)gg * A.sp_fit_stride + (int64_t)br * MD + o];
            }
            if (fast_brk && cls != bix_class) {
                // per-entry table of this thread's tile: low byte = combined (d1,d2) index, high byte = flip cost
                const double* FC = A.sp.flipc + (int64_t)cls * SS;
#pragma unroll 1
                for (int kc = 0; kc < RR * CW; ++kc) {
                    const int ir = i0 + kc / CW, jc = j0 + kc % CW;
                    unsigned short v = 0xff00;                       // out of range: flip cost 255 -> weight 0
                    if (act && ir < S && jc < S) {
                        const int d1 = sgn * ((int)tots[ir * M + 1] - (int)tots[jc * M + 1]) + d.doff;
                        const int d2 = sgn * ((int)tots[ir * M + 2] - (int)tots[jc * M + 2]) + d.doff;
                        int fc = (int)FC[(int64_t)ir * S + jc];
                        fc = fc < RMX_FETAB - 1 ? fc : RMX_FETAB - 1;
                        v = (unsigned short)((fc << 8) | (d1 * d.D + d2));
                    }
                    bix[kc * T + t] = v;
                }
                bix_class = cls;
            }
            __syncthreads();
            if (fast_brk) {
                for (int e = t; e < GT * 256; e += T) {
                    const int g = e >> 8, ix = e & 255;
                    const int d1 = ix / d.D, d2 = ix - d1 * d.D;
                    w2[e] = d1 < d.D ? wtab[g * MD + d.D + d1] * wtab[g * MD + 2 * d.D + d2] : 0.;
                }
                __syncthreads();
#pragma unroll 1
                for (int k = 0; k < RR; ++k) {
                    double v[GT];
#pragma unroll
                    for (int g = 0; g < GT; ++g) v[g] = uin[g * Sp + k];
#pragma unroll
                    for (int c = 0; c < CW; ++c) {
                        const int code = bix[(k * CW + c) * T + t];
                        const int fc = code >> 8;
                        const double fe = fc < RMX_FETAB ? fetab[fc] : 0.;
                        const double* wp = w2 + (code & 255);
#pragma unroll
                        for (int g = 0; g < GT; ++g) acc[g][c] = fma(fe * v[g], wp[g * 256], acc[g][c]);
                    }
                }
            } else if (act) {
                const double* FE = A.sp.flipE + (int64_t)cls * SS;
                const int kmax = min(rr, S - i0);
#pragma unroll 1
                for (int k = 0; k < kmax; ++k) {
                    const int ir = i0 + k;
#pragma unroll 1
                    for (int c = 0; c < CW; ++c) {
                        const int jc = j0 + c;
                        if (jc >= S) break;
                        const double fe = FE[(int64_t)ir * S + jc];
                        double w[GT];
#pragma unroll
                        for (int g = 0; g < GT; ++g) w[g] = fe * uin[g * Sp + k];
                        for (int m = 1; m < M; ++m) {
                            const int dd = m * d.D + sgn * ((int)tots[ir * M + m] - (int)tots[jc * M + m]) + d.doff;
#pragma unroll
                            for (int g = 0; g < GT; ++g) w[g] *= wtab[g * MD + dd];
                        }
#pragma unroll
                        for (int cc = 0; cc < CW; ++cc)
                            if (cc == c) {
#pragma unroll
                                for (int g = 0; g < GT; ++g) acc[g][cc] += w[g];
                            }
                    }
                }
            }
            const int d0 = A.rows.d0[row];
#pragma unroll
            for (int g = 0; g < GT; ++g) {
                const double w0 = wtab[g * MD + d0 + d.doff];
#pragma unroll
                for (int c = 0; c < CW; ++c) acc[g][c] *= w0;
            }
        }
        if (act) {
            double* pp = part + q * QS + jg;
#pragma unroll
            for (int g = 0; g < GT; ++g)
#pragma unroll
                for (int c = 0; c < CW; ++c) pp[(g * CW + c) * NJ] = acc[g][c];
        }
        __syncthreads();

        // ---- epilogue: combine the Q slices, rescale by the power of two that brought the PREVIOUS message's
        // largest entry into [1, 2) (exact; its exponent goes to the log normaliser), multiply by f_n
        double* uout = u + (cur ^ 1) * GT * Sp;
        const int* redp = redi + cur * GT * NCH;
        int* redn = redi + (cur ^ 1) * GT * NCH;
#pragma unroll
        for (int e = 0; e < RMX_EPI; ++e) {
            if (sl_r[e] >= 0) {                                    // warp-uniform (GT*SJ is a multiple of 32)
                int be = redp[sl_rb[e]];
                if (REG) {                                         // S <= 128: at most 4 chunks
#pragma unroll
                    for (int c2 = 1; c2 < 4; ++c2)
                        if (c2 < NCH) be = max(be, redp[sl_rb[e] + c2]);
                } else {
                    for (int c2 = 1; c2 < NCH; ++c2) be = max(be, redp[sl_rb[e] + c2]);
                }
                double un = 0.;
                if (sl_ok[e]) {
                    const double* pq = part + sl_p[e];
                    double a0 = pq[0], a1 = 0.;
                    if (QT > 0) {
#pragma unroll
                        for (int qq = 1; qq < (QT > 0 ? QT : 1); qq += 2) {
                            a1 += pq[qq * QS];
                            if (qq + 1 < QT) a0 += pq[(qq + 1) * QS];
                        }
                    } else {
                        for (int qq = 1; qq < Q; ++qq) a0 += pq[qq * QS];
                    }
                    double v = (a0 + a1) * pow2_scale(be);
                    const int64_t o = sl_g[e] + (int64_t)n * S;
                    if (dir == 0) {
                        v *= fpre[e];
                        un = v;
                        A.ahat[o] = v;
                    } else {
                        un = v * fpre[e];
                        A.bhat[o] = v;
                    }
                    uout[sl_u[e]] = un;
                }
                const int mx = __reduce_max_sync(0xffffffffu, dexp_field(un));
                if ((t & 31) == 0) redn[sl_r[e]] = mx;
                if (sl_lead[e]) {
                    esum[e] += be - 1023;
                    mfsum[e] += sl_mf[e][prev];
                    if (be <= 0 || be >= 2047) bad = 1;
                }
            }
        }
        cur ^= 1;
        __syncthreads();
    }
    // ---- the last message's own sum closes the forward log normaliser
#pragma unroll
    for (int e = 0; e < RMX_EPI; ++e)
        if (sl_r[e] >= 0) {
            const double w = warp_sum(sl_ok[e] ? u[cur * GT * Sp + sl_u[e]] : 0.);
            if ((t & 31) == 0) redd[sl_r[e]] = w;
        }
    __syncthreads();
#pragma unroll
    for (int e = 0; e < RMX_EPI; ++e)
        if (sl_lead[e]) {
            double slast = 0.;
            for (int c2 = 0; c2 < NCH; ++c2) slast += redd[sl_rb[e] + c2];
            const int g = gbase + sl_rb[e] / NCH;
            const int last = dir == 0 ? s1 : s0;
            if (dir == 0) {
                const double logz = log(slast) + 0.693147180559945309417 * (double)esum[e] + mfsum[e] + sl_mf[e][last];
                A.chain_logz[(int64_t)g * A.ch.NC + chain] = logz;
                if (isnan(logz)) bad = 1;
            }
            if (!(slast > 0.) || isinf(slast)) bad = 1;
            if (bad) atomicOr(A.dflags, RMX_F_MSG_NAN);
        }
}

// posterior marginals: post[g,n] = normalise(ahat[g,n] .* bhat[g,n]), one warp per (fit, segment)
__global__ void __launch_bounds__(256) k_post(RmxDims d, int g0, const double* __restrict__ ahat,
                                              const double* __restrict__ bhat, double* __restrict__ post,
                                              int* __restrict__ dflags) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    int64_t base = ((int64_t)RMX_FIT() * d.N + warp) * d.S;
    double z = 0.;
    for (int s = lane; s < d.S; s += 32) z = __dadd_rn(z, __dmul_rn(ahat[base + s], bhat[base + s]));
    z = warp_sum(z);
    bool bad = !(z > 0.);
    for (int s = lane; s < d.S; s += 32) {
        double p = __ddiv_rn(__dmul_rn(ahat[base + s], bhat[base + s]), z);
        post[base + s] = p;
        bad |= isnan(p);
    }
    if (bad) atomicOr(dflags, RMX_F_POST_NAN);
}

// ---------------------------------------------------------------------------------------------
// pair weights of row n: w_ij = a[n,i] * E_n[i,j] * f[n+1,j] * b[n+1,j]   (joint = w / sum w)
// ---------------------------------------------------------------------------------------------
struct PairCtx {
    const double* a;     // smem [S]
    const double* g;     // smem [S]
    const double* wtab;  // smem [M*D] (breakend rows)
    const int8_t* tots;  // smem [S*M]
    const double* E;     // class Emat (plain rows) or flipE (breakend rows), or nullptr for all-ones
    int kind;            // 0 ones (telomere / uniform), 1 plain, 2 breakend
    int d0;
};

__device__ __forceinline__ double pair_w(const PairCtx& c, const RmxDims& d, int i, int j) {
    double e = 1.;
    if (c.kind == 1) e = c.E[(int64_t)i * d.S + j];
    else if (c.kind == 2) e = brk_E(c.wtab, c.tots, d.M, d.D, d.doff, c.d0, i, j, c.E[(int64_t)i * d.S + j]);
    return c.a[i] * e * c.g[j];
}

// loads a, g, wtab, tots into shared memory and fills ctx (all threads must call)
__device__ __forceinline__ void pair_setup(PairCtx& c, unsigned char* smem_raw, const RmxDims& d, const RmxRows& rows,
                                           const RmxTransSpec& sp, const int8_t* tab, int uniform, int n,
                                           const double* ahat, const double* bhat, const double* fexp) {
    // ahat / bhat / fexp and sp.pd / sp.wexp point at the fit's own slice
    const int S = d.S, M = d.M;
    double* a = reinterpret_cast<double*>(smem_raw);
    double* g = a + S;
    double* wtab = g + S;
    int8_t* tots = reinterpret_cast<int8_t*>(wtab + M * d.D);
    int cls = rows.tclass[n], br = rows.brow[n];
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        a[s] = uniform ? 1. / S : ahat[(int64_t)n * S + s];
        g[s] = uniform ? 1. / S : fexp[(int64_t)(n + 1) * S + s] * bhat[(int64_t)(n + 1) * S + s];
    }
    for (int e = threadIdx.x; e < S * M; e += blockDim.x) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(tab[(s * M + m) * 2] + tab[(s * M + m) * 2 + 1]);
    }
    c.kind = 0;
    c.E = nullptr;
    c.d0 = rows.d0[n];
    if (!uniform && cls >= 0) {
        if (br >= 0) {
            for (int e = threadIdx.x; e < M * d.D; e += blockDim.x) wtab[e] = sp.wexp[(int64_t)br * M * d.D + e];
            c.kind = 2;
            c.E = sp.flipE + (int64_t)cls * S * S;
        } else {
            c.kind = 1;
            c.E = sp.Emat + (int64_t)cls * S * S;
        }
    }
    c.a = a;
    c.g = g;
    c.wtab = wtab;
    c.tots = tots;
    __syncthreads();
}

__device__ __forceinline__ double block_sum_all(double v, double* red) {
    // deterministic CTA sum, result broadcast to all threads; red has >= 32 doubles
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    double s = 0.;
    int nw = (blockDim.x + 31) >> 5;
    for (int w = 0; w < nw; ++w) s += red[w];
    return s;
}

// K-hist: per breakend row r and clone m, hist[r][m][slot] = sum_{ij: tot_i[m]-tot_j[m] = d} joint[n,i,j]
// (bpmodel.pyx:628-633).  Accumulated in 2^-62 fixed point with integer shared-memory atomics so the
// result does not depend on the order in which threads arrive.
__global__ void __launch_bounds__(256) k_hist(RmxDims d, int g0, RmxRows rows, RmxTransSpec sp, int64_t sp_fit_stride,
                                              const int8_t* __restrict__ tab, int uniform,
                                              const double* __restrict__ ahat_all, const double* __restrict__ bhat_all,
                                              const double* __restrict__ fexp_all, double* __restrict__ hist_all,
                                              int* __restrict__ dflags) {
    const int g = RMX_FIT();
    const int64_t NSo = (int64_t)g * d.N * d.S;
    const double *ahat = ahat_all + NSo, *bhat = bhat_all + NSo, *fexp = fexp_all + NSo;
    double* hist = hist_all + (int64_t)g * sp_fit_stride;
    sp.pd += (int64_t)g * sp_fit_stride;
    sp.wexp += (int64_t)g * sp_fit_stride;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red[32];
    __shared__ unsigned long long hacc[RMX_MAXM * 64];
    const int S = d.S, M = d.M;
    int r = blockIdx.x;
    int n = rows.brk_rows[r];
    PairCtx c;
    pair_setup(c, smem_raw, d, rows, sp, tab, uniform, n, ahat, bhat, fexp);
    for (int e = threadIdx.x; e < M * d.D; e += blockDim.x) hacc[e] = 0ull;
    const int64_t SS = (int64_t)S * S;
    double z = 0.;
    for (int64_t idx = threadIdx.x; idx < SS; idx += blockDim.x) {
        int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
        z += pair_w(c, d, i, j);
    }
    z = block_sum_all(z, red);
    if (!(z > 0.) || isinf(z)) {
        if (threadIdx.x == 0) atomicOr(dflags, RMX_F_JOINT_NAN);
    }
    const double scale = 4611686018427387904.0 / z;   // 2^62 / Z
    const int nrm_d = c.d0;
    for (int64_t idx = threadIdx.x; idx < SS; idx += blockDim.x) {
        int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
        unsigned long long qv = __double2ull_rn(pair_w(c, d, i, j) * scale);
        if (qv) {
            atomicAdd(&hacc[nrm_d + d.doff], qv);
            for (int m = 1; m < M; ++m)
                atomicAdd(&hacc[m * d.D + (int)c.tots[i * M + m] - (int)c.tots[j * M + m] + d.doff], qv);
        }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < M * d.D; e += blockDim.x)
        hist[(int64_t)r * M * d.D + e] = (double)hacc[e] * 2.168404344971008868e-19;   // 2^-62
}

// K-hist, state spaces up to 128: no atomics.  The order of the states by each tumour clone's total copy number
// (ord) and the start of every total's group in that order (gstart) depend on the state table only and are built
// once per batch (k_hist_prep).  Per breakend row: (1) the S x S pair weights go to shared memory once (warp per
// row i, lanes along j; Z by a fixed-order CTA sum); (2) warp per (clone m, total t): column sums over the rows of
// that group, colsum[m][t][j] = sum_{i: tot_i[m] = t} W[i][j]; (3) warp per (m, slot):
// hist[m][slot] = sum_j colsum[m][tot_j[m] + slot - doff][j] / Z.  Every sum has a fixed order.
#define HIST2_TV 16     // totals per clone covered (cn_max + 2 <= 16)
#define HIST2_SP 128    // padded state count
__global__ void k_hist_prep(const int8_t* __restrict__ tab, int S, int M, uint8_t* __restrict__ ord,
                            int* __restrict__ gstart) {
    for (int e = threadIdx.x; e < M * S; e += blockDim.x) {
        const int m = e / S, j = e - m * S;
        if (m == 0) { ord[e] = (uint8_t)j; continue; }
        const int tj = tab[(j * M + m) * 2] + tab[(j * M + m) * 2 + 1];
        int rank = 0;
        for (int k = 0; k < S; ++k) {
            const int tk = tab[(k * M + m) * 2] + tab[(k * M + m) * 2 + 1];
            rank += (tk < tj || (tk == tj && k < j)) ? 1 : 0;
        }
        ord[m * S + rank] = (uint8_t)j;
    }
    for (int e = threadIdx.x; e < M * (HIST2_TV + 1); e += blockDim.x) {
        const int m = e / (HIST2_TV + 1), tv = e - m * (HIST2_TV + 1);
        int cnt = 0;
        for (int k = 0; k < S; ++k) cnt += (m > 0 && tab[(k * M + m) * 2] + tab[(k * M + m) * 2 + 1] < tv) ? 1 : 0;
        gstart[e] = cnt;
    }
}

// pair weights of the row in place over the staged transition factor: lanes keep their columns' g_j and table
// offsets in registers, the row's a_i and totals are warp-uniform; same operation order as pair_w / brk_E.
// MM = number of clones at compile time (0: any, loops not unrolled).
template <int MM>
__device__ __forceinline__ double hist2_weights(const PairCtx& c, const RmxDims& d, double* __restrict__ W, int lane,
                                                int wid, int nw) {
    const int S = d.S, M = MM > 0 ? MM : d.M;
    constexpr int MR = MM > 0 ? MM : RMX_MAXM;
    double gj[HIST2_SP / 32];
    int bj[HIST2_SP / 32][MR];
#pragma unroll
    for (int kk = 0; kk < HIST2_SP / 32; ++kk) {
        const int j = lane + 32 * kk;
        gj[kk] = j < S ? c.g[j] : 0.;
#pragma unroll
        for (int m = 1; m < MR; ++m)
            if (m < M) bj[kk][m] = m * d.D + d.doff - (j < S ? (int)c.tots[j * M + m] : 0);
    }
    const double w0 = c.kind == 2 ? c.wtab[c.d0 + d.doff] : 1.;
    double z = 0.;
    for (int i = wid; i < S; i += nw) {
        const double ai = c.a[i];
        int ti[MR];
#pragma unroll
        for (int m = 1; m < MR; ++m)
            if (m < M) ti[m] = (int)c.tots[i * M + m];
#pragma unroll
        for (int kk = 0; kk < HIST2_SP / 32; ++kk) {
            const int j = lane + 32 * kk;
            if (j < S) {
                double e = 1.;
                if (c.kind == 1) {
                    e = W[i * S + j];
                } else if (c.kind == 2) {
                    e = W[i * S + j] * w0;
#pragma unroll
                    for (int m = 1; m < MR; ++m)
                        if (m < M) e *= c.wtab[bj[kk][m] + ti[m]];
                }
                const double w = ai * e * gj[kk];
                W[i * S + j] = w;
                z += w;
            }
        }
    }
    return z;
}

__global__ void __launch_bounds__(256, 2) k_hist2(RmxDims d, int g0, RmxRows rows, RmxTransSpec sp, int64_t sp_fit_stride,
                                                 const int8_t* __restrict__ tab, const uint8_t* __restrict__ ord,
                                                 const int* __restrict__ gstart, int uniform,
                                                 const double* __restrict__ ahat_all, const double* __restrict__ bhat_all,
                                                 const double* __restrict__ fexp_all, double* __restrict__ hist_all,
                                                 int* __restrict__ dflags) {
    const int g = RMX_FIT();
    const int64_t NSo = (int64_t)g * d.N * d.S;
    const double *ahat = ahat_all + NSo, *bhat = bhat_all + NSo, *fexp = fexp_all + NSo;
    double* hist = hist_all + (int64_t)g * sp_fit_stride;
    sp.pd += (int64_t)g * sp_fit_stride;
    sp.wexp += (int64_t)g * sp_fit_stride;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red[32];
    const int S = d.S, M = d.M;
    const int r = blockIdx.x;
    const int n = rows.brk_rows[r];
    PairCtx c;
    pair_setup(c, smem_raw, d, rows, sp, tab, uniform, n, ahat, bhat, fexp);
    size_t off = (sizeof(double) * (2 * (size_t)S + (size_t)M * d.D) + (size_t)S * M + 15) & ~(size_t)15;
    double* W = reinterpret_cast<double*>(smem_raw + off);              // [S][S]
    double* colsum = W + (size_t)S * S;                                 // [M-1][HIST2_TV][HIST2_SP]
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5, nw = blockDim.x >> 5;
    // the row's S x S transition factor (class matrix, shared by every CTA: L2) lands in W asynchronously, one
    // round trip for the whole matrix instead of one per pair; the weights are then formed in place
    if (c.kind != 0) {
        const unsigned wbase = (unsigned)__cvta_generic_to_shared(W);
        for (int e = t; e < S * S; e += blockDim.x)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(wbase + 8u * (unsigned)e), "l"(c.E + e) : "memory");
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    double z;
    switch (M) {
    case 2: z = hist2_weights<2>(c, d, W, lane, wid, nw); break;
    case 3: z = hist2_weights<3>(c, d, W, lane, wid, nw); break;
    case 4: z = hist2_weights<4>(c, d, W, lane, wid, nw); break;
    default: z = hist2_weights<0>(c, d, W, lane, wid, nw); break;
    }
    z = block_sum_all(z, red);      // includes the barrier that publishes W
    if (!(z > 0.) || isinf(z)) {
        if (t == 0) atomicOr(dflags, RMX_F_JOINT_NAN);
    }
    for (int q = wid; q < (M - 1) * HIST2_TV; q += nw) {
        const int m = 1 + q / HIST2_TV, tv = q - (m - 1) * HIST2_TV;
        const int k0 = gstart[m * (HIST2_TV + 1) + tv], k1 = gstart[m * (HIST2_TV + 1) + tv + 1];
        double acc[HIST2_SP / 32] = {0., 0., 0., 0.};
        for (int k = k0; k < k1; ++k) {
            const double* Wi = W + (int)ord[m * S + k] * S;
#pragma unroll
            for (int kk = 0; kk < HIST2_SP / 32; ++kk)
                if (lane + 32 * kk < S) acc[kk] += Wi[lane + 32 * kk];
        }
#pragma unroll
        for (int kk = 0; kk < HIST2_SP / 32; ++kk) colsum[q * HIST2_SP + lane + 32 * kk] = acc[kk];
    }
    __syncthreads();
    const double inv = 1. / z;
    for (int q = wid; q < (M - 1) * d.D; q += nw) {
        const int m = 1 + q / d.D, slot = q - (m - 1) * d.D;
        const int dd = slot - d.doff;
        double v = 0.;
        for (int j = lane; j < S; j += 32) {
            const int tv = (int)c.tots[j * M + m] + dd;
            if (tv >= 0 && tv < HIST2_TV) v += colsum[((m - 1) * HIST2_TV + tv) * HIST2_SP + j];
        }
        v = warp_sum(v);
        if (lane == 0) hist[((int64_t)r * M + m) * d.D + slot] = v * inv;
    }
    for (int e = t; e < d.D; e += blockDim.x) hist[(int64_t)r * M * d.D + e] = (e == c.d0 + d.doff) ? 1. : 0.;
}

// K-hist, third generation, any state space (S = 97: 1.32 -> 0.87 ms per 48-fit launch against k_hist2; S = 165: 5.0 ->
// 2.2 ms and S = 355: 23.5 -> 6.8 ms against k_hist): no atomics, no S x S buffer.  A thread
// owns a column j and walks the rows grouped by the tuple of their tumour-clone totals ("classes"; order and class
// starts built on the host once per batch): within a class the per-fit factor is constant, so a class costs its
// sum_i a_i flipE[i][j] (one L2 load, coalesced along j, and one FMA per pair), one factor and one add per clone into
// colsum[m][total][j] (the thread's own column: no conflicts).  The histogram slot of (m, delta) is then the sum over
// the columns of colsum[m][tot_j[m] + delta][j], as in k_hist2.  ONE pass over the S^2 pairs.  (The first-generation kernel adds every pair to M shared-memory counters with 64-bit atomics: 256
// threads on at most D addresses per clone -- 23.5 ms per 48-fit launch at S = 355, 5.0 ms at S = 165.)
#define HIST4_T 384
__global__ void __launch_bounds__(HIST4_T, 2) k_hist4(RmxDims d, int g0, RmxRows rows, RmxTransSpec sp, int64_t sp_fit_stride,
                                                   const int8_t* __restrict__ tab, const uint16_t* __restrict__ ord_g,
                                                   const int* __restrict__ gstart_g, int TVN, int NGS,
                                                   const double* __restrict__ ahat_all, const double* __restrict__ bhat_all,
                                                   const double* __restrict__ fexp_all, double* __restrict__ hist_all,
                                                   int* __restrict__ dflags) {
    const int g = RMX_FIT();
    const int64_t NSo = (int64_t)g * d.N * d.S;
    const double *ahat = ahat_all + NSo, *bhat = bhat_all + NSo, *fexp = fexp_all + NSo;
    double* hist = hist_all + (int64_t)g * sp_fit_stride;
    sp.pd += (int64_t)g * sp_fit_stride;
    sp.wexp += (int64_t)g * sp_fit_stride;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red[32];
    const int S = d.S, M = d.M, D = d.D;
    const int r = blockIdx.x;
    const int n = rows.brk_rows[r];
    PairCtx c;
    pair_setup(c, smem_raw, d, rows, sp, tab, 0, n, ahat, bhat, fexp);      // a, g, wtab, tots (includes the barrier)
    size_t off = (sizeof(double) * (2 * (size_t)S + (size_t)M * D) + (size_t)S * M + 15) & ~(size_t)15;
    double* colsum = reinterpret_cast<double*>(smem_raw + off);              // [M-1][TVN][S]
    uint16_t* ord = reinterpret_cast<uint16_t*>(colsum + (size_t)(M - 1) * TVN * S);     // [M][S]
    int* gstart = reinterpret_cast<int*>(ord + (((size_t)M * S + 1) & ~(size_t)1));    // [NCLS + 2] class starts
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5, nw = blockDim.x >> 5;
    for (int e = t; e < M * S; e += blockDim.x) ord[e] = ord_g[e];
    for (int e = t; e < NGS; e += blockDim.x) gstart[e] = gstart_g[e];
    __syncthreads();
    const double w0 = c.kind == 2 ? c.wtab[c.d0 + d.doff] : 1.;
    // ord row 0: the states ordered by the tuple of their tumour-clone totals ("classes"); gstart row 0: the class
    // starts, gstart[0][TVN] = number of classes is stored by the host in the last used entry + 1 ... see api.cu.
    const uint16_t* co = ord;                 // [S] class-ordered states
    const int* cs = gstart;                   // cs[0] = number of classes NC, cs[1 .. NC + 1] = class starts
    const int NCLS = cs[0];
    for (int j = t; j < S; j += blockDim.x) {
        const double gj = c.g[j];
        int bj[RMX_MAXM];
#pragma unroll
        for (int m = 1; m < RMX_MAXM; ++m) bj[m] = (m < M) ? m * D + d.doff - (int)c.tots[j * M + m] : 0;
        const double* Ej = c.E + j;
        for (int m = 1; m < M; ++m)
            for (int tv = 0; tv < TVN; ++tv) colsum[((size_t)(m - 1) * TVN + tv) * S + j] = 0.;
        for (int cl = 0; cl < NCLS; ++cl) {
            // the per-fit factor depends on (i, j) through the totals only: it is the same for every row of a class
            double acc = 0.;
            const int k1 = cs[cl + 2];
#pragma unroll 4
            for (int k = cs[cl + 1]; k < k1; ++k) {
                const int i = co[k];
                acc = fma(c.a[i], c.kind == 0 ? 1. : __ldg(Ej + (int64_t)i * S), acc);
            }
            const int irep = co[cs[cl + 1]];
            double e = acc;
            if (c.kind == 2) {
                e *= w0;
#pragma unroll
                for (int m = 1; m < RMX_MAXM; ++m)
                    if (m < M) e *= c.wtab[bj[m] + (int)c.tots[irep * M + m]];
            }
            e *= gj;
#pragma unroll
            for (int m = 1; m < RMX_MAXM; ++m)
                if (m < M) colsum[((size_t)(m - 1) * TVN + (int)c.tots[irep * M + m]) * S + j] += e;
        }
    }
    __syncthreads();
    // Z: every pair once, from the first tumour clone's column sums in a fixed order
    double z = 0.;
    for (int j = t; j < S; j += blockDim.x)
        for (int tv = 0; tv < TVN; ++tv) z += colsum[(size_t)tv * S + j];
    z = block_sum_all(z, red);
    if (!(z > 0.) || isinf(z)) {
        if (t == 0) atomicOr(dflags, RMX_F_JOINT_NAN);
    }
    const double inv = 1. / z;
    for (int q = wid; q < (M - 1) * D; q += nw) {
        const int m = 1 + q / D, slot = q - (m - 1) * D;
        const int dd = slot - d.doff;
        double v = 0.;
        for (int j = lane; j < S; j += 32) {
            const int tv = (int)c.tots[j * M + m] + dd;
            if (tv >= 0 && tv < TVN) v += colsum[((size_t)(m - 1) * TVN + tv) * S + j];
        }
        v = warp_sum(v);
        if (lane == 0) hist[((int64_t)r * M + m) * D + slot] = v * inv;
    }
    for (int e = t; e < D; e += blockDim.x) hist[(int64_t)r * M * D + e] = (e == c.d0 + d.doff) ? 1. : 0.;
}

// update_p_breakpoint (bpmodel.pyx:964-983): one CTA per (fit, breakpoint), one thread per breakpoint
// state (B <= 64).  Each thread accumulates its log q in the reference's order (breakend rows ascending,
// clones, copy-number differences); thread 0 then does the two sequential sums of _exp_normalize
// (bpmodel.pyx:120-128) so the result does not depend on a reduction tree.
__global__ void __launch_bounds__(64) k_brk_update(RmxDims d, int g0, const int* __restrict__ krow_ptr,
                                                   const int* __restrict__ krow_idx, const int* __restrict__ brk_rows,
                                                   const int* __restrict__ borient, const int8_t* __restrict__ brk,
                                                   const double* __restrict__ hist_all, int64_t hist_fit_stride,
                                                   double pen, int tm, double* __restrict__ p_brk_all, int mode,
                                                   double* __restrict__ logq_all) {
    // mode 0: the whole update; mode 1: only this model's partial log q (its own breakend rows) to logq_all
    // (G,K,B); mode 2: log q taken from logq_all (the sum over the shards of a chromosome-sharded fit) and
    // normalised.  Modes 1 + 2 around an allreduce are mode 0 for a model that holds a subset of the chains.
    __shared__ double buf[64];
    __shared__ double stat;
    const int g = RMX_FIT();
    const int k = blockIdx.x, b = threadIdx.x;
    const double* hist = hist_all + (int64_t)g * hist_fit_stride;
    double* p = p_brk_all + ((int64_t)g * d.K + k) * d.B;
    const double mult = -pen;
    double acc = 0.;
    double* lq = logq_all ? logq_all + ((int64_t)g * d.K + k) * d.B : nullptr;
    if (b < d.B && mode == 2) {
        acc = lq[b];
        buf[b] = acc;
    } else if (b < d.B) {
        for (int e = krow_ptr[k]; e < krow_ptr[k + 1]; ++e) {
            const int r = krow_idx[e];
            const int o = borient[brk_rows[r]];
            for (int m = 0; m < d.M; ++m) {
                const double* h = hist + ((int64_t)r * d.M + m) * d.D;
                const int bs = brk[b * d.M + m];
                for (int slot = 0; slot < d.D; ++slot)
                    acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(mult, h[slot]), ftrans(slot - d.doff - o * bs, tm)));
            }
        }
        buf[b] = acc;
        if (mode == 1) lq[b] = acc;
    }
    if (mode == 1) return;
    __syncthreads();
    if (b == 0) {
        double mx = -INFINITY;
        for (int i = 0; i < d.B; ++i) if (buf[i] > mx) mx = buf[i];
        double ps = 0.;
        for (int i = 0; i < d.B; ++i) ps += exp(buf[i] - mx);
        stat = log(ps) + mx;
    }
    __syncthreads();
    const double e = b < d.B ? exp(acc - stat) : 0.;
    __syncthreads();
    if (b < d.B) buf[b] = e;
    __syncthreads();
    if (b == 0) {
        double z = 0.;
        for (int i = 0; i < d.B; ++i) z += buf[i];
        stat = z;
    }
    __syncthreads();
    if (b < d.B) p[b] = e / stat;
}

// exact log transition T_spec[n,i,j] for pair sums / materialisation / Viterbi
struct LogTCtx {
    const double* Tm;    // class Tmat (plain)
    const double* fc;    // class flipc (breakend)
    const double* pdr;   // pd of the breakend row
    const int8_t* tots;
    int kind;            // 0 zero, 1 plain, 2 breakend
    int d0;
    double pen;
};

__device__ __forceinline__ void logT_setup(LogTCtx& c, const RmxDims& d, const RmxRows& rows, const RmxTransSpec& sp,
                                           const int8_t* tots, int n) {
    int cls = rows.tclass[n], br = rows.brow[n];
    c.kind = 0;
    c.tots = tots;
    c.d0 = rows.d0[n];
    c.pen = sp.pen;
    c.Tm = c.fc = c.pdr = nullptr;
    if (cls >= 0) {
        if (br >= 0) {
            c.kind = 2;
            c.fc = sp.flipc + (int64_t)cls * d.S * d.S;
            c.pdr = sp.pd + (int64_t)br * d.M * d.D;
        } else {
            c.kind = 1;
            c.Tm = sp.Tmat + (int64_t)cls * d.S * d.S;
        }
    }
}

__device__ __forceinline__ double logT(const LogTCtx& c, const RmxDims& d, int i, int j) {
    if (c.kind == 1) return c.Tm[(int64_t)i * d.S + j];
    if (c.kind == 2) return brk_logT(c.pdr, c.tots, d.M, d.D, d.doff, c.d0, i, j, c.fc[(int64_t)i * d.S + j], c.pen);
    return 0.;
}

// ELBO term sum_ij joint[n,i,j] * (T_cached - T_used)[n,i,j] of the breakend rows from the histograms update_p_cn
// already made: on a breakend row both transitions are -pen * (sum_m pd[m][d_m(i,j)] + flipc[i,j]) with the same
// flipc, so the difference depends on the pair only through the clones' total-copy differences d_m, and
//   sum_ij joint * (T_c - T_u) = -pen * sum_m sum_d hist[m][d] * (pd_c[m][d] - pd_u[m][d]).
// 45 terms per row instead of two passes over S^2 pairs (k_pairsum).  One thread per (fit, breakend row).
__global__ void k_pairsum_hist(RmxDims d, int g0, int R, const int* __restrict__ brk_rows, const int* __restrict__ tclass,
                               const double* __restrict__ hist_all, const double* __restrict__ pd_c_all,
                               const double* __restrict__ pd_u_all, int64_t fit_stride, double pen,
                               double* __restrict__ rowsum_all, int64_t rowsum_fit_stride) {
    const int g = RMX_FIT();
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const int MD = d.M * d.D;
    const double* h = hist_all + (int64_t)g * fit_stride + (int64_t)r * MD;
    const double* pc = pd_c_all + (int64_t)g * fit_stride + (int64_t)r * MD;
    const double* pu = pd_u_all + (int64_t)g * fit_stride + (int64_t)r * MD;
    double acc = 0.;
    if (tclass[brk_rows[r]] >= 0)            // a telomere row's transitions are identically zero
        for (int e = 0; e < MD; ++e) acc += h[e] * (pc[e] - pu[e]);
    rowsum_all[(int64_t)g * rowsum_fit_stride + r] = -pen * acc;
}

// rowsum[n] = sum_ij joint[n,i,j] * T_spec[n,i,j]; joint from (joint_sp, messages), T from t_sp.
// rowlist != nullptr: only those rows (the breakend rows), rowsum indexed by position in the list, and the
// value is sum joint * (T_tsp - T_t2sp) -- the only rows where cached and used transitions differ.
__global__ void __launch_bounds__(256) k_pairsum(RmxDims d, int g0, RmxRows rows, RmxTransSpec jsp, RmxTransSpec tsp,
                                                 RmxTransSpec t2sp, int diff, int64_t sp_fit_stride,
                                                 const int* __restrict__ rowlist, const int8_t* __restrict__ tab,
                                                 int uniform, const double* __restrict__ ahat_all,
                                                 const double* __restrict__ bhat_all, const double* __restrict__ fexp_all,
                                                 double* __restrict__ rowsum_all, int64_t rowsum_fit_stride) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red[32];
    const int g = RMX_FIT();
    const int64_t NSo = (int64_t)g * d.N * d.S;
    const double *ahat = ahat_all + NSo, *bhat = bhat_all + NSo, *fexp = fexp_all + NSo;
    double* rowsum = rowsum_all + (int64_t)g * rowsum_fit_stride;
    jsp.pd += (int64_t)g * sp_fit_stride; jsp.wexp += (int64_t)g * sp_fit_stride;
    tsp.pd += (int64_t)g * sp_fit_stride; tsp.wexp += (int64_t)g * sp_fit_stride;
    t2sp.pd += (int64_t)g * sp_fit_stride; t2sp.wexp += (int64_t)g * sp_fit_stride;
    const int n = rowlist ? rowlist[blockIdx.x] : blockIdx.x;
    const int orow = blockIdx.x;
    LogTCtx lt, lt2;
    // T of the row under t_sp; rows whose T is identically zero contribute nothing
    if (rows.tclass[n] < 0) {
        if (threadIdx.x == 0) rowsum[orow] = 0.;
        return;
    }
    PairCtx c;
    pair_setup(c, smem_raw, d, rows, jsp, tab, uniform, n, ahat, bhat, fexp);
    logT_setup(lt, d, rows, tsp, c.tots, n);
    logT_setup(lt2, d, rows, t2sp, c.tots, n);
    const int S = d.S;
    const int64_t SS = (int64_t)S * S;
    double z = 0., num = 0.;
    for (int64_t idx = threadIdx.x; idx < SS; idx += blockDim.x) {
        int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
        double w = pair_w(c, d, i, j);
        z += w;
        double tv = logT(lt, d, i, j);
        if (diff) tv -= logT(lt2, d, i, j);
        num += w * tv;
    }
    z = block_sum_all(z, red);
    num = block_sum_all(num, red);
    if (threadIdx.x == 0) rowsum[orow] = num / z;
}

// ---------------------------------------------------------------------------------------------
// Viterbi per chain (bpmodel.pyx:1296-1333): first-max tie breaking (strict >, lowest index wins).
// Telomere rows are all-zero so the lattice of a chain only differs from the reference's by an
// additive constant; the traceback at a telomere is argmax of the lattice row itself.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void logT_from(LogTCtx& c, const RmxDims& d, const RmxTransSpec& sp, const int8_t* tots,
                                          int cls, int br, int d0) {
    // logT_setup with the row's class / breakend row / normal-copy difference already loaded
    c.kind = 0;
    c.tots = tots;
    c.d0 = d0;
    c.pen = sp.pen;
    c.Tm = c.fc = c.pdr = nullptr;
    if (cls >= 0) {
        if (br >= 0) {
            c.kind = 2;
            c.fc = sp.flipc + (int64_t)cls * d.S * d.S;
            c.pdr = sp.pd + (int64_t)br * d.M * d.D;
        } else {
            c.kind = 1;
            c.Tm = sp.Tmat + (int64_t)cls * d.S * d.S;
        }
    }
}

// One column of a lattice row on the general path (telomere rows, breakend rows, classes that are not staged).
__device__ __noinline__ double vit_column_general(RmxDims d, RmxTransSpec sp, const int8_t* tots, int cls, int br, int dd0,
                                                  const double* pc, int jj) {
    LogTCtx lt;
    logT_from(lt, d, sp, tots, cls, br, dd0);
    double vmax = -INFINITY;
    if (lt.kind == 0) {
        for (int i = 0; i < d.S; ++i)
            if (pc[i] > vmax) vmax = pc[i];
    } else {
        for (int i = 0; i < d.S; ++i) {
            double v = __dadd_rn(pc[i], logT(lt, d, i, jj));
            if (v > vmax) vmax = v;
        }
    }
    return vmax;
}

// One CTA decodes the chains of bin blockIdx.x (bins of about equal total length, RmxChains::bin_*) of `fpc` groups
// of FR fits: the groups share the class transition matrices staged in shared memory (classes 0 .. staged-1), each
// group has blockDim.x / fpc threads (thread per state) and its own named barrier, so the groups advance
// independently.  The staged matrices are stored transposed, one row per destination state with a row stride
// SP = roundup(S, 8) + 2 doubles (SP / 2 odd: the 16-byte loads of a quarter warp fall in distinct bank groups); a
// thread walks its row with 128-bit loads and uses each matrix value for the lattice rows of its FR fits (the matrix
// stream is the shared-memory traffic of this kernel: FR = 2 halves it), eight pairs per fit and iteration into four
// independent running maxima.  Pad entries are -inf in the lattice row and 0 in the matrix, so they never win.  The
// lattice rows are double buffered: one barrier per segment.  Row descriptors and the next frame values are loaded
// one segment ahead; the traceback of fit f is done by warp f of the group without block barriers.
template <int FR, int MAXT>
__global__ void __launch_bounds__(MAXT) k_viterbi(RmxDims d, int g0, int gn, RmxChains ch, RmxRows rows, RmxTransSpec sp,
                                                  int64_t sp_fit_stride, const int8_t* __restrict__ tab,
                                                  const double* __restrict__ F_all, double* __restrict__ lattice_all,
                                                  int* __restrict__ path_all, int staged, int fpc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = d.S, M = d.M;
    const int SL = (S + 7) & ~7, SP = SL + 2;
    const int tpf = blockDim.x / fpc;                     // threads per group, a multiple of 32
    const int grp = threadIdx.x / tpf, t = threadIdx.x - grp * tpf;
    double* prev_all = reinterpret_cast<double*>(smem_raw);                      // [fpc][FR][2][SP]
    double* Tsm = prev_all + (size_t)fpc * FR * 2 * SP;                         // [staged][S][SP], transposed
    int8_t* tots = reinterpret_cast<int8_t*>(Tsm + (size_t)staged * S * SP);    // [S*M]
    for (int e = threadIdx.x; e < S * M; e += blockDim.x) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(tab[(s * M + m) * 2] + tab[(s * M + m) * 2 + 1]);
    }
    for (int e = threadIdx.x; e < staged * S * S; e += blockDim.x) {
        const int c = e / (S * S), r = e - c * S * S, i = r / S, j = r - i * S;
        Tsm[((size_t)c * S + j) * SP + i] = sp.Tmat[e];
    }
    for (int e = threadIdx.x; e < staged * S * (SP - S); e += blockDim.x) {
        const int row = e / (SP - S), i = S + (e - row * (SP - S));
        Tsm[(size_t)row * SP + i] = 0.;
    }
    for (int e = threadIdx.x; e < fpc * FR * 2 * SP; e += blockDim.x) prev_all[e] = -INFINITY;
    __syncthreads();
    const int lf0 = ((int)blockIdx.y * fpc + grp) * FR;   // first fit of the group within the launched range
    if (lf0 >= gn) return;
    // a group at the end of the range may hold fewer than FR fits: the missing ones repeat the last fit's inputs and
    // store nothing
    const int64_t NS = (int64_t)d.N * d.S;
    const double* F[FR];
    double* lattice[FR];
    bool live[FR];
#pragma unroll
    for (int f = 0; f < FR; ++f) {
        live[f] = lf0 + f < gn;
        const int lf = live[f] ? lf0 + f : gn - 1;
        F[f] = F_all + (int64_t)(g0 + lf) * NS;
        lattice[f] = lattice_all + (int64_t)lf * NS;      // scratch holds the launched fits only
    }
    double* prev = prev_all + (size_t)grp * FR * 2 * SP;  // fit f, buffer c at prev + (f * 2 + c) * SP
    const int c_lo = ch.bin_ptr[blockIdx.x], c_hi = ch.bin_ptr[blockIdx.x + 1];
#define VIT_BAR() asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(tpf) : "memory")
    int cur = 0;
    for (int ci = c_lo; ci < c_hi; ++ci) {
        const int chain = ch.bin_idx[ci];
        const int s0 = ch.start[chain], s1 = ch.end[chain];
#pragma unroll
        for (int f = 0; f < FR; ++f)
            for (int jj = t; jj < S; jj += tpf) {
                double v = F[f][(int64_t)s0 * S + jj];
                if (live[f]) lattice[f][(int64_t)s0 * S + jj] = v;
                prev[(f * 2 + cur) * SP + jj] = v;
            }
        VIT_BAR();
        // row n-1 -> n; descriptors of the next row and the next frame values (first column) one segment ahead
        int cls = s0 < s1 ? rows.tclass[s0] : -1, br = s0 < s1 ? rows.brow[s0] : -1, dd0 = s0 < s1 ? rows.d0[s0] : 0;
        int64_t off = (int64_t)(s0 + 1) * S + t;                // this thread's first column of row n
        double fcur[FR];
#pragma unroll
        for (int f = 0; f < FR; ++f) fcur[f] = (s0 < s1 && t < S) ? F[f][off] : 0.;
        for (int n = s0 + 1; n <= s1; ++n, off += S) {
            const bool fast = cls >= 0 && br < 0 && cls < staged;
            const int cls0 = cls, br0 = br, dd00 = dd0;
            double fn0[FR];
#pragma unroll
            for (int f = 0; f < FR; ++f) fn0[f] = fcur[f];
            if (n < s1) {
                cls = rows.tclass[n]; br = rows.brow[n]; dd0 = rows.d0[n];
                if (t < S) {
#pragma unroll
                    for (int f = 0; f < FR; ++f) fcur[f] = F[f][off + S];
                }
            }
            if (fast) {
                if (t < S) {            // staged classes have S <= tpf: one column per thread
                    // the value of a maximum does not depend on the order (ties only matter in the traceback)
                    const double2* pv = reinterpret_cast<const double2*>(prev + cur * SP);
                    const double2* tv = reinterpret_cast<const double2*>(Tsm + ((size_t)cls0 * S + t) * SP);
                    const double2* pe = pv + SL / 2;
                    double mx[FR][4];
#pragma unroll
                    for (int f = 0; f < FR; ++f)
#pragma unroll
                        for (int k = 0; k < 4; ++k) mx[f][k] = -INFINITY;
#pragma unroll 1
                    for (; pv < pe; pv += 4, tv += 4) {
                        const double2 t0 = tv[0], t1 = tv[1], t2 = tv[2], t3 = tv[3];
#pragma unroll
                        for (int f = 0; f < FR; ++f) {
                            const double2* pf = pv + f * SP;     // fit f's buffers start 2 * SP doubles further
                            const double2 p0 = pf[0], p1 = pf[1], p2 = pf[2], p3 = pf[3];
                            double v;
                            v = __dadd_rn(p0.x, t0.x); if (v > mx[f][0]) mx[f][0] = v;
                            v = __dadd_rn(p0.y, t0.y); if (v > mx[f][1]) mx[f][1] = v;
                            v = __dadd_rn(p1.x, t1.x); if (v > mx[f][2]) mx[f][2] = v;
                            v = __dadd_rn(p1.y, t1.y); if (v > mx[f][3]) mx[f][3] = v;
                            v = __dadd_rn(p2.x, t2.x); if (v > mx[f][0]) mx[f][0] = v;
                            v = __dadd_rn(p2.y, t2.y); if (v > mx[f][1]) mx[f][1] = v;
                            v = __dadd_rn(p3.x, t3.x); if (v > mx[f][2]) mx[f][2] = v;
                            v = __dadd_rn(p3.y, t3.y); if (v > mx[f][3]) mx[f][3] = v;
                        }
                    }
#pragma unroll
                    for (int f = 0; f < FR; ++f) {
                        if (mx[f][1] > mx[f][0]) mx[f][0] = mx[f][1];
                        if (mx[f][3] > mx[f][2]) mx[f][2] = mx[f][3];
                        if (mx[f][2] > mx[f][0]) mx[f][0] = mx[f][2];
                        const double o = __dadd_rn(mx[f][0], fn0[f]);
                        prev[(f * 2 + (cur ^ 1)) * SP + t] = o;
                        if (live[f]) lattice[f][off] = o;
                    }
                }
            } else {
#pragma unroll
                for (int f = 0; f < FR; ++f) {
                    RmxTransSpec spf = sp;
                    const int64_t fo = (int64_t)(g0 + (live[f] ? lf0 + f : gn - 1)) * sp_fit_stride;
                    spf.pd += fo;
                    spf.wexp += fo;
                    for (int jj = t; jj < S; jj += tpf) {
                        const double fn = jj == t ? fn0[f] : F[f][(int64_t)n * S + jj];
                        const double o = __dadd_rn(
                            vit_column_general(d, spf, tots, cls0, br0, dd00, prev + (f * 2 + cur) * SP, jj), fn);
                        prev[(f * 2 + (cur ^ 1)) * SP + jj] = o;
                        if (live[f]) lattice[f][(int64_t)n * S + jj] = o;
                    }
                }
            }
            VIT_BAR();
            cur ^= 1;
        }
    }
#undef VIT_BAR
    // traceback of fit f by warp f of the group: no barriers on the sequential path; the lattice row and the row
    // descriptors of the next step are loaded before the current step's reduction
    const int wf = t >> 5;
    if (wf >= FR || lf0 + wf >= gn) return;
    const int lane = t & 31;
    const double* lat = lattice_all + (int64_t)(lf0 + wf) * NS;
    int* path = path_all + (int64_t)(g0 + lf0 + wf) * d.N;
    sp.pd += (int64_t)(g0 + lf0 + wf) * sp_fit_stride;
    sp.wexp += (int64_t)(g0 + lf0 + wf) * sp_fit_stride;
    for (int ci = c_lo; ci < c_hi; ++ci) {
        const int chain = ch.bin_idx[ci];
        const int s0 = ch.start[chain], s1 = ch.end[chain];
        constexpr int KP = 4;                 // rows of up to 128 states are prefetched in registers
        const bool pre = S <= 32 * KP;
        int nxt = 0;
        double lv[KP] = {0., 0., 0., 0.};
        int pc = -1, pb = -1, pd0 = 0;
        if (pre) {
#pragma unroll
            for (int k = 0; k < KP; ++k) lv[k] = (lane + 32 * k < S) ? lat[(int64_t)s1 * S + lane + 32 * k] : 0.;
        }
        for (int n = s1; n >= s0; --n) {
            // on entry: lv = lattice row n, (pc, pb, pd0) = descriptors of row n (transition n -> n+1, n < s1)
            LogTCtx lt;
            const double* Ts = nullptr;
            lt.kind = 0;
            double cur_row[KP];
#pragma unroll
            for (int k = 0; k < KP; ++k) cur_row[k] = lv[k];
            if (n < s1) {
                logT_from(lt, d, sp, tots, pc, pb, pd0);
                if (lt.kind == 1 && pc < staged) Ts = Tsm + ((size_t)pc * S + nxt) * SP;
            }
            int nc = -1, nb = -1, nd0 = 0;
            if (n > s0) {       // the next step's inputs, in flight during this step's reduction
                if (pre) {
#pragma unroll
                    for (int k = 0; k < KP; ++k)
                        lv[k] = (lane + 32 * k < S) ? lat[(int64_t)(n - 1) * S + lane + 32 * k] : 0.;
                }
                nc = rows.tclass[n - 1]; nb = rows.brow[n - 1]; nd0 = rows.d0[n - 1];
            }
            double bv = -INFINITY;
            int bi = 0x7fffffff;
            if (pre) {
#pragma unroll
                for (int k = 0; k < KP; ++k) {
                    const int i = lane + 32 * k;
                    if (i < S) {
                        double v = cur_row[k];
                        if (n < s1) {
                            if (Ts) v = __dadd_rn(v, Ts[i]);
                            else v = __dadd_rn(v, logT(lt, d, i, nxt));
                        }
                        if (v > bv) { bv = v; bi = i; }   // ascending i per lane: first max kept
                    }
                }
            } else {
                for (int i = lane; i < S; i += 32) {
                    double v = lat[(int64_t)n * S + i];
                    if (n < s1) {
                        if (Ts) v = __dadd_rn(v, Ts[i]);
                        else v = __dadd_rn(v, logT(lt, d, i, nxt));
                    }
                    if (v > bv) { bv = v; bi = i; }
                }
            }
            // warp argmax with lowest-index tie break
            for (int o = 16; o > 0; o >>= 1) {
                double ov = __shfl_xor_sync(0xffffffffu, bv, o);
                int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (bi == 0x7fffffff) bi = 0;   // all -inf / NaN: _argmax returns 0
            if (lane == 0) path[n] = bi;
            nxt = bi;
            pc = nc; pb = nb; pd0 = nd0;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// materialisation of the reference's (N-1,S,S) attributes, for inspection / tests only
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_materialize_T(RmxDims d, RmxRows rows, RmxTransSpec sp,
                                                       const int8_t* __restrict__ tab, int64_t row0,
                                                       double* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = d.S, M = d.M;
    int8_t* tots = reinterpret_cast<int8_t*>(smem_raw);
    int n = (int)(row0 + blockIdx.x);
    for (int e = threadIdx.x; e < S * M; e += blockDim.x) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(tab[(s * M + m) * 2] + tab[(s * M + m) * 2 + 1]);
    }
    __syncthreads();
    LogTCtx lt;
    logT_setup(lt, d, rows, sp, tots, n);
    const int64_t SS = (int64_t)S * S;
    for (int64_t idx = threadIdx.x; idx < SS; idx += blockDim.x) {
        int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
        out[(int64_t)blockIdx.x * SS + idx] = logT(lt, d, i, j);
    }
}

__global__ void __launch_bounds__(256) k_materialize_joint(RmxDims d, RmxRows rows, RmxTransSpec sp,
                                                           const int8_t* __restrict__ tab, int uniform,
                                                           const double* __restrict__ ahat,
                                                           const double* __restrict__ bhat,
                                                           const double* __restrict__ fexp, int64_t row0,
                                                           double* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red[32];
    int n = (int)(row0 + blockIdx.x);
    PairCtx c;
    pair_setup(c, smem_raw, d, rows, sp, tab, uniform, n, ahat, bhat, fexp);
    const int S = d.S;
    const int64_t SS = (int64_t)S * S;
    double z = 0.;
    for (int64_t idx = threadIdx.x; idx < SS; idx += blockDim.x) {
        int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
        z += pair_w(c, d, i, j);
    }
    z = block_sum_all(z, red);
    for (int64_t idx = threadIdx.x; idx < SS; idx += blockDim.x) {
        int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
        out[(int64_t)blockIdx.x * SS + idx] = pair_w(c, d, i, j) / z;
    }
}

// ---------------------------------------------------------------------------------------------
// Stateless module functions on caller-supplied (N-1,S,S) log_transmat, log domain, reference
// summation order (bpmodel.pyx:1213-1246, :1249-1292, :1296-1333).  grid = 2 (forward, backward).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_sum_product_log(int64_t N, int S, const double* __restrict__ F,
                                                          const double* __restrict__ Tm, double* __restrict__ alphas,
                                                          double* __restrict__ betas) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* prev = reinterpret_cast<double*>(smem_raw);
    const int64_t SS = (int64_t)S * S;
    const int t = threadIdx.x;
    if (blockIdx.x == 0) {
        for (int j = t; j < S; j += blockDim.x) { double v = F[j]; alphas[j] = v; prev[j] = v; }
        __syncthreads();
        for (int64_t n = 1; n < N; ++n) {
            const double* Tn = Tm + (n - 1) * SS;
            double out[4]; int cnt = 0;
            for (int j = t; j < S; j += blockDim.x) {
                double vmax = -INFINITY;
                for (int i = 0; i < S; ++i) { double v = prev[i] + Tn[(int64_t)i * S + j]; if (v > vmax) vmax = v; }
                double ps = 0.;
                for (int i = 0; i < S; ++i) ps += exp(prev[i] + Tn[(int64_t)i * S + j] - vmax);
                out[cnt++] = log(ps) + vmax + F[n * S + j];
            }
            __syncthreads();
            cnt = 0;
            for (int j = t; j < S; j += blockDim.x) { prev[j] = out[cnt]; alphas[n * S + j] = out[cnt]; ++cnt; }
            __syncthreads();
        }
    } else {
        for (int j = t; j < S; j += blockDim.x) { betas[(N - 1) * S + j] = 0.; prev[j] = F[(N - 1) * S + j]; }
        __syncthreads();
        for (int64_t n = N - 2; n >= 0; --n) {
            const double* Tn = Tm + n * SS;
            double out[4]; int cnt = 0;
            for (int i = t; i < S; i += blockDim.x) {
                double vmax = -INFINITY;
                for (int j = 0; j < S; ++j) { double v = Tn[(int64_t)i * S + j] + prev[j]; if (v > vmax) vmax = v; }
                double ps = 0.;
                for (int j = 0; j < S; ++j) ps += exp(Tn[(int64_t)i * S + j] + prev[j] - vmax);
                out[cnt++] = log(ps) + vmax;
            }
            __syncthreads();
            cnt = 0;
            for (int i = t; i < S; i += blockDim.x) { betas[n * S + i] = out[cnt]; prev[i] = F[n * S + i] + out[cnt]; ++cnt; }
            __syncthreads();
        }
    }
}

__device__ __forceinline__ double logsumpair(double x, double y) {
    double vmax = fmax(x, y);
    return log(exp(x - vmax) + exp(y - vmax)) + vmax;
}

__global__ void __launch_bounds__(1024) k_sum_product_2p(int64_t N, int S, const double* __restrict__ F,
                                                         double* __restrict__ alphas, double* __restrict__ betas,
                                                         double log_e0, double log_e1) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red[32];
    double* prev = reinterpret_cast<double*>(smem_raw);
    const int t = threadIdx.x;
    const bool fwd = blockIdx.x == 0;
    if (fwd) { for (int j = t; j < S; j += blockDim.x) { double v = F[j]; alphas[j] = v; prev[j] = v; } }
    else { for (int j = t; j < S; j += blockDim.x) { betas[(N - 1) * S + j] = 0.; prev[j] = F[(N - 1) * S + j]; } }
    __syncthreads();
    for (int64_t step = 1; step < N; ++step) {
        int64_t n = fwd ? step : N - 1 - step;
        double m = -INFINITY;
        for (int j = t; j < S; j += blockDim.x) m = fmax(m, prev[j]);
        m = warp_max(m);
        __syncthreads();
        if ((t & 31) == 0) red[t >> 5] = m;
        __syncthreads();
        m = -INFINITY;
        for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) m = fmax(m, red[w]);
        double ps = 0.;
        for (int j = t; j < S; j += blockDim.x) ps += exp(prev[j] - m);
        ps = block_sum_all(ps, red);
        double notrans = log_e0 + log(ps) + m;
        double out[4]; int cnt = 0;
        for (int j = t; j < S; j += blockDim.x) out[cnt++] = logsumpair(notrans, log_e1 + prev[j]);
        __syncthreads();
        cnt = 0;
        for (int j = t; j < S; j += blockDim.x) {
            if (fwd) { double v = out[cnt] + F[n * S + j]; alphas[n * S + j] = v; prev[j] = v; }
            else { betas[n * S + j] = out[cnt]; prev[j] = F[n * S + j] + out[cnt]; }
            ++cnt;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(1024) k_max_product(int64_t N, int S, const double* __restrict__ F,
                                                      const double* __restrict__ Tm, double* __restrict__ lattice,
                                                      int64_t* __restrict__ path, double* __restrict__ logprob) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double rv[32];
    __shared__ int ri[32];
    __shared__ int s_next;
    double* prev = reinterpret_cast<double*>(smem_raw);
    const int64_t SS = (int64_t)S * S;
    const int t = threadIdx.x;
    for (int j = t; j < S; j += blockDim.x) { double v = F[j]; lattice[j] = v; prev[j] = v; }
    __syncthreads();
    for (int64_t n = 1; n < N; ++n) {
        const double* Tn = Tm + (n - 1) * SS;
        double out[4]; int cnt = 0;
        for (int j = t; j < S; j += blockDim.x) {
            double vmax = -INFINITY;
            for (int i = 0; i < S; ++i) { double v = __dadd_rn(prev[i], Tn[(int64_t)i * S + j]); if (v > vmax) vmax = v; }
            out[cnt++] = __dadd_rn(vmax, F[n * S + j]);
        }
        __syncthreads();
        cnt = 0;
        for (int j = t; j < S; j += blockDim.x) { prev[j] = out[cnt]; lattice[n * S + j] = out[cnt]; ++cnt; }
        __syncthreads();
    }
    for (int64_t n = N - 1; n >= 0; --n) {
        int nxt = (n < N - 1) ? s_next : 0;
        double bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int i = t; i < S; i += blockDim.x) {
            double v = lattice[n * S + i];
            if (n < N - 1) v = __dadd_rn(v, Tm[n * SS + (int64_t)i * S + nxt]);
            if (v > bv) { bv = v; bi = i; }
        }
        for (int o = 16; o > 0; o >>= 1) {
            double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        __syncthreads();
        if ((t & 31) == 0) { rv[t >> 5] = bv; ri[t >> 5] = bi; }
        __syncthreads();
        if (t == 0) {
            int nw = (blockDim.x + 31) >> 5;
            double fv = rv[0];
            int fi = ri[0];
            for (int w = 1; w < nw; ++w)
                if (rv[w] > fv || (rv[w] == fv && ri[w] < fi)) { fv = rv[w]; fi = ri[w]; }
            if (fi == 0x7fffffff) fi = 0;
            path[n] = fi;
            s_next = fi;
            if (n == N - 1) *logprob = lattice[n * S + fi];
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
static inline int cdiv64(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

void rmx_launch_build_class(cudaStream_t st, const RmxTables& T, const RmxDims& d, int C, const int8_t* class_nrm,
                            double pen, int tm, double* Tmat, double* Emat, double* flipc, double* flipE) {
    if (C <= 0) return;
    dim3 grid(cdiv64((int64_t)d.S * d.S, 256), C);
    k_build_class<<<grid, 256, 0, st>>>(T.tab, d.S, d.M, class_nrm, pen, tm, Tmat, Emat, flipc, flipE);
}

void rmx_launch_prep_brk(cudaStream_t st, const RmxDims& d, int g0, int gn, int R, const int* brk_rows, const int* bidx,
                         const int* borient, const int8_t* brk, const double* p_brk, double pen, int tm,
                         int64_t fit_stride, double* pd, double* wexp) {
    if (R <= 0 || gn <= 0) return;
    k_prep_brk<<<dim3(R, gn), 64, 0, st>>>(d, g0, brk_rows, bidx, borient, brk, p_brk, pen, tm, fit_stride, pd, wexp);
}

template <int RR, int CW, int GT, int QT, bool REG>
static int launch_chain_cfg(cudaStream_t st, ChainArgs& A, int Q, int rr) {
    const RmxDims& d = A.d;
    const int NJ = (d.S + CW - 1) / CW;
    A.Q = Q; A.NJ = NJ; A.rr = rr;
    A.ntiles = (A.gn + GT - 1) / GT;
    A.fast_brk = 1;
    const int Sp = Q * rr, SJ = (d.S + 31) & ~31, NCH = SJ >> 5;
    const int threads = ((Q * NJ + 31) / 32) * 32;
    if (threads > (REG ? 256 : 512) || RMX_EPI * threads < GT * SJ) return 2;
    const bool fast = REG && d.M == 3 && d.D * d.D <= 256;
    size_t smem = sizeof(double) * ((size_t)2 * GT * Sp + (size_t)Q * GT * CW * NJ + (size_t)GT * d.M * d.D + GT * NCH +
                                    (fast ? GT * 256 + RMX_FETAB : 0)) +
                  sizeof(int) * 2 * GT * NCH + (fast ? sizeof(unsigned short) * RR * CW * threads : 0) +
                  (size_t)d.S * d.M + 16;
    auto kern = k_chain<RR, CW, GT, QT, REG>;
    if (smem > 48 * 1024) {
        if (smem > 227 * 1024) return 2;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 1;
    }
    kern<<<A.ch.NC * 2 * A.ntiles, threads, smem, st>>>(A);
    return 0;
}

template <int GT>
static int launch_chain_gt(cudaStream_t st, ChainArgs& A) {
    const int S = A.d.S;
    if (S <= 16) return launch_chain_cfg<4, 4, GT, 4, true>(st, A, 4, 4);
    if (S <= 64) return launch_chain_cfg<8, 4, GT, 8, true>(st, A, 8, 8);
    if (S <= 104) return launch_chain_cfg<13, 4, GT, 8, true>(st, A, 8, 13);
    if (S <= 128) return launch_chain_cfg<16, 4, GT, 8, true>(st, A, 8, 16);
    // large state spaces: E streamed from L2 every step
    const int NJ = (S + 3) / 4;
    int Q = 512 / NJ;
    if (Q > 8) Q = 8;
    if (Q < 1) return 2;
    const int rr = (S + Q - 1) / Q;
    return launch_chain_cfg<1, 4, GT, 0, false>(st, A, Q, rr);
}

// returns 0 on success, 1 if the shared-memory configuration failed, 2 if the state space is unsupported
int rmx_launch_chain(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const RmxRows& rows,
                     const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp, const double* mF,
                     double* ahat, double* bhat, double* chain_logz, int* dflags) {
    if (ch.NC <= 0 || gn <= 0) return 0;
    ChainArgs A;
    A.d = d; A.ch = ch; A.rows = rows; A.sp = sp; A.sp_fit_stride = sp_fit_stride; A.tab = T.tab;
    A.g0 = g0; A.gn = gn; A.fexp = fexp; A.mF = mF; A.ahat = ahat; A.bhat = bhat; A.chain_logz = chain_logz;
    A.dflags = dflags;
    const int SJ = (d.S + 31) & ~31;
    if (gn >= 3 && SJ <= 256) return launch_chain_gt<4>(st, A);
    if (gn >= 2 && SJ <= 512) return launch_chain_gt<2>(st, A);
    return launch_chain_gt<1>(st, A);
}

void rmx_launch_post(cudaStream_t st, const RmxDims& d, int g0, int gn, const double* ahat, const double* bhat,
                     double* post, int* dflags) {
    if (d.N <= 0 || gn <= 0) return;
    k_post<<<dim3(cdiv64((int64_t)d.N * 32, 256), gn), 256, 0, st>>>(d, g0, ahat, bhat, post, dflags);
}

static size_t pair_smem(const RmxDims& d) {
    return sizeof(double) * (2 * (size_t)d.S + (size_t)d.M * d.D) + (size_t)d.S * d.M + 16;
}

int rmx_hist2_ok(const RmxDims& d) { return d.S <= 128 && d.cn_max + 2 <= HIST2_TV && d.M >= 2; }
int rmx_hist2_gstart_count(const RmxDims& d) { return d.M * (HIST2_TV + 1); }
void rmx_launch_hist_prep(cudaStream_t st, const RmxTables& T, const RmxDims& d, uint8_t* hord, int* hgstart) {
    k_hist_prep<<<1, 256, 0, st>>>(T.tab, d.S, d.M, hord, hgstart);
}

void rmx_launch_hist(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, int R, const RmxRows& rows,
                     const RmxTransSpec& sp, int64_t sp_fit_stride, int uniform, const double* ahat, const double* bhat,
                     const double* fexp, double* hist, int* dflags, const uint8_t* hord, const int* hgstart,
                     const uint16_t* hord16, const int* hgs16, int TVN, int NGS) {
    if (R <= 0 || gn <= 0) return;
    // fitted models, any state space: class-wise column sums (k_hist4).  RMX_HIST2 / RMX_HIST1 force the older kernels
    // (A/B measurements only).
    static const bool no_hist4 = getenv("RMX_HIST1") != nullptr || getenv("RMX_HIST2") != nullptr;
    if (!uniform && !no_hist4 && hord16 != nullptr && TVN > 0 && d.M >= 2) {
        const size_t base4 = (pair_smem(d) + 15) & ~(size_t)15;
        const size_t smem4 = base4 + sizeof(double) * (size_t)(d.M - 1) * TVN * d.S +
                             sizeof(uint16_t) * (((size_t)d.M * d.S + 1) & ~(size_t)1) + sizeof(int) * (size_t)NGS + 16;
        if (smem4 <= 200 * 1024) {
            static RmxPerDeviceOnce attr4_once;
            if (attr4_once.need()) cudaFuncSetAttribute(k_hist4, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            k_hist4<<<dim3(R, gn), std::min(HIST4_T, ((d.S + 31) / 32) * 32), smem4, st>>>(
                d, g0, rows, sp, sp_fit_stride, T.tab, hord16, hgs16, TVN, NGS, ahat, bhat, fexp, hist, dflags);
            return;
        }
    }
    // S <= 128 and totals below HIST2_TV: the pair weights staged in shared memory (constructor state, A/B)
    static const bool no_hist2 = getenv("RMX_HIST1") != nullptr;
    const size_t base = (pair_smem(d) + 15) & ~(size_t)15;
    const size_t smem2 = base + sizeof(double) * ((size_t)d.S * d.S + (size_t)(d.M - 1) * HIST2_TV * HIST2_SP) + 16;
    if (!no_hist2 && hord != nullptr && rmx_hist2_ok(d) && smem2 <= 200 * 1024) {
        static RmxPerDeviceOnce attr_once;
        if (attr_once.need()) {
            cudaFuncSetAttribute(k_hist2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        }
        k_hist2<<<dim3(R, gn), 256, smem2, st>>>(d, g0, rows, sp, sp_fit_stride, T.tab, hord, hgstart, uniform, ahat, bhat,
                                                 fexp, hist, dflags);
        return;
    }
    k_hist<<<dim3(R, gn), 256, pair_smem(d), st>>>(d, g0, rows, sp, sp_fit_stride, T.tab, uniform, ahat, bhat, fexp, hist,
                                                  dflags);
}

void rmx_launch_brk_update(cudaStream_t st, const RmxDims& d, int g0, int gn, const int* krow_ptr, const int* krow_idx,
                           const int* brk_rows, const int* borient, const int8_t* brk, const double* hist,
                           int64_t hist_fit_stride, double pen, int tm, double* p_brk, int mode, double* logq) {
    if (d.K <= 0 || gn <= 0) return;
    k_brk_update<<<dim3(d.K, gn), 64, 0, st>>>(d, g0, krow_ptr, krow_idx, brk_rows, borient, brk, hist, hist_fit_stride,
                                               pen, tm, p_brk, mode, logq);
}

void rmx_launch_pairsum(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, const RmxRows& rows,
                        const RmxTransSpec& joint_sp, const RmxTransSpec& t_sp, const RmxTransSpec& t2_sp, int diff,
                        int64_t sp_fit_stride, const int* rowlist, int nrows, int uniform, const double* ahat,
                        const double* bhat, const double* fexp, double* rowsum, int64_t rowsum_fit_stride) {
    if (nrows <= 0 || gn <= 0) return;
    k_pairsum<<<dim3(nrows, gn), 256, pair_smem(d), st>>>(d, g0, rows, joint_sp, t_sp, t2_sp, diff, sp_fit_stride, rowlist,
                                                         T.tab, uniform, ahat, bhat, fexp, rowsum, rowsum_fit_stride);
}

void rmx_launch_pairsum_hist(cudaStream_t st, const RmxDims& d, int g0, int gn, int R, const RmxRows& rows,
                             const double* hist, const RmxTransSpec& cached, const RmxTransSpec& used,
                             int64_t sp_fit_stride, double* rowsum, int64_t rowsum_fit_stride) {
    if (R <= 0 || gn <= 0) return;
    k_pairsum_hist<<<dim3((R + 127) / 128, gn), 128, 0, st>>>(d, g0, R, rows.brk_rows, rows.tclass, hist, cached.pd, used.pd,
                                                             sp_fit_stride, cached.pen, rowsum, rowsum_fit_stride);
}

void rmx_launch_viterbi(cudaStream_t st, const RmxTables& T, const RmxDims& d, int g0, int gn, const RmxChains& ch,
                        const RmxRows& rows, const RmxTransSpec& sp, int64_t sp_fit_stride, const double* F,
                        double* lattice, int* path, int nclass) {
    if (ch.NB <= 0 || gn <= 0) return;
    int tpf = ((d.S + 31) / 32) * 32;
    if (tpf > 1024) tpf = 1024;
    // class matrices of the plain rows in shared memory, as many as fit (class ids are dense from 0)
    int staged = 0;
    const size_t SP = (size_t)((d.S + 7) & ~7) + 2;       // padded row of a transposed class matrix (see k_viterbi)
    while (staged < nclass && sizeof(double) * (size_t)(staged + 1) * d.S * SP <= 176 * 1024) ++staged;
    // FR = 2 (two fits per thread: the matrix stream is shared by the two) measures the same as FR = 1 at config 2 and
    // stays an A/B knob; it needs the traceback warps (warp f of a group traces fit f)
    int fr = 1;
    const int maxt = tpf <= 512 ? 512 : 1024;   // 512-thread CTAs compile to 128 registers (no spills)
    {
        static const char* env2 = getenv("RMX_VITERBI_FR");
        if (env2 && atoi(env2) == 2 && staged > 0 && gn >= 2 && tpf >= 64 && tpf <= 256) fr = 2;
    }
    int fpc = maxt / tpf;                 // groups per CTA: they share the staged class matrices
    if (fpc > 4 / fr) fpc = 4 / fr;
    {
        static const char* env = getenv("RMX_VITERBI_FPC");     // A/B measurements only
        if (env && atoi(env) >= 1 && atoi(env) < fpc) fpc = atoi(env);
    }
    const int groups = (gn + fr - 1) / fr;
    if (fpc > groups) fpc = groups;
    size_t smem = sizeof(double) * ((size_t)fpc * fr * 2 * SP + (size_t)staged * d.S * SP) + (size_t)d.S * d.M + 16;
    static RmxPerDeviceOnce attr_once;
    if (attr_once.need()) {
        cudaFuncSetAttribute(k_viterbi<1, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024);
        cudaFuncSetAttribute(k_viterbi<1, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024);
        cudaFuncSetAttribute(k_viterbi<2, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024);
    }
    const dim3 grid(ch.NB, (groups + fpc - 1) / fpc);
#define VIT_ARGS d, g0, gn, ch, rows, sp, sp_fit_stride, T.tab, F, lattice, path, staged, fpc
    if (fr == 2) k_viterbi<2, 512><<<grid, tpf * fpc, smem, st>>>(VIT_ARGS);
    else if (maxt == 512) k_viterbi<1, 512><<<grid, tpf * fpc, smem, st>>>(VIT_ARGS);
    else k_viterbi<1, 1024><<<grid, tpf * fpc, smem, st>>>(VIT_ARGS);
#undef VIT_ARGS
}

void rmx_launch_materialize_T(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxRows& rows,
                              const RmxTransSpec& sp, int64_t row0, int64_t nrows, double* out) {
    if (nrows <= 0) return;
    k_materialize_T<<<(int)nrows, 256, (size_t)d.S * d.M + 16, st>>>(d, rows, sp, T.tab, row0, out);
}

void rmx_launch_materialize_joint(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxRows& rows,
                                  const RmxTransSpec& sp, int uniform, const double* ahat, const double* bhat,
                                  const double* fexp, int64_t row0, int64_t nrows, double* out) {
    if (nrows <= 0) return;
    k_materialize_joint<<<(int)nrows, 256, pair_smem(d), st>>>(d, rows, sp, T.tab, uniform, ahat, bhat, fexp, row0, out);
}

void rmx_launch_sum_product_log(cudaStream_t st, int64_t N, int S, const double* F, const double* Tm, double* alphas,
                                double* betas) {
    int threads = ((S + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    k_sum_product_log<<<2, threads, sizeof(double) * S, st>>>(N, S, F, Tm, alphas, betas);
}

void rmx_launch_sum_product_2p(cudaStream_t st, int64_t N, int S, const double* F, double* alphas, double* betas,
                               double e0, double e1) {
    int threads = ((S + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    k_sum_product_2p<<<2, threads, sizeof(double) * S, st>>>(N, S, F, alphas, betas, log(e0), log(e1));
}

void rmx_launch_max_product(cudaStream_t st, int64_t N, int S, const double* F, const double* Tm, double* lattice,
                            int64_t* path, double* logprob) {
    int threads = ((S + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    k_max_product<<<1, threads, sizeof(double) * S, st>>>(N, S, F, Tm, lattice, path, logprob);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free Viterbi decode (rmx_max_product_noalloc).  k_max_product needs an (N, S) lattice of running maxima
// as scratch; here that scratch is a file-scope device array of RMX_NOALLOC_MAX_CELLS doubles (16 MiB, held by every
// GPU that loads the module until the process exits) instead of a cudaMalloc per call.  In host code the array's
// name is only a stand-in for the symbol: its device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and that address is what the kernel receives.  The kernel writes
// every lattice cell it later reads, so nothing has to be cleared between calls; all calls on one device share the
// array, so two calls whose kernels overlap on the GPU (different streams, no dependency) race.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_noalloc_lattice[RMX_NOALLOC_MAX_CELLS];

#define RMX_NOALLOC_MAX_DEVICES 64
static double* g_noalloc_ptr[RMX_NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mu;

const char* rmx_launch_max_product_noalloc(cudaStream_t st, int device, int64_t N, int S, const double* F,
                                           const double* Tm, int64_t* path, double* logprob) {
    if (device < 0 || device >= RMX_NOALLOC_MAX_DEVICES) return "device index out of range";
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cudaGetErrorString(e);
    double* lattice = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_lattice);
            if (e != cudaSuccess) return cudaGetErrorString(e);
            g_noalloc_ptr[device] = static_cast<double*>(p);
        }
        lattice = g_noalloc_ptr[device];
    }
    int threads = ((S + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    k_max_product<<<1, threads, sizeof(double) * S, st>>>(N, S, F, Tm, lattice, path, logprob);
    e = cudaGetLastError();
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}
