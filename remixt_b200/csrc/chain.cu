// Chain kernels: the structured part of the variational update.
//
//   k_build_class  : the S x S log-transition matrix of "plain" rows, exp'd once     (bpmodel.pyx:640-684)
//   k_prep_brk     : per-breakend-row expected transition tables under p_breakpoint  (bpmodel.pyx:658-664)
//   k_chain        : forward / backward sum-product messages, one CTA per chromosome
//                    chain and direction, in the scaled probability domain            (bpmodel.pyx:1213-1246)
//   k_post         : posterior marginals                                              (bpmodel.pyx:948-950)
//   k_hist         : joint-marginal histograms by total-copy difference at breakend
//                    rows                                                             (bpmodel.pyx:619-633, 954-960)
//   k_brk_update   : update_p_breakpoint                                              (bpmodel.pyx:964-983)
//   k_pairsum      : sum_ij joint[n,i,j] * log_transmat[n,i,j] per row (ELBO terms)   (bpmodel.pyx:1052, 1112-1115)
//   k_viterbi      : max_product + traceback per chain                                (bpmodel.pyx:1296-1333)
//
// The reference materialises three (N-1,S,S) arrays; nothing here does.  A log-domain message
//   alpha[n,j] = LSE_i(alpha[n-1,i] + T[i,j]) + F[n,j]
// is evaluated as a scaled matrix-vector product  a_n = (a_{n-1}^T E) .* f_n / c_n  with
// E = exp(T) held in registers across the whole chain, f_n = exp(F[n]-max F[n]); the chain's
// log normaliser is sum_n (log c_n + max F[n]).  Chromosomes are coupled only through telomere
// rows whose log_transmat is identically zero, so chains are exactly independent.
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
__global__ void k_prep_brk(RmxDims d, const int* __restrict__ brk_rows, const int* __restrict__ bidx,
                           const int* __restrict__ borient, const int8_t* __restrict__ brk,
                           const double* __restrict__ p_brk, double pen, int tm, double* __restrict__ pd,
                           double* __restrict__ wexp) {
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
// K2: forward / backward messages.  grid = (chains, 2); thread t -> (column j = t % S, slice q = t / S);
// slice q owns input indices [q*RR, (q+1)*RR).  REG: the thread's RR entries of E live in registers for
// the whole chain (reloaded only when the row's transition class changes, which is rare); otherwise E is
// streamed from L2 each step.  Shared memory: vec[2][Sp] (double buffered message), part[Q][Sp]
// (slice partial sums), red[32], wtab[M*D] (breakend-row table), tots[S*M] (int8 totals).
// ---------------------------------------------------------------------------------------------
template <int RR, bool REG>
__global__ void __launch_bounds__(REG ? 512 : 1024)
k_chain(RmxDims d, RmxChains ch, RmxRows rows, RmxTransSpec sp, const int8_t* __restrict__ tab, int Q, int rr_rt,
        const double* __restrict__ fexp, const double* __restrict__ mF, double* __restrict__ ahat,
        double* __restrict__ bhat, double* __restrict__ cnorm, double* __restrict__ chain_logz,
        int* __restrict__ dflags) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = d.S, M = d.M;
    const int rr = REG ? RR : rr_rt;
    const int Sp = Q * rr;
    double* vec = reinterpret_cast<double*>(smem_raw);   // [2][Sp]
    double* part = vec + 2 * Sp;                          // [Q][Sp]
    double* red = part + Q * Sp;                          // [32]
    double* wtab = red + 32;                              // [M*D]
    int8_t* tots = reinterpret_cast<int8_t*>(wtab + M * d.D);   // [S*M]

    const int chain = blockIdx.x, dir = blockIdx.y;
    const int s0 = ch.start[chain], s1 = ch.end[chain];
    const int t = threadIdx.x;
    const int j = t % S, q = t / S;
    const bool act = q < Q;
    const int i0 = act ? q * rr : 0;
    const int lane = t & 31, wid = t >> 5;
    const int nwS = (S + 31) >> 5;   // warps that hold the S output lanes (t < S)
    const int64_t SS = (int64_t)S * S;

    for (int e = t; e < 2 * Sp; e += blockDim.x) vec[e] = 0.;
    for (int e = t; e < S * M; e += blockDim.x) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(tab[(s * M + m) * 2] + tab[(s * M + m) * 2 + 1]);
    }
    __syncthreads();

    double ereg[REG ? RR : 1];
    int cur_class = -1;
    int cur = 0;
    double logz = 0.;
    int bad = 0;

    // ---- first segment of the sweep
    {
        double v = 0.;
        if (dir == 0) {
            if (t < S) v = fexp[(int64_t)s0 * S + t];
            double w = warp_sum(v);
            if (lane == 0 && wid < 32) red[wid] = w;
            __syncthreads();
            double c = 0.;
            for (int w2 = 0; w2 < nwS; ++w2) c += red[w2];
            if (t < S) {
                double a = v / c;
                ahat[(int64_t)s0 * S + t] = a;
                vec[t] = a;
            }
            if (t == 0) {
                cnorm[s0] = c;
                logz = log(c) + mF[s0];
                if (!(c > 0.)) bad = 1;
            }
        } else {
            if (t < S) {
                bhat[(int64_t)s1 * S + t] = 1.;
                vec[t] = fexp[(int64_t)s1 * S + t];
            }
        }
        __syncthreads();
    }

    const int nsteps = s1 - s0;
    for (int step = 0; step < nsteps; ++step) {
        const int n = dir == 0 ? s0 + 1 + step : s1 - 1 - step;   // segment being produced
        const int row = dir == 0 ? n - 1 : n;                     // transition row n -> n+1
        const int cls = rows.tclass[row];
        const int br = rows.brow[row];
        double fn = (t < S) ? fexp[(int64_t)n * S + t] : 0.;      // issued early, used in the epilogue
        const double* vin = vec + cur * Sp;
        double acc0 = 0., acc1 = 0.;
        if (br < 0) {
            if (REG) {
                if (cls != cur_class) {
                    const double* E = sp.Emat + (int64_t)cls * SS;   // symmetric: E[i][j] == E[j][i]
#pragma unroll
                    for (int k = 0; k < RR; ++k) {
                        int i = i0 + k;
                        ereg[REG ? k : 0] = (act && i < S) ? E[(int64_t)i * S + j] : 0.;
                    }
                    cur_class = cls;
                }
#pragma unroll
                for (int k = 0; k < RR; k += 2) {
                    acc0 = fma(vin[i0 + k], ereg[REG ? k : 0], acc0);
                    if (k + 1 < RR) acc1 = fma(vin[i0 + k + 1], ereg[REG ? k + 1 : 0], acc1);
                }
            } else if (act) {
                const double* E = sp.Emat + (int64_t)cls * SS;
                int kmax = min(rr, S - i0);
                for (int k = 0; k < kmax; ++k) acc0 = fma(vin[i0 + k], E[(int64_t)(i0 + k) * S + j], acc0);
            }
        } else {
            // breakend row: E_n[i][jj] = flipE[cls][i][jj] * prod_m wexp[r][m][tot_i - tot_jj]
            for (int e = t; e < M * d.D; e += blockDim.x) wtab[e] = sp.wexp[(int64_t)br * M * d.D + e];
            __syncthreads();
            if (act) {
                const double* FE = sp.flipE + (int64_t)cls * SS;
                const int d0 = rows.d0[row];
                int kmax = min(rr, S - i0);
                for (int k = 0; k < kmax; ++k) {
                    int o = i0 + k;
                    // forward: from-state o, to-state j ; backward: from-state j, to-state o
                    double e = dir == 0 ? brk_E(wtab, tots, M, d.D, d.doff, d0, o, j, FE[(int64_t)o * S + j])
                                        : brk_E(wtab, tots, M, d.D, d.doff, d0, j, o, FE[(int64_t)o * S + j]);
                    acc0 = fma(vin[o], e, acc0);
                }
            }
        }
        if (act) part[q * Sp + j] = acc0 + acc1;
        __syncthreads();
        double v = 0.;
        if (t < S) {
            for (int qq = 0; qq < Q; ++qq) v += part[qq * Sp + t];
            if (dir == 0) v *= fn;
        }
        {
            double w = warp_sum(v);
            if (lane == 0 && wid < nwS) red[wid] = w;
        }
        __syncthreads();
        double c = 0.;
        for (int w2 = 0; w2 < nwS; ++w2) c += red[w2];
        double* vout = vec + (cur ^ 1) * Sp;
        if (t < S) {
            double o = v / c;
            if (dir == 0) {
                ahat[(int64_t)n * S + t] = o;
                vout[t] = o;
            } else {
                bhat[(int64_t)n * S + t] = o;
                vout[t] = o * fn;
            }
        }
        if (t == 0) {
            if (dir == 0) {
                cnorm[n] = c;
                logz += log(c) + mF[n];
            }
            if (!(c > 0.) || isinf(c)) bad = 1;
        }
        cur ^= 1;
        __syncthreads();
    }
    if (t == 0) {
        if (dir == 0) chain_logz[chain] = logz;
        if (bad || isnan(logz)) atomicOr(dflags, RMX_F_MSG_NAN);
    }
}

// posterior marginals: post[n] = normalise(ahat[n] .* bhat[n]); zp[n] = the normaliser
__global__ void __launch_bounds__(256) k_post(RmxDims d, const double* __restrict__ ahat, const double* __restrict__ bhat,
                                              double* __restrict__ post, double* __restrict__ zp,
                                              int* __restrict__ dflags) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.N) return;
    int64_t base = (int64_t)warp * d.S;
    double z = 0.;
    for (int s = lane; s < d.S; s += 32) z += ahat[base + s] * bhat[base + s];
    z = warp_sum(z);
    bool bad = !(z > 0.);
    for (int s = lane; s < d.S; s += 32) {
        double p = ahat[base + s] * bhat[base + s] / z;
        post[base + s] = p;
        bad |= isnan(p);
    }
    if (lane == 0) zp[warp] = z;
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
__global__ void __launch_bounds__(256) k_hist(RmxDims d, RmxRows rows, RmxTransSpec sp, const int8_t* __restrict__ tab,
                                              int uniform, const double* __restrict__ ahat,
                                              const double* __restrict__ bhat, const double* __restrict__ fexp,
                                              double* __restrict__ hist, int* __restrict__ dflags) {
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

// update_p_breakpoint (bpmodel.pyx:964-983): one thread per breakpoint, reference accumulation order.
__global__ void k_brk_update(RmxDims d, const int* __restrict__ krow_ptr, const int* __restrict__ krow_idx,
                             const int* __restrict__ brk_rows, const int* __restrict__ borient,
                             const int8_t* __restrict__ brk, const double* __restrict__ hist, double pen, int tm,
                             double* __restrict__ p_brk, double* __restrict__ logq_scratch) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= d.K) return;
    double* lq = logq_scratch + (int64_t)k * d.B;
    for (int b = 0; b < d.B; ++b) lq[b] = 0.;
    const double mult = -pen;
    for (int e = krow_ptr[k]; e < krow_ptr[k + 1]; ++e) {
        int r = krow_idx[e];
        int o = borient[brk_rows[r]];
        for (int m = 0; m < d.M; ++m) {
            const double* h = hist + ((int64_t)r * d.M + m) * d.D;
            for (int b = 0; b < d.B; ++b) {
                double acc = lq[b];
                int bs = brk[b * d.M + m];
                for (int slot = 0; slot < d.D; ++slot)
                    acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(mult, h[slot]), ftrans(slot - d.doff - o * bs, tm)));
                lq[b] = acc;
            }
        }
    }
    // _exp_normalize (bpmodel.pyx:120-128)
    double mx = -INFINITY;
    for (int b = 0; b < d.B; ++b) if (lq[b] > mx) mx = lq[b];
    double ps = 0.;
    for (int b = 0; b < d.B; ++b) ps += exp(lq[b] - mx);
    double lse = log(ps) + mx;
    double z = 0.;
    double* p = p_brk + (int64_t)k * d.B;
    for (int b = 0; b < d.B; ++b) {
        p[b] = exp(lq[b] - lse);
        z += p[b];
    }
    for (int b = 0; b < d.B; ++b) p[b] /= z;
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

// rowsum[n] = sum_ij joint[n,i,j] * T_spec[n,i,j]; joint from (joint_sp, messages), T from t_sp.
__global__ void __launch_bounds__(256) k_pairsum(RmxDims d, RmxRows rows, RmxTransSpec jsp, RmxTransSpec tsp,
                                                 const int8_t* __restrict__ tab, int uniform,
                                                 const double* __restrict__ ahat, const double* __restrict__ bhat,
                                                 const double* __restrict__ fexp, double* __restrict__ rowsum) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red[32];
    int n = blockIdx.x;
    LogTCtx lt;
    // T of the row under t_sp; rows whose T is identically zero contribute nothing
    if (rows.tclass[n] < 0) {
        if (threadIdx.x == 0) rowsum[n] = 0.;
        return;
    }
    PairCtx c;
    pair_setup(c, smem_raw, d, rows, jsp, tab, uniform, n, ahat, bhat, fexp);
    logT_setup(lt, d, rows, tsp, c.tots, n);
    const int S = d.S;
    const int64_t SS = (int64_t)S * S;
    double z = 0., num = 0.;
    for (int64_t idx = threadIdx.x; idx < SS; idx += blockDim.x) {
        int i = (int)(idx / S), j = (int)(idx - (int64_t)i * S);
        double w = pair_w(c, d, i, j);
        z += w;
        num += w * logT(lt, d, i, j);
    }
    z = block_sum_all(z, red);
    num = block_sum_all(num, red);
    if (threadIdx.x == 0) rowsum[n] = num / z;
}

// ---------------------------------------------------------------------------------------------
// Viterbi per chain (bpmodel.pyx:1296-1333): first-max tie breaking (strict >, lowest index wins).
// Telomere rows are all-zero so the lattice of a chain only differs from the reference's by an
// additive constant; the traceback at a telomere is argmax of the lattice row itself.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_viterbi(RmxDims d, RmxChains ch, RmxRows rows, RmxTransSpec sp,
                                                  const int8_t* __restrict__ tab, const double* __restrict__ F,
                                                  double* __restrict__ lattice, int* __restrict__ path) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double rv[32];
    __shared__ int ri[32];
    __shared__ int s_next;
    const int S = d.S, M = d.M;
    double* prev = reinterpret_cast<double*>(smem_raw);   // [S]
    int8_t* tots = reinterpret_cast<int8_t*>(prev + S);   // [S*M]
    const int chain = blockIdx.x;
    const int s0 = ch.start[chain], s1 = ch.end[chain];
    const int t = threadIdx.x;
    for (int e = t; e < S * M; e += blockDim.x) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(tab[(s * M + m) * 2] + tab[(s * M + m) * 2 + 1]);
    }
    for (int jj = t; jj < S; jj += blockDim.x) {
        double v = F[(int64_t)s0 * S + jj];
        lattice[(int64_t)s0 * S + jj] = v;
        prev[jj] = v;
    }
    __syncthreads();
    for (int n = s0 + 1; n <= s1; ++n) {
        LogTCtx lt;
        logT_setup(lt, d, rows, sp, tots, n - 1);
        double out[4];   // up to 4 columns per thread (S <= 4096)
        int cnt = 0;
        for (int jj = t; jj < S; jj += blockDim.x) {
            double vmax = -INFINITY;
            for (int i = 0; i < S; ++i) {
                double v = __dadd_rn(prev[i], logT(lt, d, i, jj));
                if (v > vmax) vmax = v;
            }
            out[cnt++] = __dadd_rn(vmax, F[(int64_t)n * S + jj]);
        }
        __syncthreads();
        cnt = 0;
        for (int jj = t; jj < S; jj += blockDim.x) {
            prev[jj] = out[cnt];
            lattice[(int64_t)n * S + jj] = out[cnt];
            ++cnt;
        }
        __syncthreads();
    }
    // traceback
    for (int n = s1; n >= s0; --n) {
        LogTCtx lt;
        int nxt = 0;
        if (n < s1) {
            logT_setup(lt, d, rows, sp, tots, n);
            nxt = s_next;
        }
        double bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int i = t; i < S; i += blockDim.x) {
            double v = lattice[(int64_t)n * S + i];
            if (n < s1) v = __dadd_rn(v, logT(lt, d, i, nxt));
            if (v > bv) { bv = v; bi = i; }   // ascending i per thread: first max kept
        }
        // warp argmax with lowest-index tie break
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
            if (fi == 0x7fffffff) fi = 0;   // all -inf / NaN: _argmax returns 0
            path[n] = fi;
            s_next = fi;
        }
        __syncthreads();
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

void rmx_launch_prep_brk(cudaStream_t st, const RmxDims& d, int R, const int* brk_rows, const int* bidx,
                         const int* borient, const int8_t* brk, const double* p_brk, double pen, int tm, double* pd,
                         double* wexp) {
    if (R <= 0) return;
    k_prep_brk<<<R, 64, 0, st>>>(d, brk_rows, bidx, borient, brk, p_brk, pen, tm, pd, wexp);
}

template <int RR>
static int launch_chain_reg(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch,
                            const RmxRows& rows, const RmxTransSpec& sp, int Q, const double* fexp, const double* mF,
                            double* ahat, double* bhat, double* cnorm, double* chain_logz, int* dflags) {
    int Sp = Q * RR;
    size_t smem = sizeof(double) * (2 * Sp + Q * Sp + 32 + d.M * d.D) + (size_t)d.S * d.M + 16;
    int threads = ((d.S * Q + 31) / 32) * 32;
    auto kern = k_chain<RR, true>;
    if (smem > 48 * 1024) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 1;
    }
    dim3 grid(ch.NC, 2);
    kern<<<grid, threads, smem, st>>>(d, ch, rows, sp, T.tab, Q, RR, fexp, mF, ahat, bhat, cnorm, chain_logz, dflags);
    return 0;
}

int rmx_launch_chain(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const RmxRows& rows,
                     const RmxTransSpec& sp, const double* fexp, const double* mF, double* ahat, double* bhat,
                     double* cnorm, double* chain_logz, int* dflags) {
    if (ch.NC <= 0) return 0;
    const int S = d.S;
    if (S <= 16) return launch_chain_reg<4>(st, T, d, ch, rows, sp, 4, fexp, mF, ahat, bhat, cnorm, chain_logz, dflags);
    if (S <= 64) return launch_chain_reg<16>(st, T, d, ch, rows, sp, 4, fexp, mF, ahat, bhat, cnorm, chain_logz, dflags);
    if (S <= 100) return launch_chain_reg<25>(st, T, d, ch, rows, sp, 4, fexp, mF, ahat, bhat, cnorm, chain_logz, dflags);
    if (S <= 128) return launch_chain_reg<32>(st, T, d, ch, rows, sp, 4, fexp, mF, ahat, bhat, cnorm, chain_logz, dflags);
    // generic: E streamed from L2
    int Q = 1024 / S;
    if (Q < 1) Q = 1;
    if (Q > 8) Q = 8;
    int threads, rr;
    if (S > 1024) return 2;   // would need a column loop; not a supported state-space size
    rr = (S + Q - 1) / Q;
    threads = ((S * Q + 31) / 32) * 32;
    if (threads > 1024) { Q = 1024 / S; rr = (S + Q - 1) / Q; threads = ((S * Q + 31) / 32) * 32; }
    int Sp = Q * rr;
    size_t smem = sizeof(double) * (2 * Sp + Q * Sp + 32 + d.M * d.D) + (size_t)d.S * d.M + 16;
    auto kern = k_chain<1, false>;
    if (smem > 48 * 1024) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 1;
    }
    dim3 grid(ch.NC, 2);
    kern<<<grid, threads, smem, st>>>(d, ch, rows, sp, T.tab, Q, rr, fexp, mF, ahat, bhat, cnorm, chain_logz, dflags);
    return 0;
}

void rmx_launch_post(cudaStream_t st, const RmxDims& d, const double* ahat, const double* bhat, double* post,
                     double* zp, int* dflags) {
    if (d.N <= 0) return;
    k_post<<<cdiv64((int64_t)d.N * 32, 256), 256, 0, st>>>(d, ahat, bhat, post, zp, dflags);
}

static size_t pair_smem(const RmxDims& d) {
    return sizeof(double) * (2 * (size_t)d.S + (size_t)d.M * d.D) + (size_t)d.S * d.M + 16;
}

void rmx_launch_hist(cudaStream_t st, const RmxTables& T, const RmxDims& d, int R, const RmxRows& rows,
                     const RmxTransSpec& sp, int uniform, const double* ahat, const double* bhat, const double* fexp,
                     double* hist, int* dflags) {
    if (R <= 0) return;
    k_hist<<<R, 256, pair_smem(d), st>>>(d, rows, sp, T.tab, uniform, ahat, bhat, fexp, hist, dflags);
}

void rmx_launch_brk_update(cudaStream_t st, const RmxDims& d, const int* krow_ptr, const int* krow_idx,
                           const int* brk_rows, const int* borient, const int8_t* brk, const double* hist, double pen,
                           int tm, double* p_brk, double* logq_scratch) {
    if (d.K <= 0) return;
    k_brk_update<<<cdiv64(d.K, 64), 64, 0, st>>>(d, krow_ptr, krow_idx, brk_rows, borient, brk, hist, pen, tm, p_brk,
                                                 logq_scratch);
}

void rmx_launch_pairsum(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxRows& rows,
                        const RmxTransSpec& joint_sp, const RmxTransSpec& t_sp, int uniform, const double* ahat,
                        const double* bhat, const double* fexp, double* rowsum) {
    if (d.N <= 1) return;
    k_pairsum<<<d.N - 1, 256, pair_smem(d), st>>>(d, rows, joint_sp, t_sp, T.tab, uniform, ahat, bhat, fexp, rowsum);
}

void rmx_launch_viterbi(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch,
                        const RmxRows& rows, const RmxTransSpec& sp, const double* F, double* lattice, int* path) {
    if (ch.NC <= 0) return;
    int threads = ((d.S + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    size_t smem = sizeof(double) * d.S + (size_t)d.S * d.M + 16;
    k_viterbi<<<ch.NC, threads, smem, st>>>(d, ch, rows, sp, T.tab, F, lattice, path);
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
