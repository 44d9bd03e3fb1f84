// K2 (S <= 128): forward / backward sum-product messages of EIGHT fits of an initialisation grid in lock step,
// one block barrier per segment.                                               (bpmodel.pyx:1213-1246, :921-946)
//
// In the scaled probability domain one message update is
//     out[g][j] = sum_i u[g][i] * E[i][j],       g = fit (8 per CTA), i, j = copy-number states,
// with E = exp(log_transmat) shared by every fit of the grid and by every non-breakend row of a chromosome: an
// (8 x S) x (S x S) FP64 product per segment.  With E in registers and u broadcast from shared memory the vector
// FP64 pipe starves on shared-memory bandwidth (8 B of u per CW FMAs; measured, profiles/r01_c_*), so the product
// runs on the FP64 tensor path instead: mma.sync.m8n8k4.f64 (DMMA), A = u (8 fits x 4 states, one LDS.64 per
// thread, stored in shared memory in fragment order), B = a 4 x 8 block of E held in ONE register per thread for
// the whole chain, C = 8 fits x 8 states.  Warp w owns output states 16w .. 16w+15 (two column blocks) and all
// KB = ceil(S/4) row blocks, so there is no cross-warp reduction: per segment a warp issues KB LDS.64 + 2 KB DMMA
// (four interleaved accumulator chains), then rescales by the power of two that normalised the previous message (exact; its exponent goes to the
// log normaliser), multiplies by f_n = exp(F[n] - max F[n]) (prefetched two segments ahead) and stores the message
// to ahat / bhat and, in fragment order, to shared memory for the next segment.  DMMA is IEEE FP64 FMA arithmetic.
//
// Breakend rows (2 per breakpoint, ~2 % of rows) have a per-fit matrix
//     E_n[from][to] = flipE[from][to] * prod_m wexp[g][r][m][tot_from[m] - tot_to[m]]
// which is not a shared-operand product; they take a scalar path (same thread <-> (rows, column) ownership as the
// B fragments) whose results are shuffled into the C-fragment layout so the epilogue is shared.
#include "common.cuh"
#include "kernels.h"

#define CM_GT 8
#define CM_NWMAX 16

__device__ __forceinline__ int cm_expfield(double v) { return (__double2hiint(v) >> 20) & 0x7ff; }
__device__ __forceinline__ double cm_pow2(int be) { return __hiloint2double((2046 - be) << 20, 0); }
__device__ __forceinline__ void cm_dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

struct ChainMmaArgs {
    RmxDims d;
    RmxChains ch;
    const int* rinfo;      // (N-1) packed row info: bits 0-7 class+1 (0 = telomere), 8-11 d0+8, 12.. breakend row+1
    RmxTransSpec sp;
    int64_t sp_fit_stride;
    const int8_t* tab;
    int g0, gn, ntiles;
    const double* fexp;    // (G,N,S)
    const double* mF;      // (G,N)
    double* ahat;          // (G,N,S)
    double* bhat;          // (G,N,S)
    double* chain_logz;    // (G,NC)
    int* dflags;
};

// message in fragment order: state i of fit g lives at (i/4)*32 + g*4 + i%4
__device__ __forceinline__ int cm_pos(int g, int i) { return ((i >> 2) << 5) + (g << 2) + (i & 3); }

// one CTA per SM, at most 8 warps: up to 232 registers per thread
#define CM_NH 2     // column blocks (of 8 states) per warp
#define CM_FST 4    // stages of the f ring (cp.async prefetch distance + 1)
#define CM_RIMAX 8192   // chains up to this many rows keep their row descriptors in shared memory

template <int KB>
struct CmCfg {
    static constexpr int NB = (KB + 1) / 2;                 // column blocks covering 4*KB >= S states
    static constexpr int NWARPS = (NB + CM_NH - 1) / CM_NH;
    static constexpr int MAXT = NWARPS * 32;
    // B fragments (the transition matrix) per column block: KREG in registers, the rest in a per-thread shared-memory
    // slot.  With all of them declared as registers the compiler ran out at 255 and re-loaded ~10 from L2 every
    // segment (LDG in front of the first DMMA: 24 % of the kernel's stall samples, profiles/r01_summary.md).
    static constexpr int KREG = KB < 19 ? KB : 19;
    static constexpr int KSM = KB - KREG;
};

template <int KB>
__global__ void __launch_bounds__(CmCfg<KB>::MAXT, 1) k_chain_mma(ChainMmaArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const RmxDims& d = A.d;
    const int S = d.S, M = d.M, MD = M * d.D;
    const int T = blockDim.x, NW = T >> 5;
    double* u = reinterpret_cast<double*>(smem_raw);                 // [2][KB*32]
    double* wtab = u + 2 * KB * 32;                                   // [8][MD]   breakend tables of the step
    double* redd = wtab + CM_GT * MD;                                 // [8][NWMAX] + [8] mF sums
    int* redi = reinterpret_cast<int*>(redd + CM_GT * CM_NWMAX + CM_GT);   // [2][8][NWMAX] max exponent fields
    int8_t* tots = reinterpret_cast<int8_t*>(redi + 2 * CM_GT * CM_NWMAX); // [S*M]
    constexpr int KREG = CmCfg<KB>::KREG, KSM = CmCfg<KB>::KSM;
    // [CM_NH * KSM][T] doubles after tots (16-byte aligned): this thread's overflow B fragments
    double* esm = reinterpret_cast<double*>(smem_raw + ((sizeof(double) * ((size_t)2 * KB * 32 + (size_t)CM_GT * MD + CM_GT * CM_NWMAX + CM_GT) +
                                                         sizeof(int) * 2 * CM_GT * CM_NWMAX + (size_t)S * M + 15) & ~(size_t)15)) + threadIdx.x;
    // f ring: [CM_FST stages][CM_NH][2][T] doubles after the B-fragment slots; every thread copies its own entries
    // with cp.async several segments ahead (no registers held across the copy, no barrier: own slots only)
    double* fring = esm + (size_t)CM_NH * KSM * T;
    double* fesm = fring + (size_t)CM_FST * CM_NH * 2 * T;        // [KB][T]: breakend rows' flip factors, own slots
    const unsigned fesm_s = (unsigned)__cvta_generic_to_shared(fesm);
    int* rism = reinterpret_cast<int*>(fesm + (size_t)KB * T - threadIdx.x);   // [CM_RIMAX] row info of the chain

    const int tile = blockIdx.x % A.ntiles;
    const int cd = blockIdx.x / A.ntiles;
    const int dir = cd & 1, chain = A.ch.order[cd >> 1];
    const int s0 = A.ch.start[chain], s1 = A.ch.end[chain];
    const int t = threadIdx.x, w = t >> 5, lane = t & 31;
    const int gid = lane >> 2, tig = lane & 3;     // mma fragment coordinates
    const int gbase = A.g0 + tile * CM_GT;
    const int gcount = min(CM_GT, A.g0 + A.gn - gbase);
    const int64_t SS = (int64_t)S * S;
    const int64_t NS = (int64_t)d.N * S;
    const int sgn = dir == 0 ? 1 : -1;
    // column block h of this warp: B fragment = column jb[h], rows 4*kb + tig; C fragment = fit gid, columns j0[h], +1
    const int jbase = w * (8 * CM_NH);
    const bool fit_ok = gid < gcount;
    const int gfit = gbase + min(gid, gcount - 1);
    int upos[CM_NH];
    bool ok0[CM_NH], ok1[CM_NH];
#pragma unroll
    for (int h = 0; h < CM_NH; ++h) {
        const int j0 = jbase + 8 * h + 2 * tig;
        ok0[h] = fit_ok && j0 < S;
        ok1[h] = fit_ok && j0 + 1 < S;
        upos[h] = cm_pos(gid, j0);                 // j0 even: j0, j0+1 adjacent, 16-byte aligned
    }

    for (int e = t; e < 2 * KB * 32; e += T) u[e] = 0.;
    for (int e = t; e < 2 * CM_GT * CM_NWMAX; e += T) redi[e] = ((e % CM_NWMAX) == 0) ? 1023 : 0;   // first message: max f = 1
    for (int e = t; e < S * M; e += T) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(A.tab[(s * M + m) * 2] + A.tab[(s * M + m) * 2 + 1]);
    }
    __syncthreads();

    const double* fp = A.fexp + (int64_t)gfit * NS + jbase + 2 * tig;    // + n*S + 8*h
    double* outp = (dir == 0 ? A.ahat : A.bhat) + (int64_t)gfit * NS + jbase + 2 * tig;
    const int nsteps = s1 - s0;

    // ---- first segment of the sweep: u = f (forward: ahat = f; backward: bhat = 1)
    double un[CM_NH][2];
    {
        const int64_t o = (int64_t)(dir == 0 ? s0 : s1) * S;
#pragma unroll
        for (int h = 0; h < CM_NH; ++h) {
            un[h][0] = un[h][1] = 0.;
            if (ok0[h]) { un[h][0] = fp[o + 8 * h]; outp[o + 8 * h] = dir == 0 ? un[h][0] : 1.; }
            if (ok1[h]) { un[h][1] = fp[o + 8 * h + 1]; outp[o + 8 * h + 1] = dir == 0 ? un[h][1] : 1.; }
            if (ok0[h]) *reinterpret_cast<double2*>(u + upos[h]) = make_double2(un[h][0], un[h][1]);
        }
    }
    auto seg_of = [&](int k) { return dir == 0 ? s0 + 1 + k : s1 - 1 - k; };   // segment produced at step k
    auto row_of = [&](int k) { return dir == 0 ? s0 + k : s1 - 1 - k; };       // its transition row
    // software pipeline: f CM_FST - 1 segments ahead through the shared-memory ring, row info two rows ahead
    int ri0 = 0, ri1 = 0;
    const unsigned fring_s = (unsigned)__cvta_generic_to_shared(fring);
    auto f_issue = [&](int k) {
        if (k < nsteps) {
            const int64_t o = (int64_t)seg_of(k) * S;
            const unsigned base = fring_s + 8u * (unsigned)(((k % CM_FST) * CM_NH * 2) * T);
#pragma unroll
            for (int h = 0; h < CM_NH; ++h) {
                if (ok0[h]) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(base + 8u * (unsigned)((h * 2) * T)), "l"(fp + o + 8 * h) : "memory");
                if (ok1[h]) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(base + 8u * (unsigned)((h * 2 + 1) * T)), "l"(fp + o + 8 * h + 1) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int e = 0; e < CM_FST * CM_NH * 2; ++e) fring[e * T] = 0.;     // entries outside the state space stay zero
#pragma unroll
    for (int k = 0; k < CM_FST - 1; ++k) f_issue(k);
    const bool ri_sm = nsteps <= CM_RIMAX;
    if (ri_sm)
        for (int k = t; k < nsteps; k += T) rism[k] = A.rinfo[row_of(k)];
    __syncthreads();
    auto ri_at = [&](int k) { return ri_sm ? rism[k] : A.rinfo[row_of(k)]; };
    if (nsteps > 0) ri0 = ri_at(0);
    if (nsteps > 1) ri1 = ri_at(1);

    double ereg[CM_NH][KREG];
    int cur_class = -1;
    int cur = 0, esum = 0, bad = 0;

    for (int step = 0; step < nsteps; ++step) {
        const int64_t on = (int64_t)seg_of(step) * S;
        const int ri = ri0;
        ri0 = ri1;
        if (step + 2 < nsteps) ri1 = ri_at(step + 2);
        f_issue(step + CM_FST - 1);       // into the ring stage the previous step consumed
        const int cls = (ri & 0xff) - 1;
        const int br = (ri >> 12) - 1;
        // exponents of the largest entry of the input message of fit gid, per warp (written before the last barrier)
        const int4* rp = reinterpret_cast<const int4*>(redi + (cur * CM_GT + gid) * CM_NWMAX);
        int4 r0 = rp[0], r1 = rp[1];
        if (CmCfg<KB>::NWARPS > 8) {        // more than eight warps: fold the upper half of the per-warp maxima in
            const int4 r2 = rp[2], r3 = rp[3];
            r0.x = max(r0.x, r2.x); r0.y = max(r0.y, r2.y); r0.z = max(r0.z, r2.z); r0.w = max(r0.w, r2.w);
            r1.x = max(r1.x, r3.x); r1.y = max(r1.y, r3.y); r1.z = max(r1.z, r3.z); r1.w = max(r1.w, r3.w);
        }
        const double* uc = u + cur * KB * 32;

        double c[CM_NH][2];
#pragma unroll
        for (int h = 0; h < CM_NH; ++h) c[h][0] = c[h][1] = 0.;
        if (br < 0) {
            if (cls != cur_class) {
                const double* E = A.sp.Emat + (int64_t)cls * SS;
#pragma unroll
                for (int h = 0; h < CM_NH; ++h)
#pragma unroll
                    for (int kb = 0; kb < KB; ++kb) {
                        const int i = kb * 4 + tig, jb = jbase + 8 * h + gid;
                        double ev = (i < S && jb < S) ? E[(int64_t)i * S + jb] : 0.;
                        asm volatile("" : "+d"(ev));     // opaque: keep the value, do not re-load it every segment
                        if (kb < KREG) ereg[h][kb < KREG ? kb : 0] = ev;
                        else esm[(h * KSM + (kb - KREG)) * T] = ev;
                    }
                cur_class = cls;
            }
            double e[CM_NH][2];
#pragma unroll
            for (int h = 0; h < CM_NH; ++h) e[h][0] = e[h][1] = 0.;
#pragma unroll
            for (int kb = 0; kb < KB; ++kb) {
                const double a = uc[kb * 32 + lane];
#pragma unroll
                for (int h = 0; h < CM_NH; ++h) {
                    const double eb = kb < KREG ? ereg[h][kb < KREG ? kb : 0] : esm[(h * KSM + (kb - KREG)) * T];
                    if (kb & 1) cm_dmma(e[h][0], e[h][1], a, eb);
                    else cm_dmma(c[h][0], c[h][1], a, eb);
                }
            }
#pragma unroll
            for (int h = 0; h < CM_NH; ++h) { c[h][0] += e[h][0]; c[h][1] += e[h][1]; }
        } else {
            // ---- breakend row: thread = (column jb, rows 4*kb + tig), all eight fits
            for (int e = t; e < CM_GT * MD; e += T) {
                const int g = e / MD, o = e - g * MD;
                const int gg = gbase + min(g, gcount - 1);
                wtab[e] = A.sp.wexp[(int64_t)gg * A.sp_fit_stride + (int64_t)br * MD + o];
            }
            __syncthreads();
            const int d0 = ((ri >> 8) & 15) - 8 + d.doff;     // the normal-clone factor does not depend on the states
#pragma unroll 1
            for (int h = 0; h < CM_NH; ++h) {
                const int jb = jbase + 8 * h + gid;
                double sum[CM_GT];
#pragma unroll
                for (int g = 0; g < CM_GT; ++g) sum[g] = 0.;
                if (jb < S) {
                    // this thread's KB entries of the row's flip factor: one round trip (cp.async into own slots)
                    // instead of KB dependent L2 loads
                    const double* FE = A.sp.flipE + (int64_t)cls * SS;
#pragma unroll
                    for (int kb = 0; kb < KB; ++kb) {
                        const int i = kb * 4 + tig;
                        if (i < S)
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(fesm_s + 8u * (unsigned)(kb * T)),
                                         "l"(FE + (int64_t)i * S + jb) : "memory");
                    }
                    asm volatile("cp.async.commit_group;" ::: "memory");
                    asm volatile("cp.async.wait_all;" ::: "memory");
#pragma unroll 1
                    for (int kb = 0; kb < KB; ++kb) {
                        const int i = kb * 4 + tig;
                        if (i >= S) break;
                        const double fe = fesm[kb * T];
                        double v[CM_GT];
#pragma unroll
                        for (int g = 0; g < CM_GT; ++g) v[g] = fe * uc[kb * 32 + g * 4 + tig];
                        for (int m = 1; m < M; ++m) {
                            const int dd = m * d.D + sgn * ((int)tots[i * M + m] - (int)tots[jb * M + m]) + d.doff;
#pragma unroll
                            for (int g = 0; g < CM_GT; ++g) v[g] *= wtab[g * MD + dd];
                        }
#pragma unroll
                        for (int g = 0; g < CM_GT; ++g) sum[g] += v[g];
                    }
                }
                double x0 = 0., x1 = 0.;
#pragma unroll
                for (int g = 0; g < CM_GT; ++g) {
                    double sg = sum[g];
                    sg += __shfl_xor_sync(0xffffffffu, sg, 1);
                    sg += __shfl_xor_sync(0xffffffffu, sg, 2);
                    sg *= wtab[g * MD + d0];
                    // into the C-fragment layout: (fit gid, columns 2*tig, 2*tig+1 of the block) <- column lanes
                    const double y0 = __shfl_sync(0xffffffffu, sg, (2 * tig) * 4);
                    const double y1 = __shfl_sync(0xffffffffu, sg, (2 * tig + 1) * 4);
                    if (g == gid) { x0 = y0; x1 = y1; }
                }
                if (h == 0) { c[0][0] = x0; c[0][1] = x1; }
                else { c[CM_NH - 1][0] = x0; c[CM_NH - 1][1] = x1; }
            }
        }
        // ---- epilogue (C-fragment layout)
        double f[CM_NH][2];
        asm volatile("cp.async.wait_group %0;" ::"n"(CM_FST - 1) : "memory");     // this step's f has landed
        {
            const double* fr = fring + ((step % CM_FST) * CM_NH * 2) * T;
#pragma unroll
            for (int h = 0; h < CM_NH; ++h) { f[h][0] = fr[(h * 2) * T]; f[h][1] = fr[(h * 2 + 1) * T]; }
        }
        int be = max(max(max(r0.x, r0.y), max(r0.z, r0.w)), max(max(r1.x, r1.y), max(r1.z, r1.w)));
        esum += be - 1023;
        if (be <= 0 || be >= 2047) bad = 1;
        const double sc = cm_pow2(be);
        int mx = 0;
#pragma unroll
        for (int h = 0; h < CM_NH; ++h) {
            double v0 = c[h][0] * sc, v1 = c[h][1] * sc;
            if (dir == 0) {
                v0 *= f[h][0]; v1 *= f[h][1];
                un[h][0] = v0; un[h][1] = v1;
            } else {
                un[h][0] = v0 * f[h][0]; un[h][1] = v1 * f[h][1];
            }
            if (!ok0[h]) un[h][0] = 0.;
            if (!ok1[h]) un[h][1] = 0.;
            if (ok0[h]) outp[on + 8 * h] = v0;
            if (ok1[h]) outp[on + 8 * h + 1] = v1;
            if (ok0[h]) *reinterpret_cast<double2*>(u + (cur ^ 1) * KB * 32 + upos[h]) = make_double2(un[h][0], un[h][1]);
            mx = max(mx, max(cm_expfield(un[h][0]), cm_expfield(un[h][1])));
        }
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        if (tig == 0) redi[((cur ^ 1) * CM_GT + gid) * CM_NWMAX + w] = mx;
        cur ^= 1;
        __syncthreads();
    }

    // ---- log normaliser of the forward sweep: log(sum of the last message) + ln2 * esum + sum_n max F[n]
    {
        double sv = 0.;
#pragma unroll
        for (int h = 0; h < CM_NH; ++h) sv += un[h][0] + un[h][1];
        sv += __shfl_xor_sync(0xffffffffu, sv, 1);
        sv += __shfl_xor_sync(0xffffffffu, sv, 2);
        if (tig == 0) redd[gid * CM_NWMAX + w] = sv;
        for (int g = w; g < gcount; g += NW) {
            const double* mf = A.mF + (int64_t)(gbase + g) * d.N;
            double ms = 0.;
            for (int n = s0 + lane; n <= s1; n += 32) ms += mf[n];
            ms = warp_sum(ms);
            if (lane == 0) redd[CM_GT * CM_NWMAX + g] = ms;
        }
    }
    __syncthreads();
    if (w == 0 && tig == 0 && fit_ok) {
        double slast = 0.;
        for (int c2 = 0; c2 < NW; ++c2) slast += redd[gid * CM_NWMAX + c2];
        if (dir == 0) {
            const double logz = log(slast) + 0.693147180559945309417 * (double)esum + redd[CM_GT * CM_NWMAX + gid];
            A.chain_logz[(int64_t)(gbase + gid) * A.ch.NC + chain] = logz;
            if (isnan(logz)) bad = 1;
        }
        if (!(slast > 0.) || isinf(slast)) bad = 1;
        if (bad) atomicOr(A.dflags, RMX_F_MSG_NAN);
    }
}

template <int KB>
static int launch_chain_mma_cfg(cudaStream_t st, ChainMmaArgs& A) {
    const RmxDims& d = A.d;
    const int NB = (d.S + 7) / 8;
    const int NW = (NB + CM_NH - 1) / CM_NH;
    const int threads = NW * 32;
    if (threads > CmCfg<KB>::MAXT || NW > CM_NWMAX || 4 * KB < d.S) return 2;
    size_t smem = sizeof(double) * ((size_t)2 * KB * 32 + (size_t)CM_GT * d.M * d.D + CM_GT * CM_NWMAX + CM_GT) +
                  sizeof(int) * 2 * CM_GT * CM_NWMAX + (size_t)d.S * d.M + 16;
    smem = ((smem + 15) & ~(size_t)15) + sizeof(double) * ((size_t)CM_NH * CmCfg<KB>::KSM + (size_t)CM_FST * CM_NH * 2 + KB) * threads +
           sizeof(int) * CM_RIMAX + 16;
    auto kern = k_chain_mma<KB>;
    if (smem > 48 * 1024) {
        if (smem > 227 * 1024) return 2;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 1;
    }
    A.ntiles = (A.gn + CM_GT - 1) / CM_GT;
    kern<<<A.ch.NC * 2 * A.ntiles, threads, smem, st>>>(A);
    return cudaPeekAtLastError() == cudaSuccess ? 0 : 1;
}

// returns 0 on success, 1 if the configuration or the launch failed, 2 if this kernel does not cover the shape
int rmx_launch_chain_mma(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                         const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                         const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags) {
    if (ch.NC <= 0 || gn <= 0) return 0;
    if (d.S > 128 || rinfo == nullptr) return 2;
    ChainMmaArgs A;
    A.d = d; A.ch = ch; A.rinfo = rinfo; A.sp = sp; A.sp_fit_stride = sp_fit_stride; A.tab = T.tab;
    A.g0 = g0; A.gn = gn; A.ntiles = 1; A.fexp = fexp; A.mF = mF; A.ahat = ahat; A.bhat = bhat;
    A.chain_logz = chain_logz; A.dflags = dflags;
    if (d.S <= 16) return launch_chain_mma_cfg<4>(st, A);
    if (d.S <= 32) return launch_chain_mma_cfg<8>(st, A);
    if (d.S <= 64) return launch_chain_mma_cfg<16>(st, A);
    if (d.S <= 100) return launch_chain_mma_cfg<25>(st, A);
    return launch_chain_mma_cfg<32>(st, A);
}
