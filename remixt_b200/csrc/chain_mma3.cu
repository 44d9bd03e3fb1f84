// K2 for state spaces of 98 .. 168 states (config 3: max copy number 8, S = 165): forward / backward sum-product
// messages of EIGHT fits in lock step on the FP64 tensor path (mma.sync.m8n8k4.f64), one block barrier per segment.
//                                                                              (bpmodel.pyx:1213-1246, :921-946)
//
// chain_mma2.cu keeps the whole transition matrix of a class in registers as B fragments (97 x 97 doubles = 75 KB over
// four warps).  At S = 165 the matrix is 218 KB: it fits neither the register file next to everything else nor shared
// memory next to the message buffers, so here it lives in BOTH: seven DMMA warps own three column blocks each
// (21 x 8 = 168 columns); of the KB = 42 row blocks of a column block the first KR stay in registers, the rest sit in
// per-thread shared-memory slots ([fragment][thread]: conflict-free 8-byte loads, one per DMMA, issued under the
// 16-cycle DMMA interval).  The streamed kernel this replaces (chain.cu) re-reads the matrix from L2 every segment.
//
// Everything else is chain_mma2.cu's scheme: messages scaled by exact powers of two, f through a cp.async ring,
// bookkeeping inside the DMMA stream, the epilogue of a column block under the next block's DMMAs, breakend rows on the
// tensor path through total classes.  Two differences on breakend rows: the class-ordered flipE fragments are read
// from global memory (L2 hits; 1.8 % of rows -- they do not fit shared memory here), and the table of total-difference
// tuples is indexed with 16 bits (17 x 17 = 289 tuples at max copy number 8).
#include "common.cuh"
#include "kernels.h"

#define C3_GT 8          // fits per CTA (the M dimension of m8n8k4)
#define C3_NW 8          // slots of the per-warp exponent maxima
#define C3_FST 3         // stages of the f ring (cp.async prefetch distance + 1)
#define C3_RIMAX 4096    // chains up to this many rows keep their row descriptors in shared memory

__device__ __forceinline__ int c3_expfield(double v) { return (__double2hiint(v) >> 20) & 0x7ff; }
__device__ __forceinline__ double c3_pow2(int be) { return __hiloint2double((2046 - be) << 20, 0); }
__device__ __forceinline__ void c3_dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void c3_cp8(unsigned dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}

struct Chain3Args {
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
    const uint8_t* bc;     // [perm (4 KBB)] [kbcls (KBB)] [scls (S)] [didx8 (BC, BC), unused here] [ddel (ND, M)]
    const uint16_t* didx;  // (BC, BC): which difference tuple is class ci minus class cj
    int BC, KBB, ND;
    const uint16_t* cord;  // k_chain_mma_s: states in class order (S)
    const int* cstart;     // k_chain_mma_s: [number of classes, class starts ..., S]
};

// message in fragment order: state i of fit g lives at (i/4)*32 + g*4 + i%4
__device__ __forceinline__ int c3_pos(int g, int i) { return ((i >> 2) << 5) + (g << 2) + (i & 3); }

// shared-memory layout, in bytes from the base (every block 16-byte aligned)
struct C3Layout {
    size_t u, wtab, redd, redi, tots, fring, esm, Wd, bct, didx, rism, total;
};
__host__ __device__ inline size_t c3_al(size_t x) { return (x + 15) & ~(size_t)15; }
__host__ __device__ inline C3Layout c3_layout(int KB, int T, int NSLOT, int NOVF, int S, int M, int MD, int BC, int KBB, int ND) {
    C3Layout L;
    size_t o = 0;
    L.u = o;     o = c3_al(o + sizeof(double) * 2 * KB * 32);
    L.wtab = o;  o = c3_al(o + sizeof(double) * C3_GT * MD);
    L.redd = o;  o = c3_al(o + sizeof(double) * (C3_GT * C3_NW + C3_GT));
    L.redi = o;  o = c3_al(o + sizeof(int) * 2 * C3_GT * C3_NW);
    L.tots = o;  o = c3_al(o + (size_t)S * M);
    L.fring = o; o = c3_al(o + sizeof(double) * C3_FST * NSLOT * T);
    L.esm = o;   o = c3_al(o + sizeof(double) * (size_t)NOVF * T);
    L.Wd = o;    o = c3_al(o + sizeof(double) * C3_GT * ND);
    L.bct = o;   o = c3_al(o + (size_t)5 * KBB + S + (size_t)BC * BC + (size_t)ND * M);
    L.didx = o;  o = c3_al(o + sizeof(uint16_t) * BC * BC);
    L.rism = o;  o = c3_al(o + sizeof(int) * C3_RIMAX);
    L.total = o;
    return L;
}

template <int KB, int NW, int NBW, int KR>
__global__ void __launch_bounds__(NW * 32, 1) k_chain_mma3(Chain3Args A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int T = NW * 32, NSLOT = 2 * NBW, KO = KB - KR, NOVF = NBW * KO;
    const RmxDims& d = A.d;
    const int S = d.S, M = d.M, MD = M * d.D;
    const int BC = A.BC, KBB = A.KBB, ND = A.ND;
    const C3Layout L = c3_layout(KB, T, NSLOT, NOVF, S, M, MD, BC, KBB, ND);
    double* u = reinterpret_cast<double*>(smem_raw + L.u);             // [2][KB*32]
    double* wtab = reinterpret_cast<double*>(smem_raw + L.wtab);       // [8][MD]   breakend tables of the step
    double* redd = reinterpret_cast<double*>(smem_raw + L.redd);       // [8][C3_NW] + [8] mF sums
    int* redi = reinterpret_cast<int*>(smem_raw + L.redi);             // [2][8][C3_NW] max exponent fields
    int8_t* tots = reinterpret_cast<int8_t*>(smem_raw + L.tots);       // [S*M]
    double* fring = reinterpret_cast<double*>(smem_raw + L.fring) + threadIdx.x;   // [C3_FST][NSLOT][T], own column
    double* esm = reinterpret_cast<double*>(smem_raw + L.esm) + threadIdx.x;       // [NOVF][T], own column
    double* Wd = reinterpret_cast<double*>(smem_raw + L.Wd);           // [8][ND]
    uint8_t* bct = smem_raw + L.bct;
    const uint8_t* bperm = bct;                   // row of the class-ordered layout -> state (255 = padding)
    const uint8_t* kbcls = bct + 4 * KBB;         // class of a row block
    const uint8_t* scls = kbcls + KBB;            // class of a state
    const uint8_t* ddel = scls + S + BC * BC;     // (ND, M): the tuple's per-clone differences + 128
    const int bct_n = 5 * KBB + S + BC * BC + ND * M;
    uint16_t* didx = reinterpret_cast<uint16_t*>(smem_raw + L.didx);   // (BC, BC)
    int* rism = reinterpret_cast<int*>(smem_raw + L.rism);             // [C3_RIMAX] row info of the chain

    const int tile = blockIdx.x % A.ntiles;
    const int cd = blockIdx.x / A.ntiles;
    const int dir = cd & 1, chain = A.ch.order[cd >> 1];
    const int s0 = A.ch.start[chain], s1 = A.ch.end[chain];
    const int t = threadIdx.x, w = t >> 5, lane = t & 31;
    const int gid = lane >> 2, tig = lane & 3;     // mma fragment coordinates
    const int gbase = A.g0 + tile * C3_GT;
    const int gcount = min(C3_GT, A.g0 + A.gn - gbase);
    const int64_t SS = (int64_t)S * S;
    const int64_t NS = (int64_t)d.N * S;
    const int sgn = dir == 0 ? 1 : -1;
    const int jbase = w * (8 * NBW);
    const bool fit_ok = gid < gcount;
    const int gfit = gbase + min(gid, gcount - 1);
    // C fragment = fit gid, columns j0[h], j0[h] + 1 of column block h
    int upos[NBW];
    bool ok0[NBW], ok1[NBW];
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        const int j0 = jbase + 8 * h + 2 * tig;
        ok0[h] = fit_ok && j0 < S;
        ok1[h] = fit_ok && j0 + 1 < S;
        upos[h] = c3_pos(gid, j0);                 // j0 even: j0, j0+1 adjacent, 16-byte aligned
    }

    for (int e = t; e < 2 * KB * 32; e += T) u[e] = 0.;
    for (int e = t; e < 2 * C3_GT * C3_NW; e += T) redi[e] = ((e % C3_NW) == 0) ? 1023 : 0;   // first message: max f = 1
    for (int e = t; e < S * M; e += T) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(A.tab[(s * M + m) * 2] + A.tab[(s * M + m) * 2 + 1]);
    }
    for (int e = t; e < bct_n; e += T) bct[e] = A.bc[e];
    for (int e = t; e < BC * BC; e += T) didx[e] = A.didx[e];
    __syncthreads();

    const int nsteps = s1 - s0;
    const int seg_first = dir == 0 ? s0 : s1;
    // this thread's row pointers: f source and message destination at its first column, advanced one row per segment
    const double* fp = A.fexp + (int64_t)gfit * NS + (int64_t)seg_first * S + jbase + 2 * tig;
    double* outp = (dir == 0 ? A.ahat : A.bhat) + (int64_t)gfit * NS + (int64_t)seg_first * S + jbase + 2 * tig;
    const int rowstep = sgn * S;

    // ---- first segment of the sweep: u = f (forward: ahat = f; backward: bhat = 1)
    double un[NBW][2];
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        un[h][0] = un[h][1] = 0.;
        if (ok0[h]) { un[h][0] = fp[8 * h]; outp[8 * h] = dir == 0 ? un[h][0] : 1.; }
        if (ok1[h]) { un[h][1] = fp[8 * h + 1]; outp[8 * h + 1] = dir == 0 ? un[h][1] : 1.; }
        if (ok0[h]) *reinterpret_cast<double2*>(u + upos[h]) = make_double2(un[h][0], un[h][1]);
    }
    auto row_of = [&](int k) { return dir == 0 ? s0 + k : s1 - 1 - k; };       // transition row of step k
    // software pipeline: f C3_FST - 1 segments ahead through the shared-memory ring, row info two rows ahead
    const unsigned fring_s = (unsigned)__cvta_generic_to_shared(fring);
    const double* fsrc = fp + rowstep;        // f row of step 0's segment; advanced by f_issue
    int f_next = 0;
    auto f_issue = [&]() {
        if (f_next < nsteps) {
            const unsigned base = fring_s + 8u * (unsigned)(((f_next % C3_FST) * NSLOT) * T);
#pragma unroll
            for (int h = 0; h < NBW; ++h) {
                if (ok0[h]) c3_cp8(base + 8u * (unsigned)((h * 2) * T), fsrc + 8 * h);
                if (ok1[h]) c3_cp8(base + 8u * (unsigned)((h * 2 + 1) * T), fsrc + 8 * h + 1);
            }
            fsrc += rowstep;
        }
        ++f_next;
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int e = 0; e < C3_FST * NSLOT; ++e) fring[e * T] = 0.;     // entries outside the state space stay zero
#pragma unroll
    for (int k = 0; k < C3_FST - 1; ++k) f_issue();
    const bool ri_sm = nsteps <= C3_RIMAX;
    if (ri_sm)
        for (int k = t; k < nsteps; k += T) rism[k] = A.rinfo[row_of(k)];
    __syncthreads();
    auto ri_at = [&](int k) { return ri_sm ? rism[k] : A.rinfo[row_of(k)]; };
    int ri0 = nsteps > 0 ? ri_at(0) : 0, ri1 = nsteps > 1 ? ri_at(1) : 0;

    double ereg[NBW][KR];       // B fragments of row blocks 0 .. KR-1; row blocks KR .. KB-1 in esm[(h * KO + kb - KR) * T]
    int cur_class = -1;
    // offsets of this thread's output columns' total classes into didx (breakend rows): column index forward, row
    // index backward
    int cjo[NBW][2];
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        const int j0 = jbase + 8 * h + 2 * tig;
        cjo[h][0] = (j0 < S ? scls[j0] : 0) * (dir == 0 ? 1 : BC);
        cjo[h][1] = (j0 + 1 < S ? scls[j0 + 1] : 0) * (dir == 0 ? 1 : BC);
    }
    int cur = 0, esum = 0, bad = 0;

    for (int step = 0; step < nsteps; ++step) {
        outp += rowstep;
        const int ri = ri0;
        ri0 = ri1;
        const int cls = (ri & 0xff) - 1;
        const int br = (ri >> 12) - 1;
        const double* uc = u + cur * KB * 32;
        double c[NBW][2], f[NBW][2];
        int be = 1023;
#pragma unroll
        for (int h = 0; h < NBW; ++h) c[h][0] = c[h][1] = f[h][0] = f[h][1] = 0.;

        // bookkeeping of the segment, placed inside the DMMA block by the callers below
        auto chores = [&](int part) {
            if (part == 0) {
                // exponent of the largest entry of the input message of fit gid (written before the last barrier)
                const int4* rp = reinterpret_cast<const int4*>(redi + (cur * C3_GT + gid) * C3_NW);
                const int4 r0 = rp[0], r1 = rp[1];
                be = max(max(max(r0.x, r0.y), max(r0.z, r0.w)), max(max(r1.x, r1.y), max(r1.z, r1.w)));
            } else if (part == 1) {
                asm volatile("cp.async.wait_group %0;" ::"n"(C3_FST - 2) : "memory");     // this step's f has landed
                const double* fr = fring + ((step % C3_FST) * NSLOT) * T;
#pragma unroll
                for (int h = 0; h < NBW; ++h) { f[h][0] = fr[(h * 2) * T]; f[h][1] = fr[(h * 2 + 1) * T]; }
            } else {
                f_issue();          // into the ring stage the previous step consumed
                if (step + 2 < nsteps) ri1 = ri_at(step + 2);
            }
        };

        int mx = 0;
        // epilogue of one column block (C-fragment layout): scale by the power of two that normalised the input message,
        // times f, store the message and pass it on in fragment order
        auto finish_block = [&](int h, double c0, double c1) {
            const double sc = c3_pow2(be);
            double v0 = c0 * sc, v1 = c1 * sc;
            if (dir == 0) {
                v0 *= f[h][0]; v1 *= f[h][1];
                un[h][0] = v0; un[h][1] = v1;
            } else {
                un[h][0] = v0 * f[h][0]; un[h][1] = v1 * f[h][1];
            }
            if (!ok0[h]) un[h][0] = 0.;
            if (!ok1[h]) un[h][1] = 0.;
            if (ok0[h]) outp[8 * h] = v0;
            if (ok1[h]) outp[8 * h + 1] = v1;
            if (ok0[h]) *reinterpret_cast<double2*>(u + (cur ^ 1) * KB * 32 + upos[h]) = make_double2(un[h][0], un[h][1]);
            mx = max(mx, max(c3_expfield(un[h][0]), c3_expfield(un[h][1])));
        };

        if (br < 0) {
            if (cls != cur_class) {
                const double* E = A.sp.Emat + (int64_t)cls * SS;
#pragma unroll
                for (int h = 0; h < NBW; ++h) {
                    const int jb = jbase + 8 * h + gid;
#pragma unroll
                    for (int kb = 0; kb < KR; ++kb) {
                        const int i = kb * 4 + tig;
                        double ev = (i < S && jb < S) ? E[(int64_t)i * S + jb] : 0.;
                        asm volatile("" : "+d"(ev));     // opaque: keep the value, do not re-load it every segment
                        ereg[h][kb] = ev;
                    }
                    for (int kb = KR; kb < KB; ++kb) {
                        const int i = kb * 4 + tig;
                        esm[(h * KO + kb - KR) * T] = (i < S && jb < S) ? E[(int64_t)i * S + jb] : 0.;
                    }
                }
                cur_class = cls;
            }
            // row blocks outermost: one A fragment (one shared-memory load instead of NBW) feeds the DMMAs of all NBW column
            // blocks, and consecutive DMMAs belong to different accumulators by construction (chain_mma2.cu: ptxas had
            // re-ordered a block's even / odd row blocks into two dependent chains).  Every accumulator sums the row blocks
            // it summed before: same values bit for bit.
            double acc[NBW][2][2];
#pragma unroll
            for (int h = 0; h < NBW; ++h) acc[h][0][0] = acc[h][0][1] = acc[h][1][0] = acc[h][1][1] = 0.;
#pragma unroll
            for (int kb = 0; kb < KB; ++kb) {
                const double a = uc[kb * 32 + lane];
#pragma unroll
                for (int h = 0; h < NBW; ++h) {
                    const double bf = kb < KR ? ereg[h][kb < KR ? kb : 0] : esm[(h * KO + (kb < KR ? 0 : kb - KR)) * T];
                    c3_dmma(acc[h][kb & 1][0], acc[h][kb & 1][1], a, bf);
                }
                if (kb == 2) chores(0);
                if (kb == 6) chores(1);
                if (kb == 12) chores(2);
            }
#pragma unroll
            for (int h = 0; h < NBW; ++h) finish_block(h, acc[h][0][0] + acc[h][1][0], acc[h][0][1] + acc[h][1][1]);
        } else {
            // ---- breakend row: per-fit matrix  E_g[i][j] = flipE[i][j] * prod_m wexp[g][m][tot_i[m] - tot_j[m]].
            // The per-fit factor depends on (i, j) only through the states' tumour-clone totals.  With the states grouped
            // by those totals into BC classes, each class padded to whole row blocks of four,
            //     out[g][j] = w0[g] * sum_ci  W[g][ci - class(j)] * ( sum_{i in ci} u[g][i] flipE[i][j] ),
            // and the inner sums are ONE DMMA per (row block, column block) with the shared operand flipE (see
            // chain_mma2.cu); the flipE fragments come from global memory (the class matrices stay in L2).
            chores(0);
            chores(1);
            chores(2);
            for (int e = t; e < C3_GT * MD; e += T) {
                const int g = e / MD, o = e - g * MD;
                const int gg = gbase + min(g, gcount - 1);
                wtab[e] = A.sp.wexp[(int64_t)gg * A.sp_fit_stride + (int64_t)br * MD + o];
            }
            const double* FE = A.sp.flipE + (int64_t)cls * SS;
            __syncthreads();
            for (int e = t; e < C3_GT * ND; e += T) {     // the factor of every distinct tuple of total differences
                const int g = e / ND, q = e - g * ND;
                double v = 1.;
                for (int m = 1; m < M; ++m) v *= wtab[g * MD + m * d.D + ((int)ddel[q * M + m] - 128) + d.doff];
                Wd[e] = v;
            }
            __syncthreads();
            const int d0 = ((ri >> 8) & 15) - 8 + d.doff;     // the normal-clone factor does not depend on the states
            const double w0 = wtab[gid * MD + d0];
            const double* wd = Wd + gid * ND;
            // difference tuple of (input class ci, output class cj): forward ci - cj, backward cj - ci
            const int dstr = dir == 0 ? BC : 1;
            double o[NBW][2];
#pragma unroll
            for (int h = 0; h < NBW; ++h) o[h][0] = o[h][1] = 0.;
            const double* ug = uc + gid * 4;
#pragma unroll 2
            for (int kb = 0; kb < KBB; ++kb) {
                const int i = bperm[4 * kb + tig];
                const bool iok = i < S;
                const double a = ug[iok ? ((i >> 2) << 5) + (i & 3) : 0];      // padded rows have a zero B fragment
                const uint16_t* dr = didx + (int)kbcls[kb] * dstr;
                const double* fer = FE + (int64_t)(iok ? i : 0) * S + jbase + gid;
#pragma unroll
                for (int h = 0; h < NBW; ++h) {
                    const double bf = (iok && jbase + 8 * h + gid < S) ? __ldg(fer + 8 * h) : 0.;
                    const double w0h = wd[dr[cjo[h][0]]], w1h = wd[dr[cjo[h][1]]];
                    double p0 = 0., p1 = 0.;
                    c3_dmma(p0, p1, a, bf);
                    o[h][0] = fma(p0, w0h, o[h][0]);
                    o[h][1] = fma(p1, w1h, o[h][1]);
                }
            }
#pragma unroll
            for (int h = 0; h < NBW; ++h) { c[h][0] = o[h][0] * w0; c[h][1] = o[h][1] * w0; }
        }
        // ---- what is left of the step
        esum += be - 1023;
        if (be <= 0 || be >= 2047) bad = 1;
        if (br >= 0) {
#pragma unroll
            for (int h = 0; h < NBW; ++h) finish_block(h, c[h][0], c[h][1]);
        }
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        if (tig == 0) redi[((cur ^ 1) * C3_GT + gid) * C3_NW + w] = mx;
        cur ^= 1;
        __syncthreads();
    }

    // ---- log normaliser of the forward sweep: log(sum of the last message) + ln2 * esum + sum_n max F[n]
    {
        double sv = 0.;
#pragma unroll
        for (int h = 0; h < NBW; ++h) sv += un[h][0] + un[h][1];
        sv += __shfl_xor_sync(0xffffffffu, sv, 1);
        sv += __shfl_xor_sync(0xffffffffu, sv, 2);
        if (tig == 0) redd[gid * C3_NW + w] = sv;
        for (int g = w; g < gcount; g += NW) {
            const double* mf = A.mF + (int64_t)(gbase + g) * d.N;
            double ms = 0.;
            for (int n = s0 + lane; n <= s1; n += 32) ms += mf[n];
            ms = warp_sum(ms);
            if (lane == 0) redd[C3_GT * C3_NW + g] = ms;
        }
    }
    __syncthreads();
    if (w == 0 && tig == 0 && fit_ok) {
        double slast = 0.;
        for (int c2 = 0; c2 < NW; ++c2) slast += redd[gid * C3_NW + c2];
        if (dir == 0) {
            const double logz = log(slast) + 0.693147180559945309417 * (double)esum + redd[C3_GT * C3_NW + gid];
            A.chain_logz[(int64_t)(gbase + gid) * A.ch.NC + chain] = logz;
            if (isnan(logz)) bad = 1;
        }
        if (!(slast > 0.) || isinf(slast)) bad = 1;
        if (bad) atomicOr(A.dflags, RMX_F_MSG_NAN);
    }
}

template <int KB, int NW, int NBW, int KR>
static int launch_chain_mma3_cfg(cudaStream_t st, Chain3Args& A) {
    const RmxDims& d = A.d;
    const C3Layout L = c3_layout(KB, NW * 32, 2 * NBW, NBW * (KB - KR), d.S, d.M, d.M * d.D, A.BC, A.KBB, A.ND);
    auto kern = k_chain_mma3<KB, NW, NBW, KR>;
    if (L.total > 227 * 1024) return 2;
    if (L.total > 48 * 1024) {
        static RmxPerDeviceOnce once;
        static size_t set_for = 0;
        if (once.need() || set_for < L.total) {
            if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) return 1;
            set_for = 227 * 1024;
        }
    }
    A.ntiles = (A.gn + C3_GT - 1) / C3_GT;
    kern<<<A.ch.NC * 2 * A.ntiles, NW * 32, L.total, st>>>(A);
    return cudaPeekAtLastError() == cudaSuccess ? 0 : 1;
}

// returns 0 on success, 1 if the configuration or the launch failed, 2 if this kernel does not cover the shape
int rmx_launch_chain_mma3(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                          const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                          const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags,
                          const uint8_t* bc, const uint16_t* didx, int BC, int KBB, int ND) {
    if (ch.NC <= 0 || gn <= 0) return 0;
    if (rinfo == nullptr || bc == nullptr || didx == nullptr || BC <= 0 || ND <= 0 || ND > 1024) return 2;
    Chain3Args A;
    A.d = d; A.ch = ch; A.rinfo = rinfo; A.sp = sp; A.sp_fit_stride = sp_fit_stride; A.tab = T.tab;
    A.g0 = g0; A.gn = gn; A.ntiles = 1; A.fexp = fexp; A.mF = mF; A.ahat = ahat; A.bhat = bhat;
    A.chain_logz = chain_logz; A.dflags = dflags;
    A.bc = bc; A.didx = didx; A.BC = BC; A.KBB = KBB; A.ND = ND;
    A.cord = nullptr; A.cstart = nullptr;
    if (d.S > 128 && d.S <= 168) return launch_chain_mma3_cfg<42, 7, 3, 24>(st, A);     // max copy number 8, M = 3: S = 165
    return 2;
}

// ---------------------------------------------------------------------------------------------------------------------
// K2 for 169 .. 360 states (the reference's DEFAULT max copy number 12: S = 355 with two tumour clones): the class
// matrix is 1 MB and lives in L2 only, but it is still a tensor-path product.  The L2-streamed kernel this replaces
// (chain.cu) serves FOUR fits per matrix read with scalar FMAs; here a CTA serves EIGHT fits per read: eight DMMA warps
// own six column blocks each (48 x 8 = 384 columns) and load their B fragments (8 bytes per lane, 32-byte sectors fully
// used) straight from L2 under the DMMA stream -- no transition operand on chip at all, 540 loads + 540 DMMAs per warp
// and segment.  Messages, scaling, f ring and bookkeeping as in k_chain_mma3.  Breakend rows (1.8 % of rows) take a
// scalar path: every thread sums its own two columns per block over all S input states.
// ---------------------------------------------------------------------------------------------------------------------
struct C3sLayout {
    size_t u, wtab, redd, redi, tots, fring, rism, total;
};
__host__ __device__ inline C3sLayout c3s_layout(int KB, int T, int NSLOT, int S, int M, int MD) {
    C3sLayout L;
    size_t o = 0;
    L.u = o;     o = c3_al(o + sizeof(double) * 2 * KB * 32);
    L.wtab = o;  o = c3_al(o + sizeof(double) * C3_GT * MD);
    L.redd = o;  o = c3_al(o + sizeof(double) * (C3_GT * C3_NW + C3_GT));
    L.redi = o;  o = c3_al(o + sizeof(int) * 2 * C3_GT * C3_NW);
    L.tots = o;  o = c3_al(o + (size_t)S * M);
    L.fring = o; o = c3_al(o + sizeof(double) * C3_FST * NSLOT * T);
    L.rism = o;  o = c3_al(o + sizeof(int) * C3_RIMAX);
    L.total = o;
    return L;
}

template <int KB, int NW, int NBW>
__global__ void __launch_bounds__(NW * 32, 1) k_chain_mma_s(Chain3Args A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int T = NW * 32, NSLOT = 2 * NBW;
    const RmxDims& d = A.d;
    const int S = d.S, M = d.M, MD = M * d.D;
    const C3sLayout L = c3s_layout(KB, T, NSLOT, S, M, MD);
    double* u = reinterpret_cast<double*>(smem_raw + L.u);             // [2][KB*32]
    double* wtab = reinterpret_cast<double*>(smem_raw + L.wtab);       // [8][MD]   breakend tables of the step
    double* redd = reinterpret_cast<double*>(smem_raw + L.redd);       // [8][C3_NW] + [8] mF sums
    int* redi = reinterpret_cast<int*>(smem_raw + L.redi);             // [2][8][C3_NW] max exponent fields
    int8_t* tots = reinterpret_cast<int8_t*>(smem_raw + L.tots);       // [S*M]
    double* fring = reinterpret_cast<double*>(smem_raw + L.fring) + threadIdx.x;   // [C3_FST][NSLOT][T], own column
    int* rism = reinterpret_cast<int*>(smem_raw + L.rism);             // [C3_RIMAX] row info of the chain

    const int tile = blockIdx.x % A.ntiles;
    const int cd = blockIdx.x / A.ntiles;
    const int dir = cd & 1, chain = A.ch.order[cd >> 1];
    const int s0 = A.ch.start[chain], s1 = A.ch.end[chain];
    const int t = threadIdx.x, w = t >> 5, lane = t & 31;
    const int gid = lane >> 2, tig = lane & 3;     // mma fragment coordinates
    const int gbase = A.g0 + tile * C3_GT;
    const int gcount = min(C3_GT, A.g0 + A.gn - gbase);
    const int64_t SS = (int64_t)S * S;
    const int64_t NS = (int64_t)d.N * S;
    const int sgn = dir == 0 ? 1 : -1;
    const int jbase = w * (8 * NBW);
    const bool fit_ok = gid < gcount;
    const int gfit = gbase + min(gid, gcount - 1);
    int upos[NBW];
    bool ok0[NBW], ok1[NBW];
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        const int j0 = jbase + 8 * h + 2 * tig;
        ok0[h] = fit_ok && j0 < S;
        ok1[h] = fit_ok && j0 + 1 < S;
        upos[h] = c3_pos(gid, j0);
    }
    // column blocks of this warp inside the state space (warp-uniform): the others issue nothing
    int nbw = 0;
#pragma unroll
    for (int h = 0; h < NBW; ++h) nbw += (jbase + 8 * h < S) ? 1 : 0;

    for (int e = t; e < 2 * KB * 32; e += T) u[e] = 0.;
    for (int e = t; e < 2 * C3_GT * C3_NW; e += T) redi[e] = ((e % C3_NW) == 0) ? 1023 : 0;   // first message: max f = 1
    for (int e = t; e < S * M; e += T) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(A.tab[(s * M + m) * 2] + A.tab[(s * M + m) * 2 + 1]);
    }
    __syncthreads();

    const int nsteps = s1 - s0;
    const int seg_first = dir == 0 ? s0 : s1;
    const double* fp = A.fexp + (int64_t)gfit * NS + (int64_t)seg_first * S + jbase + 2 * tig;
    double* outp = (dir == 0 ? A.ahat : A.bhat) + (int64_t)gfit * NS + (int64_t)seg_first * S + jbase + 2 * tig;
    const int rowstep = sgn * S;

    double un[NBW][2];
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        un[h][0] = un[h][1] = 0.;
        if (ok0[h]) { un[h][0] = fp[8 * h]; outp[8 * h] = dir == 0 ? un[h][0] : 1.; }
        if (ok1[h]) { un[h][1] = fp[8 * h + 1]; outp[8 * h + 1] = dir == 0 ? un[h][1] : 1.; }
        if (ok0[h]) *reinterpret_cast<double2*>(u + upos[h]) = make_double2(un[h][0], un[h][1]);
    }
    auto row_of = [&](int k) { return dir == 0 ? s0 + k : s1 - 1 - k; };
    const unsigned fring_s = (unsigned)__cvta_generic_to_shared(fring);
    const double* fsrc = fp + rowstep;
    int f_next = 0;
    auto f_issue = [&]() {
        if (f_next < nsteps) {
            const unsigned base = fring_s + 8u * (unsigned)(((f_next % C3_FST) * NSLOT) * T);
#pragma unroll
            for (int h = 0; h < NBW; ++h) {
                if (ok0[h]) c3_cp8(base + 8u * (unsigned)((h * 2) * T), fsrc + 8 * h);
                if (ok1[h]) c3_cp8(base + 8u * (unsigned)((h * 2 + 1) * T), fsrc + 8 * h + 1);
            }
            fsrc += rowstep;
        }
        ++f_next;
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int e = 0; e < C3_FST * NSLOT; ++e) fring[e * T] = 0.;
#pragma unroll
    for (int k = 0; k < C3_FST - 1; ++k) f_issue();
    const bool ri_sm = nsteps <= C3_RIMAX;
    if (ri_sm)
        for (int k = t; k < nsteps; k += T) rism[k] = A.rinfo[row_of(k)];
    __syncthreads();
    auto ri_at = [&](int k) { return ri_sm ? rism[k] : A.rinfo[row_of(k)]; };
    int ri0 = nsteps > 0 ? ri_at(0) : 0, ri1 = nsteps > 1 ? ri_at(1) : 0;
    int cur = 0, esum = 0, bad = 0;

    for (int step = 0; step < nsteps; ++step) {
        outp += rowstep;
        const int ri = ri0;
        ri0 = ri1;
        const int cls = (ri & 0xff) - 1;
        const int br = (ri >> 12) - 1;
        const double* uc = u + cur * KB * 32;
        double c[NBW][2], f[NBW][2];
        int be = 1023;
#pragma unroll
        for (int h = 0; h < NBW; ++h) c[h][0] = c[h][1] = f[h][0] = f[h][1] = 0.;
        {   // bookkeeping of the segment (the loads below hide it)
            const int4* rp = reinterpret_cast<const int4*>(redi + (cur * C3_GT + gid) * C3_NW);
            const int4 r0 = rp[0], r1 = rp[1];
            be = max(max(max(r0.x, r0.y), max(r0.z, r0.w)), max(max(r1.x, r1.y), max(r1.z, r1.w)));
        }
        if (br < 0) {
            // B fragment of (row block kb, column block h): E[4 kb + tig][jbase + 8 h + gid].  Row blocks outermost: the A
            // fragment is read once per row block, the six B fragments sit at constant offsets from ONE pointer that
            // advances four rows per block -- 2.5 instructions per DMMA (the first version of this loop spent 18 on
            // predicates and 64-bit index products and was issue-bound at 53 k cycles per segment against the DMMA
            // floor of 17 k; ncu: 10 k warp instructions per warp and segment, issue slots 36 % busy).  Columns beyond
            // the state space are read (inside the allocation, padded) and feed output columns that are never stored;
            // rows beyond it exist in the last row block only and are zeroed there.
            const double* Ep = A.sp.Emat + (int64_t)cls * SS + (int64_t)tig * S + jbase + gid;
            const int kfull = S >> 2;
            // register pipeline: the loads of block kb + 8 are issued before the DMMAs of block kb (left to itself the
            // compiler put each load six instructions ahead of its DMMA: 46 % of all stall samples long-scoreboard)
            double q0[NBW], q1[NBW], q2[NBW], q3[NBW], q4[NBW], q5[NBW], q6[NBW], q7[NBW];
            auto ldblk = [&](double* q, const double* p) {
#pragma unroll
                for (int h = 0; h < NBW; ++h) q[h] = (h < nbw) ? __ldg(p + 8 * h) : 0.;
            };
            auto mmablk = [&](const double* q, int kb) {
                const double a = uc[kb * 32 + lane];
#pragma unroll
                for (int h = 0; h < NBW; ++h)
                    if (h < nbw) c3_dmma(c[h][0], c[h][1], a, q[h]);
            };
            const int64_t rs = 4 * (int64_t)S;
            // eight row blocks deep (four were not enough: ncu put 27 % of all stall samples on the first DMMA of a
            // block waiting for its operand -- under this load an L2 hit outlasts four blocks of DMMAs, ~770 cycles)
            const int k4 = kfull & ~7;
            if (k4 > 0) {
                ldblk(q0, Ep); ldblk(q1, Ep + rs); ldblk(q2, Ep + 2 * rs); ldblk(q3, Ep + 3 * rs);
                ldblk(q4, Ep + 4 * rs); ldblk(q5, Ep + 5 * rs); ldblk(q6, Ep + 6 * rs); ldblk(q7, Ep + 7 * rs);
                const double* En = Ep + 8 * rs;
                for (int kb = 0; kb < k4; kb += 8) {
                    const bool more = kb + 8 < k4;
                    mmablk(q0, kb);
                    if (more) ldblk(q0, En);
                    mmablk(q1, kb + 1);
                    if (more) ldblk(q1, En + rs);
                    mmablk(q2, kb + 2);
                    if (more) ldblk(q2, En + 2 * rs);
                    mmablk(q3, kb + 3);
                    if (more) ldblk(q3, En + 3 * rs);
                    mmablk(q4, kb + 4);
                    if (more) ldblk(q4, En + 4 * rs);
                    mmablk(q5, kb + 5);
                    if (more) ldblk(q5, En + 5 * rs);
                    mmablk(q6, kb + 6);
                    if (more) ldblk(q6, En + 6 * rs);
                    mmablk(q7, kb + 7);
                    if (more) ldblk(q7, En + 7 * rs);
                    En += 8 * rs;
                }
            }
            Ep += (int64_t)k4 * rs;
            for (int kb = k4; kb < kfull; ++kb) {
                ldblk(q0, Ep);
                mmablk(q0, kb);
                Ep += rs;
            }
            if (kfull * 4 < S) {
                const double a = uc[kfull * 32 + lane];
                const bool rin = kfull * 4 + tig < S;
#pragma unroll
                for (int h = 0; h < NBW; ++h)
                    if (h < nbw) c3_dmma(c[h][0], c[h][1], a, rin ? __ldg(Ep + 8 * h) : 0.);
            }
        } else {
            // ---- breakend row, scalar: out[g][j] = w0 sum_i u[g][i] flipE[i][j] prod_m wexp[g][m][tot_i[m] - tot_j[m]]
            // (backward sweep: the summed state is the LATER one, so the difference is tot_j[m] - tot_i[m])
            for (int e = t; e < C3_GT * MD; e += T) {
                const int g = e / MD, o = e - g * MD;
                const int gg = gbase + min(g, gcount - 1);
                wtab[e] = A.sp.wexp[(int64_t)gg * A.sp_fit_stride + (int64_t)br * MD + o];
            }
            __syncthreads();
            const double* FE = A.sp.flipE + (int64_t)cls * SS;
            const int d0 = ((ri >> 8) & 15) - 8 + d.doff;
            const double* wg = wtab + gid * MD;
            const double w0 = wg[d0];
            const double* ug = uc + gid * 4;
            // Rows in class order (states grouped by the tuple of their tumour-clone totals: A.cord / A.cstart, the
            // tables of k_hist4): the per-fit factor is constant within a class, so a class costs its
            // sum_i u[g][i] flipE[i][j] -- one L2 load and one FMA per pair -- and ONE factor per output column.
            // (Per pair with two table look-ups each, this path was 25 % of the launch for 1.8 % of the rows; the
            // twelve columns of a thread side by side were slower still: eight per-fit tables hit at once.)
            const uint16_t* co = A.cord;
            const int* cs = A.cstart;
            const int NCLS = cs[0];
            int bj[NBW][2][RMX_MAXM];
            const double* fcol[NBW][2];
#pragma unroll
            for (int h = 0; h < NBW; ++h)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int j = min(jbase + 8 * h + 2 * tig + q, S - 1);      // columns outside: a valid address, never stored
                    fcol[h][q] = FE + j;
#pragma unroll
                    for (int m = 1; m < RMX_MAXM; ++m) bj[h][q][m] = (m < M) ? m * d.D + d.doff - sgn * (int)tots[j * M + m] : 0;
                }
            double o[NBW][2];
#pragma unroll
            for (int h = 0; h < NBW; ++h) o[h][0] = o[h][1] = 0.;
            for (int cl = 0; cl < NCLS; ++cl) {
                double acc[NBW][2];
#pragma unroll
                for (int h = 0; h < NBW; ++h) acc[h][0] = acc[h][1] = 0.;
                const int k0 = __ldg(cs + cl + 1), k1 = __ldg(cs + cl + 2);
                for (int k = k0; k < k1; ++k) {
                    const int i = __ldg(co + k);
                    const double ui = ug[((i >> 2) << 5) + (i & 3)];
                    const int64_t ro = (int64_t)i * S;
#pragma unroll
                    for (int h = 0; h < NBW; ++h)
                        if (h < nbw) {
                            acc[h][0] = fma(ui, __ldg(fcol[h][0] + ro), acc[h][0]);
                            acc[h][1] = fma(ui, __ldg(fcol[h][1] + ro), acc[h][1]);
                        }
                }
                const int irep = __ldg(co + k0);
                int ti[RMX_MAXM];
#pragma unroll
                for (int m = 1; m < RMX_MAXM; ++m) ti[m] = (m < M) ? sgn * (int)tots[irep * M + m] : 0;
#pragma unroll
                for (int h = 0; h < NBW; ++h)
                    if (h < nbw) {
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
                            double e = acc[h][q];
#pragma unroll
                            for (int m = 1; m < RMX_MAXM; ++m)
                                if (m < M) e *= wg[bj[h][q][m] + ti[m]];
                            o[h][q] += e;
                        }
                    }
            }
#pragma unroll
            for (int h = 0; h < NBW; ++h) { c[h][0] = o[h][0] * w0; c[h][1] = o[h][1] * w0; }
        }
        // ---- the segment's f (landed long ago), the next copies, the next row descriptor
        asm volatile("cp.async.wait_group %0;" ::"n"(C3_FST - 2) : "memory");
        {
            const double* fr = fring + ((step % C3_FST) * NSLOT) * T;
#pragma unroll
            for (int h = 0; h < NBW; ++h) { f[h][0] = fr[(h * 2) * T]; f[h][1] = fr[(h * 2 + 1) * T]; }
        }
        f_issue();
        if (step + 2 < nsteps) ri1 = ri_at(step + 2);
        esum += be - 1023;
        if (be <= 0 || be >= 2047) bad = 1;
        int mx = 0;
        const double sc = c3_pow2(be);
#pragma unroll
        for (int h = 0; h < NBW; ++h) {
            double v0 = c[h][0] * sc, v1 = c[h][1] * sc;
            if (dir == 0) {
                v0 *= f[h][0]; v1 *= f[h][1];
                un[h][0] = v0; un[h][1] = v1;
            } else {
                un[h][0] = v0 * f[h][0]; un[h][1] = v1 * f[h][1];
            }
            if (!ok0[h]) un[h][0] = 0.;
            if (!ok1[h]) un[h][1] = 0.;
            if (ok0[h]) outp[8 * h] = v0;
            if (ok1[h]) outp[8 * h + 1] = v1;
            if (ok0[h]) *reinterpret_cast<double2*>(u + (cur ^ 1) * KB * 32 + upos[h]) = make_double2(un[h][0], un[h][1]);
            mx = max(mx, max(c3_expfield(un[h][0]), c3_expfield(un[h][1])));
        }
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        if (tig == 0) redi[((cur ^ 1) * C3_GT + gid) * C3_NW + w] = mx;
        cur ^= 1;
        __syncthreads();
    }

    {
        double sv = 0.;
#pragma unroll
        for (int h = 0; h < NBW; ++h) sv += un[h][0] + un[h][1];
        sv += __shfl_xor_sync(0xffffffffu, sv, 1);
        sv += __shfl_xor_sync(0xffffffffu, sv, 2);
        if (tig == 0) redd[gid * C3_NW + w] = sv;
        for (int g = w; g < gcount; g += NW) {
            const double* mf = A.mF + (int64_t)(gbase + g) * d.N;
            double ms = 0.;
            for (int n = s0 + lane; n <= s1; n += 32) ms += mf[n];
            ms = warp_sum(ms);
            if (lane == 0) redd[C3_GT * C3_NW + g] = ms;
        }
    }
    __syncthreads();
    if (w == 0 && tig == 0 && fit_ok) {
        double slast = 0.;
        for (int c2 = 0; c2 < NW; ++c2) slast += redd[gid * C3_NW + c2];
        if (dir == 0) {
            const double logz = log(slast) + 0.693147180559945309417 * (double)esum + redd[C3_GT * C3_NW + gid];
            A.chain_logz[(int64_t)(gbase + gid) * A.ch.NC + chain] = logz;
            if (isnan(logz)) bad = 1;
        }
        if (!(slast > 0.) || isinf(slast)) bad = 1;
        if (bad) atomicOr(A.dflags, RMX_F_MSG_NAN);
    }
}

template <int KB, int NW, int NBW>
static int launch_chain_mma_s_cfg(cudaStream_t st, Chain3Args& A) {
    const RmxDims& d = A.d;
    const C3sLayout L = c3s_layout(KB, NW * 32, 2 * NBW, d.S, d.M, d.M * d.D);
    auto kern = k_chain_mma_s<KB, NW, NBW>;
    if (L.total > 227 * 1024) return 2;
    if (L.total > 48 * 1024) {
        static RmxPerDeviceOnce once;
        if (once.need())
            if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) return 1;
    }
    A.ntiles = (A.gn + C3_GT - 1) / C3_GT;
    kern<<<A.ch.NC * 2 * A.ntiles, NW * 32, L.total, st>>>(A);
    return cudaPeekAtLastError() == cudaSuccess ? 0 : 1;
}

// returns 0 on success, 1 if the configuration or the launch failed, 2 if this kernel does not cover the shape
int rmx_launch_chain_mma_s(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                           const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                           const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags,
                           const uint16_t* cord, const int* cstart) {
    if (ch.NC <= 0 || gn <= 0) return 0;
    if (rinfo == nullptr || d.M > RMX_MAXM || cord == nullptr || cstart == nullptr) return 2;
    Chain3Args A;
    A.d = d; A.ch = ch; A.rinfo = rinfo; A.sp = sp; A.sp_fit_stride = sp_fit_stride; A.tab = T.tab;
    A.g0 = g0; A.gn = gn; A.ntiles = 1; A.fexp = fexp; A.mF = mF; A.ahat = ahat; A.bhat = bhat;
    A.chain_logz = chain_logz; A.dflags = dflags;
    A.bc = nullptr; A.didx = nullptr; A.BC = 0; A.KBB = 0; A.ND = 0;
    A.cord = cord; A.cstart = cstart;
    if (d.S > 168 && d.S <= 360) return launch_chain_mma_s_cfg<90, 8, 6>(st, A);     // max copy number 9 .. 12, M = 3
    return 2;
}
