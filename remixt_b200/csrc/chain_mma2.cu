// K2, second generation (S <= 97): forward / backward sum-product messages of EIGHT fits of an initialisation grid in
// lock step on the FP64 tensor path (mma.sync.m8n8k4.f64), one block barrier per segment.
//                                                                              (bpmodel.pyx:1213-1246, :921-946)
//
// What the first generation (chain_mma.cu) measured (profiles/r02_chain.md): of its 3640 cycles per segment only 1300
// are DMMA issue.  Seven warps own two column blocks each (13 blocks at S = 97), so two of the four SM sub-partitions
// queue 100 DMMAs (1600 cycles at 16 per m8n8k4) while one queues 50; and all the per-segment bookkeeping -- the
// scale exponent, the f prefetch, row descriptors, ~250 integer instructions per warp -- runs BEFORE or AFTER the
// DMMA block of a warp that issues one DMMA every 16 cycles and is otherwise idle.  Dependent DMMAs are not the limit
// (26 cycles back to back, scripts/micro/dmma_latency.cu).
//
// Here:
//   * four DMMA warps, one per sub-partition, own NBW column blocks each (3 at S = 97: 75 DMMAs = 1200 cycles, the same
//     on every sub-partition); the transition matrix is NBW * KB B-fragments per thread, all in registers;
//   * S = 32 NBW + 1 (97 = 96 + 1): the last state's column is a 25-term dot product per fit on the vector pipe, done by
//     a fifth warp next to the DMMA warps instead of a 13th column block that would put 100 DMMAs on one sub-partition;
//   * the bookkeeping of a segment sits INSIDE its DMMA block (the warp has 15 free issue slots per DMMA): reading the
//     previous message's exponent, the f values of this segment from the cp.async ring, issuing the ring copy of
//     segment + 3, the next row descriptor.  What remains after the last DMMA is the dependent tail only: add the two
//     accumulators, scale, multiply by f, store, exponent field, two shuffles, barrier;
//   * pointers advance by one row per segment instead of 64-bit index products.
// The arithmetic (scaling by exact powers of two, order of the accumulations over row blocks) is that of chain_mma.cu.
#include "common.cuh"
#ifdef C2_PROFILE
#include <cstdio>
#endif
#include "kernels.h"

#define C2_GT 8          // fits per CTA (the M dimension of m8n8k4)
#define C2_NW 20         // slots of a fit's exponent maxima: one per (warp, thread of the quad), up to 5 warps
#define C2_FST 4         // stages of the f ring (cp.async prefetch distance + 1)
#define C2_RIMAX 8192    // chains up to this many rows keep their row descriptors in shared memory

__device__ __forceinline__ double c2_pow2(int be) { return __hiloint2double((2046 - be) << 20, 0); }
__device__ __forceinline__ void c2_dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void c2_cp8(unsigned dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}

struct Chain2Args {
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
    // breakend rows: the states grouped by their tumour-clone totals ("total classes"), see the breakend path below
    const uint8_t* bc;     // [perm (4 KBB)] [kbcls (KBB)] [scls (S)] [didx (BC, BC)] [ddel (ND, M)]
    int BC, KBB, ND;
};

// message in fragment order: state i of fit g lives at (i/4)*32 + g*4 + i%4
__device__ __forceinline__ int c2_pos(int g, int i) { return ((i >> 2) << 5) + (g << 2) + (i & 3); }

template <int KB, int NBW, bool LONE>
struct C2Cfg {
    static constexpr int NWARPS = LONE ? 5 : 4;
    static constexpr int T = NWARPS * 32;
    static constexpr int NSLOT = 2 * NBW;       // f values per thread and segment (the helper warp uses slot 0)
};

// The sweep of one (chain, direction, tile of 8 fits).  The direction is a template parameter: every "forward or
// backward" choice of the per-segment code (what is stored, which way the pointers walk, how the class-difference table
// is indexed) is folded at compile time instead of costing selects and moves on the serial path of the one warp a
// sub-partition has.
template <int KB, int NBW, bool LONE, int DIR>
__device__ __forceinline__ void c2_sweep(const Chain2Args& A, unsigned char* smem_raw) {
    using Cfg = C2Cfg<KB, NBW, LONE>;
    constexpr int T = Cfg::T, NSLOT = Cfg::NSLOT;
    const RmxDims& d = A.d;
    const int S = d.S, M = d.M, MD = M * d.D;
    double* u = reinterpret_cast<double*>(smem_raw);                  // [2][KB*32]
    double* wtab = u + 2 * KB * 32;                                    // [8][MD]   breakend tables of the step
    double* redd = wtab + C2_GT * MD;                                  // [8][C2_NW] + [8] mF sums
    unsigned* redi = reinterpret_cast<unsigned*>(redd + C2_GT * C2_NW + C2_GT);  // [2][8][C2_NW] largest high words (as unsigned: a NaN is the largest)
    int8_t* tots = reinterpret_cast<int8_t*>(redi + 2 * C2_GT * C2_NW);   // [S*M]
    // f ring: [C2_FST stages][NSLOT][T] doubles, every thread copies its own entries with cp.async (own slots only)
    double* fring = reinterpret_cast<double*>(smem_raw + ((sizeof(double) * ((size_t)2 * KB * 32 + (size_t)C2_GT * MD + C2_GT * C2_NW + C2_GT) +
                                                           sizeof(int) * 2 * C2_GT * C2_NW + (size_t)S * M + 15) & ~(size_t)15)) + threadIdx.x;
    // breakend rows: this thread's flipE fragments in total-class row order [NBW * KBB][128] (DMMA threads), the per-row
    // factors Wd [8][ND] of the ND distinct total-difference tuples, the class tables, then the row info of the chain
    const int BC = A.BC, KBB = A.KBB, ND = A.ND;
    double* fes = fring + (size_t)C2_FST * NSLOT * T - threadIdx.x;           // base of the block
    double* Wd = fes + (size_t)NBW * KBB * 128;
    uint8_t* bct = reinterpret_cast<uint8_t*>(Wd + (size_t)C2_GT * ND);
    const uint8_t* bperm = bct;                   // row of the class-ordered layout -> state (255 = padding)
    const uint8_t* kbcls = bct + 4 * KBB;         // class of a row block
    const uint8_t* scls = kbcls + KBB;            // class of a state
    const uint8_t* didx = scls + S;               // (BC, BC): which difference tuple is class ci minus class cj
    const uint8_t* ddel = didx + BC * BC;         // (ND, M): the tuple's per-clone differences + 128
    const int bct_n = 5 * KBB + S + BC * BC + ND * M;
    int* rism = reinterpret_cast<int*>(bct + ((bct_n + 15) & ~15));            // [C2_RIMAX] row info of the chain

    const int tile = blockIdx.x % A.ntiles;
    const int cd = blockIdx.x / A.ntiles;
    constexpr int dir = DIR;
    const int chain = A.ch.order[cd >> 1];
    const int s0 = A.ch.start[chain], s1 = A.ch.end[chain];
    const int t = threadIdx.x, w = t >> 5, lane = t & 31;
    const int gid = lane >> 2, tig = lane & 3;     // mma fragment coordinates
    const bool helper = LONE && w == 4;
    const int jl = 32 * NBW;                         // the lone last state (LONE)
    const int gbase = A.g0 + tile * C2_GT;
    const int gcount = min(C2_GT, A.g0 + A.gn - gbase);
    const int64_t SS = (int64_t)S * S;
    const int64_t NS = (int64_t)d.N * S;
    constexpr int sgn = dir == 0 ? 1 : -1;
    const int jbase = helper ? jl : w * (8 * NBW);
    const bool fit_ok = gid < gcount;
    const int gfit = gbase + min(gid, gcount - 1);
    // DMMA warps: C fragment = fit gid, columns j0[h], j0[h] + 1 of column block h; helper: fit gid, column jl (tig 0).
    // Rows of a partial tile beyond its last fit repeat that fit (gfit is clamped): they are computed like the others and
    // only their stores to global memory are dropped, so the per-segment code carries no per-fit masking.  With a lone
    // last state (LONE) the 32 NBW columns of the DMMA warps are all inside the state space and no column masking either.
    // position of column block 0 in the message buffer (j0 even: j0, j0 + 1 adjacent, 16-byte aligned); block h is 64
    // doubles further (jbase is a multiple of 8), so one address register serves all blocks
    const int upos0 = c2_pos(gid, (helper ? 0 : jbase) + 2 * tig);
    bool ok0[NBW], ok1[NBW];
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        const int j0 = jbase + 8 * h + 2 * tig;
        ok0[h] = !helper && (LONE || j0 < S);
        ok1[h] = !helper && (LONE || j0 + 1 < S);
    }
    const bool hok = helper && tig == 0;

    for (int e = t; e < 2 * KB * 32; e += T) u[e] = 0.;
    // the exponent exchange: every thread leaves the largest HIGH WORD of the message entries it produced (the entries
    // are non-negative, so the integer order of the high words is the order of the values) in its own slot of its fit;
    // the reader takes the maximum of the fit's slots and its exponent field.  No shuffles on the serial path.
    for (int e = t; e < 2 * C2_GT * C2_NW; e += T) redi[e] = ((e % C2_NW) == 0) ? 0x3ff00000 : 0;   // first message: max f = 1
    for (int e = t; e < S * M; e += T) {
        int s = e / M, m = e - s * M;
        tots[e] = (m == 0) ? 0 : (int8_t)(A.tab[(s * M + m) * 2] + A.tab[(s * M + m) * 2 + 1]);
    }
    for (int e = t; e < bct_n; e += T) bct[e] = A.bc[e];
    __syncthreads();

    const int nsteps = s1 - s0;
    const int seg_first = dir == 0 ? s0 : s1;
    // this thread's row pointers: f source and message destination at its first column, advanced one row per segment
    const double* fp = A.fexp + (int64_t)gfit * NS + (int64_t)seg_first * S + jbase + (helper ? 0 : 2 * tig);
    double* outp = (dir == 0 ? A.ahat : A.bhat) + (int64_t)gfit * NS + (int64_t)seg_first * S + jbase + (helper ? 0 : 2 * tig);
    const int rowstep = sgn * S;

    // ---- first segment of the sweep: u = f (forward: ahat = f; backward: bhat = 1)
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        double u0 = 0., u1 = 0.;
        if (ok0[h]) u0 = fp[8 * h];
        if (ok1[h]) u1 = fp[8 * h + 1];
        if (ok0[h] && fit_ok) outp[8 * h] = dir == 0 ? u0 : 1.;
        if (ok1[h] && fit_ok) outp[8 * h + 1] = dir == 0 ? u1 : 1.;
        if (ok0[h]) *reinterpret_cast<double2*>(u + (upos0 + 64 * h)) = make_double2(u0, u1);
    }
    if (hok) {
        const double u0 = fp[0];
        if (fit_ok) outp[0] = dir == 0 ? u0 : 1.;
        u[c2_pos(gid, jl)] = u0;
    }
    auto row_of = [&](int k) { return dir == 0 ? s0 + k : s1 - 1 - k; };       // transition row of step k
    // software pipeline: f C2_FST - 1 segments ahead through the shared-memory ring, row info two rows ahead
    const unsigned fring_s = (unsigned)__cvta_generic_to_shared(fring);
    const double* fsrc = fp + rowstep;        // f row of step 0's segment; advanced by f_issue
    int f_next = 0;
    auto f_issue = [&]() {
        if (f_next < nsteps) {
            const unsigned base = fring_s + 8u * (unsigned)(((f_next % C2_FST) * NSLOT) * T);
#pragma unroll
            for (int h = 0; h < NBW; ++h) {
                if (ok0[h]) c2_cp8(base + 8u * (unsigned)((h * 2) * T), fsrc + 8 * h);
                if (ok1[h]) c2_cp8(base + 8u * (unsigned)((h * 2 + 1) * T), fsrc + 8 * h + 1);
            }
            if (hok) c2_cp8(base, fsrc);
            fsrc += rowstep;
        }
        ++f_next;
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int e = 0; e < C2_FST * NSLOT; ++e) fring[e * T] = 0.;     // entries outside the state space stay zero
#pragma unroll
    for (int k = 0; k < C2_FST - 1; ++k) f_issue();
    const bool ri_sm = nsteps <= C2_RIMAX;
    if (ri_sm)
        for (int k = t; k < nsteps; k += T) rism[k] = A.rinfo[row_of(k)];
    __syncthreads();
    auto ri_at = [&](int k) { return ri_sm ? rism[k] : A.rinfo[row_of(k)]; };
    int ri0 = nsteps > 0 ? ri_at(0) : 0, ri1 = nsteps > 1 ? ri_at(1) : 0;

    double ereg[NBW][KB];       // DMMA warps: B fragments; helper: ereg[0][kb] = E[4 kb + tig][jl],
                                // ereg[1][kb] = flipE[4 kb + tig][jl] (NBW >= 2 whenever there is a helper)
    int cur_class = -1, cur_bclass = -1;
    // offsets of this thread's output columns' total classes into didx (breakend rows): column index forward, row
    // index backward
    int cjo[NBW][2];
#pragma unroll
    for (int h = 0; h < NBW; ++h) {
        const int j0 = jbase + 8 * h + 2 * tig;
        cjo[h][0] = ((!helper && j0 < S) ? scls[j0] : 0) * (dir == 0 ? 1 : BC);
        cjo[h][1] = ((!helper && j0 + 1 < S) ? scls[j0 + 1] : 0) * (dir == 0 ? 1 : BC);
    }
    int cur = 0, esum = 0, bad = 0;

#ifdef C2_PROFILE
    unsigned pt0 = 0, pd1 = 0, pd2 = 0, pd3 = 0, pd4 = 0, pd5 = 0, pn = 0;
#ifdef C2_PROFILE2
    unsigned pe1 = 0, pe2 = 0, pe3 = 0;
#endif
    auto pclk = []() { unsigned v; asm volatile("mov.u32 %0, %%clock;" : "=r"(v) :: "memory"); return v; };
    pt0 = pclk();
#endif
    for (int step = 0; step < nsteps; ++step) {
        outp += rowstep;
        const int ri = ri0;
        ri0 = ri1;
        const int cls = (ri & 0xff) - 1;
        const int br = (ri >> 12) - 1;
        const double* uc = u + cur * KB * 32;
        double c[NBW][2], f[NBW][2];      // f: this segment's f values times the scale of the step
        int be = 1023;
        double sc = 1.;
#pragma unroll
        for (int h = 0; h < NBW; ++h) c[h][0] = c[h][1] = f[h][0] = f[h][1] = 0.;

        // bookkeeping of the segment, placed inside the DMMA block by the callers below
        auto chores = [&](int part) {
            if (part == 0) {
                // exponent of the largest entry of the input message of fit gid (written before the last barrier)
                const uint4* rp = reinterpret_cast<const uint4*>(redi + (cur * C2_GT + gid) * C2_NW);
                unsigned mh = 0u;
#pragma unroll
                for (int q = 0; q < C2_NW / 4; ++q) {
                    const uint4 r = rp[q];
                    mh = max(mh, max(max(r.x, r.y), max(r.z, r.w)));
                }
                be = (int)((mh >> 20) & 0x7ffu);
                sc = c2_pow2(be);
                asm volatile("" : "+d"(sc));        // here, under the DMMAs, not in the epilogue
                esum += be - 1023;
                if (be <= 0 || be >= 2047) bad = 1;
            } else if (part == 1) {
                asm volatile("cp.async.wait_group %0;" ::"n"(C2_FST - 2) : "memory");     // this step's f has landed
                const double* fr = fring + ((step % C2_FST) * NSLOT) * T;
                // f times the scale (a power of two: exact), so that the epilogue is one multiplication per entry
#pragma unroll
                for (int h = 0; h < NBW; ++h) {
                    f[h][0] = fr[(h * 2) * T] * sc;
                    f[h][1] = fr[(h * 2 + 1) * T] * sc;
                    asm volatile("" : "+d"(f[h][0]), "+d"(f[h][1]));
                }
            } else {
                f_issue();          // into the ring stage the previous step consumed
                if (step + 2 < nsteps) ri1 = ri_at(step + 2);
            }
        };

        unsigned mx = 0u;
        // epilogue of one column block (C-fragment layout): scale by the power of two that normalised the input message,
        // times f, store the message and pass it on in fragment order
        auto finish_block = [&](int h, double c0, double c1) {
            double u0 = c0 * f[h][0], u1 = c1 * f[h][1];      // the message passed on; forward it is also what is stored
            if (!ok0[h]) u0 = 0.;
            if (!ok1[h]) u1 = 0.;
#ifndef C2_NOSTG
            if (ok0[h] && fit_ok) outp[8 * h] = dir == 0 ? u0 : c0 * sc;
            if (ok1[h] && fit_ok) outp[8 * h + 1] = dir == 0 ? u1 : c1 * sc;
#endif
            if (ok0[h]) *reinterpret_cast<double2*>(u + (cur ^ 1) * KB * 32 + (upos0 + 64 * h)) = make_double2(u0, u1);
            mx = max(mx, max((unsigned)__double2hiint(u0), (unsigned)__double2hiint(u1)));
        };

        if (br < 0) {
            if (cls != cur_class) {
                const double* E = A.sp.Emat + (int64_t)cls * SS;
                if (!helper) {
#pragma unroll
                    for (int h = 0; h < NBW; ++h)
#pragma unroll
                        for (int kb = 0; kb < KB; ++kb) {
                            const int i = kb * 4 + tig, jb = jbase + 8 * h + gid;
                            double ev = (i < S && jb < S) ? E[(int64_t)i * S + jb] : 0.;
                            asm volatile("" : "+d"(ev));     // opaque: keep the value, do not re-load it every segment
                            ereg[h][kb] = ev;
                        }
                } else {
#pragma unroll
                    for (int kb = 0; kb < KB; ++kb) {
                        const int i = kb * 4 + tig;
                        double ev = i < S ? E[(int64_t)i * S + jl] : 0.;
                        asm volatile("" : "+d"(ev));
                        ereg[0][kb] = ev;
                    }
                }
                cur_class = cls;
            }
            if (!helper) {
                // Row blocks outermost: one A fragment (one shared-memory load) feeds the DMMAs of all NBW column blocks,
                // so consecutive DMMAs belong to different accumulators by construction.  (Column block after column
                // block, with the even and odd row blocks of a block on two accumulators, let the epilogue of block h
                // run under the DMMAs of block h + 1 -- but ptxas re-ordered the DMMAs of a block into its two dependent
                // chains, 26 instead of 16 cycles per DMMA; a last block lagging behind the others was clumped the same
                // way.)  The bookkeeping sits under the first DMMAs; the NBW epilogues are independent of each other and
                // pipeline after the last DMMA.  Every accumulator sums the row blocks it summed before (even / odd row
                // blocks of a column block, added at the end): same values bit for bit.
                double acc[NBW][2][2];
#pragma unroll
                for (int h = 0; h < NBW; ++h) acc[h][0][0] = acc[h][0][1] = acc[h][1][0] = acc[h][1][1] = 0.;
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) {
                    const double a = uc[kb * 32 + lane];
#ifdef C2_PROFILE
                    if (kb == 0) { asm volatile("" :: "d"(a)); unsigned q = pclk(); pd1 += q - pt0; pt0 = q; }
#ifdef C2_PROFILE2
                    if (kb == 4) { unsigned q = pclk(); pe1 += q - pt0; }
                    if (kb == 8) { unsigned q = pclk(); pe2 += q - pt0; }
                    if (kb == 14) { unsigned q = pclk(); pe3 += q - pt0; }
#endif
#endif
#pragma unroll
                    for (int h = 0; h < NBW; ++h) c2_dmma(acc[h][kb & 1][0], acc[h][kb & 1][1], a, ereg[h][kb]);
                    if (kb == (KB > 6 ? 2 : 0)) chores(0);
                    if (kb == (KB > 12 ? 6 : 0)) chores(1);
                    if (kb == (KB > 18 ? 12 : 0)) chores(2);
                }
#ifdef C2_PROFILE
                { unsigned q = pclk(); pd2 += q - pt0; pt0 = q; ++pn; }
#endif
#pragma unroll
                for (int h = 0; h < NBW; ++h) finish_block(h, acc[h][0][0] + acc[h][1][0], acc[h][0][1] + acc[h][1][1]);
#ifdef C2_PROFILE
                { asm volatile("" :: "r"(mx)); unsigned q = pclk(); pd3 += q - pt0; pt0 = q; }
#endif
            } else {
                // the lone last state: out[g][jl] = sum_i u[g][i] E[i][jl], lane (g, q) takes rows i = q mod 4
                chores(0);
                chores(1);
                chores(2);
                double a0 = 0., a1 = 0.;
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) {
                    const double a = uc[kb * 32 + lane];
                    if (kb & 1) a1 = fma(a, ereg[0][kb], a1);
                    else a0 = fma(a, ereg[0][kb], a0);
                }
                a0 += a1;
                a0 += __shfl_xor_sync(0xffffffffu, a0, 1);
                a0 += __shfl_xor_sync(0xffffffffu, a0, 2);
                c[0][0] = a0;
            }
        } else {
            // ---- breakend row: per-fit matrix  E_g[i][j] = flipE[i][j] * prod_m wexp[g][m][tot_i[m] - tot_j[m]].
            // The per-fit factor depends on (i, j) only through the states' tumour-clone totals.  With the states grouped
            // by those totals into BC classes (29 at S = 97), each class padded to whole row blocks of four,
            //     out[g][j] = w0[g] * sum_ci  W[g][ci - class(j)] * ( sum_{i in ci} u[g][i] flipE[i][j] ),
            // and the inner sums are ONE DMMA per (row block, column block) with the shared operand flipE -- the tensor
            // path again, KBB / KB times the DMMAs of a plain row plus one multiply-add per DMMA output, instead of
            // 8 S^2 scalar products with three table lookups each (19 % of the first version of this kernel).
            chores(0);
            chores(1);
            chores(2);
            for (int e = t; e < C2_GT * MD; e += T) {
                const int g = e / MD, o = e - g * MD;
                const int gg = gbase + min(g, gcount - 1);
                wtab[e] = A.sp.wexp[(int64_t)gg * A.sp_fit_stride + (int64_t)br * MD + o];
            }
            const double* FE = A.sp.flipE + (int64_t)cls * SS;
            if (cls != cur_bclass) {        // this thread's flipE fragments, rows in class order (once per chain and class)
                if (!helper) {
                    for (int h = 0; h < NBW; ++h) {
                        const int jb = jbase + 8 * h + gid;
                        for (int kb = 0; kb < KBB; ++kb) {
                            const int i = bperm[4 * kb + tig];
                            fes[(size_t)(h * KBB + kb) * 128 + t] = (i < S && jb < S) ? FE[(int64_t)i * S + jb] : 0.;
                        }
                    }
                }
                if (LONE && helper) {
#pragma unroll
                    for (int kb = 0; kb < KB; ++kb) {
                        const int i = kb * 4 + tig;
                        double ev = i < S ? FE[(int64_t)i * S + jl] : 0.;
                        asm volatile("" : "+d"(ev));
                        ereg[NBW > 1 ? 1 : 0][kb] = ev;
                    }
                }
                cur_bclass = cls;
            }
            __syncthreads();
            for (int e = t; e < C2_GT * ND; e += T) {     // the factor of every distinct tuple of total differences
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
            if (!helper) {
                double o[NBW][2];
#pragma unroll
                for (int h = 0; h < NBW; ++h) o[h][0] = o[h][1] = 0.;
                const double* fb = fes + t;
                const double* ug = uc + gid * 4;
                const int hstride = KBB * 128;
#pragma unroll 2
                for (int kb = 0; kb < KBB; ++kb) {
                    const int i = bperm[4 * kb + tig];
                    const double a = ug[i < S ? ((i >> 2) << 5) + (i & 3) : 0];      // padded rows have a zero B fragment
                    const uint8_t* dr = didx + (int)kbcls[kb] * dstr;
#pragma unroll
                    for (int h = 0; h < NBW; ++h) {
                        const double w0h = wd[dr[cjo[h][0]]], w1h = wd[dr[cjo[h][1]]];
                        double p0 = 0., p1 = 0.;
                        c2_dmma(p0, p1, a, fb[h * hstride]);
                        o[h][0] = fma(p0, w0h, o[h][0]);
                        o[h][1] = fma(p1, w1h, o[h][1]);
                    }
                    fb += 128;
                }
#pragma unroll
                for (int h = 0; h < NBW; ++h) { c[h][0] = o[h][0] * w0; c[h][1] = o[h][1] * w0; }
            } else {
                // lone column: lane (g, q) sums the rows i = q mod 4 of its own fit
                const uint8_t* dc = didx + (int)scls[jl] * (dir == 0 ? 1 : BC);
                double a0 = 0.;
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) {
                    const int i = kb * 4 + tig;
                    if (i < S) a0 = fma(ereg[NBW > 1 ? 1 : 0][kb] * uc[kb * 32 + lane], wd[dc[(int)scls[i] * dstr]], a0);
                }
                a0 += __shfl_xor_sync(0xffffffffu, a0, 1);
                a0 += __shfl_xor_sync(0xffffffffu, a0, 2);
                c[0][0] = a0 * w0;
            }
        }
        // ---- what is left of the step
        if (!helper) {
            if (br >= 0) {
#pragma unroll
                for (int h = 0; h < NBW; ++h) finish_block(h, c[h][0], c[h][1]);
            }
        } else {
            double u0 = c[0][0] * f[0][0];
            if (!hok) u0 = 0.;
            if (hok) {
                if (fit_ok) outp[0] = dir == 0 ? u0 : c[0][0] * sc;
                u[(cur ^ 1) * KB * 32 + c2_pos(gid, jl)] = u0;
            }
            mx = (unsigned)__double2hiint(u0);
        }
        redi[((cur ^ 1) * C2_GT + gid) * C2_NW + 4 * w + tig] = mx;
        cur ^= 1;
#ifdef C2_PROFILE
        { unsigned q = pclk(); if (br < 0 && !helper) pd4 += q - pt0; pt0 = q; }
#endif
        __syncthreads();
#ifdef C2_PROFILE
        { unsigned q = pclk(); if (br < 0 && !helper) pd5 += q - pt0; pt0 = q; }
#endif
    }
#ifdef C2_PROFILE
    if (blockIdx.x < 2 * A.ntiles && tile == 0 && lane == 0 && !helper)
        printf("c2prof dir %d warp %d rows %u: head %.1f dmma %.1f epi %.1f tail %.1f barrier %.1f cycles per plain row\n", dir, w, pn,
               (double)pd1 / pn, (double)pd2 / pn, (double)pd3 / pn, (double)pd4 / pn, (double)pd5 / pn);
#ifdef C2_PROFILE2
    if (blockIdx.x < 2 * A.ntiles && tile == 0 && lane == 0 && !helper)
        printf("c2prof2 dir %d warp %d: DMMA phase up to row block 4: %.1f, 8: %.1f, 14: %.1f, 25: %.1f\n", dir, w, (double)pe1 / pn,
               (double)pe2 / pn, (double)pe3 / pn, (double)pd2 / pn);
#endif
#endif

    // ---- log normaliser of the forward sweep: log(sum of the last message) + ln2 * esum + sum_n max F[n]
    {
        // the last message, read back from its shared-memory copy (entries outside the state space are zero there):
        // keeping it in registers through the sweep costs twelve registers the per-segment code needs
        const double* ul = u + cur * KB * 32;
        double sv = 0.;
        if (!helper) {
#pragma unroll
            for (int h = 0; h < NBW; ++h) {
                if (ok0[h]) {
                    const double2 q = *reinterpret_cast<const double2*>(ul + (upos0 + 64 * h));
                    sv += q.x + q.y;
                }
            }
        } else {
            sv = tig == 0 ? ul[c2_pos(gid, jl)] : 0.;
        }
        sv += __shfl_xor_sync(0xffffffffu, sv, 1);
        sv += __shfl_xor_sync(0xffffffffu, sv, 2);
        if (tig == 0) redd[gid * C2_NW + w] = sv;
        for (int g = w; g < gcount; g += Cfg::NWARPS) {
            const double* mf = A.mF + (int64_t)(gbase + g) * d.N;
            double ms = 0.;
            for (int n = s0 + lane; n <= s1; n += 32) ms += mf[n];
            ms = warp_sum(ms);
            if (lane == 0) redd[C2_GT * C2_NW + g] = ms;
        }
    }
    __syncthreads();
    if (w == 0 && tig == 0 && fit_ok) {
        double slast = 0.;
        for (int c2 = 0; c2 < Cfg::NWARPS; ++c2) slast += redd[gid * C2_NW + c2];
        if (dir == 0) {
            const double logz = log(slast) + 0.693147180559945309417 * (double)esum + redd[C2_GT * C2_NW + gid];
            A.chain_logz[(int64_t)(gbase + gid) * A.ch.NC + chain] = logz;
            if (isnan(logz)) bad = 1;
        }
        if (!(slast > 0.) || isinf(slast)) bad = 1;
        if (bad) atomicOr(A.dflags, RMX_F_MSG_NAN);
    }
}

template <int KB, int NBW, bool LONE>
__global__ void __launch_bounds__(C2Cfg<KB, NBW, LONE>::T, 1) k_chain_mma2(Chain2Args A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if (((blockIdx.x / A.ntiles) & 1) == 0) c2_sweep<KB, NBW, LONE, 0>(A, smem_raw);
    else c2_sweep<KB, NBW, LONE, 1>(A, smem_raw);
}

template <int KB, int NBW, bool LONE>
static int launch_chain_mma2_cfg(cudaStream_t st, Chain2Args& A) {
    using Cfg = C2Cfg<KB, NBW, LONE>;
    const RmxDims& d = A.d;
    size_t smem = sizeof(double) * ((size_t)2 * KB * 32 + (size_t)C2_GT * d.M * d.D + C2_GT * C2_NW + C2_GT) +
                  sizeof(int) * 2 * C2_GT * C2_NW + (size_t)d.S * d.M + 16;
    const size_t bct_n = (size_t)5 * A.KBB + d.S + (size_t)A.BC * A.BC + (size_t)A.ND * d.M;
    smem = ((smem + 15) & ~(size_t)15) + sizeof(double) * ((size_t)C2_FST * Cfg::NSLOT * Cfg::T + (size_t)NBW * A.KBB * 128 +
                                                             (size_t)C2_GT * A.ND) +
           ((bct_n + 15) & ~(size_t)15) + sizeof(int) * C2_RIMAX + 16;
    auto kern = k_chain_mma2<KB, NBW, LONE>;
    if (smem > 48 * 1024) {
        if (smem > 227 * 1024) return 2;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 1;
    }
    A.ntiles = (A.gn + C2_GT - 1) / C2_GT;
    kern<<<A.ch.NC * 2 * A.ntiles, Cfg::T, smem, st>>>(A);
    return cudaPeekAtLastError() == cudaSuccess ? 0 : 1;
}

// returns 0 on success, 1 if the configuration or the launch failed, 2 if this kernel does not cover the shape
int rmx_launch_chain_mma2(cudaStream_t st, const RmxTables& T, const RmxDims& d, const RmxChains& ch, const int* rinfo,
                          const RmxTransSpec& sp, int64_t sp_fit_stride, int g0, int gn, const double* fexp,
                          const double* mF, double* ahat, double* bhat, double* chain_logz, int* dflags,
                          const uint8_t* bc, int BC, int KBB, int ND) {
    if (ch.NC <= 0 || gn <= 0) return 0;
    if (rinfo == nullptr || bc == nullptr || BC <= 0 || ND <= 0 || ND > 256 || KBB > 48) return 2;
    Chain2Args A;
    A.d = d; A.ch = ch; A.rinfo = rinfo; A.sp = sp; A.sp_fit_stride = sp_fit_stride; A.tab = T.tab;
    A.g0 = g0; A.gn = gn; A.ntiles = 1; A.fexp = fexp; A.mF = mF; A.ahat = ahat; A.bhat = bhat;
    A.chain_logz = chain_logz; A.dflags = dflags;
    A.bc = bc; A.BC = BC; A.KBB = KBB; A.ND = ND;
    if (d.S == 97) return launch_chain_mma2_cfg<25, 3, true>(st, A);      // M = 3, max copy number 6 (BASELINE configs 1, 2, 4, 5)
    if (d.S > 64 && d.S <= 96) return launch_chain_mma2_cfg<24, 3, false>(st, A);
    if (d.S > 32 && d.S <= 64) return launch_chain_mma2_cfg<16, 2, false>(st, A);
    if (d.S > 16 && d.S <= 32) return launch_chain_mma2_cfg<8, 1, false>(st, A);
    return 2;
}
