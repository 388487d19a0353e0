// tp_tri.cuh -- kernel TRI: the exhaustive scorer for k' >= 6 on up to 80 + p free devices.
// Included by tp_tri_m*_d*.cu, one translation unit per (lane-vector width M, walked depth DW).
//
// Same contract and the same outer structure as kernel SPLIT (tp_split.cuh): positions are the
// canonical order reversed, a subset is j_1 > ... > j_k'; the top p positions A belong to a LANE
// (32 consecutive A's of one pivot per warp, V[j] = sum_{a in A} W[a][j] in shared memory), the
// next DW positions are walked by the WARP (T[j] = u[j] + sum over walked b of W[b][j]).  What
// differs is everything below: the last THREE positions b > jx > jy are scored against a table
// the block builds once per request,
//
//     S3[b][jx][jy] = W[b][jx] + W[b][jy] + W[jx][jy]          (the score inside the triple)
//
// so that a lane needs nothing per leaf b: its vector VT[j] = V[j] + T[j] is built once per walked
// prefix (not once per leaf, as SPLIT's P = V + T + W[b] is) and
//
//     score(A, walked, b, jx, jy) = base + VT[b] + VT[jx] + VT[jy] + S3[b][jx][jy].
//
// Layout that makes one row cost nothing beyond its pairs: rows jx = 2h and 2h+1 of leaf b are
// interleaved in one 32-bit word per jy (low half: row 2h, high half: row 2h+1; TP_BIG where
// jy >= jx or the row does not exist), and the lane keeps VT[jy] SPLATTED (the same 16-bit value in
// both halves of a register, one register per position).  Then
//
//     acc = min(acc, VT[jy] + word(h, jy))        one VIADDMNMX.U16x2 = two subsets, for jy <= 2h
//     run = min(run, acc + (VT[2h] | VT[2h+1] << 16))   ONE closing op for TWO rows (+ one PRMT)
//
// where SPLIT pays a PRMT and a VIADDMNMX per row, and an odd row wastes half an op only once per
// row pair.  Leaves 2g+1 and 2g+2 have the same row pairs 0..g and are scored in one pass: two
// independent chains off the same lane registers, two table rows per load group.
//
// Shared memory decides how many leaves get a table: 2 * sum_{g<G} off(g+1) 16-byte chunks, 100 KB
// for G = 32 (leaves 1..64, which is every leaf for n - p <= 67).  Leaves above the cap (0.06 % of
// the subsets of the 8-of-80 search) go to the exact scalar re-walk that also serves candidates
// (tp_split_slow), which prunes whole rows against the bound.
#pragma once
#include "tp_split.cuh"

#ifndef TP_TRI_WARPS
#define TP_TRI_WARPS 16
#endif
#define TP_TRI_THREADS (TP_TRI_WARPS * 32)
#ifndef TP_TRI_GC
#define TP_TRI_GC 24 /* leaf pairs with a table, at most: rows below 2 GC -> 2 GC lane registers */
#endif
#define TP_TRI_NREG (2 * TP_TRI_GC)


// Row pair h of a leaf's table holds entries jy = 0..2h, padded to 16-byte chunks of four.
__host__ __device__ constexpr int tp_tri_chunks(int h) { return (h >> 1) + 1; }
// ... and starts at this chunk: sum of tp_tri_chunks(h') over h' < h.
__host__ __device__ constexpr int tp_tri_off(int h) {
    return (h & 1) ? ((h >> 1) + 1) * ((h >> 1) + 1) : (h >> 1) * ((h >> 1) + 1);
}
// Leaves 2g+1 and 2g+2 both have row pairs 0..g, tp_tri_off(g+1) chunks each; pair g starts here.
__host__ __device__ constexpr int tp_tri_pair_base(int g) {
    int s = 0;
    for (int i = 0; i < g; i++) s += tp_tri_off(i + 1);
    return 2 * s;
}
static __host__ __device__ inline size_t tp_tri_table_bytes(int gcap) { return (size_t)tp_tri_pair_base(gcap) * 16; }

// Shared-memory loads by 32-bit shared address + immediate: the hot function is not inlined, and
// a generic pointer handed across a call would turn every table load into a 64-bit generic LD.
__device__ __forceinline__ uint4 tp_lds128(unsigned addr) {
    uint4 v;
    asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint4 tp_lds128v(unsigned addr) { // volatile: executed where it stands, never kept alive
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ unsigned tp_lds32(unsigned addr) {
    unsigned v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ unsigned tp_lds16(unsigned addr) {
    unsigned short v;
    asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
}

// Two "acc = min(acc, vt + word)" steps.  With TP_TRI_MIN3 they are IMAD + IMAD + VIMNMX3.U16x2: the adds go to
// the FMA pipe (a multiply by a register that holds 1 keeps ptxas from turning them back into ALU adds) and the
// half-rate ALU pipe sees one instruction for two steps instead of two VIADDMNMX.  As a microbenchmark that is
// 1.77 x the add-min rate (scripts/int_peaks.py, variants 6 / 7: 16.4 T against 29.0 T steps/s); in this kernel
// it is 12-17 % SLOWER (8 of 64: 0.358 against 0.321 ms, 8 of 80: 1.86 against 1.58 ms): three issue slots
// for two steps instead of two, and the kernel as a whole is short of issue slots, not of ALU cycles.  Off.
#ifndef TP_TRI_MIN3
#define TP_TRI_MIN3 0
#endif
__device__ __forceinline__ unsigned tp_step2(unsigned acc, unsigned v0, unsigned w0, unsigned v1, unsigned w1, unsigned one) {
#if TP_TRI_MIN3
    unsigned t0, t1;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(t0) : "r"(w0), "r"(one), "r"(v0));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(t1) : "r"(w1), "r"(one), "r"(v1));
    return __vimin3_u16x2(t0, t1, acc);
#else
    (void)one;
    return __viaddmin_u16x2(v1, w1, __viaddmin_u16x2(v0, w0, acc));
#endif
}

// One row pair of both leaves of a pass.  H is a compile-time constant: every register index and
// every table offset below is an immediate.  tA, tB: shared addresses of the two leaves' tables.
template <int H>
__device__ __forceinline__ void tp_tri_rowpair(const unsigned (&VT)[TP_TRI_NREG], unsigned tA, unsigned tB,
                                               unsigned &runA, unsigned &runB, unsigned one) {
    constexpr int NCH = tp_tri_chunks(H), OFF = tp_tri_off(H);
    unsigned accA = TP_BIG2 + TP_BIG2, accB = TP_BIG2 + TP_BIG2;
#pragma unroll
    for (int c = 0; c < NCH; c++) {
        const uint4 a = tp_lds128(tA + (OFF + c) * 16), b = tp_lds128(tB + (OFF + c) * 16);
        if (4 * c + 1 <= 2 * H) {
            accA = tp_step2(accA, VT[4 * c + 0], a.x, VT[4 * c + 1], a.y, one);
            accB = tp_step2(accB, VT[4 * c + 0], b.x, VT[4 * c + 1], b.y, one);
        } else {
            accA = __viaddmin_u16x2(VT[4 * c + 0], a.x, accA);
            accB = __viaddmin_u16x2(VT[4 * c + 0], b.x, accB);
        }
        if (4 * c + 3 <= 2 * H) {
            accA = tp_step2(accA, VT[4 * c + 2], a.z, VT[4 * c + 3], a.w, one);
            accB = tp_step2(accB, VT[4 * c + 2], b.z, VT[4 * c + 3], b.w, one);
        } else if (4 * c + 2 <= 2 * H) {
            accA = __viaddmin_u16x2(VT[4 * c + 2], a.z, accA);
            accB = __viaddmin_u16x2(VT[4 * c + 2], b.z, accB);
        }
    }
    const unsigned pk = __byte_perm(VT[2 * H], VT[2 * H + 1], 0x5410); // VT[2H] | VT[2H+1] << 16
    runA = __viaddmin_u16x2(accA, pk, runA);
    runB = __viaddmin_u16x2(accB, pk, runB);
}

// The generic pass enters its rows through a switch, so every row pair is a basic block of its
// own and the compiler cannot start a row pair's loads during the one before it.  With
// TP_TRI_PREFETCH the first chunk of row pair H - 1 (both leaves) is loaded by hand while row pair H
// computes, into two alternating pairs of buffers.
#ifndef TP_TRI_PREFETCH
#define TP_TRI_PREFETCH 1
#endif
// Chunks are consumed row pair by row pair (H descending, c ascending); chunk (H, c) sits in buffer
// set (S(H) + c) & 1, S(H) = chunks of the row pairs above H, so that consecutive chunks alternate.
__host__ __device__ constexpr int tp_tri_buf(int h, int c) {
    return (tp_tri_off(TP_TRI_GC) - tp_tri_off(h + 1) + c) & 1;
}
__device__ __forceinline__ unsigned tp_pack_lo(unsigned a, unsigned b) { // a.lo | b.lo << 16; volatile: computed
    unsigned r;                                                          // where it stands, not hoisted out of the
    asm volatile("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(r) : "r"(a), "r"(b)); // pass loop into 24 more registers
    return r;
}
template <int H>
__device__ __forceinline__ void tp_tri_rowpair_pf(const unsigned (&VT)[TP_TRI_NREG], unsigned tA, unsigned tB,
                                                  unsigned &runA, unsigned &runB, uint4 (&bA)[2], uint4 (&bB)[2],
                                                  unsigned one) {
    constexpr int NCH = tp_tri_chunks(H), OFF = tp_tri_off(H);
    unsigned accA = TP_BIG2 + TP_BIG2, accB = TP_BIG2 + TP_BIG2;
#pragma unroll
    for (int c = 0; c < NCH; c++) {
        constexpr int dummy = 0;
        (void)dummy;
        const int cur = tp_tri_buf(H, c);
        // the next chunk (of this row pair, or the first of the one below) starts before this one is used
        if (c + 1 < NCH) {
            bA[cur ^ 1] = tp_lds128v(tA + (OFF + c + 1) * 16);
            bB[cur ^ 1] = tp_lds128v(tB + (OFF + c + 1) * 16);
        } else if (H > 0) {
            bA[cur ^ 1] = tp_lds128v(tA + tp_tri_off(H > 0 ? H - 1 : 0) * 16);
            bB[cur ^ 1] = tp_lds128v(tB + tp_tri_off(H > 0 ? H - 1 : 0) * 16);
        }
        const uint4 a = bA[cur], b = bB[cur];
        if (4 * c + 1 <= 2 * H) {
            accA = tp_step2(accA, VT[4 * c + 0], a.x, VT[4 * c + 1], a.y, one);
            accB = tp_step2(accB, VT[4 * c + 0], b.x, VT[4 * c + 1], b.y, one);
        } else {
            accA = __viaddmin_u16x2(VT[4 * c + 0], a.x, accA);
            accB = __viaddmin_u16x2(VT[4 * c + 0], b.x, accB);
        }
        if (4 * c + 3 <= 2 * H) {
            accA = tp_step2(accA, VT[4 * c + 2], a.z, VT[4 * c + 3], a.w, one);
            accB = tp_step2(accB, VT[4 * c + 2], b.z, VT[4 * c + 3], b.w, one);
        } else if (4 * c + 2 <= 2 * H) {
            accA = __viaddmin_u16x2(VT[4 * c + 2], a.z, accA);
            accB = __viaddmin_u16x2(VT[4 * c + 2], b.z, accB);
        }
    }
    const unsigned pk = tp_pack_lo(VT[2 * H], VT[2 * H + 1]); // VT[2H] | VT[2H+1] << 16
    runA = __viaddmin_u16x2(accA, pk, runA);
    runB = __viaddmin_u16x2(accB, pk, runB);
}

#if TP_TRI_PREFETCH
#define TP_TRI_CASE(H)                                                                               \
    case H:                                                                                          \
        tp_tri_rowpair_pf<(H)>(VT, tA, tB, runA, runB, bufA, bufB, one);                             \
        [[fallthrough]];
#else
#define TP_TRI_CASE(H)                                   \
    case H:                                              \
        tp_tri_rowpair<(H)>(VT, tA, tB, runA, runB, one);     \
        [[fallthrough]];
#endif
#define TP_TRI_CASE8(B)                                                                          \
    TP_TRI_CASE(B + 7) TP_TRI_CASE(B + 6) TP_TRI_CASE(B + 5) TP_TRI_CASE(B + 4) TP_TRI_CASE(B + 3) \
    TP_TRI_CASE(B + 2) TP_TRI_CASE(B + 1) TP_TRI_CASE(B + 0)

// Eight positions of VT = V + T, splatted.
template <int C>
__device__ __forceinline__ void tp_tri_vt_chunk(unsigned (&VT)[TP_TRI_NREG], unsigned v4, unsigned t4) {
    const uint4 v = tp_lds128v(v4 + C * 512); // lane-private: [chunk][lane]
    const uint4 t = tp_lds128v(t4 + C * 16);
    const unsigned x = v.x + t.x, y = v.y + t.y, z = v.z + t.z, w = v.w + t.w;
    VT[8 * C + 0] = __byte_perm(x, 0, 0x1010);
    VT[8 * C + 1] = __byte_perm(x, 0, 0x3232);
    VT[8 * C + 2] = __byte_perm(y, 0, 0x1010);
    VT[8 * C + 3] = __byte_perm(y, 0, 0x3232);
    VT[8 * C + 4] = __byte_perm(z, 0, 0x1010);
    VT[8 * C + 5] = __byte_perm(z, 0, 0x3232);
    VT[8 * C + 6] = __byte_perm(w, 0, 0x1010);
    VT[8 * C + 7] = __byte_perm(w, 0, 0x3232);
}
#define TP_TRI_VT_CASE(C)                  \
    case (C) + 1:                          \
        tp_tri_vt_chunk<(C)>(VT, va, tl);  \
        [[fallthrough]];

// What the hot loop does not need in registers lives here: the block's tables (written once per
// request) and, per warp, the item it is on.  The slow path rebuilds its view from these, so that
// the leaf loop carries the 64 lane registers and little else.
// Phase timestamps of block 0 (tuning aid, compiled only with -DTP_TRI_TIMING; read by tp_tri_times_m<M>_d<D>).
#ifdef TP_TRI_TIMING
static __device__ unsigned long long tp_tri_times[16];
#define TP_TRI_STAMP(i)                                                      \
    do {                                                                     \
        if (blockIdx.x == 0 && threadIdx.x == 0) {                           \
            unsigned long long t__;                                          \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));          \
            tp_tri_times[i] = t__;                                           \
        }                                                                    \
    } while (0)
// slots 9 / 10: earliest entry and latest exit over ALL blocks (10 must be zeroed by the reader between runs)
#define TP_TRI_STAMP_ALL(i, op)                                              \
    do {                                                                     \
        if (threadIdx.x == 0) {                                              \
            unsigned long long t__;                                          \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));          \
            op(&tp_tri_times[i], t__);                                       \
        }                                                                    \
    } while (0)
#else
#define TP_TRI_STAMP(i) ((void)0)
#define TP_TRI_STAMP_ALL(i, op) ((void)0)
#endif

struct TpTriBlock {
    TpSplitShared sh;
    TpSmem s;
    const uint4 *tbl;       // the S3 table: leaf pairs 0..gcap-1
    const uint16_t *t0;     // u in reversed positions (level 0 of T)
    unsigned char *warp0;   // per-warp areas
    int per_warp;
    int gcap;
    int c_req, jt_hi, n_entries, b_star;
    unsigned total_items;
    unsigned *g_bound_inv; // see tp_tri_setup
};
static __shared__ TpTriBlock tp_tri_blk;
static __shared__ int tp_tri_bound;                          // best Family score seen by the block
static __shared__ uint16_t tp_tri_pair_start[TP_TRI_GC + 1]; // first chunk of every leaf pair
// per leaf pair g, in bytes: table of leaf 2g+1, distance to the table of leaf 2g+2, first chunk of row pair g
static __shared__ uint4 tp_tri_pair_info[TP_TRI_GC];
// the plan's per-pivot tables (item -> pivot lookups happen once per item)
static __shared__ uint32_t tp_tri_item_base[TP_MAX_N], tp_tri_n_chunks[TP_MAX_N];
static __shared__ uint8_t tp_tri_pieces[TP_MAX_N];
static __shared__ uint16_t tp_tri_cum[TP_MAX_N + 2];

// byte area of a warp (32 bytes): [0..7] walked positions, [8], [9] bounds on the second position of
// the item, [10] the item's pivot, [12..15] its chunk of lane sets, [16..23] subsets scored so far
#define TP_TRI_WB_JT 10
#define TP_TRI_WB_AC 12
#define TP_TRI_WB_COUNT 16
#define TP_TRI_WB_BYTES 32

// Rank -> the p-1 offsets above the pivot (colexicographic, largest first), positions one per byte,
// descending, pivot last.
template <class BinomFn>
__device__ __forceinline__ void tp_tri_unrank(BinomFn C, int mA, int p, int jt, u64 rank, u64 &a0, u64 &a1) {
    a0 = 0;
    a1 = 0;
    int cnt = 0, hi = mA;
    for (int i = p - 1; i >= 1; --i) {
        int lo = i - 1, h = hi; // C(lo,i) <= rank < C(h,i)
        while (h - lo > 1) {
            const int mid = (lo + h) >> 1;
            if (C(mid, i) <= rank) lo = mid;
            else h = mid;
        }
        rank -= C(lo, i);
        hi = lo;
        const u64 pos = (u64)(jt + 1 + lo);
        if (cnt < 8) a0 |= pos << (8 * cnt);
        else a1 |= pos << (8 * (cnt - 8));
        cnt++;
    }
    if (cnt < 8) a0 |= (u64)jt << (8 * cnt);
    else a1 |= (u64)jt << (8 * (cnt - 8));
}

// The exact re-walk of one leaf for this lane (candidates, and leaves without a table): everything
// it needs is read back from shared memory, the lane's A is unranked again.
template <int M, int DW>
__device__ __noinline__ void tp_tri_slow(int b_leaf, int base, TpKey *best) {
    const TpTriBlock &blk = tp_tri_blk;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *wb = blk.warp0 + (size_t)warp * blk.per_warp;
    TpSplitWarp w;
    w.tstack = (uint16_t *)wb;
    w.va = (uint32_t *)(wb + (size_t)(DW + 1) * M * 2);
    w.bstack = (uint8_t *)(w.va + (M / 8) * 32 * 4);
    w.lane = lane;
    const int jt = w.bstack[TP_TRI_WB_JT];
    const unsigned ac = *(const unsigned *)(w.bstack + TP_TRI_WB_AC);
    const int kcols = blk.sh.k + 1;
    const u64 *binom = blk.s.binom;
    auto C = [&](int a, int b) -> u64 { return (b < 0 || a < b) ? 0ull : binom[a * kcols + b]; };
    const int mA = blk.sh.n - 1 - jt;
    const u64 NA = C(mA, blk.sh.p - 1);
    u64 rank = (u64)ac * 32 + lane;
    w.active = rank < NA;
    if (!w.active) return;
    tp_tri_unrank(C, mA, blk.sh.p, jt, rank, w.a0, w.a1);
    tp_split_slow<M>(blk.sh, blk.s, w, DW + 1, b_leaf, b_leaf, base, &tp_tri_bound, *best);
}

// Offsets inside a warp's area: [DW+1][M] T vectors, [M/8][32] lane vectors (16 bytes each), byte area.
#define TP_TRI_VA_OFF(M, DW) ((unsigned)((DW) + 1) * (M) * 2u)
#define TP_TRI_WB_OFF(M, DW) (TP_TRI_VA_OFF(M, DW) + (unsigned)((M) / 8) * 512u)


struct TpTriHot { // what the walk keeps in registers: 32-bit shared addresses and a few numbers
    unsigned wb;  // the warp's area
    unsigned wp;  // plain weight table
    unsigned tbl; // S3 table
    int rsp;
    int gcap_active; // table cap, negated for a lane without a real A
    int lane;
};

__device__ __forceinline__ void tp_sts32(unsigned addr, unsigned v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void tp_sts8(unsigned addr, unsigned v) {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned tp_lds32v(unsigned addr) { // volatile: the walk rewrites what it reads
    unsigned v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ unsigned tp_lds16v(unsigned addr) {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ unsigned tp_lds8v(unsigned addr) {
    unsigned v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

// Row pairs H..0 of a pass, straight-line.
template <int H>
struct TpTriRows {
    static __device__ __forceinline__ void run(const unsigned (&VT)[TP_TRI_NREG], unsigned tA, unsigned tB,
                                               unsigned &runA, unsigned &runB, unsigned one) {
        tp_tri_rowpair<H>(VT, tA, tB, runA, runB, one);
        if constexpr (H > 0) TpTriRows<H - 1>::run(VT, tA, tB, runA, runB, one);
    }
};

// Leaf pairs below TP_TRI_GT are straight-line code with compile-time positions (their tables sit
// at constant offsets, their own VT terms are registers): an average leaf holds only ~60-100
// pairs, so for the small passes the work AROUND the rows (dispatch, table lookup, the lane's own
// terms, loop) would otherwise weigh as much as the rows.  The passes below TP_TRI_GB form one
// basic block with a single test at its end, so that their table loads can all be issued ahead.
// (More straight-line passes cost instruction-cache reach: 12 passes are ~27 KB of code.)
#ifndef TP_TRI_GT
#define TP_TRI_GT 8
#endif
#ifndef TP_TRI_GB
#define TP_TRI_GB 4
#endif

// A compile-time pass: leaves 2G+2 (A) and 2G+1 (B), folded into the running minimum xm; passes at
// or above TP_TRI_GB, and pass 0, end with the test of what has accumulated.
#define TP_TRI_PASS(G)                                                                                  \
    case G:                                                                                             \
        if constexpr ((G) < TP_TRI_GT) {                                                                \
            constexpr unsigned OFFB = (unsigned)tp_tri_pair_base(G) * 16u;                              \
            constexpr unsigned OFFA = OFFB + (unsigned)tp_tri_off((G) + 1) * 16u;                       \
            unsigned runA = TP_BIG2 + TP_BIG2, runB = TP_BIG2 + TP_BIG2;                                \
            TpTriRows<(G)>::run(VT, tbl + OFFA, tbl + OFFB, runA, runB, one);                           \
            xm = __vimin3_u16x2(xm, runA + VT[2 * (G) + 2], runB + VT[2 * (G) + 1]);                    \
            if constexpr ((G) >= TP_TRI_GB || (G) == 0) {                                               \
                if (__vminu2(xm, room2) != room2) {                                                     \
                    g_to = (G);                                                                         \
                    g = min(g, (G) >= TP_TRI_GB ? (G) : TP_TRI_GB - 1);                                 \
                    goto rare;                                                                          \
                }                                                                                       \
                xm = 0xFFFFFFFFu;                                                                       \
            }                                                                                           \
        }                                                                                               \
        [[fallthrough]];

// All leaves b in [lo, hi) of one walked prefix.  va, tl, tbl: shared addresses of the lane's vectors
// (chunk c at va + 512 c), of the prefix's T vector and of the S3 table; base: the score of
// everything above the triple (c_req, inside A, A x walked, inside walked, unary terms); gcap_active:
// table cap, negated for a lane without a real A.  Inlined into the kernel itself (only an entry
// function may use all 128 registers; everything cold around it is a call).
template <int M, int DW>
__device__ __forceinline__ void tp_tri_leaves(unsigned va, unsigned tl, unsigned tbl, int gcap_active, int hi, int lo,
                                              int base, TpKey *best) {
    // The lane registers: positions below hi (at most the 64 that have registers); the rest is never
    // read.  They are rebuilt after every call of the slow path rather than kept alive across it.
    unsigned VT[TP_TRI_NREG];
    const int vt_chunks = min((hi + 7) >> 3, TP_TRI_NREG / 8);
    const int gcap = abs(gcap_active);
    auto vt_at = [&](int j) -> int { // (V + T)[j] of this lane: two 16-bit loads
        return (int)tp_lds16(va + (unsigned)((j >> 3) * 512 + (j & 7) * 2)) + (int)tp_lds16(tl + 2u * (unsigned)j);
    };
    const unsigned one = (unsigned)(*(volatile int *)&tp_tri_blk.gcap >> 30) + 1u; // 1, but not to the compiler
    const int room = *(volatile int *)&tp_tri_bound - base + 1; // a total below this is a candidate
    // (a lane without a real A never reports: its threshold is 0)
    const unsigned room2 = gcap_active > 0 ? (unsigned)min(max(room, 0), 0xFFFF) * 0x10001u : 0u;
    int g = (hi - 2) >> 1; // leaves 2g+2 (A; may be hi, not a leaf of this range) and 2g+1 (B; may be lo - 1)
    int g_to = 0;          // on the way to `rare`: passes g .. g_to reported
    unsigned xA = 0, xB = 0;
    for (;;) { // (re-entered after a call of the slow path)
        switch (vt_chunks) {
#if TP_TRI_GC > 24
            TP_TRI_VT_CASE(7) TP_TRI_VT_CASE(6)
#endif
            TP_TRI_VT_CASE(5) TP_TRI_VT_CASE(4) TP_TRI_VT_CASE(3) TP_TRI_VT_CASE(2) TP_TRI_VT_CASE(1) TP_TRI_VT_CASE(0)
        default: break;
        }
        // ---- the generic pass: big leaves, ranges that do not reach down to leaf 2, and a first
        // pass whose leaf A is not part of the range ----------------------------------------------
        for (; 2 * g + 2 >= lo && (g >= TP_TRI_GT || lo > 2 || 2 * g + 2 >= hi); --g) {
            const int bA = 2 * g + 2, bB = 2 * g + 1;
            g_to = g;
            if (g >= gcap) { // no table up here: the exact re-walk prunes rows against the bound
                xA = xB = 0u;
                goto rare;
            }
            const uint4 pi = tp_tri_pair_info[g];
            const unsigned tB = tbl + pi.x, tA = tB + pi.y;
            unsigned runA = TP_BIG2 + TP_BIG2, runB = TP_BIG2 + TP_BIG2;
#if TP_TRI_PREFETCH
            const unsigned off_g = pi.z;
            uint4 bufA[2], bufB[2]; // (the entry chunk goes to both sets: which one row pair g reads is a compile-time fact)
            bufA[0] = bufA[1] = tp_lds128v(tA + off_g);
            bufB[0] = bufB[1] = tp_lds128v(tB + off_g);
#endif
            switch (g) {
#if TP_TRI_GC > 24
                TP_TRI_CASE8(24)
#endif
                TP_TRI_CASE8(16) TP_TRI_CASE8(8) TP_TRI_CASE8(0)
            default: break;
            }
            // the leaves' own terms (bA == hi is not a leaf, and can be one past the lane vector)
            const int vbA = vt_at(min(bA, hi - 1)), vbB = vt_at(bB);
            xA = bA < hi ? runA + (unsigned)vbA * 0x10001u : 0xFFFFFFFFu;
            xB = bB >= lo ? runB + (unsigned)vbB * 0x10001u : 0xFFFFFFFFu;
            if (__vminu2(__vminu2(xA, xB), room2) != room2) goto rare;
        }
        if (2 * g + 2 < lo) return;
        { // ---- leaves 2g+2 .. 2 in one straight line (all of them leaves of the range) -----------
            unsigned xm = 0xFFFFFFFFu;
            xA = xB = 0u; // should a block report, every leaf of it is re-walked
            switch (g) {
                TP_TRI_PASS(15) TP_TRI_PASS(14) TP_TRI_PASS(13) TP_TRI_PASS(12)
                TP_TRI_PASS(11) TP_TRI_PASS(10) TP_TRI_PASS(9) TP_TRI_PASS(8) TP_TRI_PASS(7) TP_TRI_PASS(6)
                TP_TRI_PASS(5) TP_TRI_PASS(4) TP_TRI_PASS(3) TP_TRI_PASS(2) TP_TRI_PASS(1) TP_TRI_PASS(0)
            default: break;
            }
            return;
        }
    rare: // passes g .. g_to reported (or have no table): re-walk their leaves exactly, then go on below
        for (; g >= g_to; --g) {
            if (__vminu2(xA, room2) != room2 && 2 * g + 2 < hi && 2 * g + 2 >= lo)
                tp_tri_slow<M, DW>(2 * g + 2, base + vt_at(2 * g + 2), best);
            if (__vminu2(xB, room2) != room2 && 2 * g + 1 >= lo)
                tp_tri_slow<M, DW>(2 * g + 1, base + vt_at(2 * g + 1), best);
        }
        if (2 * g + 2 < lo) return;
    }
}

// The warp-uniform walk over the DW positions above the triple; level L picks b in [lo, hi).
template <int M, int DW, int L>
struct TpTriWalk {
    static __device__ __forceinline__ void run(const TpTriHot &w, int hi, int lo, int lb, int sb, TpKey *best) {
        const unsigned tl = w.wb + (unsigned)(L * M * 2);              // T_L
        const unsigned va = w.wb + TP_TRI_VA_OFF(M, DW) + 16u * w.lane; // this lane's V, chunk c at + 512 c
        if constexpr (L == DW) {
            if (hi > lo) tp_tri_leaves<M, DW>(va, tl, w.tbl, w.gcap_active, hi, lo, lb + sb, best);
        } else {
            const unsigned bst = w.wb + TP_TRI_WB_OFF(M, DW);
            for (int b = hi - 1; b >= lo; --b) {
                const unsigned pk = tp_lds32v(va + (unsigned)((b >> 3) * 512 + ((b >> 1) & 3) * 4));
                const int lb2 = lb + (int)((b & 1) ? (pk >> 16) : (pk & 0xFFFF));
                const int sb2 = sb + (int)tp_lds16v(tl + 2u * b);
                // T_{L+1} = T_L + W[b] on the positions below b (the rest is never read)
                const unsigned dst = tl + M * 2, rw = w.wp + (unsigned)(b * w.rsp) * 2u;
                __syncwarp(); // lanes still in the slow path of an earlier leaf read T_{L+1} and the byte area
                if (w.lane == 0) tp_sts8(bst + L, (unsigned)b);
                // (the same trip count for every lane: a lane-dependent one leaves the warp split in two
                // groups that would each run the whole leaf stage below on their own)
#pragma unroll
                for (int w0 = 0; w0 < M / 2; w0 += 32) {
                    const int wd = w0 + w.lane;
                    if (wd < M / 2) tp_sts32(dst + 4u * wd, tp_lds32v(tl + 4u * wd) + tp_lds32v(rw + 4u * wd));
                }
                __syncwarp();
                // below the walked positions sit three more, so level L + 1 starts at DW + 2 - (L + 1)
                if constexpr (L == 0) // the item may hold only a sub-slice of the second position
                    TpTriWalk<M, DW, 1>::run(w, min(b, (int)tp_lds8v(bst + 9)), max(DW + 1, (int)tp_lds8v(bst + 8)), lb2, sb2,
                                             best);
                else
                    TpTriWalk<M, DW, L + 1>::run(w, b, DW + 1 - L, lb2, sb2, best);
            }
        }
    }
};

// Build the S3 table of this request from the plain table wp (reversed positions), one leaf per
// warp at a time, one 16-byte chunk per lane.  hmap[c] = row pair of chunk c (the same for every leaf).
static __device__ void tp_tri_build_table(uint4 *tbl, const uint16_t *pair_base, const uint8_t *hmap, int gcap,
                                          const uint16_t *wp, int rsp) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    // every turn of a warp takes one leaf from the light end of the list and one from the heavy end
    auto fill = [&](int lf) { // leaf lf + 1: 2g+1 first, then 2g+2
        const int g = lf >> 1, b = lf + 1;
        const int nchunks = tp_tri_off(g + 1);
        uint4 *dst = tbl + pair_base[g] + ((lf & 1) ? nchunks : 0);
        const uint16_t *rb = wp + b * rsp;
        for (int c = lane; c < nchunks; c += 32) {
            const int h = hmap[c], e0 = (c - tp_tri_off(h)) * 4;
            const int r0 = 2 * h, r1 = 2 * h + 1;
            const bool has1 = r1 < b;
            const uint16_t *w0 = wp + r0 * rsp, *w1 = wp + r1 * rsp;
            const unsigned c0 = rb[r0], c1 = has1 ? rb[r1] : 0;
            unsigned out[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int e = e0 + i;
                const unsigned be = rb[e];
                const unsigned lo = e < r0 ? w0[e] + be + c0 : TP_BIG;
                const unsigned hi = (has1 && e < r1) ? w1[e] + be + c1 : TP_BIG;
                out[i] = lo | (hi << 16);
            }
            dst[c] = make_uint4(out[0], out[1], out[2], out[3]);
        }
    };
    for (int i = warp; i < gcap; i += nw) {
        fill(i);
        fill(2 * gcap - 1 - i);
    }
}

// Per-warp shared memory: T vectors of the DW + 1 levels, the lane vectors, the byte area.
template <int M, int DW>
static __host__ __device__ inline size_t tp_tri_per_warp() {
    return (size_t)(DW + 1) * M * 2 + (size_t)(M / 8) * 32 * 16 + TP_TRI_WB_BYTES;
}

// Set-up of one work item (pivot, chunk of 32 lane sets, slice) for the calling warp: the lane's A,
// its vector V in shared memory, the slice bounds in the warp's byte area.  A call (like everything
// cold), reading its constants back from shared memory.  Returns false for an empty slice.
struct TpTriItem {
    int b0_lo, b0_hi, sA, active;
};
template <int M, int DW>
__device__ __noinline__ bool tp_tri_item_setup(unsigned item, TpTriItem *out) {
    const TpTriBlock &blk = tp_tri_blk;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blk.sh.n, p = blk.sh.p, q = blk.sh.q, rsp = blk.sh.rsp;
    const uint16_t *wp = blk.sh.wp, *t0 = blk.t0;
    unsigned char *wb = blk.warp0 + (size_t)warp * blk.per_warp;
    uint32_t *va = (uint32_t *)(wb + (size_t)(DW + 1) * M * 2); // [M/8][32][4]
    uint8_t *bstack = (uint8_t *)(va + (M / 8) * 32 * 4);       // TP_TRI_WB_BYTES
    const int kcols = blk.sh.k + 1;
    const u64 *binom = blk.s.binom;
    auto C = [&](int a, int b) -> u64 { return (b < 0 || a < b) ? 0ull : binom[a * kcols + b]; };
    // ---- item -> (pivot, chunk of A's, slice of the first position below the lane set) ----
    int e = 0;
    const int n_entries = blk.n_entries;
    for (int bs = 0; bs < n_entries; bs += 32) {
        const int idx = bs + lane;
        const bool ok = idx < n_entries && tp_tri_item_base[idx] <= item;
        const unsigned m = __ballot_sync(0xFFFFFFFFu, ok);
        if (m) e = bs + 31 - __clz(m);
    }
    const int jt = blk.jt_hi - e;
    const unsigned local = item - tp_tri_item_base[e];
    const unsigned nchunks = tp_tri_n_chunks[e];
    const unsigned slice = local / nchunks;
    const unsigned ac = local - slice * nchunks;
    TpSlice sl{q - 1, jt, 0, TP_MAX_N, false};
    {
        auto cum = [&](int b) -> int { return tp_tri_cum[b]; };
        if (!tp_split_slice(C, cum, jt, q, (int)tp_tri_pieces[e], blk.b_star, (int)slice, sl)) return false; // empty
    }
    __syncwarp(); // the previous item's slow-path lanes are done with the warp's areas
    if (lane == 0 && blk.g_bound_inv) { // what other blocks found so far, and what this one did
        const int theirs = (int)(0x7FFFFFFFu - *(volatile unsigned *)blk.g_bound_inv);
        const int ours = *(volatile int *)&tp_tri_bound;
        if (theirs < ours) atomicMin(&tp_tri_bound, theirs);
        else if (ours < theirs) atomicMax(blk.g_bound_inv, 0x7FFFFFFFu - (unsigned)ours);
    }
    {
        u64 per_lane; // sets below the lane set that this item scores per lane
        if (sl.cut_again) per_lane = C(sl.b1_hi, q - 1) - C(sl.b1_lo, q - 1);
        else per_lane = C(sl.b0_hi, q) - C(sl.b0_lo, q);
        if (lane == 0) {
            const u64 na = C(n - 1 - jt, p - 1);
            *(u64 *)(bstack + TP_TRI_WB_COUNT) += per_lane * (u64)min((u64)32, na - (u64)ac * 32);
            bstack[8] = (uint8_t)sl.b1_lo;
            bstack[9] = (uint8_t)sl.b1_hi;
            bstack[TP_TRI_WB_JT] = (uint8_t)jt;
            *(unsigned *)(bstack + TP_TRI_WB_AC) = ac;
        }
    }
    // ---- this lane's A ----------------------------------------------------------------------
    const int mA = n - 1 - jt;
    const u64 NA = C(mA, p - 1);
    u64 rank = (u64)ac * 32 + lane;
    const int active = rank < NA;
    if (!active) rank = NA - 1;
    u64 a0, a1;
    tp_tri_unrank(C, mA, p, jt, rank, a0, a1);
    auto elem = [&](int i) { return (int)(((i < 8 ? a0 : a1) >> (8 * (i & 7))) & 0xFF); };
    // ---- V = sum of the A rows (positions < jt) -> shared memory; sA = score inside A ----
    int sA = blk.c_req;
    {
        int rowoff[TP_SPLIT_MAX_P];
#pragma unroll
        for (int i = 0; i < TP_SPLIT_MAX_P; i++) {
            if (i >= p) break;
            rowoff[i] = elem(i) * rsp;
        }
        const int nch = (jt + 7) >> 3;
        for (int c = 0; c < nch; c++) {
            uint4 v = make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int i = 0; i < TP_SPLIT_MAX_P; i++) {
                if (i >= p) break;
                const uint4 r = ((const uint4 *)(wp + rowoff[i]))[c];
                v.x += r.x;
                v.y += r.y;
                v.z += r.z;
                v.w += r.w;
            }
            ((uint4 *)va)[c * 32 + lane] = v;
        }
#pragma unroll
        for (int i = 0; i < TP_SPLIT_MAX_P; i++) {
            if (i >= p) break;
            sA += t0[elem(i)];
            for (int i2 = 0; i2 < i; i2++) sA += wp[rowoff[i] + elem(i2)];
        }
    }
    __syncwarp();
    out->b0_lo = sl.b0_lo;
    out->b0_hi = sl.b0_hi;
    out->sA = sA;
    out->active = active;
    return true;
}

// Exact seed, spread over (mask of whole groups, partial group) units: every Family member is one
// unit (h == n_groups stands for "no partial group").  Thread `first` of `stride` takes units
// first, first + stride, ...: the whole block for a batch request, the whole GRID for a single
// request (each block then scores a 1/gridDim share and publishes its best through global memory).
static __device__ __noinline__ int tp_tri_family_seed(const TpSmem s, int n_groups, int k, int c_req, unsigned first,
                                                      unsigned stride) {
    int best = 0x7FFFFFFF;
    if (n_groups > TP_SEED_MAX_GROUPS) return best;
    const unsigned per = (unsigned)n_groups + 1u, total = (1u << n_groups) * per;
    for (unsigned unit = first; unit < total; unit += stride) {
        const unsigned mask = unit / per;
        const int h = (int)(unit - mask * per);
        if (h < n_groups && ((mask >> h) & 1u)) continue;
        int size = 0;
        for (unsigned m = mask; m && size <= k; m &= m - 1) size += s.gsize[__ffs(m) - 1];
        const int r = k - size; // devices still to take: a proper prefix of group h
        if (r < 0 || (h == n_groups ? r != 0 : (r == 0 || s.gsize[h] <= r))) continue;
        int pos[TP_MAX_K];
        int cnt = 0;
        for (unsigned m = mask; m; m &= m - 1) {
            const int g = __ffs(m) - 1;
            for (int t = 0; t < s.gsize[g]; t++) pos[cnt++] = s.gstart[g] + t;
        }
        for (int t = 0; t < r; t++) pos[cnt++] = s.gstart[h] + t; // the prefix (order does not matter here)
        int sco = c_req;
        for (int i = 0; i < cnt; i++) {
            sco += s.u[pos[i]];
            for (int j = 0; j < i; j++) sco += s.w[pos[j] * s.stride + pos[i]];
        }
        best = min(best, sco);
    }
    return best;
}

// Once per request and block: stage the request, build the tables, seed the bound.  Two
// barriers: after (1) the descriptor's small tables + (2) the weights in canonical order (gathered
// from global memory straight into the byte table and the 16-bit reversed table), and after (3)
// the S3 table and the seeds.  g_bound_inv: a global word shared by the grid (single request) through
// which blocks exchange the best Family score they know, stored as 0x7FFFFFFF - score so that a
// zeroed word means "none"; null for a batch request, where the block owns the request.
template <int M, int DW>
__device__ __noinline__ void tp_tri_setup(const uint8_t *__restrict__ W, int nd, const u64 *__restrict__ binom_g,
                                          const TpReq *__restrict__ rq, const TpPlan *__restrict__ plan,
                                          unsigned *g_bound_inv) {
    unsigned char *const smem = tp_split_smem;
    __shared__ uint8_t s_hmap[tp_tri_off(TP_TRI_GC)];
    const int n = rq->n, k = rq->k, n_req = rq->n_req;
    const int gcap = plan->tri_gcap;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, tid = threadIdx.x, nt = blockDim.x;
    // layout: the S3 table first (+ a zero the compiler cannot see, so that its base lives in a
    // register and rows are addressed [base + constant]), then the staged request, the plain
    // table, u, and the per-warp areas
    uint4 *tbl = (uint4 *)smem + (plan->p >> 24);
    size_t off = tp_tri_table_bytes(gcap);
    const TpSmem s = tp_smem_carve(smem + off, n, k);
    off += (tp_smem_bytes(n, k) + 15) & ~(size_t)15;
    const int rsp = tp_split_rsp(n);
    uint16_t *wp = (uint16_t *)(smem + off);
    off += (size_t)n * rsp * 2;
    uint16_t *t0 = (uint16_t *)(smem + off); // u in reversed positions, all n of them (level 0 of T)
    off += (size_t)TP_MAX_N * 2;
    const size_t per_warp = tp_tri_per_warp<M, DW>();
    unsigned char *warp_base = smem + off + (size_t)warp * per_warp;
    uint16_t *tstack = (uint16_t *)warp_base; // [DW+1][M]
    uint8_t *bstack = warp_base + (size_t)(DW + 1) * M * 2 + (size_t)(M / 8) * 32 * 16;
    TP_TRI_STAMP(1);
    // ---- (1) everything that is a plain copy or a formula ------------------------------------
    if (tid == 0) {
        tp_tri_blk.sh = TpSplitShared{wp, nullptr, n, k, plan->p, plan->q, rq->dbits, rsp, plan->min_w};
        tp_tri_blk.s = s;
        tp_tri_blk.tbl = tbl;
        tp_tri_blk.t0 = t0;
        tp_tri_blk.warp0 = smem + off;
        tp_tri_blk.per_warp = (int)per_warp;
        tp_tri_blk.gcap = gcap;
        tp_tri_blk.c_req = rq->c_req;
        tp_tri_blk.jt_hi = plan->jt_hi;
        tp_tri_blk.n_entries = plan->n_entries;
        tp_tri_blk.b_star = plan->b_star;
        tp_tri_blk.total_items = plan->total_items;
        tp_tri_blk.g_bound_inv = g_bound_inv;
        tp_tri_bound = 0x7FFFFFFF;
    }
    if (lane == 0) *(u64 *)(bstack + TP_TRI_WB_COUNT) = 0;
    for (int i = tid; i < n; i += nt) {
        s.group[i] = rq->group_of[i];
        s.gstart[i] = rq->gstart[i];
        s.gsize[i] = rq->gsize[i];
    }
    for (int a = warp; a <= n; a += TP_TRI_WARPS)
        for (int j = lane; j <= k; j += 32) s.binom[a * (k + 1) + j] = binom_g[a * TP_BINOM_COLS + j];
    for (int i = tid; i < plan->n_entries; i += nt) {
        tp_tri_item_base[i] = plan->item_base[i];
        tp_tri_n_chunks[i] = plan->n_chunks[i];
        tp_tri_pieces[i] = plan->n_pieces[i];
    }
    for (int i = tid; i < TP_MAX_N + 2; i += nt) tp_tri_cum[i] = plan->cum[i];
    for (int g = tid; g <= gcap; g += nt) tp_tri_pair_start[g] = (uint16_t)tp_tri_pair_base(g);
    for (int g = tid; g < gcap; g += nt)
        tp_tri_pair_info[g] = make_uint4(16u * tp_tri_pair_base(g), 16u * tp_tri_off(g + 1), 16u * tp_tri_off(g), 0u);
    for (int h = tid; h < gcap; h += nt)
        for (int c = tp_tri_off(h); c < tp_tri_off(h + 1); c++) s_hmap[c] = (uint8_t)h;
    TP_TRI_STAMP(2);
    // ---- (2) weights: canonical byte table (seeds) and 16-bit reversed table (everything else); the
    // canonical order is read from the descriptor itself, so that no barrier separates (1) and (2) ----
    for (int r = warp; r < n; r += TP_TRI_WARPS) { // one row per warp pass: no divisions
        const uint8_t *wrow = W + (int)rq->canon[r] * nd;
        uint16_t *prow = wp + (n - 1 - r) * rsp;
        for (int c = lane; c < rsp; c += 32) { // reversed column n - 1 - c holds canonical column c
            const unsigned v = (c < n && c != r) ? wrow[rq->canon[c]] : 0u;
            if (c < n) {
                s.w[r * s.stride + c] = (uint8_t)v;
                prow[n - 1 - c] = (uint16_t)v;
            } else {
                prow[c] = 0; // padding (it is summed)
            }
        }
    }
    for (int i = tid; i < TP_MAX_N; i += nt) {
        unsigned u = 0;
        if (i < n) {
            const uint8_t *wrow = W + (int)rq->canon[i] * nd;
            for (int r = 0; r < n_req; r++) u += wrow[rq->req[r]];
            s.u[i] = (uint16_t)u;
            t0[n - 1 - i] = (uint16_t)u;
        } else {
            t0[i] = 0;
        }
    }
    __syncthreads();
    TP_TRI_STAMP(3);
    // ---- (3) seeds (a few threads; Family members all) and the S3 table (everybody) ----------
    if (tid == nt - 1 || tid == nt - 2) {
        // canonical positions {0..k-1}, and whole groups from the largest down
        int left = k, sco = rq->c_req, cnt = 0;
        int pos[TP_MAX_K];
        if (tid == nt - 1) {
            for (int i = 0; i < k; i++) pos[cnt++] = i;
            left = 0;
        } else {
            for (int g = rq->n_groups - 1; g >= 0 && left > 0; g--) {
                const int take = min((int)s.gsize[g], left);
                for (int t = 0; t < take; t++) pos[cnt++] = s.gstart[g] + t;
                left -= take;
            }
        }
        for (int i = 0; i < cnt; i++) {
            sco += s.u[pos[i]];
            for (int j = 0; j < i; j++) sco += s.w[pos[j] * s.stride + pos[i]];
        }
        if (left == 0) atomicMin(&tp_tri_bound, sco);
    }
    {
        // A share per block only where the search is long enough for the other blocks' shares to arrive
        // early in it (until then the two cheap seeds above bound it); a small search is mostly over by
        // then, and with a loose bound most of its leaves would go through the slow path.
        const bool spread = g_bound_inv && rq->total > (1ull << 28);
        const unsigned first = spread ? blockIdx.x * (unsigned)nt + tid : (unsigned)tid;
        const unsigned stride = spread ? gridDim.x * (unsigned)nt : (unsigned)nt;
        const int sco = tp_tri_family_seed(s, rq->n_groups, k, rq->c_req, first, stride);
        if (sco != 0x7FFFFFFF) atomicMin(&tp_tri_bound, sco);
    }
    TP_TRI_STAMP(4);
    tp_tri_build_table(tbl, tp_tri_pair_start, s_hmap, gcap, wp, rsp);
    TP_TRI_STAMP(5);
    for (int i = lane; i < M; i += 32) tstack[i] = t0[i];
    __syncthreads();
    if (tid == 0 && g_bound_inv) atomicMax(g_bound_inv, 0x7FFFFFFFu - (unsigned)tp_tri_bound);
}

// Single request (batch.count == 0): the whole grid shares the request and its item counter, the
// last block reduces the partials.  Batch: every block takes whole requests off batch.order.
template <int M, int DW>
__global__ void __launch_bounds__(TP_TRI_THREADS, 1)
tp_exhaustive_tri(const uint8_t *__restrict__ W, int nd, const u64 *__restrict__ binom_g,
                  const TpReq *__restrict__ reqs, const TpPlan *__restrict__ plans, int shard_idx, int shard_cnt,
                  TpScratch sc, unsigned *work_counter, TpBatch batch) {
    __shared__ unsigned s_entry, s_items;
    for (int iter = 0;; iter++) {
        int slot = 0, plan = 0, out = 0;
        unsigned *wc = work_counter;
        if (batch.count > 0) {
            __syncthreads(); // everybody is done with the previous request's shared memory
            if (threadIdx.x == 0) {
                s_entry = atomicAdd(batch.grab, 1u);
                s_items = 0;
            }
            __syncthreads();
            if (s_entry >= (unsigned)batch.count) break;
            const TpBatchEntry en = batch.order[s_entry];
            slot = en.slot;
            plan = en.plan;
            out = en.out;
            shard_idx = en.share_idx;
            shard_cnt = en.share_cnt;
            wc = &s_items;
        } else if (iter > 0) {
            break;
        }
        TP_TRI_STAMP(0);
        TP_TRI_STAMP_ALL(9, atomicMin);
        TP_TRI_STAMP_ALL(11, atomicMax);
        tp_tri_setup<M, DW>(W, nd, binom_g, reqs + slot, plans + plan, batch.count > 0 ? nullptr : work_counter + 1);
        TP_TRI_STAMP(6);
        TP_TRI_STAMP_ALL(12, atomicMax);
        TpKey best = {TP_KEY_INF_HI, TP_KEY_INF_HI}; // (its address goes to the callees: it lives in the frame)
        const int lane = threadIdx.x & 31;
        // the warp's areas, by 32-bit shared address
        const unsigned wb = (unsigned)__cvta_generic_to_shared(tp_tri_blk.warp0) +
                            (unsigned)(threadIdx.x >> 5) * (unsigned)tp_tri_per_warp<M, DW>();
        for (;;) {
            unsigned t_idx = 0;
            if (lane == 0) t_idx = atomicAdd(wc, 1u);
            t_idx = __shfl_sync(0xFFFFFFFFu, t_idx, 0);
            // shard s of P owns item i when (i + i / P) mod P == s (see tp_split.cuh)
            const unsigned rot = (unsigned)shard_idx + (unsigned)shard_cnt - t_idx % (unsigned)shard_cnt;
            const u64 item64 = (u64)t_idx * (unsigned)shard_cnt + rot % (unsigned)shard_cnt;
            if (item64 >= tp_tri_blk.total_items) break;
            TpTriItem it;
            if (!tp_tri_item_setup<M, DW>((unsigned)item64, &it)) continue;
            const TpTriHot hot{wb, (unsigned)__cvta_generic_to_shared(tp_tri_blk.sh.wp),
                               (unsigned)__cvta_generic_to_shared(tp_tri_blk.tbl), tp_tri_blk.sh.rsp,
                               it.active ? tp_tri_blk.gcap : -tp_tri_blk.gcap, lane};
            if constexpr (DW == 0) // the slice bounds the top of the triple itself
                TpTriWalk<M, 0, 0>::run(hot, it.b0_hi, max(it.b0_lo, 2), it.sA, 0, &best);
            else
                TpTriWalk<M, DW, 0>::run(hot, it.b0_hi, max(it.b0_lo, DW + 2), it.sA, 0, &best);
            __syncwarp();
        }
        __syncwarp();
        TP_TRI_STAMP(7);
        TP_TRI_STAMP_ALL(13, atomicMax);
        u64 counted = 0;
        if (lane == 0) {
            unsigned clo, chi;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(clo), "=r"(chi) : "r"(wb + TP_TRI_WB_OFF(M, DW) + TP_TRI_WB_COUNT));
            counted = ((u64)chi << 32) | clo;
        }
        if (batch.count > 0) {
            tp_block_reduce(best, counted);
            if (threadIdx.x == 0) {
                batch.results[out].key = best;
                batch.results[out].scored = counted;
            }
        } else {
            // the last block to finish re-arms the two words of work_counter for the next launch (no memset in
            // the launcher: one stream operation less per search)
            if (tp_block_finish(best, counted, sc) && threadIdx.x == 0) {
                work_counter[0] = 0u;
                work_counter[1] = 0u;
            }
            TP_TRI_STAMP(8);
            TP_TRI_STAMP_ALL(10, atomicMax);
        }
    }
}

template <int M, int DW>
static size_t tp_tri_smem_bytes_md(int n, int k, int gcap) {
    size_t off = tp_tri_table_bytes(gcap) + ((tp_smem_bytes(n, k) + 15) & ~(size_t)15);
    off += (size_t)n * tp_split_rsp(n) * 2 + (size_t)TP_MAX_N * 2;
    off += (size_t)TP_TRI_WARPS * tp_tri_per_warp<M, DW>();
    return off;
}

// h describes the request (single mode) or carries the largest n, k and table cap of the batch.
template <int M, int DW>
static cudaError_t launch_tri_md(const uint8_t *W, int nd, const u64 *binom, const TpReq *d_rq, const TpReq &h,
                                 const TpPlan *d_plan, int gcap, int shard_idx, int shard_cnt, const TpScratch &sc,
                                 unsigned *d_counter, int sm_count, cudaStream_t st, TpLaunchInfo *info,
                                 const TpBatch &batch) {
    const size_t smem = tp_tri_smem_bytes_md<M, DW>(h.n, h.k, gcap);
    cudaError_t e = cudaFuncSetAttribute(tp_exhaustive_tri<M, DW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    if (batch.count > 0) { // (a single request's counter words are re-armed by the kernel itself)
        e = cudaMemsetAsync(batch.grab, 0, sizeof(unsigned), st);
        if (e != cudaSuccess) return e;
    }
    int blocks = sm_count;
    if (batch.count > 0 && batch.count < blocks) blocks = batch.count;
    tp_exhaustive_tri<M, DW><<<blocks, TP_TRI_THREADS, smem, st>>>(W, nd, binom, d_rq, d_plan, shard_idx, shard_cnt,
                                                                   sc, d_counter, batch);
    if (info) {
        info->launches = 1;
        info->blocks = blocks;
        info->threads = TP_TRI_THREADS;
    }
    return cudaGetLastError();
}

#ifdef TP_TRI_TIMING
#define TP_TRI_TIMES_READER(M, D)                                                                        \
    extern "C" int tp_tri_times_m##M##_d##D(unsigned long long *out) {                                   \
        int rc = (int)cudaMemcpyFromSymbol(out, tp_tri_times, sizeof(tp_tri_times));                     \
        unsigned long long reset[16];                                                                    \
        for (int i = 0; i < 16; i++) reset[i] = i == 9 ? ~0ull : 0ull;                                   \
        cudaMemcpyToSymbol(tp_tri_times, reset, sizeof(reset));                                          \
        return rc;                                                                                       \
    }
#else
#define TP_TRI_TIMES_READER(M, D)
#endif
// One translation unit per (M, DW) defines these for its instantiation.
#define TP_DEFINE_TRI_LAUNCHER(M, D)                                                                             \
    cudaError_t tp_launch_tri_m##M##_d##D(const uint8_t *W, int nd, const unsigned long long *binom,            \
                                          const TpReq *d_rq, const TpReq &h, const TpPlan *d_plan, int gcap,    \
                                          int si, int sn, const TpScratch &sc, unsigned *d_counter, int sms,    \
                                          cudaStream_t st, TpLaunchInfo *info, const TpBatch &batch) {          \
        return launch_tri_md<M, D>(W, nd, binom, d_rq, h, d_plan, gcap, si, sn, sc, d_counter, sms, st, info,   \
                                   batch);                                                                       \
    }                                                                                                            \
    size_t tp_tri_smem_m##M##_d##D(int n, int k, int gcap) { return tp_tri_smem_bytes_md<M, D>(n, k, gcap); }   \
    TP_TRI_TIMES_READER(M, D)
