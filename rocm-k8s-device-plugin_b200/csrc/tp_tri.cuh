// tp_tri.cuh -- kernel TRI: the exhaustive scorer for k' >= 6 on up to 80 + p free devices.
// Included by tp_tri_m*_d*.cu, one translation unit per (lane-vector width M, walked depth DW >= 1).
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
// The two 16-bit halves of every lane register belong to TWO PREFIXES: the last walked position
// takes the values b0 and b0 - 1 at once (VT.lo = V + T(.., b0), VT.hi = V + T(.., b0 - 1)), and a
// table word holds the same S3 entry in both halves.  Then, for leaf b and row jx,
//
//     acc = min(acc, VT[jy] + word(b, jx, jy))     one VIADDMNMX.U16x2 = the same triple for both prefixes
//     run = min(run, acc + VT[jx])                 the closing op of the row: no repacking of anything
//
// Every prefix-level cost (the T update, building VT, the leaf loop's set-up and tests) is paid once per
// two prefixes, rows are exactly as long as they are (no interleaved row pairs with a diagonal and a missing
// half), and a leaf is a pass of its own (no pairs of leaves with an odd one out).  Leaf b0 - 1 exists for the
// first prefix only: T(.., b0 - 1) holds TP_BIG at position b0 - 1, which keeps the other half out of every
// minimum.  A last single value of the walked range runs with its upper half switched off (threshold 0).
//
// A leaf's table is its rows 1 .. b-1 back to back (row jx: jx words at word C(jx,2)); rows are consumed in
// pairs (2H+1, 2H), top down, as two independent chains.  An odd leaf has no row b = 2H+1 in its top pair: b
// words of TP_BIG follow its table, so that the pair's code needs no special case.
//
// Shared memory decides how many leaves get a table: C(L+1,3) words for leaves up to L, 76 KB for L = 48
// (every leaf for n - p <= 50).  Leaves above the cap (0.06 % of the subsets of the 8-of-80 search) go to
// the exact scalar re-walk that also serves candidates (tp_split_slow), which prunes whole rows against the bound.
#pragma once
#include <type_traits>
#include "tp_split.cuh"

#ifndef TP_TRI_WARPS
#define TP_TRI_WARPS 16
#endif
#define TP_TRI_THREADS (TP_TRI_WARPS * 32)
#ifndef TP_TRI_GC
#define TP_TRI_GC 24 /* row pairs with code: leaves up to 2 GC have a table, positions below 2 GC a lane register */
#endif
#define TP_TRI_NREG (2 * TP_TRI_GC)
#define TP_TRI_LMAX (2 * TP_TRI_GC) /* the largest leaf with a table, at most */

__host__ __device__ constexpr int tp_tri_c2(int x) { return x * (x - 1) / 2; }
// Words of leaf b's table: rows 1 .. b-1, b words of TP_BIG after an odd leaf, rounded up to 16 bytes.
__host__ __device__ constexpr int tp_tri_leaf_words(int b) { return (tp_tri_c2(b) + ((b & 1) ? b : 0) + 3) & ~3; }
// ... and where it starts (in words): leaves 2, 3, ... back to back.
__host__ __device__ constexpr int tp_tri_leaf_off(int b) {
    int s = 0;
    for (int i = 2; i < b; i++) s += tp_tri_leaf_words(i);
    return s;
}
// (the same sums as a table, for the kernel's set-up)
struct TpTriLeafOffsets {
    uint16_t words[TP_TRI_LMAX + 2];
    constexpr TpTriLeafOffsets() : words() {
        int s = 0;
        for (int b = 2; b < TP_TRI_LMAX + 2; b++) {
            words[b] = (uint16_t)s;
            s += tp_tri_leaf_words(b);
        }
    }
};
static __constant__ TpTriLeafOffsets tp_tri_leaf_offsets = TpTriLeafOffsets();
static __host__ __device__ inline size_t tp_tri_table_bytes(int gcap) { return (size_t)tp_tri_leaf_off(2 * gcap + 1) * 4; }
// Row jx inside a leaf: 16-byte chunks c0 .. c1 hold its words (the first and the last one shared with its neighbours).
__host__ __device__ constexpr int tp_tri_row_c0(int jx) { return tp_tri_c2(jx) / 4; }
__host__ __device__ constexpr int tp_tri_row_nch(int jx) { return jx < 1 ? 0 : (tp_tri_c2(jx) + jx - 1) / 4 - tp_tri_row_c0(jx) + 1; }
// Rows of one parity form a stream of chunks, top row first; this is where row jx's first chunk stands in it
// (its low bit picks the buffer of the software pipeline below).
__host__ __device__ constexpr int tp_tri_row_seq(int jx) {
    int s = 0;
    for (int r = jx + 2; r < TP_TRI_LMAX; r += 2) s += tp_tri_row_nch(r);
    return s;
}

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

// f(integral_constant<int, I>) for I in [I0, I1): compile-time indices for registers and table offsets
template <int I0, int I1, class F>
__device__ __forceinline__ void tp_static_for(F &&f) {
    if constexpr (I0 < I1) {
        f(std::integral_constant<int, I0>{});
        tp_static_for<I0 + 1, I1>(f);
    }
}

// The words of chunk C (of a leaf's table) that belong to row JX, against the lane registers.
template <int JX, int C>
__device__ __forceinline__ void tp_tri_row_chunk(const unsigned (&VT)[TP_TRI_NREG], const uint4 &q, unsigned &acc) {
    constexpr int J0 = 4 * C - tp_tri_c2(JX); // jy of the chunk's first word
    if constexpr (J0 + 0 >= 0 && J0 + 0 < JX) acc = __viaddmin_u16x2(VT[J0 + 0], q.x, acc);
    if constexpr (J0 + 1 >= 0 && J0 + 1 < JX) acc = __viaddmin_u16x2(VT[J0 + 1], q.y, acc);
    if constexpr (J0 + 2 >= 0 && J0 + 2 < JX) acc = __viaddmin_u16x2(VT[J0 + 2], q.z, acc);
    if constexpr (J0 + 3 >= 0 && J0 + 3 < JX) acc = __viaddmin_u16x2(VT[J0 + 3], q.w, acc);
}

// One row of a leaf whose table starts at shared address tl, loads where they are needed (the
// straight-line small leaves: the compiler schedules across rows and leaves).
template <int JX>
__device__ __forceinline__ void tp_tri_row(const unsigned (&VT)[TP_TRI_NREG], unsigned tl, unsigned &run) {
    unsigned acc = TP_BIG2 + TP_BIG2;
    tp_static_for<0, tp_tri_row_nch(JX)>([&](auto ic) {
        constexpr int C = tp_tri_row_c0(JX) + decltype(ic)::value;
        const uint4 q = tp_lds128(tl + C * 16);
        tp_tri_row_chunk<JX, C>(VT, q, acc);
    });
    run = __viaddmin_u16x2(acc, VT[JX], run);
}

// Rows 2H+1 (chain A) and 2H (chain B) of the leaf at tl, software-pipelined: the generic pass enters its rows
// through a switch, so every row pair is a basic block of its own and the compiler cannot start a row pair's
// loads during the one before it.  Each chain keeps two buffers; chunk i of row jx sits in buffer
// (tp_tri_row_seq(jx) + i) & 1 and the next chunk of the chain (of this row, or the first of row jx - 2) is
// loaded by hand before this one is used.
template <int H>
__device__ __forceinline__ void tp_tri_rowpair_pf(const unsigned (&VT)[TP_TRI_NREG], unsigned tl, unsigned &runA,
                                                  unsigned &runB, uint4 (&bA)[2], uint4 (&bB)[2]) {
    constexpr int JA = 2 * H + 1, JB = 2 * H;
    constexpr int NA = tp_tri_row_nch(JA), NB = tp_tri_row_nch(JB);
    constexpr int SA = tp_tri_row_seq(JA), SB = tp_tri_row_seq(JB);
    unsigned accA = TP_BIG2 + TP_BIG2, accB = TP_BIG2 + TP_BIG2;
    tp_static_for<0, (NA > NB ? NA : NB)>([&](auto ic) {
        constexpr int I = decltype(ic)::value;
        if constexpr (I < NA) {
            constexpr int CUR = (SA + I) & 1;
            if constexpr (I + 1 < NA) bA[CUR ^ 1] = tp_lds128v(tl + (tp_tri_row_c0(JA) + I + 1) * 16);
            else if constexpr (H > 0) bA[CUR ^ 1] = tp_lds128v(tl + tp_tri_row_c0(JA - 2) * 16);
            tp_tri_row_chunk<JA, tp_tri_row_c0(JA) + I>(VT, bA[CUR], accA);
        }
        if constexpr (I < NB) {
            constexpr int CUR = (SB + I) & 1;
            if constexpr (I + 1 < NB) bB[CUR ^ 1] = tp_lds128v(tl + (tp_tri_row_c0(JB) + I + 1) * 16);
            else if constexpr (H > 1) bB[CUR ^ 1] = tp_lds128v(tl + tp_tri_row_c0(JB - 2) * 16);
            tp_tri_row_chunk<JB, tp_tri_row_c0(JB) + I>(VT, bB[CUR], accB);
        }
    });
    runA = __viaddmin_u16x2(accA, VT[JA], runA);
    if constexpr (H > 0) runB = __viaddmin_u16x2(accB, VT[JB], runB);
}

#define TP_TRI_CASE(H)                                                                         \
    case H:                                                                                    \
        if constexpr ((H) < TP_TRI_GC) tp_tri_rowpair_pf<(H)>(VT, tl, runA, runB, bufA, bufB); \
        [[fallthrough]];
#define TP_TRI_CASE8(B)                                                                          \
    TP_TRI_CASE(B + 7) TP_TRI_CASE(B + 6) TP_TRI_CASE(B + 5) TP_TRI_CASE(B + 4) TP_TRI_CASE(B + 3) \
    TP_TRI_CASE(B + 2) TP_TRI_CASE(B + 1) TP_TRI_CASE(B + 0)

// Eight positions of VT: low halves V + T', high halves V + T'' (T'' stands TP_TRI_T2 bytes after T').
template <int C>
__device__ __forceinline__ void tp_tri_vt_chunk(unsigned (&VT)[TP_TRI_NREG], unsigned v4, unsigned t4, unsigned t2_off) {
    const uint4 v = tp_lds128v(v4 + C * 512); // lane-private: [chunk][lane]
    const uint4 t = tp_lds128v(t4 + C * 16);
    const uint4 s = tp_lds128v(t4 + t2_off + C * 16);
    const unsigned x0 = v.x + t.x, x1 = v.y + t.y, x2 = v.z + t.z, x3 = v.w + t.w;
    const unsigned y0 = v.x + s.x, y1 = v.y + s.y, y2 = v.z + s.z, y3 = v.w + s.w;
    VT[8 * C + 0] = __byte_perm(x0, y0, 0x5410);
    VT[8 * C + 1] = __byte_perm(x0, y0, 0x7632);
    VT[8 * C + 2] = __byte_perm(x1, y1, 0x5410);
    VT[8 * C + 3] = __byte_perm(x1, y1, 0x7632);
    VT[8 * C + 4] = __byte_perm(x2, y2, 0x5410);
    VT[8 * C + 5] = __byte_perm(x2, y2, 0x7632);
    VT[8 * C + 6] = __byte_perm(x3, y3, 0x5410);
    VT[8 * C + 7] = __byte_perm(x3, y3, 0x7632);
}
#define TP_TRI_VT_CASE(C)                                                            \
    case (C) + 1:                                                                    \
        if constexpr (8 * (C) < TP_TRI_NREG) tp_tri_vt_chunk<(C)>(VT, va, t1, 2u * M); \
        [[fallthrough]];

// What the hot loop does not need in registers lives here: the block's tables (written once per
// request) and, per warp, the item it is on.  The slow path rebuilds its view from these, so that
// the leaf loop carries the lane registers and little else.
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
    const uint32_t *tbl;    // the S3 table: leaves 2 .. 2 gcap
    const uint16_t *t0;     // u in reversed positions (level 0 of T)
    unsigned char *warp0;   // per-warp areas
    int per_warp;
    int gcap;
    int c_req, jt_hi, n_entries, b_star;
    unsigned total_items;
    unsigned *g_bound_inv; // see tp_tri_setup
};
static __shared__ TpTriBlock tp_tri_blk;
static __shared__ int tp_tri_bound; // best Family score seen by the block
// per leaf b, in bytes: its table, and inside it the first chunk of the two rows of its top pair (2H+1, 2H; H = (b-1)/2)
static __shared__ uint4 tp_tri_leaf_info[TP_TRI_LMAX + 1];
// the plan's per-pivot tables (item -> pivot lookups happen once per item)
static __shared__ uint32_t tp_tri_item_base[TP_MAX_N], tp_tri_n_chunks[TP_MAX_N];
static __shared__ uint8_t tp_tri_pieces[TP_MAX_N];
static __shared__ uint16_t tp_tri_cum[TP_MAX_N + 2];
// Slice boundaries, computed once per request (tp_split_cut costs a 64-bit division and a binary search, and an
// item needs two of them): the np + 1 boundaries of pivot entry e start at tp_tri_cut_base[e]; the leaf cuts of the
// pair with head b (one walked position: tp_split_bstar) are tp_tri_lcuts[b][0 .. parts].  tp_tri_cuts_ok == 0: the
// boundaries did not fit, items compute theirs.
#define TP_TRI_MAX_CUTS 1536
#define TP_TRI_MAX_HEADS (TP_TRI_MAX_M + 8)
static __shared__ uint16_t tp_tri_cut_base[TP_MAX_N + 1];
static __shared__ uint8_t tp_tri_cuts[TP_TRI_MAX_CUTS];
static __shared__ uint8_t tp_tri_lcuts[TP_TRI_MAX_HEADS][TP_SPLIT_MAX_SUB + 1];
static __shared__ int tp_tri_cuts_ok;
static __shared__ int tp_tri_next_leaf; // table build: the next leaf to fill (counted down from the largest)
// C(m, i) for i = 1 .. 4, saturated: unranking a lane set of up to five positions in 32-bit arithmetic
static __shared__ uint32_t tp_tri_c32[4][TP_MAX_N + 1];

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

// The same for p - 1 <= 4 and ranks below 2^32, from the 32-bit table (the per-item path).
__device__ __forceinline__ void tp_tri_unrank32(int mA, int p, int jt, unsigned rank, u64 &a0, u64 &a1) {
    unsigned lo32 = 0, hi32 = 0;
    int cnt = 0, hi = mA;
    for (int i = p - 1; i >= 1; --i) {
        const uint32_t *row = tp_tri_c32[i - 1];
        int lo; // the largest one with C(lo,i) <= rank (it is below hi)
        if (i <= 3) { // closed-form estimate, then a step or two along the table
            lo = i == 1 ? (int)rank
                        : i == 2 ? (int)((1.0f + sqrtf(1.0f + 8.0f * (float)rank)) * 0.5f) : (int)cbrtf(6.0f * (float)rank) + 1;
            lo = min(max(lo, i - 1), hi - 1);
            while (row[lo] > rank) lo--;
            while (lo + 1 < hi && row[lo + 1] <= rank) lo++;
        } else {
            lo = i - 1;
            int h = hi; // C(lo,i) <= rank < C(h,i)
            while (h - lo > 1) {
                const int mid = (lo + h) >> 1;
                if (row[mid] <= rank) lo = mid;
                else h = mid;
            }
        }
        rank -= row[lo];
        hi = lo;
        const unsigned pos = (unsigned)(jt + 1 + lo);
        if (cnt < 4) lo32 |= pos << (8 * cnt);
        else hi32 |= pos << (8 * (cnt - 4));
        cnt++;
    }
    if (cnt < 4) lo32 |= (unsigned)jt << (8 * cnt);
    else hi32 |= (unsigned)jt << (8 * (cnt - 4));
    a0 = ((u64)hi32 << 32) | lo32;
    a1 = 0;
}

// Offsets inside a warp's area: T vectors of the DW walk levels, then T' and T'' of the prefix pair ([DW+2][M]),
// [M/8][32] lane vectors (16 bytes each), byte area.
#define TP_TRI_VA_OFF(M, DW) ((unsigned)((DW) + 2) * (M) * 2u)
#define TP_TRI_WB_OFF(M, DW) (TP_TRI_VA_OFF(M, DW) + (unsigned)((M) / 8) * 512u)

// The exact re-walk of one leaf for this lane and one prefix of the pair (half 0: last walked position as stored,
// T'; half 1: one below, T''): candidates, and leaves without a table.  Everything it needs is read back from
// shared memory, the lane's A is unranked again.
template <int M, int DW>
__device__ __noinline__ void tp_tri_slow(int b_leaf, int base, int half, TpKey *best) {
    const TpTriBlock &blk = tp_tri_blk;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *wb = blk.warp0 + (size_t)warp * blk.per_warp;
    const uint8_t *area = wb + TP_TRI_WB_OFF(M, DW);
    uint8_t walked[8];
#pragma unroll
    for (int i = 0; i < 8; i++) walked[i] = area[i];
    walked[DW - 1] = (uint8_t)(walked[DW - 1] - half);
    TpSplitWarp w;
    w.tstack = (uint16_t *)wb + half * M; // level DW of this view: T' or T''
    w.va = (uint32_t *)(wb + TP_TRI_VA_OFF(M, DW));
    w.bstack = walked;
    w.lane = lane;
    const int jt = area[TP_TRI_WB_JT];
    const unsigned ac = *(const unsigned *)(area + TP_TRI_WB_AC);
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

// Leaves up to TP_TRI_LT are straight-line code with compile-time positions (their tables sit at
// constant offsets, their own VT terms are registers): an average leaf holds only ~60-100 pairs, so
// for the small leaves the work AROUND the rows (dispatch, table lookup, the lane's own terms, loop)
// would otherwise weigh as much as the rows.  The leaves below TP_TRI_LB form one basic block with a
// single test at its end, so that their table loads can all be issued ahead.
// (More straight-line leaves cost instruction-cache reach.)
#ifndef TP_TRI_LT
#define TP_TRI_LT 14
#endif
#ifndef TP_TRI_LB
#define TP_TRI_LB 8
#endif

// Rows JX .. 1 of the leaf at tl, odd rows on chain A and even rows on chain B.
template <int JX>
struct TpTriRows {
    static __device__ __forceinline__ void run(const unsigned (&VT)[TP_TRI_NREG], unsigned tl, unsigned &runA,
                                               unsigned &runB) {
        if constexpr (JX >= 1) {
            tp_tri_row<JX>(VT, tl, (JX & 1) ? runA : runB);
            TpTriRows<JX - 1>::run(VT, tl, runA, runB);
        }
    }
};

// A straight-line leaf B, folded into the running minimum xm; leaves at or above TP_TRI_LB, and leaf 2,
// end with the test of what has accumulated.
#define TP_TRI_PASS(B)                                                                                  \
    case B:                                                                                             \
        if constexpr ((B) <= TP_TRI_LT) {                                                               \
            unsigned runA = TP_BIG2 + TP_BIG2, runB = TP_BIG2 + TP_BIG2;                                \
            TpTriRows<(B) - 1>::run(VT, tbl + (unsigned)tp_tri_leaf_off(B) * 4u, runA, runB);           \
            xm = __vimin3_u16x2(xm, runA + VT[(B)], runB + VT[(B)]);                                    \
            if constexpr ((B) >= TP_TRI_LB || (B) == 2) {                                               \
                if (__vminu2(xm, room2) != room2) {                                                     \
                    b_to = (B);                                                                         \
                    b = min(b, (B) >= TP_TRI_LB ? (B) : TP_TRI_LB - 1);                                 \
                    goto rare;                                                                          \
                }                                                                                       \
                xm = 0xFFFFFFFFu;                                                                       \
            }                                                                                           \
        }                                                                                               \
        [[fallthrough]];

// ---- the quad stage: the values below TP_TRI_QB of the last walked position, without walking them ------------
// Most prefixes are small (on 64 devices 42 % of the prefix pairs have their walked value below 9, with less
// packed work than the ~330 instructions a pair costs before its first row), so for the values b0 < TP_TRI_QB the
// walked position joins the table: S4 = S3[b][jx][jy] + W[b0][b] + W[b0][jx] + W[b0][jy], low half for b0, high
// half for b0 - 1 (TP_BIG where b = b0 - 1, and for the head 3, which has no partner), one table per pair with an
// odd head b0 = 3, 5, .. < TP_TRI_QB, leaves 2 .. b0 - 1 in the row layout of the S3 leaves.  The lane registers
// then hold VU = V + T_L SPLATTED (they do not depend on b0 any more: built once for the whole range), and
//     score = base + VU[b0] + VU[b] + VU[jx] + VU[jy] + S4 .
// A pair that reports is walked again by the ordinary code (T', T'', the leaf loop), which finds the candidate.
#ifndef TP_TRI_QB
#define TP_TRI_QB 8 /* even; 0 switches the stage off.  Measured 0 / 8 / 10 / 12 / 16: 8 of 80 1.445 / 1.414 / 1.447 /
                      1.437 / 1.503 ms, block-per-request 8-of-48 searches 1.24 / 1.31 / 1.31 / 1.33 / 1.32e13 subsets/s,
                      8 of 64 unchanged (0.282 ms): the stage's code competes for the same instruction cache */
#endif
__host__ __device__ constexpr int tp_tri_q_leaf_words(int b) { return (tp_tri_c2(b) + 3) & ~3; }
// word offset of leaf b in the table of the pair with the odd head b0
__host__ __device__ constexpr int tp_tri_q_off(int b0, int b) {
    int s = 0;
    for (int h = 3; h < b0; h += 2)
        for (int l = 2; l < h; l++) s += tp_tri_q_leaf_words(l);
    for (int l = 2; l < b; l++) s += tp_tri_q_leaf_words(l);
    return s;
}
#define TP_TRI_Q_WORDS (TP_TRI_QB >= 4 ? tp_tri_q_off(TP_TRI_QB + 1, 2) : 4)
static __shared__ __align__(16) uint32_t tp_tri_q4[TP_TRI_Q_WORDS];

// One pair (B0, B0 - 1): the best total below the walked position per half, the pair's own terms added.
template <int B0>
__device__ __forceinline__ unsigned tp_tri_qpair(const unsigned (&Q)[TP_TRI_NREG], unsigned qt) {
    unsigned xp = 0xFFFFFFFFu;
    tp_static_for<2, B0>([&](auto ib) {
        constexpr int B = decltype(ib)::value;
        unsigned runA = TP_BIG2 + TP_BIG2, runB = TP_BIG2 + TP_BIG2;
        TpTriRows<B - 1>::run(Q, qt + (unsigned)tp_tri_q_off(B0, B) * 4u, runA, runB);
        xp = __vimin3_u16x2(xp, runA + Q[B], runB + Q[B]);
    });
    return xp + __byte_perm(Q[B0], Q[B0 - 1], 0x5410);
}
#define TP_TRI_QCASE(B0)                                                        \
    case B0:                                                                    \
        if constexpr ((B0) < TP_TRI_QB) {                                       \
            if (__vminu2(tp_tri_qpair<(B0)>(Q, qt), room2) != room2) bits |= 1u << ((B0) >> 1); \
        }                                                                       \
        [[fallthrough]];
// The pairs with heads q_top, q_top - 2, .., 3; bit (head >> 1) of the result: this lane has a candidate in that pair.
// va, tl: this lane's V and the level's T vector; room2: the threshold in both halves.
__device__ __forceinline__ unsigned tp_tri_quads(unsigned va, unsigned tl, unsigned room2, int q_top) {
    unsigned Q[TP_TRI_NREG]; // positions 0 .. TP_TRI_QB - 1 are used
    {
        const uint4 v0 = tp_lds128v(va), v1 = tp_lds128v(va + 512), t0 = tp_lds128v(tl), t1 = tp_lds128v(tl + 16);
        const unsigned x[8] = {v0.x + t0.x, v0.y + t0.y, v0.z + t0.z, v0.w + t0.w,
                               v1.x + t1.x, v1.y + t1.y, v1.z + t1.z, v1.w + t1.w};
#pragma unroll
        for (int m = 0; m < 8; m++) {
            Q[2 * m] = __byte_perm(x[m], 0, 0x1010);
            Q[2 * m + 1] = __byte_perm(x[m], 0, 0x3232);
        }
    }
    const unsigned qt = (unsigned)__cvta_generic_to_shared(tp_tri_q4);
    unsigned bits = 0;
    switch (q_top) {
        TP_TRI_QCASE(15) TP_TRI_QCASE(13) TP_TRI_QCASE(11) TP_TRI_QCASE(9) TP_TRI_QCASE(7) TP_TRI_QCASE(5) TP_TRI_QCASE(3)
    default: break;
    }
    return bits;
}

// The S4 tables of the quad stage from the plain table wp (reversed positions): a thread per (pair, leaf).
static __device__ void tp_tri_build_quads(const uint16_t *wp, int rsp, int n_pos) {
    for (int c = threadIdx.x; c < (TP_TRI_QB / 2) * TP_TRI_QB; c += blockDim.x) {
        const int h = 3 + 2 * (c / TP_TRI_QB), b = c % TP_TRI_QB; // head of the pair, leaf
        if (h >= TP_TRI_QB || b < 2 || b >= h) continue;
        uint32_t *dst = tp_tri_q4 + tp_tri_q_off(h, b);
        const bool hi_ok = h - 1 >= 3 && b < h - 1 && h < n_pos; // the second prefix of the pair exists and has this leaf
        const uint16_t *r0 = wp + h * rsp, *r1 = wp + (h - 1) * rsp, *rb = wp + b * rsp;
        int w = 0;
        for (int jx = 1; jx < b; jx++)
            for (int jy = 0; jy < jx; jy++, w++) {
                const unsigned s3 = (unsigned)rb[jx] + rb[jy] + wp[jx * rsp + jy];
                const unsigned lo = h < n_pos ? s3 + r0[b] + r0[jx] + r0[jy] : TP_BIG;
                const unsigned hi = hi_ok ? s3 + r1[b] + r1[jx] + r1[jy] : TP_BIG;
                dst[w] = lo | (hi << 16);
            }
        for (; w < tp_tri_q_leaf_words(b); w++) dst[w] = TP_BIG2;
    }
}

// All leaves b in [lo_leaf, hi_leaf) of a pair of walked prefixes, hi the walked value of the first one (leaves are
// below it; the second prefix is without leaf hi - 1); the range is all of [2, hi) except in an item that holds one
// walked value cut by its leaves.  va, t1, tbl:
// shared addresses of the lane's vectors (chunk c at va + 512 c), of the pair's T vectors (T' at t1, T'' 2 M bytes
// on) and of the S3 table; base0, base1: the score of everything above the triple (c_req, inside A, A x walked,
// inside walked, unary terms) for the two prefixes, base1 < 0 when there is no second prefix; gcap_active: table
// cap, negated for a lane without a real A.  Inlined into the kernel itself (only an entry function may use all
// 128 registers; everything cold around it is a call).
template <int M, int DW>
__device__ __forceinline__ void tp_tri_leaves(unsigned va, unsigned t1, unsigned tbl, int gcap_active, int hi, int hi_leaf,
                                              int lo_leaf, int base0, int base1, TpKey *best) {
    // The lane registers: positions below hi (at most the 48 that have registers); the rest is never
    // read.  They are rebuilt after every call of the slow path rather than kept alive across it.
    unsigned VT[TP_TRI_NREG];
    const int vt_chunks = min((hi_leaf + 7) >> 3, TP_TRI_NREG / 8);
    const int lmax = 2 * abs(gcap_active);
    auto vt2_at = [&](int j) -> unsigned { // (V + T')[j] | (V + T'')[j] << 16 of this lane: three 16-bit loads
        const unsigned v = tp_lds16(va + (unsigned)((j >> 3) * 512 + (j & 7) * 2));
        return (v + tp_lds16(t1 + 2u * (unsigned)j)) | ((v + tp_lds16(t1 + 2u * M + 2u * (unsigned)j)) << 16);
    };
    // a total below its half of room2 is a candidate (real totals stay below TP_BIG, switched-off ones do not);
    // a lane without a real A, and the missing second prefix, never report: threshold 0
    const int bound = *(volatile int *)&tp_tri_bound;
    const unsigned room_lo = (unsigned)min(max(bound - base0 + 1, 0), (int)TP_BIG - 1);
    const unsigned room_hi = base1 >= 0 ? (unsigned)min(max(bound - base1 + 1, 0), (int)TP_BIG - 1) : 0u;
    const unsigned room2 = gcap_active > 0 ? (room_lo | (room_hi << 16)) : 0u;
    int b = hi_leaf - 1; // the next leaf
    int b_to = 2;   // on the way to `rare`: leaves b .. b_to reported
    unsigned x = 0; // what they reported (0: all of them, both prefixes)
    for (;;) {      // (re-entered after a call of the slow path)
        switch (vt_chunks) {
#if TP_TRI_GC > 24
            TP_TRI_VT_CASE(7) TP_TRI_VT_CASE(6)
#endif
            TP_TRI_VT_CASE(5) TP_TRI_VT_CASE(4) TP_TRI_VT_CASE(3) TP_TRI_VT_CASE(2) TP_TRI_VT_CASE(1) TP_TRI_VT_CASE(0)
        default: break;
        }
        // ---- the generic pass: one leaf above TP_TRI_LT, rows in pairs through a switch ----------
        for (; b > TP_TRI_LT && b >= lo_leaf; --b) {
            b_to = b;
            if (b > lmax) { // no table up here: the exact re-walk prunes rows against the bound
                x = 0u;
                goto rare;
            }
            const uint4 li = tp_tri_leaf_info[b];
            const unsigned tl = tbl + li.x;
            unsigned runA = TP_BIG2 + TP_BIG2, runB = TP_BIG2 + TP_BIG2;
            uint4 bufA[2], bufB[2]; // (the entry chunks go to both buffers: which one the top pair reads is a compile-time fact)
            bufA[0] = bufA[1] = tp_lds128v(tl + li.y);
            bufB[0] = bufB[1] = tp_lds128v(tl + li.z);
            switch ((b - 1) >> 1) {
#if TP_TRI_GC > 24
                TP_TRI_CASE8(24)
#endif
                TP_TRI_CASE8(16) TP_TRI_CASE8(8) TP_TRI_CASE8(0)
            default: break;
            }
            x = __vminu2(runA, runB) + vt2_at(b);
            if (__vminu2(x, room2) != room2) goto rare;
        }
        if (b < lo_leaf) return;
        { // ---- leaves b .. 2 in one straight line (a range that starts among them rescans the few below it) ----
            unsigned xm = 0xFFFFFFFFu;
            x = 0u; // should a block report, every leaf of it is re-walked
            switch (b) {
                TP_TRI_PASS(16) TP_TRI_PASS(15) TP_TRI_PASS(14) TP_TRI_PASS(13) TP_TRI_PASS(12)
                TP_TRI_PASS(11) TP_TRI_PASS(10) TP_TRI_PASS(9) TP_TRI_PASS(8) TP_TRI_PASS(7) TP_TRI_PASS(6)
                TP_TRI_PASS(5) TP_TRI_PASS(4) TP_TRI_PASS(3) TP_TRI_PASS(2)
            default: break;
            }
            return;
        }
    rare: // leaves b .. b_to reported (or have no table): re-walk them exactly, then go on below
        for (; b >= b_to; --b) {
            const unsigned vtb = vt2_at(b);
            if (min(x & 0xFFFFu, room2 & 0xFFFFu) != (room2 & 0xFFFFu))
                tp_tri_slow<M, DW>(b, base0 + (int)(vtb & 0xFFFFu), 0, best);
            if (min(x >> 16, room2 >> 16) != (room2 >> 16) && b < hi - 1)
                tp_tri_slow<M, DW>(b, base1 + (int)(vtb >> 16), 1, best);
        }
        if (b < lo_leaf) return;
    }
}

// The warp-uniform walk over the DW positions above the triple; level L picks b in [lo, hi).  The last
// level takes its values two at a time.
template <int M, int DW, int L>
struct TpTriWalk {
    static __device__ __forceinline__ void run(const TpTriHot &w, int hi, int lo, int lb, int sb, TpKey *best) {
        static_assert(DW >= 1 && L < DW, "the pair of prefixes is made at the last walked level");
        const unsigned tl = w.wb + (unsigned)(L * M * 2);              // T_L
        const unsigned va = w.wb + TP_TRI_VA_OFF(M, DW) + 16u * w.lane; // this lane's V, chunk c at + 512 c
        const unsigned bst = w.wb + TP_TRI_WB_OFF(M, DW);
        auto v_at = [&](int j) -> int { // V[j] of this lane
            const unsigned pk = tp_lds32v(va + (unsigned)((j >> 3) * 512 + ((j >> 1) & 3) * 4));
            return (int)((j & 1) ? (pk >> 16) : (pk & 0xFFFFu));
        };
        if constexpr (L == DW - 1) {
            const unsigned t1 = w.wb + (unsigned)(DW * M * 2); // T', and T'' after it
            // an item that holds a single first walked value may be cut by the position below it: here, the leaves
            const int lo_leaf = L == 0 ? max(2, (int)tp_lds8v(bst + 8)) : 2;
            const int cap_leaf = L == 0 ? (int)tp_lds8v(bst + 9) : TP_MAX_N;
            // The values below TP_TRI_QB go through the quad stage first (a range that starts at 3 only: the lowest
            // slice of a pivot); the pairs it reports are walked again below, after the rest of the range.
            unsigned qmask = 0;
            int lo_cur = lo;
            if (TP_TRI_QB >= 4 && lo <= 3 && hi >= 4) {
                const int top = min(hi, TP_TRI_QB) - 1, q_top = (top & 1) ? top : top - 1; // the largest odd head
                const int room = *(volatile int *)&tp_tri_bound - (lb + sb) + 1;
                const unsigned room2 = w.gcap_active > 0 ? (unsigned)min(max(room, 0), (int)TP_BIG - 1) * 0x10001u : 0u;
                qmask = __reduce_or_sync(0xFFFFFFFFu, tp_tri_quads(va, tl, room2, q_top));
                lo_cur = q_top + 1;
            }
            bool again = false; // walking the reported pairs of the quad stage
            for (int b = hi - 1;; b -= 2) {
                if (!again && b < lo_cur) again = true;
                if (again) {
                    if (!qmask) break;
                    const int t = 31 - __clz((int)qmask);
                    qmask ^= 1u << t;
                    b = 2 * t + 1;
                    lo_cur = t == 1 ? 3 : b - 1; // (the head 3 has no partner)
                }
                const int lo = lo_cur;
                const int b1 = b - 1; // (a position even when it is below lo: lo >= 3)
                const int base0 = lb + v_at(b) + sb + (int)tp_lds16v(tl + 2u * b);
                const int base1 = b1 >= lo ? lb + v_at(b1) + sb + (int)tp_lds16v(tl + 2u * b1) : -1;
                // T' = T_L + W[b], T'' = T_L + W[b - 1] on the positions below (the rest is never read), with
                // T''[b - 1] = TP_BIG: leaf b - 1 belongs to the first prefix only
                const unsigned rw0 = w.wp + (unsigned)(b * w.rsp) * 2u, rw1 = w.wp + (unsigned)(b1 * w.rsp) * 2u;
                __syncwarp(); // lanes still in the slow path of an earlier leaf read T', T'' and the byte area
                if (w.lane == 0) tp_sts8(bst + L, (unsigned)b);
                // (the same trip count for every lane: a lane-dependent one leaves the warp split in two
                // groups that would each run the whole leaf stage below on their own)
#pragma unroll
                for (int w0 = 0; w0 < M / 2; w0 += 32) {
                    const int wd = w0 + w.lane;
                    if (wd < M / 2) {
                        const unsigned t = tp_lds32v(tl + 4u * wd);
                        tp_sts32(t1 + 4u * wd, t + tp_lds32v(rw0 + 4u * wd));
                        unsigned u = t + tp_lds32v(rw1 + 4u * wd);
                        if (wd == (b1 >> 1)) u = (b1 & 1) ? ((u & 0xFFFFu) | (TP_BIG << 16)) : ((u & 0xFFFF0000u) | TP_BIG);
                        tp_sts32(t1 + 2u * M + 4u * wd, u);
                    }
                }
                __syncwarp();
                tp_tri_leaves<M, DW>(va, t1, w.tbl, w.gcap_active, b, min(b, cap_leaf), lo_leaf, base0, base1, best);
            }
        } else {
            for (int b = hi - 1; b >= lo; --b) {
                const int lb2 = lb + v_at(b);
                const int sb2 = sb + (int)tp_lds16v(tl + 2u * b);
                // T_{L+1} = T_L + W[b] on the positions below b (the rest is never read)
                const unsigned dst = tl + M * 2, rw = w.wp + (unsigned)(b * w.rsp) * 2u;
                __syncwarp(); // lanes still in the slow path of an earlier leaf read the T vectors and the byte area
                if (w.lane == 0) tp_sts8(bst + L, (unsigned)b);
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

// Build the S3 table of this request from the plain table wp (reversed positions): a warp per leaf, a 16-byte
// chunk per lane and turn.
static __device__ void tp_tri_build_table(uint32_t *tbl, int gcap, const uint16_t *wp, int rsp) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    auto fill = [&](int b) {
        uint4 *dst = (uint4 *)(tbl + (tp_tri_leaf_info[b].x >> 2));
        const int real = tp_tri_c2(b), chunks = tp_tri_leaf_words(b) >> 2;
        const uint16_t *rb = wp + b * rsp;
        for (int c = lane; c < chunks; c += 32) {
            const int w = 4 * c;
            int jx = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)w)) * 0.5f); // the row of word w: C(jx,2) <= w < C(jx+1,2)
            if (tp_tri_c2(jx) > w) jx--;
            else if (tp_tri_c2(jx + 1) <= w) jx++;
            int jy = w - tp_tri_c2(jx);
            unsigned out[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                unsigned v = TP_BIG; // (after the rows: the missing top row of an odd leaf, and padding)
                if (w + i < real) v = (unsigned)rb[jx] + rb[jy] + wp[jx * rsp + jy];
                out[i] = v * 0x10001u;
                if (++jy == jx) {
                    jx++;
                    jy = 0;
                }
            }
            dst[c] = make_uint4(out[0], out[1], out[2], out[3]);
        }
    };
    // warps take leaves off a counter, heaviest first (they arrive at different times: seeds, slice boundaries)
    const int top = 2 * gcap;
    for (;;) {
        int i = 0;
        if (lane == 0) i = atomicAdd(&tp_tri_next_leaf, 1);
        i = __shfl_sync(0xFFFFFFFFu, i, 0);
        if (top - i < 2) break;
        fill(top - i);
    }
    (void)warp;
    (void)nw;
}

// Per-warp shared memory: T vectors (DW walk levels + the pair), the lane vectors, the byte area.
template <int M, int DW>
static __host__ __device__ inline size_t tp_tri_per_warp() {
    return (size_t)(DW + 2) * M * 2 + (size_t)(M / 8) * 32 * 16 + TP_TRI_WB_BYTES;
}

// The lane vector V (positions below 8 nch) and the score inside A for a lane set of P <= 4 positions (one per
// byte of a32), by 32-bit shared addresses: the per-item path of every search up to k' = 10.
template <int P>
__device__ __forceinline__ int tp_tri_lane_vector(unsigned a32, unsigned wp_s, unsigned t0_s, unsigned va_lane, int rsp,
                                                  int nch, int c_req) {
    unsigned row[P], pos[P];
#pragma unroll
    for (int i = 0; i < P; i++) {
        pos[i] = (a32 >> (8 * i)) & 0xFFu;
        row[i] = wp_s + pos[i] * (unsigned)rsp * 2u;
    }
    for (int c = 0; c < nch; c++) {
        uint4 v = tp_lds128(row[0] + 16u * c);
#pragma unroll
        for (int i = 1; i < P; i++) {
            const uint4 r = tp_lds128(row[i] + 16u * c);
            v.x += r.x;
            v.y += r.y;
            v.z += r.z;
            v.w += r.w;
        }
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(va_lane + 512u * c), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                     : "memory");
    }
    int sA = c_req;
#pragma unroll
    for (int i = 0; i < P; i++) {
        sA += (int)tp_lds16(t0_s + 2u * pos[i]);
#pragma unroll
        for (int i2 = 0; i2 < i; i2++) sA += (int)tp_lds16(row[i] + 2u * pos[i2]);
    }
    return sA;
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
    uint32_t *va = (uint32_t *)(wb + TP_TRI_VA_OFF(M, DW));     // [M/8][32][4]
    uint8_t *bstack = (uint8_t *)(va + (M / 8) * 32 * 4);       // TP_TRI_WB_BYTES
    const int kcols = blk.sh.k + 1;
    const u64 *binom = blk.s.binom;
    auto C = [&](int a, int b) -> u64 { return (b < 0 || a < b) ? 0ull : binom[a * kcols + b]; };
    // the grid-wide bound word: read now, looked at after the slice arithmetic (a global load's latency apart)
    unsigned theirs_inv = 0;
    if (lane == 0 && blk.g_bound_inv) asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(theirs_inv) : "l"(blk.g_bound_inv));
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
        const int np = (int)tp_tri_pieces[e];
        if (tp_tri_cuts_ok && (int)slice < np) { // the boundaries are in the tables
            const int cb = tp_tri_cut_base[e] + (int)slice;
            sl.b0_lo = tp_tri_cuts[cb];
            sl.b0_hi = tp_tri_cuts[cb + 1];
            if (sl.b0_lo >= sl.b0_hi) return false; // empty
        } else if (tp_tri_cuts_ok && q == 4) { // a pair (b, b - 1) of single values, cut by the leaves (tp_split_slice)
            const int b_eff = tp_split_bstar(jt, blk.b_star, true), t = (int)slice - np;
            int lo = 0, hi = (jt - b_eff) >> 1;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if ((int)tp_tri_cum[b_eff + 1 + 2 * mid] <= t) lo = mid;
                else hi = mid;
            }
            const int b = b_eff + 1 + 2 * lo;
            const int nsub = (int)tp_tri_cum[b + 2] - (int)tp_tri_cum[b], sub = t - (int)tp_tri_cum[b];
            sl.b0_lo = b - 1;
            sl.b0_hi = b + 1;
            if (nsub > 1) {
                sl.b1_lo = tp_tri_lcuts[b][sub];
                sl.b1_hi = tp_tri_lcuts[b][sub + 1];
                sl.cut_again = true;
                if (sl.b1_lo >= sl.b1_hi) return false;
            }
        } else if (!tp_split_slice(C, cum, jt, q, np, blk.b_star, (int)slice, sl, true)) {
            return false; // empty
        }
    }
    __syncwarp(); // the previous item's slow-path lanes are done with the warp's areas
    if (lane == 0 && blk.g_bound_inv) { // what other blocks found so far, and what this one did
        const int theirs = (int)(0x7FFFFFFFu - theirs_inv);
        const int ours = *(volatile int *)&tp_tri_bound;
        if (theirs < ours) atomicMin(&tp_tri_bound, theirs);
        else if (ours < theirs) atomicMax(blk.g_bound_inv, 0x7FFFFFFFu - (unsigned)ours);
    }
    {
        const u64 per_lane = tp_split_slice_count(C, sl, q); // sets below the lane set that this item scores per lane
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
    if (p <= 5 && NA < 0xFFFFFFFFull) tp_tri_unrank32(mA, p, jt, (unsigned)rank, a0, a1);
    else tp_tri_unrank(C, mA, p, jt, rank, a0, a1);
    auto elem = [&](int i) { return (int)(((i < 8 ? a0 : a1) >> (8 * (i & 7))) & 0xFF); };
    // ---- V = sum of the A rows (positions < jt) -> shared memory; sA = score inside A ----
    int sA = blk.c_req;
    const int nch = (min(jt, sl.b0_hi) + 7) >> 3; // (nothing at or above the slice's top is ever read)
    if (p >= 2 && p <= 4) {
        const unsigned wp_s = (unsigned)__cvta_generic_to_shared(wp), t0_s = (unsigned)__cvta_generic_to_shared(t0);
        const unsigned va_lane = (unsigned)__cvta_generic_to_shared(va) + 16u * lane;
        const unsigned a32 = (unsigned)a0;
        sA = p == 4 ? tp_tri_lane_vector<4>(a32, wp_s, t0_s, va_lane, rsp, nch, sA)
                    : p == 3 ? tp_tri_lane_vector<3>(a32, wp_s, t0_s, va_lane, rsp, nch, sA)
                             : tp_tri_lane_vector<2>(a32, wp_s, t0_s, va_lane, rsp, nch, sA);
    } else {
        int rowoff[TP_SPLIT_MAX_P];
#pragma unroll
        for (int i = 0; i < TP_SPLIT_MAX_P; i++) {
            if (i >= p) break;
            rowoff[i] = elem(i) * rsp;
        }
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
    const int n = rq->n, k = rq->k, n_req = rq->n_req;
    const int gcap = plan->tri_gcap;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, tid = threadIdx.x, nt = blockDim.x;
    // layout: the S3 table first (+ a zero the compiler cannot see, so that its base lives in a
    // register and rows are addressed [base + constant]), then the staged request, the plain
    // table, u, and the per-warp areas
    uint32_t *tbl = (uint32_t *)smem + (plan->p >> 24);
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
    uint16_t *tstack = (uint16_t *)warp_base; // [DW+2][M]
    uint8_t *bstack = warp_base + TP_TRI_WB_OFF(M, DW);
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
        tp_tri_next_leaf = 0;
    }
    if (lane == 0) *(u64 *)(bstack + TP_TRI_WB_COUNT) = 0;
    for (int i = tid; i < n; i += nt) {
        s.group[i] = rq->group_of[i];
        s.gstart[i] = rq->gstart[i];
        s.gsize[i] = rq->gsize[i];
    }
    { // (two independent global loads per turn: the block waits for the slowest thread of this phase)
        const int kc = k + 1, cells = (n + 1) * kc;
        for (int i = tid; i < cells; i += 2 * nt) {
            const int i2 = i + nt, a = i / kc, a2 = i2 / kc;
            const u64 x = binom_g[a * TP_BINOM_COLS + (i - a * kc)];
            const u64 y = i2 < cells ? binom_g[a2 * TP_BINOM_COLS + (i2 - a2 * kc)] : 0ull;
            s.binom[i] = x;
            if (i2 < cells) s.binom[i2] = y;
        }
    }
    for (int i = tid; i < plan->n_entries; i += nt) {
        tp_tri_item_base[i] = plan->item_base[i];
        tp_tri_n_chunks[i] = plan->n_chunks[i];
        tp_tri_pieces[i] = plan->n_pieces[i];
    }
    for (int i = tid; i < TP_MAX_N + 2; i += nt) tp_tri_cum[i] = plan->cum[i];
    for (int i = tid; i < 4 * (TP_MAX_N + 1); i += nt) {
        const int col = i / (TP_MAX_N + 1) + 1, m = i % (TP_MAX_N + 1);
        const u64 c = col < TP_BINOM_COLS ? binom_g[m * TP_BINOM_COLS + col] : ~0ull;
        tp_tri_c32[col - 1][m] = c > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)c;
    }
    if (warp == TP_TRI_WARPS - 1) { // where every pivot entry's slice boundaries start: a running sum of n_pieces + 1
        int run = 0;
        for (int bs = 0; bs < plan->n_entries; bs += 32) {
            const int idx = bs + lane;
            const int v = idx < plan->n_entries ? (int)plan->n_pieces[idx] + 1 : 0;
            int incl = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                if (lane >= d) incl += up;
            }
            if (idx < plan->n_entries) tp_tri_cut_base[idx] = (uint16_t)min(run + incl - v, 0xFFFF);
            run += __shfl_sync(0xFFFFFFFFu, incl, 31);
        }
        if (lane == 0) {
            tp_tri_cut_base[plan->n_entries] = (uint16_t)min(run, 0xFFFF);
            tp_tri_cuts_ok = run <= TP_TRI_MAX_CUTS && n <= TP_TRI_MAX_HEADS;
        }
    }
    for (int b = 2 + tid; b <= 2 * gcap; b += nt) {
        const int h = (b - 1) >> 1;
        tp_tri_leaf_info[b] = make_uint4(4u * tp_tri_leaf_offsets.words[b], 16u * tp_tri_row_c0(2 * h + 1), 16u * tp_tri_row_c0(2 * h), 0u);
    }
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
    if (tp_tri_cuts_ok) { // the slice boundaries of every pivot entry, and the leaf cuts of the pairs
        const int kcols = k + 1, q = tp_tri_blk.sh.q, n_entries = tp_tri_blk.n_entries, b_star = tp_tri_blk.b_star;
        const int jt_hi = tp_tri_blk.jt_hi;
        const u64 *binom = s.binom;
        auto C64 = [&](int a, int b) -> u64 { return (b < 0 || a < b) ? 0ull : binom[a * kcols + b]; };
        // (q <= 4: the 32-bit table serves, and neither the search nor the division in tp_split_cut is 64 bits wide)
        auto C32 = [&](int a, int b) -> u64 { return b == 0 ? 1ull : (u64)tp_tri_c32[b - 1][a]; };
        const int total = tp_tri_cut_base[n_entries];
        for (int idx = tid; idx < total; idx += nt) {
            int lo = 0, hi = n_entries; // the entry whose range holds idx
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if ((int)tp_tri_cut_base[mid] <= idx) lo = mid;
                else hi = mid;
            }
            const int jt = jt_hi - lo, np = (int)tp_tri_pieces[lo], i = idx - (int)tp_tri_cut_base[lo];
            const int b_eff = tp_split_bstar(jt, b_star, q == 4), jlow = jt < b_eff ? jt : b_eff;
            tp_tri_cuts[idx] = (uint8_t)(q <= 4 ? tp_split_cut(C32, jlow, q, np, i, true) : tp_split_cut(C64, jlow, q, np, i, true));
        }
        if (q == 4)
            for (int idx = tid; idx < TP_TRI_MAX_HEADS * (TP_SPLIT_MAX_SUB + 1); idx += nt) {
                const int b = idx / (TP_SPLIT_MAX_SUB + 1), sub = idx - b * (TP_SPLIT_MAX_SUB + 1);
                if (b <= b_star || b >= jt_hi || b + 2 > TP_MAX_N + 1) continue;
                const int nsub = (int)tp_tri_cum[b + 2] - (int)tp_tri_cum[b];
                if (nsub > 1 && sub <= nsub) tp_tri_lcuts[b][sub] = (uint8_t)tp_split_cut(C32, b, q - 1, nsub, sub);
            }
    }
    TP_TRI_STAMP(4);
    if (TP_TRI_QB >= 4) tp_tri_build_quads(wp, rsp, n);
    tp_tri_build_table(tbl, gcap, wp, rsp);
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
    cudaError_t e = tp_opt_in_max_smem<tp_exhaustive_tri<M, DW>>(smem);
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
