// tp_split.cuh -- kernel SPLIT: the fast exhaustive scorer.  Included by tp_split_m*.cu, one
// translation unit per lane-vector width M so that they compile in parallel.
//
// Same contract as tp_exhaustive_lex (all C(n,k') subsets scored, Family + RefKey winner), but
// laid out so that a warp's 32 lanes never diverge and the pair weights are touched by
// 16-byte broadcast loads instead of per-lane byte gathers:
//
//   * positions are the canonical order reversed (position j = canonical n-1-j), a subset is
//     j_1 > j_2 > ... > j_k';
//   * the top p positions A = {j_1..j_p} belong to a LANE (32 consecutive A's of one pivot
//     jt = j_p per warp); V[j] = sum_{a in A} W[a][j], j < jt, is the lane's private vector
//     (shared memory, 16 bytes per lane per 8 positions);
//   * the next D = q-2 positions B'' are walked by the WARP (uniform nested loops); the
//     vector T[j] = u[j] + sum_{b in B''} W[b][j] lives in shared memory, one per level;
//   * the last two positions jx > jy are an unrolled triangular nest over registers
//     P[j] = V[j] + T[j] + W[b_last][j], two 16-bit positions per register:
//     min_{jy<jx} P[jy] + Wm[jx][jy] is one VIADDMNMX.U16x2 per register (two subsets per
//     lane-instruction) fed by 16-byte broadcast loads of row jx; one more VIADDMNMX adds
//     P[jx].  Rows of Wm carry TP_BIG at jy >= jx, so the triangular mask costs nothing;
//   * only the packed minimum is compared with the best Family score so far; the rare leaf
//     that beats it is re-walked by the slow path, which applies Family/RefKey exactly as the
//     LEX kernel does (tp_family_key).
//
// Work = items (pivot jt, chunk of 32 A's, slice of the first B'' position), heaviest first,
// handed out by an atomic counter; shard s of P takes the items i with (i + i/P) mod P == s.
#pragma once
#include "tp_device.cuh"

// Dynamic shared memory of the kernel.  The masked table sits at its very start, so that the
// unrolled rows address it with immediates.
extern __shared__ __align__(16) unsigned char tp_split_smem[];

struct TpSplitShared { // block-shared tables (pointers into dynamic shared memory)
    const uint16_t *wp; // [n][rsp] plain weights, reversed positions (all n columns: A sits above M)
    const uint16_t *wm; // [n][M+8] row jx: W for jy < jx, TP_BIG for jy >= jx
    int n, k, p, q, dbits, rsp;
    int min_w; // smallest weight between two free devices (prunes the slow path)
};

// Row stride (halfwords) of the plain table: covers all n positions, a multiple of 8 whose
// 16-byte chunk count is odd so that consecutive rows start in different bank groups.
static __host__ __device__ inline int tp_split_rsp(int n) {
    const int r8 = (n + 7) & ~7;
    return r8 + (((r8 / 8) & 1) ? 16 : 8);
}

template <int M>
struct TpSplitCfg {
    static constexpr int RS = M + 8;   // row stride of the masked table in halfwords
    static constexpr int RS4 = RS / 8; // ... in uint4
    static constexpr int NREG = M / 2;
    static constexpr int NCHUNK = M / 8;
};

// One row of the triangular nest: all pairs (JX, jy < JX).  Registers past the last valid jy
// are not touched at all (JX is a compile-time constant), so only an odd JX wastes half a lane-op.
// The first 16 bytes of the row arrive in `cur`, loaded while the previous row was computing;
// this row in turn starts the load of row JX-1 into `nxt`, so the shared-memory latency of the
// short rows (where the leaves spend their time) overlaps the arithmetic.
template <int M, int JX>
__device__ __forceinline__ void tp_split_row(const unsigned (&P)[TP_SPLIT_PREGS], const uint4 *__restrict__ wm4,
                                             unsigned &run, const uint4 &cur, uint4 &nxt) {
    constexpr int NFULL = JX / 8, REM = JX % 8;
    if constexpr (JX > 1) nxt = wm4[(JX - 1) * TpSplitCfg<M>::RS4];
    const unsigned px2 = __byte_perm(P[JX / 2], 0, (JX & 1) ? 0x3232 : 0x1010);
    unsigned acc = TP_BIG2 + TP_BIG2;
    if constexpr (NFULL >= 1) {
        acc = __viaddmin_u16x2(P[0], cur.x, acc);
        acc = __viaddmin_u16x2(P[1], cur.y, acc);
        acc = __viaddmin_u16x2(P[2], cur.z, acc);
        acc = __viaddmin_u16x2(P[3], cur.w, acc);
#pragma unroll
        for (int c = 1; c < NFULL; c++) {
            const uint4 r = wm4[JX * TpSplitCfg<M>::RS4 + c];
            acc = __viaddmin_u16x2(P[4 * c + 0], r.x, acc);
            acc = __viaddmin_u16x2(P[4 * c + 1], r.y, acc);
            acc = __viaddmin_u16x2(P[4 * c + 2], r.z, acc);
            acc = __viaddmin_u16x2(P[4 * c + 3], r.w, acc);
        }
        if constexpr (REM > 4) {
            const uint4 r = wm4[JX * TpSplitCfg<M>::RS4 + NFULL];
            acc = __viaddmin_u16x2(P[4 * NFULL + 0], r.x, acc);
            acc = __viaddmin_u16x2(P[4 * NFULL + 1], r.y, acc);
            acc = __viaddmin_u16x2(P[4 * NFULL + 2], r.z, acc);
            if constexpr (REM > 6) acc = __viaddmin_u16x2(P[4 * NFULL + 3], r.w, acc);
        } else if constexpr (REM > 0) {
            const uint2 r = ((const uint2 *)wm4)[(JX * TpSplitCfg<M>::RS4 + NFULL) * 2];
            acc = __viaddmin_u16x2(P[4 * NFULL + 0], r.x, acc);
            if constexpr (REM > 2) acc = __viaddmin_u16x2(P[4 * NFULL + 1], r.y, acc);
        }
    } else { // the whole row sits in the prefetched chunk
        acc = __viaddmin_u16x2(P[0], cur.x, acc);
        if constexpr (REM > 2) acc = __viaddmin_u16x2(P[1], cur.y, acc);
        if constexpr (REM > 4) acc = __viaddmin_u16x2(P[2], cur.z, acc);
        if constexpr (REM > 6) acc = __viaddmin_u16x2(P[3], cur.w, acc);
    }
    run = __viaddmin_u16x2(acc, px2, run);
}

#define TP_ROW_CASE(J)                                                                  \
    case J:                                                                             \
        if constexpr ((J) < M && (J) >= 1) {                                            \
            if constexpr ((J) & 1) tp_split_row<M, (J)>(P, wm4, run, buf_odd, buf_even); \
            else tp_split_row<M, (J)>(P, wm4, run, buf_even, buf_odd);                  \
        }                                                                               \
        [[fallthrough]];
#define TP_ROW_CASE8(B)                                                                              \
    TP_ROW_CASE(B + 7) TP_ROW_CASE(B + 6) TP_ROW_CASE(B + 5) TP_ROW_CASE(B + 4) TP_ROW_CASE(B + 3) \
    TP_ROW_CASE(B + 2) TP_ROW_CASE(B + 1) TP_ROW_CASE(B + 0)

// min over all pairs jx > jy with jx <= start of P[jx] + P[jy] + W[jx][jy], packed per half.
template <int M>
__device__ __forceinline__ unsigned tp_split_pairs(const unsigned (&P)[TP_SPLIT_PREGS], const uint4 *__restrict__ wm4,
                                                   int start) {
    unsigned run = TP_BIG2 + TP_BIG2;
    uint4 buf_even = wm4[start * TpSplitCfg<M>::RS4], buf_odd = buf_even; // first chunk of the entry row
    switch (start) {
        TP_ROW_CASE8(120) TP_ROW_CASE8(112) TP_ROW_CASE8(104) TP_ROW_CASE8(96) TP_ROW_CASE8(88) TP_ROW_CASE8(80)
        TP_ROW_CASE8(72) TP_ROW_CASE8(64) TP_ROW_CASE8(56) TP_ROW_CASE8(48) TP_ROW_CASE8(40) TP_ROW_CASE8(32)
        TP_ROW_CASE8(24) TP_ROW_CASE8(16) TP_ROW_CASE8(8) TP_ROW_CASE8(0)
    default: break;
    }
    return run;
}

// P = V + T (+ masked row of the leaf position) for the chunks that hold a position < limit:
// entered at the highest such chunk, falling through to chunk 0.
template <int M, bool HAS_ROW, int C>
__device__ __forceinline__ void tp_split_p_chunk(unsigned (&P)[TP_SPLIT_PREGS], const uint4 *__restrict__ v4,
                                                 const uint4 *__restrict__ t4, const uint4 *__restrict__ wb4) {
    if constexpr (C < M / 8) {
        const uint4 v = v4[C * 32]; // lane-private: [chunk][lane]
        const uint4 t = t4[C];
        if constexpr (HAS_ROW) {
            const uint4 w = wb4[C];
            P[4 * C + 0] = v.x + t.x + w.x;
            P[4 * C + 1] = v.y + t.y + w.y;
            P[4 * C + 2] = v.z + t.z + w.z;
            P[4 * C + 3] = v.w + t.w + w.w;
        } else {
            P[4 * C + 0] = v.x + t.x;
            P[4 * C + 1] = v.y + t.y;
            P[4 * C + 2] = v.z + t.z;
            P[4 * C + 3] = v.w + t.w;
        }
    }
}
#define TP_P_CASE(C)                                     \
    case (C) + 1:                                        \
        tp_split_p_chunk<M, HAS_ROW, (C)>(P, v4, t4, wb4); \
        [[fallthrough]];
template <int M, bool HAS_ROW>
__device__ __forceinline__ void tp_split_make_p(unsigned (&P)[TP_SPLIT_PREGS], const uint4 *__restrict__ v4,
                                                const uint4 *__restrict__ t4, const uint4 *__restrict__ wb4,
                                                int limit) {
    switch ((limit + 7) >> 3) { // number of chunks
        TP_P_CASE(15) TP_P_CASE(14) TP_P_CASE(13) TP_P_CASE(12) TP_P_CASE(11) TP_P_CASE(10) TP_P_CASE(9) TP_P_CASE(8)
        TP_P_CASE(7) TP_P_CASE(6) TP_P_CASE(5) TP_P_CASE(4) TP_P_CASE(3) TP_P_CASE(2) TP_P_CASE(1) TP_P_CASE(0)
    default: break;
    }
}

// ---- two leaves at once -----------------------------------------------------------------------
// Consecutive leaf positions b and b-1 of one lane share everything except the masked row that
// enters P, so they are scored together: one set of row loads, address arithmetic and dispatch
// feeds two independent VIADDMNMX chains (Pa for b, Pb for b-1).  Row jx = b-1 is not a pair of
// the second leaf and excludes itself: Pb carries TP_BIG at every position >= b-1.  Limited to
// positions < TP_SPLIT_DUAL_M so that both vectors fit in registers for every M.
#define TP_SPLIT_DUAL_M 64
#define TP_SPLIT_PB (TP_SPLIT_DUAL_M / 2) /* the second leaf's vector lives in the upper half of the register array */

template <int M, int JX>
__device__ __forceinline__ void tp_split_row2(const unsigned (&Pa)[TP_SPLIT_PREGS],
                                              const uint4 *__restrict__ wm4, unsigned &runa, unsigned &runb,
                                              const uint4 &cur, uint4 &nxt) {
    constexpr int NFULL = JX / 8, REM = JX % 8;
    if constexpr (JX > 1) nxt = wm4[(JX - 1) * TpSplitCfg<M>::RS4];
    constexpr unsigned SEL = (JX & 1) ? 0x3232 : 0x1010;
    const unsigned pxa = __byte_perm(Pa[JX / 2], 0, SEL), pxb = __byte_perm(Pa[TP_SPLIT_PB + JX / 2], 0, SEL);
    unsigned acca = TP_BIG2 + TP_BIG2, accb = TP_BIG2 + TP_BIG2;
#define TP_ROW2_OP(r, w)                          \
    do {                                          \
        acca = __viaddmin_u16x2(Pa[r], (w), acca); \
        accb = __viaddmin_u16x2(Pa[TP_SPLIT_PB + r], (w), accb); \
    } while (0)
    if constexpr (NFULL >= 1) {
        TP_ROW2_OP(0, cur.x);
        TP_ROW2_OP(1, cur.y);
        TP_ROW2_OP(2, cur.z);
        TP_ROW2_OP(3, cur.w);
#pragma unroll
        for (int c = 1; c < NFULL; c++) {
            const uint4 r = wm4[JX * TpSplitCfg<M>::RS4 + c];
            TP_ROW2_OP(4 * c + 0, r.x);
            TP_ROW2_OP(4 * c + 1, r.y);
            TP_ROW2_OP(4 * c + 2, r.z);
            TP_ROW2_OP(4 * c + 3, r.w);
        }
        if constexpr (REM > 4) {
            const uint4 r = wm4[JX * TpSplitCfg<M>::RS4 + NFULL];
            TP_ROW2_OP(4 * NFULL + 0, r.x);
            TP_ROW2_OP(4 * NFULL + 1, r.y);
            TP_ROW2_OP(4 * NFULL + 2, r.z);
            if constexpr (REM > 6) TP_ROW2_OP(4 * NFULL + 3, r.w);
        } else if constexpr (REM > 0) {
            const uint2 r = ((const uint2 *)wm4)[(JX * TpSplitCfg<M>::RS4 + NFULL) * 2];
            TP_ROW2_OP(4 * NFULL + 0, r.x);
            if constexpr (REM > 2) TP_ROW2_OP(4 * NFULL + 1, r.y);
        }
    } else {
        TP_ROW2_OP(0, cur.x);
        if constexpr (REM > 2) TP_ROW2_OP(1, cur.y);
        if constexpr (REM > 4) TP_ROW2_OP(2, cur.z);
        if constexpr (REM > 6) TP_ROW2_OP(3, cur.w);
    }
#undef TP_ROW2_OP
    runa = __viaddmin_u16x2(acca, pxa, runa);
    runb = __viaddmin_u16x2(accb, pxb, runb);
}

#define TP_ROW2_CASE(J)                                                                            \
    case J:                                                                                        \
        if constexpr ((J) < TP_SPLIT_DUAL_M && (J) >= 1) {                                         \
            if constexpr ((J) & 1) tp_split_row2<M, (J)>(Pa, wm4, runa, runb, buf_odd, buf_even); \
            else tp_split_row2<M, (J)>(Pa, wm4, runa, runb, buf_even, buf_odd);                  \
        }                                                                                          \
        [[fallthrough]];
#define TP_ROW2_CASE8(B)                                                                                \
    TP_ROW2_CASE(B + 7) TP_ROW2_CASE(B + 6) TP_ROW2_CASE(B + 5) TP_ROW2_CASE(B + 4) TP_ROW2_CASE(B + 3) \
    TP_ROW2_CASE(B + 2) TP_ROW2_CASE(B + 1) TP_ROW2_CASE(B + 0)

template <int M>
__device__ __forceinline__ void tp_split_pairs2(const unsigned (&Pa)[TP_SPLIT_PREGS],
                                                const uint4 *__restrict__ wm4, int start, unsigned &runa,
                                                unsigned &runb) {
    runa = TP_BIG2 + TP_BIG2;
    runb = TP_BIG2 + TP_BIG2;
    uint4 buf_even = wm4[start * TpSplitCfg<M>::RS4], buf_odd = buf_even;
    switch (start) {
        TP_ROW2_CASE8(56) TP_ROW2_CASE8(48) TP_ROW2_CASE8(40) TP_ROW2_CASE8(32) TP_ROW2_CASE8(24) TP_ROW2_CASE8(16)
        TP_ROW2_CASE8(8) TP_ROW2_CASE8(0)
    default: break;
    }
}

template <int M, int C>
__device__ __forceinline__ void tp_split_p2_chunk(unsigned (&Pa)[TP_SPLIT_PREGS],
                                                  const uint4 *__restrict__ v4, const uint4 *__restrict__ t4,
                                                  const uint4 *__restrict__ wa4, const uint4 *__restrict__ wb4) {
    if constexpr (C < TP_SPLIT_DUAL_M / 8) {
        const uint4 v = v4[C * 32];
        const uint4 t = t4[C];
        const uint4 a = wa4[C], b = wb4[C];
        const unsigned x = v.x + t.x, y = v.y + t.y, z = v.z + t.z, w = v.w + t.w;
        Pa[4 * C + 0] = x + a.x;
        Pa[4 * C + 1] = y + a.y;
        Pa[4 * C + 2] = z + a.z;
        Pa[4 * C + 3] = w + a.w;
        Pa[TP_SPLIT_PB + 4 * C + 0] = x + b.x;
        Pa[TP_SPLIT_PB + 4 * C + 1] = y + b.y;
        Pa[TP_SPLIT_PB + 4 * C + 2] = z + b.z;
        Pa[TP_SPLIT_PB + 4 * C + 3] = w + b.w;
    }
}
#define TP_P2_CASE(C)                                      \
    case (C) + 1:                                          \
        tp_split_p2_chunk<M, (C)>(Pa, v4, t4, wa4, wb4);   \
        [[fallthrough]];
template <int M>
__device__ __forceinline__ void tp_split_make_p2(unsigned (&Pa)[TP_SPLIT_PREGS],
                                                 const uint4 *__restrict__ v4, const uint4 *__restrict__ t4,
                                                 const uint4 *__restrict__ wa4, const uint4 *__restrict__ wb4,
                                                 int limit) {
    switch ((limit + 7) >> 3) {
        TP_P2_CASE(7) TP_P2_CASE(6) TP_P2_CASE(5) TP_P2_CASE(4) TP_P2_CASE(3) TP_P2_CASE(2) TP_P2_CASE(1) TP_P2_CASE(0)
    default: break;
    }
}

// ---- the six smallest leaves at once ----------------------------------------------------------
// A leaf at position b holds only C(b,2) pairs, and most leaves are small (the average leaf
// position is about n / (k'-1)), so for them the per-leaf work around the rows (dispatch, lane
// vector loads, threshold) outweighs the rows.  Leaves 2..7 live entirely in the first 16-byte
// chunk of every vector: this block scores all six with compile-time positions -- V+T loaded once,
// seven row chunks loaded once, no dispatch -- and returns the packed minimum of
// res[b] = (min over the leaf's pairs) + (V+T)[b].
template <int M>
__device__ __forceinline__ unsigned tp_split_tail(const uint4 *__restrict__ v4, const uint4 *__restrict__ t4,
                                                  const uint4 *__restrict__ wm4, unsigned (&res)[8],
                                                  unsigned (&vt)[4]) {
    const uint4 v = v4[0], t = t4[0];
    vt[0] = v.x + t.x;
    vt[1] = v.y + t.y;
    vt[2] = v.z + t.z;
    vt[3] = v.w + t.w;
    unsigned row[8][4];
#pragma unroll
    for (int j = 1; j < 8; j++) {
        const uint4 r = wm4[j * TpSplitCfg<M>::RS4]; // positions 0..7 of row j, TP_BIG at jy >= j
        row[j][0] = r.x;
        row[j][1] = r.y;
        row[j][2] = r.z;
        row[j][3] = r.w;
    }
    unsigned m = 0xFFFFFFFFu;
#pragma unroll
    for (int b = 2; b < 8; b++) {
        unsigned Pb[4];
#pragma unroll
        for (int r = 0; r < (b + 1) / 2; r++) Pb[r] = vt[r] + row[b][r];
        unsigned run = TP_BIG2 + TP_BIG2;
#pragma unroll
        for (int jx = 1; jx < b; jx++) {
            unsigned acc = TP_BIG2 + TP_BIG2;
#pragma unroll
            for (int r = 0; r < (jx + 1) / 2; r++) acc = __viaddmin_u16x2(Pb[r], row[jx][r], acc);
            run = __viaddmin_u16x2(acc, __byte_perm(Pb[jx / 2], 0, (jx & 1) ? 0x3232 : 0x1010), run);
        }
        res[b] = run + __byte_perm(vt[b / 2], 0, (b & 1) ? 0x3232 : 0x1010);
        m = __vminu2(m, res[b]);
    }
    return m;
}

// Leaves 8..11 the same way (two chunks per vector): rows are loaded once for the four leaves.
template <int M>
__device__ __forceinline__ unsigned tp_split_mid(const uint4 *__restrict__ v4, const uint4 *__restrict__ t4,
                                                 const uint4 *__restrict__ wm4, const unsigned (&vt0)[4],
                                                 unsigned (&res)[4], unsigned (&vt1)[4]) {
    constexpr int RS4 = TpSplitCfg<M>::RS4;
    const uint4 v = v4[32], t = t4[1];
    vt1[0] = v.x + t.x;
    vt1[1] = v.y + t.y;
    vt1[2] = v.z + t.z;
    vt1[3] = v.w + t.w;
    unsigned Pb[4][6], run[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { // leaf b = 8 + i: positions 0..7, and 8..b-1 from the second chunk
        const int b = 8 + i;
        const uint4 r0 = wm4[b * RS4];
        Pb[i][0] = vt0[0] + r0.x;
        Pb[i][1] = vt0[1] + r0.y;
        Pb[i][2] = vt0[2] + r0.z;
        Pb[i][3] = vt0[3] + r0.w;
        if (i > 0) {
            const uint2 r1 = ((const uint2 *)(wm4 + b * RS4 + 1))[0];
            Pb[i][4] = vt1[0] + r1.x;
            if (i > 2) Pb[i][5] = vt1[1] + r1.y;
        }
        run[i] = TP_BIG2 + TP_BIG2;
    }
#pragma unroll
    for (int jx = 10; jx >= 1; jx--) {
        unsigned row[6];
        {
            const uint4 r = wm4[jx * RS4];
            row[0] = r.x;
            row[1] = r.y;
            row[2] = r.z;
            row[3] = r.w;
            if (jx > 8) {
                const uint2 r1 = ((const uint2 *)(wm4 + jx * RS4 + 1))[0];
                row[4] = r1.x;
                row[5] = r1.y;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (8 + i > jx) {
                unsigned acc = TP_BIG2 + TP_BIG2;
#pragma unroll
                for (int r = 0; r < (jx + 1) / 2; r++) acc = __viaddmin_u16x2(Pb[i][r], row[r], acc);
                run[i] = __viaddmin_u16x2(acc, __byte_perm(Pb[i][jx / 2], 0, (jx & 1) ? 0x3232 : 0x1010), run[i]);
            }
        }
    }
    unsigned m = 0xFFFFFFFFu;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        res[i] = run[i] + __byte_perm(vt1[i / 2], 0, (i & 1) ? 0x3232 : 0x1010);
        m = __vminu2(m, res[i]);
    }
    return m;
}

struct TpSplitWarp { // per-warp state (uniform unless noted)
    uint16_t *tstack;   // [D+1][M] level vectors
    uint32_t *va;       // [M/8][32][4] lane-private V, packed pairs, 16 bytes per lane per chunk
    uint8_t *bstack;    // [TP_SPLIT_MAX_D] chosen B'' positions; [8], [9]: the item's bounds on the second one
    u64 a0, a1;         // lane: A positions, one per byte, descending
    int active;         // lane: this lane holds a real A
    int lane;
};

__device__ __forceinline__ int tp_split_va_get(const uint32_t *va, int lane, int j) {
    const unsigned pk = va[((j >> 3) * 32 + lane) * 4 + ((j >> 1) & 3)];
    return (j & 1) ? (int)(pk >> 16) : (int)(pk & 0xFFFF);
}

// The fixed part of a leaf (A, B'', leaf position; canonical ascending = positions descending)
// against tp_family_alive: most leaves that merely tie with the bound die here, before any pair
// of theirs is looked at.  The pairs come from positions < limit.
static __device__ __noinline__ bool tp_split_leaf_alive(const TpSplitShared &sh, const TpSmem &s, const TpSplitWarp &w,
                                                        int d_used, int b_leaf, int limit) {
    u64 f0 = 0, f1 = 0;
    int cnt = 0;
    auto push = [&](int pos) {
        const u64 c = (u64)(sh.n - 1 - pos);
        if (cnt < 8) f0 |= c << (8 * cnt);
        else f1 |= c << (8 * (cnt - 8));
        cnt++;
    };
    for (int i = 0; i < sh.p; i++) push((int)(((i < 8 ? w.a0 : w.a1) >> (8 * (i & 7))) & 0xFF));
    for (int i = 0; i + 1 < d_used; i++) push(w.bstack[i]);
    if (d_used > 0) push(b_leaf);
    return tp_family_alive(f0, f1, cnt, sh.n - limit, s.group, s.gstart, s.gsize);
}

// Re-walk one leaf for one lane with scalar arithmetic and apply Family/RefKey to every pair
// that scores no worse than the bound.  base = c_req + sA + cross(A,B'') + score inside B''.
template <int M>
// (The three descriptors come by value: handing out their addresses would pin the caller's copies
// to local memory and make the hot loop reload them after every possible call.)
__device__ __noinline__ void tp_split_slow(const TpSplitShared sh, const TpSmem s, const TpSplitWarp w,
                                           int d_used, int b_leaf, int limit, int base, int *s_bound,
                                           TpKey &best) {
    const int n = sh.n;
    const uint16_t *t = w.tstack + (d_used > 0 ? d_used - 1 : 0) * M;
    auto pval = [&](int j) {
        int v = tp_split_va_get(w.va, w.lane, j) + t[j];
        if (d_used > 0) v += sh.wp[b_leaf * sh.rsp + j];
        return v;
    };
    if (!tp_split_leaf_alive(sh, s, w, d_used, b_leaf, limit)) return;
    // Stage the lane's P values once, then visit only rows that can still reach the bound:
    // row jx needs P[jx] + min_{jy<jx} P[jy] + (smallest pair weight) <= bound - base.
    int pv[TP_MAX_N];
    for (int j = 0; j < limit; j++) pv[j] = pval(j);
    int pmin = pv[0];
    for (int jx = 1; jx < limit; jx++) {
        const int px = base + pv[jx];
        if (px + pmin + sh.min_w <= *(volatile int *)s_bound) {
            const uint16_t *wrow = sh.wp + jx * sh.rsp;
            for (int jy = 0; jy < jx; jy++) {
                const int sco = px + pv[jy] + wrow[jy];
                if (sco > *(volatile int *)s_bound) continue;
                // canonical ascending = positions descending: A, then B'', then jx, jy
                u64 p0 = 0, p1 = 0;
                int cnt = 0;
                auto push = [&](int pos) {
                    const u64 c = (u64)(n - 1 - pos);
                    if (cnt < 8) p0 |= c << (8 * cnt);
                    else p1 |= c << (8 * (cnt - 8));
                    cnt++;
                };
                for (int i = 0; i < sh.p; i++) push((int)(((i < 8 ? w.a0 : w.a1) >> (8 * (i & 7))) & 0xFF));
                for (int i = 0; i + 1 < d_used; i++) push(w.bstack[i]); // written before a __syncwarp
                if (d_used > 0) push(b_leaf);
                push(jx);
                push(jy);
                TpKey cand;
                if (tp_family_key(p0, p1, sh.k, sco, sh.dbits, s.group, s.gstart, s.gsize, cand)) {
                    if (tp_key_less(cand, best)) best = cand;
                    atomicMin(s_bound, sco);
                }
            }
        }
        pmin = min(pmin, pv[jx]);
    }
}

// The warp-uniform walk over B'' = (b_0 > ... > b_{D-1}), one template level per position so
// that every level's state is a named register.  Level L picks b in [lo, hi).
template <int M, int D, int L>
struct TpSplitWalk {
    static __device__ __forceinline__ void run(const TpSplitShared &sh, const TpSmem &s, TpSplitWarp &w,
                                               unsigned (&P)[TP_SPLIT_PREGS], int hi, int lo, int lb, int sb,
                                               int *s_bound, TpKey &best) {
        using Cfg = TpSplitCfg<M>;
        const uint16_t *tl = w.tstack + L * M;
        if constexpr (L < D - 1) {
            for (int b = hi - 1; b >= lo; --b) {
                const int lb2 = lb + tp_split_va_get(w.va, w.lane, b);
                const int sb2 = sb + tl[b];
                // T_{L+1} = T_L + W[b] on the positions below b (the rest is never read)
                uint32_t *dst = (uint32_t *)(w.tstack + (L + 1) * M);
                const uint32_t *src = (const uint32_t *)tl, *rw = (const uint32_t *)(sh.wp + b * sh.rsp);
                __syncwarp(); // lanes still in the slow path of an earlier leaf read T_{L+1} and bstack
                if (w.lane == 0) w.bstack[L] = (uint8_t)b;
                for (int wd = w.lane; wd < M / 2; wd += 32) dst[wd] = src[wd] + rw[wd];
                __syncwarp();
                if constexpr (L == 0) // the item may hold only a sub-slice of the second position
                    TpSplitWalk<M, D, 1>::run(sh, s, w, P, min(b, (int)w.bstack[9]), max(sh.q - 2, (int)w.bstack[8]), lb2,
                                              sb2, s_bound, best);
                else
                    TpSplitWalk<M, D, L + 1>::run(sh, s, w, P, b, sh.q - 2 - L, lb2, sb2, s_bound, best);
            }
        } else {
            // leaf level: b is the last walked position, the pairs live below it
            // (sh.wm is the start of shared memory; going through the field rather than the symbol keeps
            // the table base in a register instead of re-deriving the shared window address per load)
            const uint4 *wm4 = (const uint4 *)sh.wm;
            const uint4 *v4 = (const uint4 *)w.va + w.lane;
            const int room = *(volatile int *)s_bound - lb - sb + 1; // > score of (b, pair) <=> no candidate
            // (a lane without a real A never reports: its threshold is 0)
            const unsigned room2 = w.active ? (unsigned)min(max(room, 0), 0xFFFF) * 0x10001u : 0u;
            int b = hi - 1;
            while (b >= lo) {
                if ((b | 1) <= TP_SPLIT_DUAL_M) {
                    // Leaves top and top-1 together, top odd: the two positions then share one packed
                    // word of V and of T, so one add yields both leaves' own terms and one packed
                    // compare decides for both.  An even first leaf drags the dummy top = hi along
                    // and a last one the dummy top-1 < lo; both are turned away on the rare path.
                    const bool tail = lo == 2 && hi > 7; // leaves 2..7 are all there: scored in one block below
                    const bool mid = tail && hi > 11;    // ... and so are leaves 8..11
                    const int stop = mid ? 13 : tail ? 9 : lo;
                    for (int top = b | 1; top >= stop; top -= 2) {
                        tp_split_make_p2<M>(P, v4, (const uint4 *)tl, (const uint4 *)(sh.wm + top * Cfg::RS),
                                            (const uint4 *)(sh.wm + (top - 1) * Cfg::RS), top);
                        unsigned runa, runb;
                        tp_split_pairs2<M>(P, wm4, top - 1, runa, runb);
                        const int h = top >> 1;
                        const unsigned vt = ((const uint32_t *)v4)[(h >> 2) * 128 + (h & 3)] + ((const uint32_t *)tl)[h];
                        const unsigned ta = runa + __byte_perm(vt, 0, 0x3232); // + (V+T)[top] in both halves
                        const unsigned tb = runb + __byte_perm(vt, 0, 0x1010); // + (V+T)[top-1]
                        if (__vminu2(__vminu2(ta, tb), room2) != room2) {
                            if (__vminu2(ta, room2) != room2 && top < hi)
                                tp_split_slow<M>(sh, s, w, D, top, top, lb + sb + (int)(vt >> 16), s_bound, best);
                            if (__vminu2(tb, room2) != room2 && top - 1 >= lo)
                                tp_split_slow<M>(sh, s, w, D, top - 1, top - 1, lb + sb + (int)(vt & 0xFFFF), s_bound,
                                                 best);
                        }
                    }
                    if (tail) {
                        unsigned res[8], vt[4];
                        const unsigned m = tp_split_tail<M>(v4, (const uint4 *)tl, wm4, res, vt);
                        if (mid) {
                            unsigned res1[4], vt1[4];
                            const unsigned m1 = tp_split_mid<M>(v4, (const uint4 *)tl, wm4, vt, res1, vt1);
                            if (__vminu2(m1, room2) != room2) {
#pragma unroll
                                for (int i = 3; i >= 0; i--)
                                    if (__vminu2(res1[i], room2) != room2)
                                        tp_split_slow<M>(sh, s, w, D, 8 + i, 8 + i,
                                                         lb + sb + (int)((vt1[i / 2] >> (16 * (i & 1))) & 0xFFFF), s_bound,
                                                         best);
                            }
                        }
                        if (__vminu2(m, room2) != room2) {
#pragma unroll
                            for (int lb7 = 7; lb7 >= 2; lb7--)
                                if (__vminu2(res[lb7], room2) != room2)
                                    tp_split_slow<M>(sh, s, w, D, lb7, lb7,
                                                     lb + sb + (int)((vt[lb7 / 2] >> (16 * (lb7 & 1))) & 0xFFFF), s_bound,
                                                     best);
                        }
                    }
                    break;
                }
                if constexpr (M > TP_SPLIT_DUAL_M) { // positions above 64: one leaf at a time
                    tp_split_make_p<M, true>(P, v4, (const uint4 *)tl, (const uint4 *)(sh.wm + b * Cfg::RS), b);
                    const unsigned run = tp_split_pairs<M>(P, wm4, b - 1);
                    const int vab = tp_split_va_get(w.va, w.lane, b);
                    const int thr = min(max(room - vab - (int)tl[b], 0), 0xFFFF);
                    const unsigned thr2 = (unsigned)thr * 0x10001u;
                    if (__vminu2(run, thr2) != thr2 && w.active)
                        tp_split_slow<M>(sh, s, w, D, b, b, lb + vab + sb + tl[b], s_bound, best);
                }
                --b;
            }
        }
    }
};

// Exact seed for requests with few groups: the block scores every Family member (whole groups by
// bitmask, plus at most one lowest-NodeId prefix of another group), so the scan starts from the
// optimal Family score and only ties -- and non-members below it -- ever reach the slow path.
// Without it a search that is all "first leaves" (thousands of warps starting at once from the
// two cheap seeds) spends most of its time re-walking leaves that a better bound would skip.
#define TP_SEED_MAX_GROUPS 14
static __device__ __noinline__ void tp_split_family_seed(const TpSmem s, int n_groups, int k, int c_req, int *s_bound) {
    if (n_groups > TP_SEED_MAX_GROUPS) return;
    const unsigned n_masks = 1u << n_groups;
    int best = 0x7FFFFFFF;
    if (n_masks <= blockDim.x) {
        // Few groups (the usual node: 8 GPUs): one thread per mask would leave the few masks that
        // carry members with all the scoring; spread (mask, partial group) pairs instead, at most
        // one member per pair.  h == n_groups stands for "no partial group".
        const unsigned per = (unsigned)n_groups + 1u;
        for (unsigned unit = threadIdx.x; unit < n_masks * per; unit += blockDim.x) {
            const unsigned mask = unit / per;
            const int h = (int)(unit - mask * per);
            if (h < n_groups && ((mask >> h) & 1u)) continue;
            int pos[TP_MAX_K];
            int cnt = 0;
            bool over = false;
            for (unsigned m = mask; m; m &= m - 1) {
                const int g = __ffs(m) - 1;
                const int gz = s.gsize[g], gs = s.gstart[g];
                if (cnt + gz > k) {
                    over = true;
                    break;
                }
                for (int t = 0; t < gz; t++) pos[cnt++] = gs + t;
            }
            if (over) continue;
            const int r = k - cnt;
            if (h == n_groups ? r != 0 : (r == 0 || s.gsize[h] <= r)) continue;
            for (int t = 0; t < r; t++) pos[cnt++] = s.gstart[h] + t; // the prefix (order does not matter here)
            int sco = c_req;
            for (int i = 0; i < cnt; i++) {
                sco += s.u[pos[i]];
                for (int j = 0; j < i; j++) sco += s.w[pos[j] * s.stride + pos[i]];
            }
            best = min(best, sco);
        }
        if (best != 0x7FFFFFFF) atomicMin(s_bound, best);
        return;
    }
    for (unsigned mask = threadIdx.x; mask < n_masks; mask += blockDim.x) {
        int pos[TP_MAX_K];
        int cnt = 0;
        bool over = false;
        for (unsigned m = mask; m; m &= m - 1) {
            const int g = __ffs(m) - 1;
            const int gz = s.gsize[g], gs = s.gstart[g];
            if (cnt + gz > k) {
                over = true;
                break;
            }
            for (int t = 0; t < gz; t++) pos[cnt++] = gs + t;
        }
        if (over) continue;
        int base = c_req;
        for (int i = 0; i < cnt; i++) {
            base += s.u[pos[i]];
            for (int j = 0; j < i; j++) base += s.w[pos[j] * s.stride + pos[i]];
        }
        if (cnt == k) {
            best = min(best, base);
            continue;
        }
        const int r = k - cnt; // devices still to take: a proper prefix of one more group
        for (int h = 0; h < n_groups; h++) {
            if (((mask >> h) & 1u) || s.gsize[h] <= r) continue;
            const int hs = s.gstart[h];
            int sco = base;
            for (int t = 0; t < r; t++) {
                const int x = hs + t;
                sco += s.u[x];
                for (int i = 0; i < cnt; i++) sco += s.w[pos[i] * s.stride + x];
                for (int t2 = 0; t2 < t; t2++) sco += s.w[(hs + t2) * s.stride + x];
            }
            best = min(best, sco);
        }
    }
    if (best != 0x7FFFFFFF) atomicMin(s_bound, best);
}

// The search of one request by the calling block: items are handed out through *work_counter
// (a global word shared by the grid, or a shared-memory word when the block owns the request).
// Not inlined on purpose: the kernel's request loop must not compete with it for registers.
template <int M, int D>
__device__ __noinline__ void tp_split_search(const uint8_t *__restrict__ W, int nd,
                                                const u64 *__restrict__ binom_g, const TpReq *__restrict__ rq,
                                                const TpPlan *__restrict__ plan, int shard_idx, int shard_cnt,
                                                unsigned *work_counter, TpKey &best, u64 &counted) {
    using Cfg = TpSplitCfg<M>;
    unsigned char *const smem = tp_split_smem;
    __shared__ int s_bound;
    const int n = rq->n, k = rq->k, p = plan->p, q = plan->q;
    const int n_entries = plan->n_entries;
    const unsigned total_items = plan->total_items;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // layout: masked table first (its row chunks are then compile-time offsets from the start of
    // shared memory), then the staged request, the plain table, u, and the per-warp areas
    // (+ a zero the compiler cannot see: the base then lives in a register and rows are addressed
    // [base + constant]; from the bare symbol every load re-derives the shared window address)
    uint16_t *wm = (uint16_t *)smem + (plan->p >> 24);
    size_t off = (size_t)n * Cfg::RS * 2;
    const TpSmem s = tp_smem_carve(smem + off, n, k);
    tp_stage_request(s, W, nd, binom_g, rq, k, &s_bound);
    off += (tp_smem_bytes(n, k) + 15) & ~(size_t)15;
    const int rsp = tp_split_rsp(n);
    uint16_t *wp = (uint16_t *)(smem + off);
    off += (size_t)n * rsp * 2;
    uint16_t *t0 = (uint16_t *)(smem + off); // u in reversed positions, all n of them (level 0 of T)
    off += (size_t)TP_MAX_N * 2;
    const size_t per_warp = (size_t)(D + 1) * M * 2 + (size_t)Cfg::NCHUNK * 32 * 16 + 16;
    unsigned char *warp_base = smem + off + (size_t)warp * per_warp;
    uint16_t *tstack = (uint16_t *)warp_base;                             // [D+1][M]
    uint32_t *va = (uint32_t *)(warp_base + (size_t)(D + 1) * M * 2);     // [M/8][32][4]
    uint8_t *bstack = (uint8_t *)(va + Cfg::NCHUNK * 32 * 4);             // [<=16]
    for (int j = warp; j < n; j += TP_SPLIT_WARPS) { // one table row per warp pass: no divisions
        const uint8_t *src = s.w + (n - 1 - j) * s.stride + (n - 1);
        for (int jj = lane; jj < rsp; jj += 32) // plain table, padding 0 (it is summed)
            wp[j * rsp + jj] = (jj < n && jj != j) ? (uint16_t)src[-jj] : (uint16_t)0;
        for (int jj = lane; jj < Cfg::RS; jj += 32) // masked table, padding BIG
            wm[j * Cfg::RS + jj] = (jj < j && jj < n) ? (uint16_t)src[-jj] : (uint16_t)TP_BIG;
    }
    for (int i = threadIdx.x; i < TP_MAX_N; i += blockDim.x) t0[i] = i < n ? s.u[n - 1 - i] : (uint16_t)0;
    if (threadIdx.x == 0 && rq->n_groups > TP_SEED_MAX_GROUPS) { // second seed where the exact one does not run:
        // whole groups from the largest down (also a Family member)
        int left = k, sco = rq->c_req, cnt = 0;
        int pos[TP_MAX_K];
        for (int g = rq->n_groups - 1; g >= 0 && left > 0; g--) {
            const int take = min((int)s.gsize[g], left);
            for (int t = 0; t < take; t++) pos[cnt++] = s.gstart[g] + t;
            left -= take;
        }
        for (int i = 0; i < cnt; i++) {
            sco += s.u[pos[i]];
            for (int j = 0; j < i; j++) sco += s.w[pos[j] * s.stride + pos[i]];
        }
        if (left == 0 && sco < s_bound) atomicMin(&s_bound, sco);
    }
    tp_split_family_seed(s, rq->n_groups, k, rq->c_req, &s_bound);
    __syncthreads();
    for (int i = lane; i < M; i += 32) tstack[i] = t0[i];
    __syncwarp();

    // the plan's per-pivot tables, block-shared (item -> pivot lookups happen once per item)
    __shared__ uint32_t s_item_base[TP_MAX_N], s_chunks[TP_MAX_N];
    __shared__ uint8_t s_pieces[TP_MAX_N];
    __shared__ uint16_t s_cum[TP_MAX_N + 2];
    for (int i = threadIdx.x; i < plan->n_entries; i += blockDim.x) {
        s_item_base[i] = plan->item_base[i];
        s_chunks[i] = plan->n_chunks[i];
        s_pieces[i] = plan->n_pieces[i];
    }
    for (int i = threadIdx.x; i < TP_MAX_N + 2; i += blockDim.x) s_cum[i] = plan->cum[i];
    __syncthreads();
    TpSplitShared sh{wp, wm, n, k, p, q, rq->dbits, rsp, plan->min_w};
    const int c_req = rq->c_req;
    const int kcols = k + 1;
    auto C = [&](int a, int b) -> u64 { return (b < 0 || a < b) ? 0ull : s.binom[a * kcols + b]; };
    TpSplitWarp wst{tstack, va, bstack, 0, 0, 0, lane};

    for (;;) {
        unsigned t_idx = 0;
        if (lane == 0) t_idx = atomicAdd(work_counter, 1u);
        t_idx = __shfl_sync(0xFFFFFFFFu, t_idx, 0);
        // Shard s of P owns item i when (i + i / P) mod P == s: the residue a shard takes rotates from
        // one group of P items to the next, so no periodic pattern of the item list (slices of one
        // pivot, chunks of one slice) lands on the same shard every time.
        const unsigned rot = (unsigned)shard_idx + (unsigned)shard_cnt - t_idx % (unsigned)shard_cnt;
        const u64 item64 = (u64)t_idx * (unsigned)shard_cnt + rot % (unsigned)shard_cnt;
        if (item64 >= total_items) break;
        const unsigned item = (unsigned)item64;
        // ---- item -> (pivot, chunk of A's, slice of the first B'' position) -----------------
        int e = 0;
        {
            const int ne = n_entries;
            for (int base = 0; base < ne; base += 32) {
                const int idx = base + lane;
                const bool ok = idx < ne && s_item_base[idx] <= item;
                const unsigned m = __ballot_sync(0xFFFFFFFFu, ok);
                if (m) e = base + 31 - __clz(m);
            }
        }
        const int jt = plan->jt_hi - e;
        const unsigned local = item - s_item_base[e];
        const unsigned nchunks = s_chunks[e];
        const unsigned slice = local / nchunks;
        const unsigned ac = local - slice * nchunks;
        TpSlice sl{q - 1, jt, 0, TP_MAX_N, false};
        if (D > 0) {
            auto cum = [&](int b) -> int { return s_cum[b]; };
            // (plan->b_star is read here, per item, rather than held in a register across the search)
            if (!tp_split_slice(C, cum, jt, q, (int)s_pieces[e], plan->b_star, (int)slice, sl)) continue; // empty
        }
        const int b0_lo = sl.b0_lo, b0_hi = sl.b0_hi;
        // Everything else the slice decides is settled here, so that only b0_lo / b0_hi stay live over
        // the lane set-up below: the count of what the item scores, and the bounds on the second walked
        // position, which the walk reads from the warp's byte area once per b0.  (All lanes left the
        // previous item through a __syncwarp, nobody reads that area any more.)
        {
            u64 per_lane; // B sets this item scores per lane
            if (D == 0) per_lane = C(jt, 2);
            else if (sl.cut_again) per_lane = C(sl.b1_hi, q - 1) - C(sl.b1_lo, q - 1);
            else per_lane = C(b0_hi, q) - C(b0_lo, q);
            if (lane == 0) {
                const u64 na = C(n - 1 - jt, p - 1);
                counted += per_lane * (u64)min((u64)32, na - (u64)ac * 32);
                bstack[8] = (uint8_t)sl.b1_lo;
                bstack[9] = (uint8_t)sl.b1_hi;
            }
        }

        // ---- this lane's A: rank -> p-1 offsets above the pivot, colexicographic unrank --------
        // (largest first; the smallest offset varies fastest, so neighbouring lanes read
        // neighbouring rows).  a0/a1 hold the positions one per byte, descending, pivot last.
        const int mA = n - 1 - jt;
        const u64 NA = C(mA, p - 1);
        u64 rank = (u64)ac * 32 + lane;
        const int active = rank < NA;
        if (!active) rank = NA - 1;
        u64 a0 = 0, a1 = 0;
        {
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
        auto elem = [&](int i) { return (int)(((i < 8 ? a0 : a1) >> (8 * (i & 7))) & 0xFF); };
        // ---- V = sum of the A rows (positions < jt) -> shared memory; sA = score inside A ----
        __syncwarp(); // the previous item's slow-path lanes are done with va
        int sA = c_req;
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
        wst.a0 = a0;
        wst.a1 = a1;
        wst.active = active;

        unsigned P[TP_SPLIT_PREGS];
        if constexpr (D == 0) {
            tp_split_make_p<M, false>(P, (const uint4 *)va + lane, (const uint4 *)tstack, nullptr, jt);
            const unsigned run = tp_split_pairs<M>(P, (const uint4 *)tp_split_smem, jt - 1);
            const int thr = min(max(*(volatile int *)&s_bound - sA + 1, 0), 0xFFFF);
            const unsigned thr2 = (unsigned)thr * 0x10001u;
            if (__vminu2(run, thr2) != thr2 && active)
                tp_split_slow<M>(sh, s, wst, 0, 0, jt, sA, &s_bound, best);
        } else {
            TpSplitWalk<M, D, 0>::run(sh, s, wst, P, b0_hi, max(b0_lo, q - 1), sA, 0, &s_bound, best);
        }
        __syncwarp();
    }
}

// Single request (batch.count == 0): the whole grid shares the request and its item counter, the
// last block reduces the partials.  Batch: every block takes whole requests off batch.order.
template <int M, int D>
__global__ void __launch_bounds__(TP_SPLIT_THREADS, 2)
tp_exhaustive_split(const uint8_t *__restrict__ W, int nd, const u64 *__restrict__ binom_g,
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
        TpKey best = {TP_KEY_INF_HI, TP_KEY_INF_HI};
        u64 counted = 0;
        tp_split_search<M, D>(W, nd, binom_g, reqs + slot, plans + plan, shard_idx, shard_cnt, wc, best, counted);
        if (batch.count > 0) {
            tp_block_reduce(best, counted);
            if (threadIdx.x == 0) {
                batch.results[out].key = best;
                batch.results[out].scored = counted;
            }
        } else {
            // (the launcher's memset arms the work counter; the last block also leaves it at zero, which is
            // what kernel TRI, sharing the word, expects to find)
            if (tp_block_finish(best, counted, sc) && threadIdx.x == 0) *work_counter = 0u;
        }
    }
}

static size_t tp_split_smem_bytes(int n, int k, int M, int D) {
    size_t off = (tp_smem_bytes(n, k) + 15) & ~(size_t)15;
    off += (size_t)n * tp_split_rsp(n) * 2 + (size_t)n * (M + 8) * 2 + (size_t)TP_MAX_N * 2;
    off += (size_t)TP_SPLIT_WARPS * ((size_t)(D + 1) * M * 2 + (size_t)(M / 8) * 32 * 16 + 16);
    return off;
}

// h describes the request (single mode) or carries the largest n and k of the batch (batch mode).
template <int M, int D>
static cudaError_t launch_split_md(const uint8_t *W, int nd, const u64 *binom, const TpReq *d_rq, const TpReq &h,
                                   const TpPlan *d_plan, int shard_idx, int shard_cnt, const TpScratch &sc,
                                   unsigned *d_counter, int sm_count, cudaStream_t st, TpLaunchInfo *info,
                                   const TpBatch &batch) {
    const size_t smem = tp_split_smem_bytes(h.n, h.k, M, D);
    cudaError_t e = tp_opt_in_max_smem<tp_exhaustive_split<M, D>>(smem);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(batch.count > 0 ? batch.grab : d_counter, 0, sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    int blocks = sm_count * 2;
    if (batch.count > 0 && batch.count < blocks) blocks = batch.count;
    tp_exhaustive_split<M, D><<<blocks, TP_SPLIT_THREADS, smem, st>>>(W, nd, binom, d_rq, d_plan, shard_idx,
                                                                      shard_cnt, sc, d_counter, batch);
    if (info) {
        info->launches = 1;
        info->blocks = blocks;
        info->threads = TP_SPLIT_THREADS;
    }
    return cudaGetLastError();
}

// One translation unit per (M, D) defines this for its instantiation.
#define TP_DEFINE_SPLIT_LAUNCHER(M, D)                                                                           \
    cudaError_t tp_launch_split_m##M##_d##D(const uint8_t *W, int nd, const unsigned long long *binom,          \
                                            const TpReq *d_rq, const TpReq &h, const TpPlan *d_plan, int si,    \
                                            int sn, const TpScratch &sc, unsigned *d_counter, int sms,          \
                                            cudaStream_t st, TpLaunchInfo *info, const TpBatch &batch) {        \
        return launch_split_md<M, D>(W, nd, binom, d_rq, h, d_plan, si, sn, sc, d_counter, sms, st, info, batch); \
    }
