// tp_split.cuh -- kernel SPLIT: the fast exhaustive scorer (included by tp_kernels.cu).
//
// Same contract as tp_exhaustive_lex (all C(n,k') subsets scored, Family + RefKey winner), but
// laid out so that a warp's 32 lanes never diverge and the pair weights are touched by
// 16-byte broadcast loads instead of per-lane byte gathers:
//
//   * positions are the canonical order reversed (position j = canonical n-1-j), a subset is
//     j_1 > j_2 > ... > j_k';
//   * the top p positions A = {j_1..j_p} belong to a LANE (32 consecutive A's of one pivot
//     jt = j_p per warp); the lane keeps V[j] = sum_{a in A} W[a][j], j < jt, in registers,
//     packed two 16-bit positions per register;
//   * the next D = q-2 positions B'' are walked by the WARP (uniform nested loops); the
//     vector T[j] = u[j] + sum_{b in B''} W[b][j] lives in shared memory, one per level;
//   * the last two positions jx > jy are an unrolled triangular nest over registers
//     P[j] = V[j] + T[j]:  min_{jy<jx} P[jy] + Wm[jx][jy]  is four VIADDMNMX.U16x2 per
//     16-byte row chunk (two subsets per lane-instruction), then one more VIADDMNMX adds P[jx].
//     Rows of Wm carry TP_BIG at jy >= jx, so the triangular mask costs nothing.
//   * only the packed minimum is compared with the best Family score so far; the rare leaf
//     that beats it is re-walked by the slow path, which applies Family/RefKey exactly as the
//     LEX kernel does (tp_family_key).
//
// Work = items (pivot jt, chunk of 32 A's, slice of the first B'' position), heaviest first,
// handed out by an atomic counter; shard s of P takes items s, s+P, ...
#pragma once

struct TpSplitShared { // block-shared tables (pointers into dynamic shared memory)
    const uint16_t *wp; // [n][rsp] plain weights, reversed positions (all n columns: A sits above M)
    const uint16_t *wm; // [n][M+8] row jx: W for jy < jx, TP_BIG for jy >= jx
    int n, k, p, q, dbits, rsp;
};

// Row stride (halfwords) of the plain table: covers all n positions, a multiple of 8 whose
// 16-byte chunk count is odd so that consecutive rows start in different bank groups.
__host__ __device__ inline int tp_split_rsp(int n) {
    const int r8 = (n + 7) & ~7;
    return r8 + (((r8 / 8) & 1) ? 16 : 8);
}

template <int M>
struct TpSplitCfg {
    static constexpr int RS = M + 8;   // row stride in halfwords (one extra 16-byte chunk skews the banks)
    static constexpr int RS4 = RS / 8; // ... in uint4
    static constexpr int NREG = M / 2;
    static constexpr int NCHUNK = M / 8;
};

// One row of the triangular nest: all pairs (JX, jy < JX).
template <int M, int JX>
__device__ __forceinline__ void tp_split_row(const unsigned (&P)[M / 2], const uint4 *__restrict__ wm4,
                                             unsigned &run) {
    constexpr int NC = (JX + 7) / 8;
    const unsigned px2 = __byte_perm(P[JX / 2], 0, (JX & 1) ? 0x3232 : 0x1010);
    unsigned acc0 = TP_BIG2 + TP_BIG2, acc1 = TP_BIG2 + TP_BIG2;
#pragma unroll
    for (int c = 0; c < NC; c++) {
        const uint4 r = wm4[JX * TpSplitCfg<M>::RS4 + c];
        acc0 = __viaddmin_u16x2(P[4 * c + 0], r.x, acc0);
        acc1 = __viaddmin_u16x2(P[4 * c + 1], r.y, acc1);
        acc0 = __viaddmin_u16x2(P[4 * c + 2], r.z, acc0);
        acc1 = __viaddmin_u16x2(P[4 * c + 3], r.w, acc1);
    }
    acc0 = __vminu2(acc0, acc1);
    run = __viaddmin_u16x2(acc0, px2, run);
}

#define TP_ROW_CASE(J)                                                \
    case J:                                                           \
        if constexpr ((J) < M && (J) >= 1) tp_split_row<M, (J)>(P, wm4, run); \
        [[fallthrough]];
#define TP_ROW_CASE8(B)                                                                              \
    TP_ROW_CASE(B + 7) TP_ROW_CASE(B + 6) TP_ROW_CASE(B + 5) TP_ROW_CASE(B + 4) TP_ROW_CASE(B + 3) \
    TP_ROW_CASE(B + 2) TP_ROW_CASE(B + 1) TP_ROW_CASE(B + 0)

// min over all pairs jx > jy with jx <= start of P[jx] + P[jy] + W[jx][jy], packed per half.
template <int M>
__device__ __forceinline__ unsigned tp_split_pairs(const unsigned (&P)[M / 2], const uint4 *__restrict__ wm4,
                                                   int start) {
    unsigned run = TP_BIG2 + TP_BIG2;
    switch (start) {
        TP_ROW_CASE8(120) TP_ROW_CASE8(112) TP_ROW_CASE8(104) TP_ROW_CASE8(96) TP_ROW_CASE8(88) TP_ROW_CASE8(80)
        TP_ROW_CASE8(72) TP_ROW_CASE8(64) TP_ROW_CASE8(56) TP_ROW_CASE8(48) TP_ROW_CASE8(40) TP_ROW_CASE8(32)
        TP_ROW_CASE8(24) TP_ROW_CASE8(16) TP_ROW_CASE8(8) TP_ROW_CASE8(0)
    default: break;
    }
    return run;
}

// P = V + T (+ masked row of the leaf position b) for every chunk holding a position < limit.
template <int M, bool HAS_ROW>
__device__ __forceinline__ void tp_split_make_p(unsigned (&P)[M / 2], const unsigned (&V)[M / 2],
                                                const uint4 *__restrict__ t4, const uint4 *__restrict__ wb4,
                                                int limit) {
#pragma unroll
    for (int c = 0; c < M / 8; c++) {
        if (8 * c < limit) {
            const uint4 t = t4[c];
            if (HAS_ROW) {
                const uint4 w = wb4[c];
                P[4 * c + 0] = V[4 * c + 0] + t.x + w.x;
                P[4 * c + 1] = V[4 * c + 1] + t.y + w.y;
                P[4 * c + 2] = V[4 * c + 2] + t.z + w.z;
                P[4 * c + 3] = V[4 * c + 3] + t.w + w.w;
            } else {
                P[4 * c + 0] = V[4 * c + 0] + t.x;
                P[4 * c + 1] = V[4 * c + 1] + t.y;
                P[4 * c + 2] = V[4 * c + 2] + t.z;
                P[4 * c + 3] = V[4 * c + 3] + t.w;
            }
        }
    }
}

struct TpSplitWarp { // per-warp state that the slow path needs (all uniform unless noted)
    const uint16_t *tstack; // [D+1][M] level vectors
    const uint32_t *va;     // [M/2][32] lane-private V, packed pairs
    uint8_t *bstack;        // [TP_SPLIT_MAX_D] chosen B'' positions
    u64 a0, a1;             // lane: A positions, one per byte, descending
    int active;             // lane: this lane holds a real A
};

// Re-walk one leaf for one lane with scalar arithmetic and apply Family/RefKey to every pair
// that scores no worse than the bound.  base = c_req + sA + cross(A,B'') + score inside B''.
template <int M>
__device__ __noinline__ void tp_split_slow(const TpSplitShared &sh, const TpSmem &s, const TpSplitWarp &w,
                                           int d_used, int b_leaf, int limit, int base, int *s_bound,
                                           TpKey &best) {
    const int lane = threadIdx.x & 31, n = sh.n;
    const uint16_t *t = w.tstack + (d_used > 0 ? d_used - 1 : 0) * M;
    auto pval = [&](int j) {
        const unsigned pk = w.va[(j >> 1) * 32 + lane];
        int v = (j & 1) ? (pk >> 16) : (pk & 0xFFFF);
        v += t[j];
        if (d_used > 0) v += sh.wp[b_leaf * sh.rsp + j];
        return v;
    };
    for (int jx = 1; jx < limit; jx++) {
        const int px = base + pval(jx);
        for (int jy = 0; jy < jx; jy++) {
            const int sco = px + pval(jy) + sh.wp[jx * sh.rsp + jy];
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
}

template <int M, int D>
__global__ void __launch_bounds__(TP_SPLIT_THREADS, 2)
tp_exhaustive_split(const uint8_t *__restrict__ W, int nd, const u64 *__restrict__ binom_g,
                    const TpReq *__restrict__ rq, const TpPlan *__restrict__ plan, int shard_idx, int shard_cnt,
                    TpScratch sc, unsigned *work_counter) {
    using Cfg = TpSplitCfg<M>;
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int s_bound;
    const int n = rq->n, k = rq->k, p = plan->p, q = plan->q;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const TpSmem s = tp_smem_carve(smem, n, k);
    tp_stage_request(s, W, nd, binom_g, rq, k, &s_bound);

    // ---- block-shared 16-bit tables in reversed positions ------------------------------------
    size_t off = (tp_smem_bytes(n, k) + 15) & ~(size_t)15;
    const int rsp = tp_split_rsp(n);
    uint16_t *wp = (uint16_t *)(smem + off);
    off += (size_t)n * rsp * 2;
    uint16_t *wm = (uint16_t *)(smem + off);
    off += (size_t)n * Cfg::RS * 2;
    uint16_t *t0 = (uint16_t *)(smem + off); // u in reversed positions, all n of them (level 0 of T)
    off += (size_t)TP_MAX_N * 2;
    unsigned char *warp_base = smem + off + (size_t)warp * (((size_t)(D + 1) * M * 2) + (size_t)Cfg::NREG * 32 * 4 + 16);
    uint16_t *tstack = (uint16_t *)warp_base;                               // [D+1][M]
    uint32_t *va = (uint32_t *)(warp_base + (size_t)(D + 1) * M * 2);       // [M/2][32]
    uint8_t *bstack = (uint8_t *)(va + Cfg::NREG * 32);                     // [<=16]
    for (int i = threadIdx.x; i < n * rsp; i += blockDim.x) { // plain table, padding 0 (it is summed)
        const int j = i / rsp, jj = i - j * rsp;
        wp[i] = (jj < n && jj != j) ? (uint16_t)s.w[(n - 1 - j) * s.stride + (n - 1 - jj)] : (uint16_t)0;
    }
    for (int i = threadIdx.x; i < n * Cfg::RS; i += blockDim.x) { // masked table, padding BIG
        const int j = i / Cfg::RS, jj = i - j * Cfg::RS;
        wm[i] = (jj < j && jj < n) ? (uint16_t)s.w[(n - 1 - j) * s.stride + (n - 1 - jj)] : (uint16_t)TP_BIG;
    }
    for (int i = threadIdx.x; i < TP_MAX_N; i += blockDim.x) t0[i] = i < n ? s.u[n - 1 - i] : (uint16_t)0;
    if (threadIdx.x == 0) { // second seed: whole groups from the largest down (also a Family member)
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
        if (left == 0 && sco < s_bound) s_bound = sco;
    }
    __syncthreads();
    for (int i = lane; i < M; i += 32) tstack[i] = t0[i];
    __syncwarp();

    TpSplitShared sh{wp, wm, n, k, p, q, rq->dbits, rsp};
    const uint4 *wm4 = (const uint4 *)wm;
    const int c_req = rq->c_req;
    TpKey best = {TP_KEY_INF_HI, TP_KEY_INF_HI};
    u64 counted = 0;
    const int kcols = k + 1;
    auto C = [&](int a, int b) -> u64 { return (b < 0 || a < b) ? 0ull : s.binom[a * kcols + b]; };

    for (;;) {
        unsigned t_idx = 0;
        if (lane == 0) t_idx = atomicAdd(work_counter, 1u);
        t_idx = __shfl_sync(0xFFFFFFFFu, t_idx, 0);
        const u64 item64 = (u64)t_idx * (unsigned)shard_cnt + (unsigned)shard_idx;
        if (item64 >= plan->total_items) break;
        const unsigned item = (unsigned)item64;
        // ---- item -> (pivot, chunk of A's, slice of the first B'' position) -----------------
        int e = 0;
        {
            const int ne = plan->n_entries;
            int found = 0;
            for (int base = 0; base < ne; base += 32) {
                const int idx = base + lane;
                const bool ok = idx < ne && plan->item_base[idx] <= item;
                const unsigned m = __ballot_sync(0xFFFFFFFFu, ok);
                if (m) found = base + 31 - __clz(m);
            }
            e = found;
        }
        const int jt = plan->jt_hi - e;
        const unsigned local = item - plan->item_base[e];
        const unsigned nchunks = plan->n_chunks[e];
        const int piece = (int)(local / nchunks);
        const unsigned ac = local - (unsigned)piece * nchunks;
        const int b0_lo = plan->cuts[e][piece], b0_hi = plan->cuts[e][piece + 1];

        // ---- this lane's A: rank -> p-1 offsets above the pivot (lexicographic unrank) -------
        const int mA = n - 1 - jt;
        const u64 NA = C(mA, p - 1);
        u64 rank = (u64)ac * 32 + lane;
        const int active = rank < NA;
        if (!active) rank = NA - 1;
        int apos[TP_SPLIT_MAX_P];
        {
            int a = 0;
#pragma unroll
            for (int i = 0; i < TP_SPLIT_MAX_P; i++) {
                apos[i] = 0;
                if (i < p - 1) {
                    for (;; a++) {
                        const u64 cnt = C(mA - a - 1, p - 1 - i - 1);
                        if (rank < cnt) break;
                        rank -= cnt;
                    }
                    apos[i] = jt + 1 + a;
                    a++;
                }
            }
        }
        // descending list for the slow path: largest offsets first, pivot last
        u64 a0 = 0, a1 = 0;
        {
            int cnt = 0;
#pragma unroll
            for (int i = TP_SPLIT_MAX_P - 1; i >= 0; i--) {
                if (i < p - 1) {
                    if (cnt < 8) a0 |= (u64)apos[i] << (8 * cnt);
                    else a1 |= (u64)apos[i] << (8 * (cnt - 8));
                    cnt++;
                }
            }
            if (cnt < 8) a0 |= (u64)jt << (8 * cnt);
            else a1 |= (u64)jt << (8 * (cnt - 8));
        }
        // ---- V = sum of the A rows (positions < jt), sA = score inside A ---------------------
        unsigned V[Cfg::NREG];
#pragma unroll
        for (int i = 0; i < Cfg::NREG; i++) V[i] = 0;
        int sA = c_req;
#pragma unroll
        for (int i = 0; i < TP_SPLIT_MAX_P; i++) {
            if (i < p) {
                const int a = (i < p - 1) ? apos[i] : jt;
                const uint4 *row = (const uint4 *)(wp + a * rsp);
#pragma unroll
                for (int c = 0; c < Cfg::NCHUNK; c++) {
                    if (8 * c < jt) {
                        const uint4 r = row[c];
                        V[4 * c + 0] += r.x;
                        V[4 * c + 1] += r.y;
                        V[4 * c + 2] += r.z;
                        V[4 * c + 3] += r.w;
                    }
                }
                sA += t0[a];
#pragma unroll
                for (int i2 = 0; i2 < TP_SPLIT_MAX_P; i2++)
                    if (i2 < i) sA += wp[a * rsp + apos[i2]]; // i2 < i <= p-1: apos[i2] is a real element
            }
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < Cfg::NREG; i++)
            if (2 * i < jt) va[i * 32 + lane] = V[i]; // words holding a position < jt (uniform test)
        __syncwarp();

        TpSplitWarp wst{tstack, va, bstack, a0, a1, active};
        // number of B sets this item scores per lane
        u64 per_lane;
        if (D == 0) per_lane = C(jt, 2);
        else per_lane = C(b0_hi, q) - C(b0_lo, q);
        if (lane == 0) counted += per_lane * (u64)min((u64)32, NA - (u64)ac * 32);

        unsigned P[Cfg::NREG];
        if constexpr (D == 0) {
            tp_split_make_p<M, false>(P, V, (const uint4 *)tstack, nullptr, jt);
            const unsigned run = tp_split_pairs<M>(P, wm4, jt - 1);
            const int thr = max(*(volatile int *)&s_bound - sA + 1, 0);
            const unsigned thr2 = (unsigned)min(thr, 0xFFFF) * 0x10001u;
            if (__vminu2(run, thr2) != thr2 && active)
                tp_split_slow<M>(sh, s, wst, 0, 0, jt, sA, &s_bound, best);
        } else {
            // ---- uniform nested walk over B'' = (b_0 > b_1 > ... > b_{D-1}) ------------------
            // level l picks b_l in [q-1-l, hi_l); levels 0..D-2 extend the T stack, the last one is the leaf.
            int bcur[TP_SPLIT_MAX_D];   // uniform
            int lbs[TP_SPLIT_MAX_D + 1]; // lane: sA + cross(A, b_0..b_{l-1})
            int sbs[TP_SPLIT_MAX_D + 1]; // uniform: score inside b_0..b_{l-1}
            lbs[0] = sA;
            sbs[0] = 0;
            int level = 0;
            bcur[0] = b0_hi;
#pragma unroll 1
            for (;;) {
                // advance at `level`
                int b = -1;
#pragma unroll
                for (int l = 0; l < D; l++)
                    if (l == level) b = --bcur[l];
                const int lo = (level == 0) ? max(b0_lo, q - 1) : (q - 1 - level);
                if (b < lo) {
                    if (level == 0) break;
                    level--;
                    continue;
                }
                const uint16_t *tl = tstack + level * M;
                int lb_next = 0, sb_next = 0;
                {
                    const unsigned pk = va[(b >> 1) * 32 + lane];
                    const int vab = (b & 1) ? (int)(pk >> 16) : (int)(pk & 0xFFFF);
#pragma unroll
                    for (int l = 0; l < D; l++)
                        if (l == level) {
                            lb_next = lbs[l] + vab;
                            sb_next = sbs[l] + tl[b];
                        }
                }
                if (level < D - 1) {
                    // T_{level+1} = T_level + W[b] on positions < b
                    uint16_t *tn = tstack + (level + 1) * M;
                    const uint32_t *src = (const uint32_t *)tl, *rw = (const uint32_t *)(wp + b * rsp);
                    uint32_t *dst = (uint32_t *)tn;
                    __syncwarp(); // lanes still in the slow path of an earlier leaf read T_{level+1} and bstack
                    if (lane == 0) bstack[level] = (uint8_t)b;
                    for (int wd = lane; wd < M / 2; wd += 32) dst[wd] = src[wd] + rw[wd];
                    __syncwarp();
#pragma unroll
                    for (int l = 0; l < D; l++)
                        if (l == level) {
                            lbs[l + 1] = lb_next;
                            sbs[l + 1] = sb_next;
                            if (l + 1 < D) bcur[l + 1] = b;
                        }
                    level++;
                    continue;
                }
                // ---- leaf: pairs below b ----------------------------------------------------
                if (b < 2) continue;
                tp_split_make_p<M, true>(P, V, (const uint4 *)tl, (const uint4 *)(wm + b * Cfg::RS), b);
                const unsigned run = tp_split_pairs<M>(P, wm4, b - 1);
                const int base = lb_next + sb_next;                   // lane
                const int thr = max(*(volatile int *)&s_bound - base + 1, 0);
                const unsigned thr2 = (unsigned)min(thr, 0xFFFF) * 0x10001u;
                if (__vminu2(run, thr2) != thr2 && active)
                    tp_split_slow<M>(sh, s, wst, D, b, b, base, &s_bound, best);
            }
        }
        __syncwarp();
    }
    tp_block_finish(best, counted, sc); // the work counter is re-armed by the launcher's memset
}

static size_t tp_split_smem_bytes(int n, int k, int M, int D) {
    size_t off = (tp_smem_bytes(n, k) + 15) & ~(size_t)15;
    off += (size_t)n * tp_split_rsp(n) * 2 + (size_t)n * (M + 8) * 2 + (size_t)TP_MAX_N * 2;
    off += (size_t)TP_SPLIT_WARPS * ((size_t)(D + 1) * M * 2 + (size_t)(M / 2) * 32 * 4 + 16);
    return off;
}

template <int M, int D>
static cudaError_t launch_split_md(const uint8_t *W, int nd, const u64 *binom, const TpReq *d_rq, const TpReq &h,
                                   const TpPlan *d_plan, int shard_idx, int shard_cnt, const TpScratch &sc,
                                   unsigned *d_counter, int sm_count, cudaStream_t st, TpLaunchInfo *info) {
    const size_t smem = tp_split_smem_bytes(h.n, h.k, M, D);
    cudaError_t e = cudaFuncSetAttribute(tp_exhaustive_split<M, D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(d_counter, 0, sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    const int blocks = sm_count * 2;
    tp_exhaustive_split<M, D><<<blocks, TP_SPLIT_THREADS, smem, st>>>(W, nd, binom, d_rq, d_plan, shard_idx,
                                                                      shard_cnt, sc, d_counter);
    if (info) {
        info->launches = 1;
        info->blocks = blocks;
        info->threads = TP_SPLIT_THREADS;
    }
    return cudaGetLastError();
}

template <int M>
static cudaError_t launch_split_m(int D, const uint8_t *W, int nd, const u64 *binom, const TpReq *d_rq,
                                  const TpReq &h, const TpPlan *d_plan, int si, int sn, const TpScratch &sc,
                                  unsigned *d_counter, int sms, cudaStream_t st, TpLaunchInfo *info) {
    switch (D) {
    case 0: return launch_split_md<M, 0>(W, nd, binom, d_rq, h, d_plan, si, sn, sc, d_counter, sms, st, info);
    case 1: return launch_split_md<M, 1>(W, nd, binom, d_rq, h, d_plan, si, sn, sc, d_counter, sms, st, info);
    case 2: return launch_split_md<M, 2>(W, nd, binom, d_rq, h, d_plan, si, sn, sc, d_counter, sms, st, info);
    case 3: return launch_split_md<M, 3>(W, nd, binom, d_rq, h, d_plan, si, sn, sc, d_counter, sms, st, info);
    case 4: return launch_split_md<M, 4>(W, nd, binom, d_rq, h, d_plan, si, sn, sc, d_counter, sms, st, info);
    }
    return cudaErrorInvalidValue;
}

cudaError_t tp_launch_exhaustive_split(const uint8_t *d_weight, int n_devices, const u64 *d_binom,
                                       const TpReq *d_rq, const TpReq &h, const TpPlan *d_plan, const TpPlan &hp,
                                       int shard_idx, int shard_cnt, const TpScratch &sc, unsigned *d_counter,
                                       int sm_count, cudaStream_t st, TpLaunchInfo *info) {
    const int D = hp.q - 2, need = h.n - hp.p; // largest pivot = number of positions a lane keeps
    if (need <= 64)
        return launch_split_m<64>(D, d_weight, n_devices, d_binom, d_rq, h, d_plan, shard_idx, shard_cnt, sc,
                                  d_counter, sm_count, st, info);
    if (need <= 96)
        return launch_split_m<96>(D, d_weight, n_devices, d_binom, d_rq, h, d_plan, shard_idx, shard_cnt, sc,
                                  d_counter, sm_count, st, info);
    return cudaErrorInvalidValue;
}
