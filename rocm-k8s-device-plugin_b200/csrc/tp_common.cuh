// tp_common.cuh -- types shared by the host side of the C ABI and the sm_100a kernels.
//
// Vocabulary (follows the reference, internal/pkg/allocator/device.go):
//   device     one KFD node with a render minor (a GPU or a partition), index 0..n-1
//   group      the partitions of one physical GPU (DevicePartitions, device.go:75-80)
//   free       available and not required (filterPartitions, device.go:310-351)
//   canonical  order of the free devices: groups in filterPartitions order
//              (fewest free first, then ParentId), NodeId ascending inside a group
//   k'         size - len(required): how many free devices a candidate holds
//   Family     the sets the reference's BFS can emit (device.go:375-440): per group
//              nothing / the whole group / for at most one group a lowest-NodeId prefix
//   RefKey     (m, sigma): groups touched, then whole groups ascending + the partial
//              group: the order in which the BFS appends candidates
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "tp_allocator.h"

#define TP_MAX_N TP_MAX_DEVICES
#define TP_MAX_K TP_MAX_SUBSET
#define TP_SIGMA_BITS 96 /* bits of the packed group sequence: 32 in key.hi + 64 in key.lo */
#define TP_KEY_SCORE_SHIFT 40
#define TP_KEY_GROUPS_SHIFT 32

// One prepared request, as the kernels read it (device resident, ~1 KB).
struct TpReq {
    int32_t n;        // free devices
    int32_t k;        // k'
    int32_t n_groups; // groups with at least one free device
    int32_t n_req;    // required devices
    int32_t dbits;    // bits per sigma digit = ceil(log2(n_groups)), >= 1
    int32_t c_req;    // weight among the required devices (every pair once)
    uint64_t total;   // C(n, k)
    uint8_t canon[TP_MAX_N];    // canonical position -> device index
    uint8_t group_of[TP_MAX_N]; // canonical position -> group position
    uint8_t gstart[TP_MAX_N];   // group position -> first canonical position
    uint8_t gsize[TP_MAX_N];    // group position -> free devices in the group
    uint8_t req[TP_MAX_N];      // required device indices, request order
};

// 128-bit key, compared as (hi, lo); see tp_key in tp_allocator.h.
struct TpKey {
    unsigned long long hi, lo;
};

// "no candidate": larger than any real key (scores < 2^15) and still positive as int64, so a
// signed 64-bit MIN all-reduce (ncclInt64 / torch.int64) orders it like the unsigned compare.
#define TP_KEY_INF_HI 0x7FFFFFFFFFFFFFFFull

__host__ __device__ __forceinline__ bool tp_key_less(const TpKey &a, const TpKey &b) {
    return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo);
}

// 0:1 | score:23 | m:8 | sigma digits, most significant first, dbits each, in a 96-bit field.
__host__ __device__ __forceinline__ void tp_key_begin(TpKey &k, int score, int m) {
    k.hi = ((unsigned long long)(unsigned)score << TP_KEY_SCORE_SHIFT) |
           ((unsigned long long)(unsigned)m << TP_KEY_GROUPS_SHIFT);
    k.lo = 0;
}
__host__ __device__ __forceinline__ void tp_key_put(TpKey &k, int i, int dbits, unsigned digit) {
    int shift = TP_SIGMA_BITS - (i + 1) * dbits; // bit position inside the 96-bit field
    if (shift >= 64) {
        k.hi |= (unsigned long long)digit << (shift - 64);
    } else if (shift + dbits <= 64) {
        k.lo |= (unsigned long long)digit << shift;
    } else { // straddles the two words
        k.lo |= (unsigned long long)digit << shift;
        k.hi |= (unsigned long long)digit >> (64 - shift);
    }
}
__host__ __device__ __forceinline__ unsigned tp_key_get(const TpKey &k, int i, int dbits) {
    int shift = TP_SIGMA_BITS - (i + 1) * dbits;
    unsigned long long v;
    if (shift >= 64)
        v = k.hi >> (shift - 64);
    else if (shift + dbits <= 64)
        v = k.lo >> shift;
    else
        v = (k.lo >> shift) | (k.hi << (64 - shift));
    return (unsigned)(v & ((1ull << dbits) - 1));
}
__host__ __device__ __forceinline__ int tp_key_score(const TpKey &k) { return (int)(k.hi >> TP_KEY_SCORE_SHIFT); }
__host__ __device__ __forceinline__ int tp_key_groups(const TpKey &k) { return (int)((k.hi >> TP_KEY_GROUPS_SHIFT) & 0xFF); }

// Per-launch scratch in global memory: one partial winner per block, reduced by the
// last block to finish (ticket), so one launch yields the final key with no second kernel.
// Winner exchange between the ranks of a multi-process run (one process per GPU), fused into the
// search kernel's last block: it stores (key, count, epoch) into slot [epoch & 1][rank] of EVERY
// rank's buffer -- plain st.global through peer mappings (CUDA IPC), 32 bytes per peer over NVLink --
// then waits until the `world` slots of its own buffer carry this epoch and takes their minimum,
// so that every rank's kernel ends with the global winner and no collective kernel follows.
// Two slots per rank (epoch parity): a rank cannot be two exchanges ahead of another, since it
// leaves an exchange only after all ranks have published for it.  world == 0: no exchange.
#define TP_EX_SLOT_WORDS 4 /* key.hi, key.lo, subsets scored, epoch (written last) */
struct TpExchangeDev {
    unsigned long long *peer[TP_MAX_GPUS]; // every rank's buffer, own included: [2][world][TP_EX_SLOT_WORDS]
    int rank, world;
    unsigned long long epoch; // > 0, the same on all ranks for one exchange
};

struct TpScratch {
    TpKey *partial;               // [max_blocks]
    unsigned long long *scored;   // running count of subsets scored
    unsigned int *ticket;         // blocks finished
    TpKey *out;                   // final key
    unsigned long long *out_scored;
    TpExchangeDev ex;
};

#define TP_MAX_BLOCKS 4096

// Batch mode (tp_allocate_batch): one block per request.  `slot` indexes the batch's request and
// result arrays, `plan` the (SPLIT-only) plan array.
struct TpBatchEntry {
    int32_t slot;       // request
    int32_t plan;       // its plan (SPLIT)
    int32_t out;        // where in `results` this entry reports
    int16_t share_idx;  // heavy requests are cut into shares (items share_idx, share_idx + share_cnt, ..)
    int16_t share_cnt;  // that different blocks take; the host min-reduces the shares
};
struct TpBatchOut {
    TpKey key;
    unsigned long long scored;
    unsigned long long pad;
};
struct TpBatch { // count == 0: single-request mode (all blocks share request 0, cross-block finish)
    const TpBatchEntry *order;
    TpBatchOut *results;
    unsigned *grab; // next entry of `order` to hand out
    int32_t count;
    int32_t pad;
};

// ---- kernel SPLIT (tp_split.cuh): constants and the host-built work plan ----------------------
#define TP_BIG 0x3000u
#define TP_BIG2 0x30003000u
#define TP_SPLIT_PREGS 64 /* lane registers holding packed P values: 128 positions, or 2 x 64 for two leaves */
#define TP_SPLIT_WARPS 8
#define TP_SPLIT_THREADS (TP_SPLIT_WARPS * 32)
#define TP_SPLIT_MAX_PIECES 64
#define TP_SPLIT_MAX_P 12
#define TP_SPLIT_MAX_D 4
#define TP_SPLIT_SCORE_LIMIT 12000 /* every partial sum must stay below TP_BIG */

// Work plan: items = (pivot, chunk of 32 lane sets, slice).  Item index inside pivot entry e:
// slice * n_chunks[e] + chunk.  The slices of a pivot jt are
//   * n_pieces[e] equal-work slices of the first walked position b0 over [q-1, min(jt, b_star)), then
//   * every single value b0 in [b_star, jt), cut into cum[b0+1] - cum[b0] equal-work parts of the
//     second walked position b1: one b0 is the finest grain of the first kind, and at the top
//     pivots one b0 still holds tens of thousands of subsets per lane (one warp, serial), which
//     becomes the tail of a shard or of a mid-size search.
struct TpPlan {
    int32_t p, q;         // |A|, |B''| + 2
    int32_t jt_hi;        // pivots run from jt_hi = n - p down to q
    int32_t n_entries;    // jt_hi - q + 1
    uint32_t total_items;
    int32_t min_w;        // smallest weight between two free devices
    uint32_t item_base[TP_MAX_N]; // entry e <-> pivot jt_hi - e
    uint32_t n_chunks[TP_MAX_N];  // chunks of 32 A's
    uint8_t n_pieces[TP_MAX_N];   // slices of the first walked position below b_star, see tp_split_cut
    int32_t b_star;               // values b0 >= b_star are items of their own, cut again (see cum)
    uint16_t cum[TP_MAX_N + 2];   // cum[b] = sub-slices of the single values b_star .. b-1
    int32_t tri_gcap;             // kernel TRI: leaf pairs (leaves 2g+1, 2g+2) that have a table in shared memory
};
#define TP_SPLIT_MAX_SUB 16

struct TpSlice {
    int b0_lo, b0_hi; // first walked position in [b0_lo, b0_hi)
    int b1_lo, b1_hi; // second walked position, further restricted when cut_again
    bool cut_again;
};

// Slice i of np of the first walked position b0 in [q-1, jt): boundaries put about the same number
// of subsets (C(b,q) counts those with b0 < b) in every slice; a slice can be empty, since one
// b0 value is the finest grain.  Used by the kernel (item -> slice) and by the host (counting).
// pair (kernel TRI, which takes the values of its last walked position two at a time from the top): an inner
// boundary keeps an even distance to the top of the range, so that only the lowest slice can end on a single value.
template <class BinomFn>
__host__ __device__ inline int tp_split_cut(BinomFn C, int jt, int q, int np, int i, bool pair = false) {
    if (i <= 0) return q - 1;
    if (i >= np) return jt;
    const unsigned long long tot = C(jt, q); // (a 32-bit division where it fits: this runs once per work item)
    const unsigned long long target = (tot >> 25) ? tot * (unsigned long long)i / (unsigned long long)np
                                                  : (unsigned long long)((unsigned)tot * (unsigned)i / (unsigned)np);
    int lo = q - 1, hi = jt; // smallest b with C(b,q) >= target
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (C(mid, q) >= target) hi = mid;
        else lo = mid + 1;
    }
    if (pair && ((jt - lo) & 1)) lo++;
    return lo;
}

// Subsets per lane in a slice (q positions below the lane set).
template <class BinomFn>
__host__ __device__ inline unsigned long long tp_split_slice_count(BinomFn C, const TpSlice &sl, int q) {
    if (!sl.cut_again) return C(sl.b0_hi, q) - C(sl.b0_lo, q);
    unsigned long long sum = 0; // every first value of the slice, cut by the position below it
    for (int b0 = sl.b0_lo; b0 < sl.b0_hi; b0++) {
        const int hi = sl.b1_hi < b0 ? sl.b1_hi : b0, lo = sl.b1_lo < b0 ? sl.b1_lo : b0;
        sum += C(hi, q - 1) - C(lo, q - 1);
    }
    return sum;
}

// Kernel TRI with one walked position (q == 4) takes that position's values two at a time, so there the values at
// and above b_star become items in PAIRS (b, b - 1), b = jt - 1, jt - 3, ...: the pairs start at b_star or one above
// it, whichever leaves an even number of values below the pivot, and cum() sums the parts of the pair heads of
// one parity (cum(b + 2) = cum(b) + parts of head b).
__host__ __device__ inline int tp_split_bstar(int jt, int b_star, bool pair_b0) {
    if (!pair_b0 || jt <= b_star) return b_star;
    return b_star + ((jt - b_star) & 1);
}
// Slices of pivot jt beyond its np pieces: the parts of the single values (pairs) [b_star, jt).
template <class CumFn>
__host__ __device__ inline int tp_split_top_slices(CumFn cum, int jt, int b_star, bool pair_b0) {
    if (pair_b0) return jt - tp_split_bstar(jt, b_star, true) >= 2 ? (int)cum(jt + 1) : 0;
    return jt > b_star ? (int)cum(jt) : 0;
}

// Slice number -> ranges of the first two walked positions; false when the slice is empty.
// Shared by the kernel (item -> work) and the host (counting what a shard scores).
template <class BinomFn, class CumFn>
__host__ __device__ inline bool tp_split_slice(BinomFn C, CumFn cum, int jt, int q, int np, int b_star, int slice,
                                               TpSlice &o, bool pair = false) {
    o.b1_lo = 0;
    o.b1_hi = TP_MAX_N;
    o.cut_again = false;
    const bool pair_b0 = pair && q == 4;
    const int b_eff = tp_split_bstar(jt, b_star, pair_b0);
    if (slice < np) {
        const int jlow = jt < b_eff ? jt : b_eff;
        o.b0_lo = tp_split_cut(C, jlow, q, np, slice, pair);
        o.b0_hi = tp_split_cut(C, jlow, q, np, slice + 1, pair);
        return o.b0_lo < o.b0_hi;
    }
    if (pair_b0) {
        const int t = slice - np;
        int lo = 0, hi = (jt - b_eff) >> 1; // heads b_eff + 1 + 2 i: the largest i with cum(head) <= t
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if ((int)cum(b_eff + 1 + 2 * mid) <= t) lo = mid;
            else hi = mid;
        }
        const int b = b_eff + 1 + 2 * lo;
        const int nsub = (int)cum(b + 2) - (int)cum(b), sub = t - (int)cum(b);
        o.b0_lo = b - 1;
        o.b0_hi = b + 1;
        if (nsub > 1) { // cut by the position below the pair: the top of the triple
            o.b1_lo = tp_split_cut(C, b, q - 1, nsub, sub);
            o.b1_hi = tp_split_cut(C, b, q - 1, nsub, sub + 1);
            o.cut_again = true;
            return o.b1_lo < o.b1_hi;
        }
        return true;
    }
    const int t = slice - np;
    int lo = b_star, hi = jt; // largest b in [b_star, jt) with cum(b) <= t
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if ((int)cum(mid) <= t) lo = mid;
        else hi = mid;
    }
    const int nsub = (int)cum(lo + 1) - (int)cum(lo), sub = t - (int)cum(lo);
    o.b0_lo = lo;
    o.b0_hi = lo + 1;
    if (nsub > 1) {
        o.b1_lo = tp_split_cut(C, lo, q - 1, nsub, sub, pair);
        o.b1_hi = tp_split_cut(C, lo, q - 1, nsub, sub + 1, pair);
        o.cut_again = true;
        return o.b1_lo < o.b1_hi;
    }
    return true;
}
