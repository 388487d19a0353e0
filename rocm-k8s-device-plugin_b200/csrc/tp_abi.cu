// tp_abi.cu -- host side of the C ABI declared in include/tp_allocator.h.
//
// Mirrors the non-search parts of BestEffortPolicy.Allocate
// (internal/pkg/allocator/besteffort_policy.go:88-151): the ordered validation checks,
// the two short-circuits, filterPartitions' grouping and group order
// (device.go:310-351), and the NodeId -> output order mapping (:141-148).  The search
// itself (device.go:353-442 + the argmin) always runs on the GPU; there is no CPU path.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <functional>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "tp_kernels.cuh"

typedef unsigned long long u64;

namespace {

struct Ctx { // one in-flight call: stream, events, pinned staging, device request + scratch
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    TpReq *h_req = nullptr; // pinned; a TpPlan follows it in the same buffer (h_plan / d_plan)
    TpReq *d_req = nullptr;
    TpPlan *h_plan = nullptr, *d_plan = nullptr;
    struct Out {
        TpKey key;
        u64 scored;
    } *h_out = nullptr; // pinned
    unsigned char *d_scratch = nullptr;
    TpScratch scratch{};
    unsigned *d_counter = nullptr; // work counter of kernel SPLIT (a word of d_scratch)
    // latency path: the kernel writes key, count and a flag straight into this mapped host buffer
    struct Fast {
        TpKey key;
        u64 scored;
        volatile unsigned flag;
        unsigned pad;
    } *h_fast = nullptr;
    void *d_fast = nullptr;
    unsigned epoch = 0;
};

// Everything the host derives from one request before (and after) the GPU search.
struct Prep {
    TpReq rq;
    TpPlan plan;         // valid when kernel == TP_KERNEL_SPLIT
    int status = 0;      // negative tp_status, or 0
    int shortcut = 0;    // 1: return available as given; 2: return required as given
    int n_req = 0;
    int kernel = TP_KERNEL_LEX;
    bool family_only = false; // not enumerable and not within the FAMILY kernel's limits either
    u64 family_masks = 0;     // kernel FAMILY: 2^groups bitmasks to visit
    int shards = 1;           // how many shards the search will be split into (sizes the SPLIT items)
    bool batch = false;       // prepared for tp_allocate_batch: a block, not the grid, takes each share
    int share_shift = 24;     // ... and a request above 2^share_shift subsets is cut into shares
};

// fn(t) for t in [0, nthreads): the calling thread takes t == 0
template <class F>
void run_threads(int nthreads, F fn) {
    std::vector<std::thread> pool;
    for (int t = 1; t < nthreads; t++) pool.emplace_back(fn, t);
    fn(0);
    for (std::thread &t : pool) t.join();
}

inline bool uses_plan(int kernel) { return kernel == TP_KERNEL_SPLIT || kernel == TP_KERNEL_TRI; }

// tp_allocate_batch cuts a heavy SPLIT / TRI request into this many shares, each taken by one block.
inline int batch_shares(int shift, u64 total) {
    return total <= (1ull << shift) ? 1 : (int)std::min<u64>(16, (total >> shift) + 1);
}

} // namespace

namespace {
// One thread per extra GPU: launches and collects that GPU's shard while the caller does the same on its own.
struct GpuWorker {
    std::mutex mu;
    std::condition_variable cv;
    std::deque<std::function<void()>> jobs;
    std::atomic<int> n_jobs{0}; // (what the worker looks at while it spins between two calls)
    bool stop = false;
    std::thread th;
    GpuWorker() : th([this] { loop(); }) {}
    ~GpuWorker() {
        {
            std::lock_guard<std::mutex> lk(mu);
            stop = true;
        }
        cv.notify_one();
        th.join();
    }
    void submit(std::function<void()> f) {
        {
            std::lock_guard<std::mutex> lk(mu);
            jobs.push_back(std::move(f));
            n_jobs.fetch_add(1, std::memory_order_release);
        }
        cv.notify_one();
    }
    void loop() {
        // After a job the worker stays awake for a moment: calls come in bursts (one per container of an RPC, one
        // RPC after another during a pod storm), and a futex wake-up is 10-20 us of a 300 us call.
        auto spin_until = std::chrono::steady_clock::now();
        for (;;) {
            while (n_jobs.load(std::memory_order_acquire) == 0 && std::chrono::steady_clock::now() < spin_until) {
#if defined(__x86_64__) || defined(__i386__)
                __builtin_ia32_pause();
#endif
            }
            std::function<void()> f;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || !jobs.empty(); });
                if (jobs.empty()) return;
                f = std::move(jobs.front());
                jobs.pop_front();
                n_jobs.fetch_sub(1, std::memory_order_relaxed);
            }
            f();
            spin_until = std::chrono::steady_clock::now() + std::chrono::microseconds(200);
        }
    }
};
} // namespace

struct tp_policy {
    int n = 0, n_groups = 0;
    // multi-GPU: a complete policy per extra GPU (same topology) and the thread that drives it
    std::vector<tp_policy *> peers;
    std::vector<GpuWorker *> workers;
    u64 multi_min = 1ull << 31;
    std::vector<int32_t> node_id, group, parent_rank;
    std::vector<uint8_t> weight;
    bool weights_empty = true; // len(p2pWeights) == 0
    bool host_only = false;    // created with TP_DEVICE_NONE: no CUDA state at all
    int cuda_device = 0, sm_count = 148;
    int kernel = 0;
    u64 exhaustive_limit = 0;
    // development knobs, read from the environment ONCE in tp_create (never on the request path):
    // TP_SPLIT_Q (walk split), TP_SPLIT_CAP / TP_SPLIT_CAP1 (item sizes), TP_BATCH_TRACE (phase timings),
    // TP_BATCH_SHARE_SHIFT / TP_BATCH_ITEMS_PER_WARP (batch work plan)
    int tune_q = 0;
    u64 tune_cap = 0, tune_cap1 = 0;
    bool batch_trace = false;
    uint8_t *d_weight = nullptr;
    u64 *d_binom = nullptr;
    u64 binom[TP_MAX_N + 1][TP_BINOM_COLS];
    std::vector<Ctx> ctxs;
    std::vector<int> free_ctx;
    std::mutex mu;
    std::condition_variable cv;
    std::string last_cuda_error;
    // tp_allocate_batch: kBatchSlots batches in flight per policy, each with buffers (grown on demand), a copy
    // stream and kBatchSide kernel streams of its own
    static constexpr int kBatchSlots = 2, kBatchSide = 4;
    struct Batch {
        cudaStream_t stream = nullptr;          // H2D of the waves, D2H of the results
        cudaStream_t side[kBatchSide] = {};     // the kernel groups, round robin
        cudaEvent_t ev_in[2] = {}, ev_done[kBatchSide] = {};
        size_t cap_reqs = 0, cap_plans = 0;
        TpReq *h_reqs = nullptr, *d_reqs = nullptr;
        TpPlan *h_plans = nullptr, *d_plans = nullptr;
        TpBatchEntry *h_order = nullptr, *d_order = nullptr;
        TpBatchOut *h_results = nullptr, *d_results = nullptr;
        unsigned *d_grab = nullptr; // one counter per kernel group
    } batches[kBatchSlots];
    std::mutex batch_mu;
    std::condition_variable batch_cv;
    std::vector<int> free_batches{1, 0};
    // batch work plan: a request above 2^shift subsets is cut into (total >> shift) + 1 shares (at most 16), each
    // taken by one block (0: by the number of heavy requests in the batch -- many of them balance the SMs without
    // being cut, and every share pays the block's table build again); a share is cut into about
    // batch_items_per_warp items per warp
    int batch_share_shift = 0;
    int batch_items_per_warp = 4;
    bool batch_light_first = false; // wave order (tuning aid): the light requests first
};

namespace {

#define TP_CUDA(p, call)                                                                      \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            std::lock_guard<std::mutex> lk__((p)->mu);                                        \
            (p)->last_cuda_error = std::string(#call) + ": " + cudaGetErrorString(e__);       \
            return TP_ERR_CUDA;                                                               \
        }                                                                                     \
    } while (0)

struct DeviceGuard { // restores the caller's current device
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
    }
    ~DeviceGuard() {
        int cur = -1;
        cudaGetDevice(&cur);
        if (prev >= 0 && cur != prev) cudaSetDevice(prev);
    }
};

struct CtxLease {
    tp_policy *p;
    int idx;
    explicit CtxLease(tp_policy *pol) : p(pol) {
        std::unique_lock<std::mutex> lk(p->mu);
        p->cv.wait(lk, [&] { return !p->free_ctx.empty(); });
        idx = p->free_ctx.back();
        p->free_ctx.pop_back();
    }
    ~CtxLease() {
        {
            std::lock_guard<std::mutex> lk(p->mu);
            p->free_ctx.push_back(idx);
        }
        p->cv.notify_one();
    }
    Ctx &ctx() { return p->ctxs[idx]; }
};

int init_ctx(tp_policy *p, Ctx &c) {
    TP_CUDA(p, cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    TP_CUDA(p, cudaEventCreate(&c.ev0));
    TP_CUDA(p, cudaEventCreate(&c.ev1));
    static_assert(sizeof(TpReq) % 16 == 0, "the plan follows the request in one buffer");
    TP_CUDA(p, cudaHostAlloc((void **)&c.h_req, sizeof(TpReq) + sizeof(TpPlan), cudaHostAllocDefault));
    TP_CUDA(p, cudaHostAlloc((void **)&c.h_out, sizeof(Ctx::Out), cudaHostAllocDefault));
    TP_CUDA(p, cudaMalloc((void **)&c.d_req, sizeof(TpReq) + sizeof(TpPlan)));
    TP_CUDA(p, cudaHostAlloc((void **)&c.h_fast, sizeof(Ctx::Fast), cudaHostAllocMapped));
    memset((void *)c.h_fast, 0, sizeof(Ctx::Fast));
    TP_CUDA(p, cudaHostGetDevicePointer(&c.d_fast, (void *)c.h_fast, 0));
    c.h_plan = (TpPlan *)((unsigned char *)c.h_req + sizeof(TpReq));
    c.d_plan = (TpPlan *)((unsigned char *)c.d_req + sizeof(TpReq));
    const size_t partial_bytes = sizeof(TpKey) * TP_MAX_BLOCKS;
    const size_t total = partial_bytes + 64;
    TP_CUDA(p, cudaMalloc((void **)&c.d_scratch, total));
    TP_CUDA(p, cudaMemset(c.d_scratch, 0, total));
    unsigned char *tail = c.d_scratch + partial_bytes;
    c.scratch.partial = (TpKey *)c.d_scratch;
    c.scratch.out = (TpKey *)tail;                    // 16 B
    c.scratch.out_scored = (u64 *)(tail + 16);        // 8 B (contiguous with out: one D2H copy)
    c.scratch.scored = (u64 *)(tail + 32);
    c.scratch.ticket = (unsigned int *)(tail + 40);
    c.d_counter = (unsigned *)(tail + 48);
    return 0;
}

void batch_release(tp_policy::Batch &b) {
    if (b.h_reqs) cudaFreeHost(b.h_reqs);
    if (b.h_plans) cudaFreeHost(b.h_plans);
    if (b.h_order) cudaFreeHost(b.h_order);
    if (b.h_results) cudaFreeHost(b.h_results);
    if (b.d_reqs) cudaFree(b.d_reqs);
    if (b.d_plans) cudaFree(b.d_plans);
    if (b.d_order) cudaFree(b.d_order);
    if (b.d_results) cudaFree(b.d_results);
    if (b.d_grab) cudaFree(b.d_grab);
    if (b.stream) cudaStreamDestroy(b.stream);
    for (cudaStream_t st : b.side)
        if (st) cudaStreamDestroy(st);
    for (cudaEvent_t ev : b.ev_in)
        if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : b.ev_done)
        if (ev) cudaEventDestroy(ev);
}

void free_ctx(Ctx &c) {
    if (c.stream) cudaStreamDestroy(c.stream);
    if (c.ev0) cudaEventDestroy(c.ev0);
    if (c.ev1) cudaEventDestroy(c.ev1);
    if (c.h_req) cudaFreeHost(c.h_req);
    if (c.h_out) cudaFreeHost(c.h_out);
    if (c.h_fast) cudaFreeHost((void *)c.h_fast);
    if (c.d_req) cudaFree(c.d_req);
    if (c.d_scratch) cudaFree(c.d_scratch);
}

// Work plan of kernels SPLIT and TRI for a prepared request (out.rq filled): sets out.kernel.
// Returns false when the request is outside what those kernels take (k' < 4, too many positions
// per lane, scores that would not fit their 16-bit packed arithmetic); the LEX kernel then serves it.
// TRI (three positions against the triple table) takes k' >= 6 on n - p <= 80 positions; SPLIT the rest.
bool build_plan(const tp_policy *p, Prep &out) {
    const TpReq &rq = out.rq;
    const int n = rq.n, k = rq.k;
    if (k < 4 || k > TP_MAX_K) return false;
    int q = k >= 6 ? std::min(k - 4, TP_SPLIT_MAX_D + 2) : 2;
    if (p->tune_q) q = std::max(2, std::min({p->tune_q, k - 1, TP_SPLIT_MAX_D + 2})); // tuning aid
    int pp = k - q, tri_gcap = 0;
    if (k >= 6 && p->kernel != TP_KERNEL_SPLIT && !p->tune_q) {
        const int tq = std::max(4, std::min(k - 4, TP_TRI_MAX_DW + 3)), tpp = k - tq; // >= 1 walked position
        if (tpp <= TP_SPLIT_MAX_P && n - tpp >= tq && (tri_gcap = tp_tri_fit(n, k, tpp, tq)) > 0) {
            q = tq;
            pp = tpp;
        }
    }
    if (pp > TP_SPLIT_MAX_P || n - pp > TP_SPLIT_MAX_M || n - pp < q) return false;
    int max_w = 0, max_u = 0, min_w = 255;
    for (int i = 0; i < n; i++) {
        const uint8_t *row = &p->weight[(size_t)rq.canon[i] * p->n];
        int u = 0;
        for (int j = 0; j < n; j++) {
            max_w = std::max<int>(max_w, row[rq.canon[j]]);
            if (j != i) min_w = std::min<int>(min_w, row[rq.canon[j]]);
        }
        for (int r = 0; r < rq.n_req; r++) u += row[rq.req[r]];
        max_u = std::max(max_u, u);
    }
    if (k * (k - 1) / 2 * max_w + k * max_u > TP_SPLIT_SCORE_LIMIT) return false;

    TpPlan &pl = out.plan;
    memset(&pl, 0, sizeof pl);
    pl.p = pp;
    pl.q = q;
    pl.jt_hi = n - pp;
    pl.n_entries = n - pp - q + 1;
    pl.min_w = min_w;
    // Slice so that the longest item (one warp, serial: ~2.6 ns per subset and lane) stays a few
    // percent of what one GPU has to do, while an item is worth >= 2K subsets per lane: the
    // per-item set-up (unrank, lane vector) is latency-bound and costs throughput when items get
    // small (measured sweep on C4: 3.12 ms at 1K, 2.86 ms at 64K; at 8 shards 64K items leave a
    // 0.17 ms tail on a 0.37 ms shard).
    u64 cap0 = std::min<u64>(std::max<u64>(2048, rq.total / (400000ull * (u64)std::max(out.shards, 1))), 65536);
    // In a batch the warps of one block (8 for SPLIT, 16 for TRI) share a share: a handful of items per warp
    // balances them, and larger items spend less on the per-item set-up (unrank, lane vectors).
    const u64 batch_lanes = (u64)batch_shares(out.share_shift, rq.total) * 32u * (tri_gcap > 0 ? 16u : (u64)TP_SPLIT_WARPS);
    if (out.batch)
        cap0 = std::min<u64>(std::max<u64>(2048, rq.total / (batch_lanes * (u64)p->batch_items_per_warp)), 65536);
    // Kernel TRI pays more per item (it builds its lane registers per walked prefix, and an item that
    // covers only part of a prefix runs the generic pass), so its items are larger: the longest one
    // (the heaviest run first) about 45 % of what a warp has to do in the whole search.  Measured on 8 of 64:
    // 0.400 ms at 4 K, 0.352 at 11 K, 0.314 at 64 K, 0.449 at 128 K subsets per lane and item; slowest of
    // 8 shards of 8 of 80: 0.289 ms at 8 K, 0.255 at 32 K, 0.333 at 48 K.
    if (tri_gcap > 0 && !out.batch)
        cap0 = std::min<u64>(std::max<u64>(2048, rq.total / (110000ull * (u64)std::max(out.shards, 1))), 262144);
    if (p->tune_cap) cap0 = std::max<u64>(64, p->tune_cap); // tuning aid
    // Single values of b0 become items of their own, cut by b1 (D >= 2 only) -- only where it matters:
    // when one value alone is more than a quarter of what a lane has to do in this search (shard);
    // restricted b1 ranges lose the compile-time leaf blocks and add per-item set-up.
    pl.b_star = TP_MAX_N + 1;
    const bool pair_b0 = tri_gcap > 0 && q == 4; // kernel TRI pairs the values of its one walked position (tp_split_bstar)
    if (q > 3) {
        const u64 lanes = (u64)p->sm_count * 2 * TP_SPLIT_THREADS;
        u64 cap1 = std::max<u64>(cap0, rq.total / (lanes * (u64)std::max(out.shards, 1) * 4));
        // Kernel TRI on a shard: the pairs of single values (the items that end a shard, ~2 items per warp at 8
        // shards) are cut finer than the pieces -- slowest shard of C4, cap1 = cap0 / (2/3 cap0) / (1/2 cap0):
        // 4 shards 0.447 / 0.413 / 0.429 ms, 8 shards 0.282 / 0.280 / 0.260 ms (scripts/shard_sweep3.py).
        if (pair_b0 && !out.batch && out.shards >= 4)
            cap1 = std::max<u64>(out.shards >= 8 ? cap0 / 2 : cap0 * 2 / 3, rq.total / (lanes * (u64)out.shards * 4));
        if (out.batch) cap1 = std::max<u64>(cap0, rq.total / (batch_lanes * 4));
        if (p->tune_cap1) cap1 = std::max<u64>(64, p->tune_cap1); // tuning aid
        // (pairs: from a quarter of the cap on, so that the equal-work pieces below b_star are made of grains well
        // below their target -- a piece of two pairs where one was meant is the longest item of the search)
        int bs = q - 1;
        while (bs < n && (pair_b0 ? 8 * p->binom[bs][q - 1] : p->binom[bs][q - 1]) <= cap1) bs++;
        pl.b_star = bs;
        pl.cum[bs] = 0;
        if (pair_b0) {
            pl.cum[bs + 1] = 0;
            if (bs + 2 <= TP_MAX_N + 1) pl.cum[bs + 2] = 0;
            for (int b = bs + 1; b + 2 <= TP_MAX_N + 1; b++) { // head b of the pair (b, b - 1)
                const u64 work = b <= n ? p->binom[b][q - 1] + p->binom[b - 1][q - 1] : 0;
                const u64 nsub = std::min<u64>(std::max<u64>((work + cap1 - 1) / cap1, 1), TP_SPLIT_MAX_SUB);
                pl.cum[b + 2] = (uint16_t)(pl.cum[b] + nsub);
            }
        } else {
            for (int b = bs; b <= n && b <= TP_MAX_N; b++) {
                const u64 nsub = std::min<u64>(std::max<u64>((p->binom[b][q - 1] + cap1 - 1) / cap1, 1), TP_SPLIT_MAX_SUB);
                pl.cum[b + 1] = (uint16_t)(pl.cum[b] + nsub);
            }
        }
    }
    u64 items = 0;
    for (int e = 0; e < pl.n_entries; e++) {
        const int jt = pl.jt_hi - e;
        const u64 na = p->binom[n - 1 - jt][pp - 1];
        const u64 chunks = (na + 31) / 32;
        if (chunks == 0 || chunks > 0xFFFFFFFFull) return false;
        int np = 1; // slices of the first walked position, each about `cap0` subsets per lane
        const int jlow = std::min(jt, tp_split_bstar(jt, (int)pl.b_star, pair_b0));
        if (q > 2) np = (int)std::min<u64>(std::max<u64>((p->binom[jlow][q] + cap0 - 1) / cap0, 1), TP_SPLIT_MAX_PIECES);
        auto cum = [&](int b) -> int { return pl.cum[b]; };
        const u64 top = (u64)tp_split_top_slices(cum, jt, (int)pl.b_star, pair_b0); // parts of the single values (pairs)
        pl.item_base[e] = (uint32_t)items;
        pl.n_chunks[e] = (uint32_t)chunks;
        pl.n_pieces[e] = (uint8_t)np;
        items += chunks * ((u64)np + top);
        if (items > 0xFFFFFFF0ull) return false;
    }
    pl.total_items = (uint32_t)items;
    pl.tri_gcap = tri_gcap;
    out.kernel = tri_gcap > 0 ? TP_KERNEL_TRI : TP_KERNEL_SPLIT;
    return true;
}

// The nine ordered checks and two short-circuits of Allocate (besteffort_policy.go:90-124),
// then filterPartitions (device.go:310-351) as canonical order + group tables.
void prepare(const tp_policy *p, const int32_t *avail, int n_avail, const int32_t *req, int n_req, int size,
             Prep &out, int shards = 1, bool batch = false) {
    out.shards = shards;
    out.batch = batch;
    out.status = 0;
    out.shortcut = 0;
    out.n_req = n_req;
    out.family_only = false;
    if (size <= 0) { out.status = TP_ERR_INVALID_SIZE; return; }                   // :90-92
    if (n_avail < size) { out.status = TP_ERR_INVALID_AVAILABLE; return; }         // :94-96
    if (n_req > size) { out.status = TP_ERR_INVALID_REQUIRED; return; }            // :98-100
    if (n_req > n_avail) { out.status = TP_ERR_INVALID_REQ_AVAILABLE; return; }    // :102-104
    if (p->n == 0) { out.status = TP_ERR_INVALID_INIT; return; }                   // :106-108
    // (a NULL array with a positive count would be read by the short-circuits' copies: refuse it first)
    if ((n_avail > 0 && !avail) || (n_req > 0 && !req)) { out.status = TP_ERR_BAD_ARGUMENT; return; }
    if (n_avail == size) { out.shortcut = 1; return; }                             // :110-112
    if (n_req == size) { out.shortcut = 2; return; }                               // :114-116
    if (p->weights_empty) { out.status = TP_ERR_INVALID_INIT; return; }            // :118-120

    // setContainsAll (:122-124) is evaluated on ids before any lookup; an index outside
    // the policy is an id the reference's devicesMap does not know (it would nil-deref).
    uint8_t state[TP_MAX_N] = {0}; // 1 available, 2 required
    for (int i = 0; i < n_req; i++) {
        bool found = false;
        for (int j = 0; j < n_avail; j++)
            if (avail[j] == req[i]) { found = true; break; }
        if (!found) { out.status = TP_ERR_NO_CANDIDATE; return; }
    }
    for (int i = 0; i < n_avail; i++) {
        if (avail[i] < 0 || avail[i] >= p->n || state[avail[i]]) { out.status = TP_ERR_BAD_INDEX; return; }
        state[avail[i]] = 1;
    }
    for (int i = 0; i < n_req; i++) {
        if (state[req[i]] == 2) { out.status = TP_ERR_BAD_INDEX; return; }
        state[req[i]] = 2;
    }

    TpReq &rq = out.rq;
    memset(&rq, 0, sizeof rq);
    // free devices per group, NodeId ascending (device.go:322-333)
    std::vector<std::vector<int>> members(p->n_groups);
    for (int d = 0; d < p->n; d++)
        if (state[d] == 1) members[p->group[d]].push_back(d);
    std::vector<int> order;
    for (int g = 0; g < p->n_groups; g++) {
        if (members[g].empty()) continue;
        std::sort(members[g].begin(), members[g].end(),
                  [&](int a, int b) { return p->node_id[a] < p->node_id[b]; });
        order.push_back(g);
    }
    // (free count, ParentId) ascending (device.go:342-349); group index breaks the ties the
    // reference leaves to Go's map order.
    std::sort(order.begin(), order.end(), [&](int a, int b) {
        if (members[a].size() != members[b].size()) return members[a].size() < members[b].size();
        if (p->parent_rank[a] != p->parent_rank[b]) return p->parent_rank[a] < p->parent_rank[b];
        return a < b;
    });
    int n = 0;
    for (size_t pos = 0; pos < order.size(); pos++) {
        const std::vector<int> &m = members[order[pos]];
        rq.gstart[pos] = (uint8_t)n;
        rq.gsize[pos] = (uint8_t)m.size();
        for (int d : m) {
            rq.group_of[n] = (uint8_t)pos;
            rq.canon[n++] = (uint8_t)d;
        }
    }
    rq.n = n;
    rq.k = size - n_req;
    rq.n_groups = (int)order.size();
    rq.n_req = n_req;
    int dbits = 1;
    while ((1 << dbits) < rq.n_groups) dbits++;
    rq.dbits = dbits;
    int c_req = 0;
    for (int i = 0; i < n_req; i++) {
        rq.req[i] = (uint8_t)req[i];
        for (int j = 0; j < i; j++) c_req += p->weight[(size_t)req[j] * p->n + req[i]];
    }
    rq.c_req = c_req;
    if (rq.k > n) { out.status = TP_ERR_NO_CANDIDATE; return; } // unreachable: n_avail > size
    rq.total = rq.k <= TP_MAX_K ? p->binom[n][rq.k] : ~0ull; // saturating
    const u64 limit = p->exhaustive_limit ? p->exhaustive_limit : (1ull << 38);
    if (rq.k > TP_MAX_K || rq.total > limit || rq.k * dbits > TP_SIGMA_BITS) out.family_only = true;
    out.kernel = TP_KERNEL_LEX;
    // Family-only search: when the space cannot be enumerated (or when forced, for cross-checks).
    const bool family_fits = rq.n_groups <= TP_FAMILY_MAX_GROUPS && std::min(rq.n_groups, rq.k) * dbits <= TP_SIGMA_BITS;
    if ((out.family_only || p->kernel == TP_KERNEL_FAMILY) && family_fits) {
        out.kernel = TP_KERNEL_FAMILY;
        out.family_only = false;
        out.family_masks = 1ull << rq.n_groups;
        return;
    }
    if (p->kernel == TP_KERNEL_AUTO && rq.total <= TP_FAST_MAX_SUBSETS) return; // small: LEX (latency path), no plan
    if (!out.family_only && p->kernel != TP_KERNEL_LEX) build_plan(p, out); // sets out.kernel on success
}

// Advisory mode (tp_global_optimum): every free device becomes a group of its own, so Family holds
// every subset and RefKey orders equal scores by canonical position; the canonical order itself
// stays that of the real groups.  Re-decides the kernel for the new tables.
int ungroup(const tp_policy *p, Prep &pr) {
    TpReq &rq = pr.rq;
    for (int i = 0; i < rq.n; i++) {
        rq.group_of[i] = (uint8_t)i;
        rq.gstart[i] = (uint8_t)i;
        rq.gsize[i] = 1;
    }
    rq.n_groups = rq.n;
    int dbits = 1;
    while ((1 << dbits) < rq.n_groups) dbits++;
    rq.dbits = dbits;
    const u64 limit = p->exhaustive_limit ? p->exhaustive_limit : (1ull << 38);
    if (rq.k > TP_MAX_K || rq.total > limit || rq.k * dbits > TP_SIGMA_BITS) return TP_ERR_TOO_LARGE;
    pr.family_only = false;
    pr.family_masks = 0;
    pr.kernel = TP_KERNEL_LEX;
    if (p->kernel == TP_KERNEL_AUTO && rq.total <= TP_FAST_MAX_SUBSETS) return 0;
    if (p->kernel != TP_KERNEL_LEX) build_plan(p, pr);
    return 0;
}

// Winner key -> device indices in the reference's output order (besteffort_policy.go:141-148):
// groups in sequence order, NodeId ascending inside a group, required last in request order.
int decode(const Prep &pr, const TpKey &key, const int32_t *req, int32_t *out) {
    const TpReq &rq = pr.rq;
    if (key.hi == TP_KEY_INF_HI) return TP_ERR_NO_CANDIDATE;
    const int m = tp_key_groups(key);
    int remaining = rq.k, n_out = 0;
    for (int i = 0; i < m; i++) {
        const int g = (int)tp_key_get(key, i, rq.dbits);
        if (g >= rq.n_groups) return TP_ERR_BAD_ARGUMENT;
        const int take = std::min<int>(rq.gsize[g], remaining);
        for (int t = 0; t < take; t++) out[n_out++] = rq.canon[rq.gstart[g] + t];
        remaining -= take;
    }
    if (remaining != 0) return TP_ERR_BAD_ARGUMENT;
    for (int i = 0; i < pr.n_req; i++) out[n_out++] = req[i];
    return n_out;
}

void shard_range(u64 total, int idx, int cnt, u64 &lo, u64 &hi) {
    lo = (u64)(((unsigned __int128)total * (unsigned)idx) / (unsigned)cnt);
    hi = (u64)(((unsigned __int128)total * (unsigned)(idx + 1)) / (unsigned)cnt);
}

// Stage request (and plan) in the call's pinned buffer and start the copy to the device.
cudaError_t upload_request(Ctx &c, const Prep &pr) {
    *c.h_req = pr.rq;
    size_t bytes = sizeof(TpReq);
    if (uses_plan(pr.kernel)) {
        *c.h_plan = pr.plan;
        bytes += sizeof(TpPlan);
    }
    return cudaMemcpyAsync(c.d_req, c.h_req, bytes, cudaMemcpyHostToDevice, c.stream);
}

// Does shard (idx of cnt) hold any work?  LEX shards by rank range, SPLIT by rotating item stride.
bool shard_nonempty(const Prep &pr, int idx, int cnt) {
    if (uses_plan(pr.kernel)) return (u64)idx < pr.plan.total_items;
    u64 lo, hi;
    shard_range(pr.kernel == TP_KERNEL_FAMILY ? pr.family_masks : pr.rq.total, idx, cnt, lo, hi);
    return hi > lo;
}

// Subsets shard (idx of cnt) scores in one launch.
u64 shard_subsets(const tp_policy *p, const Prep &pr, int idx, int cnt) {
    if (pr.kernel == TP_KERNEL_FAMILY) return 0; // Family members are counted by the kernel only
    if (!uses_plan(pr.kernel)) {
        u64 lo, hi;
        shard_range(pr.rq.total, idx, cnt, lo, hi);
        return hi - lo;
    }
    if (cnt == 1) return pr.rq.total;
    const TpPlan &pl = pr.plan;
    const int n = pr.rq.n;
    u64 sum = 0;
    for (int e = 0; e < pl.n_entries; e++) {
        const int jt = pl.jt_hi - e;
        const u64 na = p->binom[n - 1 - jt][pl.p - 1];
        auto Cb = [&](int a, int b) -> u64 { return p->binom[a][b]; };
        auto cum = [&](int b) -> int { return pl.cum[b]; };
        const int np = pl.n_pieces[e];
        const bool pair = pr.kernel == TP_KERNEL_TRI;
        const int n_slices = np + tp_split_top_slices(cum, jt, (int)pl.b_star, pair && pl.q == 4);
        for (int slice = 0; slice < n_slices; slice++) {
            TpSlice sl;
            u64 per_lane = p->binom[jt][2]; // q == 2: every pair below the pivot
            if (pl.q > 2) {
                if (!tp_split_slice(Cb, cum, jt, pl.q, np, pl.b_star, slice, sl, pair)) continue;
                per_lane = tp_split_slice_count(Cb, sl, pl.q);
            }
            for (u64 ac = 0; ac < pl.n_chunks[e]; ac++) {
                const u64 item = pl.item_base[e] + (u64)slice * pl.n_chunks[e] + ac;
                if ((item + item / (unsigned)cnt) % (unsigned)cnt != (unsigned)idx) continue; // as tp_split_search
                sum += per_lane * std::min<u64>(32, na - ac * 32);
            }
        }
    }
    return sum;
}

// Enqueue the search kernel chosen by prepare() for shard (idx of cnt) on `st`, result into sc.out.
cudaError_t launch_search(tp_policy *p, Ctx &c, const Prep &pr, int idx, int cnt, const TpScratch &sc,
                          cudaStream_t st, TpLaunchInfo *li) {
    (void)cudaGetLastError(); // the launchers report cudaGetLastError(): an older, unrelated error must not be taken for theirs
    if (pr.kernel == TP_KERNEL_TRI)
        return tp_launch_exhaustive_tri(p->d_weight, p->n, p->d_binom, c.d_req, pr.rq, c.d_plan, pr.plan, idx, cnt, sc,
                                        c.d_counter, p->sm_count, st, li);
    if (pr.kernel == TP_KERNEL_SPLIT)
        return tp_launch_exhaustive_split(p->d_weight, p->n, p->d_binom, c.d_req, pr.rq, c.d_plan, pr.plan, idx, cnt,
                                          sc, c.d_counter, p->sm_count, st, li);
    u64 lo, hi;
    if (pr.kernel == TP_KERNEL_FAMILY) {
        shard_range(pr.family_masks, idx, cnt, lo, hi);
        return tp_launch_family(p->d_weight, p->n, p->d_binom, c.d_req, pr.rq, lo, hi, sc, p->sm_count, st, li);
    }
    shard_range(pr.rq.total, idx, cnt, lo, hi);
    return tp_launch_exhaustive_lex(p->d_weight, p->n, p->d_binom, c.d_req, pr.rq, lo, hi, sc, p->sm_count, st, li);
}

// Latency path for small searches (BASELINE configs C1/C2): one single-block launch with the
// descriptor as a kernel parameter; the result lands in mapped host memory and is picked up by
// polling its flag, so the call makes no memcpy, records no event and never blocks in the driver.
int run_search_fast(tp_policy *p, Ctx &c, const Prep &pr, TpKey *key, tp_result *info) {
    const unsigned epoch = ++c.epoch ? c.epoch : ++c.epoch; // never 0
    (void)cudaGetLastError();
    TP_CUDA(p, tp_launch_exhaustive_lex_small(p->d_weight, p->n, p->d_binom, pr.rq, c.d_fast, epoch, c.stream));
    unsigned spins = 0;
    while (c.h_fast->flag != epoch) {
        if ((++spins & 0xFFFF) == 0) { // a failed launch would never raise the flag
            cudaError_t e = cudaStreamQuery(c.stream);
            if (e != cudaSuccess && e != cudaErrorNotReady) TP_CUDA(p, e);
            if (e == cudaSuccess && c.h_fast->flag != epoch) TP_CUDA(p, cudaStreamSynchronize(c.stream));
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    key->hi = c.h_fast->key.hi;
    key->lo = c.h_fast->key.lo;
    if (info) {
        info->searched = 1;
        info->kernel = TP_KERNEL_LEX;
        info->subsets_scored = c.h_fast->scored;
        info->subsets_total = pr.rq.total;
        info->kernel_ms = 0.f; // not timed on this path (no events)
        info->gpu_launches = 1;
        info->key.hi = key->hi;
        info->key.lo = key->lo;
        if (key->hi != TP_KEY_INF_HI) {
            info->score = tp_key_score(*key);
            info->groups_touched = tp_key_groups(*key);
        }
    }
    return 0;
}

// Upload the prepared request, run the search over shard (idx of cnt) `iters` times, fetch the key.
int run_search(tp_policy *p, Ctx &c, const Prep &pr, int shard_idx, int shard_cnt, int iters, bool upload,
               TpKey *key, tp_result *info, float *total_ms) {
    if (pr.family_only) return TP_ERR_TOO_LARGE; // neither enumerable nor within kernel FAMILY's limits
    const TpReq &rq = pr.rq;
    const bool nonempty = shard_nonempty(pr, shard_idx, shard_cnt);
    if (upload) TP_CUDA(p, upload_request(c, pr));
    TpLaunchInfo li{};
    int launches = 0;
    TP_CUDA(p, cudaEventRecord(c.ev0, c.stream));
    for (int it = 0; it < iters; it++) {
        if (nonempty) {
            TP_CUDA(p, launch_search(p, c, pr, shard_idx, shard_cnt, c.scratch, c.stream, &li));
            launches += li.launches;
        }
    }
    TP_CUDA(p, cudaEventRecord(c.ev1, c.stream));
    if (nonempty) {
        TP_CUDA(p, cudaMemcpyAsync(c.h_out, c.scratch.out, sizeof(Ctx::Out), cudaMemcpyDeviceToHost, c.stream));
    } else {
        c.h_out->key = TpKey{TP_KEY_INF_HI, TP_KEY_INF_HI};
        c.h_out->scored = 0;
    }
    TP_CUDA(p, cudaStreamSynchronize(c.stream));
    float ms = 0.f;
    TP_CUDA(p, cudaEventElapsedTime(&ms, c.ev0, c.ev1));
    if (total_ms) *total_ms = ms;
    *key = c.h_out->key;
    if (info) {
        info->searched = 1;
        info->kernel = pr.kernel;
        info->subsets_scored = c.h_out->scored;
        info->subsets_total = rq.total;
        info->kernel_ms = ms / (iters > 0 ? iters : 1);
        info->gpu_launches = launches;
        info->key.hi = key->hi;
        info->key.lo = key->lo;
        if (key->hi != TP_KEY_INF_HI) {
            info->score = tp_key_score(*key);
            info->groups_touched = tp_key_groups(*key);
        }
    }
    return 0;
}

// One request on all of the policy's GPUs: shard s of N runs on GPU s (N = 1 + extra GPUs), each GPU driven by
// its own thread (the caller's for the first one), every shard's packed winner comes back to the host, the
// minimum wins.  The shards are the ones tp_search_shard hands out, so the result is that of one GPU.
int allocate_multi(tp_policy *p, const Prep &pr, const int32_t *req, int32_t *out, tp_result *info) {
    const int N = 1 + (int)p->peers.size(); // (pr: prepared with the work plan sized for N shards)
    std::vector<TpKey> keys(N, TpKey{TP_KEY_INF_HI, TP_KEY_INF_HI});
    std::vector<tp_result> infos(N);
    std::vector<int> rcs(N, 0);
    std::mutex mu;
    std::condition_variable cv;
    int pending = N - 1;
    std::atomic<int> pending_seen{N - 1}; // (the caller spins on this one before it sleeps on the condition)
    auto shard = [&](tp_policy *q, int s) {
        DeviceGuard guard(q->cuda_device);
        CtxLease lease(q);
        memset(&infos[s], 0, sizeof(tp_result));
        rcs[s] = run_search(q, lease.ctx(), pr, s, N, 1, true, &keys[s], &infos[s], nullptr);
    };
    for (int s = 1; s < N; s++)
        p->workers[s - 1]->submit([&, s] {
            shard(p->peers[s - 1], s);
            pending_seen.fetch_sub(1, std::memory_order_release);
            std::lock_guard<std::mutex> lk(mu);
            if (--pending == 0) cv.notify_one();
        });
    shard(p, 0);
    { // the other shards end within microseconds of this one: look before sleeping
        const auto give_up = std::chrono::steady_clock::now() + std::chrono::microseconds(500);
        while (pending_seen.load(std::memory_order_acquire) != 0 && std::chrono::steady_clock::now() < give_up) {
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
        }
        std::unique_lock<std::mutex> lk(mu); // (the workers still touch mu and cv after the counter: wait for that too)
        cv.wait(lk, [&] { return pending == 0; });
    }
    TpKey key{TP_KEY_INF_HI, TP_KEY_INF_HI};
    tp_result sum{};
    for (int s = 0; s < N; s++) {
        if (rcs[s]) {
            if (s > 0 && rcs[s] == TP_ERR_CUDA) { // the text lives in the peer's handle: bring it home
                std::lock_guard<std::mutex> lk(p->mu);
                p->last_cuda_error = "GPU " + std::to_string(p->peers[s - 1]->cuda_device) + ": " +
                                     p->peers[s - 1]->last_cuda_error;
            }
            return rcs[s];
        }
        if (tp_key_less(keys[s], key)) key = keys[s];
        sum.subsets_scored += infos[s].subsets_scored;
        sum.gpu_launches += infos[s].gpu_launches;
        sum.kernel_ms = std::max(sum.kernel_ms, infos[s].kernel_ms);
    }
    if (info) {
        *info = sum;
        info->searched = 1;
        info->kernel = pr.kernel;
        info->subsets_total = pr.rq.total;
        info->key.hi = key.hi;
        info->key.lo = key.lo;
        if (key.hi != TP_KEY_INF_HI) {
            info->score = tp_key_score(key);
            info->groups_touched = tp_key_groups(key);
        }
    }
    return decode(pr, key, req, out);
}

} // namespace

extern "C" {

const char *tp_strerror(int status) {
    switch (status) {
    case TP_OK: return "ok";
    // the reference's strings, besteffort_policy.go:36-43 and :73
    case TP_ERR_INVALID_SIZE: return "allocation size can not be negative";
    case TP_ERR_INVALID_AVAILABLE: return "available devices count less than allocation size";
    case TP_ERR_INVALID_REQUIRED: return "must_include devices size is more than allocation size";
    case TP_ERR_INVALID_REQ_AVAILABLE: return "must_include length should be less than or equal to avilable device size";
    case TP_ERR_INVALID_INIT: return "Init method must be called before Allocate";
    case TP_ERR_NO_CANDIDATE: return "No candidate subset found with matching criteria";
    case TP_ERR_INIT_FAILED: return "Besteffort Policy init failed to initialize p2pWeights";
    case TP_ERR_BAD_INDEX: return "unknown or duplicate device index";
    case TP_ERR_BAD_ARGUMENT: return "bad argument";
    case TP_ERR_TOO_LARGE: return "search space too large";
    case TP_ERR_STATIC_SCRATCH:
        return "request does not fit the static scratch (more than 2^38 subsets, or size minus required above 16)";
    case TP_ERR_CUDA: return "CUDA error";
    }
    return "unknown status";
}

const char *tp_last_cuda_error(const tp_policy *p) { return p ? p->last_cuda_error.c_str() : ""; }

int tp_create(const tp_topology *t, const tp_options *o, tp_policy **out) {
    if (!t || !out) return TP_ERR_BAD_ARGUMENT;
    *out = nullptr;
    if (t->n_devices < 0 || t->n_devices > TP_MAX_N || t->n_groups < 0 || t->n_groups > t->n_devices)
        return TP_ERR_BAD_ARGUMENT;
    const int n = t->n_devices;
    if (n > 0 && (!t->node_id || !t->group || !t->parent_rank || !t->weight)) return TP_ERR_BAD_ARGUMENT;
    tp_policy *p = new (std::nothrow) tp_policy();
    if (!p) return TP_ERR_BAD_ARGUMENT;
    p->n = n;
    p->n_groups = t->n_groups;
    p->node_id.assign(t->node_id, t->node_id + n);
    p->group.assign(t->group, t->group + n);
    p->parent_rank.assign(t->parent_rank, t->parent_rank + t->n_groups);
    p->weight.assign(t->weight, t->weight + (size_t)n * n);
    for (int i = 0; i < n; i++)
        if (p->group[i] < 0 || p->group[i] >= p->n_groups) { delete p; return TP_ERR_BAD_ARGUMENT; }
    p->weights_empty = true;
    for (uint8_t w : p->weight)
        if (w) { p->weights_empty = false; break; }
    // Init fails when no pair weight could be read (besteffort_policy.go:72-74)
    if (n > 0 && p->weights_empty) { delete p; return TP_ERR_INIT_FAILED; }
    for (int a = 0; a <= TP_MAX_N; a++)
        for (int j = 0; j < TP_BINOM_COLS; j++) {
            if (j == 0) p->binom[a][j] = 1;
            else if (a == 0) p->binom[a][j] = 0;
            else {
                u64 s = p->binom[a - 1][j - 1] + p->binom[a - 1][j];
                if (s < p->binom[a - 1][j]) s = ~0ull;
                p->binom[a][j] = s;
            }
        }
    p->kernel = o ? o->kernel : 0;
    p->exhaustive_limit = o ? o->exhaustive_limit : 0;
    if (const char *env = getenv("TP_SPLIT_Q")) p->tune_q = atoi(env);
    if (const char *env = getenv("TP_SPLIT_CAP")) p->tune_cap = strtoull(env, nullptr, 10);
    if (const char *env = getenv("TP_SPLIT_CAP1")) p->tune_cap1 = strtoull(env, nullptr, 10);
    p->batch_trace = getenv("TP_BATCH_TRACE") != nullptr;
    if (const char *env = getenv("TP_BATCH_SHARE_SHIFT")) p->batch_share_shift = std::max(20, std::min(40, atoi(env)));
    if (const char *env = getenv("TP_BATCH_ITEMS_PER_WARP")) p->batch_items_per_warp = std::max(1, atoi(env));
    int dev = o ? o->cuda_device : -1;
    if (dev == TP_DEVICE_NONE) { // host-only handle: validation, sharding arithmetic and tp_decode work,
        p->host_only = true;     // every entry point that would search returns TP_ERR_CUDA
        p->cuda_device = -1;
        *out = p;
        return TP_OK;
    }
    if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) { delete p; return TP_ERR_CUDA; }
    {
        int count = 0;
        if (cudaGetDeviceCount(&count) != cudaSuccess || dev >= count) {
            (void)cudaGetLastError(); // (leave no stale error behind for the caller's, or our own, next launch check)
            p->last_cuda_error = "no such CUDA device: " + std::to_string(dev);
            delete p;
            return TP_ERR_CUDA;
        }
    }
    p->cuda_device = dev;
    int rc = [&]() -> int {
        DeviceGuard guard(dev);
        int sms = 0;
        TP_CUDA(p, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        p->sm_count = sms > 0 ? sms : 148;
        TP_CUDA(p, cudaMalloc((void **)&p->d_weight, std::max<size_t>((size_t)n * n, 16)));
        TP_CUDA(p, cudaMemcpy(p->d_weight, p->weight.data(), (size_t)n * n, cudaMemcpyHostToDevice));
        TP_CUDA(p, cudaMalloc((void **)&p->d_binom, sizeof(p->binom)));
        TP_CUDA(p, cudaMemcpy(p->d_binom, p->binom, sizeof(p->binom), cudaMemcpyHostToDevice));
        const int inflight = (o && o->max_inflight > 0) ? o->max_inflight : 8;
        p->ctxs.resize(inflight);
        for (int i = 0; i < inflight; i++) {
            int r = init_ctx(p, p->ctxs[i]);
            if (r) return r;
            p->free_ctx.push_back(i);
        }
        return 0;
    }();
    if (rc) {
        tp_destroy(p);
        return rc;
    }
    // the other GPUs: a complete single-GPU policy each (same topology, same tuning), and its worker thread
    const int extra = o ? o->n_extra_devices : 0;
    if (extra < 0 || extra > TP_MAX_GPUS) { tp_destroy(p); return TP_ERR_BAD_ARGUMENT; }
    if (o && o->multi_gpu_min_subsets) p->multi_min = o->multi_gpu_min_subsets;
    for (int i = 0; i < extra; i++) {
        const int d = o->extra_devices[i];
        if (d < 0) { tp_destroy(p); return TP_ERR_BAD_ARGUMENT; }
        tp_options o2 = *o;
        o2.cuda_device = d;
        o2.n_extra_devices = 0;
        tp_policy *peer = nullptr;
        rc = tp_create(t, &o2, &peer);
        if (rc) { // e.g. an ordinal that does not exist: Init fails, the plugin runs without the policy
            tp_destroy(p);
            return rc;
        }
        p->peers.push_back(peer);
        p->workers.push_back(new GpuWorker());
    }
    *out = p;
    return TP_OK;
}

void tp_destroy(tp_policy *p) {
    if (!p) return;
    for (GpuWorker *w : p->workers) delete w; // (joins: no shard is in flight any more)
    for (tp_policy *q : p->peers) tp_destroy(q);
    if (p->host_only) {
        delete p;
        return;
    }
    {
        DeviceGuard guard(p->cuda_device);
        for (Ctx &c : p->ctxs) free_ctx(c);
        for (tp_policy::Batch &b : p->batches) batch_release(b);
        if (p->d_weight) cudaFree(p->d_weight);
        if (p->d_binom) cudaFree(p->d_binom);
    }
    delete p;
}

int tp_allocate(tp_policy *p, const int32_t *avail, int32_t n_avail, const int32_t *req, int32_t n_req,
                int32_t size, int32_t *out, tp_result *info) {
    if (!p || !out) return TP_ERR_BAD_ARGUMENT;
    if (info) memset(info, 0, sizeof *info);
    Prep pr;
    const int gpus = 1 + (int)p->peers.size();
    prepare(p, avail, n_avail, req, n_req, size, pr, gpus); // (a multi-GPU policy plans for all its GPUs first)
    if (pr.status) return pr.status;
    if (pr.shortcut == 1) { memcpy(out, avail, sizeof(int32_t) * size); return size; } // :110-112, as given
    if (pr.shortcut == 2) { memcpy(out, req, sizeof(int32_t) * size); return size; }   // :114-116
    if (p->host_only) return TP_ERR_CUDA; // there is no CPU search
    if (gpus > 1) {
        if (!pr.family_only && pr.kernel != TP_KERNEL_FAMILY && pr.rq.total >= p->multi_min)
            return allocate_multi(p, pr, req, out, info);
        prepare(p, avail, n_avail, req, n_req, size, pr, 1); // a small search: one GPU, items sized for it
    }
    DeviceGuard guard(p->cuda_device);
    CtxLease lease(p);
    TpKey key;
    const bool fast = pr.kernel != TP_KERNEL_FAMILY && !pr.family_only && pr.rq.total <= TP_FAST_MAX_SUBSETS &&
                      p->kernel == TP_KERNEL_AUTO;
    int rc = fast ? run_search_fast(p, lease.ctx(), pr, &key, info)
                  : run_search(p, lease.ctx(), pr, 0, 1, 1, true, &key, info, nullptr);
    if (rc) return rc;
    return decode(pr, key, req, out);
}

int tp_global_optimum(tp_policy *p, const int32_t *avail, int32_t n_avail, const int32_t *req, int32_t n_req,
                      int32_t size, int32_t *out, tp_result *info) {
    if (!p || !out) return TP_ERR_BAD_ARGUMENT;
    if (info) memset(info, 0, sizeof *info);
    Prep pr;
    prepare(p, avail, n_avail, req, n_req, size, pr);
    if (pr.status) return pr.status;
    if (pr.shortcut == 1) { memcpy(out, avail, sizeof(int32_t) * size); return size; }
    if (pr.shortcut == 2) { memcpy(out, req, sizeof(int32_t) * size); return size; }
    if (int rc = ungroup(p, pr)) return rc;
    if (p->host_only) return TP_ERR_CUDA;
    DeviceGuard guard(p->cuda_device);
    CtxLease lease(p);
    TpKey key;
    const bool fast = pr.rq.total <= TP_FAST_MAX_SUBSETS && p->kernel == TP_KERNEL_AUTO;
    int rc = fast ? run_search_fast(p, lease.ctx(), pr, &key, info)
                  : run_search(p, lease.ctx(), pr, 0, 1, 1, true, &key, info, nullptr);
    if (rc) return rc;
    return decode(pr, key, req, out);
}

namespace {

constexpr int kBatchMaxGroups = 64;

struct BatchLease { // one of the policy's batch slots, for the length of a tp_allocate_batch call
    tp_policy *p;
    int idx;
    explicit BatchLease(tp_policy *pol) : p(pol) {
        std::unique_lock<std::mutex> lk(p->batch_mu);
        p->batch_cv.wait(lk, [&] { return !p->free_batches.empty(); });
        idx = p->free_batches.back();
        p->free_batches.pop_back();
    }
    ~BatchLease() {
        {
            std::lock_guard<std::mutex> lk(p->batch_mu);
            p->free_batches.push_back(idx);
        }
        p->batch_cv.notify_one();
    }
    tp_policy::Batch &slot() { return p->batches[idx]; }
};

int batch_reserve(tp_policy *p, tp_policy::Batch &b, size_t n_reqs, size_t n_plans) {
    if (!b.stream) {
        TP_CUDA(p, cudaStreamCreateWithFlags(&b.stream, cudaStreamNonBlocking));
        for (cudaStream_t &st : b.side) TP_CUDA(p, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        for (cudaEvent_t &ev : b.ev_in) TP_CUDA(p, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        for (cudaEvent_t &ev : b.ev_done) TP_CUDA(p, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        TP_CUDA(p, cudaMalloc((void **)&b.d_grab, 2 * kBatchMaxGroups * sizeof(unsigned)));
    }
    if (n_reqs > b.cap_reqs) {
        const size_t cap = std::max<size_t>(n_reqs, 256) * 5 / 4;
        if (b.h_reqs) cudaFreeHost(b.h_reqs);
        if (b.h_order) cudaFreeHost(b.h_order);
        if (b.h_results) cudaFreeHost(b.h_results);
        if (b.d_reqs) cudaFree(b.d_reqs);
        if (b.d_order) cudaFree(b.d_order);
        if (b.d_results) cudaFree(b.d_results);
        b.h_reqs = nullptr; b.h_order = nullptr; b.h_results = nullptr;
        b.d_reqs = nullptr; b.d_order = nullptr; b.d_results = nullptr;
        b.cap_reqs = 0;
        TP_CUDA(p, cudaHostAlloc((void **)&b.h_reqs, cap * sizeof(TpReq), cudaHostAllocDefault));
        TP_CUDA(p, cudaHostAlloc((void **)&b.h_order, cap * sizeof(TpBatchEntry), cudaHostAllocDefault));
        TP_CUDA(p, cudaHostAlloc((void **)&b.h_results, cap * sizeof(TpBatchOut), cudaHostAllocDefault));
        TP_CUDA(p, cudaMalloc((void **)&b.d_reqs, cap * sizeof(TpReq)));
        TP_CUDA(p, cudaMalloc((void **)&b.d_order, cap * sizeof(TpBatchEntry)));
        TP_CUDA(p, cudaMalloc((void **)&b.d_results, cap * sizeof(TpBatchOut)));
        b.cap_reqs = cap;
    }
    if (n_plans > b.cap_plans) {
        const size_t cap = std::max<size_t>(n_plans, 64) * 5 / 4;
        if (b.h_plans) cudaFreeHost(b.h_plans);
        if (b.d_plans) cudaFree(b.d_plans);
        b.h_plans = nullptr; b.d_plans = nullptr;
        b.cap_plans = 0;
        TP_CUDA(p, cudaHostAlloc((void **)&b.h_plans, cap * sizeof(TpPlan), cudaHostAllocDefault));
        TP_CUDA(p, cudaMalloc((void **)&b.d_plans, cap * sizeof(TpPlan)));
        b.cap_plans = cap;
    }
    return 0;
}

struct BatchGroup {
    int kernel, a, b, c; // LEX: (k', -, -); SPLIT / TRI: (p, q, lane-vector class)
    int max_n = 0, max_k = 0, max_gcap = 0;
    u64 work = 0;             // subsets of its members
    std::vector<int> members; // request indices
};

} // namespace

// (a failed call may leave earlier launches of this batch in flight: wait for them before the slot is handed back)
#define TP_BATCH_CUDA(call)                                                                   \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            {                                                                                 \
                std::lock_guard<std::mutex> lk__(p->mu);                                      \
                p->last_cuda_error = std::string(#call) + ": " + cudaGetErrorString(e__);     \
            }                                                                                 \
            for (cudaStream_t st__ : b.side) cudaStreamSynchronize(st__);                     \
            cudaStreamSynchronize(b.stream);                                                  \
            (void)cudaGetLastError();                                                         \
            return TP_ERR_CUDA;                                                               \
        }                                                                                     \
    } while (0)

// All requests of one RPC in a handful of launches.  Requests are prepared on the host (in parallel),
// grouped by the kernel instantiation they need, and each group runs as ONE persistent launch in which
// a block owns a whole request, or one share of a heavy one (heaviest first).  The batch goes in two
// WAVES: the requests that search longest (k' >= 6) are prepared, staged and launched first, and the
// host prepares the rest while the GPU works on them; groups run on a few streams, so that the light
// groups fill the SMs the heavy group's last blocks leave.  Requests the reference answers without
// searching never reach the GPU; the rare ones outside the batch kernels (Family-only) go through
// tp_allocate.  Up to kBatchSlots calls run concurrently on one policy, each on buffers of its own.
int tp_allocate_batch(tp_policy *p, int32_t n_requests, const tp_request *rqs, int32_t *out, int32_t out_stride,
                      int32_t *status, tp_result *infos) {
    if (!p || n_requests < 0 || (n_requests && (!rqs || !out || !status))) return TP_ERR_BAD_ARGUMENT;
    if (n_requests < 4 || p->host_only) { // not worth a batch: answer one by one
        for (int i = 0; i < n_requests; i++) {
            if (rqs[i].size > out_stride) { status[i] = TP_ERR_BAD_ARGUMENT; continue; }
            status[i] = tp_allocate(p, rqs[i].available, rqs[i].n_available, rqs[i].required, rqs[i].n_required,
                                    rqs[i].size, out + (size_t)i * out_stride, infos ? infos + i : nullptr);
        }
        return TP_OK;
    }
    const bool trace = p->batch_trace; // phase timings on stderr (tuning aid)
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point t0) {
        return std::chrono::duration<double, std::milli>(now() - t0).count();
    };
    const auto t_start = now();
    // ---- which wave, and upper bounds on what the batch needs (before anything is prepared) ----
    std::vector<int> wave[2];
    size_t out_bound = 0, plan_bound = 0;
    int share_shift = p->batch_share_shift;
    if (share_shift == 0) {
        int n_heavy = 0;
        for (int i = 0; i < n_requests; i++) n_heavy += rqs[i].size - rqs[i].n_required >= 6;
        share_shift = n_heavy >= 4 * p->sm_count ? 28 : n_heavy >= p->sm_count ? 26 : 24;
    }
    for (int i = 0; i < n_requests; i++) {
        const tp_request &r = rqs[i];
        const int k1 = r.size - r.n_required, n1 = r.n_available - r.n_required;
        const bool heavy = k1 >= 6 && n1 > k1;
        wave[(heavy ? 0 : 1) ^ (p->batch_light_first ? 1 : 0)].push_back(i);
        if (k1 >= 4 && n1 > k1) {
            plan_bound++;
            out_bound += (n1 <= TP_MAX_N && k1 < TP_BINOM_COLS) ? (size_t)batch_shares(share_shift, p->binom[n1][k1]) : 16;
        } else {
            out_bound++;
        }
    }
    if (wave[0].empty()) wave[0].swap(wave[1]);
    std::vector<Prep> preps(n_requests);
    std::vector<int> singles;       // served by tp_allocate after the batch
    std::vector<int> searched;      // requests in a batch group, both waves
    std::vector<int> out_base(n_requests, -1), n_shares(n_requests, 1);
    std::vector<BatchGroup> groups; // of the wave at hand
    int launches = 0;
    size_t n_slots = 0, n_plans = 0, n_order = 0, n_out = 0, n_grab = 0;
    unsigned side_used = 0, side_next = 0;
    double t_prepare[2] = {0, 0}, t_stage[2] = {0, 0};
    BatchLease lease(p);
    tp_policy::Batch &b = lease.slot();
    DeviceGuard guard(p->cuda_device);
    for (int w = 0; w < 2; w++) {
        const std::vector<int> &idx = wave[w];
        if (idx.empty()) continue;
        const auto t_wave = now();
        const int n_w = (int)idx.size();
        {
            const unsigned hw = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
            const int nthreads = (int)std::min<unsigned>(hw, (unsigned)(n_w / 64 + 1));
            run_threads(nthreads, [&](int t) {
                for (int j = t; j < n_w; j += nthreads) {
                    const tp_request &r = rqs[idx[j]];
                    preps[idx[j]].share_shift = share_shift;
                    prepare(p, r.available, r.n_available, r.required, r.n_required, r.size, preps[idx[j]], 1, true);
                }
            });
        }
        t_prepare[w] = ms_since(t_wave);
        groups.clear();
        for (int i : idx) {
            const Prep &pr = preps[i];
            if (infos) memset(&infos[i], 0, sizeof(tp_result));
            if (rqs[i].size > out_stride && !pr.status) { status[i] = TP_ERR_BAD_ARGUMENT; continue; }
            if (pr.status) { status[i] = pr.status; continue; }
            if (pr.shortcut) {
                memcpy(out + (size_t)i * out_stride, pr.shortcut == 1 ? rqs[i].available : rqs[i].required,
                       sizeof(int32_t) * rqs[i].size);
                status[i] = rqs[i].size;
                continue;
            }
            if (pr.family_only || pr.kernel == TP_KERNEL_FAMILY) { singles.push_back(i); continue; }
            // a LEX request gets ONE block in a batch group; a large one (scores beyond the packed kernels'
            // 16-bit range can hold up to 2^38 subsets) would keep that block, and the whole RPC, busy for minutes
            if (pr.kernel == TP_KERNEL_LEX && pr.rq.total > (1ull << 24)) { singles.push_back(i); continue; }
            BatchGroup key;
            key.kernel = pr.kernel;
            if (uses_plan(pr.kernel)) {
                const int need = pr.rq.n - pr.plan.p;
                key.a = pr.plan.p; key.b = pr.plan.q; key.c = need <= 64 ? 64 : need <= 80 ? 80 : need <= 96 ? 96 : 128;
            } else {
                key.a = pr.rq.k; key.b = 0; key.c = 0;
            }
            size_t g = 0;
            for (; g < groups.size(); g++)
                if (groups[g].kernel == key.kernel && groups[g].a == key.a && groups[g].b == key.b && groups[g].c == key.c) break;
            if (g == groups.size()) {
                if ((int)groups.size() == kBatchMaxGroups) { singles.push_back(i); continue; }
                groups.push_back(key);
            }
            groups[g].members.push_back(i);
            groups[g].work += pr.rq.total;
            groups[g].max_n = std::max(groups[g].max_n, pr.rq.n);
            groups[g].max_k = std::max(groups[g].max_k, pr.rq.k);
            if (pr.kernel == TP_KERNEL_TRI) groups[g].max_gcap = std::max(groups[g].max_gcap, (int)pr.plan.tri_gcap);
        }
        if (groups.empty()) continue;
        if (!b.stream || out_bound > b.cap_reqs || (size_t)n_requests > b.cap_reqs || plan_bound > b.cap_plans) {
            // (only ever before the first launch of this call: the bounds cover both waves)
            int rc = batch_reserve(p, b, std::max<size_t>((size_t)n_requests, out_bound), plan_bound);
            if (rc) return rc;
        }
        std::sort(groups.begin(), groups.end(), [](const BatchGroup &x, const BatchGroup &y) { return x.work > y.work; });
        // Heavy requests are cut into shares taken by different blocks (the same item striding as
        // multi-GPU shards), so that one big search does not become the tail of its group.
        const size_t slot0 = n_slots, plan0 = n_plans, order0 = n_order;
        std::vector<size_t> first(groups.size()), n_entries(groups.size());
        std::vector<std::pair<int, int>> placed; // (request, slot) of this wave
        for (size_t g = 0; g < groups.size(); g++) {
            BatchGroup &grp = groups[g];
            std::sort(grp.members.begin(), grp.members.end(),
                      [&](int x, int y) { return preps[x].rq.total > preps[y].rq.total; }); // heaviest first
            first[g] = n_order;
            for (int i : grp.members) {
                const int S = uses_plan(grp.kernel) ? batch_shares(share_shift, preps[i].rq.total) : 1;
                if (n_out + (size_t)S > b.cap_reqs || n_order + (size_t)S > b.cap_reqs ||
                    (uses_plan(grp.kernel) && n_plans >= b.cap_plans)) { // (the bounds above make this unreachable)
                    singles.push_back(i);
                    continue;
                }
                const int slot = (int)n_slots++;
                int plan_idx = 0;
                if (uses_plan(grp.kernel)) {
                    plan_idx = (int)n_plans++;
                    b.h_plans[plan_idx] = preps[i].plan;
                }
                placed.emplace_back(i, slot);
                out_base[i] = (int)n_out;
                n_shares[i] = S;
                for (int sidx = 0; sidx < S; sidx++)
                    b.h_order[n_order++] = TpBatchEntry{slot, plan_idx, (int)n_out + sidx, (int16_t)sidx, (int16_t)S};
                n_out += (size_t)S;
                searched.push_back(i);
            }
            n_entries[g] = n_order - first[g];
        }
        { // descriptors into the pinned array (~2 KB per request): worth a few threads
            const unsigned hw = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
            const int n_p = (int)placed.size();
            const int nthreads = (int)std::min<unsigned>(hw, (unsigned)(n_p / 512 + 1));
            run_threads(nthreads, [&](int t) {
                for (int j = t; j < n_p; j += nthreads) b.h_reqs[placed[j].second] = preps[placed[j].first].rq;
            });
        }
        t_stage[w] = ms_since(t_wave) - t_prepare[w];
        TP_BATCH_CUDA(cudaMemcpyAsync(b.d_reqs + slot0, b.h_reqs + slot0, (n_slots - slot0) * sizeof(TpReq),
                                      cudaMemcpyHostToDevice, b.stream));
        if (n_plans > plan0)
            TP_BATCH_CUDA(cudaMemcpyAsync(b.d_plans + plan0, b.h_plans + plan0, (n_plans - plan0) * sizeof(TpPlan),
                                          cudaMemcpyHostToDevice, b.stream));
        TP_BATCH_CUDA(cudaMemcpyAsync(b.d_order + order0, b.h_order + order0, (n_order - order0) * sizeof(TpBatchEntry),
                                      cudaMemcpyHostToDevice, b.stream));
        TP_BATCH_CUDA(cudaEventRecord(b.ev_in[w], b.stream));
        (void)cudaGetLastError();
        for (size_t g = 0; g < groups.size(); g++) {
            const BatchGroup &grp = groups[g];
            if (n_entries[g] == 0) continue;
            const unsigned sidx = side_next++ % tp_policy::kBatchSide;
            cudaStream_t st = b.side[sidx];
            TP_BATCH_CUDA(cudaStreamWaitEvent(st, b.ev_in[w], 0));
            side_used |= 1u << sidx;
            TpBatch tb{b.d_order + first[g], b.d_results, b.d_grab + n_grab++, (int32_t)n_entries[g], 0};
            TpLaunchInfo li{};
            if (grp.kernel == TP_KERNEL_TRI)
                TP_BATCH_CUDA(tp_launch_batch_tri(p->d_weight, p->n, p->d_binom, b.d_reqs, b.d_plans, grp.a, grp.b,
                                                  grp.max_n, grp.max_k, grp.max_gcap, tb, p->sm_count, st, &li));
            else if (grp.kernel == TP_KERNEL_SPLIT)
                TP_BATCH_CUDA(tp_launch_batch_split(p->d_weight, p->n, p->d_binom, b.d_reqs, b.d_plans, grp.a, grp.b,
                                                    grp.max_n, grp.max_k, tb, p->sm_count, st, &li));
            else
                TP_BATCH_CUDA(tp_launch_batch_lex(p->d_weight, p->n, p->d_binom, b.d_reqs, grp.a, grp.max_n, tb,
                                                  p->sm_count, st, &li));
            launches += li.launches;
        }
    }
    const double t_host = ms_since(t_start);
    if (!searched.empty()) {
        for (int sidx = 0; sidx < tp_policy::kBatchSide; sidx++)
            if (side_used >> sidx & 1u) {
                TP_BATCH_CUDA(cudaEventRecord(b.ev_done[sidx], b.side[sidx]));
                TP_BATCH_CUDA(cudaStreamWaitEvent(b.stream, b.ev_done[sidx], 0));
            }
        TP_BATCH_CUDA(cudaMemcpyAsync(b.h_results, b.d_results, n_out * sizeof(TpBatchOut), cudaMemcpyDeviceToHost,
                                      b.stream));
        TP_BATCH_CUDA(cudaStreamSynchronize(b.stream));
    }
    if (trace)
        fprintf(stderr,
                "tp_allocate_batch: %d requests, %d launches: wave 0 (%zu) prepare %.2f stage %.2f ms, wave 1 (%zu) "
                "prepare %.2f stage %.2f ms, host done at %.2f ms, GPU done at %.2f ms\n",
                n_requests, launches, wave[0].size(), t_prepare[0], t_stage[0], wave[1].size(), t_prepare[1],
                t_stage[1], t_host, ms_since(t_start));
    for (int i : searched) {
        TpKey key = b.h_results[out_base[i]].key; // min over the request's shares
        u64 scored = b.h_results[out_base[i]].scored;
        for (int sidx = 1; sidx < n_shares[i]; sidx++) {
            const TpBatchOut &part = b.h_results[out_base[i] + sidx];
            if (tp_key_less(part.key, key)) key = part.key;
            scored += part.scored;
        }
        status[i] = decode(preps[i], key, rqs[i].required, out + (size_t)i * out_stride);
        if (infos) {
            tp_result &r = infos[i];
            r.searched = 1;
            r.kernel = preps[i].kernel;
            r.subsets_scored = scored;
            r.subsets_total = preps[i].rq.total;
            r.gpu_launches = launches; // launches of the whole batch (shared by its requests)
            r.key.hi = key.hi;
            r.key.lo = key.lo;
            if (key.hi != TP_KEY_INF_HI) {
                r.score = tp_key_score(key);
                r.groups_touched = tp_key_groups(key);
            }
        }
    }
    for (int i : singles)
        status[i] = tp_allocate(p, rqs[i].available, rqs[i].n_available, rqs[i].required, rqs[i].n_required,
                                rqs[i].size, out + (size_t)i * out_stride, infos ? infos + i : nullptr);
    return TP_OK;
}
#undef TP_BATCH_CUDA

int tp_search_resident(tp_policy *p, const int32_t *avail, int32_t n_avail, const int32_t *req, int32_t n_req,
                       int32_t size, int32_t shard_index, int32_t shard_count, int32_t iters, float *ms,
                       tp_key *key, tp_result *info) {
    if (!p || !key || shard_count <= 0 || shard_index < 0 || shard_index >= shard_count || iters <= 0)
        return TP_ERR_BAD_ARGUMENT;
    if (info) memset(info, 0, sizeof *info);
    Prep pr;
    prepare(p, avail, n_avail, req, n_req, size, pr, shard_count);
    if (pr.status) return pr.status;
    if (pr.shortcut) return 1;
    if (p->host_only) return TP_ERR_CUDA; // there is no CPU search
    DeviceGuard guard(p->cuda_device);
    CtxLease lease(p);
    Ctx &c = lease.ctx();
    // upload first, outside the timed region
    TP_CUDA(p, upload_request(c, pr));
    TP_CUDA(p, cudaStreamSynchronize(c.stream));
    TpKey k;
    int rc = run_search(p, c, pr, shard_index, shard_count, iters, false, &k, info, ms);
    if (rc) return rc;
    key->hi = k.hi;
    key->lo = k.lo;
    return 0;
}

int tp_search_shard(tp_policy *p, const int32_t *avail, int32_t n_avail, const int32_t *req, int32_t n_req,
                    int32_t size, int32_t shard_index, int32_t shard_count, tp_key *key, tp_result *info) {
    if (!p || !key || shard_count <= 0 || shard_index < 0 || shard_index >= shard_count) return TP_ERR_BAD_ARGUMENT;
    if (info) memset(info, 0, sizeof *info);
    Prep pr;
    prepare(p, avail, n_avail, req, n_req, size, pr, shard_count);
    if (pr.status) return pr.status;
    if (pr.shortcut) return 1;
    if (p->host_only) return TP_ERR_CUDA; // there is no CPU search
    DeviceGuard guard(p->cuda_device);
    CtxLease lease(p);
    TpKey k;
    int rc = run_search(p, lease.ctx(), pr, shard_index, shard_count, 1, true, &k, info, nullptr);
    if (rc) return rc;
    key->hi = k.hi;
    key->lo = k.lo;
    return 0;
}

int tp_shard_subsets(tp_policy *p, const int32_t *avail, int32_t n_avail, const int32_t *req, int32_t n_req,
                     int32_t size, int32_t shard_index, int32_t shard_count, uint64_t *subsets, int32_t *kernel) {
    if (!p || !subsets || shard_count <= 0 || shard_index < 0 || shard_index >= shard_count) return TP_ERR_BAD_ARGUMENT;
    Prep pr;
    prepare(p, avail, n_avail, req, n_req, size, pr, shard_count);
    if (pr.status) return pr.status;
    if (pr.shortcut) return 1;
    if (pr.family_only) return TP_ERR_TOO_LARGE;
    *subsets = shard_subsets(p, pr, shard_index, shard_count);
    if (kernel) *kernel = pr.kernel;
    return TP_OK;
}

int tp_decode(tp_policy *p, const int32_t *avail, int32_t n_avail, const int32_t *req, int32_t n_req, int32_t size,
              const tp_key *key, int32_t *out) {
    if (!p || !key || !out) return TP_ERR_BAD_ARGUMENT;
    Prep pr;
    prepare(p, avail, n_avail, req, n_req, size, pr);
    if (pr.status) return pr.status;
    if (pr.shortcut == 1) { memcpy(out, avail, sizeof(int32_t) * size); return size; }
    if (pr.shortcut == 2) { memcpy(out, req, sizeof(int32_t) * size); return size; }
    TpKey k{key->hi, key->lo};
    return decode(pr, k, req, out);
}

struct tp_prepared {
    int ctx_idx;
    Prep pr;
    int shard_index, shard_count;
    u64 subsets; // subsets this shard scores per launch
};

int tp_prepare(tp_policy *p, const int32_t *avail, int32_t n_avail, const int32_t *req, int32_t n_req, int32_t size,
               int32_t shard_index, int32_t shard_count, tp_prepared **out) {
    if (!p || !out || shard_count <= 0 || shard_index < 0 || shard_index >= shard_count) return TP_ERR_BAD_ARGUMENT;
    *out = nullptr;
    tp_prepared *q = new (std::nothrow) tp_prepared();
    if (!q) return TP_ERR_BAD_ARGUMENT;
    prepare(p, avail, n_avail, req, n_req, size, q->pr, shard_count);
    int rc = q->pr.status ? q->pr.status : q->pr.shortcut ? 1 : q->pr.family_only ? TP_ERR_TOO_LARGE
             : p->host_only ? TP_ERR_CUDA : 0;
    if (rc) {
        delete q;
        return rc;
    }
    q->shard_index = shard_index;
    q->shard_count = shard_count;
    q->subsets = shard_subsets(p, q->pr, shard_index, shard_count);
    {
        std::unique_lock<std::mutex> lk(p->mu);
        p->cv.wait(lk, [&] { return !p->free_ctx.empty(); });
        q->ctx_idx = p->free_ctx.back();
        p->free_ctx.pop_back();
    }
    Ctx &c = p->ctxs[q->ctx_idx];
    DeviceGuard guard(p->cuda_device);
    cudaError_t e = upload_request(c, q->pr);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    if (e != cudaSuccess) {
        tp_prepared_free(p, q);
        return TP_ERR_CUDA;
    }
    *out = q;
    return TP_OK;
}

int tp_launch(tp_policy *p, tp_prepared *q, void *cuda_stream, uint64_t *d_out) {
    if (!p || !q || !d_out) return TP_ERR_BAD_ARGUMENT;
    Ctx &c = p->ctxs[q->ctx_idx];
    DeviceGuard guard(p->cuda_device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    TpScratch sc = c.scratch;
    sc.out = (TpKey *)d_out;
    sc.out_scored = (u64 *)(d_out + 2);
    if (!shard_nonempty(q->pr, q->shard_index, q->shard_count)) { // empty shard: no candidate, nothing scored
        static const u64 none[3] = {TP_KEY_INF_HI, TP_KEY_INF_HI, 0};
        TP_CUDA(p, cudaMemcpyAsync(d_out, none, sizeof none, cudaMemcpyHostToDevice, st));
        return 0;
    }
    TpLaunchInfo li{};
    TP_CUDA(p, launch_search(p, c, q->pr, q->shard_index, q->shard_count, sc, st, &li));
    return li.launches;
}

struct tp_exchange {
    int device = 0, rank = 0, world = 0;
    unsigned long long *local = nullptr;          // [2][world][TP_EX_SLOT_WORDS]
    unsigned long long *peer[TP_MAX_GPUS] = {};   // local for `rank`, IPC mappings for the others
    bool connected = false;
    unsigned long long epoch = 0;
};

int tp_exchange_create(tp_policy *p, int32_t rank, int32_t world, tp_exchange **out, void *handle_out) {
    if (!p || !out || !handle_out || world < 1 || world > TP_MAX_GPUS || rank < 0 || rank >= world) return TP_ERR_BAD_ARGUMENT;
    if (p->host_only) return TP_ERR_CUDA;
    static_assert(sizeof(cudaIpcMemHandle_t) == TP_IPC_HANDLE_BYTES, "IPC handle size");
    tp_exchange *x = new (std::nothrow) tp_exchange();
    if (!x) return TP_ERR_BAD_ARGUMENT;
    x->device = p->cuda_device;
    x->rank = rank;
    x->world = world;
    DeviceGuard guard(x->device);
    const size_t bytes = (size_t)2 * world * TP_EX_SLOT_WORDS * sizeof(unsigned long long);
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaMalloc((void **)&x->local, bytes);
    if (e == cudaSuccess) e = cudaMemset(x->local, 0, bytes);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, x->local);
    if (e != cudaSuccess) {
        if (x->local) cudaFree(x->local);
        delete x;
        std::lock_guard<std::mutex> lk(p->mu);
        p->last_cuda_error = std::string("tp_exchange_create: ") + cudaGetErrorString(e);
        return TP_ERR_CUDA;
    }
    memcpy(handle_out, &h, sizeof h);
    x->peer[rank] = x->local;
    *out = x;
    return TP_OK;
}

int tp_exchange_connect(tp_exchange *x, const void *handles) {
    if (!x || !handles) return TP_ERR_BAD_ARGUMENT;
    DeviceGuard guard(x->device);
    for (int r = 0; r < x->world; r++) {
        if (r == x->rank || x->peer[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const unsigned char *)handles + (size_t)r * TP_IPC_HANDLE_BYTES, sizeof h);
        void *ptr = nullptr;
        if (cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) return TP_ERR_CUDA;
        x->peer[r] = (unsigned long long *)ptr;
    }
    x->connected = true;
    return TP_OK;
}

void tp_exchange_destroy(tp_exchange *x) {
    if (!x) return;
    DeviceGuard guard(x->device);
    for (int r = 0; r < x->world; r++)
        if (r != x->rank && x->peer[r]) cudaIpcCloseMemHandle(x->peer[r]);
    if (x->local) cudaFree(x->local);
    delete x;
}

int tp_launch_exchange(tp_policy *p, tp_prepared *q, tp_exchange *x, void *cuda_stream, uint64_t *d_out) {
    if (!p || !q || !x || !d_out || !x->connected || x->world != q->shard_count || x->rank != q->shard_index)
        return TP_ERR_BAD_ARGUMENT;
    Ctx &c = p->ctxs[q->ctx_idx];
    DeviceGuard guard(p->cuda_device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    TpScratch sc = c.scratch;
    sc.out = (TpKey *)d_out;
    sc.out_scored = (u64 *)(d_out + 2);
    for (int r = 0; r < x->world; r++) sc.ex.peer[r] = x->peer[r];
    sc.ex.rank = x->rank;
    sc.ex.world = x->world;
    sc.ex.epoch = ++x->epoch;
    TpLaunchInfo li{};
    if (!shard_nonempty(q->pr, q->shard_index, q->shard_count)) {
        // a shard without work still has to show up in the exchange: one block of the LEX kernel over an empty range
        TP_CUDA(p, tp_launch_exchange_only(sc, st));
        return 1;
    }
    TP_CUDA(p, launch_search(p, c, q->pr, q->shard_index, q->shard_count, sc, st, &li));
    return li.launches;
}

uint64_t tp_prepared_subsets(const tp_prepared *q) { return q ? q->subsets : 0; }

void tp_prepared_free(tp_policy *p, tp_prepared *q) {
    if (!p || !q) return;
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->free_ctx.push_back(q->ctx_idx);
    }
    p->cv.notify_one();
    delete q;
}

// tp_allocate for callers that may not allocate (see the long comment in tp_kernels.cu and the
// declaration in tp_allocator.h).  Host-side preparation is the same code tp_allocate uses,
// run on a throw-away, host-only policy object built from the caller's topology arrays.
int tp_allocate_static(const tp_topology *t, const uint8_t *d_weight, const int32_t *avail, int32_t n_avail,
                       const int32_t *req, int32_t n_req, int32_t size, void *cuda_stream, uint64_t *d_out,
                       int32_t *out, tp_result *info) {
    if (!t || !out) return TP_ERR_BAD_ARGUMENT;
    if (info) memset(info, 0, sizeof *info);
    const int n = t->n_devices;
    if (n < 0 || n > TP_MAX_N || t->n_groups < 0 || t->n_groups > n) return TP_ERR_BAD_ARGUMENT;
    if (n > 0 && (!t->node_id || !t->group || !t->parent_rank || !t->weight)) return TP_ERR_BAD_ARGUMENT;
    tp_policy host; // host fields only: no stream, no device memory, never handed to tp_destroy
    host.n = n;
    host.n_groups = t->n_groups;
    host.node_id.assign(t->node_id, t->node_id + n);
    host.group.assign(t->group, t->group + n);
    host.parent_rank.assign(t->parent_rank, t->parent_rank + t->n_groups);
    host.weight.assign(t->weight, t->weight + (size_t)n * n);
    for (int i = 0; i < n; i++)
        if (host.group[i] < 0 || host.group[i] >= host.n_groups) return TP_ERR_BAD_ARGUMENT;
    host.weights_empty = true;
    for (uint8_t w : host.weight)
        if (w) { host.weights_empty = false; break; }
    for (int a = 0; a <= TP_MAX_N; a++)
        for (int j = 0; j < TP_BINOM_COLS; j++) {
            if (j == 0) host.binom[a][j] = 1;
            else if (a == 0) host.binom[a][j] = 0;
            else {
                u64 s = host.binom[a - 1][j - 1] + host.binom[a - 1][j];
                if (s < host.binom[a - 1][j]) s = ~0ull;
                host.binom[a][j] = s;
            }
        }
    host.kernel = TP_KERNEL_LEX; // the static scratch is laid out for the LEX kernel only
    Prep pr;
    prepare(&host, avail, n_avail, req, n_req, size, pr);
    if (pr.status) return pr.status;
    if (pr.shortcut == 1) { memcpy(out, avail, sizeof(int32_t) * size); return size; }
    if (pr.shortcut == 2) { memcpy(out, req, sizeof(int32_t) * size); return size; }
    if (!d_weight || !d_out) return TP_ERR_BAD_ARGUMENT;
    // The static array has room for a fixed number of block partials and this entry point only
    // drives the exhaustive kernel: anything larger is refused, not silently truncated.
    if (pr.family_only || pr.rq.total > TP_STATIC_MAX_SUBSETS) return TP_ERR_STATIC_SCRATCH;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    TpLaunchInfo li{};
    (void)cudaGetLastError();
    cudaError_t e = tp_launch_exhaustive_lex_static(d_weight, n, &host.binom[0][0], pr.rq, 0, pr.rq.total,
                                                    (unsigned long long *)d_out, st, &li);
    u64 h_out[3] = {TP_KEY_INF_HI, TP_KEY_INF_HI, 0};
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_out, d_out, sizeof h_out, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return TP_ERR_CUDA;
    TpKey key{h_out[0], h_out[1]};
    if (info) {
        info->searched = 1;
        info->kernel = TP_KERNEL_LEX;
        info->subsets_scored = h_out[2];
        info->subsets_total = pr.rq.total;
        info->gpu_launches = li.launches;
        info->key.hi = key.hi;
        info->key.lo = key.lo;
        if (key.hi != TP_KEY_INF_HI) {
            info->score = tp_key_score(key);
            info->groups_touched = tp_key_groups(key);
        }
    }
    return decode(pr, key, req, out);
}

int tp_descriptor_bytes(int32_t with_plan) { return (int)(sizeof(TpReq) + (with_plan ? sizeof(TpPlan) : 0)); }

void tp_set_batch_tuning(tp_policy *p, int32_t share_shift, int32_t items_per_warp, int32_t light_first) {
    if (!p) return;
    p->batch_share_shift = share_shift > 0 ? std::max(20, std::min(40, (int)share_shift)) : 0;
    p->batch_items_per_warp = items_per_warp > 0 ? items_per_warp : 4;
    p->batch_light_first = light_first != 0;
}

void tp_set_tuning(tp_policy *p, int32_t walk_q, uint64_t item_cap, uint64_t item_cap1) {
    if (!p) return;
    p->tune_q = walk_q;
    p->tune_cap = item_cap;
    p->tune_cap1 = item_cap1;
}

int tp_measure_int_peak(int32_t cuda_device, int32_t variant, double *gops) {
    if (!gops) return TP_ERR_BAD_ARGUMENT;
    int dev = cuda_device;
    if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return TP_ERR_CUDA;
    DeviceGuard guard(dev);
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return TP_ERR_CUDA;
    unsigned *d_out = nullptr;
    cudaEvent_t e0, e1;
    if (cudaMalloc((void **)&d_out, 64) != cudaSuccess) return TP_ERR_CUDA;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8, iters = 1 << 14;
    double best = 0;
    cudaError_t e = tp_launch_int_peak(variant, blocks, 256, d_out, 0); // warm-up
    for (int rep = 0; rep < 5 && e == cudaSuccess; rep++) {
        cudaEventRecord(e0, 0);
        e = tp_launch_int_peak(variant, blocks, iters, d_out, 0);
        cudaEventRecord(e1, 0);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double ops = 16.0 * 256.0 * blocks * iters;
        if (ms > 0) best = std::max(best, ops / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    *gops = best;
    return e == cudaSuccess ? TP_OK : TP_ERR_CUDA;
}

int tp_build_weights(int32_t n, const int32_t *group, const int32_t *numa, const int32_t *link, uint8_t *out,
                     int32_t cuda_device) {
    if (n <= 0 || n > TP_MAX_N || !group || !numa || !link || !out) return TP_ERR_BAD_ARGUMENT;
    int dev = cuda_device;
    if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return TP_ERR_CUDA;
    DeviceGuard guard(dev);
    int32_t *d_group = nullptr, *d_numa = nullptr, *d_link = nullptr;
    uint8_t *d_out = nullptr;
    const size_t nn = (size_t)n * n;
    cudaError_t e = cudaMalloc((void **)&d_group, n * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_numa, n * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_link, nn * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_out, nn);
    if (e == cudaSuccess) e = cudaMemcpy(d_group, group, n * sizeof(int32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_numa, numa, n * sizeof(int32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_link, link, nn * sizeof(int32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = tp_launch_build_weights(n, d_group, d_numa, d_link, d_out, 0);
    if (e == cudaSuccess) e = cudaMemcpy(out, d_out, nn, cudaMemcpyDeviceToHost);
    cudaFree(d_group);
    cudaFree(d_numa);
    cudaFree(d_link);
    cudaFree(d_out);
    return e == cudaSuccess ? TP_OK : TP_ERR_CUDA;
}

} // extern "C"
