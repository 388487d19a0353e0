// tp_device.cuh -- device helpers shared by every kernel translation unit: the Family / RefKey
// evaluation of one subset, the 128-bit min reductions, the cross-block finish, and the staging
// of a request into shared memory.  Internal linkage: each .cu gets its own copy.
#pragma once
#include "tp_kernels.cuh"

// Opt `Kernel` in to the largest dynamic shared memory the current device allows (once per device) and check `need`
// against it.  The attribute is per function, not per launch: setting it to what each launch needs lets two
// host threads undo each other's value between the setting and the launch (two tp_allocate_batch calls, or two
// tp_allocate calls with different requests, then fail with "invalid argument").  A value that never changes
// cannot be undone.
#include <atomic>
template <auto Kernel>
static cudaError_t tp_opt_in_max_smem(size_t need) {
    static std::atomic<int> limit[64]; // per device ordinal; 0 = not asked yet
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    int lim = limit[dev].load(std::memory_order_acquire);
    if (lim == 0) {
        cudaFuncAttributes fa;
        if ((e = cudaFuncGetAttributes(&fa, Kernel)) != cudaSuccess) return e;
        int optin = 0;
        if ((e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev)) != cudaSuccess) return e;
        lim = optin - (int)fa.sharedSizeBytes;
        if (lim <= 0) return cudaErrorInvalidValue;
        if ((e = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, lim)) != cudaSuccess) return e;
        limit[dev].store(lim, std::memory_order_release); // (two threads may both get here: the same value)
    }
    return need <= (size_t)lim ? cudaSuccess : cudaErrorInvalidValue;
}


typedef unsigned long long u64;

// ---------------------------------------------------------------------------------------
// Family membership + RefKey of one scored subset (slow path, reached only when a subset
// scores no worse than the best Family member seen so far).  The subset arrives packed,
// one canonical position per byte, ascending.
// ---------------------------------------------------------------------------------------
static __device__ __noinline__ bool tp_family_key(u64 p0, u64 p1, int k, int score, int dbits,
                                           const uint8_t *s_group, const uint8_t *s_gstart,
                                           const uint8_t *s_gsize, TpKey &out) {
    auto elem = [&](int i) { return (int)(((i < 8 ? p0 : p1) >> ((i & 7) * 8)) & 0xFF); };
    TpKey key;
    tp_key_begin(key, score, 0);
    int n_whole = 0, m = 0, partial = -1, i = 0;
    while (i < k) {
        const int e = elem(i);
        const int g = s_group[e];
        int j = i + 1;
        while (j < k && s_group[elem(j)] == g) j++;
        const int cnt = j - i;
        m++;
        if (cnt == s_gsize[g]) {
            tp_key_put(key, n_whole++, dbits, (unsigned)g);
        } else {
            if (partial >= 0) return false; // two partial groups: the BFS never emits this
            if (e != s_gstart[g]) return false;
            for (int t = 1; t < cnt; t++)
                if (elem(i + t) != e + t) return false; // not the lowest-NodeId prefix
            partial = g;
        }
        i = j;
    }
    if (partial >= 0) tp_key_put(key, n_whole, dbits, (unsigned)partial);
    key.hi |= (u64)m << TP_KEY_GROUPS_SHIFT;
    out = key;
    return true;
}

// Can a subset still be a Family member when its `cnt` lowest canonical positions are the packed
// ones and every remaining member lies at a canonical position >= first_future?  A hole below a
// chosen member of a group can no longer be filled, and a group that ends before first_future
// stays as it is: two such prefixes, or one non-prefix, rule the whole leaf out (necessary
// condition only: tp_family_key decides).
static __device__ __noinline__ bool tp_family_alive(u64 p0, u64 p1, int cnt, int first_future,
                                                    const uint8_t *s_group, const uint8_t *s_gstart,
                                                    const uint8_t *s_gsize) {
    auto elem = [&](int i) { return (int)(((i < 8 ? p0 : p1) >> ((i & 7) * 8)) & 0xFF); };
    int closed_partials = 0, i = 0;
    while (i < cnt) {
        const int e = elem(i);
        const int g = s_group[e];
        int j = i + 1;
        while (j < cnt && s_group[elem(j)] == g) j++;
        const int c = j - i, gs = s_gstart[g], gz = s_gsize[g];
        if (c != gz) {
            if (e != gs) return false;
            for (int t = 1; t < c; t++)
                if (elem(i + t) != e + t) return false;
            if (gs + gz <= first_future && ++closed_partials > 1) return false;
        }
        i = j;
    }
    return true;
}

static __device__ __forceinline__ TpKey tp_warp_min_key(TpKey k) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        TpKey o;
        o.hi = __shfl_xor_sync(0xFFFFFFFFu, k.hi, off);
        o.lo = __shfl_xor_sync(0xFFFFFFFFu, k.lo, off);
        if (tp_key_less(o, k)) k = o;
    }
    return k;
}

// Block-wide min of the per-thread winners, then the cross-block step: every block stores
// its partial, the last one to arrive (ticket) reduces all partials into scratch.out and
// re-arms the scratch for the next launch.  Returns true in every thread of that last block.
static __device__ bool tp_block_finish(TpKey best, u64 counted, const TpScratch &sc) {
    __shared__ TpKey s_red[32];
    __shared__ u64 s_cnt[32];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    best = tp_warp_min_key(best);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) counted += __shfl_xor_sync(0xFFFFFFFFu, counted, off);
    if (lane == 0) {
        s_red[warp] = best;
        s_cnt[warp] = counted;
    }
    __syncthreads();
    if (warp == 0) {
        TpKey k = {TP_KEY_INF_HI, TP_KEY_INF_HI};
        u64 c = 0;
        if (lane < nwarps) {
            k = s_red[lane];
            c = s_cnt[lane];
        }
        k = tp_warp_min_key(k);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, off);
        if (lane == 0) {
            sc.partial[blockIdx.x] = k;
            atomicAdd(sc.scored, c);
            __threadfence();
            const unsigned t = atomicAdd(sc.ticket, 1u);
            s_last = (t == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (!s_last) return false;
    __threadfence();
    TpKey k = {TP_KEY_INF_HI, TP_KEY_INF_HI};
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
        TpKey o = sc.partial[i];
        if (tp_key_less(o, k)) k = o;
    }
    k = tp_warp_min_key(k);
    if (lane == 0) s_red[warp] = k;
    __syncthreads();
    if (warp == 0) {
        TpKey f = {TP_KEY_INF_HI, TP_KEY_INF_HI};
        if (lane < nwarps) f = s_red[lane];
        f = tp_warp_min_key(f);
        u64 total = 0;
        if (lane == 0) total = atomicExch(sc.scored, 0ull);
        if (sc.ex.world > 1) { // fused exchange (see TpExchangeDev): lane r talks to rank r
            f.hi = __shfl_sync(0xFFFFFFFFu, f.hi, 0);
            f.lo = __shfl_sync(0xFFFFFFFFu, f.lo, 0);
            total = __shfl_sync(0xFFFFFFFFu, total, 0);
            const int world = sc.ex.world;
            const size_t base = (size_t)(sc.ex.epoch & 1) * world;
            TpKey k = {TP_KEY_INF_HI, TP_KEY_INF_HI};
            u64 c = 0;
            if (lane < world) {
                volatile u64 *dst = sc.ex.peer[lane] + (base + sc.ex.rank) * TP_EX_SLOT_WORDS;
                dst[0] = f.hi;
                dst[1] = f.lo;
                dst[2] = total;
                __threadfence_system();
                dst[3] = sc.ex.epoch;
                volatile u64 *src = sc.ex.peer[sc.ex.rank] + (base + lane) * TP_EX_SLOT_WORDS;
                u64 t0, t1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
                bool ok = true;
                while (src[3] != sc.ex.epoch) { // a rank that never publishes must not hang the others
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                    if (t1 - t0 > 5000000000ull) { // 5 s
                        ok = false;
                        break;
                    }
                }
                __threadfence_system();
                if (ok) {
                    k.hi = src[0];
                    k.lo = src[1];
                    c = src[2];
                } else {
                    c = ~0ull >> 1; // poison: the count can no longer add up
                }
            }
            f = tp_warp_min_key(k);
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, off);
            total = c;
        }
        if (lane == 0) {
            *sc.out = f;
            *sc.out_scored = total;
            *sc.ticket = 0u;
        }
    }
    return true;
}

// Block-wide (min key, sum count); the result is valid in thread 0.  Ends with a __syncthreads,
// so shared memory may be reused right after.
static __device__ void tp_block_reduce(TpKey &best, u64 &counted) {
    __shared__ TpKey s_red[32];
    __shared__ u64 s_cnt[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    best = tp_warp_min_key(best);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) counted += __shfl_xor_sync(0xFFFFFFFFu, counted, off);
    if (lane == 0) {
        s_red[warp] = best;
        s_cnt[warp] = counted;
    }
    __syncthreads();
    if (warp == 0) {
        TpKey k = {TP_KEY_INF_HI, TP_KEY_INF_HI};
        u64 c = 0;
        if (lane < nwarps) {
            k = s_red[lane];
            c = s_cnt[lane];
        }
        best = tp_warp_min_key(k);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, off);
        counted = c;
    }
    __syncthreads();
}

// Stage the request into shared memory: pair weights in canonical order, the unary term
// u[x] = sum over required r of W[x][r], and the group tables.  Returns the seed bound:
// the score of canonical positions {0..k-1}, which is always a Family member.
struct TpSmem {
    u64 *binom;       // [(n+1) * (K+1)]
    uint16_t *u;      // [n]
    uint8_t *w;       // [n * stride]
    uint8_t *group, *gstart, *gsize;
    int stride;
};

static __device__ __forceinline__ TpSmem tp_smem_carve(unsigned char *smem, int n, int K) {
    TpSmem s;
    s.stride = (n + 3) & ~3;
    s.binom = (u64 *)smem;
    s.u = (uint16_t *)(s.binom + (n + 1) * (K + 1));
    s.w = (uint8_t *)(s.u + ((n + 7) & ~7));
    s.group = s.w + n * s.stride;
    s.gstart = s.group + n;
    s.gsize = s.gstart + n;
    return s;
}

static __host__ __device__ size_t tp_smem_bytes(int n, int K) {
    int stride = (n + 3) & ~3;
    return (size_t)(n + 1) * (K + 1) * 8 + (size_t)((n + 7) & ~7) * 2 + (size_t)n * stride + 3 * (size_t)n + 16;
}

static __device__ void tp_stage_request(const TpSmem &s, const uint8_t *__restrict__ W, int nd,
                                 const u64 *__restrict__ binom_g, const TpReq *__restrict__ rq, int K,
                                 int *s_bound) {
    // Rows by warp, columns by lane: no integer division, and the canonical order is read from
    // shared memory instead of twice per element from the descriptor.
    __shared__ uint8_t s_canon[TP_MAX_N];
    const int n = rq->n, tid = threadIdx.x, nt = blockDim.x;
    const int warp = tid >> 5, lane = tid & 31, nw = nt >> 5;
    for (int i = tid; i < n; i += nt) s_canon[i] = rq->canon[i];
    for (int a = warp; a <= n; a += nw)
        for (int j = lane; j <= K; j += 32) s.binom[a * (K + 1) + j] = binom_g[a * TP_BINOM_COLS + j];
    const bool tiny = n <= 16; // the latency path (often a single warp): a flat loop, no round trip through s_canon
    if (tiny) {
        for (int i = tid; i < n * n; i += nt) {
            const int r = i / n, c = i - r * n;
            s.w[r * s.stride + c] = (r == c) ? 0 : W[(int)rq->canon[r] * nd + rq->canon[c]];
        }
    }
    __syncthreads();
    if (!tiny) {
        for (int r = warp; r < n; r += nw) {
            const uint8_t *wrow = W + (int)s_canon[r] * nd;
            for (int c = lane; c < n; c += 32) s.w[r * s.stride + c] = (r == c) ? 0 : wrow[s_canon[c]];
        }
    }
    for (int i = tid; i < n; i += nt) {
        const uint8_t *wrow = W + (int)s_canon[i] * nd;
        int u = 0;
        for (int r = 0; r < rq->n_req; r++) u += wrow[rq->req[r]];
        s.u[i] = (uint16_t)u;
        s.group[i] = rq->group_of[i];
        s.gstart[i] = rq->gstart[i];
        s.gsize[i] = rq->gsize[i];
    }
    __syncthreads();
    if (tid == 0) {
        int sc = rq->c_req;
        for (int i = 0; i < K; i++) {
            sc += s.u[i];
            for (int j = 0; j < i; j++) sc += s.w[j * s.stride + i];
        }
        *s_bound = sc;
    }
    __syncthreads();
}
