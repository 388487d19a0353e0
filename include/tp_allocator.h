/*
 * tp_allocator.h -- C ABI of the B200 topology-aware preferred-allocation policy.
 *
 * Drop-in boundary for ONE path of the reference device plugin: the best-effort
 * device selection behind the kubelet GetPreferredAllocation RPC, i.e. the Go
 * interface
 *
 *     type Policy interface {                       // internal/pkg/allocator/allocator.go:27-30
 *         Init(devs []*Device, topoDir string) error
 *         Allocate(available, required []string, size int) ([]string, error)
 *     }
 *
 * as implemented by BestEffortPolicy (internal/pkg/allocator/besteffort_policy.go:45-151,
 * device.go:254-442).  A Go policy keeps the string<->index maps and the sysfs
 * parsers (device.go:107-252) and calls these entry points through cgo; see
 * INTEGRATION.md for the stub.  Plain pointers and sizes only; every entry
 * point returns 0 / a count on success or a negative tp_status; nothing throws
 * or aborts across the boundary.  There is no CPU fallback: every request the
 * reference would search runs on the GPU, and all entry points fail with
 * TP_ERR_CUDA when no device is usable.
 *
 * Devices are named by INDEX into the arrays given to tp_create (0..n-1).
 */
#ifndef TP_ALLOCATOR_H
#define TP_ALLOCATOR_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TP_MAX_DEVICES 128 /* devices per policy (KFD nodes with a render minor) */
#define TP_MAX_SUBSET 16   /* largest size - len(required) the exhaustive kernels take */

typedef enum tp_status {
    TP_OK = 0,
    /* 1:1 with the reference's error strings, besteffort_policy.go:36-43 */
    TP_ERR_INVALID_SIZE = -1,          /* invalidSize,         check at :90-92   */
    TP_ERR_INVALID_AVAILABLE = -2,     /* invalidAvailable,    check at :94-96   */
    TP_ERR_INVALID_REQUIRED = -3,      /* invalidRequired,     check at :98-100  */
    TP_ERR_INVALID_REQ_AVAILABLE = -4, /* invalidReqAvailable, check at :102-104 */
    TP_ERR_INVALID_INIT = -5,          /* invalidInit,         checks at :106-108, :118-120 */
    TP_ERR_NO_CANDIDATE = -6,          /* noCandidateFound,    check at :122-124 */
    TP_ERR_INIT_FAILED = -7,           /* "Besteffort Policy init failed to initialize p2pWeights", :72-74 */
    /* not in the reference */
    TP_ERR_BAD_INDEX = -20,   /* unknown / duplicate device index (the reference nil-derefs, :61-68) */
    TP_ERR_BAD_ARGUMENT = -21,
    TP_ERR_TOO_LARGE = -22,   /* search space beyond what any kernel here enumerates */
    TP_ERR_STATIC_SCRATCH = -23, /* tp_allocate_static: request needs more than the fixed scratch holds */
    TP_ERR_CUDA = -30 /* any CUDA failure, on any of the policy's GPUs; text: tp_last_cuda_error */
} tp_status;

/*
 * What BestEffortPolicy.Init leaves behind (besteffort_policy.go:70-86), index based:
 *   node_id      Device.NodeId (device.go:58); distinct.
 *   group        dense index of Device.DevId: same physical GPU => same value
 *                (groupPartitionsByDevId, device.go:287-304).
 *   parent_rank  per group: rank of its ParentId under Go's string `<`, which
 *                orders groups with equally many free partitions (device.go:342-349).
 *   weight       n*n, row major, symmetric: p2pWeights[min][max] (device.go:135-157, :214),
 *                0 where the map has no entry (Go zero value, device.go:266).
 * All arrays are copied.
 */
typedef struct tp_topology {
    int32_t n_devices;
    int32_t n_groups;
    const int32_t *node_id;
    const int32_t *group;
    const int32_t *parent_rank;
    const uint8_t *weight;
} tp_topology;

#define TP_DEVICE_NONE (-2) /* tp_options.cuda_device: host-only handle (validation, tp_decode); searches fail */
#define TP_MAX_GPUS 16
/*
 * The reference policy lives in ONE process (constructed at internal/pkg/manager/manager.go:101, called at
 * internal/pkg/amdgpu/amdgpu.go:300-319), so more than one GPU has to be usable from one tp_policy, one call:
 * with n_extra_devices > 0 the policy also sets itself up on extra_devices[], and a tp_allocate whose search
 * holds at least multi_gpu_min_subsets subsets is cut into 1 + n_extra_devices shards (the same disjoint
 * slices tp_search_shard hands out), one per GPU, launched in parallel by per-GPU worker threads; every GPU
 * returns its packed winner to the host, which takes the minimum.  No collective library is involved: the
 * exchange is 16 bytes per GPU.  Smaller searches, batches and the other entry points stay on cuda_device.
 */
typedef struct tp_options {
    int32_t cuda_device;        /* ordinal; -1 = the current device; TP_DEVICE_NONE = no device */
    int32_t kernel;             /* 0 = auto; else force a kernel variant (tests / profiling) */
    uint64_t exhaustive_limit;  /* 0 = default; above this many subsets only Family members are enumerated */
    int32_t max_inflight;       /* 0 = default (8): concurrent tp_allocate calls before callers queue */
    int32_t n_extra_devices;    /* 0 = single GPU */
    int32_t extra_devices[TP_MAX_GPUS]; /* ordinals of the other GPUs (an ordinal may repeat: its shards then queue
                                         * on that GPU, which is how a one-GPU box tests this path) */
    uint64_t multi_gpu_min_subsets;     /* 0 = default (2^31) */
} tp_options;

/* Packed winner: (score, groups touched, group sequence) -- smaller is better; the
 * order is the order in which the reference's BFS emits candidates
 * (device.go:375-440) under the strict `<` of besteffort_policy.go:133-140. */
typedef struct tp_key {
    uint64_t hi; /* 0:1 | score:23 | groups:8 | first 32 bits of the group sequence */
    uint64_t lo; /* rest of the group sequence; 0 whenever (size-len(required))*ceil(log2 groups) <= 32 */
} tp_key;

typedef struct tp_result {
    int32_t score;           /* DeviceSet.TotalWeight of the winner (device.go:67-73) */
    int32_t groups_touched;
    int32_t searched;        /* 0: answered by a short-circuit (besteffort_policy.go:110-116), no GPU work */
    int32_t kernel;          /* kernel variant that ran */
    uint64_t subsets_scored; /* counted on the device */
    uint64_t subsets_total;  /* C(free, size-len(required)) */
    float kernel_ms;         /* CUDA-event time of the search kernel(s) on the call's stream (multi-GPU: the slowest GPU) */
    int32_t gpu_launches;    /* kernels launched for the request, over all GPUs */
    tp_key key;
} tp_result;

typedef struct tp_request {
    const int32_t *available; /* AvailableDeviceIDs as indices */
    int32_t n_available;
    const int32_t *required;  /* MustIncludeDeviceIDs as indices */
    int32_t n_required;
    int32_t size;             /* AllocationSize */
} tp_request;

typedef struct tp_policy tp_policy;

/* BestEffortPolicy.Init (besteffort_policy.go:70-86) minus the sysfs parse, which stays
 * with the caller: copies the topology to the GPU and sets up streams/buffers. */
int tp_create(const tp_topology *topo, const tp_options *opts, tp_policy **out);
void tp_destroy(tp_policy *p);

/* BestEffortPolicy.Allocate (besteffort_policy.go:88-151).  Writes `size` device
 * indices to out in the reference's output order (:141-148) and returns size, or a
 * negative tp_status.  Reentrant and blocking; info may be NULL. */
int tp_allocate(tp_policy *p, const int32_t *available, int32_t n_available, const int32_t *required,
                int32_t n_required, int32_t size, int32_t *out, tp_result *info);

/* Advisory mode (off the allocation path; no counterpart in the reference): the minimum of
 * TotalWeight over ALL subsets of the request, i.e. the same search without the restriction to
 * the sets the reference's BFS can emit (device.go:375-440).  Same arguments, validation,
 * short-circuits and return values as tp_allocate; out receives one optimal set (ties: the
 * first in canonical order; free devices in canonical order, required last) and info->score
 * its weight, to be compared with tp_allocate's info->score: equal on fully linked topologies,
 * lower where links are missing and the reference's candidate family misses the optimum.
 * Limits: (size - n_required) * ceil(log2 free devices) <= 96 and an enumerable subset
 * count (tp_options.exhaustive_limit), else TP_ERR_TOO_LARGE. */
int tp_global_optimum(tp_policy *p, const int32_t *available, int32_t n_available, const int32_t *required,
                      int32_t n_required, int32_t size, int32_t *out, tp_result *info);

/* The ContainerRequests loop of AMDGPUKFDImpl.GetPreferredAllocation
 * (internal/pkg/amdgpu/amdgpu.go:307-317) as one call.  out is n_requests rows of
 * out_stride int32; status[i] receives what tp_allocate would have returned. */
int tp_allocate_batch(tp_policy *p, int32_t n_requests, const tp_request *requests, int32_t *out,
                      int32_t out_stride, int32_t *status, tp_result *infos);

/* Multi-GPU: score only shard `shard_index` of `shard_count` disjoint, equally heavy slices of
 * the subset space (rank ranges, or a stride of the work-item list) and return the shard's
 * packed winner; the caller min-reduces keys
 * across shards (one ncclMin all-reduce on key.hi when key.lo is always 0) and
 * decodes the winner with tp_decode.  Requests answered by a short-circuit return
 * 1 and leave key untouched (call tp_allocate for them). */
int tp_search_shard(tp_policy *p, const int32_t *available, int32_t n_available, const int32_t *required,
                    int32_t n_required, int32_t size, int32_t shard_index, int32_t shard_count,
                    tp_key *key, tp_result *info);
int tp_decode(tp_policy *p, const int32_t *available, int32_t n_available, const int32_t *required,
              int32_t n_required, int32_t size, const tp_key *key, int32_t *out);

/* Host-only arithmetic of the sharding (works on a TP_DEVICE_NONE handle; no counterpart in the reference):
 * how many subsets shard `shard_index` of `shard_count` scores for this request, and with which kernel.
 * Over all shards the counts add up to C(free, size - n_required).  Returns 0, 1 for a request answered
 * by a short-circuit (nothing to score), or a negative tp_status. */
int tp_shard_subsets(tp_policy *p, const int32_t *available, int32_t n_available, const int32_t *required,
                     int32_t n_required, int32_t size, int32_t shard_index, int32_t shard_count,
                     uint64_t *subsets, int32_t *kernel);

/* Measurement hook: the request is prepared and uploaded once, then the search
 * kernel is launched `iters` times back to back on the policy's stream between two
 * CUDA events; *ms receives the total device time.  No host<->device copy in the
 * timed region.  Same result contract as tp_search_shard. */
int tp_search_resident(tp_policy *p, const int32_t *available, int32_t n_available, const int32_t *required,
                       int32_t n_required, int32_t size, int32_t shard_index, int32_t shard_count,
                       int32_t iters, float *ms, tp_key *key, tp_result *info);

/* Stream-ordered form of tp_search_shard for callers that own the stream and the device
 * memory (torch, NCCL): tp_prepare validates the request and uploads its descriptor;
 * tp_launch enqueues the search on `cuda_stream` (a cudaStream_t) and leaves
 *   d_out[0] = key.hi, d_out[1] = key.lo, d_out[2] = subsets scored
 * in DEVICE memory d_out (3 x uint64), with no host synchronisation, so that a
 * min-all-reduce of d_out[0] can be enqueued right behind it.  TP_KEY_NONE in d_out[0]
 * means the shard holds no candidate.  One launch at a time per tp_prepared. */
#define TP_KEY_NONE 0x7FFFFFFFFFFFFFFFull
typedef struct tp_prepared tp_prepared;
int tp_prepare(tp_policy *p, const int32_t *available, int32_t n_available, const int32_t *required,
               int32_t n_required, int32_t size, int32_t shard_index, int32_t shard_count, tp_prepared **out);
int tp_launch(tp_policy *p, tp_prepared *q, void *cuda_stream, uint64_t *d_out);
uint64_t tp_prepared_subsets(const tp_prepared *q); /* subsets this shard scores per launch */
void tp_prepared_free(tp_policy *p, tp_prepared *q);

/*
 * Multi-process runs (one process per GPU, e.g. under torch.distributed.run): the exchange of the shards'
 * winners fused into the search kernel instead of a collective launched behind it.  (No counterpart in the
 * reference; the single-process form of the same thing is tp_options.extra_devices.)  Every rank
 *   1. tp_exchange_create  -- a small buffer in its GPU's memory, exported as a CUDA IPC handle (64 bytes),
 *   2. gathers all ranks' handles by whatever means it has (a host all-gather),
 *   3. tp_exchange_connect -- maps the other ranks' buffers (peer access over NVLink),
 * and then launches with tp_launch_exchange instead of tp_launch: the kernel's last block stores its shard's
 * packed winner into every rank's buffer and waits for the others', so that d_out holds the GLOBAL minimum
 * (and the global count of subsets scored) on every rank when the kernel ends.  All ranks must make the same
 * sequence of tp_launch_exchange calls (an exchange is matched by its sequence number).  A rank that never
 * arrives makes the others give up after 5 s with TP_KEY_NONE in d_out[0].
 */
#define TP_IPC_HANDLE_BYTES 64
typedef struct tp_exchange tp_exchange;
int tp_exchange_create(tp_policy *p, int32_t rank, int32_t world, tp_exchange **out, void *handle_out /* 64 bytes */);
int tp_exchange_connect(tp_exchange *x, const void *handles /* world x 64 bytes, rank order */);
int tp_launch_exchange(tp_policy *p, tp_prepared *q, tp_exchange *x, void *cuda_stream, uint64_t *d_out);
void tp_exchange_destroy(tp_exchange *x);

/*
 * tp_allocate without a tp_policy and without creating any memory.  (This entry point has no
 * counterpart in the reference: the Go policy never leaves the host heap.)
 *
 * Meant for callers that cannot ask the CUDA runtime for memory when a request arrives: the
 * plugin shares its GPU with tenant jobs, so HBM may already be full (cudaMalloc fails) and
 * cudaMalloc/cudaFree wait for the device to go idle (they stall behind the tenants' kernels);
 * the same holds for hosts confined to memory another runtime handed them.  The caller brings
 *   topo       the host arrays of tp_create (read during the call only),
 *   d_weight   topo->weight already in DEVICE memory (n_devices^2 bytes),
 *   d_out      3 x uint64 of DEVICE memory: receives key.hi, key.lo, subsets scored,
 *   out        host array, `size` device indices in the reference's output order,
 * and the call makes no cudaMalloc, cudaFree, cudaHostAlloc, stream or event: its working
 * memory is a fixed 36 KiB array compiled into the library (held by every GPU that loads it,
 * until the process exits).  Consequences of that fixed array:
 *   - limit: size - n_required <= TP_MAX_SUBSET and at most 2^38 candidate subsets; beyond
 *     that the call returns TP_ERR_STATIC_SCRATCH and does no GPU work;
 *   - the array is shared: calls whose GPU work overlaps on one device corrupt each other,
 *     so run them one after another (the call itself blocks until its result is on the host);
 *   - the first call on each GPU looks up the array's address, loads the code of the sixteen
 *     kernel variants it may launch later (one per size - n_required; loading code takes
 *     device memory, so it is done up front, not on a later request) and uploads a small constant
 *     table, synchronising `cuda_stream`: make that first call early, before any stream capture
 *     and while memory is still to be had;
 *   - it always runs the plain rank-range kernel (LEX), not the fast ones behind tp_allocate: an
 *     8-of-80 request that tp_allocate answers in ~1.6 ms takes ~220 ms here.  It is a fallback for
 *     the moment allocation is impossible, not a second fast path.
 * Runs on the CURRENT device, which must own d_weight, d_out and cuda_stream (a cudaStream_t,
 * NULL for the default stream).  Same return values as tp_allocate otherwise.
 */
int tp_allocate_static(const tp_topology *topo, const uint8_t *d_weight, const int32_t *available,
                       int32_t n_available, const int32_t *required, int32_t n_required, int32_t size,
                       void *cuda_stream, uint64_t *d_out, int32_t *out, tp_result *info);

/* Bytes of the prepared request a search uploads to the GPU (with_plan: plus the work plan of the SPLIT / TRI
 * kernels): what bench.py counts as host-to-device traffic per request.  The way back is 24 bytes. */
int tp_descriptor_bytes(int32_t with_plan);

/* Development hook (tests, tuning sweeps; no counterpart in the reference): override the work plan of the
 * exhaustive kernels for requests prepared after the call -- walk_q = positions below the lane set (forces
 * kernel SPLIT), item_cap / item_cap1 = subsets per lane per work item / per cut value; 0 = default.  The
 * same three values are read ONCE, in tp_create, from the environment variables TP_SPLIT_Q, TP_SPLIT_CAP and
 * TP_SPLIT_CAP1; nothing on the request path reads the environment.  Not thread-safe against running calls. */
void tp_set_tuning(tp_policy *p, int32_t walk_q, uint64_t item_cap, uint64_t item_cap1);

/* Development hook for tp_allocate_batch (no counterpart in the reference): a request above 2^share_shift
 * subsets is cut into shares taken by different blocks (default: 24, 26 or 28 by the number of heavy requests in
 * the batch), a share into about items_per_warp work items per warp (default 4); light_first != 0 swaps the two waves of the batch (the light requests go to the
 * GPU first).  0 = default.  The first two are also read ONCE, in tp_create, from TP_BATCH_SHARE_SHIFT and
 * TP_BATCH_ITEMS_PER_WARP.  Not thread-safe against running calls. */
void tp_set_batch_tuning(tp_policy *p, int32_t share_shift, int32_t items_per_warp, int32_t light_first);

/* Roofline denominator: dependency-free integer-ALU microbenchmark on the given device.
 * variant 0: IADD3 only (alu pipe); 1: IADD3 + IMAD mix (alu + fma pipes); 2: packed
 * 16x2 min (VIMNMX).  *gops receives 10^9 lane-operations per second. */
int tp_measure_int_peak(int32_t cuda_device, int32_t variant, double *gops);

/* calculatePairWeight (device.go:135-157) for all pairs on the GPU: weight[i*n+j] from
 * per-device group / NUMA node and an n*n matrix of KFD link types (io_links/p2p_links
 * `type`; 11 = XGMI, 2 = PCIe; -1 = no link entry => weight 0, the Go map's zero value). */
int tp_build_weights(int32_t n, const int32_t *group, const int32_t *numa_node, const int32_t *link_type,
                     uint8_t *weight_out, int32_t cuda_device);

const char *tp_strerror(int status);
/* Last CUDA error text seen by this policy (diagnostics). */
const char *tp_last_cuda_error(const tp_policy *p);

#ifdef __cplusplus
}
#endif
#endif
