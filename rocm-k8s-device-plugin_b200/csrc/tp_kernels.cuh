// tp_kernels.cuh -- launcher interface between the C ABI (tp_abi.cu) and the kernels.
#pragma once
#include "tp_common.cuh"

enum TpKernelId {
    TP_KERNEL_AUTO = 0,
    TP_KERNEL_LEX = 1,    // tp_exhaustive_lex: one rank range per thread, lexicographic walk
    TP_KERNEL_SPLIT = 2,  // tp_exhaustive_split: lane sets x warp-uniform walk x unrolled pair nest
    TP_KERNEL_FAMILY = 3, // tp_family: Family members only (search spaces too large to enumerate)
    TP_KERNEL_TRI = 4,    // tp_exhaustive_tri: lane sets x warp-uniform walk x triple table (k' >= 6, n - p <= 80)
};

struct TpLaunchInfo {
    int launches;
    int blocks;
    int threads;
};

// Kernel SPLIT: shard s of P scores the plan's items i with (i + i/P) mod P == s; d_counter is a zeroable word
// of scratch used to hand out items.  Needs n - plan.p <= TP_SPLIT_MAX_M.
#define TP_SPLIT_MAX_M 128
cudaError_t tp_launch_exhaustive_split(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                                       const TpReq *d_rq, const TpReq &h_rq, const TpPlan *d_plan,
                                       const TpPlan &h_plan, int shard_idx, int shard_cnt, const TpScratch &scratch,
                                       unsigned *d_counter, int sm_count, cudaStream_t stream, TpLaunchInfo *info);
// Batch mode: one block per request of `batch.order`; d_reqs / d_plans are the batch arrays, (p, q,
// max_n, max_k) describe the group (all its requests share p and q, hence the instantiation).
cudaError_t tp_launch_batch_split(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                                  const TpReq *d_reqs, const TpPlan *d_plans, int p, int q, int max_n, int max_k,
                                  const TpBatch &batch, int sm_count, cudaStream_t stream, TpLaunchInfo *info);
cudaError_t tp_launch_batch_lex(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                                const TpReq *d_reqs, int k, int max_n, const TpBatch &batch, int sm_count,
                                cudaStream_t stream, TpLaunchInfo *info);
// per-instantiation launchers, defined in tp_split_m<M>_d<D>.cu (D = q - 2)
#define TP_DECLARE_SPLIT_LAUNCHER(M, D)                                                                          \
    cudaError_t tp_launch_split_m##M##_d##D(const uint8_t *W, int nd, const unsigned long long *binom,          \
                                            const TpReq *d_rq, const TpReq &h, const TpPlan *d_plan, int si,    \
                                            int sn, const TpScratch &sc, unsigned *d_counter, int sms,          \
                                            cudaStream_t st, TpLaunchInfo *info, const TpBatch &batch);
#define TP_DECLARE_SPLIT_LAUNCHERS(M)                                                                  \
    TP_DECLARE_SPLIT_LAUNCHER(M, 0) TP_DECLARE_SPLIT_LAUNCHER(M, 1) TP_DECLARE_SPLIT_LAUNCHER(M, 2)    \
    TP_DECLARE_SPLIT_LAUNCHER(M, 3) TP_DECLARE_SPLIT_LAUNCHER(M, 4)
TP_DECLARE_SPLIT_LAUNCHERS(64)
TP_DECLARE_SPLIT_LAUNCHERS(80)
TP_DECLARE_SPLIT_LAUNCHERS(96)
TP_DECLARE_SPLIT_LAUNCHERS(128)

// Kernel TRI (tp_tri.cuh): same plan, shards and scratch as SPLIT; needs n - plan.p <= TP_TRI_MAX_M, plan.q in
// [3, 6] and plan.tri_gcap from tp_tri_fit.
#define TP_TRI_MAX_M 80
#define TP_TRI_MAX_DW 3
cudaError_t tp_launch_exhaustive_tri(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                                     const TpReq *d_rq, const TpReq &h_rq, const TpPlan *d_plan,
                                     const TpPlan &h_plan, int shard_idx, int shard_cnt, const TpScratch &scratch,
                                     unsigned *d_counter, int sm_count, cudaStream_t stream, TpLaunchInfo *info);
cudaError_t tp_launch_batch_tri(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                                const TpReq *d_reqs, const TpPlan *d_plans, int p, int q, int max_n, int max_k,
                                int max_gcap, const TpBatch &batch, int sm_count, cudaStream_t stream,
                                TpLaunchInfo *info);
// Largest table cap (leaf pairs) whose shared memory fits beside everything else the kernel stages for a
// request with n free devices, k' = k, |A| = p and q positions below the lane set; 0 = kernel TRI cannot run.
int tp_tri_fit(int n, int k, int p, int q);
int tp_tri_gc(); // TP_TRI_GC of the build (defined next to the kernel, tp_tri_m64_d1.cu)
#define TP_DECLARE_TRI_LAUNCHER(M, D)                                                                            \
    cudaError_t tp_launch_tri_m##M##_d##D(const uint8_t *W, int nd, const unsigned long long *binom,            \
                                          const TpReq *d_rq, const TpReq &h, const TpPlan *d_plan, int gcap,    \
                                          int si, int sn, const TpScratch &sc, unsigned *d_counter, int sms,    \
                                          cudaStream_t st, TpLaunchInfo *info, const TpBatch &batch);           \
    size_t tp_tri_smem_m##M##_d##D(int n, int k, int gcap);
#define TP_DECLARE_TRI_LAUNCHERS(M) \
    TP_DECLARE_TRI_LAUNCHER(M, 1) TP_DECLARE_TRI_LAUNCHER(M, 2) TP_DECLARE_TRI_LAUNCHER(M, 3)
TP_DECLARE_TRI_LAUNCHERS(64)
TP_DECLARE_TRI_LAUNCHERS(80)

// Score ranks [lo, hi) of the C(n, k') lexicographic subset order of request *rq (device
// pointer; h_rq is the host copy used for launch geometry) and leave the winner in scratch.out.
cudaError_t tp_launch_exhaustive_lex(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                                     const TpReq *d_rq, const TpReq &h_rq, unsigned long long lo,
                                     unsigned long long hi, const TpScratch &scratch, int sm_count,
                                     cudaStream_t stream, TpLaunchInfo *info);

// Latency path: whole request in ONE block, descriptor by value, winner + flag written to mapped
// host memory (layout: key.hi, key.lo, scored, flag:u32).  For totals <= TP_FAST_MAX_SUBSETS.
#define TP_FAST_MAX_SUBSETS 32768
cudaError_t tp_launch_exhaustive_lex_small(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                                           const TpReq &h_rq, void *d_out_mapped, unsigned epoch,
                                           cudaStream_t stream);

// Family members only: one thread per bitmask over the groups with free devices, masks [lo, hi).
#define TP_FAMILY_MAX_GROUPS 24
cudaError_t tp_launch_family(const uint8_t *d_weight, int n_devices, const unsigned long long *d_binom,
                             const TpReq *d_rq, const TpReq &h_rq, unsigned long long mask_lo,
                             unsigned long long mask_hi, const TpScratch &scratch, int sm_count,
                             cudaStream_t stream, TpLaunchInfo *info);

// A rank whose shard is empty still takes part in the fused exchange (TpExchangeDev): "no candidate, nothing scored".
cudaError_t tp_launch_exchange_only(const TpScratch &scratch, cudaStream_t stream);

cudaError_t tp_launch_build_weights(int n, const int32_t *d_group, const int32_t *d_numa, const int32_t *d_link,
                                    uint8_t *d_out, cudaStream_t stream);

#define TP_BINOM_COLS (TP_MAX_K + 1)

// LEX search whose scratch is the library's file-scope __device__ array (see tp_kernels.cu):
// current device, caller-owned d_weight / d_out, nothing allocated.
#define TP_STATIC_MAX_SUBSETS (1ull << 38)
cudaError_t tp_launch_exhaustive_lex_static(const uint8_t *d_weight, int n_devices,
                                            const unsigned long long *h_binom, const TpReq &h_rq,
                                            unsigned long long lo, unsigned long long hi,
                                            unsigned long long *d_out, cudaStream_t stream, TpLaunchInfo *info);

// 16 * 256 * blocks * iters lane-operations per launch.
cudaError_t tp_launch_int_peak(int variant, int blocks, int iters, unsigned *d_out, cudaStream_t stream);
