/*
 * oracle/ref_policy.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C) of the reference's best-effort allocation policy,
 * /root/reference/internal/pkg/allocator/{device.go,besteffort_policy.go}.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this library; the product path
 * (rocm-k8s-device-plugin_b200/csrc, host/) never links or calls it.
 *
 * Parity status: PINNED.  The Go toolchain is absent in the build container
 * (`go`, `gccgo` not found), so the reference cannot be executed; instead the
 * restatement is checked against every known-answer test the reference holds
 * for this path (besteffort_policy_test.go:25-216, device_test.go:80-169):
 * see tests/test_oracle_golden.py.
 */
#ifndef REF_POLICY_H
#define REF_POLICY_H

#ifdef __cplusplus
extern "C" {
#endif

#define REF_ID_LEN 64

/* Mirrors allocator.Device (device.go:56-65); only the four fields the
 * policy reads are kept (Id, NodeId, NumaNode, DevId). */
typedef struct ref_device {
    char id[REF_ID_LEN];
    int node_id;
    int numa_node;
    char dev_id[REF_ID_LEN];
} ref_device;

/* Error classes; ref_strerror() returns the reference's exact strings
 * (besteffort_policy.go:36-43). */
enum {
    REF_OK = 0,
    REF_ERR_INVALID_SIZE = -1,          /* "allocation size can not be negative" */
    REF_ERR_INVALID_AVAILABLE = -2,     /* "available devices count less than allocation size" */
    REF_ERR_INVALID_REQUIRED = -3,      /* "must_include devices size is more than allocation size" */
    REF_ERR_INVALID_REQ_AVAILABLE = -4, /* "must_include length should be less than or equal to avilable device size" */
    REF_ERR_INVALID_INIT = -5,          /* "Init method must be called before Allocate" */
    REF_ERR_NO_CANDIDATE = -6,          /* "No candidate subset found with matching criteria" */
    REF_ERR_INIT_FAILED = -7,           /* "Besteffort Policy init failed to initialize p2pWeights" */
    REF_ERR_EMPTY_DEVICES = -8,         /* "Devices list is empty. Unable to calculate pair wise weights" */
    REF_ERR_PANIC = -9                  /* the Go code would nil-deref (unknown id / no candidate) */
};

typedef struct ref_policy ref_policy;

typedef struct ref_result {
    int score;        /* TotalWeight of the chosen DeviceSet */
    int n_candidates; /* len(subsetsFinal) */
    int searched;     /* 1 if getCandidateDeviceSubsets ran, 0 on a short-circuit */
} ref_result;

ref_policy *ref_policy_new(void);
void ref_policy_free(ref_policy *p);

/* BestEffortPolicy.Init (besteffort_policy.go:70-86): weights parsed from a
 * KFD sysfs tree (device.go:107-252).  topo_dir must be given explicitly. */
int ref_policy_init(ref_policy *p, const ref_device *devs, int n, const char *topo_dir);

/* Same as Init but the p2pWeights map is given as (from,to,weight) triples
 * (from < to, node ids) instead of parsed from sysfs: used on machines where
 * the fixture trees are not present (the GPU box). */
int ref_policy_init_pairs(ref_policy *p, const ref_device *devs, int n,
                          const int *from, const int *to, const int *weight, int n_pairs);

/* BestEffortPolicy.Allocate (besteffort_policy.go:88-151).  out receives up
 * to `size` ids (REF_ID_LEN bytes each); returns the number written or a
 * negative error. */
int ref_policy_allocate(ref_policy *p, const char *const *avail, int n_avail,
                        const char *const *req, int n_req, int size,
                        char *out, ref_result *res);

/* Introspection used by the known-answer tests. */
int ref_policy_weight_rows(const ref_policy *p);               /* len(p2pWeights) */
int ref_policy_group_count(const ref_policy *p);               /* len(devicePartitions) */
int ref_policy_pair_count(const ref_policy *p);                /* number of (from,to) entries */
int ref_policy_weight(const ref_policy *p, int from, int to);  /* p2pWeights[from][to], 0 if absent */
int ref_policy_has_pair(const ref_policy *p, int from, int to);
int ref_policy_max_node(const ref_policy *p);

const char *ref_strerror(int code);

#ifdef __cplusplus
}
#endif
#endif
