/* oracle/ref_internal.h -- TEST INFRASTRUCTURE: shared between ref_policy.c and exhaustive.c */
#ifndef REF_INTERNAL_H
#define REF_INTERNAL_H
#include "ref_policy.h"

/* DevicePartitions, device.go:75-80 */
typedef struct partitions {
    char parent_id[REF_ID_LEN];
    char dev_id[REF_ID_LEN];
    int *ids; /* node ids in Init order */
    int n_ids;
} partitions;

struct ref_policy {
    ref_device *devices;
    int n_devices;      /* len(b.devices); 0 until a successful Init */
    partitions *parts;  /* b.devicePartitions, in first-appearance order */
    int n_parts;
    /* b.p2pWeights: map[int]map[int]int, dense over node ids [0, dim) */
    int dim;
    int *w;                 /* dim*dim, row = from, col = to */
    unsigned char *present; /* dim*dim */
    unsigned char *row_present;
};

typedef struct fpart { /* filtered DevicePartitions, device.go:335-339 */
    const char *parent_id;
    int *ids;
    int n_ids;
    int order; /* first-appearance index: deterministic stand-in for map order */
} fpart;


int ref_weights_get(const ref_policy *p, int from, int to);
fpart *ref_filter_partitions(const ref_policy *p, const int *avail_nodes, int n_avail,
                             const int *req_nodes, int n_req, int *n_out);
const ref_device *ref_device_by_id(const ref_policy *p, const char *id);
#endif
