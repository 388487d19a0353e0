/*
 * oracle/ref_policy.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see ref_policy.h).
 *
 * Restatement in plain C of the reference policy.  Each function names the
 * reference lines it follows (paths relative to
 * /root/reference/internal/pkg/allocator/).  Go maps are modelled as dense
 * arrays indexed by KFD node id; Go's random map-iteration order never
 * reaches the result except where noted.
 */
#define _GNU_SOURCE
#include "ref_policy.h"

#include <dirent.h>
#include <errno.h>
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* device.go:38-54 */
enum {
    sameDevIdWeight = 10,
    xgmiLinkWeight = 10,
    sameNumaNodeWeight = 10,
    differentDevIdWeight = 20,
    differentNumaNodeWeight = 20,
    pcieLinkWeight = 40,
    otherLinkWeight = 50
};

#include "ref_internal.h"

const char *ref_strerror(int code) {
    switch (code) {
    case REF_OK: return "";
    case REF_ERR_INVALID_SIZE: return "allocation size can not be negative";
    case REF_ERR_INVALID_AVAILABLE: return "available devices count less than allocation size";
    case REF_ERR_INVALID_REQUIRED: return "must_include devices size is more than allocation size";
    case REF_ERR_INVALID_REQ_AVAILABLE:
        return "must_include length should be less than or equal to avilable device size";
    case REF_ERR_INVALID_INIT: return "Init method must be called before Allocate";
    case REF_ERR_NO_CANDIDATE: return "No candidate subset found with matching criteria";
    case REF_ERR_INIT_FAILED: return "Besteffort Policy init failed to initialize p2pWeights";
    case REF_ERR_EMPTY_DEVICES: return "Devices list is empty. Unable to calculate pair wise weights";
    case REF_ERR_PANIC: return "reference would panic (nil pointer dereference)";
    }
    return "unknown";
}

ref_policy *ref_policy_new(void) { return (ref_policy *)calloc(1, sizeof(ref_policy)); }

static void free_parts(ref_policy *p) {
    for (int i = 0; i < p->n_parts; i++) free(p->parts[i].ids);
    free(p->parts);
    p->parts = NULL;
    p->n_parts = 0;
}

void ref_policy_free(ref_policy *p) {
    if (!p) return;
    free_parts(p);
    free(p->devices);
    free(p->w);
    free(p->present);
    free(p->row_present);
    free(p);
}

/* ---- p2pWeights map model ---------------------------------------------- */

static int weights_alloc(ref_policy *p, const ref_device *devs, int n) {
    int maxid = 0;
    for (int i = 0; i < n; i++)
        if (devs[i].node_id > maxid) maxid = devs[i].node_id;
    if (p->w && p->dim > maxid) return 0; /* Init on a used policy keeps old entries, like the Go map */
    int dim = maxid + 1;
    int *w = (int *)calloc((size_t)dim * dim, sizeof(int));
    unsigned char *pr = (unsigned char *)calloc((size_t)dim * dim, 1);
    unsigned char *rp = (unsigned char *)calloc((size_t)dim, 1);
    if (!w || !pr || !rp) return -1;
    for (int f = 0; f < p->dim; f++) {
        rp[f] = p->row_present[f];
        for (int t = 0; t < p->dim; t++) {
            w[(size_t)f * dim + t] = p->w[(size_t)f * p->dim + t];
            pr[(size_t)f * dim + t] = p->present[(size_t)f * p->dim + t];
        }
    }
    free(p->w);
    free(p->present);
    free(p->row_present);
    p->w = w;
    p->present = pr;
    p->row_present = rp;
    p->dim = dim;
    return 0;
}

static void weights_set(ref_policy *p, int from, int to, int weight) {
    p->row_present[from] = 1; /* p2pWeights[from] = make(map[int]int), device.go:211-213 */
    p->present[(size_t)from * p->dim + to] = 1;
    p->w[(size_t)from * p->dim + to] = weight; /* last write wins, device.go:214 */
}

/* p2pWeights[from][to] with Go's zero value for an absent key (device.go:266) */
int ref_weights_get(const ref_policy *p, int from, int to) {
    if (from < 0 || to < 0 || from >= p->dim || to >= p->dim) return 0;
    return p->w[(size_t)from * p->dim + to];
}

int ref_policy_weight_rows(const ref_policy *p) {
    int c = 0;
    for (int i = 0; i < p->dim; i++) c += p->row_present[i];
    return c;
}
int ref_policy_pair_count(const ref_policy *p) {
    int c = 0;
    for (size_t i = 0; i < (size_t)p->dim * p->dim; i++) c += p->present[i];
    return c;
}
int ref_policy_group_count(const ref_policy *p) { return p->n_parts; }
int ref_policy_weight(const ref_policy *p, int from, int to) { return ref_weights_get(p, from, to); }
int ref_policy_has_pair(const ref_policy *p, int from, int to) {
    if (from < 0 || to < 0 || from >= p->dim || to >= p->dim) return 0;
    return p->present[(size_t)from * p->dim + to];
}
int ref_policy_max_node(const ref_policy *p) { return p->dim - 1; }

/* ---- sysfs parsing, device.go:107-252 ---------------------------------- */

static int is_go_space(char c) { /* regexp \s == [\t\n\f\r ] */
    return c == ' ' || c == '\t' || c == '\n' || c == '\f' || c == '\r';
}

/* strconv.ParseInt(s, 0, 32) for a string of decimal digits (all the regex
 * can capture): base 0 means a leading "0" selects octal.  Returns 0 on
 * success, -1 on syntax/range error. */
static int go_parse_int_base0_32(const char *s, int len, int *out) {
    int base = 10, i = 0;
    if (len > 1 && s[0] == '0') {
        base = 8;
        i = 1;
    }
    long long v = 0;
    for (; i < len; i++) {
        int d = s[i] - '0';
        if (d < 0 || d >= base) return -1;
        v = v * base + d;
        if (v > 2147483647LL) return -1;
    }
    *out = (int)v;
    return 0;
}

/* Unanchored match of `<key>\s(\d+)` on one line, leftmost match
 * (regexp.FindStringSubmatch, device.go:117).  Returns 1 if matched. */
static int match_key(const char *line, const char *key, const char **digits, int *ndigits) {
    size_t kl = strlen(key);
    for (const char *p = strstr(line, key); p; p = strstr(p + 1, key)) {
        const char *q = p + kl;
        if (!is_go_space(*q)) continue;
        q++;
        int n = 0;
        while (q[n] >= '0' && q[n] <= '9') n++;
        if (n == 0) continue;
        *digits = q;
        *ndigits = n;
        return 1;
    }
    return 0;
}

/* fetchTopoProperties, device.go:107-133.  Returns 0 ok, 1 open error (res =
 * {0}), 2 parse error (res = nil in Go). */
static int fetch_topo_properties(const char *path, const char *const *keys, int nkeys, int *res) {
    for (int i = 0; i < nkeys; i++) res[i] = 0;
    FILE *f = fopen(path, "r");
    if (!f) return 1;
    char *line = NULL;
    size_t cap = 0;
    ssize_t len;
    int rc = 0;
    while ((len = getline(&line, &cap, f)) >= 0) {
        /* bufio.ScanLines strips "\n" and one trailing "\r" */
        if (len > 0 && line[len - 1] == '\n') line[--len] = 0;
        if (len > 0 && line[len - 1] == '\r') line[--len] = 0;
        for (int i = 0; i < nkeys; i++) {
            const char *d;
            int nd;
            if (!match_key(line, keys[i], &d, &nd)) continue;
            int v;
            if (go_parse_int_base0_32(d, nd, &v) != 0) {
                rc = 2;
                goto done;
            }
            res[i] = v; /* later lines overwrite earlier ones */
        }
    }
done:
    free(line);
    fclose(f);
    return rc;
}

static int cmp_str(const void *a, const void *b) {
    return strcmp(*(const char *const *)a, *(const char *const *)b);
}

/* filepath.Glob(dir + "/[0-9]*"): names starting with a digit, sorted as
 * byte strings ("0","1","10","11",...,"2",...).  A missing directory gives an
 * empty list and no error, as in Go. */
static char **glob_digit_entries(const char *dir, int *count) {
    *count = 0;
    DIR *d = opendir(dir);
    if (!d) return NULL;
    char **names = NULL;
    int n = 0, cap = 0;
    struct dirent *e;
    while ((e = readdir(d)) != NULL) {
        if (e->d_name[0] < '0' || e->d_name[0] > '9') continue;
        if (n == cap) {
            cap = cap ? cap * 2 : 64;
            names = (char **)realloc(names, (size_t)cap * sizeof(char *));
        }
        names[n++] = strdup(e->d_name);
    }
    closedir(d);
    if (n) qsort(names, (size_t)n, sizeof(char *), cmp_str);
    *count = n;
    return names;
}

static void free_names(char **names, int n) {
    for (int i = 0; i < n; i++) free(names[i]);
    free(names);
}

/* calculatePairWeight, device.go:135-157 */
static int calculate_pair_weight(const ref_device *from, const ref_device *to, int link_type) {
    int weight = 0;
    weight += strcmp(from->dev_id, to->dev_id) == 0 ? sameDevIdWeight : differentDevIdWeight;
    if (link_type == 11)
        weight += xgmiLinkWeight;
    else if (link_type == 2)
        weight += pcieLinkWeight;
    else
        weight += otherLinkWeight;
    weight += from->numa_node == to->numa_node ? sameNumaNodeWeight : differentNumaNodeWeight;
    return weight;
}

static int node_is_device(const ref_device *devs, int n, int node) {
    for (int i = 0; i < n; i++)
        if (devs[i].node_id == node) return 1;
    return 0;
}

/* scanAndPopulatePeerWeights, device.go:159-218 */
static void scan_and_populate_peer_weights(ref_policy *p, const char *from_path,
                                           const ref_device *devs, int n) {
    static const char *const keys[3] = {"node_from", "node_to", "type"};
    static const char *const subdirs[2] = {"io_links", "p2p_links"}; /* io first, then p2p: :160-168 */
    for (int s = 0; s < 2; s++) {
        char dir[PATH_MAX];
        snprintf(dir, sizeof dir, "%s/%s", from_path, subdirs[s]);
        int cnt;
        char **names = glob_digit_entries(dir, &cnt);
        for (int i = 0; i < cnt; i++) {
            char prop[PATH_MAX];
            snprintf(prop, sizeof prop, "%s/%s/properties", dir, names[i]);
            int vals[3];
            if (fetch_topo_properties(prop, keys, 3, vals) != 0) continue; /* :172-174 */
            int from, to;
            if (vals[0] < vals[1]) { /* :181-188 */
                from = vals[0];
                to = vals[1];
            } else {
                from = vals[1];
                to = vals[0];
            }
            if (!node_is_device(devs, n, from)) continue; /* lookupNodes, :189-194 */
            if (!node_is_device(devs, n, to)) continue;
            const ref_device *from_dev = NULL, *to_dev = NULL;
            int found = 0;
            for (int d = 0; d < n; d++) { /* :196-208 */
                if (devs[d].node_id == from) from_dev = &devs[d];
                if (devs[d].node_id == to) to_dev = &devs[d];
                if (from_dev && to_dev) {
                    found = 1;
                    break;
                }
            }
            if (found) weights_set(p, from, to, calculate_pair_weight(from_dev, to_dev, vals[2]));
        }
        free_names(names, cnt);
    }
}

/* fetchAllPairWeights, device.go:220-252 */
static int fetch_all_pair_weights(ref_policy *p, const ref_device *devs, int n, const char *folder) {
    if (n == 0) return REF_ERR_EMPTY_DEVICES;
    static const char *const key[1] = {"drm_render_minor"};
    int cnt;
    char **names = glob_digit_entries(folder, &cnt);
    for (int i = 0; i < cnt; i++) {
        char path[PATH_MAX], prop[PATH_MAX];
        snprintf(path, sizeof path, "%s/%s", folder, names[i]);
        snprintf(prop, sizeof prop, "%s/properties", path);
        int val[1];
        int rc = fetch_topo_properties(prop, key, 1, val);
        if (rc != 0 || val[0] <= 0) continue; /* CPU nodes: drm_render_minor 0, :242 */
        scan_and_populate_peer_weights(p, path, devs, n);
    }
    free_names(names, cnt);
    return REF_OK;
}

/* groupPartitionsByDevId, device.go:287-304 */
static void group_partitions_by_dev_id(ref_policy *p, const ref_device *devs, int n) {
    free_parts(p);
    p->parts = (partitions *)calloc((size_t)(n > 0 ? n : 1), sizeof(partitions));
    for (int i = 0; i < n; i++) {
        int g = -1;
        for (int j = 0; j < p->n_parts; j++)
            if (strcmp(p->parts[j].dev_id, devs[i].dev_id) == 0) {
                g = j;
                break;
            }
        if (g < 0) {
            g = p->n_parts++;
            snprintf(p->parts[g].dev_id, REF_ID_LEN, "%s", devs[i].dev_id);
            p->parts[g].ids = (int *)calloc((size_t)n, sizeof(int));
        }
        if (strstr(devs[i].id, "amdgpu_xcp") == NULL)
            snprintf(p->parts[g].parent_id, REF_ID_LEN, "%s", devs[i].id);
        p->parts[g].ids[p->parts[g].n_ids++] = devs[i].node_id;
    }
}

static int finish_init(ref_policy *p, const ref_device *devs, int n, int err) {
    if (ref_policy_weight_rows(p) == 0) return REF_ERR_INIT_FAILED; /* besteffort_policy.go:72-74 */
    if (err == REF_OK) {                                            /* :75-85 */
        free(p->devices);
        p->devices = (ref_device *)malloc((size_t)n * sizeof(ref_device));
        memcpy(p->devices, devs, (size_t)n * sizeof(ref_device));
        p->n_devices = n;
        group_partitions_by_dev_id(p, devs, n);
    }
    return err;
}

/* BestEffortPolicy.Init, besteffort_policy.go:70-86 */
int ref_policy_init(ref_policy *p, const ref_device *devs, int n, const char *topo_dir) {
    if (n > 0 && weights_alloc(p, devs, n) != 0) return REF_ERR_INIT_FAILED;
    int err = fetch_all_pair_weights(p, devs, n, topo_dir);
    return finish_init(p, devs, n, err);
}

int ref_policy_init_pairs(ref_policy *p, const ref_device *devs, int n, const int *from,
                          const int *to, const int *weight, int n_pairs) {
    if (n > 0 && weights_alloc(p, devs, n) != 0) return REF_ERR_INIT_FAILED;
    int err = n == 0 ? REF_ERR_EMPTY_DEVICES : REF_OK;
    if (n > 0)
        for (int i = 0; i < n_pairs; i++)
            if (from[i] >= 0 && to[i] >= 0 && from[i] < p->dim && to[i] < p->dim)
                weights_set(p, from[i], to[i], weight[i]);
    return finish_init(p, devs, n, err);
}

/* ---- candidate generation, device.go:254-442 --------------------------- */

/* DeviceSet, device.go:67-73 (LastIdx is dead and omitted) */
typedef struct device_set {
    int *ids;
    int size;
    int total_weight;
    int *parent_ids;
    int n_parents;
} device_set;

typedef struct set_list {
    device_set *v;
    size_t n, cap, head;
} set_list;

static void list_push(set_list *l, device_set s) {
    if (l->n == l->cap) {
        l->cap = l->cap ? l->cap * 2 : 256;
        l->v = (device_set *)realloc(l->v, l->cap * sizeof(device_set));
    }
    l->v[l->n++] = s;
}

static void list_free(set_list *l) {
    for (size_t i = 0; i < l->n; i++) {
        free(l->v[i].ids);
        free(l->v[i].parent_ids);
    }
    free(l->v);
}

static device_set set_new(const int *ids, int n_ids, const int *parents, int n_parents, int weight,
                          int cap) {
    device_set s;
    s.ids = (int *)malloc((size_t)(cap > 0 ? cap : 1) * sizeof(int));
    memcpy(s.ids, ids, (size_t)n_ids * sizeof(int));
    s.size = n_ids;
    s.total_weight = weight;
    s.parent_ids = (int *)malloc((size_t)(n_parents > 0 ? n_parents : 1) * sizeof(int));
    memcpy(s.parent_ids, parents, (size_t)n_parents * sizeof(int));
    s.n_parents = n_parents;
    return s;
}

/* addDeviceToSubsetAndUpdateWeight, device.go:254-273 (in place: the Go code
 * returns a fresh copy, callers here own `s` exclusively) */
static void set_add(const ref_policy *p, device_set *s, int dev_id) {
    int w = s->total_weight;
    for (int i = 0; i < s->size; i++) {
        int d = s->ids[i], from, to;
        if (d < dev_id) {
            from = d;
            to = dev_id;
        } else {
            from = dev_id;
            to = d;
        }
        w += ref_weights_get(p, from, to);
    }
    s->ids[s->size++] = dev_id;
    s->total_weight = w;
}

static int cmp_int(const void *a, const void *b) {
    int x = *(const int *)a, y = *(const int *)b;
    return (x > y) - (x < y);
}

/* sort.Slice comparator of device.go:342-349; `order` only breaks ties the
 * reference leaves to Go's map iteration + unstable sort. */
static int cmp_fpart(const void *a, const void *b) {
    const fpart *x = (const fpart *)a, *y = (const fpart *)b;
    if (x->n_ids != y->n_ids) return x->n_ids < y->n_ids ? -1 : 1;
    int c = strcmp(x->parent_id, y->parent_id);
    if (c) return c;
    return (x->order > y->order) - (x->order < y->order);
}

static int contains(const int *v, int n, int x) {
    for (int i = 0; i < n; i++)
        if (v[i] == x) return 1;
    return 0;
}

/* filterPartitions, device.go:310-351 */
fpart *ref_filter_partitions(const ref_policy *p, const int *avail_nodes, int n_avail,
                                const int *req_nodes, int n_req, int *n_out) {
    fpart *out = (fpart *)calloc((size_t)(p->n_parts > 0 ? p->n_parts : 1), sizeof(fpart));
    int n = 0;
    for (int g = 0; g < p->n_parts; g++) {
        const partitions *ps = &p->parts[g];
        int *ids = (int *)malloc((size_t)(ps->n_ids > 0 ? ps->n_ids : 1) * sizeof(int));
        int c = 0;
        for (int i = 0; i < ps->n_ids; i++) {
            int id = ps->ids[i];
            if (contains(req_nodes, n_req, id)) continue;
            if (contains(avail_nodes, n_avail, id)) ids[c++] = id;
        }
        if (c > 0) {
            qsort(ids, (size_t)c, sizeof(int), cmp_int);
            out[n].parent_id = ps->parent_id;
            out[n].ids = ids;
            out[n].n_ids = c;
            out[n].order = g;
            n++;
        } else {
            free(ids);
        }
    }
    qsort(out, (size_t)n, sizeof(fpart), cmp_fpart);
    *n_out = n;
    return out;
}

/* getCandidateDeviceSubsets, device.go:353-442.  avail_nodes must already be
 * sorted by NodeId (:362-364; irrelevant to the candidates, kept for form). */
static int get_candidate_device_subsets(const ref_policy *p, const int *avail_nodes, int n_avail,
                                        const int *req_nodes, int n_req, int size,
                                        set_list *final) {
    if (size <= 0) return REF_ERR_INVALID_SIZE;
    if (n_avail < size) return REF_ERR_INVALID_AVAILABLE;

    int n_parts;
    fpart *parts = ref_filter_partitions(p, avail_nodes, n_avail, req_nodes, n_req, &n_parts);
    int new_size = size - n_req;
    int cap = size + 1;
    set_list temp = {0};

    for (int idx = 0; idx < n_parts; idx++) { /* seeds, :375-402 */
        device_set s = set_new(&parts[idx].ids[0], 1, &idx, 1, 0, cap);
        if (new_size == 1) {
            for (int r = 0; r < n_req; r++) set_add(p, &s, req_nodes[r]);
            list_push(final, s);
            continue;
        }
        int fulfilled = 0;
        for (int i = 1; i < parts[idx].n_ids; i++) {
            set_add(p, &s, parts[idx].ids[i]);
            if (i == new_size - 1) {
                fulfilled = 1;
                break;
            }
        }
        if (fulfilled) {
            for (int r = 0; r < n_req; r++) set_add(p, &s, req_nodes[r]);
            list_push(final, s);
        } else {
            list_push(&temp, s);
        }
    }

    while (temp.head < temp.n) { /* FIFO BFS, :405-440 */
        device_set cur = temp.v[temp.head++];
        if (cur.n_parents == n_parts) continue;
        for (int idx = 0; idx < n_parts; idx++) {
            if (contains(cur.parent_ids, cur.n_parents, idx)) continue;
            int *parents = (int *)malloc((size_t)(cur.n_parents + 1) * sizeof(int));
            memcpy(parents, cur.parent_ids, (size_t)cur.n_parents * sizeof(int));
            parents[cur.n_parents] = idx;
            device_set s = set_new(cur.ids, cur.size, parents, cur.n_parents + 1, cur.total_weight, cap);
            free(parents);
            int done = 0;
            for (int i = 0; i < parts[idx].n_ids; i++) {
                set_add(p, &s, parts[idx].ids[i]);
                if (s.size == new_size) {
                    for (int r = 0; r < n_req; r++) set_add(p, &s, req_nodes[r]);
                    list_push(final, s);
                    done = 1;
                    break;
                }
            }
            if (!done) { /* devset.Size < newSize, :436-438 */
                list_push(&temp, s);
                cur = temp.v[temp.head - 1]; /* realloc may have moved the array */
            }
        }
    }
    list_free(&temp);
    for (int i = 0; i < n_parts; i++) free(parts[i].ids);
    free(parts);
    return REF_OK;
}

/* ---- Allocate, besteffort_policy.go:88-151 ----------------------------- */

const ref_device *ref_device_by_id(const ref_policy *p, const char *id) {
    /* devicesMap: later duplicates overwrite earlier ones (:77-79) */
    const ref_device *hit = NULL;
    for (int i = 0; i < p->n_devices; i++)
        if (strcmp(p->devices[i].id, id) == 0) hit = &p->devices[i];
    return hit;
}

static void copy_ids(char *out, const char *const *ids, int n) {
    for (int i = 0; i < n; i++) snprintf(out + (size_t)i * REF_ID_LEN, REF_ID_LEN, "%s", ids[i]);
}

/* setContainsAll, device.go:88-105 */
static int set_contains_all(const char *const *set, int n_set, const char *const *sub, int n_sub) {
    if (n_sub > n_set) return 0;
    for (int i = 0; i < n_sub; i++) {
        int found = 0;
        for (int j = 0; j < n_set; j++)
            if (strcmp(set[j], sub[i]) == 0) {
                found = 1;
                break;
            }
        if (!found) return 0;
    }
    return 1;
}

typedef struct node_dev {
    int node_id;
    const ref_device *dev;
} node_dev;

static int cmp_node_dev(const void *a, const void *b) {
    const node_dev *x = (const node_dev *)a, *y = (const node_dev *)b;
    return (x->node_id > y->node_id) - (x->node_id < y->node_id);
}

int ref_policy_allocate(ref_policy *p, const char *const *avail, int n_avail,
                        const char *const *req, int n_req, int size, char *out, ref_result *res) {
    if (res) memset(res, 0, sizeof *res);
    if (size <= 0) return REF_ERR_INVALID_SIZE;               /* :90-92 */
    if (n_avail < size) return REF_ERR_INVALID_AVAILABLE;     /* :94-96 */
    if (n_req > size) return REF_ERR_INVALID_REQUIRED;        /* :98-100 */
    if (n_req > n_avail) return REF_ERR_INVALID_REQ_AVAILABLE; /* :102-104 */
    if (p->n_devices == 0) return REF_ERR_INVALID_INIT;       /* :106-108 */
    if (n_avail == size) {                                    /* :110-112 returns availableIds itself */
        copy_ids(out, avail, n_avail);
        return size;
    }
    if (n_req == size) { /* :114-116 */
        copy_ids(out, req, n_req);
        return size;
    }
    if (ref_policy_weight_rows(p) == 0) return REF_ERR_INVALID_INIT;              /* :118-120 */
    if (!set_contains_all(avail, n_avail, req, n_req)) return REF_ERR_NO_CANDIDATE; /* :122-124 */

    node_dev *available = (node_dev *)malloc((size_t)n_avail * sizeof(node_dev));
    int *avail_nodes = (int *)malloc((size_t)n_avail * sizeof(int));
    int *req_nodes = (int *)malloc((size_t)(n_req > 0 ? n_req : 1) * sizeof(int));
    int rc = REF_OK;
    for (int i = 0; i < n_avail; i++) { /* getDevicesFromIds, :61-68 */
        const ref_device *d = ref_device_by_id(p, avail[i]);
        if (!d) {
            rc = REF_ERR_PANIC; /* nil entry dereferenced at device.go:362-364 */
            goto out;
        }
        available[i].node_id = d->node_id;
        available[i].dev = d;
    }
    for (int i = 0; i < n_req; i++) {
        const ref_device *d = ref_device_by_id(p, req[i]);
        if (!d) {
            rc = REF_ERR_PANIC;
            goto out;
        }
        req_nodes[i] = d->node_id;
    }
    qsort(available, (size_t)n_avail, sizeof(node_dev), cmp_node_dev); /* device.go:362-364 */
    for (int i = 0; i < n_avail; i++) avail_nodes[i] = available[i].node_id;

    set_list final = {0};
    rc = get_candidate_device_subsets(p, avail_nodes, n_avail, req_nodes, n_req, size, &final);
    if (rc == REF_OK) {
        int best_score = 2147483647; /* math.MaxInt32, :133 */
        const device_set *candidate = NULL;
        for (size_t i = 0; i < final.n; i++) { /* strict <, first wins, :135-140 */
            if (final.v[i].total_weight < best_score) {
                candidate = &final.v[i];
                best_score = final.v[i].total_weight;
            }
        }
        if (!candidate) {
            rc = REF_ERR_PANIC; /* candidate.Ids on nil, :141 */
        } else {
            int n_out = 0;
            for (int i = 0; i < candidate->size; i++) /* :141-148 */
                for (int j = 0; j < n_avail; j++)
                    if (available[j].node_id == candidate->ids[i]) {
                        snprintf(out + (size_t)n_out * REF_ID_LEN, REF_ID_LEN, "%s", available[j].dev->id);
                        n_out++;
                        break;
                    }
            if (res) {
                res->score = candidate->total_weight;
                res->n_candidates = (int)final.n;
                res->searched = 1;
            }
            rc = n_out;
        }
    }
    list_free(&final);
out:
    free(available);
    free(avail_nodes);
    free(req_nodes);
    return rc;
}
