/*
 * oracle/exhaustive.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see ref_policy.h).
 *
 * CPU statement of the rule the GPU path implements:
 *
 *     argmin over S in Family(avail, req, k) of (Score(S), RefKey(S))
 *
 * evaluated by scoring EVERY k'-subset of the free devices.  It exists to
 * show, on CPU and before any CUDA runs, that "exhaustive + Family + RefKey"
 * returns exactly what the reference's BFS (ref_policy.c, device.go:353-442,
 * besteffort_policy.go:133-148) returns, and to serve as the timed CPU
 * baseline for the metric "subsets scored/sec" (OpenMP over rank ranges).
 *
 *   Family : per same-GPU partition group (filterPartitions order,
 *            device.go:310-351) a set takes nothing, the whole group, or --
 *            for at most one group -- its lowest-NodeId prefix.  These are
 *            the sets the seed loop and the FIFO BFS can emit (:375-440).
 *   RefKey : (m, sigma); m = groups touched, sigma = positions of the whole
 *            groups ascending, then the partial group.  This is the order in
 *            which the BFS appends to subsetsFinal, and the strict `<` at
 *            besteffort_policy.go:136 keeps the first minimum.
 *
 * Enumeration is lexicographic over the canonical device order (groups in
 * filterPartitions order, NodeId ascending inside); the score of the last
 * position comes from a per-level vector V[l][x] = u[x] + sum_{j<l} W[c_j][x],
 * so the innermost loop is one load, one add, one compare per subset.
 */
#define _GNU_SOURCE
#include "ref_internal.h"

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define EXH_MAX_N 256
#define EXH_MAX_K 32

typedef struct ref_exh_result {
    int score;
    int m;                       /* groups touched by the winner */
    unsigned long long scored;   /* subsets scored */
    unsigned long long total;    /* C(n_free, k') */
    unsigned long long slow;     /* Family/RefKey evaluations */
    int n_free, k_free, n_groups;
    int threads;
    /* The winner in the product's wire format (tp_key, include/tp_allocator.h), restated here
     * from that header's description: hi = 0:1 | score:23 | groups:8 | first 32 bits of the group
     * sequence, lo = the remaining 64 bits; digits are ceil(log2(groups)) bits, first digit
     * most significant, in a 96-bit field.  Lets CPU-only tests exercise the multi-rank
     * min-reduce + tp_decode path with keys produced by the oracle. */
    unsigned long long key_hi, key_lo;
} ref_exh_result;

typedef struct exh_key {
    int score;
    int m;
    int sigma[EXH_MAX_K];
    int partial_count; /* devices taken from the partial group (0 = none) */
    int valid;
} exh_key;

typedef struct exh_ctx {
    int n, k;                 /* free devices, k' */
    int w[EXH_MAX_N][EXH_MAX_N];
    int u[EXH_MAX_N];         /* sum of weights to the required devices */
    int c_req;                /* weight among required devices */
    int group_of[EXH_MAX_N];  /* canonical index -> group position */
    int g_start[EXH_MAX_N], g_size[EXH_MAX_N];
    int n_groups;
    unsigned long long binom[EXH_MAX_N + 1][EXH_MAX_K + 1];
} exh_ctx;

static int key_less(const exh_key *a, const exh_key *b) {
    if (!b->valid) return a->valid;
    if (!a->valid) return 0;
    if (a->score != b->score) return a->score < b->score;
    if (a->m != b->m) return a->m < b->m;
    for (int i = 0; i < a->m; i++)
        if (a->sigma[i] != b->sigma[i]) return a->sigma[i] < b->sigma[i];
    return 0;
}

/* Family membership and RefKey of the sorted canonical subset c[0..k). */
static int family_key(const exh_ctx *x, const int *c, int score, exh_key *out) {
    int n_whole = 0, partial = -1, partial_cnt = 0, m = 0;
    int i = 0;
    while (i < x->k) {
        int g = x->group_of[c[i]];
        int j = i;
        while (j < x->k && x->group_of[c[j]] == g) j++;
        int cnt = j - i;
        m++;
        if (cnt == x->g_size[g]) {
            out->sigma[n_whole++] = g;
        } else {
            if (partial >= 0) return 0; /* two partial groups: not emitted by the BFS */
            for (int t = 0; t < cnt; t++)
                if (c[i + t] != x->g_start[g] + t) return 0; /* not the lowest-NodeId prefix */
            partial = g;
            partial_cnt = cnt;
        }
        i = j;
    }
    if (partial >= 0) out->sigma[n_whole] = partial;
    out->m = m;
    out->score = score;
    out->partial_count = partial_cnt;
    out->valid = 1;
    return 1;
}

static void unrank_lex(const exh_ctx *x, unsigned long long r, int *c) {
    int a = 0;
    for (int i = 0; i < x->k; i++) {
        for (;; a++) {
            unsigned long long cnt = x->binom[x->n - a - 1][x->k - i - 1];
            if (r < cnt) break;
            r -= cnt;
        }
        c[i] = a++;
    }
}

typedef struct exh_stats {
    unsigned long long scored, slow;
} exh_stats;

/* Score ranks [lo, hi) of the lexicographic order. */
static void enum_range(const exh_ctx *x, unsigned long long lo, unsigned long long hi,
                       exh_key *best, exh_stats *st) {
    const int n = x->n, k = x->k;
    int c[EXH_MAX_K];
    int ps[EXH_MAX_K + 1];
    static _Thread_local int V[EXH_MAX_K][EXH_MAX_N];
    unrank_lex(x, lo, c);
    unsigned long long remaining = hi - lo;
    int level = 0; /* lowest level whose V/ps must be rebuilt */
    ps[0] = x->c_req;
    for (int i = 0; i < n; i++) V[0][i] = x->u[i];
    int first_x = c[k - 1];
    while (remaining) {
        for (int l = level; l < k - 1; l++) {
            ps[l + 1] = ps[l] + V[l][c[l]];
            const int *wr = x->w[c[l]];
            for (int i = c[l] + 1; i < n; i++) V[l + 1][i] = V[l][i] + wr[i];
        }
        const int base = ps[k - 1];
        const int *v = V[k - 1];
        int xi = first_x;
        for (; xi < n && remaining; xi++, remaining--) {
            int s = base + v[xi];
            st->scored++;
            if (!best->valid || s <= best->score) {
                exh_key cand;
                c[k - 1] = xi;
                st->slow++;
                if (family_key(x, c, s, &cand) && key_less(&cand, best)) *best = cand;
            }
        }
        if (!remaining) break;
        int j = k - 2; /* lexicographic successor of the prefix */
        while (j >= 0 && c[j] == n - k + j) j--;
        if (j < 0) break;
        c[j]++;
        for (int t = j + 1; t < k; t++) c[t] = c[t - 1] + 1;
        first_x = c[k - 1];
        level = j;
    }
}

/* Build the canonical order / weights for one request.  Returns 0 or a REF_ERR. */
static int build_ctx(const ref_policy *p, const int *avail_nodes, int n_avail, const int *req_nodes,
                     int n_req, int size, exh_ctx *x, int *canon_nodes) {
    int n_parts;
    fpart *parts = ref_filter_partitions(p, avail_nodes, n_avail, req_nodes, n_req, &n_parts);
    int n = 0;
    for (int g = 0; g < n_parts; g++) {
        x->g_start[g] = n;
        x->g_size[g] = parts[g].n_ids;
        for (int i = 0; i < parts[g].n_ids; i++) {
            if (n >= EXH_MAX_N) return REF_ERR_INVALID_AVAILABLE;
            x->group_of[n] = g;
            canon_nodes[n++] = parts[g].ids[i];
        }
        free(parts[g].ids);
    }
    free(parts);
    x->n_groups = n_parts;
    x->n = n;
    x->k = size - n_req;
    if (x->k > EXH_MAX_K || x->k < 1 || x->k > n) return REF_ERR_PANIC;
    for (int i = 0; i < n; i++) {
        for (int j = 0; j < n; j++) {
            int a = canon_nodes[i], b = canon_nodes[j];
            x->w[i][j] = i == j ? 0 : ref_weights_get(p, a < b ? a : b, a < b ? b : a);
        }
        int u = 0;
        for (int r = 0; r < n_req; r++) {
            int a = canon_nodes[i], b = req_nodes[r];
            u += ref_weights_get(p, a < b ? a : b, a < b ? b : a);
        }
        x->u[i] = u;
    }
    int c_req = 0; /* required are added one after another: every req-req pair counts (device.go:254-273) */
    for (int r = 0; r < n_req; r++)
        for (int q = 0; q < r; q++) {
            int a = req_nodes[q], b = req_nodes[r];
            c_req += ref_weights_get(p, a < b ? a : b, a < b ? b : a); /* d < devId ? (d,devId) : (devId,d) */
        }
    x->c_req = c_req;
    for (int i = 0; i <= n; i++)
        for (int j = 0; j <= EXH_MAX_K; j++) {
            if (j == 0)
                x->binom[i][j] = 1;
            else if (i == 0)
                x->binom[i][j] = 0;
            else {
                unsigned long long s = x->binom[i - 1][j - 1] + x->binom[i - 1][j];
                if (s < x->binom[i - 1][j]) s = ~0ULL; /* saturate */
                x->binom[i][j] = s;
            }
        }
    return 0;
}

/*
 * Same contract as ref_policy_allocate, answered by exhaustive search.
 * rank_lo/rank_hi bound the lexicographic rank range that is scored
 * (rank_hi == 0: everything); a partial range is only meaningful for timing.
 * n_threads <= 0 uses all OpenMP threads.
 */
static int exhaustive_impl(const ref_policy *p, const char *const *avail, int n_avail,
                           const char *const *req, int n_req, int size, char *out,
                           ref_exh_result *res, int n_threads, unsigned long long rank_lo,
                           unsigned long long rank_hi, int unconstrained) {
    if (res) memset(res, 0, sizeof *res);
    if (size <= 0) return REF_ERR_INVALID_SIZE;
    if (n_avail < size) return REF_ERR_INVALID_AVAILABLE;
    if (n_req > size) return REF_ERR_INVALID_REQUIRED;
    if (n_req > n_avail) return REF_ERR_INVALID_REQ_AVAILABLE;
    if (p->n_devices == 0) return REF_ERR_INVALID_INIT;
    if (n_avail == size) {
        for (int i = 0; i < size; i++) snprintf(out + (size_t)i * REF_ID_LEN, REF_ID_LEN, "%s", avail[i]);
        return size;
    }
    if (n_req == size) {
        for (int i = 0; i < size; i++) snprintf(out + (size_t)i * REF_ID_LEN, REF_ID_LEN, "%s", req[i]);
        return size;
    }
    if (ref_policy_weight_rows(p) == 0) return REF_ERR_INVALID_INIT;
    for (int i = 0; i < n_req; i++) {
        int found = 0;
        for (int j = 0; j < n_avail; j++)
            if (strcmp(avail[j], req[i]) == 0) found = 1;
        if (!found) return REF_ERR_NO_CANDIDATE;
    }
    int *avail_nodes = (int *)malloc(sizeof(int) * (size_t)n_avail);
    int *req_nodes = (int *)malloc(sizeof(int) * (size_t)(n_req + 1));
    int *canon = (int *)malloc(sizeof(int) * EXH_MAX_N);
    exh_ctx *x = (exh_ctx *)calloc(1, sizeof(exh_ctx));
    int rc = 0;
    for (int i = 0; i < n_avail && !rc; i++) {
        const ref_device *d = ref_device_by_id(p, avail[i]);
        if (!d) rc = REF_ERR_PANIC; else avail_nodes[i] = d->node_id;
    }
    for (int i = 0; i < n_req && !rc; i++) {
        const ref_device *d = ref_device_by_id(p, req[i]);
        if (!d) rc = REF_ERR_PANIC; else req_nodes[i] = d->node_id;
    }
    if (!rc) rc = build_ctx(p, avail_nodes, n_avail, req_nodes, n_req, size, x, canon);
    if (rc) goto done;
    if (unconstrained) { /* every free device its own group: Family = all subsets */
        for (int i = 0; i < x->n; i++) {
            x->group_of[i] = i;
            x->g_start[i] = i;
            x->g_size[i] = 1;
        }
        x->n_groups = x->n;
    }

    unsigned long long total = x->binom[x->n][x->k];
    if (rank_hi == 0 || rank_hi > total) rank_hi = total;
    if (rank_lo > rank_hi) rank_lo = rank_hi;
    unsigned long long span = rank_hi - rank_lo;
    unsigned long long chunk = 1ULL << 20;
    unsigned long long n_chunks = (span + chunk - 1) / chunk;
    /* Seed: the first k' devices of the canonical order (lexicographic rank 0) are
     * always a Family member (whole leading groups + a prefix of the next one), so
     * its score bounds the answer from above and lets the scan skip the Family /
     * RefKey evaluation for every subset that scores worse. */
    exh_key best;
    memset(&best, 0, sizeof best);
    {
        int c0[EXH_MAX_K], s0 = x->c_req;
        for (int i = 0; i < x->k; i++) {
            c0[i] = i;
            s0 += x->u[i];
            for (int j = 0; j < i; j++) s0 += x->w[j][i];
        }
        family_key(x, c0, s0, &best);
    }
    const exh_key seed = best;
    unsigned long long scored = 0, slow = 0;
    int threads = 1;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
    threads = n_threads > 0 ? n_threads : omp_get_max_threads();
#else
    (void)n_threads;
#endif
#pragma omp parallel
    {
        exh_key local = seed;
        exh_stats st = {0, 0};
#pragma omp for schedule(dynamic, 1) nowait
        for (long long ci = 0; ci < (long long)n_chunks; ci++) {
            unsigned long long lo = rank_lo + (unsigned long long)ci * chunk;
            unsigned long long hi = lo + chunk < rank_hi ? lo + chunk : rank_hi;
            enum_range(x, lo, hi, &local, &st);
        }
#pragma omp critical
        {
            if (key_less(&local, &best)) best = local;
            scored += st.scored;
            slow += st.slow;
        }
    }
    if (res) {
        res->scored = scored;
        res->total = total;
        res->slow = slow;
        res->n_free = x->n;
        res->k_free = x->k;
        res->n_groups = x->n_groups;
        res->threads = threads;
    }
    if (!best.valid) {
        rc = REF_ERR_PANIC;
        goto done;
    }
    if (res) {
        res->score = best.score;
        res->m = best.m;
        int dbits = 1;
        while ((1 << dbits) < x->n_groups) dbits++;
        unsigned __int128 field = 0;
        for (int i = 0; i < best.m; i++)
            field |= (unsigned __int128)(unsigned)best.sigma[i] << (96 - (i + 1) * dbits);
        res->key_hi = ((unsigned long long)best.score << 40) | ((unsigned long long)best.m << 32) |
                      (unsigned long long)(field >> 64);
        res->key_lo = (unsigned long long)field;
    }
    /* Output order of besteffort_policy.go:141-148: candidate order = group
     * sequence, ascending NodeId inside a group, required last in request order. */
    int n_out = 0;
    for (int i = 0; i < best.m; i++) {
        int g = best.sigma[i];
        int take = (best.partial_count && i == best.m - 1) ? best.partial_count : x->g_size[g];
        for (int t = 0; t < take; t++) {
            int node = canon[x->g_start[g] + t];
            for (int d = 0; d < p->n_devices; d++)
                if (p->devices[d].node_id == node) {
                    snprintf(out + (size_t)n_out * REF_ID_LEN, REF_ID_LEN, "%s", p->devices[d].id);
                    break;
                }
            n_out++;
        }
    }
    for (int r = 0; r < n_req; r++) snprintf(out + (size_t)(n_out++) * REF_ID_LEN, REF_ID_LEN, "%s", req[r]);
    rc = n_out;
done:
    free(avail_nodes);
    free(req_nodes);
    free(canon);
    free(x);
    return rc;
}

int ref_exhaustive_allocate(const ref_policy *p, const char *const *avail, int n_avail,
                            const char *const *req, int n_req, int size, char *out,
                            ref_exh_result *res, int n_threads, unsigned long long rank_lo,
                            unsigned long long rank_hi) {
    return exhaustive_impl(p, avail, n_avail, req, n_req, size, out, res, n_threads, rank_lo, rank_hi, 0);
}

/*
 * The unconstrained optimum (the advisory mode): the same
 * enumeration with the Family test dropped, i.e. the minimum of Score over ALL k'-subsets of
 * the free devices; ties go to the lexicographically first subset in canonical order.  The
 * reference has no such function: this is the checker of tp_global_optimum, and its score is
 * what the reference's pick (besteffort_policy.go:133-140) is compared with.
 */
int ref_exhaustive_global_optimum(const ref_policy *p, const char *const *avail, int n_avail,
                                  const char *const *req, int n_req, int size, char *out,
                                  ref_exh_result *res, int n_threads) {
    return exhaustive_impl(p, avail, n_avail, req, n_req, size, out, res, n_threads, 0, 0, 1);
}
