// abi_harness.cpp -- the call sequence of the cgo policy in INTEGRATION.md, replayed natively.
//
// The reference's policy object is created once (manager.go:101), initialised once (amdgpu.go:112) and then
// called from one goroutine per RPC (dpm/plugin.go:108-111, amdgpu.go:300-319).  No Go toolchain exists in this
// image, so this C++ program issues the same sequence against the C ABI:
//
//     tp_create -> N sequential tp_allocate -> T threads x M concurrent tp_allocate
//               -> B threads x R concurrent tp_allocate_batch calls of 64 requests each -> tp_destroy
//
// and times every call with clock_gettime(CLOCK_MONOTONIC) around the ABI call itself (SURVEY 8(d) latency
// protocol): no interpreter, no ctypes.  Beside it: the same request through the string-level policy class
// (what the Go Allocate adds: id -> index maps) and through the CPU restatement of the reference's BFS
// (oracle/, linked here as the checker and CPU baseline -- test infrastructure, not product).
// Every answer is compared with the oracle's.  Built by `make -C tests/harness` (also with -fsanitize=thread).
//
// usage: abi_harness <input file> [sequential calls] [threads] [calls per thread] [batch threads] [batches per thread]
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <time.h>
#include <vector>

#include "besteffort_policy.h"
#include "ref_policy.h"
#include "tp_allocator.h"

static double now_us() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}

struct Stats {
    double p50 = 0, p99 = 0, mean = 0, max = 0;
    size_t n = 0;
};
static Stats stats(std::vector<double> v) {
    Stats s;
    if (v.empty()) return s;
    std::sort(v.begin(), v.end());
    s.n = v.size();
    s.p50 = v[v.size() / 2];
    s.p99 = v[std::min(v.size() - 1, (size_t)(v.size() * 0.99))];
    s.max = v.back();
    for (double x : v) s.mean += x;
    s.mean /= v.size();
    return s;
}
static void print_stats(const char *name, const Stats &s, const char *tail = "") {
    printf("\"%s\": {\"calls\": %zu, \"p50_us\": %.2f, \"p99_us\": %.2f, \"mean_us\": %.2f, \"max_us\": %.2f%s}", name, s.n,
           s.p50, s.p99, s.mean, s.max, tail);
}

struct Input {
    std::string name;
    std::vector<tp::Device> devs;
    tp::PairWeights weights;
    std::vector<int> from, to, w;
    std::vector<std::string> avail, req;
    int size = 0;
};

// config <name> / devices <n> + n lines "id node numa devid" / pairs <m> + m lines "from to w" /
// request <n_avail> <n_req> <size> + one line of available ids + one line of required ids
static bool read_input(const char *path, Input &in) {
    std::ifstream f(path);
    std::string tok;
    int n = 0, m = 0, na = 0, nr = 0;
    if (!(f >> tok >> in.name) || tok != "config") return false;
    if (!(f >> tok >> n) || tok != "devices") return false;
    in.devs.resize(n);
    for (auto &d : in.devs) f >> d.Id >> d.NodeId >> d.NumaNode >> d.DevId;
    if (!(f >> tok >> m) || tok != "pairs") return false;
    in.from.resize(m), in.to.resize(m), in.w.resize(m);
    for (int i = 0; i < m; i++) {
        f >> in.from[i] >> in.to[i] >> in.w[i];
        in.weights[in.from[i]][in.to[i]] = in.w[i];
    }
    if (!(f >> tok >> na >> nr >> in.size) || tok != "request") return false;
    in.avail.resize(na), in.req.resize(nr);
    for (auto &s : in.avail) f >> s;
    for (auto &s : in.req) f >> s;
    return (bool)f;
}

int main(int argc, char **argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: %s <input> [sequential calls] [threads] [calls per thread] [batch threads] [batches per thread]\n", argv[0]);
        return 2;
    }
    Input in;
    if (!read_input(argv[1], in)) {
        fprintf(stderr, "bad input file %s\n", argv[1]);
        return 2;
    }
    const int n_seq = argc > 2 ? atoi(argv[2]) : 10000, n_thr = argc > 3 ? atoi(argv[3]) : 8,
              n_per = argc > 4 ? atoi(argv[4]) : 2000, n_bthr = argc > 5 ? atoi(argv[5]) : 3,
              n_bper = argc > 6 ? atoi(argv[6]) : 20;

    // ---- Init (tp_create through the policy class, as the Go Init would) -----------------------------
    tp::BestEffortPolicy pol;
    const double t_init = now_us();
    tp::Error err = pol.InitWithWeights(in.devs, in.weights);
    const double init_us = now_us() - t_init;
    if (!err.empty()) {
        fprintf(stderr, "Init failed: %s\n", err.c_str());
        return 1;
    }
    tp_policy *h = pol.Handle();
    std::vector<int32_t> av, rq;
    for (auto &s : in.avail) av.push_back(pol.IndexOf(s));
    for (auto &s : in.req) rq.push_back(pol.IndexOf(s));

    // ---- the checker: the reference's BFS, restated (oracle/ref_policy.c) ------------------------------
    ref_policy *ref = ref_policy_new();
    std::vector<ref_device> rdevs(in.devs.size());
    for (size_t i = 0; i < in.devs.size(); i++) {
        memset(&rdevs[i], 0, sizeof(ref_device));
        snprintf(rdevs[i].id, REF_ID_LEN, "%s", in.devs[i].Id.c_str());
        snprintf(rdevs[i].dev_id, REF_ID_LEN, "%s", in.devs[i].DevId.c_str());
        rdevs[i].node_id = in.devs[i].NodeId;
        rdevs[i].numa_node = in.devs[i].NumaNode;
    }
    if (ref_policy_init_pairs(ref, rdevs.data(), (int)rdevs.size(), in.from.data(), in.to.data(), in.w.data(),
                              (int)in.from.size()) < 0) {
        fprintf(stderr, "oracle init failed\n");
        return 1;
    }
    std::vector<const char *> cav, crq;
    for (auto &s : in.avail) cav.push_back(s.c_str());
    for (auto &s : in.req) crq.push_back(s.c_str());
    std::vector<char> ref_out((size_t)REF_ID_LEN * std::max(in.size, 1));
    ref_result rres{};
    std::vector<double> cpu;
    int ref_n = 0;
    for (int i = 0; i < std::min(n_seq, 3000); i++) {
        const double t = now_us();
        ref_n = ref_policy_allocate(ref, cav.data(), (int)cav.size(), crq.data(), (int)crq.size(), in.size, ref_out.data(), &rres);
        cpu.push_back(now_us() - t);
    }
    std::vector<std::string> want;
    for (int i = 0; i < ref_n; i++) want.emplace_back(ref_out.data() + (size_t)i * REF_ID_LEN);

    // ---- sequential tp_allocate (the cgo call) ------------------------------------------------------
    std::vector<int32_t> out(std::max(in.size, 1));
    tp_result info{};
    for (int i = 0; i < 200; i++) tp_allocate(h, av.data(), (int)av.size(), rq.data(), (int)rq.size(), in.size, out.data(), &info);
    std::vector<double> seq;
    int bad = 0;
    for (int i = 0; i < n_seq; i++) {
        const double t = now_us();
        const int rc = tp_allocate(h, av.data(), (int)av.size(), rq.data(), (int)rq.size(), in.size, out.data(), &info);
        seq.push_back(now_us() - t);
        if (rc != ref_n) bad++;
        else
            for (int j = 0; j < rc; j++) bad += in.devs[out[j]].Id != want[j];
    }
    // ---- the string-level policy (what Go's Allocate does around the cgo call) --------------------------
    std::vector<double> strs;
    for (int i = 0; i < std::min(n_seq, 3000); i++) {
        const double t = now_us();
        auto res = pol.Allocate(in.avail, in.req, in.size);
        strs.push_back(now_us() - t);
        bad += res.first != want;
    }
    // ---- concurrent callers on one policy (one goroutine per RPC) ---------------------------------------
    std::vector<std::vector<double>> per(n_thr);
    std::atomic<int> bad_conc{0};
    const double t_conc = now_us();
    {
        std::vector<std::thread> ths;
        for (int t = 0; t < n_thr; t++)
            ths.emplace_back([&, t] {
                std::vector<int32_t> o(std::max(in.size, 1));
                for (int i = 0; i < n_per; i++) {
                    const double t0 = now_us();
                    const int rc = tp_allocate(h, av.data(), (int)av.size(), rq.data(), (int)rq.size(), in.size, o.data(), nullptr);
                    per[t].push_back(now_us() - t0);
                    if (rc != ref_n) bad_conc++;
                    else
                        for (int j = 0; j < rc; j++) bad_conc += in.devs[o[j]].Id != want[j];
                }
            });
        for (auto &th : ths) th.join();
    }
    const double conc_us = now_us() - t_conc;
    std::vector<double> conc;
    for (auto &v : per) conc.insert(conc.end(), v.begin(), v.end());

    // ---- concurrent tp_allocate_batch calls on the same policy (two run at once, the others wait for a slot) ----
    std::atomic<int> bad_batch{0};
    std::vector<std::vector<double>> bper(n_bthr);
    {
        const int n_in_batch = 64;
        std::vector<std::thread> ths;
        for (int t = 0; t < n_bthr; t++)
            ths.emplace_back([&, t] {
                std::vector<tp_request> reqs(n_in_batch);
                for (auto &r : reqs) {
                    r.available = av.data();
                    r.n_available = (int32_t)av.size();
                    r.required = rq.data();
                    r.n_required = (int32_t)rq.size();
                    r.size = in.size;
                }
                const int stride = std::max(in.size, 1);
                std::vector<int32_t> o((size_t)n_in_batch * stride), st(n_in_batch);
                for (int i = 0; i < n_bper; i++) {
                    const double t0 = now_us();
                    const int rc = tp_allocate_batch(h, n_in_batch, reqs.data(), o.data(), stride, st.data(), nullptr);
                    bper[t].push_back(now_us() - t0);
                    if (rc != 0) bad_batch++;
                    for (int r = 0; r < n_in_batch; r++) {
                        if (st[r] != ref_n) bad_batch++;
                        else
                            for (int j = 0; j < ref_n; j++) bad_batch += in.devs[o[(size_t)r * stride + j]].Id != want[j];
                    }
                }
            });
        for (auto &th : ths) th.join();
    }
    std::vector<double> bat;
    for (auto &v : bper) bat.insert(bat.end(), v.begin(), v.end());

    char tail[160];
    printf("{\"config\": \"%s\", \"devices\": %zu, \"available\": %zu, \"required\": %zu, \"size\": %d, \"subsets\": %llu, "
           "\"kernel\": %d, \"score\": %d, \"cpu_candidates\": %d, \"init_us\": %.0f, \"mismatches\": %d, ",
           in.name.c_str(), in.devs.size(), in.avail.size(), in.req.size(), in.size,
           (unsigned long long)info.subsets_total, info.kernel, info.score, rres.n_candidates, init_us,
           bad + bad_conc.load() + bad_batch.load());
    print_stats("c_abi_sequential", stats(seq));
    printf(", ");
    print_stats("policy_strings_sequential", stats(strs));
    printf(", ");
    snprintf(tail, sizeof tail, ", \"threads\": %d, \"calls_per_s\": %.0f", n_thr, conc.size() / (conc_us * 1e-6));
    print_stats("c_abi_concurrent", stats(conc), tail);
    printf(", ");
    if (!bat.empty()) {
        snprintf(tail, sizeof tail, ", \"threads\": %d, \"requests_per_batch\": 64", n_bthr);
        print_stats("c_abi_batch_concurrent", stats(bat), tail);
        printf(", ");
    }
    print_stats("cpu_reference_bfs", stats(cpu));
    printf(", \"clock\": \"clock_gettime(CLOCK_MONOTONIC) around each call, native caller\"}\n");
    ref_policy_free(ref);
    return bad + bad_conc.load() + bad_batch.load() ? 1 : 0;
}
