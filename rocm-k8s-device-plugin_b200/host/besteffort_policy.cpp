// besteffort_policy.cpp -- see besteffort_policy.h.  Host logic only: sysfs parse, id<->index
// maps, the reference's validation order and error strings.  The search is tp_allocate (GPU).
#include "besteffort_policy.h"

#include <algorithm>
#include <cstring>
#include <dirent.h>
#include <fstream>
#include <regex>
#include <set>

namespace tp {

namespace {

const char *kTopoRootPath = "/sys/class/kfd/kfd/topology/nodes"; // device.go:33-35

// besteffort_policy.go:36-43
const char *kInvalidSize = "allocation size can not be negative";
const char *kInvalidAvailable = "available devices count less than allocation size";
const char *kInvalidRequired = "must_include devices size is more than allocation size";
const char *kInvalidReqAvailable = "must_include length should be less than or equal to avilable device size";
const char *kInvalidInit = "Init method must be called before Allocate";
const char *kNoCandidateFound = "No candidate subset found with matching criteria";

// filepath.Glob(dir/"[0-9]*"): entries whose name starts with a digit, sorted bytewise;
// a directory that cannot be read yields nothing and no error.
std::vector<std::string> globDigits(const std::string &dir) {
    std::vector<std::string> names;
    if (DIR *d = opendir(dir.c_str())) {
        while (dirent *e = readdir(d))
            if (e->d_name[0] >= '0' && e->d_name[0] <= '9') names.emplace_back(e->d_name);
        closedir(d);
    }
    std::sort(names.begin(), names.end());
    return names;
}

// strconv.ParseInt(s, 0, 32) on the digits the regex captured: a leading 0 means octal.
bool parseIntBase0(const std::string &s, int &out) {
    int base = 10;
    size_t i = 0;
    if (s.size() > 1 && s[0] == '0') {
        base = 8;
        i = 1;
    }
    long long v = 0;
    for (; i < s.size(); i++) {
        int d = s[i] - '0';
        if (d < 0 || d >= base) return false;
        v = v * base + d;
        if (v > 2147483647LL) return false;
    }
    out = (int)v;
    return true;
}

enum class Fetch { Ok, OpenError, ParseError };

// fetchTopoProperties (device.go:107-133): every regex against every line, last match wins.
Fetch fetchTopoProperties(const std::string &path, const std::vector<std::regex> &re, std::vector<int> &res) {
    res.assign(re.size(), 0);
    std::ifstream f(path);
    if (!f) return Fetch::OpenError;
    std::string line;
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        for (size_t i = 0; i < re.size(); i++) {
            std::smatch m;
            if (!std::regex_search(line, m, re[i])) continue;
            int v;
            if (!parseIntBase0(m[1].str(), v)) return Fetch::ParseError;
            res[i] = v;
        }
    }
    return Fetch::Ok;
}

// calculatePairWeight (device.go:135-157), weights of device.go:38-54
int calculatePairWeight(const Device &from, const Device &to, int linkType) {
    int weight = from.DevId == to.DevId ? 10 : 20;
    weight += linkType == 11 ? 10 : linkType == 2 ? 40 : 50;
    weight += from.NumaNode == to.NumaNode ? 10 : 20;
    return weight;
}

// scanAndPopulatePeerWeights (device.go:159-218)
void scanAndPopulatePeerWeights(const std::string &fromPath, const std::vector<Device> &devices,
                                const std::set<int> &lookupNodes, PairWeights &p2pWeights) {
    static const std::vector<std::regex> re = {std::regex("node_from[\\t\\n\\f\\r ]([0-9]+)"),
                                               std::regex("node_to[\\t\\n\\f\\r ]([0-9]+)"),
                                               std::regex("type[\\t\\n\\f\\r ]([0-9]+)")};
    std::vector<std::string> paths;
    for (const char *sub : {"io_links", "p2p_links"}) // io_links first (:160-168)
        for (const std::string &name : globDigits(fromPath + "/" + sub)) paths.push_back(fromPath + "/" + sub + "/" + name);
    for (const std::string &toPath : paths) {
        std::vector<int> vals;
        if (fetchTopoProperties(toPath + "/properties", re, vals) != Fetch::Ok) continue;
        const int from = std::min(vals[0], vals[1]), to = std::max(vals[0], vals[1]);
        if (!lookupNodes.count(from) || !lookupNodes.count(to)) continue;
        const Device *fromDev = nullptr, *toDev = nullptr;
        for (const Device &d : devices) {
            if (d.NodeId == from) fromDev = &d;
            if (d.NodeId == to) toDev = &d;
            if (fromDev && toDev) break;
        }
        if (fromDev && toDev) p2pWeights[from][to] = calculatePairWeight(*fromDev, *toDev, vals[2]);
    }
}

} // namespace

Error FetchAllPairWeights(const std::vector<Device> &devs, PairWeights &p2pWeights, const std::string &topoDir) {
    if (devs.empty()) return "Devices list is empty. Unable to calculate pair wise weights"; // device.go:221-225
    const std::string folder = topoDir.empty() ? kTopoRootPath : topoDir;
    std::set<int> nodeIds;
    for (const Device &d : devs) nodeIds.insert(d.NodeId);
    static const std::vector<std::regex> minor = {std::regex("drm_render_minor[\\t\\n\\f\\r ]([0-9]+)")};
    for (const std::string &name : globDigits(folder)) {
        const std::string path = folder + "/" + name;
        std::vector<int> vals;
        if (fetchTopoProperties(path + "/properties", minor, vals) != Fetch::Ok || vals[0] <= 0) continue; // :242
        scanAndPopulatePeerWeights(path, devs, nodeIds, p2pWeights);
    }
    return "";
}

BestEffortPolicy::BestEffortPolicy() {
    options_.cuda_device = -1;
}

BestEffortPolicy::~BestEffortPolicy() {
    if (handle_) tp_destroy(handle_);
}

std::unique_ptr<BestEffortPolicy> NewBestEffortPolicy() { return std::make_unique<BestEffortPolicy>(); }

Error BestEffortPolicy::Init(const std::vector<Device> &devs, const std::string &topoDir) {
    Error err = FetchAllPairWeights(devs, p2pWeights_, topoDir);
    if (p2pWeights_.empty()) return "Besteffort Policy init failed to initialize p2pWeights"; // :72-74
    if (!err.empty()) return err;
    return InitWithWeights(devs, p2pWeights_);
}

Error BestEffortPolicy::InitWithWeights(const std::vector<Device> &devs, const PairWeights &weights) {
    if (&weights != &p2pWeights_) p2pWeights_ = weights;
    if (p2pWeights_.empty()) return "Besteffort Policy init failed to initialize p2pWeights";
    if ((int)devs.size() > TP_MAX_DEVICES) return "too many devices for the GPU policy";
    devices_ = devs;
    devicesMap_.clear();
    for (size_t i = 0; i < devs.size(); i++) devicesMap_[devs[i].Id] = (int)i;

    // groupPartitionsByDevId (device.go:287-304): group index per DevId, ParentId = the
    // member whose Id lacks "amdgpu_xcp" (last such member wins).
    const int n = (int)devs.size();
    std::map<std::string, int> groupOf;
    std::vector<std::string> parentId;
    std::vector<int32_t> group(n), nodeId(n);
    std::map<int, int> indexOfNode;
    for (int i = 0; i < n; i++) {
        auto it = groupOf.find(devs[i].DevId);
        if (it == groupOf.end()) {
            it = groupOf.emplace(devs[i].DevId, (int)parentId.size()).first;
            parentId.emplace_back();
        }
        group[i] = it->second;
        if (devs[i].Id.find("amdgpu_xcp") == std::string::npos) parentId[it->second] = devs[i].Id;
        nodeId[i] = devs[i].NodeId;
        indexOfNode[devs[i].NodeId] = i;
    }
    groupCount_ = parentId.size();
    // rank of each ParentId under Go's bytewise string `<` (device.go:342-349)
    std::vector<std::string> sorted(parentId);
    std::sort(sorted.begin(), sorted.end());
    sorted.erase(std::unique(sorted.begin(), sorted.end()), sorted.end());
    std::vector<int32_t> parentRank(parentId.size());
    for (size_t g = 0; g < parentId.size(); g++)
        parentRank[g] = (int32_t)(std::lower_bound(sorted.begin(), sorted.end(), parentId[g]) - sorted.begin());

    std::vector<uint8_t> weight((size_t)n * n, 0);
    for (const auto &row : p2pWeights_)
        for (const auto &cell : row.second) {
            auto a = indexOfNode.find(row.first), b = indexOfNode.find(cell.first);
            if (a == indexOfNode.end() || b == indexOfNode.end() || a->second == b->second) continue;
            weight[(size_t)a->second * n + b->second] = (uint8_t)cell.second;
            weight[(size_t)b->second * n + a->second] = (uint8_t)cell.second;
        }
    tp_topology topo{};
    topo.n_devices = n;
    topo.n_groups = (int32_t)parentId.size();
    topo.node_id = nodeId.data();
    topo.group = group.data();
    topo.parent_rank = parentRank.data();
    topo.weight = weight.data();
    if (handle_) {
        tp_destroy(handle_);
        handle_ = nullptr;
    }
    int rc = tp_create(&topo, &options_, &handle_);
    if (rc != TP_OK) {
        devices_.clear(); // Allocate must then fail with invalidInit, as after a failed Init
        return std::string("GPU policy init failed: ") + tp_strerror(rc);
    }
    return "";
}

int BestEffortPolicy::IndexOf(const std::string &id) const {
    auto it = devicesMap_.find(id);
    return it == devicesMap_.end() ? -1 : it->second;
}

Error BestEffortPolicy::mapIds(const std::vector<std::string> &ids, std::vector<int32_t> &out) const {
    out.clear();
    for (const std::string &id : ids) out.push_back(IndexOf(id)); // unknown -> -1 -> TP_ERR_BAD_INDEX
    return "";
}

namespace {
// setContainsAll (device.go:88-105) on the id strings
bool setContainsAll(const std::vector<std::string> &set, const std::vector<std::string> &subset) {
    if (subset.size() > set.size()) return false;
    for (const std::string &s : subset)
        if (std::find(set.begin(), set.end(), s) == set.end()) return false;
    return true;
}

Error statusToError(int rc, tp_policy *h) {
    if (rc >= 0) return "";
    if (rc == TP_ERR_CUDA && h) return std::string("CUDA error: ") + tp_last_cuda_error(h);
    return tp_strerror(rc);
}
} // namespace

std::pair<std::vector<std::string>, Error> BestEffortPolicy::Allocate(const std::vector<std::string> &availableIds,
                                                                      const std::vector<std::string> &requiredIds,
                                                                      int size, tp_result *info) const {
    std::vector<std::string> outset;
    if (info) *info = tp_result{};
    // Same checks, same order as besteffort_policy.go:90-124: the count checks and the two
    // short-circuits look at ids only, so they are answered before any id is resolved.
    if (size <= 0) return {outset, kInvalidSize};
    if ((int)availableIds.size() < size) return {outset, kInvalidAvailable};
    if ((int)requiredIds.size() > size) return {outset, kInvalidRequired};
    if (requiredIds.size() > availableIds.size()) return {outset, kInvalidReqAvailable};
    if (devices_.empty()) return {outset, kInvalidInit};
    if ((int)availableIds.size() == size) return {availableIds, ""};
    if ((int)requiredIds.size() == size) return {requiredIds, ""};
    if (p2pWeights_.empty() || !handle_) return {outset, kInvalidInit};
    if (!setContainsAll(availableIds, requiredIds)) return {outset, kNoCandidateFound};

    std::vector<int32_t> av, rq;
    mapIds(availableIds, av);
    mapIds(requiredIds, rq);
    std::vector<int32_t> out(size);
    int rc = tp_allocate(handle_, av.data(), (int)av.size(), rq.data(), (int)rq.size(), size, out.data(), info);
    if (rc < 0) return {outset, statusToError(rc, handle_)};
    for (int i = 0; i < rc; i++) outset.push_back(devices_[out[i]].Id);
    return {outset, ""};
}

Error BestEffortPolicy::SearchShard(const std::vector<std::string> &availableIds,
                                    const std::vector<std::string> &requiredIds, int size, int shardIndex,
                                    int shardCount, tp_key *key, bool *shortCircuit, tp_result *info) const {
    if (!handle_) return kInvalidInit;
    std::vector<int32_t> av, rq;
    mapIds(availableIds, av);
    mapIds(requiredIds, rq);
    int rc = tp_search_shard(handle_, av.data(), (int)av.size(), rq.data(), (int)rq.size(), size, shardIndex,
                             shardCount, key, info);
    if (shortCircuit) *shortCircuit = rc == 1;
    return statusToError(rc, handle_);
}

std::pair<std::vector<std::string>, Error> BestEffortPolicy::Decode(const std::vector<std::string> &availableIds,
                                                                    const std::vector<std::string> &requiredIds,
                                                                    int size, const tp_key &key) const {
    std::vector<std::string> outset;
    if (!handle_) return {outset, kInvalidInit};
    std::vector<int32_t> av, rq;
    mapIds(availableIds, av);
    mapIds(requiredIds, rq);
    std::vector<int32_t> out(std::max(size, 1));
    int rc = tp_decode(handle_, av.data(), (int)av.size(), rq.data(), (int)rq.size(), size, &key, out.data());
    if (rc < 0) return {outset, statusToError(rc, handle_)};
    for (int i = 0; i < rc; i++) {
        if (out[i] < 0 || out[i] >= (int)devices_.size()) return {{}, "unknown or duplicate device index"};
        outset.push_back(devices_[out[i]].Id);
    }
    return {outset, ""};
}

} // namespace tp

// ---- flat C view -------------------------------------------------------------------------
struct tph_policy {
    tp::BestEffortPolicy pol;
    std::string last_error;
};

namespace {
std::vector<tp::Device> toDevices(const tph_device *devs, int n) {
    std::vector<tp::Device> v(n > 0 ? n : 0);
    for (int i = 0; i < n; i++) {
        v[i].Id = devs[i].id;
        v[i].NodeId = devs[i].node_id;
        v[i].NumaNode = devs[i].numa_node;
        v[i].DevId = devs[i].dev_id;
    }
    return v;
}
std::vector<std::string> toStrings(const char *const *s, int n) {
    std::vector<std::string> v;
    for (int i = 0; i < n; i++) v.emplace_back(s[i]);
    return v;
}
int errorToStatus(const std::string &e) {
    if (e.empty()) return 0;
    for (int code : {TP_ERR_INVALID_SIZE, TP_ERR_INVALID_AVAILABLE, TP_ERR_INVALID_REQUIRED,
                     TP_ERR_INVALID_REQ_AVAILABLE, TP_ERR_INVALID_INIT, TP_ERR_NO_CANDIDATE, TP_ERR_INIT_FAILED,
                     TP_ERR_BAD_INDEX, TP_ERR_BAD_ARGUMENT, TP_ERR_TOO_LARGE})
        if (e == tp_strerror(code)) return code;
    if (e.rfind("CUDA error", 0) == 0) return TP_ERR_CUDA;
    if (e.rfind("GPU policy init failed", 0) == 0) return TP_ERR_CUDA;
    return TP_ERR_BAD_ARGUMENT;
}
void writeIds(const std::vector<std::string> &ids, char *out) {
    for (size_t i = 0; i < ids.size(); i++) {
        strncpy(out + i * 64, ids[i].c_str(), 63);
        out[i * 64 + 63] = 0;
    }
}
} // namespace

extern "C" {

tph_policy *tph_new(const tp_options *opts) {
    tph_policy *p = new tph_policy();
    if (opts) p->pol.SetOptions(*opts);
    return p;
}
void tph_free(tph_policy *p) { delete p; }

int tph_init(tph_policy *p, const tph_device *devs, int n, const char *topo_dir) {
    p->last_error = p->pol.Init(toDevices(devs, n), topo_dir ? topo_dir : "");
    return errorToStatus(p->last_error);
}

int tph_init_pairs(tph_policy *p, const tph_device *devs, int n, const int32_t *from, const int32_t *to,
                   const int32_t *weight, int n_pairs) {
    tp::PairWeights w;
    for (int i = 0; i < n_pairs; i++) w[from[i]][to[i]] = weight[i];
    p->last_error = p->pol.InitWithWeights(toDevices(devs, n), w);
    return errorToStatus(p->last_error);
}

int tph_allocate(tph_policy *p, const char *const *avail, int n_avail, const char *const *req, int n_req, int size,
                 char *out_ids, tp_result *info) {
    auto res = p->pol.Allocate(toStrings(avail, n_avail), toStrings(req, n_req), size, info);
    p->last_error = res.second;
    if (!res.second.empty()) return errorToStatus(res.second);
    writeIds(res.first, out_ids);
    return (int)res.first.size();
}

int tph_search_shard(tph_policy *p, const char *const *avail, int n_avail, const char *const *req, int n_req,
                     int size, int shard_index, int shard_count, tp_key *key) {
    bool sc = false;
    p->last_error = p->pol.SearchShard(toStrings(avail, n_avail), toStrings(req, n_req), size, shard_index,
                                       shard_count, key, &sc);
    if (!p->last_error.empty()) return errorToStatus(p->last_error);
    return sc ? 1 : 0;
}

int tph_decode(tph_policy *p, const char *const *avail, int n_avail, const char *const *req, int n_req, int size,
               const tp_key *key, char *out_ids) {
    auto res = p->pol.Decode(toStrings(avail, n_avail), toStrings(req, n_req), size, *key);
    p->last_error = res.second;
    if (!res.second.empty()) return errorToStatus(res.second);
    writeIds(res.first, out_ids);
    return (int)res.first.size();
}

int tph_weight_rows(const tph_policy *p) { return (int)p->pol.WeightRows(); }
int tph_group_count(const tph_policy *p) { return (int)p->pol.GroupCount(); }
int tph_pair_count(const tph_policy *p) {
    int c = 0;
    for (const auto &r : p->pol.Weights()) c += (int)r.second.size();
    return c;
}
int tph_pairs(const tph_policy *p, int32_t *from, int32_t *to, int32_t *weight, int cap) {
    int c = 0;
    for (const auto &r : p->pol.Weights())
        for (const auto &cell : r.second) {
            if (c < cap) {
                from[c] = r.first;
                to[c] = cell.first;
                weight[c] = cell.second;
            }
            c++;
        }
    return c;
}
int tph_index_of(const tph_policy *p, const char *id) { return p->pol.IndexOf(id); }
tp_policy *tph_handle(tph_policy *p) { return p->pol.Handle(); }
const char *tph_last_error(const tph_policy *p) { return p->last_error.c_str(); }

} // extern "C"
