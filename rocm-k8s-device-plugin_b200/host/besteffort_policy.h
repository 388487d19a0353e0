// besteffort_policy.h -- host-side mirror of the reference's allocator package surface.
//
// The reference is Go (internal/pkg/allocator); no Go toolchain exists in this image, so the
// policy object a Go maintainer would write over cgo is written here in C++
// with the same names, argument meaning and error behaviour:
//
//   allocator.Device            device.go:56-65            -> tp::Device
//   allocator.Policy            allocator.go:27-30         -> tp::Policy
//   NewBestEffortPolicy()       besteffort_policy.go:52-59 -> tp::NewBestEffortPolicy()
//   (*BestEffortPolicy).Init    besteffort_policy.go:70-86 -> tp::BestEffortPolicy::Init
//   (*BestEffortPolicy).Allocate besteffort_policy.go:88-151 -> tp::BestEffortPolicy::Allocate
//
// Init parses the KFD sysfs tree exactly as fetchAllPairWeights does (device.go:107-252) and
// hands the matrix to the GPU through tp_create; Allocate maps ids to indices, calls
// tp_allocate and maps back.  Errors are returned as strings ("" = nil), equal to the
// reference's (besteffort_policy.go:36-43).  There is no CPU search path in this class.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "tp_allocator.h"

namespace tp {

struct Device { // device.go:56-65
    std::string Id;
    int NodeId = 0;
    int NumaNode = 0;
    std::string DevId;
    int Card = 0;
    int RenderD = 0;
    std::string ComputePartitionType;
    std::string MemoryPartitionType;
};

using Error = std::string; // "" is nil

class Policy { // allocator.go:27-30
  public:
    virtual ~Policy() = default;
    virtual Error Init(const std::vector<Device> &devs, const std::string &topoDir) = 0;
    virtual std::pair<std::vector<std::string>, Error> Allocate(const std::vector<std::string> &available,
                                                                const std::vector<std::string> &required,
                                                                int size) = 0;
};

// p2pWeights: map[from]map[to]weight with from < to (device.go:181-188, :214)
using PairWeights = std::map<int, std::map<int, int>>;

// fetchAllPairWeights (device.go:220-252); topoDir "" = /sys/class/kfd/kfd/topology/nodes.
Error FetchAllPairWeights(const std::vector<Device> &devs, PairWeights &p2pWeights, const std::string &topoDir);

class BestEffortPolicy : public Policy {
  public:
    BestEffortPolicy();
    ~BestEffortPolicy() override;
    Error Init(const std::vector<Device> &devs, const std::string &topoDir) override;
    // Init for callers that already hold p2pWeights (the Go parsers stay in Go).
    Error InitWithWeights(const std::vector<Device> &devs, const PairWeights &p2pWeights);
    // Reentrant, like the reference's (its policy is read-only after Init, one goroutine per RPC calls it):
    // nothing in the object is written after Init.
    std::pair<std::vector<std::string>, Error> Allocate(const std::vector<std::string> &available,
                                                        const std::vector<std::string> &required,
                                                        int size) override {
        return Allocate(available, required, size, nullptr);
    }
    // The same with the search's counters (score, subsets scored, kernel time) returned to the caller.
    std::pair<std::vector<std::string>, Error> Allocate(const std::vector<std::string> &available,
                                                        const std::vector<std::string> &required, int size,
                                                        tp_result *info) const;

    // Shard of the search for multi-GPU runs (tp_search_shard / tp_decode).
    Error SearchShard(const std::vector<std::string> &available, const std::vector<std::string> &required,
                      int size, int shardIndex, int shardCount, tp_key *key, bool *shortCircuit,
                      tp_result *info = nullptr) const;
    std::pair<std::vector<std::string>, Error> Decode(const std::vector<std::string> &available,
                                                      const std::vector<std::string> &required, int size,
                                                      const tp_key &key) const;

    void SetOptions(const tp_options &o) { options_ = o; }
    tp_policy *Handle() const { return handle_; }
    size_t WeightRows() const { return p2pWeights_.size(); }
    size_t GroupCount() const { return groupCount_; }
    const PairWeights &Weights() const { return p2pWeights_; }
    int IndexOf(const std::string &id) const;

  private:
    Error mapIds(const std::vector<std::string> &ids, std::vector<int32_t> &out) const;
    std::vector<Device> devices_;
    std::map<std::string, int> devicesMap_; // Id -> index (later duplicates win, besteffort_policy.go:77-79)
    PairWeights p2pWeights_;
    size_t groupCount_ = 0;
    tp_options options_{};
    tp_policy *handle_ = nullptr;
};

std::unique_ptr<BestEffortPolicy> NewBestEffortPolicy();

} // namespace tp

// Flat C view of the class for ctypes-based tests and bench.py.
extern "C" {
typedef struct tph_device {
    char id[64];
    int32_t node_id;
    int32_t numa_node;
    char dev_id[64];
} tph_device;
typedef struct tph_policy tph_policy;
tph_policy *tph_new(const tp_options *opts);
void tph_free(tph_policy *p);
int tph_init(tph_policy *p, const tph_device *devs, int n, const char *topo_dir);
int tph_init_pairs(tph_policy *p, const tph_device *devs, int n, const int32_t *from, const int32_t *to,
                   const int32_t *weight, int n_pairs);
int tph_allocate(tph_policy *p, const char *const *avail, int n_avail, const char *const *req, int n_req, int size,
                 char *out_ids /* size x 64 */, tp_result *info);
int tph_search_shard(tph_policy *p, const char *const *avail, int n_avail, const char *const *req, int n_req,
                     int size, int shard_index, int shard_count, tp_key *key);
int tph_decode(tph_policy *p, const char *const *avail, int n_avail, const char *const *req, int n_req, int size,
               const tp_key *key, char *out_ids);
int tph_weight_rows(const tph_policy *p);
int tph_group_count(const tph_policy *p);
int tph_pair_count(const tph_policy *p);
int tph_pairs(const tph_policy *p, int32_t *from, int32_t *to, int32_t *weight, int cap);
int tph_index_of(const tph_policy *p, const char *id);
tp_policy *tph_handle(tph_policy *p);
const char *tph_last_error(const tph_policy *p);
}
