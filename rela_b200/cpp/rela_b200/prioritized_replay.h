// prioritized_replay.h -- rela_b200::PrioritizedReplay<DataType, Traits>: the
// public surface of the reference's rela::PrioritizedReplay<DataType>
// (rela/prioritized_replay.h:173-347) over the B200 replay.
//
//   ctor(capacity, seed, alpha, beta, prefetch)          :176-184
//   add(const std::vector<DataType>&, priority)          :186-191
//   add(const DataType& batch, priority)                 :193-200
//   sample(batchsize, device) -> (DataType, Tensor)      :202-233
//   updatePriority(priority)                             :235-245
//   size(), numAdd()                                     :247-253
//
// Traits<DataType> maps the record type to an ordered list of named leaves and
// back. It is the only place that knows the members of a transition; the
// drop-in header (dropin/rela/prioritized_replay.h) supplies traits for the
// reference's own rela::FFTransition / rela::RNNTransition.
#pragma once

#include <algorithm>
#include <memory>
#include <string>
#include <tuple>
#include <type_traits>
#include <vector>

#include "rela_b200/replay_core.h"

namespace rela_b200 {

template <class DataType>
struct TransitionTraits;  // { static constexpr bool kTimeMajor; leaves(); fromLeaves(); }

template <class DataType, class Traits = TransitionTraits<DataType>>
class PrioritizedReplay {
 public:
  PrioritizedReplay(int capacity, int seed, float alpha, float beta, int prefetch)
      : core_(capacity, seed, alpha, beta, prefetch) {
  }

  PrioritizedReplay(int capacity, int seed, float alpha, float beta, int prefetch, int device)
      : core_(capacity, seed, alpha, beta, prefetch, device) {
  }

  // one shard of a replay striped over several GPUs (ShardConfig, rela_b200/replay_core.h)
  PrioritizedReplay(int capacity, int seed, float alpha, float beta, int prefetch, int device, const ShardConfig& shard)
      : core_(capacity, seed, alpha, beta, prefetch, device, shard) {
  }

  // Elements one by one (R2D2Actor::postStep, rela/r2d2_actor.h:301).
  void add(const std::vector<DataType>& sample, const torch::Tensor& priority) {
    std::vector<Leaves> elements;
    elements.reserve(sample.size());
    for (const DataType& s : sample) elements.push_back(Traits::leaves(s));
    core_.addElements(elements, priority, Traits::kTimeMajor);
  }

  // A batch whose row i is element i (DQNActor::postStep, rela/dqn_actor.h:189).
  void add(const DataType& sample, const torch::Tensor& priority) {
    addBatched(sample, priority, std::integral_constant<bool, Traits::kTimeMajor>());
  }

  std::tuple<DataType, torch::Tensor> sample(int batchsize, const std::string& device) {
    auto res = core_.sample(batchsize, device);
    return std::make_tuple(Traits::fromLeaves(res.first, batchsize), std::move(res.second));
  }

  void updatePriority(const torch::Tensor& priority) {
    core_.updatePriority(priority);
  }

  int size() const {
    return core_.size();
  }

  int numAdd() const {
    return core_.numAdd();
  }

  // Checkpoint / restore (not in the reference): see ReplayCore::save.
  void save(const std::string& path) {
    core_.save(path);
  }

  void load(const std::string& path) {
    core_.load(path);
  }

  ReplayCore& core() {
    return core_;
  }

 private:
  void addBatched(const DataType& sample, const torch::Tensor& priority, std::false_type) {
    core_.addBatch(Traits::leaves(sample), priority);
  }

  // Time-major records have no row-per-element layout: split with the record's
  // own index(i), as the reference's add does (rela/prioritized_replay.h:193-200).
  void addBatched(const DataType& sample, const torch::Tensor& priority, std::true_type) {
    std::vector<DataType> vec;
    const int n = (int)priority.size(0);
    vec.reserve(n);
    for (int i = 0; i < n; ++i) vec.push_back(sample.index(i));
    add(vec, priority);
  }

  ReplayCore core_;
};

// helpers for traits over std::unordered_map<std::string, torch::Tensor>
template <class Dict>
inline void appendDict(Leaves& out, const char* group, const Dict& d) {
  std::vector<const typename Dict::value_type*> items;
  items.reserve(d.size());
  for (const auto& kv : d) items.push_back(&kv);
  std::sort(items.begin(), items.end(),
            [](const typename Dict::value_type* a, const typename Dict::value_type* b) { return a->first < b->first; });
  for (const auto* kv : items) out.push_back(Leaf{std::string(group) + "." + kv->first, kv->second});
}

inline bool leafInGroup(const std::string& name, const char* group, std::string* key) {
  const std::string g = std::string(group) + ".";
  if (name.compare(0, g.size(), g) != 0) return false;
  *key = name.substr(g.size());
  return true;
}

}  // namespace rela_b200
