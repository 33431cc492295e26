// replay_core.h -- C++ host layer of the B200 replay, above the C ABI
// (include/rela_b200.h) and below the PrioritizedReplay<T> template
// (rela_b200/prioritized_replay.h).
//
// The reference stores an element as a struct of tensor dictionaries
// (rela/types.h:18-73) and is schema-agnostic. The host layer flattens a
// transition into an ordered list of named leaves; the list of the first add
// fixes the schema (names, dtypes, per-element shapes) of the HBM slabs.
// PyTorch appears here only as the tensor hand-off: it allocates the batch
// tensors that rr_sample fills and supplies the caller's CUDA stream.
#pragma once

#include <atomic>
#include <cstdint>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include <torch/extension.h>

struct rr_replay;

namespace rela_b200 {

struct Leaf {
  std::string name;
  torch::Tensor t;
};
using Leaves = std::vector<Leaf>;

struct LeafSpec {
  std::string name;
  torch::ScalarType dtype;
  std::vector<int64_t> shape;  // per stored element
  uint32_t chunks;             // T: the row is T chunks, gathered time-major
  bool time_major;             // gathered [shape[0], B, shape[1:]...] instead of [B, shape...]
  uint64_t row_bytes;
  // dtype of the gathered leaf: `dtype`, or float32 / bfloat16 for a u8 leaf
  // that the gather converts to byte / 255 (setOutputFormat)
  torch::ScalarType out_dtype = torch::kByte;
  bool converted = false;
};

// One replay striped over the GPUs of one box (no
// reference counterpart). One process per GPU, SPMD: every rank constructs the
// replay with the same arguments (`capacity` is the TOTAL capacity) and calls
// add / sample / update_priority collectively, in the same order.
//   add(batch, priority)        every rank contributes n rows (same n); the
//                               block is the ranks' rows in rank order
//   sample(B, device)           learners = -1: every rank is a learner and
//                               receives B rows: rows [rank * B, (rank + 1) * B)
//                               of ONE stratified batch of world * B drawn from
//                               the whole buffer -- what one reference replay
//                               returns for sample(world * B), split over the
//                               ranks. learners = r: rank r receives all B rows,
//                               the other ranks get empty tensors.
//   update_priority(priority)   each learner passes the priorities of its rows
// With $RELA_B200_SHARDED=1 the plain constructor reads the configuration from
// the torchrun environment (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_PORT;
// $RELA_B200_LEARNER = "all" | rank), so that unmodified callers of the
// reference API (pyrela/main.py:138-144, rela/pybind.cc:37-47) get a striped replay.
struct ShardConfig {
  int rank = 0, world = 1;
  int learner = -1;    // -1: every rank is a learner (data-parallel); else the one learner's rank
  std::string group;   // rendezvous name shared by the ranks of this replay (rr_group_create)
};

class ReplayCore {
 public:
  // PrioritizedReplay(capacity, seed, alpha, beta, prefetch)
  // (rela/prioritized_replay.h:176-184). The slabs live on CUDA device
  // `device`; -1 = $RELA_B200_DEVICE, else the current device.
  ReplayCore(int capacity, int seed, float alpha, float beta, int prefetch, int device = -1);
  ReplayCore(int capacity, int seed, float alpha, float beta, int prefetch, int device, const ShardConfig& shard);
  ~ReplayCore();
  ReplayCore(const ReplayCore&) = delete;
  ReplayCore& operator=(const ReplayCore&) = delete;

  // add(const DataType& batch, priority): every leaf is [n, ...]; row i is
  // element i (FFTransition::index, rela/types.cc:48-67).
  void addBatch(const Leaves& batch, const torch::Tensor& priority);
  // add(const std::vector<DataType>&, priority): one leaf list per element.
  // time_major: leaves are [T, ...] sequences gathered into [T, B, ...]
  // (RNNTransition::makeBatch, rela/types.cc:140-167).
  void addElements(const std::vector<Leaves>& elements, const torch::Tensor& priority, bool time_major);

  // sample(batchsize, device) (rela/prioritized_replay.h:202-233)
  std::pair<Leaves, torch::Tensor> sample(int batchsize, const std::string& device);
  // updatePriority(priority) (rela/prioritized_replay.h:235-245)
  void updatePriority(const torch::Tensor& priority);

  int size() const;
  int numAdd() const;

  // ---- device-resident producers (SequenceAssembler)
  // Freeze the schema from explicit specs (no-op when already set; the specs
  // must then match).
  void setSchema(const std::vector<LeafSpec>& specs);
  // add(vector<DataType>) whose rows already live in HBM: ptrs[f * n + e] is
  // field f of element e (device pointer); ordered on the current stream.
  void addDeviceRows(int n, const std::vector<const void*>& ptrs, const torch::Tensor& priority);
  // row i of src[f] ([n, row] packed, device) -> row dst_rows[i] of dst[f]
  void scatterRows(int n, const std::vector<uint64_t>& row_bytes, const std::vector<const void*>& src,
                   const std::vector<void*>& dst, const int32_t* dst_rows_dev);
  void* stream() const;

  // Learner-side fusion (rr_set_output_format): the u8 leaf `name` comes out of
  // sample() as `dtype` (float32 or bfloat16) = byte / 255, bit-identical to
  // `leaf.float() / 255` (`.bfloat16()`), without the extra pass over the batch
  // (pyrela/net.py:46,104). dtype = uint8 switches it off. Needs the schema
  // (after the first add / load / setSchema).
  void setOutputFormat(const std::string& name, torch::ScalarType dtype);

  void sync();  // waits for the device; throws on a sticky device-side error
  // Checkpoint / restore (rr_save / rr_load): a restored replay continues
  // bit-identically. save() also writes <path>.schema (leaf names, dtypes,
  // shapes); load() takes the schema from it and needs a freshly created replay.
  void save(const std::string& path);
  void load(const std::string& path);
  rr_replay* handle() const { return h_; }
  int device() const { return device_; }
  const std::vector<LeafSpec>& schema() const { return schema_; }
  struct Sharded;  // state of the striped-replay protocol (sharded_core.cc)
  struct Issuer;   // the thread that issues batches drawn ahead and host-priority updates (issuer_core.cc)
  // priority ^ alpha for `p` = elements [offset, offset + n) of a block of `total` priorities that
  // ONE replay would raise with a single torch::pow: bit-identical to that call's slice
  static torch::Tensor raiseAtPosition(const torch::Tensor& p, int64_t offset, int64_t total, float alpha);
  int shardRank() const { return shard_.rank; }
  int shardWorld() const { return shard_.world; }
  // Striped replay, every rank a learner: which B rows of the stratified batch of world * B the
  // batch returned by the last sample() is -- rows [lastChunk() * B, (lastChunk() + 1) * B). The
  // chunks are assigned by locality (rr_route::my_chunk), so this is not always the rank. 0 otherwise.
  int lastChunk() const;

 private:
  void ensureSchema(const Leaves& first, bool per_element, bool time_major);
  void checkLeaf(size_t i, const torch::Tensor& t, bool per_element, int64_t n) const;
  torch::Tensor weightsOf(const torch::Tensor& priority, unsigned* flags) const;

  struct BatchSet {
    int64_t B = 0;
    std::vector<torch::Tensor> fields;
    torch::Tensor weights;
    // issuer mode: a cudaEvent_t recorded behind the gather that fills this set (shared by the
    // copies of the set; destroyed with the last one)
    std::shared_ptr<void> ready;
    // Shared by the pool's copy of the set and every copy the host layer itself holds (batches
    // drawn ahead that have not been handed out): use_count() > 1 = the set is in use. Copies of a
    // set share their TensorImpls, so the storages' use counts only see the CALLER's aliases.
    std::shared_ptr<char> lease;
  };
  std::vector<BatchSet> pool_;  // batch tensors reused once the caller has released them
  std::deque<BatchSet> inflight_;  // batches drawn ahead (prefetch), oldest first
  int prefetch_ = 0;
  BatchSet acquireSet(int64_t B);
  std::vector<int64_t> batchShape(const LeafSpec& s, int64_t B) const;
  std::pair<Leaves, torch::Tensor> samplePrefetched(int batchsize, const std::string& device);
  std::pair<Leaves, torch::Tensor> handOut(const BatchSet& set, const std::string& device) const;
  void init(int capacity, int seed, float alpha, float beta, int prefetch, int device);
  void applySchema(const std::vector<LeafSpec>& specs);
  std::string shardFile(const std::string& path) const;
  // issuer_core.cc: with prefetch > 0 the reference draws batches ahead on other threads
  // (std::async futures, rela/prioritized_replay.h:215-230). Here ONE issuer thread owns the
  // replay's control stream: it launches the selections / gathers of batches drawn ahead and the
  // updates whose priorities arrive in host memory, in call order, while the calling thread only
  // waits for a batch's event and hands it out. See issuer_core.cc for what stays synchronous.
  void startIssuer();
  void stopIssuer();
  void drainIssuer();  // returns when the issuer has executed everything posted; rethrows its error
  std::pair<Leaves, torch::Tensor> issuerSample(int batchsize, const std::string& device);
  void issuerUpdate(const torch::Tensor& priority);
  void issuerAdopt(BatchSet set, void* stream);  // load(): a restored batch joins the batches drawn ahead
  int issuerQueued() const;
  std::shared_ptr<void> newReadyEvent() const;
  void issueInto(BatchSet& set, int batchsize, void* stream);  // rr_prefetch_into + ready event
  // sharded_core.cc
  void startSharded();
  void stopSharded();
  void shareSlabs();
  void shardedAdd(std::vector<torch::Tensor>& rows, const torch::Tensor& w, unsigned flags);
  std::pair<Leaves, torch::Tensor> shardedSample(int batchsize, const std::string& device);
  void shardedUpdate(const torch::Tensor& w, unsigned flags);

  rr_replay* h_ = nullptr;
  void* own_stream_ = nullptr;  // rr_get_stream
  int device_ = 0;
  float alpha_ = 1.0f;
  bool host_pow_ = true;
  std::vector<LeafSpec> schema_;
  std::atomic<bool> schema_set_{false};  // written under schema_m_ (release), read without it (acquire)
  mutable std::mutex schema_m_;

  // striped replay (ShardConfig): the SPMD protocol lives in sharded_core.cc
  ShardConfig shard_;
  std::shared_ptr<Sharded> sharded_;  // shared_ptr: the type is complete only in sharded_core.cc
  std::shared_ptr<Issuer> issuer_;    // null: prefetch == 0, striped replay, or RELA_B200_ISSUER_THREAD=0
  std::mutex pool_m_;                 // pool_ / inflight_ when the issuer thread is running
};

[[noreturn]] void throwLast(int code, const char* where);

}  // namespace rela_b200
