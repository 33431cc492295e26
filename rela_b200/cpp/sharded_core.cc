// sharded_core.cc -- host side of ONE replay striped over the GPUs of one box
// (BASELINE.json configs[4]); see ShardConfig in
// rela_b200/replay_core.h for the calling convention. No reference
// counterpart: rela has one host-memory buffer. Parity is defined against that
// single buffer: the striped replay returns the ids, bytes and weights ONE
// reference replay returns for the same elements and the same draws.
//
// One process per GPU, SPMD. What is replicated, what is partitioned, and what
// crosses NVLink:
//
//   control plane   replicated: every shard holds all S weights, the eviction
//                   flags, the prefix hierarchy, sum_ and the mt19937 and applies
//                   every operation in the same order, so all shards select the
//                   same ids with no collective on the sampling path.
//   payload         partitioned: shard r stores slots [r * ceil(S / N), ...).
//   add             every rank's insert kernel stores ITS rows straight into the
//                   slab of the shard that owns each slot (peer-mapped slabs):
//                   each row crosses NVLink at most once. The block's weights
//                   are all-gathered by rr_exchange (peer stores + flags).
//   sample          every shard selects the whole batch; its gather kernel moves
//                   the rows it owns straight into the batch tensors of the
//                   learner each row belongs to (bulk stores over NVLink) and
//                   publishes a completion flag in every rank's memory.
//   update_priority the learners' priorities are all-gathered by rr_exchange;
//                   every shard applies all of them.
//
// The steady-state cycle contains no host-side communication and no collective
// library call: rr_group (shared memory) is used at set-up time only, to
// exchange the CUDA IPC handles of slabs, batch arenas and inboxes.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>

#include <c10/cuda/CUDAFunctions.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime_api.h>

#include "rela_b200.h"
#include "rela_b200/replay_core.h"

namespace rela_b200 {

namespace {

inline void check(int rc, const char* where) {
  if (rc != RR_OK) throwLast(rc, where);
}

inline void cuda(cudaError_t e, const char* where) {
  if (e != cudaSuccess) throw std::runtime_error(std::string(where) + ": " + cudaGetErrorString(e));
}

inline void* streamOf(int device) {
  return (void*)c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)device).stream();
}

constexpr size_t kInboxFloats = 1u << 20;  // largest block of weights one exchange moves (adds)
constexpr size_t kInboxHeader = 256;       // 2 parities x 16 ranks x 8-byte flags
constexpr size_t kUpdateSlots = 16384;     // largest batch (all learners) of one update_priority: kMaxBatch
// an inbox: [flags][weights of an add, parity 0][parity 1][update slots, array 0][array 1]...
constexpr size_t kInboxSlotsOff = kInboxHeader + 2 * kInboxFloats * sizeof(float);
// How many slot arrays the updates rotate through. A rank that contributes priorities needed every
// contributor's values to finish its previous update, so contributors are never more than one
// exchange apart -- but a rank WITHOUT a contribution (a shard of a replay with one learner) holds
// nobody back that way: the learner is held back only by the gathers, it issues update y after it
// took batch y, i.e. after every shard's gather of batch y, which that shard launched behind its
// selection y and therefore behind its update y - prefetch - 1. The learner can thus be `prefetch`
// exchanges ahead of a shard that is still polling: prefetch + 2 arrays keep it off that array.
inline size_t updateArrays(int prefetch) { return (size_t)std::max(2, prefetch + 2); }
inline size_t inboxBytes(int prefetch) { return kInboxSlotsOff + updateArrays(prefetch) * kUpdateSlots * sizeof(uint64_t); }
constexpr size_t kAlign = 256;

inline size_t alignUp(size_t v) { return (v + kAlign - 1) / kAlign * kAlign; }

std::atomic<int> g_group_counter{0};

}  // namespace

struct ReplayCore::Sharded {
  rr_group* group = nullptr;
  int rank = 0, world = 1, learner = -1;
  int n_learners = 1, my_learner = -1;
  std::vector<std::pair<int, void*>> opened;  // peer mappings (device, pointer)
  std::vector<void*> owned;                   // device allocations of this rank
  void* main = nullptr;                       // the replay's internal stream

  void* slab_tab = nullptr;  // device uint64[world * nfields]

  // inbox of every rank: [flags 2 x 16][data parity 0: kInboxFloats][data parity 1][update slots x 2]
  std::vector<uint8_t*> inbox;
  uint64_t xseq = 0;  // exchanges of adds (rr_exchange)
  uint64_t useq = 0;  // exchanges of update_priority (8-byte words tagged with the sequence number)
  float* stage_pinned[2] = {nullptr, nullptr};
  float* stage_dev[2] = {nullptr, nullptr};
  cudaEvent_t stage_ev[2] = {nullptr, nullptr};
  uint64_t stage_seq = 0;

  // batch arena per learner batch size: nsets sets of [flags 256 B][field 0][field 1]...
  struct Arena {
    int64_t B = 0;
    int nsets = 0;
    size_t set_stride = 0;
    std::vector<size_t> field_off;
    std::vector<uint8_t*> base;  // [world]
    uint8_t* dst_tabs = nullptr;  // device: nsets tables uint64[n_learners * nfields]
    std::vector<std::vector<void*>> signals;  // [set][world]
    std::vector<BatchSet> sets;               // learners: tensors over this rank's arena
    int next = 0;
  };
  std::map<int64_t, Arena> arenas;
  struct Drawn {
    Arena* arena;
    int set;
    int chunk;  // which B rows of the stratified batch of n_learners * B this rank holds (-1: none)
  };
  std::deque<Drawn> inflight;
  int64_t cur_B = 0;  // per-learner size of the batch awaiting its priorities
  int cur_chunk = -1;  // ... and its position in the whole batch: rows [cur_chunk * B, (cur_chunk + 1) * B)
  bool by_locality = true;  // every rank a learner: chunks go to the GPU that stores most of their rows

  int rankOfLearner(int l) const { return learner < 0 ? l : learner; }
};

void ReplayCore::startSharded() {
  if (shard_.rank < 0 || shard_.rank >= shard_.world || shard_.world > 8)
    throw std::runtime_error("sharded replay: rank " + std::to_string(shard_.rank) + " of " +
                             std::to_string(shard_.world) + " (at most 8 shards)");
  if (shard_.learner >= shard_.world) throw std::runtime_error("sharded replay: learner rank out of range");
  sharded_.reset(new Sharded());
  Sharded& sh = *sharded_;
  sh.rank = shard_.rank;
  sh.world = shard_.world;
  sh.learner = shard_.learner;
  sh.n_learners = sh.learner < 0 ? sh.world : 1;
  sh.my_learner = sh.learner < 0 ? sh.rank : (sh.rank == sh.learner ? 0 : -1);
  if (const char* e = std::getenv("RELA_B200_ROUTE_BY_LOCALITY")) sh.by_locality = std::atoi(e) != 0;
  check(rr_get_stream(h_, &sh.main), "sharded replay");
  // every replay of the job gets its own rendezvous: the ranks construct their replays in the same order
  const std::string name = shard_.group + "." + std::to_string(g_group_counter++);
  check(rr_group_create(&sh.group, name.c_str(), sh.rank, sh.world), "sharded replay (rendezvous)");

  c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
  // inbox for the all-gather of weights (rr_exchange)
  void* mine = nullptr;
  check(rr_device_alloc(device_, inboxBytes(prefetch_), &mine), "sharded replay (inbox)");
  cuda(cudaMemset(mine, 0, inboxBytes(prefetch_)), "sharded replay (inbox)");  // no slot carries a valid tag yet
  sh.owned.push_back(mine);
  uint8_t handle[64];
  check(rr_ipc_export(device_, mine, handle), "sharded replay (inbox)");
  std::vector<uint8_t> all(64 * (size_t)sh.world);
  check(rr_group_allgather(sh.group, handle, 64, all.data()), "sharded replay (inbox)");
  sh.inbox.assign((size_t)sh.world, nullptr);
  for (int r = 0; r < sh.world; ++r) {
    if (r == sh.rank) {
      sh.inbox[(size_t)r] = static_cast<uint8_t*>(mine);
    } else {
      void* p = nullptr;
      check(rr_ipc_open(device_, all.data() + 64 * (size_t)r, &p), "sharded replay (inbox)");
      sh.opened.push_back({device_, p});
      sh.inbox[(size_t)r] = static_cast<uint8_t*>(p);
    }
  }
  for (int k = 0; k < 2; ++k) {
    cuda(cudaHostAlloc(&sh.stage_pinned[k], kInboxFloats * sizeof(float), cudaHostAllocDefault), "sharded replay");
    void* d = nullptr;
    check(rr_device_alloc(device_, kInboxFloats * sizeof(float), &d), "sharded replay");
    sh.owned.push_back(d);
    sh.stage_dev[k] = static_cast<float*>(d);
    cuda(cudaEventCreateWithFlags(&sh.stage_ev[k], cudaEventDisableTiming), "sharded replay");
  }
  check(rr_group_barrier(sh.group), "sharded replay");
}

void ReplayCore::stopSharded() {
  if (!sharded_) return;
  Sharded& sh = *sharded_;
  if (h_ != nullptr) rr_sync(h_);
  // nobody may unmap or free memory a peer's kernel can still write
  if (sh.group) rr_group_barrier(sh.group);
  sh.arenas.clear();
  for (auto& o : sh.opened) rr_ipc_close(o.first, o.second);
  if (sh.group) rr_group_barrier(sh.group);
  for (void* p : sh.owned) rr_device_free(device_, p);
  for (int k = 0; k < 2; ++k) {
    if (sh.stage_pinned[k]) cudaFreeHost(sh.stage_pinned[k]);
    if (sh.stage_ev[k]) cudaEventDestroy(sh.stage_ev[k]);
  }
  if (sh.group) rr_group_destroy(sh.group);
  sharded_.reset();
}

// After rr_set_schema: every shard maps every other shard's slabs and builds the
// table the insert kernel routes rows with (slot -> owner -> slab row).
void ReplayCore::shareSlabs() {
  Sharded& sh = *sharded_;
  const size_t nf = schema_.size();
  if (nf == 0) throw std::runtime_error("sharded replay: the schema has no leaf");
  c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
  std::vector<void*> slabs(nf, nullptr);
  check(rr_slab_ptrs(h_, slabs.data(), (int)nf), "sharded replay (slabs)");
  std::vector<uint8_t> mine(64 * nf, 0), all(64 * nf * (size_t)sh.world);
  for (size_t f = 0; f < nf; ++f)
    if (slabs[f] != nullptr) check(rr_ipc_export(device_, slabs[f], mine.data() + 64 * f), "sharded replay (slabs)");
  check(rr_group_allgather(sh.group, mine.data(), mine.size(), all.data()), "sharded replay (slabs)");
  std::vector<uint64_t> tab((size_t)sh.world * nf, 0);
  for (int r = 0; r < sh.world; ++r) {
    for (size_t f = 0; f < nf; ++f) {
      if (schema_[f].row_bytes == 0) continue;
      void* p = slabs[f];
      if (r != sh.rank) {
        check(rr_ipc_open(device_, all.data() + 64 * ((size_t)r * nf + f), &p), "sharded replay (slabs)");
        sh.opened.push_back({device_, p});
      }
      tab[(size_t)r * nf + f] = (uint64_t)(uintptr_t)p;
    }
  }
  check(rr_device_alloc(device_, tab.size() * sizeof(uint64_t), &sh.slab_tab), "sharded replay (slabs)");
  sh.owned.push_back(sh.slab_tab);
  cuda(cudaMemcpy(sh.slab_tab, tab.data(), tab.size() * sizeof(uint64_t), cudaMemcpyHostToDevice), "sharded replay");
  check(rr_group_barrier(sh.group), "sharded replay");
}

namespace {
// All-gather of `n_mine` floats per contributing rank into every rank's inbox. `src` is device
// memory (or null for a rank that only listens). Returns this rank's view of the assembled block.
const float* exchange(ReplayCore::Sharded& sh, rr_replay* h, const float* src, size_t n_mine, size_t offset,
                      void* stream) {
  const uint64_t x = ++sh.xseq;
  const size_t par = (size_t)(x & 1);
  const size_t data_off = kInboxHeader + par * kInboxFloats * sizeof(float);
  void* dst[8];
  void* flg[8];
  for (int r = 0; r < sh.world; ++r) {
    dst[r] = sh.inbox[(size_t)r] + data_off + offset * sizeof(float);
    flg[r] = sh.inbox[(size_t)r] + (par * 16 + (size_t)sh.rank) * 8;
  }
  uint8_t* mine = sh.inbox[(size_t)sh.rank];
  // every rank publishes its flag, contribution or not: when the wait returns, every rank has
  // consumed the previous block of this parity and issued everything ordered before its flag
  check(rr_exchange(h, src, n_mine * sizeof(float), dst, flg, sh.world, x, mine + par * 16 * 8, sh.world, stream),
        "exchange");
  return reinterpret_cast<const float*>(mine + data_off);
}

// Weights in host memory -> a device staging buffer, on the replay's internal stream.
const float* stageHost(ReplayCore::Sharded& sh, const float* host, size_t n) {
  const size_t k = (size_t)(sh.stage_seq++ & 1);
  cuda(cudaEventSynchronize(sh.stage_ev[k]), "stage");  // the copy that last read this pinned buffer
  std::memcpy(sh.stage_pinned[k], host, n * sizeof(float));
  cuda(cudaMemcpyAsync(sh.stage_dev[k], sh.stage_pinned[k], n * sizeof(float), cudaMemcpyHostToDevice,
                       static_cast<cudaStream_t>(sh.main)),
       "stage");
  cuda(cudaEventRecord(sh.stage_ev[k], static_cast<cudaStream_t>(sh.main)), "stage");
  return sh.stage_dev[k];
}
}  // namespace

// In a striped replay priorities are always raised before they are exchanged, with torch::pow
// where they live. For host tensors that is the function the reference calls
// (rela/prioritized_replay.h:188,239) -- but ATen's CPU pow is not position independent: the
// vectorised body of a tensor goes through Sleef, its last few elements through libm's powf,
// and the two can differ in the last bit. ONE reference replay raises the whole collective
// block at once, so this rank raises its `n` values AT THEIR POSITION `offset` inside a block
// of `total` (the other positions hold 1): bit-identical weights whatever the vector width.
torch::Tensor ReplayCore::raiseAtPosition(const torch::Tensor& p, int64_t offset, int64_t total, float alpha) {
  if (alpha == 1.0f) return p;
  if (p.is_cuda()) return torch::pow(p, alpha);
  torch::Tensor full = torch::ones({total}, torch::kFloat32);
  full.narrow(0, offset, p.size(0)).copy_(p);
  return torch::pow(full, alpha).narrow(0, offset, p.size(0)).contiguous();
}

void ReplayCore::shardedAdd(std::vector<torch::Tensor>& rows, const torch::Tensor& w_in, unsigned /*flags*/) {
  Sharded& sh = *sharded_;
  const int64_t n = w_in.size(0);
  const torch::Tensor w = raiseAtPosition(w_in, sh.rank * n, n * sh.world, alpha_);
  const int64_t n_total = n * sh.world;
  if ((size_t)n_total > kInboxFloats)
    throw std::runtime_error("add: a collective block of " + std::to_string(n_total) + " rows exceeds " +
                             std::to_string(kInboxFloats));
  if (n == 0) return;
  c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
  const torch::Device home(torch::kCUDA, (c10::DeviceIndex)device_);
  void* cs = streamOf(device_);
  std::vector<const void*> ptrs(rows.size(), nullptr);
  for (size_t i = 0; i < rows.size(); ++i) {
    if (!rows[i].is_cuda() || rows[i].get_device() != device_) rows[i] = rows[i].to(home);
    rows[i] = rows[i].contiguous();
    ptrs[i] = rows[i].data_ptr();
  }
  check(rr_add_sharded_rows(h_, (int)n_total, (int)(sh.rank * n), (int)n, ptrs.data(), sh.slab_tab, cs), "add");
  // the weights follow the rows on the internal stream: a shard that sees the flag sees the rows
  const float* mine;
  void* xs = sh.main;
  if (w.is_cuda()) {
    mine = w.data_ptr<float>();
    xs = cs;
  } else {
    mine = stageHost(sh, w.data_ptr<float>(), (size_t)n);
  }
  const float* all = exchange(sh, h_, mine, (size_t)n, (size_t)(sh.rank * n), xs);
  check(rr_add_sharded_commit(h_, all, RR_PRIO_IS_WEIGHT), "add");
  // `rows` (and a device `w`) are released when this returns: later work on the caller's stream
  // must not reuse their memory before the insert has read it
  if (cs != sh.main) check(rr_stream_fence(h_, cs), "add");
}

namespace {
size_t setBytes(const std::vector<LeafSpec>& schema, int64_t B, std::vector<size_t>* off) {
  size_t o = kAlign;  // the set's completion flags
  off->clear();
  for (const LeafSpec& s : schema) {
    off->push_back(o);
    o += alignUp((size_t)s.row_bytes * (size_t)B);
  }
  return o;
}
}  // namespace

std::pair<Leaves, torch::Tensor> ReplayCore::shardedSample(int batchsize, const std::string& device) {
  Sharded& sh = *sharded_;
  const int64_t B = batchsize;
  const size_t nf = schema_.size();
  c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
  void* cs = streamOf(device_);
  auto it = sh.arenas.find(B);
  if (it == sh.arenas.end()) {
    // first batch of this size: every rank allocates the same arena layout, the ranks map each
    // other's arenas (a collective step: all ranks sample the same sizes in the same order)
    Sharded::Arena a;
    a.B = B;
    a.nsets = prefetch_ + 2;
    a.set_stride = setBytes(schema_, B, &a.field_off);
    void* mine = nullptr;
    check(rr_device_alloc(device_, a.set_stride * (size_t)a.nsets, &mine), "sample (batch arena)");
    sh.owned.push_back(mine);
    uint8_t handle[64];
    check(rr_ipc_export(device_, mine, handle), "sample (batch arena)");
    std::vector<uint8_t> all(64 * (size_t)sh.world);
    check(rr_group_allgather(sh.group, handle, 64, all.data()), "sample (batch arena)");
    a.base.assign((size_t)sh.world, nullptr);
    for (int r = 0; r < sh.world; ++r) {
      void* p = mine;
      if (r != sh.rank) {
        check(rr_ipc_open(device_, all.data() + 64 * (size_t)r, &p), "sample (batch arena)");
        sh.opened.push_back({device_, p});
      }
      a.base[(size_t)r] = static_cast<uint8_t*>(p);
    }
    const size_t tab_n = (size_t)sh.n_learners * nf;
    std::vector<uint64_t> tabs((size_t)a.nsets * tab_n, 0);
    a.signals.resize((size_t)a.nsets);
    const torch::Device home(torch::kCUDA, (c10::DeviceIndex)device_);
    for (int k = 0; k < a.nsets; ++k) {
      for (int l = 0; l < sh.n_learners; ++l)
        for (size_t f = 0; f < nf; ++f)
          tabs[(size_t)k * tab_n + (size_t)l * nf + f] =
              (uint64_t)(uintptr_t)(a.base[(size_t)sh.rankOfLearner(l)] + (size_t)k * a.set_stride + a.field_off[f]);
      for (int r = 0; r < sh.world; ++r)
        a.signals[(size_t)k].push_back(a.base[(size_t)r] + (size_t)k * a.set_stride + 8 * (size_t)sh.rank);
      BatchSet set;
      set.B = B;
      set.weights = torch::empty({B}, torch::TensorOptions().dtype(torch::kFloat32).device(home));
      if (sh.my_learner >= 0) {
        for (size_t f = 0; f < nf; ++f) {
          const LeafSpec& s = schema_[f];
          uint8_t* p = a.base[(size_t)sh.rank] + (size_t)k * a.set_stride + a.field_off[f];
          set.fields.push_back(torch::from_blob(p, batchShape(s, B), torch::TensorOptions().dtype(s.dtype).device(home)));
        }
      }
      a.sets.push_back(std::move(set));
    }
    void* dt = nullptr;
    check(rr_device_alloc(device_, tabs.size() * sizeof(uint64_t), &dt), "sample (batch arena)");
    sh.owned.push_back(dt);
    cuda(cudaMemcpy(dt, tabs.data(), tabs.size() * sizeof(uint64_t), cudaMemcpyHostToDevice), "sample (batch arena)");
    a.dst_tabs = static_cast<uint8_t*>(dt);
    check(rr_group_barrier(sh.group), "sample (batch arena)");
    it = sh.arenas.emplace(B, std::move(a)).first;
  }
  Sharded::Arena& a = it->second;
  const size_t tab_n = (size_t)sh.n_learners * nf;
  auto issue = [&](bool on_demand) {
    const int k = a.next;
    a.next = (a.next + 1) % a.nsets;
    rr_route rt;
    std::memset(&rt, 0, sizeof(rt));
    rt.n_learners = sh.n_learners;
    rt.dst_tab = a.dst_tabs + (size_t)k * tab_n * sizeof(uint64_t);
    rt.signal_ptrs = a.signals[(size_t)k].data();
    rt.n_signals = sh.world;
    rt.wait_flags = a.base[(size_t)sh.rank] + (size_t)k * a.set_stride;
    rt.wait_n = sh.world;
    rt.my_learner = sh.my_learner;
    rt.out_weights = a.sets[(size_t)k].weights.data_ptr<float>();
    rt.on_demand = on_demand ? 1 : 0;
    int chunk = sh.my_learner;
    rt.my_chunk = (sh.learner < 0 && sh.world > 1 && sh.by_locality) ? &chunk : nullptr;
    check(rr_prefetch_routed(h_, (int)(B * sh.n_learners), &rt, cs), "sample");
    sh.inflight.push_back({&a, k, chunk});
  };
  if (!sh.inflight.empty() && sh.inflight.front().arena != &a)
    throw std::runtime_error("sample: batch size differs from the prefetched one");
  if (sh.inflight.empty()) issue(prefetch_ == 0);
  // a caller on the replay's own stream with batches drawn ahead: the prefetch issued right below
  // carries the wait for this batch (RR_STREAM_OWN_DEFERRED keeps the update -> select launch chain)
  check(rr_take_prefetched(h_, (cs == sh.main && prefetch_ > 0) ? RR_STREAM_OWN_DEFERRED : cs), "sample");
  sh.cur_B = B;
  const int k = sh.inflight.front().set;
  sh.cur_chunk = sh.inflight.front().chunk;
  sh.inflight.pop_front();
  while ((int)sh.inflight.size() < prefetch_) issue(false);
  if (sh.my_learner < 0) return {Leaves(), torch::Tensor()};
  // The tensors alias a ring of prefetch + 2 batch sets in peer-visible memory: batch y stays
  // valid until this rank calls update_priority for batch y + 1 (a shard re-gathers into the set
  // of batch y for batch y + prefetch + 2, behind its selection of that batch and therefore behind
  // its update y + 1, which cannot finish before this learner's update y + 1 has started).
  const BatchSet& set = a.sets[(size_t)k];
  Leaves out;
  out.reserve(nf);
  for (size_t f = 0; f < nf; ++f) out.push_back(Leaf{schema_[f].name, set.fields[f]});
  torch::Tensor weights = set.weights;
  const torch::Device home(torch::kCUDA, (c10::DeviceIndex)device_);
  const torch::Device want(device);
  if (want != home && !(want.is_cuda() && !want.has_index())) {
    for (auto& lf : out) lf.t = lf.t.to(want);
    weights = weights.to(want);
  }
  return {std::move(out), std::move(weights)};
}

int ReplayCore::lastChunk() const { return sharded_ ? sharded_->cur_chunk : 0; }

void ReplayCore::shardedUpdate(const torch::Tensor& w_in, unsigned /*flags*/) {
  Sharded& sh = *sharded_;
  if (sh.cur_B == 0) throw std::runtime_error("update_priority: no sampled batch is outstanding");
  c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
  void* cs = streamOf(device_);
  const size_t n = (size_t)sh.cur_B;
  const size_t total = n * (size_t)sh.n_learners;
  if (total > kUpdateSlots) throw std::runtime_error("update_priority: batch too large for the peer inbox");
  torch::Tensor w;
  unsigned f = RR_PRIO_IS_WEIGHT;
  void* xs = sh.main;
  if (sh.my_learner >= 0) {
    // every learner passes the priorities of ITS rows (the other ranks' argument is ignored)
    if (w_in.size(0) != sh.cur_B)  // assert((int)sampledIds_.size() == priority.size(0)), line 237
      throw std::runtime_error("update_priority: " + std::to_string(w_in.size(0)) + " priorities for a batch of " +
                               std::to_string(sh.cur_B));
    if (w_in.is_cuda() && sh.learner < 0) {
      // every rank is a learner and sees what kind of tensor it was given (all learners must pass
      // the same kind): device priorities stay raw and are raised by the update kernel
      w = w_in;
      f = RR_PRIO_DEVICE;
      xs = cs;
    } else {
      w = raiseAtPosition(w_in, sh.cur_chunk * sh.cur_B, sh.cur_B * sh.n_learners, alpha_);
      if (w.is_cuda()) {
        f |= RR_PRIO_DEVICE;
        xs = cs;
      }
    }
  }
  // the update kernel all-gathers the learners' priorities itself (ExchangeArgs): 8-byte words
  // tagged with the sequence number, in one of prefetch + 2 slot arrays taken in turn (updateArrays)
  const uint64_t x = ++sh.useq;
  const uint32_t tag = (uint32_t)(x % 0xffffffffull) + 1u;  // never 0 (the cleared arrays)
  const size_t slots_off = kInboxSlotsOff + (size_t)(x % updateArrays(prefetch_)) * kUpdateSlots * sizeof(uint64_t);
  void* dst[8];
  for (int r = 0; r < sh.world; ++r) dst[r] = sh.inbox[(size_t)r] + slots_off;
  const bool learner = sh.my_learner >= 0;
  check(rr_update_priority_sharded(h_, (int)total, learner ? w.data_ptr<float>() : nullptr, learner ? (int)n : 0,
                                   learner ? (int)((size_t)sh.cur_chunk * n) : 0, f, dst, sh.world, tag,
                                   sh.inbox[(size_t)sh.rank] + slots_off, xs),
        "update_priority");
  sh.cur_B = 0;
  if (w.defined() && w.is_cuda() && cs != sh.main) check(rr_stream_fence(h_, cs), "update_priority");
}

}  // namespace rela_b200
