// replay_core.cc -- host layer above the C ABI; see rela_b200/replay_core.h.
#include "rela_b200/replay_core.h"

#include <algorithm>
#include <cstdlib>
#include <fstream>
#include <stdexcept>

#include <c10/cuda/CUDAFunctions.h>
#include <c10/cuda/CUDAStream.h>

#include "rela_b200.h"

namespace rela_b200 {

void throwLast(int code, const char* where) {
  std::string msg = std::string(where) + ": rela_b200 error " + std::to_string(code) + ": " + rr_last_error();
  throw std::runtime_error(msg);
}

namespace {

inline void check(int rc, const char* where) {
  if (rc != RR_OK) throwLast(rc, where);
}

int pickDevice(int device) {
  if (device >= 0) return device;
  if (const char* e = std::getenv("RELA_B200_DEVICE")) {
    std::string s(e);
    auto pos = s.find(':');
    return std::atoi(pos == std::string::npos ? s.c_str() : s.c_str() + pos + 1);
  }
  if (!torch::cuda::is_available()) return 0;  // rr_create reports the missing device
  return (int)c10::cuda::current_device();
}

inline void* streamOf(int device) {
  return (void*)c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)device).stream();
}

// A new tensor object over the same storage, built without the dispatcher.
torch::Tensor aliasOf(const torch::Tensor& base) {
  auto impl = c10::make_intrusive<c10::TensorImpl>(c10::Storage(base.storage()), base.key_set(), base.dtype());
  impl->set_sizes_contiguous(base.sizes());
  return torch::Tensor(std::move(impl));
}

uint64_t numelOf(const std::vector<int64_t>& shape) {
  uint64_t n = 1;
  for (auto s : shape) n *= (uint64_t)s;
  return n;
}

}  // namespace

ShardConfig shardConfigFromEnv() {
  ShardConfig c;
  const char* on = std::getenv("RELA_B200_SHARDED");
  if (!(on && std::atoi(on) != 0)) return c;
  auto geti = [](const char* n, int d) {
    const char* v = std::getenv(n);
    return (v && *v) ? std::atoi(v) : d;
  };
  c.world = geti("WORLD_SIZE", 1);
  c.rank = geti("RANK", 0);
  const char* l = std::getenv("RELA_B200_LEARNER");
  c.learner = (l && *l && std::string(l) != "all") ? std::atoi(l) : -1;
  const char* g = std::getenv("RELA_B200_GROUP");
  const char* port = std::getenv("MASTER_PORT");
  c.group = g ? g : (std::string("job") + (port ? port : "0"));
  return c;
}

ReplayCore::ReplayCore(int capacity, int seed, float alpha, float beta, int prefetch, int device)
    : alpha_(alpha), shard_(shardConfigFromEnv()) {
  init(capacity, seed, alpha, beta, prefetch, device);
}

ReplayCore::ReplayCore(int capacity, int seed, float alpha, float beta, int prefetch, int device,
                       const ShardConfig& shard)
    : alpha_(alpha), shard_(shard) {
  init(capacity, seed, alpha, beta, prefetch, device);
}

void ReplayCore::init(int capacity, int seed, float alpha, float beta, int prefetch, int device) {
  if (shard_.world > 1 && device < 0 && std::getenv("RELA_B200_DEVICE") == nullptr) {
    // one process per GPU: LOCAL_RANK (several ranks may share a GPU on a smaller box)
    const char* lr = std::getenv("LOCAL_RANK");
    const int n = torch::cuda::is_available() ? (int)c10::cuda::device_count() : 1;
    device = ((lr && *lr) ? std::atoi(lr) : shard_.rank) % std::max(1, n);
  }
  device_ = pickDevice(device);
  // Host priorities are raised with the function the reference calls
  // (torch::pow on the CPU, rela/prioritized_replay.h:188,239), which makes the
  // stored weights -- and through them the sampled indices -- bit-identical to
  // the reference's. RELA_B200_DEVICE_POW=1 raises them in the kernel instead
  // (fp64 pow rounded once; <= 1e-6 relative).
  const char* e = std::getenv("RELA_B200_DEVICE_POW");
  host_pow_ = !(e && std::atoi(e) != 0);
  // prefetch > 0: the futures of the reference (rela/prioritized_replay.h:215-230)
  // are batches selected and gathered ahead straight into pooled batch tensors
  // (rr_prefetch_into / rr_take_prefetched): the C ABI replay itself runs without
  // prefetch and nothing is copied when a batch is handed out.
  prefetch_ = std::max(0, prefetch);
  check(rr_create(&h_, capacity, seed, alpha, beta, 0, device_), "PrioritizedReplay");
  check(rr_get_stream(h_, &own_stream_), "PrioritizedReplay");
  if (shard_.world > 1) startSharded();
  startIssuer();
}

ReplayCore::~ReplayCore() {
  stopIssuer();
  inflight_.clear();
  stopSharded();
  if (h_ != nullptr) rr_destroy(h_);
}

int ReplayCore::size() const { return rr_size(h_); }
int ReplayCore::numAdd() const { return rr_num_add(h_); }
void ReplayCore::sync() {
  drainIssuer();
  check(rr_sync(h_), "sync");
}
// The C ABI checkpoint holds bytes; dtypes and shapes of the leaves are the host
// layer's business and go into a text sidecar (<path>.schema), one leaf per
// line: name dtype chunks time_major row_bytes ndim dims...
void ReplayCore::save(const std::string& path) {
  if (!schema_set_.load(std::memory_order_acquire)) throw std::runtime_error("save: nothing has been added yet");
  drainIssuer();
  const std::string file = shard_.world > 1 ? shardFile(path) : path;
  check(rr_save(h_, file.c_str()), "save");  // batches drawn ahead (inflight_) are saved with it
  if (shard_.rank != 0) return;
  const std::string tmp = path + ".schema.tmp";
  {
    std::ofstream f(tmp);
    f << "rela_b200-schema 1 " << schema_.size() << "\n";
    for (const LeafSpec& s : schema_) {
      f << s.name << ' ' << (int)s.dtype << ' ' << s.chunks << ' ' << (s.time_major ? 1 : 0) << ' ' << s.row_bytes << ' '
        << s.shape.size();
      for (int64_t d : s.shape) f << ' ' << d;
      f << "\n";
    }
    f.flush();
    if (!f.good()) throw std::runtime_error("save: cannot write " + tmp);
  }
  if (std::rename(tmp.c_str(), (path + ".schema").c_str()) != 0)
    throw std::runtime_error("save: cannot rename " + tmp);
}

void ReplayCore::load(const std::string& path) {
  std::ifstream f(path + ".schema");
  std::string tag;
  int version = 0;
  size_t n = 0;
  if (!(f >> tag >> version >> n) || tag != "rela_b200-schema" || version != 1 || n > 4096)
    throw std::runtime_error("load: cannot read " + path + ".schema");
  std::vector<LeafSpec> specs(n);
  for (LeafSpec& s : specs) {
    int dtype = 0, tm = 0;
    size_t nd = 0;
    if (!(f >> s.name >> dtype >> s.chunks >> tm >> s.row_bytes >> nd) || nd > 16)
      throw std::runtime_error("load: corrupt " + path + ".schema");
    s.dtype = (torch::ScalarType)dtype;
    s.time_major = tm != 0;
    s.shape.resize(nd);
    for (int64_t& d : s.shape)
      if (!(f >> d)) throw std::runtime_error("load: corrupt " + path + ".schema");
  }
  setSchema(specs);  // a replay that already has a schema must have the same one
  const std::string file = shard_.world > 1 ? shardFile(path) : path;
  check(rr_load(h_, file.c_str()), "load");
  // batches the saved run had drawn ahead: claim them into pooled batch tensors, oldest first
  const int queued = rr_prefetched(h_);
  void* stream = streamOf(device_);
  for (int k = 0; k < queued; ++k) {
    BatchSet set;
    {
      std::lock_guard<std::mutex> pl(pool_m_);
      set = acquireSet(rr_prefetched_batch(h_, k));
    }
    std::vector<void*> ptrs(schema_.size() ? schema_.size() : 1, nullptr);
    for (size_t i = 0; i < schema_.size(); ++i) ptrs[i] = set.fields[i].data_ptr();
    check(rr_restore_prefetched(h_, k, ptrs.data(), set.weights.data_ptr<float>(), stream), "load");
    if (issuer_)
      issuerAdopt(std::move(set), stream);
    else
      inflight_.push_back(std::move(set));
  }
}

std::string ReplayCore::shardFile(const std::string& path) const {
  return path + ".shard" + std::to_string(shard_.rank) + "of" + std::to_string(shard_.world);
}

void ReplayCore::applySchema(const std::vector<LeafSpec>& specs) {  // schema_m_ held
  std::vector<rr_field> fields(specs.size() ? specs.size() : 1);
  for (size_t i = 0; i < specs.size(); ++i) {
    fields[i].name = specs[i].name.c_str();
    fields[i].row_bytes = specs[i].row_bytes;
    fields[i].chunks = specs[i].chunks;
    fields[i].reserved = 0;
  }
  if (shard_.world > 1) check(rr_set_shard(h_, shard_.rank, shard_.world), "add (shard)");
  check(rr_set_schema(h_, (int)specs.size(), fields.data()), "add (schema)");
  schema_ = specs;
  if (shard_.world > 1) shareSlabs();
  schema_set_.store(true, std::memory_order_release);
}

void ReplayCore::ensureSchema(const Leaves& first, bool per_element, bool time_major) {
  std::lock_guard<std::mutex> lk(schema_m_);
  if (schema_set_.load(std::memory_order_acquire)) return;
  std::vector<LeafSpec> specs;
  for (const Leaf& lf : first) {
    if (!lf.t.defined()) throw std::runtime_error("add: leaf '" + lf.name + "' is undefined");
    LeafSpec s;
    s.name = lf.name;
    s.dtype = lf.t.scalar_type();
    auto sz = lf.t.sizes();
    if (per_element) {
      s.shape.assign(sz.begin(), sz.end());
      if (s.shape.empty()) s.shape.push_back(1);  // 0-dim seqLen (r2d2_actor.h:109)
      // sequences [T, ...] and hidden states [layers, hid] are gathered on
      // dim 1; seq_len is gathered on dim 0 (rela/types.cc:161-167)
      const bool scalar = lf.t.dim() == 0 || (lf.t.dim() == 1 && sz[0] == 1 && lf.name == "seq_len");
      s.time_major = time_major && !scalar;
      s.chunks = s.time_major ? (uint32_t)s.shape[0] : 1u;
    } else {
      if (lf.t.dim() < 1) throw std::runtime_error("add: leaf '" + lf.name + "' has no batch dimension");
      s.shape.assign(sz.begin() + 1, sz.end());
      s.chunks = 1;
      s.time_major = false;
    }
    s.row_bytes = numelOf(s.shape) * (uint64_t)c10::elementSize(s.dtype);
    if (s.chunks == 0) s.chunks = 1;
    specs.push_back(std::move(s));
  }
  applySchema(specs);
}

void ReplayCore::setSchema(const std::vector<LeafSpec>& specs) {
  std::lock_guard<std::mutex> lk(schema_m_);
  if (schema_set_.load(std::memory_order_acquire)) {
    bool same = specs.size() == schema_.size();
    for (size_t i = 0; same && i < specs.size(); ++i)
      same = specs[i].name == schema_[i].name && specs[i].dtype == schema_[i].dtype &&
             specs[i].shape == schema_[i].shape && specs[i].chunks == schema_[i].chunks;
    if (!same) throw std::runtime_error("setSchema: differs from the schema of the first add");
    return;
  }
  applySchema(specs);
}

void ReplayCore::setOutputFormat(const std::string& name, torch::ScalarType dtype) {
  drainIssuer();
  std::lock_guard<std::mutex> lk(schema_m_);
  if (!schema_set_.load(std::memory_order_acquire)) throw std::runtime_error("set_output_format: the schema is not known yet (add or load first)");
  if (!inflight_.empty() || issuerQueued() > 0) throw std::runtime_error("set_output_format: prefetched batches are pending");
  for (size_t i = 0; i < schema_.size(); ++i) {
    LeafSpec& s = schema_[i];
    if (s.name != name) continue;
    int fmt;
    if (dtype == torch::kByte) {
      fmt = RR_OUT_RAW;
    } else if (dtype == torch::kFloat32) {
      fmt = RR_OUT_F32_DIV255;
    } else if (dtype == torch::kBFloat16) {
      fmt = RR_OUT_BF16_DIV255;
    } else {
      throw std::runtime_error("set_output_format: dtype must be float32, bfloat16 or uint8");
    }
    if (s.dtype != torch::kByte) throw std::runtime_error("set_output_format: leaf '" + name + "' is not uint8");
    check(rr_set_output_format(h_, (int)i, fmt), "set_output_format");
    s.converted = fmt != RR_OUT_RAW;
    s.out_dtype = dtype;
    {
      std::lock_guard<std::mutex> pl(pool_m_);
      pool_.clear();  // pooled batch sets have the old dtype
    }
    return;
  }
  throw std::runtime_error("set_output_format: no leaf named '" + name + "'");
}

void* ReplayCore::stream() const { return streamOf(device_); }

void ReplayCore::addDeviceRows(int n, const std::vector<const void*>& ptrs, const torch::Tensor& priority) {
  unsigned flags = 0;
  torch::Tensor w = weightsOf(priority, &flags);
  if (w.size(0) != n || ptrs.size() != schema_.size() * (size_t)n)
    throw std::runtime_error("addDeviceRows: size mismatch");
  if (n == 0) return;
  check(rr_add_rows(h_, n, ptrs.data(), w.data_ptr<float>(), flags | RR_ROWS_DEVICE, streamOf(device_)), "add");
}

void ReplayCore::scatterRows(int n, const std::vector<uint64_t>& row_bytes, const std::vector<const void*>& src,
                             const std::vector<void*>& dst, const int32_t* dst_rows_dev) {
  check(rr_scatter_rows(h_, n, (int)row_bytes.size(), row_bytes.data(), src.data(), dst.data(), dst_rows_dev,
                        streamOf(device_)),
        "scatterRows");
}

void ReplayCore::checkLeaf(size_t i, const torch::Tensor& t, bool per_element, int64_t n) const {
  const LeafSpec& s = schema_[i];
  bool ok = t.defined() && t.scalar_type() == s.dtype;
  if (ok) {
    const uint64_t want = per_element ? numelOf(s.shape) : numelOf(s.shape) * (uint64_t)n;
    ok = (uint64_t)t.numel() == want && (per_element || (t.dim() >= 1 && t.size(0) == n));
  }
  if (!ok) throw std::runtime_error("add: leaf '" + s.name + "' does not match the schema of the first add");
}

torch::Tensor ReplayCore::weightsOf(const torch::Tensor& priority, unsigned* flags) const {
  if (priority.dim() != 1)  // assert(priority.dim() == 1), rela/prioritized_replay.h:187,236
    throw std::runtime_error("priority must be a 1-D tensor");
  if (priority.scalar_type() != torch::kFloat32)  // accessor<float, 1>, :59,107
    throw std::runtime_error("priority must be float32");
  torch::Tensor p = priority.detach().contiguous();
  if (shard_.world > 1) return p;  // raised where its position in the collective block is known (sharded_core.cc)
  if (p.is_cuda()) {
    if (p.get_device() != device_) p = p.to(torch::Device(torch::kCUDA, (c10::DeviceIndex)device_));
    *flags |= RR_PRIO_DEVICE;
    return p;
  }
  if (host_pow_ && alpha_ != 1.0f) {
    *flags |= RR_PRIO_IS_WEIGHT;
    return torch::pow(p, alpha_);
  }
  return p;
}

void ReplayCore::addBatch(const Leaves& batch, const torch::Tensor& priority) {
  unsigned flags = 0;
  torch::Tensor w = weightsOf(priority, &flags);
  const int64_t n = w.size(0);
  if (!schema_set_.load(std::memory_order_acquire)) ensureSchema(batch, false, false);
  if (batch.size() != schema_.size()) throw std::runtime_error("add: number of leaves differs from the schema");
  std::vector<torch::Tensor> keep(batch.size());
  std::vector<const void*> ptrs(batch.size() ? batch.size() : 1, nullptr);
  bool any_cuda = false, any_cpu = false;
  for (size_t i = 0; i < batch.size(); ++i) {
    checkLeaf(i, batch[i].t, false, n);
    keep[i] = batch[i].t.contiguous();
    (keep[i].is_cuda() ? any_cuda : any_cpu) = true;
  }
  if (shard_.world > 1) {
    shardedAdd(keep, w, flags);
    return;
  }
  if (any_cuda && any_cpu) {
    for (auto& t : keep)
      if (t.is_cuda()) t = t.cpu();
    any_cuda = false;
  }
  for (size_t i = 0; i < keep.size(); ++i) {
    if (any_cuda && keep[i].get_device() != device_)
      keep[i] = keep[i].to(torch::Device(torch::kCUDA, (c10::DeviceIndex)device_));
    ptrs[i] = keep[i].data_ptr();
  }
  if (any_cuda) flags |= RR_ROWS_DEVICE;
  void* stream = (flags & (RR_ROWS_DEVICE | RR_PRIO_DEVICE)) ? streamOf(device_) : nullptr;
  check(rr_add(h_, (int)n, ptrs.data(), nullptr, w.data_ptr<float>(), flags, stream), "add");
}

void ReplayCore::addElements(const std::vector<Leaves>& elements, const torch::Tensor& priority,
                             bool time_major) {
  unsigned flags = 0;
  torch::Tensor w = weightsOf(priority, &flags);
  const int64_t n = w.size(0);
  if ((int64_t)elements.size() != n) throw std::runtime_error("add: priority length differs from the elements");
  if (n == 0) return;
  if (!schema_set_.load(std::memory_order_acquire)) ensureSchema(elements[0], true, time_major);
  const size_t nf = schema_.size();
  std::vector<torch::Tensor> keep;
  keep.reserve(nf * (size_t)n);
  std::vector<const void*> ptrs(std::max<size_t>(nf * (size_t)n, 1), nullptr);
  bool any_cuda = false, any_cpu = false;
  for (int64_t e = 0; e < n; ++e) {
    if (elements[e].size() != nf) throw std::runtime_error("add: number of leaves differs from the schema");
    for (size_t f = 0; f < nf; ++f) {
      checkLeaf(f, elements[e][f].t, true, 1);
      keep.push_back(elements[e][f].t.contiguous());
      (keep.back().is_cuda() ? any_cuda : any_cpu) = true;
    }
  }
  if (shard_.world > 1) {
    // a striped replay moves packed blocks: stack the elements field by field
    std::vector<torch::Tensor> packed(nf);
    for (size_t f = 0; f < nf; ++f) {
      std::vector<torch::Tensor> col;
      col.reserve((size_t)n);
      for (int64_t e = 0; e < n; ++e) col.push_back(keep[(size_t)e * nf + f].reshape(schema_[f].shape));
      packed[f] = torch::stack(col, 0);
    }
    shardedAdd(packed, w, flags);
    return;
  }
  if (any_cuda && any_cpu) {
    for (auto& t : keep)
      if (t.is_cuda()) t = t.cpu();
    any_cuda = false;
  }
  for (int64_t e = 0; e < n; ++e)
    for (size_t f = 0; f < nf; ++f) ptrs[f * (size_t)n + (size_t)e] = keep[(size_t)e * nf + f].data_ptr();
  if (any_cuda) flags |= RR_ROWS_DEVICE;
  void* stream = (flags & (RR_ROWS_DEVICE | RR_PRIO_DEVICE)) ? streamOf(device_) : nullptr;
  check(rr_add_rows(h_, (int)n, ptrs.data(), w.data_ptr<float>(), flags, stream), "add");
}

std::vector<int64_t> ReplayCore::batchShape(const LeafSpec& s, int64_t B) const {
  std::vector<int64_t> shape;
  if (!s.time_major) {
    shape.push_back(B);
    shape.insert(shape.end(), s.shape.begin(), s.shape.end());
  } else {
    shape.push_back(s.shape[0]);
    shape.push_back(B);
    shape.insert(shape.end(), s.shape.begin() + 1, s.shape.end());
  }
  return shape;
}

// A batch set nobody else holds: a pooled one that neither the host layer (lease: batches drawn
// ahead) nor the caller (storages back to a use count of 1) still uses, else a new one (pooled if
// there is room).
ReplayCore::BatchSet ReplayCore::acquireSet(int64_t B) {
  for (auto& cand : pool_) {
    if (cand.B != B) continue;
    bool idle = cand.lease.use_count() <= 1 && cand.weights.storage().use_count() == 1;
    for (size_t i = 0; idle && i < cand.fields.size(); ++i) idle = cand.fields[i].storage().use_count() == 1;
    if (idle) return cand;
  }
  const torch::Device home(torch::kCUDA, (c10::DeviceIndex)device_);
  BatchSet set;
  set.B = B;
  set.weights = torch::empty({B}, torch::TensorOptions().dtype(torch::kFloat32).device(home));
  for (const LeafSpec& s : schema_)
    set.fields.push_back(torch::empty(batchShape(s, B),
                                      torch::TensorOptions().dtype(s.converted ? s.out_dtype : s.dtype).device(home)));
  if (issuer_) set.ready = newReadyEvent();
  set.lease = std::make_shared<char>(0);
  if (pool_.size() < (size_t)(4 + prefetch_)) pool_.push_back(set);
  return set;
}

// sample() with prefetch > 0 (rela/prioritized_replay.h:208-232): hand out the
// oldest batch drawn ahead (or draw one now if there is none), then draw until
// `prefetch` batches are queued again -- before this batch's updatePriority,
// like the reference's futures.
// What the caller receives: fresh tensor objects over the storages of a batch set, on the
// device the caller named (the reference assembles on the host and copies to `device`,
// rela/types.cc:34-43; here the batch is assembled in HBM and moved only if the learner asked
// for another device).
std::pair<Leaves, torch::Tensor> ReplayCore::handOut(const BatchSet& set, const std::string& device) const {
  Leaves out;
  out.reserve(schema_.size());
  for (size_t i = 0; i < schema_.size(); ++i) out.push_back(Leaf{schema_[i].name, aliasOf(set.fields[i])});
  torch::Tensor weights = aliasOf(set.weights);
  const torch::Device home(torch::kCUDA, (c10::DeviceIndex)device_);
  const torch::Device want(device);
  if (want != home && !(want.is_cuda() && !want.has_index())) {
    for (auto& lf : out) lf.t = lf.t.to(want);
    weights = weights.to(want);
  }
  return {std::move(out), std::move(weights)};
}

// sample() with prefetch > 0 (rela/prioritized_replay.h:208-232): hand out the
// oldest batch drawn ahead (or draw one now if there is none), then draw until
// `prefetch` batches are queued again -- before this batch's updatePriority,
// like the reference's futures.
std::pair<Leaves, torch::Tensor> ReplayCore::samplePrefetched(int batchsize, const std::string& device) {
  const int64_t B = batchsize;
  void* stream = streamOf(device_);
  auto issue = [&]() {
    BatchSet set = acquireSet(B);  // the copy in inflight_ keeps the set busy (storage use count 2)
    std::vector<void*> ptrs(schema_.size() ? schema_.size() : 1, nullptr);
    for (size_t i = 0; i < schema_.size(); ++i) ptrs[i] = set.fields[i].data_ptr();
    check(rr_prefetch_into(h_, batchsize, ptrs.data(), set.weights.data_ptr<float>(), stream), "sample");
    inflight_.push_back(std::move(set));
  };
  if (!inflight_.empty() && inflight_.front().B != B) throw std::runtime_error("sample: batch size differs from the prefetched one");
  if (inflight_.empty()) issue();
  // (a caller on the replay's own stream: the prefetch issued below carries the wait for this batch)
  check(rr_take_prefetched(h_, (stream == own_stream_ && prefetch_ > 0) ? RR_STREAM_OWN_DEFERRED : stream), "sample");
  BatchSet set = std::move(inflight_.front());  // (its lease keeps it out of acquireSet until the aliases exist)
  inflight_.pop_front();
  while ((int)inflight_.size() < prefetch_) issue();
  return handOut(set, device);
}

std::pair<Leaves, torch::Tensor> ReplayCore::sample(int batchsize, const std::string& device) {
  if (!schema_set_.load(std::memory_order_acquire)) {
    // the reference prints "cannot find enough samples" and asserts (:297-302)
    throw std::runtime_error("sample: the replay buffer is empty");
  }
  if (shard_.world > 1) return shardedSample(batchsize, device);
  if (issuer_) return issuerSample(batchsize, device);
  if (prefetch_ > 0 || !inflight_.empty()) return samplePrefetched(batchsize, device);
  // Batch tensors come from a small pool. The caller never receives the pooled
  // tensor objects, only fresh aliases of their storages; a set is reused once
  // every storage is back to a use count of 1, i.e. the caller has dropped the
  // batch and every view of it -- the steady state of a learner loop that
  // rebinds `batch` each step. (Handing out the pooled TensorImpl itself would
  // never come back to a count of 1: its Python wrapper outlives the rebind
  // while another C++ owner exists.) A new set is allocated after the selection
  // kernel has been issued, so that the allocation overlaps it.
  const int64_t B = batchsize;
  const torch::Device home(torch::kCUDA, (c10::DeviceIndex)device_);
  BatchSet* set = nullptr;
  for (auto& cand : pool_) {
    if (cand.B != B) continue;
    bool idle = cand.lease.use_count() <= 1 && cand.weights.storage().use_count() == 1;
    for (size_t i = 0; idle && i < cand.fields.size(); ++i) idle = cand.fields[i].storage().use_count() == 1;
    if (idle) {
      set = &cand;
      break;
    }
  }
  BatchSet fresh;
  const bool reuse = set != nullptr;
  if (!reuse) {
    set = &fresh;
    set->B = B;
    set->weights = torch::empty({B}, torch::TensorOptions().dtype(torch::kFloat32).device(home));
  }
  void* stream = streamOf(device_);
  check(rr_sample_select(h_, batchsize, set->weights.data_ptr<float>(), stream), "sample");
  if (!reuse) {
    for (const LeafSpec& s : schema_)
      set->fields.push_back(torch::empty(batchShape(s, B),
                                         torch::TensorOptions().dtype(s.converted ? s.out_dtype : s.dtype).device(home)));
  }
  std::vector<void*> ptrs(schema_.size() ? schema_.size() : 1, nullptr);
  for (size_t i = 0; i < schema_.size(); ++i) ptrs[i] = set->fields[i].data_ptr();
  check(rr_sample_gather(h_, ptrs.data(), stream), "sample");
  auto out = handOut(*set, device);
  if (!reuse && pool_.size() < 4) pool_.push_back(std::move(fresh));
  return out;
}

void ReplayCore::updatePriority(const torch::Tensor& priority) {
  if (issuer_) {
    issuerUpdate(priority);
    return;
  }
  unsigned flags = 0;
  torch::Tensor w = weightsOf(priority, &flags);
  if (shard_.world > 1) {
    shardedUpdate(w, flags);
    return;
  }
  check(rr_update_priority(h_, (int)w.size(0), w.data_ptr<float>(), flags, streamOf(device_)),
        "update_priority");
}

}  // namespace rela_b200
