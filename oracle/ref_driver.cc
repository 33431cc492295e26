// oracle/ref_driver.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A thin C-callable driver around the UNMODIFIED reference replay
// (/root/reference/rela/prioritized_replay.h + types.cc + dqn_actor.h). It is
// compiled where the sources lie (see oracle/Makefile) into
// oracle/_ref/libref_replay.so and is used (a) to pin the C restatement in
// oracle/replay_oracle.c, (b) to generate tests/golden/*, (c) as the timed CPU
// baseline (`bench.py --impl reference`, cpu_baseline.kind == "reference").
// Nothing under rela_b200/ may link or load this file.
//
// The driver itself contains no replay logic: it only turns flat byte rows into
// the reference's FFTransition / RNNTransition structs, calls the reference's
// add / sample / updatePriority, and copies results back out.
//
// Field naming (schema order is whatever the caller passes):
//   FF : "obs.<k>" "action.<k>" "reward" "terminal" "bootstrap" "next_obs.<k>"
//   RNN: "obs.<k>" "h0.<k>" "action.<k>" "reward" "terminal" "bootstrap" "seq_len"
// Per-element shapes are given WITHOUT the batch dimension. For RNN the
// per-element shape includes the leading time dimension T (h0: [L,H];
// seq_len: [1], which makes reference types.cc:167 `.squeeze(1)` valid on
// torch >= 1.5 without patching the reference).

#include <chrono>
#include <cstdint>
#include <cstring>
#include <deque>
#include <future>
#include <iostream>
#include <queue>
#include <random>
#include <string>
#include <vector>

#include <torch/extension.h>
#include <torch/script.h>
#include <torch/torch.h>

// sampledIds_ is private in the reference; the parity tests need the physical
// slot ids. All std/torch headers are already included above, so this only
// affects the reference's own classes.
#define private public
#include "rela/prioritized_replay.h"
#undef private
#include "rela/dqn_actor.h"

using namespace rela;

namespace {

struct Field {
  std::string name;
  int dtype;  // 0 u8, 1 f32, 2 i64, 3 bool, 4 i32, 5 f64
  std::vector<int64_t> shape;
  std::string group;  // obs / action / next_obs / h0 / "" (top level)
  std::string key;
};

torch::Dtype toDtype(int code) {
  switch (code) {
    case 0: return torch::kUInt8;
    case 1: return torch::kFloat32;
    case 2: return torch::kInt64;
    case 3: return torch::kBool;
    case 4: return torch::kInt32;
    case 5: return torch::kFloat64;
  }
  throw std::runtime_error("bad dtype code");
}

Field parseField(const char* name, int dtype, int ndim, const int64_t* shape) {
  Field f;
  f.name = name;
  f.dtype = dtype;
  f.shape.assign(shape, shape + ndim);
  auto dot = f.name.find('.');
  if (dot == std::string::npos) {
    f.group = "";
    f.key = f.name;
  } else {
    f.group = f.name.substr(0, dot);
    f.key = f.name.substr(dot + 1);
  }
  return f;
}

torch::Tensor wrap(const Field& f, const void* ptr, int64_t n, bool withBatch, bool alias) {
  std::vector<int64_t> shape;
  if (withBatch) shape.push_back(n);
  for (auto s : f.shape) shape.push_back(s);
  auto t = torch::from_blob(const_cast<void*>(ptr), shape, torch::TensorOptions().dtype(toDtype(f.dtype)));
  return alias ? t : t.clone();
}

void copyOut(const torch::Tensor& t, void* dst) {
  if (dst == nullptr) return;
  auto c = t.contiguous();
  std::memcpy(dst, c.data_ptr(), c.nbytes());
}

// The reference's prefetch runs sample_ on std::async threads; whether a
// prefetched sample_ or the caller's next updatePriority/add takes mSampler_
// first is a race. The tests fix the order to "prefetch completes first"
// (the common case in training, where the learner step is longer than a
// sample): wait for the queued futures before returning to the caller.
template <class Q>
void waitFutures(Q& futures) {
  if (!futures.empty()) {
    futures.front().wait();
    futures.back().wait();
  }
}

struct FFHandle {
  std::vector<Field> fields;
  std::unique_ptr<FFPrioritizedReplay> replay;
};

struct RNNHandle {
  std::vector<Field> fields;
  std::unique_ptr<RNNPrioritizedReplay> replay;
};

}  // namespace

extern "C" {

// ---------------------------------------------------------------- FF replay
void* ref_ff_create(int capacity, int seed, float alpha, float beta, int prefetch,
                    int nfields, const char* const* names, const int* dtypes,
                    const int* ndims, const int64_t* shapes /* nfields x 6 */) {
  auto* h = new FFHandle();
  for (int i = 0; i < nfields; ++i) {
    h->fields.push_back(parseField(names[i], dtypes[i], ndims[i], shapes + 6 * i));
  }
  h->replay.reset(new FFPrioritizedReplay(capacity, seed, alpha, beta, prefetch));
  return h;
}

void ref_ff_destroy(void* hv) {
  delete static_cast<FFHandle*>(hv);
}

// ptrs[f] -> [n, shape_f...] contiguous. alias != 0: the tensors stored in the
// replay alias the caller's memory (what the reference does with the actor's
// batch, types.cc:48-67); the caller must keep it alive.
void ref_ff_add(void* hv, int n, const void* const* ptrs, const float* priority, int alias) {
  auto* h = static_cast<FFHandle*>(hv);
  FFTransition batch;
  for (size_t i = 0; i < h->fields.size(); ++i) {
    const auto& f = h->fields[i];
    auto t = wrap(f, ptrs[i], n, true, alias != 0);
    if (f.group == "obs") batch.obs[f.key] = t;
    else if (f.group == "action") batch.action[f.key] = t;
    else if (f.group == "next_obs") batch.nextObs[f.key] = t;
    else if (f.name == "reward") batch.reward = t;
    else if (f.name == "terminal") batch.terminal = t;
    else if (f.name == "bootstrap") batch.bootstrap = t;
    else throw std::runtime_error("ref_ff_add: unknown field " + f.name);
  }
  auto prio = torch::from_blob(const_cast<float*>(priority), {n}, torch::kFloat32).clone();
  h->replay->add(batch, prio);
}

// Returns the number of sampled ids written (== B). out_ptrs[f] may be null
// (or out_ptrs itself null) to skip the copy-out (timing mode).
int ref_ff_sample(void* hv, int B, void* const* out_ptrs, float* out_w, int* out_ids) {
  auto* h = static_cast<FFHandle*>(hv);
  FFTransition batch;
  torch::Tensor w;
  std::tie(batch, w) = h->replay->sample(B, "cpu");
  waitFutures(h->replay->futures_);
  if (out_ptrs != nullptr) {
    for (size_t i = 0; i < h->fields.size(); ++i) {
      const auto& f = h->fields[i];
      if (f.group == "obs") copyOut(batch.obs.at(f.key), out_ptrs[i]);
      else if (f.group == "action") copyOut(batch.action.at(f.key), out_ptrs[i]);
      else if (f.group == "next_obs") copyOut(batch.nextObs.at(f.key), out_ptrs[i]);
      else if (f.name == "reward") copyOut(batch.reward, out_ptrs[i]);
      else if (f.name == "terminal") copyOut(batch.terminal, out_ptrs[i]);
      else if (f.name == "bootstrap") copyOut(batch.bootstrap, out_ptrs[i]);
    }
  }
  if (out_w != nullptr) copyOut(w, out_w);
  const auto& ids = h->replay->sampledIds_;
  if (out_ids != nullptr) {
    for (size_t i = 0; i < ids.size(); ++i) out_ids[i] = ids[i];
  }
  return (int)ids.size();
}

void ref_ff_update(void* hv, int n, const float* priority) {
  auto* h = static_cast<FFHandle*>(hv);
  auto prio = torch::from_blob(const_cast<float*>(priority), {n}, torch::kFloat32).clone();
  h->replay->updatePriority(prio);
}

int ref_ff_size(void* hv) {
  return static_cast<FFHandle*>(hv)->replay->size();
}

int ref_ff_num_add(void* hv) {
  return static_cast<FFHandle*>(hv)->replay->numAdd();
}

// Private state, for the parity tests only.
void ref_ff_state(void* hv, int* head, int* tail, int* size, int* safeSize, double* sum) {
  auto& q = static_cast<FFHandle*>(hv)->replay->storage_;
  *head = q.head_;
  *tail = q.tail_;
  *size = q.size_;
  *safeSize = q.safeSize_;
  *sum = q.sum_;
}

void ref_ff_weights(void* hv, float* out_w, unsigned char* out_evicted) {
  auto& q = static_cast<FFHandle*>(hv)->replay->storage_;
  for (int i = 0; i < q.capacity; ++i) {
    out_w[i] = q.weights_[i];
    out_evicted[i] = q.evicted_[i] ? 1 : 0;
  }
}

// Timed sample+update loop run entirely in C++ (no ctypes overhead): `cycles`
// iterations of sample(B,"cpu") + updatePriority(prio_pool[c % pool]).
// Returns seconds.
double ref_ff_cycle_loop_dev(void* hv, int B, int cycles, const float* prio_pool, int pool, const char* device);
double ref_ff_cycle_loop(void* hv, int B, int cycles, const float* prio_pool, int pool) {
  return ref_ff_cycle_loop_dev(hv, B, cycles, prio_pool, pool, "cpu");
}

// Same with the `device` argument of sample() given by the caller: "cuda:0" makes the reference
// copy the assembled batch to the learner GPU (types.cc:34-43, pageable host memory), the
// like-for-like "batch on the learner device" number.
double ref_ff_cycle_loop_dev(void* hv, int B, int cycles, const float* prio_pool, int pool, const char* device) {
  auto* h = static_cast<FFHandle*>(hv);
  const std::string dev(device);
  std::vector<torch::Tensor> prios;
  for (int p = 0; p < pool; ++p) {
    prios.push_back(torch::from_blob(const_cast<float*>(prio_pool) + (size_t)p * B, {B}, torch::kFloat32).clone());
  }
  auto t0 = std::chrono::steady_clock::now();
  for (int c = 0; c < cycles; ++c) {
    FFTransition batch;
    torch::Tensor w;
    std::tie(batch, w) = h->replay->sample(B, dev);
    if (dev != "cpu") w = w.cpu();  // the learner's read of the step's result; waits for the copies
    h->replay->updatePriority(prios[c % pool]);
  }
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// --------------------------------------------------------------- RNN replay
void* ref_rnn_create(int capacity, int seed, float alpha, float beta, int prefetch,
                     int nfields, const char* const* names, const int* dtypes,
                     const int* ndims, const int64_t* shapes /* nfields x 6 */) {
  auto* h = new RNNHandle();
  for (int i = 0; i < nfields; ++i) {
    h->fields.push_back(parseField(names[i], dtypes[i], ndims[i], shapes + 6 * i));
  }
  h->replay.reset(new RNNPrioritizedReplay(capacity, seed, alpha, beta, prefetch));
  return h;
}

void ref_rnn_destroy(void* hv) {
  delete static_cast<RNNHandle*>(hv);
}

// ptrs[f] -> n contiguous per-element rows of field f (row e at
// ptrs[f] + e * row_bytes). Elements are added through the reference's
// add(vector<RNNTransition>, priority) (prioritized_replay.h:186), the entry
// point R2D2Actor::postStep uses (r2d2_actor.h:301).
void ref_rnn_add(void* hv, int n, const void* const* ptrs, const float* priority, int alias) {
  auto* h = static_cast<RNNHandle*>(hv);
  std::vector<RNNTransition> vec(n);
  for (size_t i = 0; i < h->fields.size(); ++i) {
    const auto& f = h->fields[i];
    int64_t numel = 1;
    for (auto s : f.shape) numel *= s;
    int64_t rowBytes = numel * (int64_t)torch::elementSize(toDtype(f.dtype));
    for (int e = 0; e < n; ++e) {
      const char* p = static_cast<const char*>(ptrs[i]) + (int64_t)e * rowBytes;
      auto t = wrap(f, p, 1, false, alias != 0);
      auto& el = vec[e];
      if (f.group == "obs") el.obs[f.key] = t;
      else if (f.group == "h0") el.h0[f.key] = t;
      else if (f.group == "action") el.action[f.key] = t;
      else if (f.name == "reward") el.reward = t;
      else if (f.name == "terminal") el.terminal = t;
      else if (f.name == "bootstrap") el.bootstrap = t;
      else if (f.name == "seq_len") el.seqLen = t;
      else throw std::runtime_error("ref_rnn_add: unknown field " + f.name);
    }
  }
  auto prio = torch::from_blob(const_cast<float*>(priority), {n}, torch::kFloat32).clone();
  h->replay->add(vec, prio);
}

int ref_rnn_sample(void* hv, int B, void* const* out_ptrs, float* out_w, int* out_ids) {
  auto* h = static_cast<RNNHandle*>(hv);
  RNNTransition batch;
  torch::Tensor w;
  std::tie(batch, w) = h->replay->sample(B, "cpu");
  waitFutures(h->replay->futures_);
  if (out_ptrs != nullptr) {
    for (size_t i = 0; i < h->fields.size(); ++i) {
      const auto& f = h->fields[i];
      if (f.group == "obs") copyOut(batch.obs.at(f.key), out_ptrs[i]);
      else if (f.group == "h0") copyOut(batch.h0.at(f.key), out_ptrs[i]);
      else if (f.group == "action") copyOut(batch.action.at(f.key), out_ptrs[i]);
      else if (f.name == "reward") copyOut(batch.reward, out_ptrs[i]);
      else if (f.name == "terminal") copyOut(batch.terminal, out_ptrs[i]);
      else if (f.name == "bootstrap") copyOut(batch.bootstrap, out_ptrs[i]);
      else if (f.name == "seq_len") copyOut(batch.seqLen, out_ptrs[i]);
    }
  }
  if (out_w != nullptr) copyOut(w, out_w);
  const auto& ids = h->replay->sampledIds_;
  if (out_ids != nullptr) {
    for (size_t i = 0; i < ids.size(); ++i) out_ids[i] = ids[i];
  }
  return (int)ids.size();
}

void ref_rnn_update(void* hv, int n, const float* priority) {
  auto* h = static_cast<RNNHandle*>(hv);
  auto prio = torch::from_blob(const_cast<float*>(priority), {n}, torch::kFloat32).clone();
  h->replay->updatePriority(prio);
}

int ref_rnn_size(void* hv) {
  return static_cast<RNNHandle*>(hv)->replay->size();
}

int ref_rnn_num_add(void* hv) {
  return static_cast<RNNHandle*>(hv)->replay->numAdd();
}

double ref_rnn_cycle_loop(void* hv, int B, int cycles, const float* prio_pool, int pool) {
  auto* h = static_cast<RNNHandle*>(hv);
  std::vector<torch::Tensor> prios;
  for (int p = 0; p < pool; ++p) {
    prios.push_back(torch::from_blob(const_cast<float*>(prio_pool) + (size_t)p * B, {B}, torch::kFloat32).clone());
  }
  auto t0 = std::chrono::steady_clock::now();
  for (int c = 0; c < cycles; ++c) {
    RNNTransition batch;
    torch::Tensor w;
    std::tie(batch, w) = h->replay->sample(B, "cpu");
    h->replay->updatePriority(prios[c % pool]);
  }
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// ------------------------------------------------- actor-side n-step return
// Drives the reference MultiStepTransitionBuffer (dqn_actor.h:15-124) with
// n+1 history rows of (reward[K], terminal[K]) and pops one transition.
// reward/terminal are [n+1, K] row-major. Outputs are [K].
void ref_nstep_pop(int multiStep, int K, float gamma, const float* reward,
                   const unsigned char* terminal, float* out_reward,
                   float* out_bootstrap, unsigned char* out_terminal) {
  MultiStepTransitionBuffer buf(multiStep, K, gamma);
  for (int s = 0; s <= multiStep; ++s) {
    TensorDict obs, action;
    obs["s"] = torch::full({K, 1}, (float)s);
    action["a"] = torch::full({K}, (int64_t)s, torch::kInt64);
    buf.pushObsAndAction(obs, action);
    auto r = torch::from_blob(const_cast<float*>(reward) + (size_t)s * K, {K}, torch::kFloat32).clone();
    auto t = torch::from_blob(const_cast<unsigned char*>(terminal) + (size_t)s * K, {K}, torch::kUInt8)
                 .to(torch::kBool);
    buf.pushRewardAndTerminal(r, t);
  }
  auto tr = buf.popTransition();
  std::memcpy(out_reward, tr.reward.data_ptr<float>(), sizeof(float) * K);
  std::memcpy(out_bootstrap, tr.bootstrap.data_ptr<float>(), sizeof(float) * K);
  auto term = tr.terminal.to(torch::kUInt8);
  std::memcpy(out_terminal, term.data_ptr<unsigned char>(), K);
}

// Timed loop of the same for the CPU baseline of config 4. Returns seconds
// for `iters` pops (history refilled by one push per pop, as in the actor).
double ref_nstep_loop(int multiStep, int K, float gamma, const float* reward,
                      const unsigned char* terminal, int iters) {
  MultiStepTransitionBuffer buf(multiStep, K, gamma);
  std::vector<torch::Tensor> rs, ts;
  for (int s = 0; s <= multiStep; ++s) {
    rs.push_back(torch::from_blob(const_cast<float*>(reward) + (size_t)s * K, {K}, torch::kFloat32).clone());
    ts.push_back(torch::from_blob(const_cast<unsigned char*>(terminal) + (size_t)s * K, {K}, torch::kUInt8)
                     .to(torch::kBool));
  }
  TensorDict obs, action;
  obs["s"] = torch::zeros({K, 1});
  action["a"] = torch::zeros({K}, torch::kInt64);
  for (int s = 0; s < multiStep; ++s) {
    buf.pushObsAndAction(obs, action);
    buf.pushRewardAndTerminal(rs[s], ts[s]);
  }
  float sink = 0;
  auto t0 = std::chrono::steady_clock::now();
  for (int it = 0; it < iters; ++it) {
    buf.pushObsAndAction(obs, action);
    buf.pushRewardAndTerminal(rs[it % (multiStep + 1)], ts[it % (multiStep + 1)]);
    auto tr = buf.popTransition();
    sink += tr.reward[0].item<float>();
  }
  auto t1 = std::chrono::steady_clock::now();
  if (sink == 12345.678f) std::cout << sink << std::endl;
  return std::chrono::duration<double>(t1 - t0).count();
}

const char* ref_build_info() {
  return "reference rela @ /root/reference, flags " REF_FLAGS ", torch " REF_TORCH_VERSION;
}

}  // extern "C"
