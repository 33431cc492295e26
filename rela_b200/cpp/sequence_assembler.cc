// sequence_assembler.cc -- see rela_b200/sequence_assembler.h. Line references
// are to the reference's R2D2TransitionBuffer, rela/r2d2_actor.h.
#include "rela_b200/sequence_assembler.h"

#include <algorithm>
#include <stdexcept>

#include <c10/cuda/CUDAStream.h>

#include "rela_b200.h"

namespace rela_b200 {

namespace {

uint64_t bytesOf(const std::vector<int64_t>& shape, torch::ScalarType dt) {
  uint64_t n = 1;
  for (auto s : shape) n *= (uint64_t)s;
  return n * (uint64_t)c10::elementSize(dt);
}

bool startsWith(const std::string& s, const char* prefix) {
  return s.rfind(prefix, 0) == 0;
}

torch::Tensor deviceBytes(int64_t rows, uint64_t row_bytes, int device) {
  return torch::zeros({rows, (int64_t)std::max<uint64_t>(row_bytes, 1)},
                      torch::TensorOptions().dtype(torch::kUInt8).device(torch::kCUDA, (c10::DeviceIndex)device));
}

}  // namespace

SequenceAssembler::SequenceAssembler(ReplayCore& replay, int batchsize, int multiStep, int seqLen, int burnin,
                                     float eta)
    : replay_(replay)
    , K_(batchsize)
    , multiStep_(multiStep)
    , seqLen_(seqLen)
    , burnin_(burnin)
    , T_(burnin + seqLen + multiStep)
    , eta_(eta)
    , next_idx_(batchsize, 0)
    , len_(batchsize, 0)
    , seq_priority_(batchsize, std::vector<float>(seqLen + multiStep, 0.0f))
    , terminal_(batchsize, std::vector<uint8_t>(burnin + seqLen + multiStep, 0)) {
  if (burnin > seqLen || multiStep > seqLen)  // asserts at :26-27
    throw std::runtime_error("SequenceAssembler: burnin and multiStep must not exceed seqLen");
}

void SequenceAssembler::setup(const Leaves& step, const Leaves& hid) {
  const int dev = replay_.device();
  std::vector<LeafSpec> specs;
  auto specOf = [&](const std::string& name, torch::ScalarType dt, std::vector<int64_t> shape, bool tm) {
    LeafSpec s;
    s.name = name;
    s.dtype = dt;
    s.shape = std::move(shape);
    s.time_major = tm;
    s.chunks = tm ? (uint32_t)s.shape[0] : 1u;
    s.row_bytes = bytesOf(s.shape, dt);
    return s;
  };
  // element leaf order of RNNTraits: obs.*, h0.*, action.*, reward, terminal, bootstrap, seq_len
  std::vector<LeafSpec> obs_specs, rest_specs;
  for (const Leaf& lf : step) {
    if (startsWith(lf.name, "next_obs.")) continue;  // RNN elements carry no next_obs (rela/types.cc:104-110)
    if (lf.t.dim() < 1 || lf.t.size(0) != K_)
      throw std::runtime_error("SequenceAssembler: leaf '" + lf.name + "' is not [K, ...]");
    Field f;
    f.name = lf.name;
    f.dtype = lf.t.scalar_type();
    f.shape.assign(lf.t.sizes().begin() + 1, lf.t.sizes().end());
    f.row_bytes = bytesOf(f.shape, f.dtype);
    f.window = deviceBytes((int64_t)K_ * T_, f.row_bytes, dev);
    f.upload = deviceBytes(K_, f.row_bytes, dev);
    std::vector<int64_t> eshape{T_};
    eshape.insert(eshape.end(), f.shape.begin(), f.shape.end());
    (startsWith(f.name, "obs.") ? obs_specs : rest_specs).push_back(specOf(f.name, f.dtype, eshape, true));
    fields_.push_back(std::move(f));
  }
  std::vector<LeafSpec> hid_specs;
  for (const Leaf& lf : hid) {
    if (lf.t.dim() != 3 || lf.t.size(1) != K_)
      throw std::runtime_error("SequenceAssembler: hidden '" + lf.name + "' is not [layers, K, hid]");
    Hidden h;
    h.name = lf.name;
    h.dtype = lf.t.scalar_type();
    h.shape = {lf.t.size(0), lf.t.size(2)};
    h.row_bytes = bytesOf(h.shape, h.dtype);
    h.cur = deviceBytes(K_, h.row_bytes, dev);
    h.next = deviceBytes(K_, h.row_bytes, dev);
    hid_specs.push_back(specOf(h.name, h.dtype, h.shape, true));
    hidden_.push_back(std::move(h));
  }
  specs = obs_specs;
  specs.insert(specs.end(), hid_specs.begin(), hid_specs.end());
  specs.insert(specs.end(), rest_specs.begin(), rest_specs.end());
  specs.push_back(specOf("seq_len", torch::kFloat32, {1}, false));
  replay_.setSchema(specs);
  for (const LeafSpec& sp : specs) {
    int kind = 2, idx = 0;
    for (size_t i = 0; i < fields_.size(); ++i)
      if (fields_[i].name == sp.name) kind = 0, idx = (int)i;
    for (size_t i = 0; i < hidden_.size(); ++i)
      if (hidden_[i].name == sp.name) kind = 1, idx = (int)i;
    elem_src_.push_back({kind, idx});
  }
  seq_len_host_ = torch::empty({2 * K_}, torch::TensorOptions().dtype(torch::kFloat32).pinned_memory(true));
  seq_len_dev_ = torch::empty({2 * K_}, torch::TensorOptions().dtype(torch::kFloat32).device(torch::kCUDA, (c10::DeviceIndex)dev));
  dst_rows_host_ = torch::empty({K_}, torch::TensorOptions().dtype(torch::kInt32).pinned_memory(true));
  dst_rows_dev_ = torch::empty({K_}, torch::TensorOptions().dtype(torch::kInt32).device(torch::kCUDA, (c10::DeviceIndex)dev));
  ready_ = true;
}

void SequenceAssembler::zeroRows(int env, int from, int to) {
  if (from >= to) return;
  // FFTransition::padLike (rela/types.cc:69-80): zeros everywhere, terminal = true
  for (Field& f : fields_) {
    torch::Tensor rows = f.window.narrow(0, (int64_t)env * T_ + from, to - from);
    if (f.name == "terminal") rows.fill_(1); else rows.zero_();
  }
  for (int t = from; t < to; ++t) terminal_[env][t] = 0;
}

void SequenceAssembler::push(const Leaves& step, const torch::Tensor& priority, const Leaves& hid) {
  if (!ready_) setup(step, hid);
  if (priority.dim() != 1 || priority.size(0) != K_ || priority.scalar_type() != torch::kFloat32)
    throw std::runtime_error("SequenceAssembler::push: priority must be float32 [K]");  // assert at :32
  const int mode = priority.is_cuda() ? 1 : 0;
  if (prio_mode_ < 0) prio_mode_ = mode;
  if (mode != prio_mode_) throw std::runtime_error("SequenceAssembler::push: host and device priorities mixed");
  const torch::Device home(torch::kCUDA, (c10::DeviceIndex)replay_.device());
  const torch::Tensor prio = mode == 0 ? priority.contiguous() : torch::Tensor();
  const float* pa = mode == 0 ? prio.data_ptr<float>() : nullptr;
  if (mode == 1 && !prio_window_.defined()) {
    prio_window_ = torch::zeros({K_, seqLen_ + multiStep_}, torch::TensorOptions().dtype(torch::kFloat32).device(home));
    env_offset_ = (torch::arange(K_, torch::TensorOptions().dtype(torch::kInt64).device(home)) + 1) * burnin_;
  }
  bool pinned_src = false;  // a pinned source is copied asynchronously: the caller may reuse it after push()

  // upload this step's rows once (one H2D copy per leaf) ...
  const uint8_t* term_host = nullptr;
  torch::Tensor term_keep;
  size_t fi = 0;
  for (const Leaf& lf : step) {
    if (startsWith(lf.name, "next_obs.")) continue;
    Field& f = fields_.at(fi++);
    if (lf.name != f.name || lf.t.scalar_type() != f.dtype) throw std::runtime_error("SequenceAssembler::push: leaf order changed");
    torch::Tensor src = lf.t.contiguous();
    if (lf.name == "terminal") {
      term_keep = src.cpu().to(torch::kUInt8);
      term_host = term_keep.data_ptr<uint8_t>();
    }
    if (f.row_bytes == 0) continue;
    torch::Tensor flat = src.view(torch::kUInt8).reshape({K_, (int64_t)f.row_bytes});
    pinned_src = pinned_src || (!src.is_cuda() && src.is_pinned());
    f.upload.copy_(flat, /*non_blocking=*/true);
  }
  if (term_host == nullptr) throw std::runtime_error("SequenceAssembler::push: no 'terminal' leaf");

  // ... and decide, per env, where it goes (the reference's push loop, :35-86)
  rows_copied_.synchronize();  // the previous step's copy out of this pinned buffer
  int32_t* dst = dst_rows_host_.data_ptr<int32_t>();
  std::vector<int> finished;
  for (int i = 0; i < K_; ++i) {
    if (next_idx_[i] == 0) {
      // fresh sequence: its h0 is the hidden state of this step (:38-39), burn-in rows are padding (:44-47)
      for (size_t h = 0; h < hidden_.size(); ++h) {
        torch::Tensor row = hid[h].t.select(1, i).contiguous().view(torch::kUInt8).reshape({(int64_t)hidden_[h].row_bytes});
        pinned_src = pinned_src || (!row.is_cuda() && row.is_pinned());
        hidden_[h].cur.select(0, i).copy_(row, true);
      }
      zeroRows(i, 0, burnin_);
      next_idx_[i] = burnin_;
    } else if (len_[i] != 0 || terminal_[i][next_idx_[i] - 1]) {
      throw std::runtime_error("SequenceAssembler::push: step after a finished sequence; call popInto() first");  // :52-56
    }
    const int idx = next_idx_[i];
    if (idx == seqLen_) {  // becomes the h0 of the next window (:65-68)
      for (size_t h = 0; h < hidden_.size(); ++h) {
        torch::Tensor row = hid[h].t.select(1, i).contiguous().view(torch::kUInt8).reshape({(int64_t)hidden_[h].row_bytes});
        pinned_src = pinned_src || (!row.is_cuda() && row.is_pinned());
        hidden_[h].next.select(0, i).copy_(row, true);
      }
    }
    dst[i] = i * T_ + idx;
    terminal_[i][idx] = term_host[i];
    if (mode == 0) seq_priority_[i][idx - burnin_] = pa[i];
    ++next_idx_[i];
    if (!term_host[i] && next_idx_[i] < T_) continue;
    // sequence finished: remember its length, pad the rest (:78-85)
    len_[i] = next_idx_[i];
    if (mode == 0) {
      for (int t = next_idx_[i]; t < T_; ++t) seq_priority_[i][t - burnin_] = 0.0f;
    } else {
      finished.push_back(i);
    }
    can_pop_ = true;
  }
  dst_rows_dev_.copy_(dst_rows_host_, true);
  rows_copied_.record(c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)replay_.device()));
  if (mode == 1) {
    // this step's priorities into the device window: column idx - burnin of row i, i.e. flat index
    // (i * T + idx) - (i + 1) * burnin; then the padding behind finished sequences
    torch::Tensor p = priority.to(home).to(torch::kFloat32).contiguous();
    prio_window_.view({-1}).index_copy_(0, dst_rows_dev_.to(torch::kInt64) - env_offset_, p);
    for (int i : finished)
      if (next_idx_[i] < T_) prio_window_.select(0, i).narrow(0, next_idx_[i] - burnin_, T_ - next_idx_[i]).zero_();
  }
  std::vector<uint64_t> rb;
  std::vector<const void*> src;
  std::vector<void*> dstp;
  for (Field& f : fields_) {
    if (f.row_bytes == 0) continue;
    rb.push_back(f.row_bytes);
    src.push_back(f.upload.data_ptr());
    dstp.push_back(f.window.data_ptr());
  }
  replay_.scatterRows(K_, rb, src, dstp, dst_rows_dev_.data_ptr<int32_t>());
  // terminal padding comes after the scatter in stream order, and only covers rows behind it
  for (int i = 0; i < K_; ++i) {
    if (len_[i] != 0 && next_idx_[i] < T_) {
      zeroRows(i, next_idx_[i], T_);
      next_idx_[i] = T_;
    }
  }
  if (pinned_src) c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)replay_.device()).synchronize();
}

int SequenceAssembler::popInto() {
  if (!can_pop_) throw std::runtime_error("SequenceAssembler::popInto: nothing to pop");  // assert at :94
  struct Emit {
    int env;
    bool second;
    float len;
    std::vector<float> prio;             // [seqLen] (host priorities)
    torch::Tensor prio_dev;              // [seqLen] (device priorities)
    std::vector<torch::Tensor> spill;    // per field, second sequences only
  };
  const bool dev_prio = prio_mode_ == 1;
  const torch::Device home(torch::kCUDA, (c10::DeviceIndex)replay_.device());
  std::vector<Emit> emits;
  std::vector<int> carry;  // envs whose window continues
  const int c = burnin_ + multiStep_;
  keep_.clear();
  for (int i = 0; i < K_; ++i) {
    if (len_[i] == 0) continue;
    const int L = std::min(len_[i], burnin_ + seqLen_);  // :107
    Emit m{i, false, (float)L, std::vector<float>(seq_priority_[i].begin(), seq_priority_[i].begin() + seqLen_),
           torch::Tensor(), {}};
    if (dev_prio) m.prio_dev = prio_window_.select(0, i).narrow(0, 0, seqLen_).clone();
    emits.push_back(std::move(m));
    if (terminal_[i][L - 1]) {
      next_idx_[i] = 0;  // episode ended: fresh start (:118-121)
    } else {
      // the last burnin + multiStep steps open the next window (:123-141)
      int len2 = -1;
      std::vector<float> carried(seqLen_ + multiStep_, 0.0f);
      for (int j = burnin_; j < c; ++j) {
        carried[j - burnin_] = seq_priority_[i][seqLen_ + j - burnin_];
        if (terminal_[i][seqLen_ + j] && len2 == -1) len2 = j + 1;
      }
      if (len2 != -1) {
        // the carried steps already end the episode: emit them at once as a
        // second, padded sequence (:146-162); built from the window as it is now
        Emit e{i, true, (float)len2, std::vector<float>(carried.begin(), carried.begin() + seqLen_), torch::Tensor(), {}};
        if (dev_prio) {
          e.prio_dev = torch::zeros({seqLen_}, torch::TensorOptions().dtype(torch::kFloat32).device(home));
          e.prio_dev.narrow(0, 0, multiStep_).copy_(prio_window_.select(0, i).narrow(0, seqLen_, multiStep_));
        }
        for (Field& f : fields_) {
          torch::Tensor sp = deviceBytes(T_, f.row_bytes, replay_.device());
          if (f.name == "terminal") sp.fill_(1);  // padLike
          if (f.row_bytes) sp.narrow(0, 0, c).copy_(f.window.narrow(0, (int64_t)i * T_ + seqLen_, c), true);
          e.spill.push_back(sp);
          keep_.push_back(sp);
        }
        emits.push_back(std::move(e));
        next_idx_[i] = 0;
      } else {
        carry.push_back(i);
        for (int j = 0; j < c; ++j) terminal_[i][j] = terminal_[i][seqLen_ + j];
        for (int j = 0; j < multiStep_; ++j) seq_priority_[i][j] = carried[j];
        if (dev_prio)
          prio_window_.select(0, i).narrow(0, 0, multiStep_).copy_(prio_window_.select(0, i).narrow(0, seqLen_, multiStep_).clone());
        next_idx_[i] = c;
      }
    }
    len_[i] = 0;
  }
  can_pop_ = false;

  // one add for all finished sequences, rows read straight from the windows
  const int n = (int)emits.size();
  const size_t nf = elem_src_.size();
  std::vector<const void*> ptrs(nf * (size_t)n, nullptr);
  torch::Tensor prio = torch::empty({n}, torch::kFloat32);
  float* pp = prio.data_ptr<float>();
  len_copied_.synchronize();  // the previous pop's copy out of this pinned buffer
  float* sl = seq_len_host_.data_ptr<float>();
  for (int e = 0; e < n; ++e) {
    const Emit& m = emits[e];
    sl[e] = m.len;
    if (dev_prio) continue;
    // eta * max + (1 - eta) * mean over the sequence (pyrela/r2d2.py:103-120); len counts the burn-in
    float sum = 0.0f, mx = 0.0f;
    for (int j = 0; j < seqLen_; ++j) {
      const float v = (float)j < m.len ? m.prio[j] : 0.0f;
      sum += v;
      mx = j == 0 ? v : std::max(mx, v);
    }
    pp[e] = eta_ * mx + (1.0f - eta_) * (sum / (m.len - (float)burnin_));
  }
  seq_len_dev_.narrow(0, 0, n).copy_(seq_len_host_.narrow(0, 0, n), true);
  len_copied_.record(c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)replay_.device()));
  if (dev_prio) {
    // the same mix on the device, one warp per sequence (K9, rela_b200/csrc/actor_math.cu)
    std::vector<torch::Tensor> rows;
    rows.reserve((size_t)n);
    for (const Emit& m : emits) rows.push_back(m.prio_dev);
    torch::Tensor per_step = torch::stack(rows, 0).contiguous();
    prio = torch::empty({n}, torch::TensorOptions().dtype(torch::kFloat32).device(home));
    const int rc = rr_aggregate_priority(n, seqLen_, per_step.data_ptr<float>(), seq_len_dev_.data_ptr<float>(), eta_,
                                         burnin_, prio.data_ptr<float>(), replay_.stream());
    if (rc != RR_OK) throwLast(rc, "SequenceAssembler::popInto");
    keep_.push_back(per_step);
  }
  for (int e = 0; e < n; ++e) {
    const Emit& m = emits[e];
    for (size_t f = 0; f < nf; ++f) {
      const auto& src = elem_src_[f];
      const void* p = nullptr;
      if (src.first == 0) {
        const Field& fd = fields_[src.second];
        p = m.second ? m.spill[src.second].data_ptr()
                     : static_cast<const uint8_t*>(fd.window.data_ptr()) + (size_t)m.env * T_ * fd.row_bytes;
      } else if (src.first == 1) {
        const Hidden& h = hidden_[src.second];
        p = static_cast<const uint8_t*>((m.second ? h.next : h.cur).data_ptr()) + (size_t)m.env * h.row_bytes;
      } else {
        p = seq_len_dev_.data_ptr<float>() + e;
      }
      ptrs[f * (size_t)n + (size_t)e] = p;
    }
  }
  replay_.addDeviceRows(n, ptrs, prio);

  // now the windows may change: carry-over, through a temporary because the
  // source and destination rows can overlap when seqLen < burnin + multiStep
  for (int i : carry) {
    for (Field& f : fields_) {
      if (!f.row_bytes) continue;
      torch::Tensor tmp = f.window.narrow(0, (int64_t)i * T_ + seqLen_, c).clone();
      f.window.narrow(0, (int64_t)i * T_, c).copy_(tmp, true);
    }
    for (Hidden& h : hidden_) h.cur.select(0, i).copy_(h.next.select(0, i), true);  // batchH0_ = batchNextH0_ (:144)
  }
  return n;
}

}  // namespace rela_b200
