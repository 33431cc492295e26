// issuer_core.cc -- the issuer thread of the C++ host layer.
//
// Reference: with prefetch > 0, PrioritizedReplay::sample hands out the oldest
// future and launches new ones with std::async (rela/prioritized_replay.h:
// 215-230): the batches drawn ahead are produced by other threads while the
// learner works. The B200 replay produces a batch with two kernel launches, so
// "another thread" here is ONE issuer thread that owns the replay's control
// stream and executes, strictly in the order the learner called them,
//
//   kIssue   rr_prefetch_into   selection + gather of a batch drawn ahead, into
//                               pooled batch tensors, + an event behind the gather
//   kTake    rr_take_prefetched the oldest queued batch becomes the current one
//   kUpdate  rr_update_priority host priorities: torch::pow on the CPU (the call
//                               the reference makes, :239) and the update launch
//
// while the learner's thread only waits for the event of the batch it is handed
// (cudaStreamWaitEvent on ITS stream) and builds the tensors it returns. The
// learner's sample() + update_priority() calls then cost what their bookkeeping
// costs; the driver calls run beside the learner's own work (its device-to-host
// read of the loss, the interpreter).
//
// What stays synchronous, i.e. is checked on the calling thread before anything
// is posted, so that the caller sees the reference's errors at the call that
// caused them: sampling before the previous update_priority (:203-206), a
// priority that is not 1-D or of the wrong length (:236-237), an update without
// a sampled batch, a batch size that differs from the batches already drawn
// ahead, and sampling when no batch can be drawn at all (the first kIssue is
// waited for). Errors of later kIssue ops surface when their batch is taken
// (like an exception stored in a future); errors of kTake / kUpdate (a sticky
// device error) surface at the next call.
//
// Device-resident priorities (an extension: the reference requires .cpu()) are
// ordered on the caller's stream, so update_priority(cuda tensor) drains the
// issuer and runs on the calling thread as before.
#include <condition_variable>
#include <cstring>
#include <exception>
#include <stdexcept>
#include <thread>

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

inline void cpuRelax() {
#if defined(__x86_64__) || defined(__i386__)
  __builtin_ia32_pause();
#else
  std::this_thread::yield();
#endif
}

}  // namespace

struct ReplayCore::Issuer {
  enum Kind { kIssue, kTake, kUpdate, kStop };
  struct Op {
    Kind kind = kStop;
    int B = 0;
    void* stream = nullptr;   // the learner's current stream when it called sample()
    std::vector<float> prio;  // kUpdate: a copy of the caller's tensor (it may reuse its buffer)
  };
  struct Result {  // of one kIssue
    BatchSet set;
    std::exception_ptr err;
  };
  std::thread th;
  std::mutex m;  // q, results, err
  std::condition_variable cv_work, cv_done;
  std::deque<Op> q;
  std::deque<Result> results;  // batches drawn ahead, oldest first
  std::atomic<uint64_t> posted{0}, done{0};
  std::atomic<uint64_t> issues_done{0};
  bool sleeping = false;   // under m
  std::exception_ptr err;  // first error of a kTake / kUpdate
  void* own = nullptr;     // the replay's own stream (rr_get_stream)
  // learner-side state: touched by the thread that calls sample / update_priority only
  uint64_t handed = 0;  // kIssue results consumed
  int queued = 0;       // kIssue ops posted whose batch has not been handed out
  int64_t queued_B = 0;
  bool awaiting_update = false;
  int64_t cur_B = 0;

  void post(Op&& op) {
    std::lock_guard<std::mutex> lk(m);
    q.push_back(std::move(op));
    posted.fetch_add(1, std::memory_order_release);
    if (sleeping) cv_work.notify_one();
  }

  template <class Pred>
  void waitUntil(Pred pred) {
    for (int i = 0; i < 20000; ++i) {  // the issuer answers within microseconds
      if (pred()) return;
      cpuRelax();
    }
    std::unique_lock<std::mutex> lk(m);
    cv_done.wait(lk, pred);
  }

  void rethrowDeferred() {
    std::exception_ptr e;
    {
      std::lock_guard<std::mutex> lk(m);
      std::swap(e, err);
    }
    if (e) std::rethrow_exception(e);
  }
};

void ReplayCore::issueInto(BatchSet& set, int batchsize, void* stream) {
  std::vector<void*> ptrs(schema_.size() ? schema_.size() : 1, nullptr);
  for (size_t i = 0; i < schema_.size(); ++i) ptrs[i] = set.fields[i].data_ptr();
  check(rr_prefetch_into(h_, batchsize, ptrs.data(), set.weights.data_ptr<float>(), stream), "sample");
  if (!set.ready) set.ready = newReadyEvent();
  check(rr_record_prefetched(h_, set.ready.get()), "sample");
}

std::shared_ptr<void> ReplayCore::newReadyEvent() const {
  c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
  cudaEvent_t ev = nullptr;
  if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess)
    throw std::runtime_error("sample: cannot create a CUDA event");
  return std::shared_ptr<void>(ev, [](void* e) { cudaEventDestroy(static_cast<cudaEvent_t>(e)); });
}

int ReplayCore::issuerQueued() const { return issuer_ ? issuer_->queued : 0; }

void ReplayCore::startIssuer() {
  const char* e = std::getenv("RELA_B200_ISSUER_THREAD");
  if (prefetch_ <= 0 || shard_.world > 1 || (e && std::atoi(e) == 0)) return;
  issuer_.reset(new Issuer());
  Issuer* is = issuer_.get();
  check(rr_get_stream(h_, &is->own), "PrioritizedReplay");
  is->th = std::thread([this, is]() {
    c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
    for (;;) {
      Issuer::Op op;
      {
        // a learner loop posts every few microseconds: poll before going to sleep
        bool got = false;
        for (int spin = 0; spin < 5000 && !got; ++spin) {
          if (is->posted.load(std::memory_order_acquire) != is->done.load(std::memory_order_relaxed)) {
            std::lock_guard<std::mutex> lk(is->m);
            if (!is->q.empty()) {
              op = std::move(is->q.front());
              is->q.pop_front();
              got = true;
            }
          } else {
            cpuRelax();
          }
        }
        if (!got) {
          std::unique_lock<std::mutex> lk(is->m);
          is->sleeping = true;
          is->cv_work.wait(lk, [&] { return !is->q.empty(); });
          is->sleeping = false;
          op = std::move(is->q.front());
          is->q.pop_front();
        }
      }
      if (op.kind == Issuer::kStop) {
        std::lock_guard<std::mutex> lk(is->m);
        is->done.fetch_add(1, std::memory_order_release);
        is->cv_done.notify_all();
        return;
      }
      Issuer::Result res;
      std::exception_ptr fail;
      try {
        if (op.kind == Issuer::kIssue) {
          {
            std::lock_guard<std::mutex> lk(pool_m_);
            res.set = acquireSet(op.B);
          }
          issueInto(res.set, op.B, op.stream);
        } else if (op.kind == Issuer::kTake) {
          // the learner's stream waits for the batch's own event (issuerSample): nothing to order here
          check(rr_take_prefetched(h_, RR_STREAM_NONE), "sample");
        } else {
          unsigned flags = 0;
          const torch::Tensor p = torch::from_blob(op.prio.data(), {(int64_t)op.prio.size()}, torch::kFloat32);
          const torch::Tensor w = weightsOf(p, &flags);
          check(rr_update_priority(h_, (int)w.size(0), w.data_ptr<float>(), flags, is->own), "update_priority");
        }
      } catch (...) {
        fail = std::current_exception();
      }
      {
        std::lock_guard<std::mutex> lk(is->m);
        if (op.kind == Issuer::kIssue) {
          if (fail) res.set = BatchSet();
          res.err = fail;
          is->results.push_back(std::move(res));
          is->issues_done.fetch_add(1, std::memory_order_release);
        } else if (fail && !is->err) {
          is->err = fail;
        }
        is->done.fetch_add(1, std::memory_order_release);
        is->cv_done.notify_all();
      }
    }
  });
}

void ReplayCore::stopIssuer() {
  if (!issuer_) return;
  Issuer::Op op;
  op.kind = Issuer::kStop;
  issuer_->post(std::move(op));
  if (issuer_->th.joinable()) issuer_->th.join();
  issuer_->results.clear();
  issuer_.reset();
}

void ReplayCore::drainIssuer() {
  if (!issuer_) return;
  Issuer& is = *issuer_;
  is.waitUntil([&] { return is.done.load(std::memory_order_acquire) == is.posted.load(std::memory_order_acquire); });
  is.rethrowDeferred();
}

std::pair<Leaves, torch::Tensor> ReplayCore::issuerSample(int batchsize, const std::string& device) {
  Issuer& is = *issuer_;
  is.rethrowDeferred();
  if (is.awaiting_update)  // assert(sampledIds_.empty()), rela/prioritized_replay.h:203-206
    throw std::runtime_error("sample: previous samples' priority has not been updated");
  const int64_t B = batchsize;
  if (B <= 0) throw std::runtime_error("sample: bad batch size");
  void* stream = (void*)c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)device_).stream();
  auto issue = [&]() {
    Issuer::Op op;
    op.kind = Issuer::kIssue;
    op.B = batchsize;
    op.stream = stream;
    is.post(std::move(op));
    ++is.queued;
    is.queued_B = B;
  };
  if (is.queued > 0 && is.queued_B != B) throw std::runtime_error("sample: batch size differs from the prefetched one");
  if (is.queued == 0) issue();
  // the oldest batch drawn ahead: normally issued `prefetch` samples ago
  is.waitUntil([&] { return is.issues_done.load(std::memory_order_acquire) > is.handed; });
  Issuer::Result res;
  {
    std::lock_guard<std::mutex> lk(is.m);
    res = std::move(is.results.front());
    is.results.pop_front();
  }
  ++is.handed;
  --is.queued;
  if (res.err) std::rethrow_exception(res.err);
  {
    c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
    if (cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), static_cast<cudaEvent_t>(res.set.ready.get()), 0) !=
        cudaSuccess)
      throw std::runtime_error("sample: cudaStreamWaitEvent failed");
  }
  Issuer::Op take;
  take.kind = Issuer::kTake;
  is.post(std::move(take));
  is.awaiting_update = true;
  is.cur_B = B;
  while (is.queued < prefetch_) issue();  // before this batch's updatePriority, like the reference's futures
  return handOut(res.set, device);
}

void ReplayCore::issuerAdopt(BatchSet set, void* stream) {
  // a batch the saved run had drawn ahead (load): its bytes were restored on `stream`
  Issuer& is = *issuer_;
  if (!set.ready) set.ready = newReadyEvent();
  {
    c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_);
    if (cudaEventRecord(static_cast<cudaEvent_t>(set.ready.get()), static_cast<cudaStream_t>(stream)) != cudaSuccess)
      throw std::runtime_error("load: cannot record a CUDA event");
  }
  is.queued_B = set.B;
  Issuer::Result res;
  res.set = std::move(set);
  {
    std::lock_guard<std::mutex> lk(is.m);
    is.results.push_back(std::move(res));
  }
  is.issues_done.fetch_add(1, std::memory_order_release);
  ++is.queued;
}

void ReplayCore::issuerUpdate(const torch::Tensor& priority) {
  Issuer& is = *issuer_;
  is.rethrowDeferred();
  if (priority.dim() != 1)  // assert(priority.dim() == 1), rela/prioritized_replay.h:236
    throw std::runtime_error("priority must be a 1-D tensor");
  if (priority.scalar_type() != torch::kFloat32) throw std::runtime_error("priority must be float32");
  if (!is.awaiting_update) throw std::runtime_error("update_priority: no sampled batch is outstanding");
  if (priority.size(0) != is.cur_B)  // assert((int)sampledIds_.size() == priority.size(0)), :237
    throw std::runtime_error("update_priority: " + std::to_string(priority.size(0)) + " priorities for a batch of " +
                             std::to_string(is.cur_B));
  if (priority.is_cuda()) {
    // ordered on the caller's stream: issued by the caller, after everything posted so far
    drainIssuer();
    unsigned flags = 0;
    const torch::Tensor w = weightsOf(priority, &flags);
    check(rr_update_priority(h_, (int)w.size(0), w.data_ptr<float>(), flags,
                             (void*)c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)device_).stream()),
          "update_priority");
    is.awaiting_update = false;
    return;
  }
  const torch::Tensor p = priority.detach().contiguous();
  Issuer::Op op;
  op.kind = Issuer::kUpdate;
  op.prio.assign(p.data_ptr<float>(), p.data_ptr<float>() + p.size(0));
  is.post(std::move(op));
  is.awaiting_update = false;
}

}  // namespace rela_b200
