// sequence_assembler.h -- R2D2 sequence assembly with the sequence windows
// resident in HBM. Same state machine as the
// reference's R2D2TransitionBuffer (rela/r2d2_actor.h:10-187): per env a window
// of T = burnin + seqLen + multiStep steps, burn-in padding at episode start,
// zero padding after a terminal step, carry-over of the last burnin + multiStep
// steps into the next window, and the immediate second sequence when the
// carried steps already contain a terminal. What differs is where the bytes
// live: the reference keeps T tensor handles per env and torch::stack()s them
// on the host into a 3.5 MB RNNTransition that is then copied again by
// makeBatch; here every env step is uploaded ONCE, scattered into a device
// window [K, T, row] by the row-copy kernel (rr_scatter_rows), and a finished
// window is moved into the replay slab device-to-device (rr_add_rows with
// RR_ROWS_DEVICE).
//
// Not a drop-in for the unmodified R2D2Actor (it owns its R2D2TransitionBuffer
// by value, rela/r2d2_actor.h:346); an actor adopts it by replacing
//   r2d2Buffer_.push(transition, priority, hid);
//   if (r2d2Buffer_.canPop()) { pop; aggregate; replayBuffer_->add(batch, p); }
// with push(...) and popInto().
//
// Known reference defect, NOT reproduced: when carrying the per-step
// priorities over, r2d2_actor.h:139 reads seqPriority[seqLen + j] with j
// counted from burnin in a vector of seqLen + multiStep entries, i.e. out of
// bounds whenever burnin > 0. The assembler carries seqPriority[seqLen + j -
// burnin] (the priorities of the carried steps), which is what line 76 stored.
#pragma once

#include <vector>

#include <ATen/cuda/CUDAEvent.h>

#include "rela_b200/replay_core.h"

namespace rela_b200 {

class SequenceAssembler {
 public:
  SequenceAssembler(ReplayCore& replay, int batchsize, int multiStep, int seqLen, int burnin, float eta);

  // One step of the K envs. `step`: the FFTransition leaves without next_obs
  // (obs.*, action.*, reward, terminal, bootstrap), each [K, ...] on the host;
  // `priority`: per-step priority [K] float32; `hid`: the hidden state BEFORE
  // this step, leaves "h0.<name>" of shape [layers, K, hid].
  // A HOST priority tensor (what the reference's actor passes after `.cpu()`,
  // r2d2_actor.h:331-341) keeps the per-step values on the host and mixes them
  // there in float, like the reference. A CUDA priority tensor keeps them in a
  // device window and the eta-mix runs on the device (rr_aggregate_priority):
  // nothing of the priority path touches the host. One assembler uses one mode.
  void push(const Leaves& step, const torch::Tensor& priority, const Leaves& hid);

  bool canPop() const { return can_pop_; }

  // Moves every finished sequence into the replay (device to device), with the
  // eta-mix priority eta * max + (1 - eta) * mean over each sequence
  // (pyrela/r2d2.py:103-120). Returns the number of sequences added.
  int popInto();

  int sequenceSteps() const { return T_; }

 private:
  struct Field {
    std::string name;
    torch::ScalarType dtype;
    std::vector<int64_t> shape;  // per step
    uint64_t row_bytes;
    torch::Tensor window;  // [K * T, row] u8, device
    torch::Tensor upload;  // [K, row] u8, device (this step's rows)
  };
  struct Hidden {
    std::string name;
    std::vector<int64_t> shape;  // [layers, hid]
    torch::ScalarType dtype;
    uint64_t row_bytes;
    torch::Tensor cur, next;  // [K, row] u8, device
  };

  void setup(const Leaves& step, const Leaves& hid);
  void zeroRows(int env, int from, int to);

  ReplayCore& replay_;
  const int K_, multiStep_, seqLen_, burnin_, T_;
  const float eta_;
  bool ready_ = false, can_pop_ = false;
  std::vector<Field> fields_;
  std::vector<Hidden> hidden_;
  std::vector<int> next_idx_, len_;
  std::vector<std::vector<float>> seq_priority_;   // [K][seqLen + multiStep]
  std::vector<std::vector<uint8_t>> terminal_;     // [K][T] host copy of the terminal flags
  torch::Tensor dst_rows_host_, dst_rows_dev_;     // scatter destinations of one step
  torch::Tensor seq_len_host_, seq_len_dev_;       // [2K] float: seq_len of the sequences of one pop
  // the pinned staging buffers above are rewritten every step / pop: each remembers the copy
  // that last read it and the host waits for that copy before writing again
  at::cuda::CUDAEvent rows_copied_, len_copied_;
  int prio_mode_ = -1;                             // 0: host priorities, 1: device priorities
  torch::Tensor prio_window_;                      // [K, seqLen + multiStep] float32, device (mode 1)
  torch::Tensor env_offset_;                       // [K] int64, device: (i + 1) * burnin
  std::vector<std::pair<int, int>> elem_src_;      // schema leaf -> (0 field | 1 hidden | 2 seq_len, index)
  std::vector<torch::Tensor> keep_;                // spill windows alive until the add has been issued
};

}  // namespace rela_b200
