// seq_fixture.cc -- TEST FIXTURE (tests/test_sequence_assembler.py). Feeds the
// same stream of env steps through
//   (a) the reference's R2D2TransitionBuffer (rela/r2d2_actor.h:10-187, compiled
//       where it lies) + aggregate_priority (pyrela/r2d2.py:103-120 as torch ops)
//       + replay.add(std::vector<RNNTransition>, priority), and
//   (b) rela_b200::SequenceAssembler (device-resident windows) + popInto(),
// into two B200 replays, so that the test can compare what the learner samples.
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <torch/extension.h>

#include "rela/r2d2_actor.h"  // reference; pulls in "rela/prioritized_replay.h" = the drop-in header

#include "rela_b200/sequence_assembler.h"

namespace py = pybind11;
using rela::TensorDict;

// (obs, action, reward, terminal, bootstrap, priority, hid)
using Step = std::tuple<TensorDict, TensorDict, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, TensorDict>;

static int feedReference(std::shared_ptr<rela::RNNPrioritizedReplay> replay, int K, int multiStep, int seqLen,
                         int burnin, float eta, const std::vector<Step>& steps) {
  rela::R2D2TransitionBuffer buf(K, multiStep, seqLen, burnin);
  int added = 0;
  for (const Step& s : steps) {
    TensorDict obs = std::get<0>(s), action = std::get<1>(s), none;
    torch::Tensor reward = std::get<2>(s), terminal = std::get<3>(s), bootstrap = std::get<4>(s);
    rela::FFTransition t(obs, action, reward, terminal, bootstrap, none);
    buf.push(t, std::get<5>(s), std::get<6>(s));
    if (!buf.canPop()) continue;
    std::vector<rela::RNNTransition> batch;
    torch::Tensor prio, len;
    std::tie(batch, prio, len) = buf.popTransition();
    // aggregate_priority, pyrela/r2d2.py:113-120
    auto mask = torch::arange(0, prio.size(1)).unsqueeze(0) < len.unsqueeze(1);
    prio = prio * mask.to(torch::kFloat32);
    auto pMean = prio.sum(1) / (len - burnin);
    auto pMax = std::get<0>(prio.max(1));
    auto agg = eta * pMax + (1.0 - eta) * pMean;
    replay->add(batch, agg.to(torch::kFloat32));
    added += (int)batch.size();
  }
  return added;
}

static int feedAssembler(std::shared_ptr<rela::RNNPrioritizedReplay> replay, int K, int multiStep, int seqLen,
                         int burnin, float eta, const std::vector<Step>& steps) {
  rela_b200::SequenceAssembler asmb(replay->core(), K, multiStep, seqLen, burnin, eta);
  int added = 0;
  for (const Step& s : steps) {
    TensorDict none;
    TensorDict obs = std::get<0>(s), action = std::get<1>(s);
    torch::Tensor reward = std::get<2>(s), terminal = std::get<3>(s), bootstrap = std::get<4>(s);
    rela::FFTransition t(obs, action, reward, terminal, bootstrap, none);
    rela_b200::Leaves hid;
    rela_b200::appendDict(hid, "h0", std::get<6>(s));
    asmb.push(rela_b200::FFTraits<rela::FFTransition>::leaves(t), std::get<5>(s), hid);
    if (asmb.canPop()) added += asmb.popInto();
  }
  return added;
}

PYBIND11_MODULE(seq_fixture, m) {
  py::module_::import("rela");
  m.def("feed_reference", &feedReference);
  m.def("feed_assembler", &feedAssembler);
}
