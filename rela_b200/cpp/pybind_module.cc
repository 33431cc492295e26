// pybind_module.cc -- Python module rela_b200._C: the replay classes of the
// reference's `rela` module (rela/pybind.cc:19-59) with identical names,
// constructor arguments and methods, served by the B200 replay.
//
//   rela.FFTransition / rela.RNNTransition        pybind.cc:20-35
//   rela.FFPrioritizedReplay(capacity, seed, alpha, beta, prefetch)
//     .size() .num_add() .sample(batchsize, device) .update_priority(priority)
//                                                  pybind.cc:37-59
// Additions (the reference binds no producer side; its actors call add from
// C++): .add(...), .sync(), .handle(), .schema(), a `device` keyword, and the
// keywords of a replay striped over several GPUs (shard_rank, shard_world,
// learner, group: ShardConfig in rela_b200/replay_core.h). The actor-side
// arithmetic kernels are module functions: nstep_return, td_priority,
// aggregate_priority.
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <torch/csrc/DynamicTypes.h>
#include <torch/extension.h>

#include <c10/cuda/CUDAStream.h>

#include "rela_b200.h"
#include "rela_b200/sequence_assembler.h"
#include "rela_b200/types.h"

namespace py = pybind11;
using namespace rela_b200;

namespace {

template <class Replay>
py::list schemaOf(Replay& r) {
  py::list out;
  for (const LeafSpec& s : r.core().schema()) {
    out.append(py::make_tuple(s.name, py::reinterpret_borrow<py::object>((PyObject*)torch::getTHPDtype(s.dtype)),
                              py::cast(s.shape), s.chunks, s.row_bytes, s.time_major));
  }
  return out;
}

template <class Replay, class PyClass>
void bindCommon(PyClass& c) {
  c.def(py::init<int, int, float, float, int>(), py::arg("capacity"), py::arg("seed"), py::arg("alpha"),
        py::arg("beta"), py::arg("prefetch"))
      .def(py::init<int, int, float, float, int, int>(), py::arg("capacity"), py::arg("seed"), py::arg("alpha"),
           py::arg("beta"), py::arg("prefetch"), py::arg("device"))
      .def(py::init([](int capacity, int seed, float alpha, float beta, int prefetch, int device, int shard_rank,
                       int shard_world, int learner, const std::string& group) {
             ShardConfig sc;
             sc.rank = shard_rank;
             sc.world = shard_world;
             sc.learner = learner;
             sc.group = group;
             return std::make_shared<Replay>(capacity, seed, alpha, beta, prefetch, device, sc);
           }),
           py::arg("capacity"), py::arg("seed"), py::arg("alpha"), py::arg("beta"), py::arg("prefetch"),
           py::arg("device"), py::arg("shard_rank"), py::arg("shard_world"), py::arg("learner") = -1,
           py::arg("group") = "rela_b200")
      .def("size", &Replay::size)
      .def("num_add", &Replay::numAdd)
      .def("sample", &Replay::sample, py::call_guard<py::gil_scoped_release>())
      .def("update_priority", &Replay::updatePriority, py::call_guard<py::gil_scoped_release>())
      .def("sync", [](Replay& r) { r.core().sync(); }, py::call_guard<py::gil_scoped_release>())
      // measurement: `steps` sample + update_priority cycles issued by a C++ caller (what a C++
      // learner's loop costs, without the interpreter between the calls); priority_pool [k, B]
      .def("cycle_loop",
           [](Replay& r, int batchsize, const std::string& device, const torch::Tensor& priority_pool, int steps) {
             const int64_t k = priority_pool.size(0);
             for (int s = 0; s < steps; ++s) {
               auto batch = r.sample(batchsize, device);
               r.updatePriority(priority_pool[s % k]);
             }
           },
           py::arg("batchsize"), py::arg("device"), py::arg("priority_pool"), py::arg("steps"),
           py::call_guard<py::gil_scoped_release>())
      .def("set_output_format",
           [](Replay& r, const std::string& leaf, py::object dtype) {
             r.core().setOutputFormat(leaf, torch::python::detail::py_object_to_dtype(dtype));
           },
           py::arg("leaf"), py::arg("dtype"))
      .def("save", [](Replay& r, const std::string& path) { r.core().save(path); },
           py::call_guard<py::gil_scoped_release>())
      .def("load", [](Replay& r, const std::string& path) { r.core().load(path); },
           py::call_guard<py::gil_scoped_release>())
      .def("handle", [](Replay& r) { return (uintptr_t)r.core().handle(); })
      .def("last_chunk", [](Replay& r) { return r.core().lastChunk(); })
      .def("device", [](Replay& r) { return r.core().device(); })
      .def("schema", [](Replay& r) { return schemaOf(r); });
}

}  // namespace

namespace {

void* currentStream(const torch::Tensor& t) {
  return (void*)c10::cuda::getCurrentCUDAStream(t.device().index()).stream();
}

void checkActor(int rc, const char* where) {
  if (rc != RR_OK) throwLast(rc, where);
}

// MultiStepTransitionBuffer::popTransition arithmetic (rela/dqn_actor.h:58-106): reward float32
// [n + 1, K], terminal bool / uint8 [n + 1, K] on CUDA -> (return [K], bootstrap [K], terminal [K] bool)
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> nstepReturn(const torch::Tensor& reward,
                                                                    const torch::Tensor& terminal, double gamma,
                                                                    bool contract_fma) {
  if (!reward.is_cuda() || reward.scalar_type() != torch::kFloat32 || reward.dim() != 2 || !terminal.is_cuda() ||
      terminal.sizes() != reward.sizes())
    throw std::runtime_error("nstep_return: reward float32 [n + 1, K] and terminal [n + 1, K] on CUDA");
  const torch::Tensor r = reward.contiguous();
  torch::Tensor t = terminal.contiguous();
  if (t.scalar_type() == torch::kBool) t = t.view(torch::kUInt8);
  if (t.scalar_type() != torch::kUInt8) throw std::runtime_error("nstep_return: terminal must be bool or uint8");
  const int64_t K = r.size(1);
  torch::Tensor out_r = torch::empty({K}, r.options());
  torch::Tensor out_b = torch::empty({K}, r.options());
  torch::Tensor out_t = torch::empty({K}, t.options());
  checkActor(rr_nstep_return((int)r.size(0) - 1, (int)K, (float)gamma, r.data_ptr<float>(), t.data_ptr<uint8_t>(),
                             out_r.data_ptr<float>(), out_b.data_ptr<float>(), out_t.data_ptr<uint8_t>(),
                             contract_fma ? 1 : 0, currentStream(r)),
             "nstep_return");
  return std::make_tuple(out_r, out_b, out_t.view(torch::kBool));
}

// |reward + bootstrap * gamma^n * q_next - q_sa| (pyrela/apex.py:44-45,77-78)
torch::Tensor tdPriority(const torch::Tensor& reward, const torch::Tensor& bootstrap, const torch::Tensor& q_next,
                         const torch::Tensor& q_sa, double gamma_n) {
  const torch::Tensor a = reward.contiguous(), b = bootstrap.contiguous(), c = q_next.contiguous(),
                      d = q_sa.contiguous();
  for (const torch::Tensor* t : {&a, &b, &c, &d})
    if (!t->is_cuda() || t->scalar_type() != torch::kFloat32 || t->numel() != a.numel())
      throw std::runtime_error("td_priority: four float32 CUDA tensors of one size");
  torch::Tensor out = torch::empty({a.numel()}, a.options());
  checkActor(rr_td_priority((int)a.numel(), a.data_ptr<float>(), b.data_ptr<float>(), c.data_ptr<float>(),
                            d.data_ptr<float>(), (float)gamma_n, out.data_ptr<float>(), currentStream(a)),
             "td_priority");
  return out;
}

// eta * max + (1 - eta) * mean per sequence (pyrela/r2d2.py:103-120)
torch::Tensor aggregatePriority(const torch::Tensor& priority, const torch::Tensor& seq_len, double eta, int burn_in) {
  if (!priority.is_cuda() || priority.scalar_type() != torch::kFloat32 || priority.dim() != 2 || !seq_len.is_cuda() ||
      seq_len.scalar_type() != torch::kFloat32 || seq_len.numel() != priority.size(0))
    throw std::runtime_error("aggregate_priority: priority float32 [nseq, L] and seq_len float32 [nseq] on CUDA");
  const torch::Tensor p = priority.contiguous(), l = seq_len.contiguous();
  torch::Tensor out = torch::empty({p.size(0)}, p.options());
  checkActor(rr_aggregate_priority((int)p.size(0), (int)p.size(1), p.data_ptr<float>(), l.data_ptr<float>(),
                                   (float)eta, burn_in, out.data_ptr<float>(), currentStream(p)),
             "aggregate_priority");
  return out;
}

}  // namespace

PYBIND11_MODULE(_C, m) {
  m.doc() = "rela_b200: B200-native prioritized replay behind rela's replay API";
  // priority ^ alpha of a shard's slice at its position inside the collective block
  // (ReplayCore::raiseAtPosition, sharded_core.cc); exposed for tests/test_sharded.py
  m.def("raise_at_position", &ReplayCore::raiseAtPosition, py::arg("priority"), py::arg("offset"), py::arg("total"),
        py::arg("alpha"));
  m.def("nstep_return", &nstepReturn, py::arg("reward"), py::arg("terminal"), py::arg("gamma"),
        py::arg("contract_fma") = true);
  m.def("td_priority", &tdPriority, py::arg("reward"), py::arg("bootstrap"), py::arg("q_next"), py::arg("q_sa"),
        py::arg("gamma_n"));
  m.def("aggregate_priority", &aggregatePriority, py::arg("priority"), py::arg("seq_len"), py::arg("eta"),
        py::arg("burn_in"));

  py::class_<FFTransition, std::shared_ptr<FFTransition>>(m, "FFTransition")
      .def(py::init<>())
      .def(py::init([](TensorDict obs, TensorDict action, torch::Tensor reward, torch::Tensor terminal,
                       torch::Tensor bootstrap, TensorDict nextObs) {
             auto t = std::make_shared<FFTransition>();
             t->obs = std::move(obs);
             t->action = std::move(action);
             t->reward = std::move(reward);
             t->terminal = std::move(terminal);
             t->bootstrap = std::move(bootstrap);
             t->nextObs = std::move(nextObs);
             return t;
           }),
           py::arg("obs"), py::arg("action"), py::arg("reward"), py::arg("terminal"), py::arg("bootstrap"),
           py::arg("next_obs"))
      .def_readwrite("obs", &FFTransition::obs)
      .def_readwrite("action", &FFTransition::action)
      .def_readwrite("reward", &FFTransition::reward)
      .def_readwrite("terminal", &FFTransition::terminal)
      .def_readwrite("bootstrap", &FFTransition::bootstrap)
      .def_readwrite("next_obs", &FFTransition::nextObs);

  py::class_<RNNTransition, std::shared_ptr<RNNTransition>>(m, "RNNTransition")
      .def(py::init<>())
      .def(py::init([](TensorDict obs, TensorDict h0, TensorDict action, torch::Tensor reward,
                       torch::Tensor terminal, torch::Tensor bootstrap, torch::Tensor seqLen) {
             auto t = std::make_shared<RNNTransition>();
             t->obs = std::move(obs);
             t->h0 = std::move(h0);
             t->action = std::move(action);
             t->reward = std::move(reward);
             t->terminal = std::move(terminal);
             t->bootstrap = std::move(bootstrap);
             t->seqLen = std::move(seqLen);
             return t;
           }),
           py::arg("obs"), py::arg("h0"), py::arg("action"), py::arg("reward"), py::arg("terminal"),
           py::arg("bootstrap"), py::arg("seq_len"))
      .def_readwrite("obs", &RNNTransition::obs)
      .def_readwrite("h0", &RNNTransition::h0)
      .def_readwrite("action", &RNNTransition::action)
      .def_readwrite("reward", &RNNTransition::reward)
      .def_readwrite("terminal", &RNNTransition::terminal)
      .def_readwrite("bootstrap", &RNNTransition::bootstrap)
      .def_readwrite("seq_len", &RNNTransition::seqLen);

  py::class_<FFPrioritizedReplay, std::shared_ptr<FFPrioritizedReplay>> ff(m, "FFPrioritizedReplay");
  bindCommon<FFPrioritizedReplay>(ff);
  ff.def("add",
         [](FFPrioritizedReplay& r, const FFTransition& batch, const torch::Tensor& priority) {
           r.add(batch, priority);
         },
         py::call_guard<py::gil_scoped_release>());

  // R2D2 sequence assembly with the windows in HBM (R2D2TransitionBuffer,
  // rela/r2d2_actor.h:10-187; not bound in the reference, which uses it from C++)
  py::class_<SequenceAssembler>(m, "SequenceAssembler")
      .def(py::init([](RNNPrioritizedReplay& r, int batchsize, int multiStep, int seqLen, int burnin, float eta) {
             return new SequenceAssembler(r.core(), batchsize, multiStep, seqLen, burnin, eta);
           }),
           py::keep_alive<1, 2>(), py::arg("replay"), py::arg("batchsize"), py::arg("multi_step"),
           py::arg("seq_len"), py::arg("burn_in"), py::arg("eta"))
      .def("push",
           [](SequenceAssembler& a, const FFTransition& t, const torch::Tensor& priority, const TensorDict& hid) {
             Leaves h;
             appendDict(h, "h0", hid);
             a.push(FFTraits<FFTransition>::leaves(t), priority, h);
           })
      .def("can_pop", &SequenceAssembler::canPop)
      .def("pop_into", &SequenceAssembler::popInto);

  py::class_<RNNPrioritizedReplay, std::shared_ptr<RNNPrioritizedReplay>> rnn(m, "RNNPrioritizedReplay");
  bindCommon<RNNPrioritizedReplay>(rnn);
  rnn.def("add",
          [](RNNPrioritizedReplay& r, const std::vector<RNNTransition>& elements, const torch::Tensor& priority) {
            r.add(elements, priority);
          },
          py::call_guard<py::gil_scoped_release>());
}
