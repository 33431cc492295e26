// pybind_module.cc -- Python module rela_b200._C: the replay classes of the
// reference's `rela` module (rela/pybind.cc:19-59) with identical names,
// constructor arguments and methods, served by the B200 replay.
//
//   rela.FFTransition / rela.RNNTransition        pybind.cc:20-35
//   rela.FFPrioritizedReplay(capacity, seed, alpha, beta, prefetch)
//     .size() .num_add() .sample(batchsize, device) .update_priority(priority)
//                                                  pybind.cc:37-59
// Additions (the reference binds no producer side; its actors call add from
// C++): .add(...), .sync(), .handle(), .schema(), and a `device` keyword.
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <torch/csrc/DynamicTypes.h>
#include <torch/extension.h>

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
      .def("size", &Replay::size)
      .def("num_add", &Replay::numAdd)
      .def("sample", &Replay::sample, py::call_guard<py::gil_scoped_release>())
      .def("update_priority", &Replay::updatePriority, py::call_guard<py::gil_scoped_release>())
      .def("sync", [](Replay& r) { r.core().sync(); }, py::call_guard<py::gil_scoped_release>())
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
      .def("device", [](Replay& r) { return r.core().device(); })
      .def("schema", [](Replay& r) { return schemaOf(r); });
}

}  // namespace

PYBIND11_MODULE(_C, m) {
  m.doc() = "rela_b200: B200-native prioritized replay behind rela's replay API";

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
