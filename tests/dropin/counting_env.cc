// counting_env.cc -- TEST FIXTURE for the drop-in build (tests/dropin/build_dropin.py).
// A deterministic rela::Env (rela/env.h:13-30) whose observation encodes
// (env id, step) so that a sampled transition can be verified from its bytes:
//   obs.s[c, y, x] = (id * 7 + step * 3 + c) & 255        uint8 [C, H, W]
//   obs.tag        = [id, step]                            int64 [2]
//   reward(step)   = ((id + step) % 5) * 0.25 - 0.5
//   terminal every `period` steps.
// It plays the role atari/atari_env.h plays in the reference and is bound like
// atari/pybind.cc binds AtariEnv. The reference's VectorEnv / BasicThreadLoop /
// DQNActor run on top of it unmodified.
#include <pybind11/pybind11.h>
#include <torch/extension.h>

#include "rela/env.h"

namespace py = pybind11;

class CountingEnv : public rela::Env {
 public:
  CountingEnv(int id, int period, int channels, int height, int width)
      : id_(id), period_(period), c_(channels), h_(height), w_(width), step_(0), terminal_(true) {
  }

  rela::TensorDict reset() override {
    step_ = 0;
    terminal_ = false;
    return obs();
  }

  std::tuple<rela::TensorDict, float, bool> step(const rela::TensorDict& /*action*/) override {
    float r = (float)((id_ + step_) % 5) * 0.25f - 0.5f;
    ++step_;
    terminal_ = (step_ % period_) == 0;
    return std::make_tuple(obs(), r, terminal_);
  }

  bool terminated() const override {
    return terminal_;
  }

 private:
  rela::TensorDict obs() const {
    auto s = torch::empty({c_, h_, w_}, torch::kUInt8);
    uint8_t* p = s.data_ptr<uint8_t>();
    for (int c = 0; c < c_; ++c) {
      const uint8_t v = (uint8_t)((id_ * 7 + step_ * 3 + c) & 255);
      std::fill(p + (size_t)c * h_ * w_, p + (size_t)(c + 1) * h_ * w_, v);
    }
    auto tag = torch::empty({2}, torch::kInt64);
    tag.data_ptr<int64_t>()[0] = id_;
    tag.data_ptr<int64_t>()[1] = step_;
    return {{"s", s}, {"tag", tag}};
  }

  const int id_, period_, c_, h_, w_;
  int step_;
  bool terminal_;
};

PYBIND11_MODULE(counting_env, m) {
  py::module_::import("rela");  // rela.Env must be registered first
  py::class_<CountingEnv, rela::Env, std::shared_ptr<CountingEnv>>(m, "CountingEnv")
      .def(py::init<int, int, int, int, int>());
}
