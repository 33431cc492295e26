// actor_math.cu -- K8 (n-step return + bootstrap) and K9 (TD-error priority,
// eta-mix aggregate) for sm_100a. Element-wise / row-reduction kernels over a
// few thousand floats: launch-latency bound, written for coalesced access.
#include <cfloat>
#include <cmath>

#include "kernels.h"

namespace rr {

// MultiStepTransitionBuffer::popTransition (rela/dqn_actor.h:58-106).
// reward, terminal: [n_step + 1, K] history rows; one thread per env.
__global__ void nstep_kernel(int n_step, int K, float gamma, const float* __restrict__ reward,
                             const uint8_t* __restrict__ terminal, float* __restrict__ out_reward,
                             float* __restrict__ out_bootstrap, uint8_t* __restrict__ out_terminal, int fma) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K) return;
  float bootstrap = 1.0f;
  int last = n_step;
  for (int step = 0; step < n_step; ++step) {  // dqn_actor.h:72-85
    if (terminal[(size_t)step * K + i]) {
      bootstrap = 0.0f;
      last = step;
      break;
    }
  }
  const int initial = (bootstrap != 0.0f) ? n_step - 1 : last;  // dqn_actor.h:91
  float R = 0.0f;
  for (int step = initial; step >= 0; --step) {  // dqn_actor.h:92-96
    const float r = reward[(size_t)step * K + i];
    R = fma ? __fmaf_rn(gamma, R, r) : __fadd_rn(r, __fmul_rn(gamma, R));
  }
  out_reward[i] = R;
  out_bootstrap[i] = bootstrap;
  out_terminal[i] = terminal[i];  // terminalHistory_.front(), dqn_actor.h:66
}

// |reward + bootstrap * gamma^n * q_next - q_sa| (apex.py:44-45,77-78;
// r2d2.py:98-99), each tensor op rounded to float like the eager expression.
__global__ void td_priority_kernel(int K, const float* __restrict__ reward, const float* __restrict__ bootstrap,
                                   const float* __restrict__ q_next, const float* __restrict__ q_sa,
                                   float gamma_n, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K) return;
  const float a = __fmul_rn(bootstrap[i], gamma_n);
  const float b = __fmul_rn(a, q_next[i]);
  const float target = __fadd_rn(reward[i], b);
  out[i] = fabsf(__fsub_rn(target, q_sa[i]));
}

// aggregate_priority (r2d2.py:103-120): one warp per sequence.
__global__ void aggregate_kernel(int nseq, int L, const float* __restrict__ priority,
                                 const float* __restrict__ len, float eta, int burn_in,
                                 float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nseq) return;
  const float n = len[warp];
  float s = 0.0f, mx = -INFINITY;
  for (int j = lane; j < L; j += 32) {
    const float m = ((float)j < n) ? 1.0f : 0.0f;
    const float p = __fmul_rn(priority[(size_t)warp * L + j], m);
    s = __fadd_rn(s, p);
    mx = fmaxf(mx, p);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s = __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (lane == 0) {
    const float mean = __fdiv_rn(s, __fsub_rn(n, (float)burn_in));
    out[warp] = __fadd_rn(__fmul_rn(eta, mx), __fmul_rn(__fsub_rn(1.0f, eta), mean));
  }
}

cudaError_t launch_nstep(int n_step, int K, float gamma, const float* reward, const uint8_t* terminal,
                         float* out_reward, float* out_bootstrap, uint8_t* out_terminal, int fma,
                         cudaStream_t stream) {
  if (K <= 0) return cudaSuccess;
  nstep_kernel<<<(K + 255) / 256, 256, 0, stream>>>(n_step, K, gamma, reward, terminal, out_reward,
                                                    out_bootstrap, out_terminal, fma);
  return cudaGetLastError();
}

cudaError_t launch_td_priority(int K, const float* reward, const float* bootstrap, const float* q_next,
                               const float* q_sa, float gamma_n, float* out, cudaStream_t stream) {
  if (K <= 0) return cudaSuccess;
  td_priority_kernel<<<(K + 255) / 256, 256, 0, stream>>>(K, reward, bootstrap, q_next, q_sa, gamma_n, out);
  return cudaGetLastError();
}

cudaError_t launch_aggregate(int nseq, int L, const float* priority, const float* len, float eta,
                             int burn_in, float* out, cudaStream_t stream) {
  if (nseq <= 0) return cudaSuccess;
  const int warps_per_block = 8;
  aggregate_kernel<<<(nseq + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0, stream>>>(
      nseq, L, priority, len, eta, burn_in, out);
  return cudaGetLastError();
}

}  // namespace rr
