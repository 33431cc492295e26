// gather_convert.cu -- the gather with the learner's first pass fused in
// (SURVEY 8(f) rank 3). The reference's learner turns the sampled u8 frames
// into floats right after sample(): `s.float() / 255` (pyrela/net.py:46,104),
// a full extra pass over the batch (read 1 B + write 4 B per pixel) behind the
// gather (read 1 B + write 1 B). Here the rows of a u8 field go from the slab
// to the batch tensor as float32 (or bfloat16) = byte / 255 (evaluated the way
// torch's CUDA kernel does, see unit() below) in one kernel:
// read 1 B + write 4 B (2 B) per pixel, HBM-bound on the stores.
//
// One warp moves 512 source bytes per step: lane l loads the 4 bytes at
// 4 * (32 u + l) (coalesced 128 B), converts, and stores 4 floats (a coalesced
// 512 B line per instruction) or 4 bf16 (256 B); 4 independent load / store
// pairs per lane are in flight. Row i of the batch is slot ids[i] of the slab;
// a row of T chunks lands time-major ([T, B, chunk]) like in row_copy.cu.
#include <cstdint>

#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/rela_b200.h"
#include "kernels.h"
#include "ptx.cuh"

namespace rr {
namespace {

template <typename Out>
struct Store4;

template <>
struct Store4<float> {
  static __device__ __forceinline__ void put(float* dst, float a, float b, float c, float d) {
    *reinterpret_cast<float4*>(dst) = make_float4(a, b, c, d);
  }
};

template <>
struct Store4<__nv_bfloat16> {
  static __device__ __forceinline__ void put(__nv_bfloat16* dst, float a, float b, float c, float d) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
    uint2 v;
    v.x = *reinterpret_cast<uint32_t*>(&lo);
    v.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(dst) = v;
  }
};

// byte / 255 exactly as the learner's `x.float() / 255` evaluates ON THE GPU:
// torch's CUDA true-division kernel multiplies by the reciprocal of a scalar
// divisor (aten/src/ATen/native/cuda/BinaryDivTrueKernel.cu), so the value is
// float(b) * (1.0f / 255.0f), at most 1 ulp away from the IEEE quotient that
// the CPU path of the same expression returns.
__device__ __forceinline__ float unit(uint32_t b) { return __fmul_rn((float)b, 1.0f / 255.0f); }

template <typename Out>
__device__ __forceinline__ void convert_field(const ConvertField& f, const int32_t* __restrict__ ids, int B,
                                              uint64_t warp_id, uint64_t n_warps, int lane) {
  const uint32_t words_per_chunk = f.chunk_bytes / 4u;
  const uint32_t groups_per_chunk = (words_per_chunk + 127u) / 128u;  // a warp step covers 128 words
  const uint64_t tasks = (uint64_t)B * f.chunks * groups_per_chunk;
  Out* const dst_base = static_cast<Out*>(f.dst);
  for (uint64_t task = warp_id; task < tasks; task += n_warps) {
    const uint32_t g = (uint32_t)(task % groups_per_chunk);
    const uint64_t ct = task / groups_per_chunk;
    const uint32_t t = (uint32_t)(ct % f.chunks);
    const uint32_t b = (uint32_t)(ct / f.chunks);
    const uint32_t row = (uint32_t)ids[b];
    const uint32_t* src =
        reinterpret_cast<const uint32_t*>(f.src + (uint64_t)row * f.row_bytes + (uint64_t)t * f.chunk_bytes);
    Out* dst = dst_base + ((uint64_t)t * B + b) * f.chunk_bytes;
    uint32_t wv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t wi = g * 128u + u * 32u + lane;
      wv[u] = wi < words_per_chunk ? __ldg(src + wi) : 0u;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t wi = g * 128u + u * 32u + lane;
      if (wi < words_per_chunk)
        Store4<Out>::put(dst + (uint64_t)wi * 4u, unit(wv[u] & 0xffu), unit((wv[u] >> 8) & 0xffu),
                         unit((wv[u] >> 16) & 0xffu), unit(wv[u] >> 24));
    }
  }
}

__global__ void __launch_bounds__(256) gather_convert_kernel(const ConvertParams p) {
  pdl_wait();
  pdl_launch_dependents();  // after the wait (the replay's kernels all do: select.cu, `rng_early`)
  const int lane = threadIdx.x & 31;
  const uint64_t warp_id = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const uint64_t n_warps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  for (int fi = 0; fi < p.n; ++fi) {
    const ConvertField& f = p.f[fi];
    if (f.format == RR_OUT_F32_DIV255)
      convert_field<float>(f, p.ids, p.B, warp_id, n_warps, lane);
    else
      convert_field<__nv_bfloat16>(f, p.ids, p.B, warp_id, n_warps, lane);
  }
}

}  // namespace

cudaError_t launch_gather_convert(const ConvertParams& p, int sm_count, cudaStream_t stream) {
  // 8 CTAs of 8 warps per SM: 64 warps x 4 loads of 128 B in flight per SM
  return launch_pdl(gather_convert_kernel, dim3(sm_count * 8), dim3(256), 0, stream, p);
}

}  // namespace rr
