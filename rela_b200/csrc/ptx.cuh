// ptx.cuh -- thin inline-PTX wrappers used by the row-copy kernels (sm_100a).
// mbarrier + 1-D bulk async copies (the TMA engine's non-tensor path,
// SASS: UBLKCP / SYNCS), 128-bit streaming loads and stores.
#pragma once
#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

namespace rr {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// global -> shared bulk copy, completion signalled on `bar` (complete_tx).
// src, dst 16-byte aligned; bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// Same with an L2 cache policy (createpolicy): the gather's source rows are read once, so they
// are marked evict-first and do not push the control state out of L2.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}

__device__ __forceinline__ void bulk_g2s_hint(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar,
                                              uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// shared -> global bulk copy, tracked by the thread's bulk async-group.
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}

__device__ __forceinline__ void bulk_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// Wait until at most N of this thread's bulk groups still READ their source.
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// 128-bit streaming global load / store (no L1 allocation).
__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ void stg_stream16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// Programmatic dependent launch (PDL): let the next kernel in the stream start
// its prologue now, and block until every prerequisite grid has completed and
// its writes are visible. Both are no-ops for kernels launched without the
// programmatic-stream-serialization attribute.
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Debug timeline stamp (Ctl::tl in kernels.h): one thread per kernel event.
// stamp = globaltimer ns << 8 | tag << 4 | code; tag: low 4 bits of the batch's epoch (which batch)
#define RR_TL_STAMP_TAG(ctl, code, tag)                                                              \
  do {                                                                                               \
    if ((ctl) != nullptr && *reinterpret_cast<volatile unsigned int*>(&(ctl)->tl_on) != 0u) {       \
      const unsigned int i_ = atomicAdd(&(ctl)->tl_n, 1u) % (unsigned int)kTimelineCap;              \
      (ctl)->tl[i_] = (globaltimer_ns() << 8) | (((unsigned long long)(tag) & 15ull) << 4) |         \
                      (unsigned long long)(code);                                                    \
    }                                                                                                \
  } while (0)
#define RR_TL_STAMP(ctl, code) RR_TL_STAMP_TAG(ctl, code, 0)

// Host: opt a kernel in to `bytes` of dynamic shared memory. The attribute is
// per device (context), and several host threads may launch at once, so the
// high-water mark is kept per device in atomics.
template <typename Kernel>
inline cudaError_t ensure_dynamic_smem(Kernel kernel, size_t bytes) {
  static std::atomic<size_t> configured[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) dev = 63;
  // No "fits the default 48 KB" shortcut: the default bound on the DYNAMIC part is
  // 48 KB minus the kernel's static shared memory, so the first launch always opts in.
  if (bytes <= configured[dev].load(std::memory_order_acquire)) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return e;
  size_t cur = configured[dev].load();
  while (cur < bytes && !configured[dev].compare_exchange_weak(cur, bytes)) {
  }
  return cudaSuccess;
}

// Host: launch with the PDL attribute (RR_PDL=0 disables it); `pdl` = false: a plain stream-ordered
// launch of the same kernel (its griddepcontrol.wait then returns at once).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_if(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t stream, Args&&... args) {
  static const bool allowed = [] {
    const char* v = getenv("RR_PDL");
    return !(v && v[0] == '0');
  }();
  const bool enabled = allowed && pdl;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = enabled ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  return launch_pdl_if(true, kernel, grid, block, smem, stream, static_cast<Args&&>(args)...);
}

}  // namespace rr
