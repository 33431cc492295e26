// row_copy.cu -- K6 gather and K1 insert: the HBM-bound row mover (sm_100a).
//
// Roofline: pure data movement, bounded by HBM bandwidth. Algorithmic bytes
// per transition = 2 * row bytes (one read, one write): 113,234 B for the
// Ape-X element, 6,974,182 B for the R2D2 sequence.
//
// Shape of the work. A row is split into chunks (T per element; the gather
// writes them time-major) and chunks into 16-byte-multiple pieces of a few KB.
// Each CTA is persistent: warp 0's elected thread streams its share of pieces
// through a ring of shared-memory stages with 1-D bulk async copies
// (cp.async.bulk global->shared completing on an mbarrier, then
// shared->global tracked by bulk groups) so that `stages - 1` loads per SM
// are always in flight and no register file or LSU bandwidth is spent on the
// frames. The remaining warps move the narrow / unaligned fields (eps,
// legal_move, action, reward, ...) with plain vector loads and stores,
// grid-strided, so one launch assembles the whole batch.
#include "kernels.h"
#include "ptx.cuh"

namespace rr {

namespace {

__device__ __forceinline__ void copy_unit(uint8_t* dst, const uint8_t* src, uint32_t vec) {
  switch (vec) {
    case 16: stg_stream16(dst, ldg_stream16(src)); break;
    case 8: *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(src); break;
    case 4: *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(src); break;
    case 2: *reinterpret_cast<uint16_t*>(dst) = *reinterpret_cast<const uint16_t*>(src); break;
    default: *dst = *src; break;
  }
}

// Source / destination of (field, row b, chunk t) at byte offset `off` in the chunk.
// The destination never depends on the sampled ids, so it can be computed
// before the kernel's dependencies have resolved.
// Slot of row b: the sampled id (gather) or (start + b) % S (insert); false if
// another shard owns it.
__device__ __forceinline__ bool row_slot(const CopyParams& p, uint32_t b, uint32_t* local) {
  uint32_t slot;
  if (p.ids != nullptr) {
    slot = (uint32_t)__ldg(p.ids + b);
  } else if (p.dst_rows != nullptr) {
    slot = (uint32_t)__ldg(p.dst_rows + b);
  } else {
    slot = (uint32_t)p.insert_start + b;
    if (slot >= (uint32_t)p.slots) slot -= (uint32_t)p.slots;
  }
  *local = slot - (uint32_t)p.own_lo;  // sharded insert (dst_tab): own_lo == 0, i.e. the ring slot itself
  return slot >= (uint32_t)p.own_lo && slot < (uint32_t)p.own_hi;
}

__device__ __forceinline__ uint8_t* row_dst(const CopyParams& p, const CopyField& f, uint32_t b, uint32_t t,
                                            uint32_t off, uint32_t local) {
  if (p.ids != nullptr) {  // gather: batch [t, b, :]
    if (p.dst_tab != nullptr) {  // ... of the learner that owns batch row b
      const uint32_t chunk = b / (uint32_t)p.rows_per_dst;
      const uint32_t bl = b - chunk * (uint32_t)p.rows_per_dst;
      const uint32_t learner = p.dst_of_chunk[chunk];
      uint8_t* base = reinterpret_cast<uint8_t*>(__ldg(p.dst_tab + (size_t)learner * p.tab_cols + f.tab));
      return base + (uint64_t)t * f.dst_chunk_stride + (uint64_t)bl * f.dst_row_stride + off;
    }
    return f.dst + (uint64_t)t * f.dst_chunk_stride + (uint64_t)b * f.dst_row_stride + off;
  }
  if (p.dst_tab != nullptr) {  // insert into a striped replay: the slab of the shard that owns the slot
    const uint32_t owner = local / (uint32_t)p.rows_per_dst;
    uint8_t* base = reinterpret_cast<uint8_t*>(__ldg(p.dst_tab + (size_t)owner * p.tab_cols + f.tab));
    return base + (uint64_t)(local - owner * (uint32_t)p.rows_per_dst) * f.dst_row_stride + (uint64_t)t * f.chunk_bytes + off;
  }
  return f.dst + (uint64_t)local * f.dst_row_stride + (uint64_t)t * f.chunk_bytes + off;  // insert: slab row
}

__device__ __forceinline__ const uint8_t* row_src(const CopyParams& p, const CopyField& f, uint32_t b, uint32_t t,
                                                  uint32_t off, uint32_t local) {
  const uint64_t inner = (uint64_t)t * f.chunk_bytes + off;
  if (p.ids != nullptr) return f.src + (uint64_t)local * f.src_row_stride + inner;
  return (f.row_ptrs != nullptr ? f.row_ptrs[b] : f.src + (uint64_t)b * f.src_row_stride) + inner;
}

struct Item {
  const CopyField* f;
  uint8_t* dst;  // gather only (an insert's destination needs the slot)
  uint32_t b, t, off, bytes;
};

// Everything about an item except its source address (which needs ids[b]).
__device__ __forceinline__ Item decode_item(const CopyParams& p, uint32_t item) {
  int fi = 0;
#pragma unroll 1
  for (int k = 1; k < p.n_bulk; ++k) {
    if (item >= p.bulk[k].begin) fi = k;
  }
  const CopyField& f = p.bulk[fi];
  Item it;
  it.f = &f;
  const uint32_t local = item - f.begin;
  const uint32_t piece = local % f.pieces;
  const uint32_t ct = local / f.pieces;  // row * chunks + chunk
  it.t = ct % f.chunks;
  it.b = ct / f.chunks;
  it.off = piece * f.piece_bytes;
  it.bytes = min(f.piece_bytes, f.chunk_bytes - it.off);
  it.dst = p.ids != nullptr ? row_dst(p, f, it.b, it.t, it.off, 0u) : nullptr;
  return it;
}

__device__ __forceinline__ void simt_rows(const CopyParams& p);

}  // namespace

__global__ void __launch_bounds__(256) row_copy_kernel(const __grid_constant__ CopyParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  // layout: 32 mbarriers | stages * stage_bytes
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
  uint8_t* stage0 = smem_raw + kCopyHeaderBytes;

  const int tid = threadIdx.x;
  unsigned long long* probe = p.probe ? p.probe + (size_t)blockIdx.x * 8 : nullptr;
  if (probe && tid == 0) {
    probe[0] = globaltimer_ns();
    probe[1] = clock64();
  }
  const bool bulk_warp = (p.n_bulk > 0) && (tid < 32);
  if (bulk_warp) {
    // ---- bulk path. Lane l owns stage l and its own mbarrier: it streams the
    // CTA's items l, l + stages, ... through that stage, so up to `stages`
    // loads per CTA are in flight and the per-item address arithmetic is spread
    // over the warp. Everything that does not depend on the sampled ids
    // (barrier init, item decode, destination) runs before pdl_wait(), i.e.
    // overlapped with the tail of the selection kernel.
    const uint32_t NS = (uint32_t)p.stages;
    const bool active = (uint32_t)tid < NS;
    const uint32_t n_items = p.bulk_items;
    const uint32_t first = blockIdx.x;
    const uint32_t step = gridDim.x;
    const uint32_t mine = first < n_items ? (n_items - first + step - 1) / step : 0;
    uint8_t* stage = stage0 + (size_t)tid * p.stage_bytes;
    uint64_t* bar = &full[active ? tid : 0];
    Item it;
    it.bytes = 0;
    if (active) {
      mbar_init(bar, 1);
      mbar_fence_init();
      if ((uint32_t)tid < mine) it = decode_item(p, first + (uint32_t)tid * step);
    }
    // The trigger comes AFTER the wait: a dependent that starts early (the
    // priority update, which does not need this kernel's output) may then rely
    // on everything this kernel waited for -- the selection -- being complete.
    if (blockIdx.x == 0 && tid == 0 && p.ids != nullptr) RR_TL_STAMP_TAG(p.tl_ctl, kTlGatherEnter, p.signal_val);
    pdl_wait();
    pdl_launch_dependents();
    if (probe && tid == 0) probe[2] = clock64();
    if (blockIdx.x == 0 && tid == 0 && p.ids != nullptr) RR_TL_STAMP_TAG(p.tl_ctl, kTlGatherBegin, p.signal_val);
    if (active) {
      const uint64_t evict_first = l2_policy_evict_first();  // a gather reads every source row once
      uint32_t use = 0;
      for (uint32_t j = (uint32_t)tid; j < mine; j += NS) {
        if (j != (uint32_t)tid) it = decode_item(p, first + j * step);
        uint32_t local;
        if (!row_slot(p, it.b, &local)) continue;  // another shard's row
        if (use > 0) bulk_wait_read<0>();  // my previous store has drained the stage
        const uint8_t* src = row_src(p, *it.f, it.b, it.t, it.off, local);
        if (p.ids == nullptr) it.dst = row_dst(p, *it.f, it.b, it.t, it.off, local);
        mbar_expect_tx(bar, it.bytes);
        if (p.ids != nullptr) bulk_g2s_hint(stage, src, it.bytes, bar, evict_first); else bulk_g2s(stage, src, it.bytes, bar);
        if (probe && tid == 0 && use == 0) probe[3] = clock64();
        mbar_wait(bar, use & 1u);
        if (probe && tid == 0 && use == 0) probe[4] = clock64();
        bulk_s2g(it.dst, stage, it.bytes);
        bulk_commit();
        ++use;
      }
      if (probe && tid == 0) probe[5] = clock64();
      // the stores only have to be done READING shared memory before the CTA
      // exits; kernel completion makes them visible. With a completion signal
      // they must be complete before the flag is published.
      if (p.n_signals > 0) bulk_wait_all<0>(); else bulk_wait_read<0>();
      if (blockIdx.x == 0 && tid == 0 && p.ids != nullptr) RR_TL_STAMP_TAG(p.tl_ctl, kTlGatherEnd, p.signal_val);
      if (tid == 0 && p.ids != nullptr && p.tl_ctl != nullptr &&
          *reinterpret_cast<volatile unsigned int*>(&p.tl_ctl->tl_on) != 0u)
        atomicMax(&p.tl_ctl->tl_gmax, globaltimer_ns());
      if (probe && tid == 0) {
        probe[6] = clock64();
        probe[7] = globaltimer_ns();
      }
    }
  } else {
    // ---- narrow fields: grid-strided units over the non-bulk warps
    pdl_wait();
    pdl_launch_dependents();
    simt_rows(p);
  }
  if (p.n_signals == 0) return;
  // ---- completion: the last CTA publishes the epoch (st.release.sys) after
  // every CTA's stores have been fenced at system scope
  __shared__ int sh_last;
  __syncthreads();
  if (tid == 0) {
    __threadfence_system();
    sh_last = atomicAdd(p.done_counter, 1u) == gridDim.x - 1;
    // (the other CTAs' fenced stores happen-before their increments, which this CTA's increment
    // has observed: the release stores below publish them all, no further fence is needed)
    if (sh_last) atomicExch(p.done_counter, 0u);
  }
  __syncthreads();
  if (!sh_last) return;
  if (tid < p.n_signals)
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p.signal[tid]), "l"(p.signal_val) : "memory");
  if (tid == 0 && p.ids != nullptr) RR_TL_STAMP_TAG(p.tl_ctl, kTlGatherSignalled, p.signal_val);
  // a rank that consumes the batch: wait here for the other shards' flags, so that this kernel's
  // completion IS the batch's (the flags are in this GPU's memory; one CTA polls them)
  if (tid < p.wait_n) {
    const unsigned long long t0 = globaltimer_ns();
    // (a replay already in error -- some wait has timed out -- does not wait again: the host raises
    // at its next call, and what is queued until then drains at once)
    const bool failed = *reinterpret_cast<volatile int*>(p.wait_error) != 0;
    for (; !failed;) {
      unsigned long long v;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p.wait_flags + tid) : "memory");
      if (v >= p.signal_val) break;
      if (globaltimer_ns() - t0 > 10ull * 1000 * 1000 * 1000) {
        *p.wait_error = 2;  // RR_ERR_CUDA: a shard never signalled
        if (p.wait_host_error != nullptr) {
          *reinterpret_cast<volatile int*>(p.wait_host_error) = 2;
          __threadfence_system();
        }
        break;
      }
      __nanosleep(100);
    }
  }
  if (p.wait_n > 0) {
    __syncthreads();
    if (tid == 0 && p.ids != nullptr) RR_TL_STAMP_TAG(p.tl_ctl, kTlGatherComplete, p.signal_val);
  }
}

__global__ void wait_signals_kernel(const unsigned long long* flags, int n, unsigned long long val, int* error,
                                    int* host_error, unsigned long long timeout_ns) {
  const int lane = threadIdx.x;
  if (lane >= n) return;
  const unsigned long long t0 = globaltimer_ns();
  const bool failed = *reinterpret_cast<volatile int*>(error) != 0;  // (already in error: no further waits)
  for (; !failed;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + lane) : "memory");
    if (v >= val) break;
    if (globaltimer_ns() - t0 > timeout_ns) {
      *error = 2;  // RR_ERR_CUDA: a shard never signalled
      if (host_error != nullptr) {
        *reinterpret_cast<volatile int*>(host_error) = 2;
        __threadfence_system();
      }
      break;
    }
    __nanosleep(100);
  }
}

// Push `bytes` from `src` to n peer buffers (peer-mapped pointers) and then
// publish `val` at each peer's flag with a system-scope release: the learner's
// new priorities travel to the other shards without a collective.
struct PeerList {
  void* dst[16];
  unsigned long long* flag[16];
};

__global__ void __launch_bounds__(256) peer_put_kernel(const uint8_t* __restrict__ src, uint32_t bytes, PeerList peers,
                                                       int n, unsigned long long val) {
  const int tid = threadIdx.x;
  for (int p = 0; p < n; ++p) {
    uint8_t* dst = static_cast<uint8_t*>(peers.dst[p]);
    if ((bytes & 15u) == 0 && ((uintptr_t)src & 15u) == 0 && ((uintptr_t)dst & 15u) == 0) {
      for (uint32_t i = tid * 16u; i < bytes; i += blockDim.x * 16u)
        *reinterpret_cast<uint4*>(dst + i) = *reinterpret_cast<const uint4*>(src + i);
    } else {
      for (uint32_t i = tid; i < bytes; i += blockDim.x) dst[i] = src[i];
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid < n) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peers.flag[tid]), "l"(val) : "memory");
}

// peer_put + wait in one launch: push this rank's contribution to every rank's inbox (its own
// included), publish `val` at this rank's flag in each of them, then acquire until the flags of
// all n_wait contributors in MY inbox have reached `val`. The all-gather of new priorities /
// block weights between the shards of a striped replay, without a collective library.
__global__ void __launch_bounds__(256) exchange_kernel(const uint8_t* __restrict__ src, uint32_t bytes, PeerList peers,
                                                       int n, unsigned long long val,
                                                       const unsigned long long* wait_flags, int n_wait, int* error,
                                                       int* host_error, unsigned long long timeout_ns) {
  const int tid = threadIdx.x;
  pdl_wait();  // `src` may come from the kernel in front of this one
  pdl_launch_dependents();
  if (bytes > 0) {
    for (int p = 0; p < n; ++p) {
      uint8_t* dst = static_cast<uint8_t*>(peers.dst[p]);
      if ((bytes & 15u) == 0 && ((uintptr_t)src & 15u) == 0 && ((uintptr_t)dst & 15u) == 0) {
        for (uint32_t i = tid * 16u; i < bytes; i += blockDim.x * 16u)
          *reinterpret_cast<uint4*>(dst + i) = *reinterpret_cast<const uint4*>(src + i);
      } else {
        for (uint32_t i = tid; i < bytes; i += blockDim.x) dst[i] = src[i];
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid < n) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peers.flag[tid]), "l"(val) : "memory");
  if (tid < n_wait) {
    const unsigned long long t0 = globaltimer_ns();
    const bool failed = *reinterpret_cast<volatile int*>(error) != 0;  // (already in error: no further waits)
    for (; !failed;) {
      unsigned long long v;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(wait_flags + tid) : "memory");
      if (v >= val) break;
      if (globaltimer_ns() - t0 > timeout_ns) {
        *error = 2;  // RR_ERR_CUDA: a shard never published
        if (host_error != nullptr) {
          *reinterpret_cast<volatile int*>(host_error) = 2;
          __threadfence_system();
        }
        break;
      }
      __nanosleep(100);
    }
  }
}

cudaError_t launch_exchange(const void* src, uint32_t bytes, void* const* dst, void* const* flags, int n,
                            unsigned long long val, const unsigned long long* wait_flags, int n_wait, int* error,
                            int* host_error, unsigned long long timeout_ns, cudaStream_t stream) {
  PeerList pl;
  for (int i = 0; i < n; ++i) {
    pl.dst[i] = dst[i];
    pl.flag[i] = static_cast<unsigned long long*>(flags[i]);
  }
  return launch_pdl(exchange_kernel, dim3(1), dim3(256), 0, stream, static_cast<const uint8_t*>(src), bytes, pl, n, val,
                    wait_flags, n_wait, error, host_error, timeout_ns);
}

cudaError_t launch_peer_put(const void* src, uint32_t bytes, void* const* dst, void* const* flags, int n,
                            unsigned long long val, cudaStream_t stream) {
  PeerList pl;
  for (int i = 0; i < n; ++i) {
    pl.dst[i] = dst[i];
    pl.flag[i] = static_cast<unsigned long long*>(flags[i]);
  }
  peer_put_kernel<<<1, 256, 0, stream>>>(static_cast<const uint8_t*>(src), bytes, pl, n, val);
  return cudaGetLastError();
}

namespace {
__device__ __forceinline__ void simt_rows(const CopyParams& p) {
  const int tid = threadIdx.x;
  const uint32_t lane_base = (p.n_bulk > 0) ? 32u : 0u;
  const uint64_t per_cta = blockDim.x - lane_base;
  uint64_t u = (uint64_t)blockIdx.x * per_cta + (tid - lane_base);
  const uint64_t stride = (uint64_t)gridDim.x * per_cta;
  if (p.norm_n > 0) {
    // importance weights: raw / max (the max's low word; the high word is the selection's epoch)
    const float mx = __uint_as_float((uint32_t)__ldcg(p.norm_max));
    for (uint64_t i = u; i < (uint64_t)p.norm_n; i += stride) p.norm_out[i] = __fdiv_rn(__ldcg(p.norm_raw + i), mx);
  }
  if (p.n_simt == 0) return;
  for (; u < p.simt_units; u += stride) {
    int fi = 0;
#pragma unroll 1
    for (int k = 1; k < p.n_simt; ++k) {
      if (u >= p.simt[k].begin) fi = k;
    }
    const CopyField& f = p.simt[fi];
    const uint64_t local = u - f.begin;
    const uint32_t upc = f.chunk_bytes / f.vec;  // units per chunk
    const uint32_t v = (uint32_t)(local % upc);
    const uint64_t ct = local / upc;
    const uint32_t t = (uint32_t)(ct % f.chunks);
    const uint32_t b = (uint32_t)(ct / f.chunks);
    uint32_t row;
    if (!row_slot(p, b, &row)) continue;
    copy_unit(row_dst(p, f, b, t, v * f.vec, row), row_src(p, f, b, t, v * f.vec, row), f.vec);
  }
}
}  // namespace

cudaError_t launch_wait_signals(const unsigned long long* flags, int n, unsigned long long val, int* error,
                                int* host_error, unsigned long long timeout_ns, cudaStream_t stream) {
  wait_signals_kernel<<<1, 32, 0, stream>>>(flags, n, val, error, host_error, timeout_ns);
  return cudaGetLastError();
}

cudaError_t launch_row_copy(const CopyParams& p, const CopyLaunch& cfg, cudaStream_t stream) {
  // the kernel has a few bytes of static shared memory: opt in to what this launch needs, not to the 227 KB maximum
  cudaError_t e = ensure_dynamic_smem(row_copy_kernel, cfg.smem);
  if (e != cudaSuccess) return e;
  return launch_pdl(row_copy_kernel, dim3(cfg.grid), dim3(cfg.block), cfg.smem, stream, p);
}

}  // namespace rr
