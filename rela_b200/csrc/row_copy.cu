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
struct RowAddr {
  const uint8_t* src;
  uint8_t* dst;
};

__device__ __forceinline__ RowAddr row_addr(const CopyParams& p, const CopyField& f, const int32_t* s_ids,
                                            uint32_t b, uint32_t t, uint32_t off) {
  RowAddr a;
  const uint64_t inner = (uint64_t)t * f.chunk_bytes + off;
  if (p.ids != nullptr) {
    // gather: slab row ids[b] -> batch [t, b, :]
    a.src = f.src + (uint64_t)(uint32_t)s_ids[b] * f.src_row_stride + inner;
    a.dst = f.dst + (uint64_t)t * f.dst_chunk_stride + (uint64_t)b * f.dst_row_stride + off;
  } else {
    // insert: producer row b -> slab row (start + b) % S
    uint32_t slot = (uint32_t)p.insert_start + b;
    if (slot >= (uint32_t)p.slots) slot -= (uint32_t)p.slots;
    a.src = (f.row_ptrs != nullptr ? f.row_ptrs[b] : f.src + (uint64_t)b * f.src_row_stride) + inner;
    a.dst = f.dst + (uint64_t)slot * f.dst_row_stride + inner;
  }
  return a;
}

struct Item {
  const uint8_t* src;
  uint8_t* dst;
  uint32_t bytes;
};

__device__ __forceinline__ Item decode_item(const CopyParams& p, const int32_t* s_ids, uint32_t item) {
  int fi = 0;
#pragma unroll 1
  for (int k = 1; k < p.n_bulk; ++k) {
    if (item >= p.bulk[k].begin) fi = k;
  }
  const CopyField& f = p.bulk[fi];
  uint32_t local = item - f.begin;
  uint32_t piece = local % f.pieces;
  uint32_t ct = local / f.pieces;  // row * chunks + chunk
  uint32_t t = ct % f.chunks;
  uint32_t b = ct / f.chunks;
  uint32_t off = piece * f.piece_bytes;
  uint32_t bytes = min(f.piece_bytes, f.chunk_bytes - off);
  RowAddr a = row_addr(p, f, s_ids, b, t, off);
  return Item{a.src, a.dst, bytes};
}

}  // namespace

__global__ void __launch_bounds__(256) row_copy_kernel(const __grid_constant__ CopyParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  // layout: 32 mbarriers | ids[rows] (gather) | pad to 128 | stages * stage_bytes
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
  int32_t* s_ids = reinterpret_cast<int32_t*>(smem_raw + kCopyHeaderBytes);
  const uint32_t ids_bytes = p.ids != nullptr ? (uint32_t)p.rows * 4u : 0u;
  uint8_t* stage0 = smem_raw + ((kCopyHeaderBytes + ids_bytes + 127u) & ~127u);

  const int tid = threadIdx.x;
  unsigned long long* probe = p.probe ? p.probe + (size_t)blockIdx.x * 8 : nullptr;
  if (probe && tid == 0) {
    probe[0] = globaltimer_ns();
    probe[1] = clock64();
  }
  if (p.ids != nullptr) {
    for (int i = tid; i < p.rows; i += blockDim.x) s_ids[i] = p.ids[i];
  }
  if (tid < p.stages && p.n_bulk > 0) mbar_init(&full[tid], 1);
  if (tid < 32 && p.n_bulk > 0) mbar_fence_init();
  __syncthreads();

  const bool bulk_warp = (p.n_bulk > 0) && (tid < 32);
  if (bulk_warp) {
    // ---- bulk path. Lane l owns stage l: it streams the CTA's items
    // l, l + stages, ... through its own shared-memory stage, so up to
    // `stages` loads per CTA are in flight and the per-item address
    // arithmetic is spread over the warp instead of serialised on one thread.
    const uint32_t NS = (uint32_t)p.stages;
    if ((uint32_t)tid < NS) {
      const uint32_t n_items = p.bulk_items;
      const uint32_t first = blockIdx.x;
      const uint32_t step = gridDim.x;
      const uint32_t mine = first < n_items ? (n_items - first + step - 1) / step : 0;
      uint8_t* stage = stage0 + (size_t)tid * p.stage_bytes;
      uint64_t* bar = &full[tid];
      uint32_t use = 0;
      if (probe && tid == 0) probe[2] = clock64();
      for (uint32_t j = (uint32_t)tid; j < mine; j += NS, ++use) {
        if (use > 0) bulk_wait_read<0>();  // my previous store has drained the stage
        const Item it = decode_item(p, s_ids, first + j * step);
        mbar_expect_tx(bar, it.bytes);
        bulk_g2s(stage, it.src, it.bytes, bar);
        if (probe && tid == 0 && use == 0) probe[3] = clock64();
        mbar_wait(bar, use & 1u);
        if (probe && tid == 0 && use == 0) probe[4] = clock64();
        bulk_s2g(it.dst, stage, it.bytes);
        bulk_commit();
      }
      if (probe && tid == 0) probe[5] = clock64();
      bulk_wait_all<0>();
      __syncwarp(__activemask());
      if (probe && tid == 0) {
        probe[6] = clock64();
        probe[7] = globaltimer_ns();
      }
    }
    return;
  }

  // ---- narrow fields: grid-strided units over the non-bulk warps
  if (p.n_simt == 0) return;
  const uint32_t lane_base = (p.n_bulk > 0) ? 32u : 0u;
  const uint64_t per_cta = blockDim.x - lane_base;
  uint64_t u = (uint64_t)blockIdx.x * per_cta + (tid - lane_base);
  const uint64_t stride = (uint64_t)gridDim.x * per_cta;
  for (; u < p.simt_units; u += stride) {
    int fi = 0;
#pragma unroll 1
    for (int k = 1; k < p.n_simt; ++k) {
      if (u >= p.simt[k].begin) fi = k;
    }
    const CopyField& f = p.simt[fi];
    uint64_t local = u - f.begin;
    uint32_t upc = f.chunk_bytes / f.vec;  // units per chunk
    uint32_t v = (uint32_t)(local % upc);
    uint64_t ct = local / upc;
    uint32_t t = (uint32_t)(ct % f.chunks);
    uint32_t b = (uint32_t)(ct / f.chunks);
    RowAddr a = row_addr(p, f, s_ids, b, t, v * f.vec);
    copy_unit(a.dst, a.src, f.vec);
  }
}

cudaError_t launch_row_copy(const CopyParams& p, const CopyLaunch& cfg, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(row_copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         227 * 1024);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  row_copy_kernel<<<cfg.grid, cfg.block, cfg.smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace rr
