// kernels.h -- launch interface between the host-side replay (replay.cu) and
// the sm_100a kernels (row_copy.cu, select.cu, actor_math.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace rr {

// ---------------------------------------------------------------- row copy
// One kernel moves rows for both directions:
//   GATHER (K6): src row = slab[ids[b]], dst = batch tensor in makeBatch
//                layout [T, B, chunk] (reference rela/types.cc:8-46,140-182)
//   INSERT (K1): src row = producer batch row b, dst = slab[(start + b) % S]
//                (reference rela/prioritized_replay.h:61-66)
constexpr int kMaxCopyFields = 24;

struct CopyField {
  const uint8_t* src;        // gather: slab base; insert: producer base (or null with row_ptrs)
  uint8_t* dst;              // gather: batch tensor base; insert: slab base
  const uint8_t* const* row_ptrs;  // insert only: one device pointer per row, or null
  uint64_t src_row_stride;   // bytes between consecutive source rows
  uint64_t dst_row_stride;   // gather: chunk_bytes; insert: row_bytes
  uint64_t dst_chunk_stride; // gather: B * chunk_bytes; insert: chunk_bytes
  uint32_t chunk_bytes;
  uint32_t chunks;           // T
  uint32_t pieces;           // bulk path: pieces per chunk
  uint32_t piece_bytes;      // bulk path: bytes per piece (last piece may be shorter)
  uint32_t vec;              // simt path: bytes per unit (1,2,4,8,16)
  uint32_t begin;            // first item (bulk) / unit (simt) of this field in the flat index
};

struct CopyParams {
  CopyField bulk[kMaxCopyFields];  // fields moved by cp.async.bulk through shared memory
  CopyField simt[kMaxCopyFields];  // fields moved by ld/st (narrow or unaligned)
  int n_bulk, n_simt;
  uint32_t bulk_items;  // total bulk items
  uint64_t simt_units;  // total simt units
  const int32_t* ids;   // gather: B sampled slot ids (device); insert: null
  const int32_t* dst_rows;  // insert: explicit destination row per source row (device), or null = ring order
  int rows;             // B (gather) or n (insert)
  int insert_start;     // insert: first slot
  int slots;            // insert: S
  // Capacity sharding: this GPU's slabs hold only slots
  // [own_lo, own_hi) of the ring, slab row = slot - own_lo. Rows whose slot is
  // outside the range are skipped by both directions (another shard moves
  // them). Unsharded: [0, S).
  int own_lo, own_hi;
  int stages;           // smem ring depth
  uint32_t stage_bytes; // smem bytes per stage (>= max piece_bytes)
  unsigned long long* probe;  // debug: 8 timestamps per CTA (tools/gather_probe.cu), else null
  // Completion signal (sharded gather): when `signal` is set, the last CTA to
  // finish its stores publishes `signal_val` there with a system-scope release.
  // `signal` may be a peer-mapped pointer into the learner GPU's memory.
  unsigned int* done_counter;
  unsigned long long* signal;
  unsigned long long signal_val;
};

// shared-memory header of row_copy_kernel: barriers + per-stage piece records
constexpr uint32_t kCopyHeaderBytes = 256;
constexpr int kMaxCopyStages = 32;

struct CopyLaunch {
  int grid;
  int block;
  size_t smem;
};

cudaError_t launch_row_copy(const CopyParams& p, const CopyLaunch& cfg, cudaStream_t stream);

// Learner side of the sharded gather: one warp spins (acquire, system scope)
// until flags[0..n) have all reached `val`; raises *error after timeout_ns.
cudaError_t launch_wait_signals(const unsigned long long* flags, int n, unsigned long long val, int* error,
                                unsigned long long timeout_ns, cudaStream_t stream);

cudaError_t launch_peer_put(const void* src, uint32_t bytes, void* const* dst, void* const* flags, int n,
                            unsigned long long val, cudaStream_t stream);

// ------------------------------------------------------------- selection
// Two-level prefix structure over the S_pad weights: tiles of (1 << tile_log2)
// slots (256 ... 2048, chosen so that there are at most kMaxTiles tiles) whose
// sums are rebuilt per sample by build_hierarchy_kernel, scanned in shared
// memory by every CTA of select_kernel, and searched per target; the slot
// inside the tile is found by one warp reading the tile's weights.
constexpr int kBuildSpan = 2048;   // slots per CTA of build_hierarchy_kernel; S_pad is a multiple
constexpr int kMinTileLog2 = 8;
constexpr int kMaxTileLog2 = 11;
constexpr int kMaxTiles = 16384;   // tile prefixes held in shared memory (128 KB)
constexpr int kSelTargets = 32;    // targets per CTA of select_kernel
constexpr int kInlinePrio = 1024;    // largest update whose priorities travel as kernel parameters
constexpr int kMaxBatch = 16384;   // update_kernel stages 12 B per sampled element in shared memory

// Device-resident control block: ConcurrentQueue::sum_ and the mt19937 of
// PrioritizedReplay (reference rela/prioritized_replay.h:166,343).
struct Ctl {
  uint32_t mt[624];                // first: 16-byte aligned source of a bulk copy
  double sum;
  int mti;
  int error;                       // sticky RR_ERR_* raised by a kernel
  unsigned int done_ctas;          // select_kernel: CTAs that finished their targets
  unsigned int gather_done;        // row_copy_kernel: CTAs that finished their stores (completion signal)
  int sel_fail;                    // select_kernel: a target ran past the buffer
  unsigned long long n_exact;      // samples selected by the parallel search
  unsigned long long n_ordered;    // samples that needed the ordered walk
  unsigned long long n_popped;
  unsigned long long probe[16];    // clock64 at the phase boundaries of the last select_kernel (CTA 0)
};

// ------------------------------------------------- gather with conversion
// A u8 field gathered as float32 / bfloat16 = byte / 255 (gather_convert.cu).
struct ConvertField {
  const uint8_t* src;    // slab
  void* dst;             // batch tensor, [chunks, B, chunk_bytes] elements of the output type
  uint64_t row_bytes;
  uint32_t chunks, chunk_bytes;  // chunk_bytes % 4 == 0
  int format;            // RR_OUT_F32_DIV255 | RR_OUT_BF16_DIV255
};
constexpr int kMaxConvertFields = 4;
struct ConvertParams {
  ConvertField f[kMaxConvertFields];
  int n;
  int B;
  const int32_t* ids;
};
cudaError_t launch_gather_convert(const ConvertParams& p, int sm_count, cudaStream_t stream);

// The selection hierarchy over the weights. A tile is 256 << k slots (k =
// tile_log2 - 8) so that at most kMaxTiles tile sums have to fit the selection
// kernel's shared memory; for k > 0 the tile is made of 1 << k LEAVES of 256
// slots whose sums / minima are kept too (in global memory), so that finding a
// slot and refreshing an entry never touch more than 256 weights.
struct Tiles {
  double* l2;      // [S_pad >> tile_log2] tile sums
  float* l2min;    // [S_pad >> tile_log2] min non-zero weight per tile
  double* leaf;    // [S_pad >> 8] leaf sums; nullptr when tile_log2 == 8 (the tiles are the leaves)
  float* leafmin;  // [S_pad >> 8]
  int tile_log2;
};

struct SelectArgs {
  Ctl* ctl;
  float* w;             // [S_pad] priority^alpha per slot; dead slots hold 0
  uint8_t* evicted;     // [S]
  double* l2;           // [S_pad >> tile_log2] tile sums
  float* l2min;         // [S_pad >> tile_log2] min non-zero weight per tile
  double* leaf;         // see Tiles
  float* leafmin;
  float* wsel;          // [>= B] scratch: weight of each selected slot
  int S, S_pad;
  int tile_log2;
  int head;             // ring head at selection time
  int size;             // safeSize_: elements the walk may visit
  int size_is;          // size_ re-read before the pop (prioritized_replay.h:312)
  int pop_n;            // elements to evict after selection
  int B;
  int32_t* ids;         // out [B]
  float* is_weights;    // out [B] (caller memory)
  float beta;
  int contract_fma;
  int force_ordered;    // test hook: always take the ordered walk
};

cudaError_t launch_build_hierarchy(const SelectArgs& a, cudaStream_t stream);
cudaError_t launch_select(const SelectArgs& a, cudaStream_t stream);

// add commit (K2 + weights write + sum_, prioritized_replay.h:58-73,188)
cudaError_t launch_commit(Ctl* ctl, float* w, const Tiles& tiles, int S, int start, int n, const float* prio,
                          float alpha, int is_weight, cudaStream_t stream);

// updatePriority (K2 + K7, prioritized_replay.h:105-119,235-245)
// wait_late: see update_kernel (select.cu); 0 is always correct
cudaError_t launch_update(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids, int n,
                          const float* prio, float alpha, int is_weight, int wait_late, cudaStream_t stream);
// same, n <= kInlinePrio priorities read from HOST memory now and carried in the launch parameters
cudaError_t launch_update_inline(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids,
                                 int n, const float* prio_host, float alpha, int is_weight, int wait_late,
                                 cudaStream_t stream);

// ------------------------------------------------------------ actor math
cudaError_t launch_nstep(int n_step, int K, float gamma, const float* reward, const uint8_t* terminal,
                         float* out_reward, float* out_bootstrap, uint8_t* out_terminal, int fma,
                         cudaStream_t stream);
cudaError_t launch_td_priority(int K, const float* reward, const float* bootstrap, const float* q_next,
                               const float* q_sa, float gamma_n, float* out, cudaStream_t stream);
cudaError_t launch_aggregate(int nseq, int L, const float* priority, const float* len, float eta,
                             int burn_in, float* out, cudaStream_t stream);

}  // namespace rr
