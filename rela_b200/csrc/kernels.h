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
struct Ctl;
constexpr int kMaxCopyFields = 24;
constexpr int kMaxPeers = 8;       // GPUs of one NVSwitch box

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
  uint32_t tab;              // multi-destination gather: this field's column in CopyParams::dst_tab
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
  // Multi-destination gather (one striped replay, several learner GPUs): batch
  // row b belongs to learner b / rows_per_dst and lands in THAT learner's batch
  // tensor, whose base for field column f is dst_tab[learner * tab_cols + f]
  // (device array of pointers, local or peer-mapped). Null: one destination
  // (CopyField::dst).
  const uint64_t* dst_tab;
  int tab_cols;
  int rows_per_dst;
  // Locality-aware assignment (every shard a learner): chunk c = b / rows_per_dst of the batch goes
  // to learner dst_of_chunk[c]. The batch is sorted in ring order and so are the shards' slot
  // ranges, so most chunks can be given to the shard that stores most of their rows, and those
  // rows never cross NVLink (replay.cu, route_assignment). Without a route the table is the identity.
  uint8_t dst_of_chunk[kMaxPeers];
  // Completion signals (sharded gather): when n_signals > 0, the last CTA to
  // finish its stores publishes `signal_val` at every signal[i] with a
  // system-scope release. The pointers may be peer-mapped (the learners' memory).
  unsigned int* done_counter;
  unsigned long long* signal[kMaxPeers];
  int n_signals;
  unsigned long long signal_val;
  // ... and, on a rank that consumes the batch, then acquires until wait_flags[0..wait_n) (its own
  // memory: one flag per shard) have all reached signal_val: the kernel is complete when the
  // BATCH is, no separate wait kernel (timeout -> *wait_error / *wait_host_error = RR_ERR_CUDA).
  const unsigned long long* wait_flags;
  int wait_n;
  int* wait_error;
  int* wait_host_error;
  // Importance weights (select.cu, SelectArgs::israw): out[i] = raw[i] / max for
  // i < norm_n, by the narrow-field warps (rela/prioritized_replay.h:322).
  const float* norm_raw;
  const unsigned long long* norm_max;
  float* norm_out;
  int norm_n;
  struct Ctl* tl_ctl;  // debug timeline (Ctl::tl), gathers only; may be null
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
                                int* host_error, unsigned long long timeout_ns, cudaStream_t stream);

cudaError_t launch_peer_put(const void* src, uint32_t bytes, void* const* dst, void* const* flags, int n,
                            unsigned long long val, cudaStream_t stream);
// put + wait in one launch (the all-gather of priorities between the shards of a striped replay)
cudaError_t launch_exchange(const void* src, uint32_t bytes, void* const* dst, void* const* flags, int n,
                            unsigned long long val, const unsigned long long* wait_flags, int n_wait, int* error,
                            int* host_error, unsigned long long timeout_ns, cudaStream_t stream);

// ------------------------------------------------------------- selection
// Three-level prefix structure over the S_pad weights, all of it maintained
// incrementally by the kernels that change a weight:
//   leaf   256 consecutive slots          (fp64 sum, smallest non-zero weight)
//   group  1 << group_log2 leaves (64)    (same pair, over its leaves)
// select_kernel scans the group sums in shared memory (at most kMaxGroups of
// them: 1.3 KB at 2^21 transitions), then walks group -> leaf -> slot with one
// warp per target and two dependent L2 reads.
constexpr int kBuildSpan = 2048;   // slots per CTA of build_leaves_kernel; S_pad is a multiple
constexpr int kLeafLog2 = 8;
constexpr int kMaxGroupLog2 = 6;   // 64 leaves = 16,384 slots per group
constexpr int kMaxGroups = 2048;   // group prefixes held in shared memory (24 KB): 33.5 M slots
constexpr int kSelThreads = 256;
constexpr int kSelTargets = 8;     // targets per CTA of select_kernel: one per warp
constexpr int kInlinePrio = 1024;  // largest update whose priorities travel as kernel parameters
constexpr int kUpdStaged = 512;    // largest update every CTA stages in shared memory (update_kernel)
constexpr int kNoAllocMaxBatch = 8192;  // = RR_NOALLOC_MAX_BATCH: elements of the static work area (select.cu)
constexpr int kMaxBatch = 16384;   // update_kernel stages 12 B per sampled element in shared memory
constexpr int kTimelineCap = 1024;

// Device-resident control block: ConcurrentQueue::sum_ and the mt19937 of
// PrioritizedReplay (reference rela/prioritized_replay.h:166,343). The RNG
// state is double-buffered: selection k reads mt[k & 1] and writes the state
// after its B draws to mt[(k + 1) & 1], so no CTA of one selection can see a
// half-written state and nobody has to finish "last".
struct Ctl {
  uint32_t mt[2][624];             // first: 16-byte aligned sources of a bulk copy (2496 B each)
  double sum;
  int mti[2];
  int error;                       // sticky RR_ERR_* raised by a kernel (mirrored to a mapped host word)
  unsigned int done_ctas;          // select_kernel with an eviction: CTAs that finished their targets
  unsigned int pad_;
  // row_copy_kernel with a completion signal: CTAs that finished their stores. One counter per
  // stream a gather can run on (main, the two gather streams): gathers on different streams overlap.
  unsigned int gather_done[4];
  unsigned long long n_exact;      // samples selected by the parallel search
  unsigned long long n_ordered;    // samples that needed the ordered walk
  unsigned long long n_popped;
  unsigned long long probe[16];    // clock64 at the phase boundaries of the last select_kernel (CTA 0)
  // Debug timeline (rr_debug_timeline): while tl_on != 0, CTA 0 of every selection, update and
  // gather kernel stamps (globaltimer_ns << 8 | tag << 4 | code) at its start (after its wait) and end.
  unsigned int tl_on, tl_n;
  unsigned long long tl_umax, tl_gmax;  // latest end of any CTA of the last update / gather kernel
  unsigned long long tl[kTimelineCap];
};

// codes of Ctl::tl entries
enum { kTlSelectBegin = 1, kTlSelectEnd = 2, kTlUpdateBegin = 3, kTlUpdateEnd = 4, kTlGatherBegin = 5, kTlGatherEnd = 6, kTlUpdateAllEnd = 7, kTlGatherAllEnd = 8, kTlUpdateFlagsIn = 9,
       kTlSelectEnter = 10, kTlUpdateEnter = 11, kTlGatherEnter = 12, kTlGatherSignalled = 13, kTlGatherComplete = 14 };

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

// The selection hierarchy over the weights (see "selection" above).
struct Tiles {
  double* leaf;    // [S_pad >> 8] leaf sums
  float* leafmin;  // [S_pad >> 8] smallest non-zero weight per leaf (FLT_MAX: none)
  double* grp;     // [ng padded to 4] group sums
  float* grpmin;   // [ng padded to 4]
  int group_log2;  // leaves per group = 1 << group_log2 (<= kMaxGroupLog2)
  int ng;          // groups = S_pad >> (8 + group_log2) <= kMaxGroups
};

struct SelectArgs {
  Ctl* ctl;
  float* w;             // [S_pad] priority^alpha per slot; dead slots hold 0
  uint8_t* evicted;     // [S]
  Tiles t;
  int S, S_pad;
  int head;             // ring head at selection time
  int size;             // safeSize_: elements the walk may visit
  int size_is;          // size_ re-read before the pop (prioritized_replay.h:312)
  int pop_n;            // elements to evict after selection
  int B;
  int32_t* ids;         // out [B]
  // Importance weights (prioritized_replay.h:320-322) leave this kernel RAW,
  // (size * w / sum) ^ -beta per target, together with their maximum; the
  // division by the maximum is done by the gather that follows (row_copy.cu),
  // so no CTA has to wait for the others. The maximum is kept as
  // (epoch << 32 | float bits) under a 64-bit atomicMax: a later selection's
  // values always win over an earlier one's, so the word never needs a reset.
  float* israw;               // out [B]
  unsigned long long* ismax;  // in/out
  uint32_t epoch;             // > the epoch of every earlier selection that used `ismax`
  int rng_par;                // which Ctl::mt / mti this selection reads
  int rng_early;              // the previous selection is complete: draw before griddepcontrol.wait (select.cu)
  float beta;
  int contract_fma;
  int force_ordered;    // test hook: always take the ordered walk
  int* host_err;        // mapped pinned host word mirroring Ctl::error (device pointer)
  // A batch taken with RR_STREAM_OWN_DEFERRED: this kernel completes only after the gather of
  // THAT batch has published `wait_val` in wait_flags[0..wait_n) -- the dependency rides in the
  // kernel (one polling warp of CTA 0, after its own work) instead of a stream wait in front of
  // it, which would cut the programmatic launch chain update -> select.
  const unsigned long long* wait_flags;
  int wait_n;
  unsigned long long wait_val;
};

cudaError_t launch_build_hierarchy(const float* w, const Tiles& t, int S_pad, cudaStream_t stream);
// sm_count, gather_reserve: see launch_select (select.cu) -- whether the kernel may start early behind
// the kernel in front of it (programmatic dependent launch)
cudaError_t launch_select(const SelectArgs& a, int sm_count, size_t gather_reserve, cudaStream_t stream);

// add commit (K2 + weights write + sum_, prioritized_replay.h:58-73,188)
cudaError_t launch_commit(Ctl* ctl, float* w, const Tiles& tiles, int S, int start, int n, const float* prio,
                          float alpha, int is_weight, cudaStream_t stream);

// The all-gather of new priorities between the shards of a striped replay, fused into the update
// kernel. Every priority travels as ONE 8-byte word (epoch << 32 | float bits): CTA 0 stores this
// rank's n_mine values (the kernel's `prio` argument) at [offset, offset + n_mine) of every rank's
// slot array (dst[i], its own included) with plain 64-bit stores -- no fence, no flag: a word is
// valid when its upper half equals `epoch`. A CTA that needs element j polls word j of THIS rank's
// array (`slots`) until then. One NVLink crossing per value instead of data, fence, flag.
// The host rotates through several arrays (rela_b200/cpp/sharded_core.cc, updateArrays): a
// contributing sender can be at most one exchange ahead of another contributor (it needed that
// rank's values to finish the previous one), the single learner of a striped replay at most
// `prefetch` exchanges ahead of a shard that only receives.
struct ExchangeArgs {
  int on;                        // 0: no exchange, `prio` holds all n priorities
  int n_mine;
  unsigned long long* dst[kMaxPeers];  // every rank's slot array + offset (peer-mapped)
  int n_dst;
  unsigned int epoch;            // low 32 bits of the exchange sequence number (never 0)
  const unsigned long long* slots;     // this rank's slot array: all n priorities
  int* host_err;
};

// updatePriority (K2 + K7, prioritized_replay.h:105-119,235-245)
// early: see update_kernel (select.cu); 0 is always correct
// scratch: 12 * n bytes of device memory, needed when n > kUpdStaged (else ignored)
cudaError_t launch_update(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids, int n,
                          const float* prio, float alpha, int is_weight, int early, const ExchangeArgs* xa,
                          void* scratch, cudaStream_t stream);
// same kernel, always unstaged, its work area in a file-scope __device__ array (n <= kNoAllocMaxBatch):
// allocates nothing (rr_update_priority_noalloc)
cudaError_t launch_update_static(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids,
                                 int n, const float* prio, float alpha, int is_weight, cudaStream_t stream);
// same, n <= kInlinePrio priorities read from HOST memory now and carried in the launch parameters
cudaError_t launch_update_inline(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids,
                                 int n, const float* prio_host, float alpha, int is_weight, int early,
                                 const ExchangeArgs* xa, void* scratch, cudaStream_t stream);

// ------------------------------------------------------------ actor math
cudaError_t launch_nstep(int n_step, int K, float gamma, const float* reward, const uint8_t* terminal,
                         float* out_reward, float* out_bootstrap, uint8_t* out_terminal, int fma,
                         cudaStream_t stream);
cudaError_t launch_td_priority(int K, const float* reward, const float* bootstrap, const float* q_next,
                               const float* q_sa, float gamma_n, float* out, cudaStream_t stream);
cudaError_t launch_aggregate(int nseq, int L, const float* priority, const float* len, float eta,
                             int burn_in, float* out, cudaStream_t stream);

}  // namespace rr
