/* rela_b200.h -- C ABI of the B200-native prioritized replay.
 *
 * This is the drop-in boundary for rela's replay hot path: every entry point
 * below replaces one member of the reference's `rela::PrioritizedReplay<T>`
 * (rela/prioritized_replay.h) as bound in rela/pybind.cc:37-59, or one piece
 * of actor-side arithmetic (rela/dqn_actor.h, pyrela/apex.py, pyrela/r2d2.py).
 * Plain pointers and sizes only; no torch types. The library is
 * rela_b200/librela_b200.so, built from rela_b200/csrc/ with nvcc for
 * sm_100a. There is no CPU fallback: every call that needs the device fails
 * with RR_ERR_CUDA when no B200 is present.
 *
 * Storage model. A replay holds S = (int)(1.25 * capacity) slots
 * (prioritized_replay.h:181). An element is a record of `nfields` byte rows
 * (one per leaf tensor of FFTransition / RNNTransition, types.h:18-73). Each
 * field lives in its own HBM slab `[S, row_bytes]` (structure of arrays).
 * A field with `chunks == T > 1` is stored element-major `[T, chunk]` and is
 * gathered TIME-major into `[T, B, chunk]`, which is what
 * RNNTransition::makeBatch's stack(dim=1) produces (types.cc:140-167); with
 * `chunks == 1` the gather is `[B, row]` = FFTransition::makeBatch's
 * stack(dim=0) (types.cc:8-32).
 *
 * Threading. rr_add* may be called from many host threads at once (the
 * reference's blockAppend protocol, prioritized_replay.h:43-78, including its
 * blocking when all S slots are in flight). rr_sample and rr_update_priority
 * must come from one thread at a time (the reference holds mSampler_).
 * The control plane (selection, priority update, add commit) is issued on one
 * internal stream in the order the calls commit; gathers of batches drawn
 * ahead run on two internal gather streams beside it. rr_sample /
 * rr_update_priority hand-shake with the caller's stream by events and never
 * block the host.
 *
 * Device-side errors (a selection that runs past the last element -- the
 * reference's assert at prioritized_replay.h:297-302 -- or a shard that never
 * signals) are sticky: the kernel mirrors the code into host-visible memory and
 * every later rr_sample* / rr_prefetch* / rr_take_prefetched /
 * rr_update_priority fails with it until rr_sync has reported and cleared it.
 */
#ifndef RELA_B200_H_
#define RELA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RR_OK 0
#define RR_ERR_ARG 1    /* bad argument / schema mismatch                       */
#define RR_ERR_CUDA 2   /* CUDA failure or no device                            */
#define RR_ERR_STATE 3  /* protocol violation, e.g. sample before update        */
#define RR_ERR_EMPTY 4  /* sample from an empty buffer (reference: assert)      */
#define RR_ERR_NOMEM 5  /* slabs do not fit in HBM                              */
#define RR_ERR_OVERRUN 6 /* prefix walk ran past the last element
                            (prioritized_replay.h:297-302)                      */

/* flags for rr_add / rr_add_rows / rr_update_priority */
#define RR_ROWS_DEVICE 0x1u    /* row pointers are device pointers (default: host) */
#define RR_PRIO_DEVICE 0x2u    /* priority pointer is a device pointer             */
#define RR_PRIO_IS_WEIGHT 0x4u /* priority is already priority^alpha: skip the
                                  device pow (bit-parity mode)          */

typedef struct rr_replay rr_replay;

typedef struct rr_field {
  const char* name;   /* informational, e.g. "obs.s"; copied                  */
  uint64_t row_bytes; /* bytes of this field per stored element               */
  uint32_t chunks;    /* T; row = T chunks of row_bytes/T bytes; 1 = batch-major */
  uint32_t reserved;
} rr_field;

typedef struct rr_stats {
  uint64_t kernel_launches;   /* kernels this replay has launched             */
  uint64_t samples;           /* completed rr_sample calls                    */
  uint64_t samples_exact;     /* ... that took the parallel (exact) selection */
  uint64_t samples_ordered;   /* ... that fell back to the ordered walk       */
  uint64_t pops;              /* elements evicted                             */
  uint64_t device_errors;     /* sticky device-side error flags (RR_ERR_*)    */
  uint64_t slab_bytes;        /* HBM held by the field slabs                  */
  uint64_t h2d_bytes;         /* bytes copied host->device by rr_add*         */
} rr_stats;

/* Thread-local text of the last error returned on this thread. */
const char* rr_last_error(void);

/* PrioritizedReplay(capacity, seed, alpha, beta, prefetch)
 * (prioritized_replay.h:176-184, pybind.cc:39-43) on CUDA device `device`. */
int rr_create(rr_replay** out, int capacity, int seed, float alpha, float beta,
              int prefetch, int device);
int rr_destroy(rr_replay* r);

/* Freeze the element layout and allocate the slabs. The reference is
 * schema-agnostic (TensorDict); callers derive the fields from the first
 * batch they add. Must be called once before the first add. */
int rr_set_schema(rr_replay* r, int nfields, const rr_field* fields);

/* add(batch, priority) (prioritized_replay.h:186-200): n elements, field f of
 * element i at field_base[f] + i * row_stride[f] (row_stride == NULL: rows are
 * packed, stride = row_bytes). priority: n floats. Blocks while fewer than n
 * of the S slots are free (prioritized_replay.h:47). Host rows have been
 * copied out of the caller's memory when the call returns. `stream` matters
 * only with RR_ROWS_DEVICE / RR_PRIO_DEVICE: it is the cudaStream_t on which
 * those device buffers were produced; the insert is ordered after it and the
 * stream is made to wait until the replay has read them. */
int rr_add(rr_replay* r, int n, const void* const* field_base,
           const uint64_t* row_stride, const float* priority, unsigned flags,
           void* stream);

/* Same with one pointer per row: row_ptrs[f * n + i]
 * (add(vector<RNNTransition>), r2d2_actor.h:301). */
int rr_add_rows(rr_replay* r, int n, const void* const* row_ptrs,
                const float* priority, unsigned flags, void* stream);

/* sample(batchsize, device) (prioritized_replay.h:202-233, 258-328):
 * stratified proportional selection, eviction down to capacity, importance
 * weights, and the gather into out_fields[f] (DEVICE pointers, layout above)
 * and out_weights[batch] (device). `stream` is the caller's cudaStream_t (NULL
 * = legacy default stream): results are ordered on it when the call returns.
 * With prefetch > 0 the call returns the batch selected `prefetch` calls ago,
 * exactly like the reference's future queue. RR_ERR_STATE if the previous
 * batch's priorities were not updated (prioritized_replay.h:203-206). */
int rr_sample(rr_replay* r, int batch, void* const* out_fields,
              float* out_weights, void* stream);

/* Zero-copy prefetch for a host layer that owns the batch buffers (replay
 * created with prefetch == 0). The reference's sample() returns the oldest
 * future and launches new ones (prioritized_replay.h:215-230); its futures
 * assemble into fresh host memory. rr_prefetch_into is one such future: the
 * batch is selected and gathered NOW -- before the current batch's
 * updatePriority, like the reference -- into the caller's buffers, and queued.
 * rr_take_prefetched makes the oldest queued batch the current one (the one
 * rr_update_priority applies to) and orders its contents on `stream`; no copy.
 * rr_prefetched returns the queue length. */
int rr_prefetch_into(rr_replay* r, int batch, void* const* out_fields,
                     float* out_weights, void* stream);
int rr_take_prefetched(rr_replay* r, void* stream);
int rr_prefetched(rr_replay* r);
/* Two special `stream` values of rr_take_prefetched, for a host layer that
 * knows more than a stream order can say:
 * RR_STREAM_NONE          the batch is not ordered on any stream: the caller
 *                         waits for the event it recorded with
 *                         rr_record_prefetched (below) wherever it consumes
 *                         the batch.
 * RR_STREAM_OWN_DEFERRED  the consumer works on the replay's own stream
 *                         (rr_get_stream) AND the caller's next call is a
 *                         prefetch (rr_prefetch_into / rr_prefetch_routed):
 *                         that selection carries the wait for this batch's
 *                         gather, so work the caller puts on the own stream
 *                         AFTER that prefetch call sees the batch. (A stream
 *                         wait in front of the selection would cut the
 *                         programmatic launch chain update -> select; this
 *                         keeps it. Any other call in between falls back to
 *                         the plain stream wait.) */
#define RR_STREAM_NONE ((void*)(intptr_t)-1)
#define RR_STREAM_OWN_DEFERRED ((void*)(intptr_t)-2)
/* Record `event` (a cudaEvent_t of the caller, created on the replay's device)
 * behind the gather of the batch most recently queued by rr_prefetch_into: a
 * host layer that issues from one thread and hands batches out on another
 * makes the consumer's stream wait for this event instead of going through
 * rr_take_prefetched's hand-shake (which it then calls with the replay's own
 * stream, rr_get_stream). */
int rr_record_prefetched(rr_replay* r, void* event);

/* rr_sample in two steps, so that a host layer can allocate the batch tensors
 * while the selection kernel already runs: rr_sample_select issues the
 * selection, eviction and raw importance weights (sample_ lines 258-321);
 * rr_sample_gather issues makeBatch (line 326) into out_fields, normalises the
 * importance weights into out_weights (line 322) in the same kernel and orders
 * the results on `stream`. rr_sample == select + gather. */
int rr_sample_select(rr_replay* r, int batch, float* out_weights, void* stream);
int rr_sample_gather(rr_replay* r, void* const* out_fields, void* stream);

/* updatePriority(priority) (prioritized_replay.h:235-245, 105-119): n must be
 * the size of the batch last returned by rr_sample. Evicted ids are skipped. */
int rr_update_priority(rr_replay* r, int n, const float* priority,
                       unsigned flags, void* stream);

/* Staging primitive for device-side sequence assembly (R2D2TransitionBuffer,
 * rela/r2d2_actor.h:29-87): row i of every packed source
 * field (src_base[f] + i * row_bytes[f], device) is copied to row dst_rows[i]
 * of dst_base[f] (device), with the replay's row-copy kernel on `stream`.
 * dst_rows is a device array of n int32. */
int rr_scatter_rows(rr_replay* r, int n, int nfields, const uint64_t* row_bytes, const void* const* src_base,
                    void* const* dst_base, const int32_t* dst_rows, void* stream);

/* ---------------------------------------------------------------------------
 * One replay striped over the GPUs of one NVSwitch box (no reference
 * counterpart). One process per GPU, SPMD: every shard
 * keeps the FULL control state (weights, evicted flags, prefix hierarchy,
 * sum_, RNG: ~4.1 B per slot) and applies every add / sample / update in the
 * same order, so all shards select the same ids -- bit-identical to the
 * single-buffer reference -- with no collective in the sampling path. Only the
 * payload slabs are partitioned: shard `rank` stores slots
 * [rank * ceil(S / world), ...). Call rr_set_shard before rr_set_schema. The
 * host side of the protocol is rela_b200/cpp/sharded_core.cc. */
int rr_set_shard(rr_replay* r, int rank, int world);
int rr_shard_range(const rr_replay* r, int* lo, int* hi);
/* Device base pointers of the field slabs (for rr_ipc_export): out[nfields]. */
int rr_slab_ptrs(rr_replay* r, void** out, int cap);

/* Where the rows of a batch drawn from a striped replay go. The batch of
 * `batch_total` rows is split evenly over n_learners learner GPUs: rows
 * [l * B, (l + 1) * B), B = batch_total / n_learners, belong to learner l and
 * are stored by whichever shard owns their slot straight into THAT learner's
 * batch tensors (bulk stores over NVLink). n_learners == 1: one learner gets
 * the whole batch; n_learners == world: every GPU draws its share of one
 * stratified batch of world * B (a data-parallel learner; equivalent to ONE
 * reference replay sampling world * B and splitting the batch). */
typedef struct rr_route {
  int n_learners;
  const void* dst_tab;       /* DEVICE array uint64[n_learners * nfields]: base of field f of
                                learner l's batch tensors (local or peer-mapped pointers)      */
  void* const* signal_ptrs;  /* host array [n_signals]: this shard's 8-byte completion flag in
                                each rank's memory (peer-mapped); published (st.release.sys)
                                with the selection epoch after this shard's stores           */
  int n_signals;
  const void* wait_flags;    /* device array of wait_n flags in THIS rank's memory, one per
                                shard, or NULL if this rank does not consume the batch        */
  int wait_n;
  int my_learner;            /* this rank's index among the learners, -1 if none             */
  float* out_weights;        /* this learner's B importance weights (device), or NULL        */
  int on_demand;             /* != 0: the batch is taken right away (no prefetch): its gather
                                stays on the internal stream behind its selection             */
  int* my_chunk;             /* host, out; NULL: chunk c = rows [c * B, (c + 1) * B) goes to
                                learner c. Non-NULL (every shard a learner, n_learners == world,
                                the same choice on every rank): the library assigns the chunks
                                BY LOCALITY -- chunk c goes to learner learner_of_chunk[c] of
                                rr_route_assignment(head, size, S, world) for the ring range the
                                selection draws from -- and stores here which chunk THIS learner
                                receives (-1 without one). The batch is sorted in ring order and
                                so are the shards' slot ranges: most rows of a chunk are stored
                                on the GPU that receives it and never cross NVLink. The
                                learners' batches are still the N chunks of ONE stratified
                                batch of batch_total; only who gets which chunk changes.       */
} rr_route;

/* The assignment rr_prefetch_routed chooses (pure host arithmetic, no device needed): `head` and
 * `size` describe the live ring range after the selection's eviction, `slots` = S, shard r stores
 * slots [r * ceil(S / world), ...); learner_of_chunk[world] receives the bijection chunk ->
 * learner under which most slots of the range belong to the shard that receives their chunk.
 * Every shard computes it from replicated values, so all shards agree without communicating.
 * No counterpart in the reference (one host buffer). */
int rr_route_assignment(int head, int size, int slots, int world, int* learner_of_chunk);

/* rr_prefetch_into for a shard of a striped replay: select batch_total ids
 * (identically on every shard), gather the rows this shard owns to their
 * learners, queue the batch. rr_take_prefetched then makes it current and, on a
 * learner, orders the batch on `stream`: a learner's gather kernel ends by
 * acquiring the wait_n flags itself, so the batch is complete when that kernel
 * is. Every rank passes wait_flags (rr_add_sharded_rows waits on them before it
 * reuses slots a queued gather may still read). */
int rr_prefetch_routed(rr_replay* r, int batch_total, const rr_route* route, void* stream);

/* rr_update_priority without any allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). Same semantics and results as
 * rr_update_priority with RR_PRIO_DEVICE (bit-identical stored weights, sum_ and hierarchy), but
 * the update kernel's work area (12 bytes per element) is a file-scope __device__ array of
 * RR_NOALLOC_MAX_BATCH elements (96 KB, held by every GPU that loads the library until exit)
 * instead of device memory the replay allocates on demand, and nothing else is allocated either:
 * no cudaMalloc / cudaFree / cudaHostAlloc, no event or stream creation.
 *   - `priority` must be device memory and `flags` must contain RR_PRIO_DEVICE (a host array would
 *     need a staging buffer); RR_PRIO_IS_WEIGHT as in rr_update_priority.
 *   - n > RR_NOALLOC_MAX_BATCH is rejected with RR_ERR_ARG (the batch stays outstanding: call
 *     rr_update_priority instead); a sharded replay is rejected with RR_ERR_STATE.
 *   - Every call on one device shares the array, and it keeps the last call's contents (the kernel
 *     writes every word it uses before reading it, so nothing is cleared). Two such updates that
 *     OVERLAP ON THE GPU -- different replays on different streams -- race; calls on one replay are
 *     serialised on its internal stream and are safe. */
#define RR_NOALLOC_MAX_BATCH 8192
int rr_update_priority_noalloc(rr_replay* r, int n, const float* priority, unsigned flags, void* stream);

/* update_priority on a striped replay: every rank applies the priorities of ALL learners, which
 * the update kernel itself all-gathers through peer memory. This rank contributes n_mine values
 * (priority_mine: host memory, or device memory with RR_PRIO_DEVICE) for elements [offset, offset +
 * n_mine) of the batch of n_total. A priority travels as one 8-byte word (epoch << 32 | float
 * bits), valid when its upper half equals `epoch`: slot_ptrs[i] = rank i's array of n_total such
 * words (peer-mapped, n_dst = world, this rank included), my_slots = this rank's own array.
 * `epoch`: the same non-zero value on every rank, different from the value of the previous
 * exchange through the same arrays. The caller rotates through several arrays per rank, enough
 * that no rank can still be polling the array a faster rank writes next: two when every rank
 * contributes (a contributor needed every other contributor's values to finish its previous
 * update), prefetch + 2 when some ranks only receive (the one learner of a striped replay is held
 * back by the gathers only and can be `prefetch` updates ahead of a shard; sharded_core.cc).
 * A rank that is no learner passes n_mine == 0 and only receives. All ranks pass the same `flags`. */
int rr_update_priority_sharded(rr_replay* r, int n_total, const float* priority_mine, int n_mine, int offset,
                               unsigned flags, void* const* slot_ptrs, int n_dst, uint32_t epoch,
                               const void* my_slots, void* stream);

/* Collective add into a striped replay. Step 1 reserves the block of n_total
 * slots (identically on every shard) and moves THIS rank's rows -- rows
 * [row0, row0 + n_local) of the block, device memory, field_base[f] packed --
 * straight into the slabs of the shards that own their slots (slab_tab: DEVICE
 * array uint64[world * nfields] of peer-mapped slab bases, rank-major).
 * RR_ERR_STATE when the ring is full (the reference blocks there; an SPMD
 * caller must sample first). Step 2, after the ranks have exchanged their
 * weights (rr_exchange, ordered after step 1 so that a shard that sees the
 * weights also sees the rows): weights_all = priority^alpha (or raw priorities
 * without RR_PRIO_IS_WEIGHT) of the whole block, device memory. */
int rr_add_sharded_rows(rr_replay* r, int n_total, int row0, int n_local, const void* const* field_base,
                        const void* slab_tab, void* stream);
int rr_add_sharded_commit(rr_replay* r, const float* weights_all, unsigned flags);

/* All-gather of small messages between shards through peer memory, one launch
 * on the replay's internal stream (ordered after `stream`): copy `bytes` from
 * `src` (device) into n peer-mapped buffers, publish `epoch` at the n flags
 * (st.release.sys), then acquire until wait_flags[0..n_wait) (this rank's
 * memory) have all reached `epoch` (10 s timeout -> sticky device error). A
 * rank with nothing to contribute passes bytes == 0, n == 0 and only waits. */
int rr_exchange(rr_replay* r, const void* src, uint64_t bytes, void* const* dst_ptrs, void* const* flag_ptrs,
                int n, uint64_t epoch, const void* wait_flags, int n_wait, void* stream);

/* Lower-level one-sided primitives (used by tests): rr_peer_put copies `bytes`
 * from `src` (device) into n peer-mapped buffers and then publishes `epoch` at
 * each peer's 8-byte flag, on `stream`. rr_wait_flags enqueues a one-warp
 * kernel on `stream` that acquires until flags[0..n) >= epoch (10 s timeout ->
 * *device_error = RR_ERR_CUDA; device_error is a DEVICE pointer). */
int rr_peer_put(int device, const void* src, uint64_t bytes, void* const* dst_ptrs, void* const* flag_ptrs, int n,
                uint64_t epoch, void* stream);
int rr_wait_flags(int device, const void* flags, int n, uint64_t epoch, int* device_error, void* stream);

/* Peer-memory plumbing: cudaMalloc'ed buffers exported as 64-byte CUDA IPC
 * handles and opened by the other shards' processes with peer access. */
int rr_device_alloc(int device, uint64_t bytes, void** out_ptr); /* zero-filled */
int rr_device_free(int device, void* ptr);
int rr_ipc_export(int device, const void* ptr, void* handle64_out);
int rr_ipc_open(int device, const void* handle64, void** out_ptr);
int rr_ipc_close(int device, void* ptr);

/* Rendezvous of the shard processes of one box, used at set-up time only (IPC
 * handle exchange): a POSIX shared-memory segment named after `name`, created
 * by rank 0. Up to 16 ranks, up to 64 KB per rank and all-gather. */
typedef struct rr_group rr_group;
int rr_group_create(rr_group** out, const char* name, int rank, int world);
int rr_group_allgather(rr_group* g, const void* mine, uint64_t bytes, void* all_out);
int rr_group_barrier(rr_group* g);
int rr_group_rank(const rr_group* g);
int rr_group_world(const rr_group* g);
int rr_group_destroy(rr_group* g);

int rr_size(const rr_replay* r);    /* safeSize_ (prioritized_replay.h:247) */
int rr_num_add(const rr_replay* r); /* numAdd_   (prioritized_replay.h:251) */

/* Wait for all device work issued so far; returns the sticky device error
 * (RR_OK if none) and clears it. */
int rr_sync(rr_replay* r);
int rr_get_stats(rr_replay* r, rr_stats* out);

/* Make `stream` (caller's cudaStream_t) wait for everything issued so far on
 * the replay's internal stream. Does not block the host. */
int rr_stream_fence(rr_replay* r, void* stream);
/* The replay's internal cudaStream_t. A caller that does its own device work
 * on this stream and passes it as `stream` gets plain stream order and none of
 * the event hand-shakes: consecutive sample / update kernels then chain by
 * programmatic dependent launch. */
int rr_get_stream(rr_replay* r, void** out_stream);
/* The reverse of rr_stream_fence: the internal stream waits for everything
 * queued so far on `stream` (inputs produced there for a call that is then
 * made in own-stream mode). Does not block the host. */
int rr_stream_wait(rr_replay* r, void* stream);

/* Gather-kernel timing for the roofline. While profiling is enabled every
 * rr_sample also keeps a copy of its sampled ids (ring of 16 sets).
 * rr_gather_bench then re-launches the gather kernel for those recorded id sets,
 * back to back on the internal stream, `reps` times over all sets, rotating
 * over n_out_sets output batches (out_fields[o * nfields + f], device), and
 * times the whole region with one CUDA-event pair: ms_per_launch is the
 * kernel's average duration without per-launch event overhead. streams == 1:
 * back to back on the internal stream; streams == 2: alternating over the two
 * gather streams, the way the gathers of batches drawn ahead are issued. */
int rr_set_profiling(rr_replay* r, int enabled, int max_batch);
int rr_gather_bench(rr_replay* r, int n_out_sets, void* const* out_fields,
                    int reps, int streams, double* ms_per_launch,
                    uint64_t* launches, uint64_t* bytes_per_launch);

/* The timed sample + update_priority loop, run inside the library (like the
 * reference arm's loop, oracle/ref_driver.cc: no per-call foreign-function
 * overhead): `steps` cycles after `warmup`, device-resident priorities
 * (prio_sets[k % n_prio_sets], `batch` floats each, flags as for
 * rr_update_priority), results into out set k % n_out_sets (out_fields[o *
 * nfields + f], out_weights[o]), everything issued with the internal stream as
 * the caller's stream. depth == 0: rr_sample + rr_update_priority, every batch
 * drawn after the previous update. depth > 0: rr_take_prefetched +
 * rr_update_priority + rr_prefetch_into with `depth` batches drawn ahead (the
 * reference's prefetch). Timed with one CUDA-event pair on the internal stream
 * after joining the gather streams. */
int rr_cycle_bench(rr_replay* r, int batch, int depth, int steps, int warmup, int n_out_sets,
                   void* const* out_fields, float* const* out_weights, int n_prio_sets,
                   const float* const* prio_sets, unsigned prio_flags, double* ms_per_step, uint64_t* launches);

/* Learner-side fusion at the batch hand-off (not in the reference; SURVEY 8(f)
 * rank 3). The reference's learner converts the sampled u8 frames right after
 * sample() with `s.float() / 255` (pyrela/net.py:46,104): a second full pass
 * over the batch. With a format set, rr_sample / rr_sample_gather write field
 * `field` as float32 (or bfloat16) = byte / 255 instead of raw bytes --
 * bit-identical to what the torch expression returns on a CUDA tensor, i.e.
 * float(byte) * (1.0f / 255.0f) (torch's CUDA kernel multiplies by the
 * reciprocal of a scalar divisor; the CPU kernel's IEEE quotient can differ
 * by 1 ulp), then rounded to bf16 -- and out_fields[field] must hold
 * batch * row_bytes ELEMENTS of that type.
 * The field's chunk size must be a multiple of 4 bytes. Not available with
 * prefetch > 0 or on a sharded replay. */
#define RR_OUT_RAW 0
#define RR_OUT_F32_DIV255 1
#define RR_OUT_BF16_DIV255 2
int rr_set_output_format(rr_replay* r, int field, int format);

/* Checkpoint / restore (not in the reference). rr_save
 * writes the ring counters, the mt19937 state, sum_, the weights, the eviction
 * flags and the live rows of this replay's slot range to `path` (blocking;
 * device -> pinned bounce buffer -> <path>.tmp, fsync, rename). No add may be
 * in flight and no sampled batch may await its priorities; batches drawn
 * AHEAD (prefetch) are saved with their ids, weights and gathered rows.
 * rr_load restores into a freshly created replay with the same capacity,
 * alpha, beta, shard and schema; sampling then continues bit-identically to
 * the run that saved (same ids, same bytes, same weights), prefetched batches
 * included: they return to the queue in library-owned buffers. A host layer
 * that owns its batch buffers (rr_prefetch_into) claims entry k with
 * rr_restore_prefetched (copied into its buffers; rr_prefetched_batch(k) is the
 * entry's batch size). A sharded replay saves / loads one file per rank and
 * must have no batch queued. */
int rr_save(rr_replay* r, const char* path);
int rr_load(rr_replay* r, const char* path);
int rr_restore_prefetched(rr_replay* r, int k, void* const* out_fields, float* out_weights, void* stream);
int rr_prefetched_batch(rr_replay* r, int k);

/* Test/debug introspection (synchronises). ids: the physical slot ids of the
 * batch last returned by rr_sample (the reference's sampledIds_). */
int rr_last_sample_ids(rr_replay* r, int n, int32_t* out_ids_host);
int rr_debug_state(rr_replay* r, int* head, int* tail, int* size,
                   int* safe_size, double* sum);
int rr_debug_weights(rr_replay* r, float* out_w_host, uint8_t* out_evicted_host);
/* The selection kernel's per-tile weight sums and smallest non-zero weights,
 * as maintained incrementally by add / update / eviction. rebuild != 0 first
 * recomputes every entry from the weights (tests compare the two bit for bit).
 * level 0: the groups the selection kernel scans (16,384 slots); level 1:
 * their 256-slot leaves.
 * Copies min(cap, tiles) entries; returns the number of tiles (< 0: error). */
int rr_debug_tiles(rr_replay* r, int rebuild, int level, double* out_sums_host,
                   float* out_mins_host, int cap, int* out_tile_slots);
/* Kernel timeline: returns (and clears) up to `cap` stamps recorded since the last call, oldest
 * first, then switches recording on / off. A stamp is (globaltimer ns << 8 | tag << 4 | code; tag = low 4 bits of the batch's epoch), written by
 * CTA 0 of each kernel after its dependency wait and at its end; codes: 1 / 2 selection begin /
 * end, 3 / 4 update, 5 / 6 gather. Returns the number of stamps (< 0: error). */
int rr_debug_timeline(rr_replay* r, int enable, uint64_t* out_host, int cap);
/* clock64 stamps at the phase boundaries of the last selection kernel (16 x u64). */
int rr_debug_select_probe(rr_replay* r, uint64_t* out16);

/* ---------------------------------------------------------------------------
 * Actor-side transition assembly (device pointers, async on `stream`).
 *
 * rr_nstep_return: MultiStepTransitionBuffer::popTransition
 * (dqn_actor.h:58-106). reward, terminal: [n_step + 1, K] history rows
 * (row 0 = oldest). Outputs [K]: discounted return, bootstrap flag, and
 * terminal (= history row 0). contract_fma != 0 evaluates r + gamma * R as one
 * fused multiply-add (what -O3 -march=native compiles dqn_actor.h:95 to).
 *
 * rr_td_priority: |reward + bootstrap * gamma_n * q_next - q_sa|
 * (apex.py:44-45,77-78; r2d2.py:98-99). gamma_n = gamma ** n_step.
 *
 * rr_aggregate_priority: eta * max + (1 - eta) * mean over each sequence
 * (r2d2.py:103-120). priority [nseq, seq_len], len [nseq] (float, includes
 * burn-in as in r2d2_actor.h:107). */
int rr_nstep_return(int n_step, int K, float gamma, const float* reward,
                    const uint8_t* terminal, float* out_reward,
                    float* out_bootstrap, uint8_t* out_terminal,
                    int contract_fma, void* stream);
int rr_td_priority(int K, const float* reward, const float* bootstrap,
                   const float* q_next, const float* q_sa, float gamma_n,
                   float* out_priority, void* stream);
int rr_aggregate_priority(int nseq, int seq_len, const float* priority,
                          const float* len, float eta, int burn_in,
                          float* out_priority, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RELA_B200_H_ */
