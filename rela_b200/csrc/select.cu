// select.cu -- K2 (p^alpha), K3 (prefix hierarchy), K4 (stratified search),
// K5 (importance weights), K7 (priority scatter-update), eviction and the add
// commit, for sm_100a. All of it is latency-bound integer / fp64 work on a few
// hundred targets; none of it is shaped into a GEMM.
//
// Bit-exactness argument (DESIGN.md "Selection"). The reference walks the ring
// from head_ accumulating `double accSum += (float) w` and takes, per target,
// the first index with accSum > 0 && accSum >= target
// (rela/prioritized_replay.h:283-307). If every partial sum is exactly
// representable in double, that walk returns the same indices as ANY
// association of the same additions. A sufficient, cheaply checkable condition:
// all weights are multiples of u = ulp(smallest non-zero weight) and the total
// is < 2^52 * u. Then every value this file computes (leaf sums, group sums,
// prefixes, differences of prefixes) is a multiple of u below 2^53 * u, hence
// exact in fp64, and the three-level search below is bit-identical to the
// sequential walk. When the condition fails (weights spanning > 29 binades
// relative to the total) the kernel takes the ordered walk instead, which is
// the reference's loop verbatim, and counts it in Ctl::n_ordered.
//
// Dead slots hold weight 0 (zeroed on eviction; the reference never reads a
// popped slot's weight again: update() skips evicted ids and append overwrites
// the slot), so physical prefixes need no live-range mask.
#include <algorithm>
#include <cfloat>
#include <cstring>
#include <cmath>
#include <mutex>

#include "../../include/rela_b200.h"
#include "kernels.h"
#include "ptx.cuh"

namespace rr {

namespace {

__device__ __forceinline__ float pow_alpha(float p, float alpha) {
  // K2: priority ^ alpha. alpha == 1 is the identity (ATen copies in that
  // case). Otherwise evaluate in fp64 and round once: correctly rounded
  // except within ~1e-8 relative of a float rounding boundary.
  if (alpha == 1.0f) return p;
  return (float)pow((double)p, (double)alpha);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// inclusive warp scan
__device__ __forceinline__ double warp_scan(double v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double y = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += y;
  }
  return v;
}

// ------------------------------------------------------------------ K3 build
// Sum (fp64) and smallest non-zero magnitude of 256 consecutive slots held as
// 8 values per lane. Every producer of a leaf entry -- the full build and the
// incremental paths of commit / update / eviction -- goes through this one
// association, so that a maintained entry is bit-identical to a rebuilt one.
__device__ __forceinline__ void reduce256(const float4 a, const float4 b, double& s, float& mn) {
  s = (((double)a.x + (double)a.y) + ((double)a.z + (double)a.w)) +
      (((double)b.x + (double)b.y) + ((double)b.z + (double)b.w));
  mn = FLT_MAX;
  const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (v[k] != 0.0f) mn = fminf(mn, fabsf(v[k]));
  s = warp_sum(s);
  mn = warp_min(mn);
}

// Same for a group entry: lane l holds the (sum, min) of leaves l and l + 32 of
// the group (0 / FLT_MAX beyond the group's last leaf).
__device__ __forceinline__ void reduce_group(double s0, double s1, float m0, float m1, double& s, float& mn) {
  s = warp_sum(s0 + s1);
  mn = warp_min(fminf(m0, m1));
}

// One warp recomputes leaf `lf` from the weights in global memory.
__device__ __forceinline__ void retile_leaf(const float* w, const Tiles& tl, int lf, int lane) {
  const float4* src = reinterpret_cast<const float4*>(w + ((size_t)lf << kLeafLog2)) + lane * 2;
  const float4 a = src[0], b = src[1];
  double s;
  float mn;
  reduce256(a, b, s, mn);
  if (lane == 0) {
    tl.leaf[lf] = s;
    tl.leafmin[lf] = mn;
  }
}

// One warp recomputes group g from its leaf entries in global memory.
__device__ __forceinline__ void regroup(const Tiles& tl, int g, int lane) {
  const int L = 1 << tl.group_log2;
  const size_t lb = (size_t)g << tl.group_log2;
  const double s0 = lane < L ? tl.leaf[lb + lane] : 0.0;
  const double s1 = lane + 32 < L ? tl.leaf[lb + 32 + lane] : 0.0;
  const float m0 = lane < L ? tl.leafmin[lb + lane] : FLT_MAX;
  const float m1 = lane + 32 < L ? tl.leafmin[lb + 32 + lane] : FLT_MAX;
  double s;
  float mn;
  reduce_group(s0, s1, m0, m1, s, mn);
  if (lane == 0) {
    tl.grp[g] = s;
    tl.grpmin[g] = mn;
  }
}

// The leaves (shift = 8) or groups (shift = 8 + group_log2) overlapping ring
// range [start, start + n) (mod S), one warp per entry. The caller has made the
// inputs (weights, resp. leaf entries) of the range visible to the CTA.
template <bool kGroups>
__device__ __forceinline__ void refresh_range(const float* w, const Tiles& tl, int S, int start, int n, int warp,
                                              int nwarps, int lane) {
  const int shift = kGroups ? kLeafLog2 + tl.group_log2 : kLeafLog2;
  const int n1 = min(n, S - start);
  for (int seg = 0; seg < 2; ++seg) {
    const int lo = seg == 0 ? start : 0;
    const int len = seg == 0 ? n1 : n - n1;
    if (len <= 0) continue;
    const int t1 = (lo + len - 1) >> shift;
    for (int t = (lo >> shift) + warp; t <= t1; t += nwarps) {
      if (kGroups) regroup(tl, t, lane); else retile_leaf(w, tl, t, lane);
    }
  }
}

// Full build: one CTA of 256 threads per 2048 slots, a warp per leaf; then one
// warp per group. Runs at creation and after a checkpoint load (and on demand
// for the debug cross-check): afterwards the entries are maintained by the
// kernels that change a weight.
__global__ void __launch_bounds__(256) build_leaves_kernel(const float* __restrict__ w, const Tiles tl) {
  pdl_wait();
  pdl_launch_dependents();
  retile_leaf(w, tl, (int)blockIdx.x * (kBuildSpan >> kLeafLog2) + (int)(threadIdx.x >> 5), (int)(threadIdx.x & 31));
}

__global__ void __launch_bounds__(256) build_groups_kernel(const Tiles tl) {
  pdl_wait();
  pdl_launch_dependents();
  const int g = (int)blockIdx.x * 8 + (int)(threadIdx.x >> 5);
  if (g < tl.ng) regroup(tl, g, (int)(threadIdx.x & 31));
}

// ------------------------------------------------------------ mt19937 (K4)
// std::mt19937 regenerated in parallel: the twist of word k needs old[k],
// old[k+1] and word k+397 (mod 624), which is old for k < 227 and already new
// after that, hence three phases + the last word.
__device__ __forceinline__ uint32_t mt_mix(uint32_t a, uint32_t b, uint32_t m) {
  uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
  uint32_t v = m ^ (y >> 1);
  if (y & 1u) v ^= 0x9908b0dfu;
  return v;
}

__device__ void mt_twist(uint32_t* mt) {
  const int tid = threadIdx.x, nthr = blockDim.x;
  uint32_t v[2] = {0u, 0u};  // blockDim.x >= 227: at most two words per thread per phase
  int n = 0;
  for (int k = tid; k < 227; k += nthr) v[n++] = mt_mix(mt[k], mt[k + 1], mt[k + 397]);
  __syncthreads();
  n = 0;
  for (int k = tid; k < 227; k += nthr) mt[k] = v[n++];
  __syncthreads();
  n = 0;
  for (int k = 227 + tid; k < 454; k += nthr) v[n++] = mt_mix(mt[k], mt[k + 1], mt[k - 227]);
  __syncthreads();
  n = 0;
  for (int k = 227 + tid; k < 454; k += nthr) mt[k] = v[n++];
  __syncthreads();
  n = 0;
  for (int k = 454 + tid; k < 623; k += nthr) v[n++] = mt_mix(mt[k], mt[k + 1], mt[k - 227]);
  __syncthreads();
  n = 0;
  for (int k = 454 + tid; k < 623; k += nthr) mt[k] = v[n++];
  __syncthreads();
  if (tid == 0) mt[623] = mt_mix(mt[623], mt[0], mt[396]);
  __syncthreads();
}

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

// Uniform draws d_lo .. d_hi-1 of a batch of B (rela/prioritized_replay.h:267,279) from the
// mt19937 state `mt` / `mti` in shared memory: libstdc++'s generate_canonical<float, 24> (one
// 32-bit draw, * 2^-32, clamp to nextafter(1, 0)). The state is twisted only as far as draw
// d_hi - 1 needs (a batch of 4096 spans 7 twists; the first CTAs need none of them), so `mt` holds
// the state after all B draws only if d_hi == B; *mti_out is the index after all B draws either
// way. Called by every thread of the CTA (the twist synchronises).
__device__ void gen_uniforms(uint32_t* mt, int mti, int B, int d_lo, int d_hi, float* out, int* mti_out) {
  const int gen_end = (mti + d_hi - 1) / 624;  // twists this range needs
  for (int gen = 0; gen <= gen_end; ++gen) {
    if (gen > 0) mt_twist(mt);
    for (int d = d_lo + (int)threadIdx.x; d < d_hi; d += (int)blockDim.x) {
      const int g = mti + d;
      if (g / 624 != gen) continue;
      const uint32_t y = mt_temper(mt[g % 624]);
      float u = __fmul_rn(__uint2float_rn(y), 2.3283064365386963e-10f);
      if (u >= 1.0f) u = 0.99999994f;  // nextafterf(1, 0)
      out[d - d_lo] = u;
    }
    __syncthreads();
  }
  *mti_out = (mti + B - 1) % 624 + 1;
}

// Target d of the batch: rand = u * segment + d * segment (the reference binary contracts
// `dist(rng_) + i * segment` into FMAs), clamped to sum - 0.2 (lines 265-280).
__device__ __forceinline__ float target_of(float u, int d, float sumf, int B, int contract_fma) {
  const float segment = __fdiv_rn(sumf, (float)B);
  const float cap = __fsub_rn(sumf, 0.2f);
  float r;
  if (contract_fma) {
    r = __fmaf_rn((float)d, segment, __fmaf_rn(u, segment, 0.0f));
  } else {
    r = __fadd_rn(__fadd_rn(__fmul_rn(u, segment), 0.0f), __fmul_rn((float)d, segment));
  }
  return (cap < r) ? cap : r;
}

}  // namespace

namespace {

// Raw importance weight of one selected element (rela/prioritized_replay.h:320-321):
// weights = w / sum; weights = pow(size * weights, -beta), in float like ATen,
// the pow evaluated in fp64 and rounded once.
__device__ __forceinline__ float is_raw(float wv, float sumf, int size_is, float beta) {
  float v = __fdiv_rn(wv, sumf);
  v = __fmul_rn((float)size_is, v);
  return (float)pow((double)v, (double)(-beta));
}

__device__ __forceinline__ void publish_max(unsigned long long* ismax, uint32_t epoch, float v) {
  // v >= +0 (or +inf): the float order is the order of the bit patterns. NaN is skipped like fmaxf does.
  if (v == v) atomicMax(ismax, ((unsigned long long)epoch << 32) | (unsigned long long)__float_as_uint(v));
}

__device__ __forceinline__ void raise_error(const SelectArgs& a, int code) {
  a.ctl->error = code;  // reference: assert(false) at lines 297-302
  if (a.host_err != nullptr) {
    *reinterpret_cast<volatile int*>(a.host_err) = code;
    __threadfence_system();
  }
}

}  // namespace

// -------------------------------------------------------------- K4/K5 select
// Grid: one CTA of 256 threads per 8 targets, one warp per target. Every CTA
// pulls the group sums and minima (1.3 KB + 0.6 KB at 2^21 transitions; at most
// 24 KB) and the mt19937 state into shared memory with three bulk async copies,
// scans the group sums, regenerates the mt19937 stream for its own targets, and
// then each warp walks group -> leaf -> slot for its target: 64 leaf sums, then
// 256 weights, both read coalesced from L2. The CTAs never wait for each other:
//   * importance weights leave RAW with an atomic max (SelectArgs::israw);
//   * the RNG state is double-buffered, CTA 0 writes the successor state;
//   * only a call that has to evict (pop_n > 0) elects a last CTA (ticket in
//     Ctl::done_ctas) to pop after every CTA's marks, like the reference.
//
// All prefix arithmetic is in fp64 on the weights themselves. Under the
// exactness condition (every weight a multiple of u = ulp(min non-zero weight),
// total < 2^52 u) every partial sum formed here is a multiple of u below
// 2^53 u, hence exact and independent of the association. Conversely, if any
// addition anywhere (leaf and group sums included) had to round, its result is
// >= 2^53 u and so is the total: the single test `total < 2^52 u` detects every
// inexact case, and those take the ordered walk.
//
// Dynamic shared memory: group prefixes (8 B x groups) | group minima (4 B x groups).
__global__ void __launch_bounds__(kSelThreads) select_kernel(const SelectArgs a) {
  extern __shared__ __align__(128) uint8_t sm_raw[];
  const Tiles& tl = a.t;
  const int ng = tl.ng;
  const int ng_pad = (ng + 3) & ~3;
  const int gl = tl.group_log2;
  const int span_log2 = kLeafLog2 + gl;  // slots per group
  double* gpre = reinterpret_cast<double*>(sm_raw);
  float* gmin = reinterpret_cast<float*>(sm_raw + (size_t)ng_pad * sizeof(double));
  __shared__ __align__(16) uint32_t mt[624];
  __shared__ __align__(8) uint64_t bar, bar_rng;
  __shared__ double red_d[kSelThreads / 32];
  __shared__ float red_f[kSelThreads / 32];
  __shared__ float s_target[kSelTargets];
  __shared__ int s_grp[kSelTargets];
  __shared__ int s_wrapped[kSelTargets];
  __shared__ double s_rem[kSelTargets];
  __shared__ double sh_offset;
  __shared__ int sh_last, sh_mti;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  constexpr int kWarps = kSelThreads / 32;
  static_assert(kWarps == kSelTargets, "one warp per target");
  Ctl* ctl = a.ctl;
  const int i0 = (int)blockIdx.x * kSelTargets;
  const int nT = min(kSelTargets, a.B - i0);
  const bool probing = blockIdx.x == 0 && tid == 0;

  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_init(&bar_rng, 1);
    mbar_fence_init();
  }
  __syncthreads();
  // ---- 0. the uniform draws need nothing but the RNG state, which only selections write. When
  //         the previous selection is known to be complete (rng_early: it is at least two kernels
  //         back on this stream and every kernel since triggers its dependents after its own wait)
  //         they are generated BEFORE the wait, while the kernel in front is still running.
  int mti0 = 0;
  if (a.rng_early) {
    if (tid == 0) {
      mbar_expect_tx(&bar_rng, 624u * 4u);
      bulk_g2s(mt, ctl->mt[a.rng_par], 624u * 4u, &bar_rng);
    }
    mti0 = ctl->mti[a.rng_par];
    mbar_wait(&bar_rng, 0);
    gen_uniforms(mt, mti0, a.B, i0, i0 + nT, s_target, &sh_mti);
  }
  if (probing) RR_TL_STAMP_TAG(ctl, kTlSelectEnter, a.epoch);
  pdl_wait();
  pdl_launch_dependents();  // after the wait: see `rng_early`, and `early` of update_kernel
  if (probing) {
    ctl->probe[0] = clock64();
    RR_TL_STAMP_TAG(ctl, kTlSelectBegin, a.epoch);
    if (*reinterpret_cast<volatile unsigned int*>(&ctl->tl_on) != 0u) {
      const unsigned long long um = atomicExch(&ctl->tl_umax, 0ull), gm = atomicExch(&ctl->tl_gmax, 0ull);
      if (um) ctl->tl[atomicAdd(&ctl->tl_n, 1u) % kTimelineCap] = (um << 8) | kTlUpdateAllEnd;
      if (gm) ctl->tl[atomicAdd(&ctl->tl_n, 1u) % kTimelineCap] = (gm << 8) | kTlGatherAllEnd;
    }
  }
  // ---- 1. group sums and minima (and the RNG state, if not yet) by bulk copies (TMA)
  if (tid == 0) {
    const uint32_t b2 = (uint32_t)ng_pad * 8u, bm = (uint32_t)ng_pad * 4u, br = a.rng_early ? 0u : 624u * 4u;
    mbar_expect_tx(&bar, b2 + bm + br);
    bulk_g2s(gpre, tl.grp, b2, &bar);
    bulk_g2s(gmin, tl.grpmin, bm, &bar);
    if (!a.rng_early) bulk_g2s(mt, ctl->mt[a.rng_par], br, &bar);
  }
  const double sum_d = ctl->sum;
  if (!a.rng_early) mti0 = ctl->mti[a.rng_par];
  // mass in front of head inside head's group: the slots of head's leaf and the
  // whole leaves of head's group that precede it (for the offset O below)
  double head_part = 0.0;
  if (warp == kWarps - 1 && a.head > 0) {
    const int last = a.head - 1;
    for (int k = ((last >> kLeafLog2) << kLeafLog2) + lane; k <= last; k += 32) head_part += (double)a.w[k];
    const int l1 = last >> kLeafLog2, l0 = (l1 >> gl) << gl;
    for (int k = l0 + lane; k < l1; k += 32) head_part += tl.leaf[k];
  }
  mbar_wait(&bar, 0);
  float mn = FLT_MAX;
  for (int i = tid; i < ng; i += kSelThreads) mn = fminf(mn, gmin[i]);
  mn = warp_min(mn);
  if (lane == 0) red_f[warp] = mn;
  __syncthreads();
  mn = warp_min(lane < kWarps ? red_f[lane] : FLT_MAX);  // every warp now holds the global min non-zero weight
  int ue = -149;
  if (mn != FLT_MAX) {
    int e;
    frexpf(mn, &e);       // mn = f * 2^e, f in [0.5, 1)
    ue = (e - 1) - 23;    // ulp of a normal float
    if (ue < -149) ue = -149;
  }
  const double unit = ldexp(1.0, ue);
  const float sumf = (float)sum_d;

  // ---- 2. this CTA's targets: uniform draws (if not yet), scaled by sum_ / B
  if (!a.rng_early) gen_uniforms(mt, mti0, a.B, i0, i0 + nT, s_target, &sh_mti);
  if (tid < nT) s_target[tid] = target_of(s_target[tid], i0 + tid, sumf, a.B, a.contract_fma);
  if (probing) ctl->probe[1] = clock64();

  // ---- 3. inclusive prefix of the group sums: contiguous chunk per thread (odd
  //         length: conflict-free 8-byte accesses), block scan of the chunk totals
  const int per = ((ng + kSelThreads - 1) / kSelThreads) | 1;
  const int c0 = min(tid * per, ng), c1 = min(c0 + per, ng);
  double run = 0.0;
  for (int i = c0; i < c1; ++i) {
    run += gpre[i];
    gpre[i] = run;
  }
  const double x = warp_scan(run, lane);
  if (lane == 31) red_d[warp] = x;
  __syncthreads();
  if (warp == 0) {
    const double t = warp_scan(lane < kWarps ? red_d[lane] : 0.0, lane);
    if (lane < kWarps) red_d[lane] = t;
  }
  __syncthreads();
  {
    const double off = (warp > 0 ? red_d[warp - 1] : 0.0) + (x - run);
    for (int i = c0; i < c1; ++i) gpre[i] += off;
  }
  __syncthreads();
  const double total = gpre[ng - 1];
  const bool exact = (total < ldexp(1.0, ue + 52)) && !a.force_ordered;  // false for NaN / Inf totals too
  if (probing) ctl->probe[2] = clock64();

  if (exact) {
    // ---- 4. offset of head: physical inclusive prefix at head - 1
    if (warp == kWarps - 1) {
      double part = 0.0;
      if (a.head > 0) {
        part = warp_sum(head_part);
        const int g = (a.head - 1) >> span_log2;
        if (g > 0) part += gpre[g - 1];
      }
      if (lane == 0) sh_offset = part;
    }
    __syncthreads();
    // ---- 5. group of each target. Both cases reduce to "first physical slot
    //   whose prefix reaches a key":
    //   unwrapped part [head, S):  P(s) - O >= t'   <=>  P(s) >= t' + O
    //   wrapped part   [0, head):  M1 + P(s) >= t'  <=>  P(s) >= t' - M1
    // with t' = max(target rounded up to a multiple of u, u) (the reference's
    // accSum > 0 && accSum >= target; every prefix is a multiple of u),
    // O = prefix at head - 1 and M1 = total - O the mass of the unwrapped part.
    if (tid < nT) {
      const double off = sh_offset;
      const double m1 = total - off;
      const float tf = s_target[tid];
      double tq = unit;
      if (tf > 0.0f) tq = fmax(ldexp(ceil(ldexp((double)tf, -ue)), ue), unit);  // exact: power-of-two scalings
      const int wrapped = !(m1 >= tq);
      const double key = wrapped ? tq - m1 : tq + off;
      int lo = 0, hi = ng;  // first group whose inclusive prefix reaches the key
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (gpre[mid] >= key) hi = mid; else lo = mid + 1;
      }
      s_grp[tid] = lo < ng ? lo : -1;
      s_rem[tid] = lo < ng ? key - (lo > 0 ? gpre[lo - 1] : 0.0) : 0.0;  // u <= rem <= group sum
      s_wrapped[tid] = wrapped;
    }
    __syncthreads();
    if (probing) ctl->probe[3] = clock64();
    // ---- 6. one warp per target: leaf inside the group (lane l holds the sums of
    //   leaves l and l + 32, two warp scans), then the slot inside the leaf (lane l
    //   owns 8 consecutive slots kept in registers; a warp scan of the lane totals
    //   finds the lane, which walks its own slots).
    const int t = warp;
    if (t < nT) {  // warp-uniform
      const int g = s_grp[t];
      double rem = s_rem[t];
      int leaf = -1;
      if (g >= 0) {
        const int L = 1 << gl;
        const size_t lb = (size_t)g << gl;
        const double x0 = lane < L ? tl.leaf[lb + lane] : 0.0;
        const double x1 = lane + 32 < L ? tl.leaf[lb + 32 + lane] : 0.0;
        const double P0 = warp_scan(x0, lane);
        const double P1 = warp_scan(x1, lane) + __shfl_sync(0xffffffffu, P0, 31);
        const unsigned h0 = __ballot_sync(0xffffffffu, lane < L && P0 >= rem);
        const unsigned h1 = __ballot_sync(0xffffffffu, lane + 32 < L && P1 >= rem);
        if (h0 != 0u) {
          const int Lh = __ffs(h0) - 1;
          rem -= __shfl_sync(0xffffffffu, P0 - x0, Lh);
          leaf = (int)lb + Lh;
        } else if (h1 != 0u) {
          const int Lh = __ffs(h1) - 1;
          rem -= __shfl_sync(0xffffffffu, P1 - x1, Lh);
          leaf = (int)lb + 32 + Lh;
        }
      }
      int slot = -1;
      float wv = 0.0f;
      if (leaf >= 0) {
        const float* wp = a.w + ((size_t)leaf << kLeafLog2) + (size_t)lane * 8;
        const float4 v0 = *reinterpret_cast<const float4*>(wp);
        const float4 v1 = *reinterpret_cast<const float4*>(wp + 4);
        const double T = (((double)v0.x + (double)v0.y) + ((double)v0.z + (double)v0.w)) +
                         (((double)v1.x + (double)v1.y) + ((double)v1.z + (double)v1.w));
        const double P = warp_scan(T, lane);
        const unsigned hit = __ballot_sync(0xffffffffu, P >= rem);
        if (hit != 0u) {
          const int Lh = __ffs(hit) - 1;
          if (lane == Lh) {
            const double rem_l = rem - (P - T);
            double acc = 0.0;
            int j = 7;
            const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            int found = 0;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              acc += (double)vv[q];
              if (!found && acc >= rem_l) {
                found = 1;
                j = q;
                wv = vv[q];
              }
            }
            if (!found) wv = vv[7];
            slot = (leaf << kLeafLog2) + Lh * 8 + j;
          }
          slot = __shfl_sync(0xffffffffu, slot, Lh);
          wv = __shfl_sync(0xffffffffu, wv, Lh);
        }
      }
      if (lane == 0) {
        if (slot >= 0 && s_wrapped[t] && slot >= a.head) slot = -1;  // ran past the wrapped part
        int logical = -1;
        if (slot >= 0) logical = slot >= a.head ? slot - a.head : slot + a.S - a.head;
        if (slot < 0 || slot >= a.S || logical >= a.size) {
          raise_error(a, RR_ERR_OVERRUN);
          slot = a.head;
          wv = a.w[slot];
          logical = 0;
        }
        a.ids[i0 + t] = slot;
        const float v = is_raw(wv, sumf, a.size_is, a.beta);
        a.israw[i0 + t] = v;
        publish_max(a.ismax, a.epoch, v);
        // getElementAndMark (line 126). A slot that this call also evicts ends
        // up evicted (the pop comes after the marks): leave it to the eviction.
        if (logical >= a.pop_n) a.evicted[slot] = 0;
      }
    }
  } else if (blockIdx.x == 0) {
    // ordered walk: the reference's loop (lines 274-308) for ALL targets, staged
    // kSelThreads weights at a time and consumed by one thread in order. The
    // targets live in a.israw (global); entry i is replaced by the selected weight
    // once target i has been served and by the raw importance weight at the end.
    volatile float* targets = a.israw;
    __shared__ float chunk[kSelThreads];
    __shared__ int sh_next_i, sh_id;
    __shared__ double sh_acc;
    __shared__ float sh_w;
    __syncthreads();
    for (int i = tid; i < 624; i += kSelThreads) mt[i] = ctl->mt[a.rng_par][i];
    if (tid == 0) {
      sh_next_i = 0;
      sh_acc = 0.0;
      sh_w = 0.0f;
      sh_id = 0;
    }
    __syncthreads();
    gen_uniforms(mt, mti0, a.B, 0, a.B, a.israw, &sh_mti);
    __threadfence_block();
    __syncthreads();
    for (int i = tid; i < a.B; i += kSelThreads) targets[i] = target_of(targets[i], i, sumf, a.B, a.contract_fma);
    __threadfence_block();
    __syncthreads();
    const int size = a.size;
    for (int base = 0; base <= size; base += kSelThreads) {
      const int idx = base + tid;
      if (idx < size) {
        int slot = a.head + idx;
        if (slot >= a.S) slot -= a.S;
        chunk[tid] = a.w[slot];
      }
      __syncthreads();
      if (tid == 0) {
        int i = sh_next_i;
        double acc = sh_acc;
        float w = sh_w;
        int id = sh_id;
        int nextIdx = base;
        const int lim = min(size, base + kSelThreads);
        while (i < a.B) {
          const double target = (double)targets[i];
          if (acc > 0.0 && acc >= target) {
            a.ids[i] = id;
            targets[i] = w;
            ++i;
            continue;
          }
          if (nextIdx == lim) break;
          id = a.head + nextIdx;
          if (id >= a.S) id -= a.S;
          w = chunk[nextIdx - base];
          acc += (double)w;
          ++nextIdx;
        }
        sh_next_i = i;
        sh_acc = acc;
        sh_w = w;
        sh_id = id;
      }
      __syncthreads();
      if (sh_next_i >= a.B) break;
    }
    if (tid == 0 && sh_next_i < a.B) {
      raise_error(a, RR_ERR_OVERRUN);
      for (int i = sh_next_i; i < a.B; ++i) {
        a.ids[i] = a.head;
        targets[i] = a.w[a.head];
      }
    }
    __threadfence_block();
    __syncthreads();
    for (int i = tid; i < a.B; i += kSelThreads) {
      const int slot = a.ids[i];
      const int logical = slot >= a.head ? slot - a.head : slot + a.S - a.head;
      if (logical >= a.pop_n) a.evicted[slot] = 0;
      const float v = is_raw(targets[i], sumf, a.size_is, a.beta);
      targets[i] = v;
      publish_max(a.ismax, a.epoch, v);
    }
  }
  if (probing) {
    ctl->probe[4] = clock64();
    RR_TL_STAMP_TAG(ctl, kTlSelectEnd, a.epoch);
  }

  // ---- 7. RNG state after the B draws: the CTA of the last targets (the only one that had to
  //         twist all the way); counters: CTA 0
  if (blockIdx.x == gridDim.x - 1) {
    __syncthreads();
    for (int i = tid; i < 624; i += kSelThreads) ctl->mt[a.rng_par ^ 1][i] = mt[i];
    if (tid == 0) ctl->mti[a.rng_par ^ 1] = sh_mti;
  }
  if (blockIdx.x == 0 && tid == 0) {
    if (exact) ctl->n_exact += 1; else ctl->n_ordered += 1;
    ctl->probe[5] = clock64();
  }
  if (blockIdx.x == 0 && tid < a.wait_n) {  // SelectArgs::wait_flags: normally long published
    const unsigned long long t0 = globaltimer_ns();
    // (once the replay is in error -- a wait somewhere has timed out -- no kernel waits again: the
    // host raises at its next call, and what is queued until then drains at once)
    const bool failed = *reinterpret_cast<volatile int*>(&ctl->error) != 0;
    for (; !failed;) {
      unsigned long long v;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(a.wait_flags + tid) : "memory");
      if (v >= a.wait_val) break;
      if (globaltimer_ns() - t0 > 10ull * 1000 * 1000 * 1000) {
        raise_error(a, RR_ERR_CUDA);  // the gather of the batch taken before this selection never finished
        break;
      }
      __nanosleep(100);
    }
  }
  if (a.pop_n <= 0) return;

  // ---- 8. eviction (blockPop, lines 84-103) after every CTA's marks, like the
  //         reference: the last CTA to get here pops
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    sh_last = atomicAdd(&ctl->done_ctas, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!sh_last) return;
  __threadfence();
  if (tid == 0) ctl->probe[6] = clock64();
  if (exact) {
    double part = 0.0;  // exact in any order under the exactness condition
    for (int j = tid; j < a.pop_n; j += kSelThreads) {
      int slot = a.head + j;
      if (slot >= a.S) slot -= a.S;
      part += (double)__ldcg(a.w + slot);
    }
    part = warp_sum(part);
    if (lane == 0) red_d[warp] = part;
    __syncthreads();
    if (tid == 0) {
      double tot = 0.0;
      for (int k = 0; k < kWarps; ++k) tot += red_d[k];
      // diff = -(w_0 + w_1 + ...) exactly, then sum_ += diff (line 96)
      ctl->sum = sum_d + -tot;
    }
  } else if (tid == 0) {
    double diff = 0.0;
    for (int j = 0; j < a.pop_n; ++j) {
      int slot = a.head + j;
      if (slot >= a.S) slot -= a.S;
      diff -= (double)__ldcg(a.w + slot);
    }
    ctl->sum = sum_d + diff;
  }
  __syncthreads();
  for (int j = tid; j < a.pop_n; j += kSelThreads) {
    int slot = a.head + j;
    if (slot >= a.S) slot -= a.S;
    a.evicted[slot] = 1;
    a.w[slot] = 0.0f;
  }
  __syncthreads();
  refresh_range<false>(a.w, tl, a.S, a.head, a.pop_n, warp, kWarps, lane);
  __syncthreads();
  refresh_range<true>(a.w, tl, a.S, a.head, a.pop_n, warp, kWarps, lane);
  if (tid == 0) {
    ctl->n_popped += (unsigned long long)a.pop_n;
    ctl->done_ctas = 0u;
    ctl->probe[7] = clock64();
  }
}

// ------------------------------------------------------------- add commit
// weights_[j] = priority^alpha for the block's slots, float block sum in index
// order, sum_ += (double) blockSum (rela/prioritized_replay.h:58-73,188).
__global__ void __launch_bounds__(1024) commit_kernel(Ctl* ctl, float* w, const Tiles tiles, int S, int start, int n,
                                                      const float* __restrict__ prio, float alpha,
                                                      int is_weight) {
  __shared__ __align__(16) float buf[1024];
  __shared__ float sh_sum;
  const int tid = threadIdx.x;
  if (tid == 0) sh_sum = 0.0f;
  for (int base = 0; base < n; base += 1024) {
    const int i = base + tid;
    float v = 0.0f;
    if (i < n) {
      const float p = prio[i];
      v = is_weight ? p : pow_alpha(p, alpha);
      int slot = start + i;
      if (slot >= S) slot -= S;
      w[slot] = v;
    }
    buf[tid] = v;
    __syncthreads();
    if (tid == 0) {
      // the reference's sum is a serial float sum in index order (line 63): one
      // dependent add per element, the operands fetched four at a time
      float s = sh_sum;
      const int m = min(1024, n - base);
      int k = 0;
      for (; k + 4 <= m; k += 4) {
        const float4 q = *reinterpret_cast<const float4*>(buf + k);
        s = __fadd_rn(s, q.x);
        s = __fadd_rn(s, q.y);
        s = __fadd_rn(s, q.z);
        s = __fadd_rn(s, q.w);
      }
      for (; k < m; ++k) s = __fadd_rn(s, buf[k]);
      sh_sum = s;
    }
    __syncthreads();
  }
  if (tid == 0) ctl->sum += (double)sh_sum;
  // the block's weights are visible to the CTA (barrier above): refresh its leaves, then its groups
  refresh_range<false>(w, tiles, S, start, n, tid >> 5, 32, tid & 31);
  __syncthreads();
  refresh_range<true>(w, tiles, S, start, n, tid >> 5, 32, tid & 31);
}

// ------------------------------------------------------ K7 priority update
// ConcurrentQueue::update (rela/prioritized_replay.h:105-119): in index order,
// skip evicted ids, diff += (float)(new - old), weights_[id] = new; a repeated
// id sees the value written by its previous occurrence. Sampled ids are
// non-decreasing in ring order (the reference's walk never rewinds), so
// repeats are always adjacent: occurrence i reads new[i-1], the last one writes.
// The double accumulation of the float differences is done in parallel when
// it is provably exact (sum |d| < 2^52 * ulp(min |d|)), else in order.
//
// CTA 0 does that bookkeeping. CTAs 1.. keep the hierarchy current, one CTA per
// touched GROUP: its warps reload the leaves of the group that hold a batch
// slot, substitute the final value of every batch slot (computed from the same
// inputs as CTA 0, so the result does not depend on whether CTA 0's store has
// landed), reduce them again, and after a CTA barrier one warp re-reduces the
// group from its 64 leaf entries. Nothing waits across CTAs.
// Candidate c = blockIdx.x - 1 is group c when there are at most n groups
// (by_group), else the group of batch element c, handled by the CTA of the
// first element that falls into it.
//
// The new priorities come from device memory or, for a learner-sized batch
// handed over in host memory, ride in the launch packet itself (PrioInline, a
// __grid_constant__ parameter read through the constant bank): no staging
// copy sits between updatePriority and the next sample on the stream.
constexpr int kUpdThreads = 256;

struct PrioInline {
  float v[kInlinePrio];
};
__device__ __forceinline__ float prio_at(const float* p, int i) { return p[i]; }
__device__ __forceinline__ float prio_at(const PrioInline& p, int i) { return p.v[i]; }

template <typename Prio>
struct PrioTraits {
  static constexpr bool kInLaunch = false;  // device memory: may be produced by the kernel in front of this one
};
template <>
struct PrioTraits<PrioInline> {
  static constexpr bool kInLaunch = true;   // part of the launch packet: readable at once
};

// `early`: the selection that wrote `ids` finished at least two kernels ago on this stream, and
// every kernel since triggers its dependents only after its own wait -- the ids (and priorities
// that ride in the launch) can be staged BEFORE griddepcontrol.wait, i.e. while the kernel in
// front of this one is still running. The host sets it only when it knows (replay.cu).
template <typename Prio>
__global__ void __launch_bounds__(kUpdThreads) update_kernel(Ctl* ctl, float* w, const uint8_t* __restrict__ evicted,
                                                             const int32_t* __restrict__ ids, int n,
                                                             const __grid_constant__ Prio prio, float alpha,
                                                             int is_weight, const Tiles tiles, int by_group,
                                                             int early, const ExchangeArgs xa, uint8_t* scratch) {
  struct EndStamp {  // debug timeline: the latest end of any CTA of this kernel
    Ctl* c;
    __device__ ~EndStamp() {
      if (threadIdx.x == 0 && *reinterpret_cast<volatile unsigned int*>(&c->tl_on) != 0u)
        atomicMax(&c->tl_umax, globaltimer_ns());
    }
  } end_stamp{ctl};
  extern __shared__ __align__(128) uint8_t sm_raw[];
  // A learner-sized batch (kUpdStaged) is staged in shared memory by every CTA. A larger one (the
  // batch of all learners of a striped replay) is not: two resident CTAs per SM must fit into the
  // shared memory a gather drawn ahead leaves free (kGatherReserve), or the CTAs polling the
  // exchange flags below keep the gathers -- whose rows the other shards are waiting for -- off
  // the SMs. The group CTAs then read ids and priorities from L2, CTA 0 works in `scratch`.
  const bool staged = scratch == nullptr;
  uint8_t* work = staged ? sm_raw : scratch;
  int32_t* s_id = reinterpret_cast<int32_t*>(work);  // slot id; -1 once it is known to be evicted
  float* s_new = reinterpret_cast<float*>(s_id + n);
  float* s_diff = s_new + n;
  __shared__ double r_sum[kUpdThreads / 32], r_abs[kUpdThreads / 32];
  __shared__ float r_min[kUpdThreads / 32];
  __shared__ int sh_exact;
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  constexpr int kWarps = kUpdThreads / 32;
  const bool prio_early = staged && early && PrioTraits<Prio>::kInLaunch && !xa.on;
  if (early && staged) {
    for (int i = tid; i < n; i += kUpdThreads) {
      s_id[i] = ids[i];
      if (prio_early) {
        const float p = prio_at(prio, i);
        s_new[i] = is_weight ? p : pow_alpha(p, alpha);
      }
    }
  }
  if (blockIdx.x == 0 && tid == 0) RR_TL_STAMP_TAG(ctl, kTlUpdateEnter, xa.epoch);
  pdl_wait();
  // the trigger comes AFTER the wait: a kernel that starts early behind this one may rely on
  // everything this kernel waited for being complete (see `early`, and select_kernel's rng_early)
  pdl_launch_dependents();
  if (blockIdx.x == 0 && tid == 0) RR_TL_STAMP_TAG(ctl, kTlUpdateBegin, xa.epoch);
  if (xa.on && blockIdx.x == 0) {
    // striped replay: all-gather of the learners' priorities through peer memory (ExchangeArgs)
    for (int p = 0; p < xa.n_dst; ++p)
      for (int i = tid; i < xa.n_mine; i += kUpdThreads) {
        const unsigned long long word =
            ((unsigned long long)xa.epoch << 32) | (unsigned long long)__float_as_uint(prio_at(prio, i));
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(xa.dst[p] + i), "l"(word) : "memory");
      }
  }
  // priority j of the exchanged block: poll its word until it carries this exchange's epoch
  auto exchanged = [&](int j) -> float {
    const unsigned long long* slot = xa.slots + j;
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
    if ((unsigned int)(v >> 32) != xa.epoch) {
      const unsigned long long t0 = globaltimer_ns();
      // (a replay already in error does not wait again: see select_kernel's wait)
      const bool failed = *reinterpret_cast<volatile int*>(&ctl->error) != 0;
      for (; !failed;) {
        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
        if ((unsigned int)(v >> 32) == xa.epoch) break;
        if (globaltimer_ns() - t0 > 10ull * 1000 * 1000 * 1000) {
          ctl->error = 2;  // RR_ERR_CUDA: a shard never published its priorities
          if (xa.host_err != nullptr) {
            *reinterpret_cast<volatile int*>(xa.host_err) = 2;
            __threadfence_system();
          }
          break;
        }
      }
    }
    return __uint_as_float((unsigned int)v);
  };
  // CTA 0 raises every priority; a group CTA raises only the few it substitutes (a batch of 4096
  // over 1280 group CTAs would otherwise pay 5 M fp64 pows for 4096 values): there s_new is raw
  const bool raise_now = blockIdx.x == 0;
  if ((staged || blockIdx.x == 0) && (!early || !prio_early)) {
    for (int i = tid; i < n; i += kUpdThreads) {
      if (!early || !staged) s_id[i] = ids[i];
      if (!prio_early) {
        const float p = xa.on ? exchanged(i) : prio_at(prio, i);
        s_new[i] = (is_weight || !raise_now) ? p : pow_alpha(p, alpha);
      }
    }
  }
  __syncthreads();
  if (xa.on && blockIdx.x == 0 && tid == 0) RR_TL_STAMP_TAG(ctl, kTlUpdateFlagsIn, xa.epoch);

  if (blockIdx.x > 0) {
    __shared__ __align__(16) float sub[kWarps][256];
    __shared__ double s_ls[64];
    __shared__ float s_lm[64];
    __shared__ unsigned long long s_mask;
    __shared__ int s_lo, s_hi;
    const int gl = tiles.group_log2;
    const int span_log2 = kLeafLog2 + gl;
    const int ncand = by_group ? tiles.ng : n;
    auto id_at = [&](int j) -> int { return staged ? s_id[j] : ids[j]; };
    // the final weight of batch element j (only ever asked for elements this CTA substitutes)
    auto weight_at = [&](int j) -> float {
      if (staged) return (is_weight || prio_early) ? s_new[j] : pow_alpha(s_new[j], alpha);
      const float p = xa.on ? exchanged(j) : prio_at(prio, j);
      return is_weight ? p : pow_alpha(p, alpha);
    };
    // the grid is capped (two CTAs per SM, replay.cu): a CTA takes candidates c, c + G, ...
    for (int c = (int)blockIdx.x - 1; c < ncand; c += (int)gridDim.x - 1) {
    __syncthreads();  // the previous candidate's shared state is no longer read
    if (tid == 0) {
      s_mask = 0ull;
      s_lo = n;
      s_hi = -1;
    }
    int g;
    if (by_group) {
      g = c;
    } else {
      g = id_at(c) >> span_log2;  // (negative for a slot an earlier candidate of this CTA found evicted)
      int dup = g < 0 ? 1 : 0;  // an earlier element of the batch in the same group: that element's CTA covers it
      for (int j = tid; j < c; j += kUpdThreads)
        if ((id_at(j) >> span_log2) == g) dup = 1;
      if (__syncthreads_or(dup)) continue;
    }
    __syncthreads();
    // leaves of the group that hold a batch slot and the index range of the batch that falls into
    // the group -- from the raw ids: the eviction flags are still on their way. (A leaf that only
    // holds evicted batch slots is re-reduced from unchanged weights: same entry, bit for bit.)
    {
      unsigned long long m = 0ull;
      int lo = n, hi = -1;
      for (int j = tid; j < n; j += kUpdThreads) {
        const int sid = id_at(j);
        if ((sid >> span_log2) == g) {
          m |= 1ull << ((sid >> kLeafLog2) & ((1 << gl) - 1));
          lo = min(lo, j);
          hi = max(hi, j);
        }
      }
      if (m != 0ull) {
        atomicOr(&s_mask, m);
        atomicMin(&s_lo, lo);
        atomicMax(&s_hi, hi);
      }
    }
    __syncthreads();
    const unsigned long long mask = s_mask;
    if (mask == 0ull) continue;
    const int jlo = s_lo, jhi = s_hi;
    const int L = 1 << gl;
    const size_t lb = (size_t)g << gl;
    // one round trip for everything that depends on the ids: the eviction flags of this group's
    // batch slots, the weights of the touched leaves (the first one per warp), the entries of the
    // untouched leaves (nobody writes those during this kernel)
    float4 pre0 = make_float4(0.f, 0.f, 0.f, 0.f), pre1 = pre0;
    int kfirst = -1;
    for (int k = warp; k < L; k += kWarps)
      if ((mask >> k) & 1ull) {
        kfirst = k;
        break;
      }
    if (kfirst >= 0) {
      const float4* src = reinterpret_cast<const float4*>(w + ((lb + kfirst) << kLeafLog2)) + lane * 2;
      pre0 = src[0];
      pre1 = src[1];
    }
    if (tid < L && !((mask >> tid) & 1ull)) {
      s_ls[tid] = tiles.leaf[lb + tid];
      s_lm[tid] = tiles.leafmin[lb + tid];
    }
    if (staged)
      for (int j = jlo + tid; j <= jhi; j += kUpdThreads) {
        const int sid = s_id[j];
        if ((sid >> span_log2) == g && evicted[sid] != 0) s_id[j] = -1;  // an evicted slot keeps weight 0
      }
    __syncthreads();
    float* mine = sub[warp];
    for (int k = warp; k < L; k += kWarps) {
      if (!((mask >> k) & 1ull)) continue;  // warp-uniform
      const int base = (int)(lb + k) << kLeafLog2;
      if (k == kfirst) {
        reinterpret_cast<float4*>(mine)[lane * 2] = pre0;
        reinterpret_cast<float4*>(mine)[lane * 2 + 1] = pre1;
      } else {
        const float4* src = reinterpret_cast<const float4*>(w + base) + lane * 2;
        reinterpret_cast<float4*>(mine)[lane * 2] = src[0];
        reinterpret_cast<float4*>(mine)[lane * 2 + 1] = src[1];
      }
      __syncwarp();
      for (int j = jlo + lane; j <= jhi; j += 32) {
        const int sid = id_at(j);
        const bool last = !(j + 1 < n && id_at(j + 1) == sid);
        if (sid >= base && sid < base + 256 && last && (staged || evicted[sid] == 0))
          mine[sid - base] = weight_at(j);
      }
      __syncwarp();
      const float4 a = reinterpret_cast<const float4*>(mine)[lane * 2];
      const float4 b = reinterpret_cast<const float4*>(mine)[lane * 2 + 1];
      __syncwarp();
      double s;
      float mn;
      reduce256(a, b, s, mn);
      if (lane == 0) {
        tiles.leaf[lb + k] = s;
        tiles.leafmin[lb + k] = mn;
        s_ls[k] = s;
        s_lm[k] = mn;
      }
    }
    __syncthreads();
    if (warp == 0) {
      double s;
      float mn;
      reduce_group(lane < L ? s_ls[lane] : 0.0, lane + 32 < L ? s_ls[lane + 32] : 0.0,
                   lane < L ? s_lm[lane] : FLT_MAX, lane + 32 < L ? s_lm[lane + 32] : FLT_MAX, s, mn);
      if (lane == 0) {
        tiles.grp[g] = s;
        tiles.grpmin[g] = mn;
      }
    }
    }  // candidates of this CTA
    return;
  }

  // ---- CTA 0: ConcurrentQueue::update itself
  float* s_old = s_diff;  // old weight, replaced by the difference below
  for (int i = tid; i < n; i += kUpdThreads) {
    const int id = s_id[i];
    const bool ev = evicted[id] != 0;  // both dependent loads are issued together
    const float old = w[id];
    s_old[i] = old;
    if (ev) s_id[i] = -1;
  }
  __syncthreads();
  double sum = 0.0, asum = 0.0;
  float mn = FLT_MAX;
  for (int i = tid; i < n; i += kUpdThreads) {
    const int id = s_id[i];
    float d = 0.0f;
    if (id >= 0) {
      const bool has_prev = i > 0 && s_id[i - 1] == id;
      const float old = has_prev ? s_new[i - 1] : s_old[i];
      d = __fsub_rn(s_new[i], old);
      sum += (double)d;
      asum += fabs((double)d);
      if (d != 0.0f) mn = fminf(mn, fabsf(d));
    }
    s_diff[i] = d;  // same thread read s_old[i] above
  }
  __syncthreads();  // all reads of the old weights are done
  for (int i = tid; i < n; i += kUpdThreads) {
    const int id = s_id[i];
    const bool last = !(i + 1 < n && s_id[i + 1] == id);
    if (id >= 0 && last) w[id] = s_new[i];
  }
  sum = warp_sum(sum);
  asum = warp_sum(asum);
  mn = warp_min(mn);
  if (lane == 0) {
    r_sum[warp] = sum;
    r_abs[warp] = asum;
    r_min[warp] = mn;
  }
  __syncthreads();
  if (warp == 0) {
    double a = lane < kWarps ? r_sum[lane] : 0.0;
    double b = lane < kWarps ? r_abs[lane] : 0.0;
    float m = lane < kWarps ? r_min[lane] : FLT_MAX;
    a = warp_sum(a);
    b = warp_sum(b);
    m = warp_min(m);
    if (lane == 0) {
      int exact = 1;
      if (m != FLT_MAX) {
        int e;
        frexpf(m, &e);
        int ulp_exp = (e - 1) - 23;
        if (ulp_exp < -149) ulp_exp = -149;
        exact = b < ldexp(1.0, ulp_exp + 52);
      }
      sh_exact = exact;
      if (exact) ctl->sum += a;
    }
  }
  __syncthreads();
  if (!sh_exact && tid == 0) {
    double diff = 0.0;
    for (int i = 0; i < n; ++i) diff += (double)s_diff[i];  // evicted entries hold +0.0f
    ctl->sum += diff;
  }
  if (tid == 0) RR_TL_STAMP_TAG(ctl, kTlUpdateEnd, xa.epoch);
}

// ----------------------------------------------------------------- launches
cudaError_t launch_build_hierarchy(const float* w, const Tiles& t, int S_pad, cudaStream_t stream) {
  cudaError_t e = launch_pdl(build_leaves_kernel, dim3(S_pad / kBuildSpan), dim3(256), 0, stream, w, t);
  if (e != cudaSuccess) return e;
  return launch_pdl(build_groups_kernel, dim3((t.ng + 7) / 8), dim3(256), 0, stream, t);
}

cudaError_t launch_select(const SelectArgs& a, int sm_count, size_t gather_reserve, cudaStream_t stream) {
  const size_t ng_pad = ((size_t)a.t.ng + 3) & ~(size_t)3;
  const size_t smem = ng_pad * (sizeof(double) + sizeof(float)) + 16;
  cudaError_t e = ensure_dynamic_smem(select_kernel, smem);
  if (e != cudaSuccess) return e;
  const int grid = (a.B + kSelTargets - 1) / kSelTargets;
  // A selection launched as a programmatic dependent becomes resident while the update in front of
  // it is still running, and waits there holding its shared memory. In a striped replay that update
  // can spend tens of microseconds waiting for the other shards' priorities; if the waiting
  // selection CTAs plus two update CTAs per SM take more than a gather drawn ahead leaves free
  // (gather_reserve), no gather CTA gets onto those SMs meanwhile -- and the rows of the next
  // batches are exactly what the other shards are waiting for. Large striped replays (group sums
  // of 12 B per 16,384 slots in every CTA; 4 or more shards at 2^21 transitions each) therefore
  // launch the selection in plain stream order; two shards are below the bound, and an unsharded
  // replay passes no bound at all (its update waits for nobody).
  const size_t per_cta = smem + 4 * 1024;  // + the kernel's static shared memory
  const size_t per_sm = (size_t)((grid + sm_count - 1) / std::max(sm_count, 1));
  const size_t update_ctas = 2 * 10 * 1024;  // update_grid: two CTAs per SM, ~9.2 KB static each
  const bool early_start = per_sm * per_cta + update_ctas <= gather_reserve;
  return launch_pdl_if(early_start, select_kernel, dim3(grid), dim3(kSelThreads), smem, stream, a);
}

cudaError_t launch_commit(Ctl* ctl, float* w, const Tiles& tiles, int S, int start, int n, const float* prio,
                          float alpha, int is_weight, cudaStream_t stream) {
  commit_kernel<<<1, 1024, 0, stream>>>(ctl, w, tiles, S, start, n, prio, alpha, is_weight);
  return cudaGetLastError();
}

namespace {
inline int update_grid(const Tiles& tiles, int n, int* by_group) {
  *by_group = tiles.ng <= n ? 1 : 0;
  // CTA 0: sum_ and weights; the others: candidate groups, at most two CTAs per SM -- with a fused
  // exchange (ExchangeArgs) every CTA polls flags until all shards have arrived, and a grid that
  // filled every thread slot of the GPU would keep the gathers on the gather streams out meanwhile
  static const int cap = [] {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return 2 * sms;
  }();
  return 1 + std::min(*by_group ? tiles.ng : n, cap - 1);
}
}  // namespace

cudaError_t launch_update(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids, int n,
                          const float* prio, float alpha, int is_weight, int early, const ExchangeArgs* xa,
                          void* scratch, cudaStream_t stream) {
  if (n > kUpdStaged && scratch == nullptr) return cudaErrorInvalidValue;
  uint8_t* sc = n > kUpdStaged ? static_cast<uint8_t*>(scratch) : nullptr;
  const size_t smem = sc ? 0 : (size_t)n * 12;
  ExchangeArgs none;
  memset(&none, 0, sizeof(none));
  cudaError_t e = ensure_dynamic_smem(update_kernel<const float*>, smem);
  if (e != cudaSuccess) return e;
  int by_group;
  const int grid = update_grid(tiles, n, &by_group);
  return launch_pdl(update_kernel<const float*>, dim3(grid), dim3(kUpdThreads), smem, stream, ctl, w, evicted, ids, n,
                    prio, alpha, is_weight, tiles, by_group, early, xa ? *xa : none, sc);
}

// ---- update without any allocation (rr_update_priority_noalloc; no counterpart in the reference,
// added for callers that cannot call cudaMalloc).
// update_kernel in its unstaged mode works in 12 bytes per batch element (ids, new weights,
// differences). The other launches take that area from device memory the replay allocates on
// demand; this one takes it from a file-scope array, which every GPU that loads the library
// holds until exit: kNoAllocMaxBatch x 12 B = 96 KB. The kernel writes every word of the area
// it uses before reading it (CTA 0: ids -> new weights -> old weights / differences), so nothing
// has to be cleared between calls; the array keeps the last call's contents. ALL calls on one
// device share it: two such updates that overlap on the GPU (different replays, different
// streams) race. In host code `g_update_scratch` is only a stand-in for the symbol: the device
// address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_update_scratch[12 * kNoAllocMaxBatch];

cudaError_t launch_update_static(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids,
                                 int n, const float* prio, float alpha, int is_weight, cudaStream_t stream) {
  if (n < 0 || n > kNoAllocMaxBatch) return cudaErrorInvalidValue;
  static std::mutex m;
  static unsigned char* area[64];  // per device
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  unsigned char* scratch;
  {
    std::lock_guard<std::mutex> lk(m);
    if (area[dev] == nullptr) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_update_scratch);
      if (e != cudaSuccess) return e;
      area[dev] = static_cast<unsigned char*>(p);
    }
    scratch = area[dev];
  }
  ExchangeArgs none;
  memset(&none, 0, sizeof(none));
  int by_group;
  const int grid = update_grid(tiles, n, &by_group);
  // no dynamic shared memory (nothing is staged), early = 0 (always correct)
  return launch_pdl(update_kernel<const float*>, dim3(grid), dim3(kUpdThreads), 0, stream, ctl, w, evicted, ids, n,
                    prio, alpha, is_weight, tiles, by_group, 0, none, scratch);
}

cudaError_t launch_update_inline(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids,
                                 int n, const float* prio_host, float alpha, int is_weight, int early,
                                 const ExchangeArgs* xa, void* scratch, cudaStream_t stream) {
  if (n > kUpdStaged && scratch == nullptr) return cudaErrorInvalidValue;
  uint8_t* sc = n > kUpdStaged ? static_cast<uint8_t*>(scratch) : nullptr;
  const int n_host = xa ? xa->n_mine : n;  // with an exchange the launch carries this rank's share only
  if (n_host > kInlinePrio) return cudaErrorInvalidValue;
  PrioInline pin;
  memcpy(pin.v, prio_host, sizeof(float) * (size_t)n_host);
  ExchangeArgs none;
  memset(&none, 0, sizeof(none));
  const size_t smem = sc ? 0 : (size_t)n * 12;
  cudaError_t e = ensure_dynamic_smem(update_kernel<PrioInline>, smem);
  if (e != cudaSuccess) return e;
  int by_group;
  const int grid = update_grid(tiles, n, &by_group);
  return launch_pdl(update_kernel<PrioInline>, dim3(grid), dim3(kUpdThreads), smem, stream, ctl, w, evicted, ids, n,
                    pin, alpha, is_weight, tiles, by_group, early, xa ? *xa : none, sc);
}

}  // namespace rr
