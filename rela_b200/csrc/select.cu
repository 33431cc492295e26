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
// is < 2^52 * u. Then every value this file computes (tile sums, prefixes,
// differences of prefixes) is a multiple of u below 2^53 * u, hence exact in
// fp64, and the two-level search below is bit-identical to the sequential walk. When the condition fails (weights spanning > 29 binades
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

#include "../../include/rela_b200.h"
#include "kernels.h"
#include "ptx.cuh"

namespace rr {

namespace {

constexpr int kSelThreads = 512;

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

// ------------------------------------------------------------------ K3 build
// Sum (fp64) and smallest non-zero magnitude of 256 consecutive slots held as
// 8 values per lane. Every producer of a tile entry -- the full build and the
// incremental retile paths of commit / update / eviction -- goes through this
// one association, so that a retiled entry is bit-identical to a rebuilt one.
__device__ __forceinline__ void reduce256(const float4 a, const float4 b, double& s, float& mn) {
  s = (((double)a.x + (double)a.y) + ((double)a.z + (double)a.w)) +
      (((double)b.x + (double)b.y) + ((double)b.z + (double)b.w));
  mn = FLT_MAX;
  const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (v[k] != 0.0f) mn = fminf(mn, fabsf(v[k]));
  s = warp_sum(s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
}

// One warp recomputes tile t (and its leaves) from the weights in global memory.
__device__ __forceinline__ void retile_warp(const float* w, const Tiles& tl, int t, int lane) {
  const int nb = 1 << (tl.tile_log2 - 8);
  double tot = 0.0;
  float m = FLT_MAX;
  for (int k = 0; k < nb; ++k) {
    const float4* src = reinterpret_cast<const float4*>(w + ((size_t)t << tl.tile_log2) + (size_t)k * 256) + lane * 2;
    const float4 a = src[0], b = src[1];
    double s;
    float mn;
    reduce256(a, b, s, mn);
    if (lane == 0 && tl.leaf != nullptr) {
      tl.leaf[(size_t)t * nb + k] = s;
      tl.leafmin[(size_t)t * nb + k] = mn;
    }
    tot += s;
    m = fminf(m, mn);
  }
  if (lane == 0) {
    tl.l2[t] = tot;
    tl.l2min[t] = m;
  }
}

// The tiles overlapping ring range [start, start + n) (mod S), one warp per
// tile. The caller has made the new weights of the range visible to the CTA.
__device__ __forceinline__ void retile_range(const float* w, const Tiles& tl, int S, int start, int n, int warp,
                                             int nwarps, int lane) {
  const int n1 = min(n, S - start);
  for (int seg = 0; seg < 2; ++seg) {
    const int lo = seg == 0 ? start : 0;
    const int len = seg == 0 ? n1 : n - n1;
    if (len <= 0) continue;
    const int t1 = (lo + len - 1) >> tl.tile_log2;
    for (int t = (lo >> tl.tile_log2) + warp; t <= t1; t += nwarps) retile_warp(w, tl, t, lane);
  }
}

__device__ __host__ __forceinline__ Tiles tiles_of(const SelectArgs& a) {
  Tiles t;
  t.l2 = a.l2;
  t.l2min = a.l2min;
  t.leaf = a.leaf;
  t.leafmin = a.leafmin;
  t.tile_log2 = a.tile_log2;
  return t;
}

// Full build: one CTA of 256 threads per 2048 slots, a warp per 256 slots;
// tiles are 256 << k slots, i.e. 1 << k consecutive warps. Runs once at
// creation (and on demand for the debug cross-check): afterwards the tile
// entries are maintained by the kernels that change a weight.
__global__ void __launch_bounds__(256) build_hierarchy_kernel(const float* __restrict__ w, const Tiles tl) {
  double* __restrict__ l2 = tl.l2;
  float* __restrict__ l2min = tl.l2min;
  const int tile_log2 = tl.tile_log2;
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  pdl_launch_dependents();
  pdl_wait();
  const float4* src = reinterpret_cast<const float4*>(w) + (size_t)blockIdx.x * (kBuildSpan / 4) + tid * 2;
  const float4 a = src[0], b = src[1];
  double s;
  float mn;
  reduce256(a, b, s, mn);
  __shared__ double ws[8];
  __shared__ float wm[8];
  if (lane == 0) {
    ws[warp] = s;
    wm[warp] = mn;
    if (tl.leaf != nullptr) {
      tl.leaf[(size_t)blockIdx.x * 8 + warp] = s;
      tl.leafmin[(size_t)blockIdx.x * 8 + warp] = mn;
    }
  }
  __syncthreads();
  const int wpt = 1 << (tile_log2 - 8);  // warps per tile
  const int tiles = 8 / wpt;             // tiles per CTA
  if (tid < tiles) {
    double tot = 0.0;
    float m = FLT_MAX;
    for (int k = 0; k < wpt; ++k) {
      tot += ws[tid * wpt + k];
      m = fminf(m, wm[tid * wpt + k]);
    }
    l2[(size_t)blockIdx.x * tiles + tid] = tot;
    l2min[(size_t)blockIdx.x * tiles + tid] = m;
  }
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

// Targets d_lo .. d_hi-1 of a batch of B draws (rela/prioritized_replay.h:265-280)
// from the mt19937 state `mt` / `mti` in shared memory: libstdc++'s
// generate_canonical<float, 24> (one 32-bit draw, * 2^-32, clamp to
// nextafter(1, 0)) then rand = u * segment + i * segment, clamped to sum - 0.2.
// On return `mt` holds the state after all B draws and *mti_out its index.
// Called by every thread of the CTA (the twist synchronises).
__device__ void gen_targets(uint32_t* mt, int mti, int B, int d_lo, int d_hi, float sumf, int contract_fma,
                            float* out, int* mti_out) {
  const float segment = __fdiv_rn(sumf, (float)B);
  const float cap = __fsub_rn(sumf, 0.2f);
  const int gen_end = (mti + B - 1) / 624;  // twists needed for the whole batch
  for (int gen = 0; gen <= gen_end; ++gen) {
    if (gen > 0) mt_twist(mt);
    for (int d = d_lo + (int)threadIdx.x; d < d_hi; d += (int)blockDim.x) {
      const int g = mti + d;
      if (g / 624 != gen) continue;
      const uint32_t y = mt_temper(mt[g % 624]);
      float u = __fmul_rn(__uint2float_rn(y), 2.3283064365386963e-10f);
      if (u >= 1.0f) u = 0.99999994f;  // nextafterf(1, 0)
      float r;
      if (contract_fma) {
        r = __fmaf_rn((float)d, segment, __fmaf_rn(u, segment, 0.0f));
      } else {
        r = __fadd_rn(__fadd_rn(__fmul_rn(u, segment), 0.0f), __fmul_rn((float)d, segment));
      }
      out[d - d_lo] = (cap < r) ? cap : r;
    }
    __syncthreads();
  }
  *mti_out = (mti + B - 1) % 624 + 1;
}

}  // namespace

// -------------------------------------------------------------- K4/K5 select
// Grid: one CTA of 512 threads per 32 targets. Every CTA pulls the tile sums
// (at most 16 K doubles, L2-resident) into shared memory with one bulk async
// copy, scans them, and regenerates the mt19937 stream for its own targets, so
// the CTAs never wait for each other; the CTA that finishes last (ticket in
// Ctl::done_ctas) evicts, turns the selected weights into importance weights
// (block-wide max) and writes the RNG state back.
//
// All prefix arithmetic is in fp64 on the weights themselves. Under the
// exactness condition (every weight a multiple of u = ulp(min non-zero weight),
// total < 2^52 u) every partial sum formed here is a multiple of u below
// 2^53 u, hence exact and independent of the association. Conversely, if any
// addition anywhere (tile sums included) had to round, its result is >= 2^53 u
// and so is the total: the single test `total < 2^52 u` detects every inexact
// case, and those take the ordered walk.
//
// Dynamic shared memory: tile prefixes (8 B x tiles) | tile minima (4 B x tiles);
// the first region is reused for the targets of the ordered walk and for the
// importance weights.
__global__ void __launch_bounds__(kSelThreads, 1) select_kernel(const SelectArgs a) {
  extern __shared__ __align__(128) uint8_t sm_raw[];
  const int tl = a.tile_log2;
  const int nt = a.S_pad >> tl;
  const int nt_pad = (nt + 3) & ~3;
  double* p2d = reinterpret_cast<double*>(sm_raw);
  float* tmin = reinterpret_cast<float*>(sm_raw + (size_t)nt_pad * sizeof(double));
  __shared__ __align__(16) uint32_t mt[624];
  __shared__ __align__(8) uint64_t bar;
  __shared__ double red_d[kSelThreads / 32];
  __shared__ float red_f[kSelThreads / 32];
  __shared__ float s_target[kSelTargets];
  __shared__ int s_tile[kSelTargets];
  __shared__ int s_wrapped[kSelTargets];
  __shared__ double s_rem[kSelTargets];
  __shared__ double sh_offset;
  __shared__ float sh_max;
  __shared__ int sh_last, sh_mti;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  constexpr int kWarps = kSelThreads / 32;
  Ctl* ctl = a.ctl;
  const int i0 = (int)blockIdx.x * kSelTargets;
  const int nT = min(kSelTargets, a.B - i0);
  const bool probing = blockIdx.x == 0 && tid == 0;

  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  pdl_launch_dependents();
  __syncthreads();
  pdl_wait();
  if (probing) ctl->probe[0] = clock64();
  // ---- 1. everything that depends on nothing but the kernel arguments:
  //         tile sums, tile minima and the RNG state by bulk copies (TMA)
  if (tid == 0) {
    const uint32_t b2 = (uint32_t)nt_pad * 8u, bm = (uint32_t)nt_pad * 4u, br = 624u * 4u;
    mbar_expect_tx(&bar, b2 + bm + br);
    bulk_g2s(p2d, a.l2, b2, &bar);
    bulk_g2s(tmin, a.l2min, bm, &bar);
    bulk_g2s(mt, ctl->mt, br, &bar);
  }
  const double sum_d = ctl->sum;
  const int mti0 = ctl->mti;
  double head_part = 0.0;  // weights in front of head inside head's tile (for the offset O)
  if (warp == kWarps - 1 && a.head > 0) {
    const int last = a.head - 1;
    for (int k = ((last >> 8) << 8) + lane; k <= last; k += 32) head_part += (double)a.w[k];
    if (tl > 8) {  // whole leaves of head's tile in front of head's leaf
      const int l0 = (last >> tl) << (tl - 8), l1 = last >> 8;
      if (l0 + lane < l1) head_part += a.leaf[l0 + lane];
    }
  }
  mbar_wait(&bar, 0);
  float mn = FLT_MAX;
  for (int i = tid; i < nt; i += kSelThreads) mn = fminf(mn, tmin[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  if (lane == 0) red_f[warp] = mn;
  __syncthreads();
  {
    float m = lane < kWarps ? red_f[lane] : FLT_MAX;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
    mn = m;  // every warp now holds the global min non-zero weight
  }
  int ue = -149;
  if (mn != FLT_MAX) {
    int e;
    frexpf(mn, &e);       // mn = f * 2^e, f in [0.5, 1)
    ue = (e - 1) - 23;    // ulp of a normal float
    if (ue < -149) ue = -149;
  }
  const double unit = ldexp(1.0, ue);
  const float sumf = (float)sum_d;

  // ---- 2. this CTA's targets (needs only the RNG state and sum_)
  gen_targets(mt, mti0, a.B, i0, i0 + nT, sumf, a.contract_fma, s_target, &sh_mti);
  if (probing) ctl->probe[1] = clock64();

  // ---- 3. inclusive prefix of the tile sums: contiguous chunk per thread (odd
  //         length: conflict-free 8-byte accesses), block scan of the chunk
  //         totals, second pass adds the chunk offset
  const int per = ((nt + kSelThreads - 1) / kSelThreads) | 1;
  const int c0 = min(tid * per, nt), c1 = min(c0 + per, nt);
  double run = 0.0;
  for (int i = c0; i < c1; ++i) {
    run += p2d[i];
    p2d[i] = run;
  }
  double x = run;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) red_d[warp] = x;
  __syncthreads();
  if (warp == 0) {
    double t = lane < kWarps ? red_d[lane] : 0.0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double y = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= o) t += y;
    }
    if (lane < kWarps) red_d[lane] = t;
  }
  __syncthreads();
  {
    const double off = (warp > 0 ? red_d[warp - 1] : 0.0) + (x - run);
    for (int i = c0; i < c1; ++i) p2d[i] += off;
  }
  __syncthreads();
  const double total = p2d[nt - 1];
  const bool exact = (total < ldexp(1.0, ue + 52)) && !a.force_ordered;  // false for NaN / Inf totals too
  if (probing) ctl->probe[2] = clock64();

  if (exact) {
    // ---- 4. offset of head: physical inclusive prefix at head - 1
    if (warp == kWarps - 1) {
      double part = 0.0;
      if (a.head > 0) {
        part = warp_sum(head_part);
        const int t = (a.head - 1) >> tl;
        if (t > 0) part += p2d[t - 1];
      }
      if (lane == 0) sh_offset = part;
    }
    __syncthreads();
    // ---- 5. tile of each target. Both cases reduce to "first physical slot
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
      int lo = 0, hi = nt;  // first tile whose inclusive prefix reaches the key
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (p2d[mid] >= key) hi = mid; else lo = mid + 1;
      }
      s_tile[tid] = lo < nt ? lo : -1;
      s_rem[tid] = lo < nt ? key - (lo > 0 ? p2d[lo - 1] : 0.0) : 0.0;  // u <= rem <= tile sum
      s_wrapped[tid] = wrapped;
    }
    __syncthreads();
    if (probing) ctl->probe[3] = clock64();
    // ---- 6. slot inside the tile: one warp per target, two targets per warp.
    //   Tiles larger than 256 slots first resolve the LEAF: lanes < 2^(tl-8)
    //   load the tile's leaf sums, a warp scan finds the leaf and the remainder
    //   inside it. Then lane l owns 8 consecutive slots of the leaf (kept in
    //   registers); a warp scan of the lane totals finds the lane, which walks
    //   its own slots.
    constexpr int kPer = kSelTargets / kWarps;
    const int nb = 1 << (tl - 8);
    int leaf_t[kPer];
    double rem_k[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int t = warp + k * kWarps;
      leaf_t[k] = t < nT ? s_tile[t] : -1;
      rem_k[k] = t < nT ? s_rem[t] : 0.0;
    }
    if (nb > 1) {
      double ls[kPer];
#pragma unroll
      for (int k = 0; k < kPer; ++k)  // the loads of both targets are issued together
        ls[k] = (leaf_t[k] >= 0 && lane < nb) ? a.leaf[(size_t)leaf_t[k] * nb + lane] : 0.0;
#pragma unroll
      for (int k = 0; k < kPer; ++k) {
        if (leaf_t[k] < 0) continue;  // warp-uniform
        double P = ls[k];
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
          const double y = __shfl_up_sync(0xffffffffu, P, o);
          if (lane >= o) P += y;
        }
        const unsigned hit = __ballot_sync(0xffffffffu, lane < nb && P >= rem_k[k]);
        if (hit != 0u) {
          const int L = __ffs(hit) - 1;
          rem_k[k] -= __shfl_sync(0xffffffffu, P - ls[k], L);
          leaf_t[k] = leaf_t[k] * nb + L;
        } else {
          leaf_t[k] = -1;
        }
      }
    }
    double T[kPer];
    float4 v0[kPer], v1[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {  // all loads of both targets are issued before the first use
      v0[k] = v1[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (leaf_t[k] >= 0) {
        const float* wp = a.w + ((size_t)leaf_t[k] << 8) + (size_t)lane * 8;
        v0[k] = *reinterpret_cast<const float4*>(wp);
        v1[k] = *reinterpret_cast<const float4*>(wp + 4);
      }
    }
#pragma unroll
    for (int k = 0; k < kPer; ++k)
      T[k] = (((double)v0[k].x + (double)v0[k].y) + ((double)v0[k].z + (double)v0[k].w)) +
             (((double)v1[k].x + (double)v1[k].y) + ((double)v1[k].z + (double)v1[k].w));
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int t = warp + k * kWarps;
      if (t >= nT) continue;  // warp-uniform
      const int leaf = leaf_t[k];
      int slot = -1;
      float wv = 0.0f;
      if (leaf >= 0) {
        const double rem = rem_k[k];
        double P = T[k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const double y = __shfl_up_sync(0xffffffffu, P, o);
          if (lane >= o) P += y;
        }
        const unsigned hit = __ballot_sync(0xffffffffu, P >= rem);
        if (hit != 0u) {
          const int L = __ffs(hit) - 1;
          if (lane == L) {
            const double rem_l = rem - (P - T[k]);
            double acc = 0.0;
            int j = 7;
            const float vv[8] = {v0[k].x, v0[k].y, v0[k].z, v0[k].w, v1[k].x, v1[k].y, v1[k].z, v1[k].w};
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
            slot = (leaf << 8) + L * 8 + j;
          }
          slot = __shfl_sync(0xffffffffu, slot, L);
          wv = __shfl_sync(0xffffffffu, wv, L);
        }
      }
      if (lane == 0) {
        if (slot >= 0 && s_wrapped[t] && slot >= a.head) slot = -1;  // ran past the wrapped part
        int logical = -1;
        if (slot >= 0) logical = slot >= a.head ? slot - a.head : slot + a.S - a.head;
        if (slot < 0 || slot >= a.S || logical >= a.size) {
          ctl->sel_fail = 1;  // reference: assert(false) at lines 297-302
          slot = a.head;
          wv = a.w[slot];
          logical = 0;
        }
        a.ids[i0 + t] = slot;
        a.wsel[i0 + t] = wv;
        // getElementAndMark (line 126). A slot that this call also evicts ends
        // up evicted (the pop comes after the marks): leave it to the eviction.
        if (logical >= a.pop_n) a.evicted[slot] = 0;
      }
    }
  } else if (blockIdx.x == 0) {
    // ordered walk: the reference's loop (lines 274-308) for ALL targets, staged
    // kSelThreads weights at a time and consumed by one thread in order
    float* targets = reinterpret_cast<float*>(sm_raw);
    __shared__ float chunk[kSelThreads];
    __shared__ int sh_next_i, sh_id;
    __shared__ double sh_acc;
    __shared__ float sh_w;
    __syncthreads();
    for (int i = tid; i < 624; i += kSelThreads) mt[i] = ctl->mt[i];
    if (tid == 0) {
      sh_next_i = 0;
      sh_acc = 0.0;
      sh_w = 0.0f;
      sh_id = 0;
    }
    __syncthreads();
    gen_targets(mt, mti0, a.B, 0, a.B, sumf, a.contract_fma, targets, &sh_mti);
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
            a.wsel[i] = w;
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
      ctl->sel_fail = 1;
      for (int i = sh_next_i; i < a.B; ++i) {
        a.ids[i] = a.head;
        a.wsel[i] = a.w[a.head];
      }
    }
    __syncthreads();
    for (int i = tid; i < a.B; i += kSelThreads) {
      const int slot = a.ids[i];
      const int logical = slot >= a.head ? slot - a.head : slot + a.S - a.head;
      if (logical >= a.pop_n) a.evicted[slot] = 0;
    }
  }
  if (probing) ctl->probe[4] = clock64();

  // ---- 7. the last CTA to get here finishes the call
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    sh_last = atomicAdd(&ctl->done_ctas, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!sh_last) return;
  __threadfence();
  if (tid == 0) ctl->probe[6] = clock64();

  // eviction (blockPop, lines 84-103): after the marks like the reference
  if (a.pop_n > 0) {
    if (exact) {
      double part = 0.0;  // exact in any order under the exactness condition
      for (int j = tid; j < a.pop_n; j += kSelThreads) {
        int slot = a.head + j;
        if (slot >= a.S) slot -= a.S;
        part += (double)a.w[slot];
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
        diff -= (double)a.w[slot];
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
    retile_range(a.w, tiles_of(a), a.S, a.head, a.pop_n, warp, kWarps, lane);
  }

  // importance weights (lines 320-322); wsel was written by other SMs: read through L2
  float* wbuf = reinterpret_cast<float*>(sm_raw);  // the tile prefixes are no longer needed
  __syncthreads();
  float mx = -INFINITY;
  for (int i = tid; i < a.B; i += kSelThreads) {
    float v = __fdiv_rn(__ldcg(a.wsel + i), sumf);
    v = __fmul_rn((float)a.size_is, v);
    v = (float)pow((double)v, (double)(-a.beta));
    wbuf[i] = v;
    mx = fmaxf(mx, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) red_f[warp] = mx;
  __syncthreads();
  if (warp == 0) {
    float m = lane < kWarps ? red_f[lane] : -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) sh_max = m;
  }
  __syncthreads();
  for (int i = tid; i < a.B; i += kSelThreads) a.is_weights[i] = __fdiv_rn(wbuf[i], sh_max);

  // RNG state after the B draws, counters
  for (int i = tid; i < 624; i += kSelThreads) ctl->mt[i] = mt[i];
  if (tid == 0) {
    ctl->mti = sh_mti;
    if (exact) ctl->n_exact += 1; else ctl->n_ordered += 1;
    ctl->n_popped += (unsigned long long)a.pop_n;
    if (*reinterpret_cast<volatile int*>(&ctl->sel_fail)) ctl->error = RR_ERR_OVERRUN;
    ctl->sel_fail = 0;
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
  __shared__ float buf[1024];
  __shared__ float sh_sum;
  const int tid = threadIdx.x;
  if (tid == 0) sh_sum = 0.0f;
  for (int base = 0; base < n; base += 1024) {
    const int i = base + tid;
    if (i < n) {
      float p = prio[i];
      float v = is_weight ? p : pow_alpha(p, alpha);
      int slot = start + i;
      if (slot >= S) slot -= S;
      w[slot] = v;
      buf[tid] = v;
    }
    __syncthreads();
    if (tid == 0) {
      float s = sh_sum;
      const int m = min(1024, n - base);
      for (int k = 0; k < m; ++k) s = __fadd_rn(s, buf[k]);
      sh_sum = s;
    }
    __syncthreads();
  }
  if (tid == 0) ctl->sum += (double)sh_sum;
  // the block's weights are visible to the CTA (barrier above): refresh its tiles
  retile_range(w, tiles, S, start, n, tid >> 5, 32, tid & 31);
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
// CTA 0 does that bookkeeping. CTAs 1.. keep the tile entries current, one
// warp per batch element: the warp of the first element of a run that falls
// into one tile reloads the tile, substitutes the final value of every batch
// slot in it (computed from the same inputs as CTA 0, so the result does not
// depend on whether CTA 0's store has landed) and reduces it again.
//
// The new priorities come from device memory or, for a learner-sized batch
// handed over in host memory, ride in the launch packet itself (PrioInline, a
// __grid_constant__ parameter read through the constant bank): no staging
// copy sits between updatePriority and the next sample on the stream.
struct PrioInline {
  float v[kInlinePrio];
};
__device__ __forceinline__ float prio_at(const float* p, int i) { return p[i]; }
__device__ __forceinline__ float prio_at(const PrioInline& p, int i) { return p.v[i]; }

template <typename Prio>
__device__ __forceinline__ void update_body(Ctl* ctl, float* w, const uint8_t* __restrict__ evicted,
                                            const int32_t* __restrict__ ids, int n, const Prio& prio, float alpha,
                                            int is_weight, const Tiles& tiles) {
  const int tile_log2 = tiles.tile_log2;
  extern __shared__ __align__(128) uint8_t sm_raw[];
  int32_t* s_id = reinterpret_cast<int32_t*>(sm_raw);  // -1 for evicted
  float* s_new = reinterpret_cast<float*>(s_id + n);
  float* s_diff = s_new + n;
  __shared__ double r_sum[32], r_abs[32];
  __shared__ float r_min[32];
  __shared__ int sh_exact;
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  if (blockIdx.x > 0) {
    __shared__ __align__(16) float sub[32][256];
    for (int i = tid; i < n; i += blockDim.x) {
      const int id = ids[i];
      const float p = prio_at(prio, i);
      const bool ev = evicted[id] != 0;
      s_new[i] = is_weight ? p : pow_alpha(p, alpha);
      s_id[i] = ev ? -1 : id;
    }
    __syncthreads();
    const int i = (blockIdx.x - 1) * 32 + warp;
    if (i >= n) return;
    const int id = s_id[i];
    if (id < 0) return;  // an evicted slot keeps weight 0: its tile does not change
    const int t = id >> tile_log2;
    if (i > 0 && s_id[i - 1] >= 0 && (s_id[i - 1] >> tile_log2) == t) return;  // covered by that element's warp
    const int nb = 1 << (tile_log2 - 8);
    float* mine = sub[warp];
    const int tbase = t << tile_log2;
    // which leaves of the tile hold a batch slot (final values only: last occurrence, not evicted)
    unsigned touched = 0u;
    for (int j = lane; j < n; j += 32) {
      const int sid = s_id[j];
      if (sid >= tbase && sid < tbase + (1 << tile_log2)) touched |= 1u << ((sid - tbase) >> 8);
    }
    touched = __reduce_or_sync(0xffffffffu, touched);
    // untouched leaves keep their entries: nobody writes them during this kernel
    double keep_s = 0.0;
    float keep_m = FLT_MAX;
    if (nb > 1 && lane < nb && !((touched >> lane) & 1u)) {
      keep_s = tiles.leaf[(size_t)t * nb + lane];
      keep_m = tiles.leafmin[(size_t)t * nb + lane];
    }
    double tot = 0.0;
    float m = FLT_MAX;
    for (int k = 0; k < nb; ++k) {
      double s;
      float mn;
      if ((touched >> k) & 1u) {
        const int base = tbase + k * 256;
        const float4* src = reinterpret_cast<const float4*>(w + base) + lane * 2;
        reinterpret_cast<float4*>(mine)[lane * 2] = src[0];
        reinterpret_cast<float4*>(mine)[lane * 2 + 1] = src[1];
        __syncwarp();
        for (int j = lane; j < n; j += 32) {
          const int sid = s_id[j];
          const bool last = !(j + 1 < n && s_id[j + 1] == sid);
          if (sid >= base && sid < base + 256 && last) mine[sid - base] = s_new[j];
        }
        __syncwarp();
        const float4 a = reinterpret_cast<const float4*>(mine)[lane * 2];
        const float4 b = reinterpret_cast<const float4*>(mine)[lane * 2 + 1];
        __syncwarp();
        reduce256(a, b, s, mn);
        if (lane == 0 && nb > 1) {
          tiles.leaf[(size_t)t * nb + k] = s;
          tiles.leafmin[(size_t)t * nb + k] = mn;
        }
      } else {
        s = __shfl_sync(0xffffffffu, keep_s, k);
        mn = __shfl_sync(0xffffffffu, keep_m, k);
      }
      tot += s;
      m = fminf(m, mn);
    }
    if (lane == 0) {
      tiles.l2[t] = tot;
      tiles.l2min[t] = m;
    }
    return;
  }
  float* s_old = s_diff;  // old weight, replaced by the difference below
  for (int i = tid; i < n; i += blockDim.x) {
    const int id = ids[i];
    const float p = prio_at(prio, i);
    const bool ev = evicted[id] != 0;  // both dependent loads are issued together
    const float old = w[id];
    s_new[i] = is_weight ? p : pow_alpha(p, alpha);
    s_id[i] = ev ? -1 : id;
    s_old[i] = old;
  }
  __syncthreads();
  double sum = 0.0, asum = 0.0;
  float mn = FLT_MAX;
  for (int i = tid; i < n; i += blockDim.x) {
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
  for (int i = tid; i < n; i += blockDim.x) {
    const int id = s_id[i];
    const bool last = !(i + 1 < n && s_id[i + 1] == id);
    if (id >= 0 && last) w[id] = s_new[i];
  }
  sum = warp_sum(sum);
  asum = warp_sum(asum);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  if (lane == 0) {
    r_sum[warp] = sum;
    r_abs[warp] = asum;
    r_min[warp] = mn;
  }
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x >> 5;
    double a = lane < nw ? r_sum[lane] : 0.0;
    double b = lane < nw ? r_abs[lane] : 0.0;
    float m = lane < nw ? r_min[lane] : FLT_MAX;
    a = warp_sum(a);
    b = warp_sum(b);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
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
}

// wait_late: the kernel in front of this one on the stream is the gather of
// the batch whose priorities are being updated (or of a batch drawn ahead).
// Nothing here reads what a gather writes, and the gather triggers its
// dependents only after its own wait for the selection, so the update runs
// BESIDE the gather and waits for it at its end instead -- its successor, which
// waits for this grid, then still sees the gather complete. The host sets it
// only when it knows the predecessor (rr_update_priority).
template <typename Prio>
__global__ void __launch_bounds__(1024) update_kernel(Ctl* ctl, float* w, const uint8_t* __restrict__ evicted,
                                                      const int32_t* __restrict__ ids, int n,
                                                      const __grid_constant__ Prio prio, float alpha,
                                                      int is_weight, const Tiles tiles, int wait_late) {
  pdl_launch_dependents();
  if (!wait_late) pdl_wait();
  update_body<Prio>(ctl, w, evicted, ids, n, prio, alpha, is_weight, tiles);
  if (wait_late) pdl_wait();
}

// ----------------------------------------------------------------- launches
cudaError_t launch_build_hierarchy(const SelectArgs& a, cudaStream_t stream) {
  return launch_pdl(build_hierarchy_kernel, dim3(a.S_pad / kBuildSpan), dim3(256), 0, stream, a.w, tiles_of(a));
}

cudaError_t launch_select(const SelectArgs& a, cudaStream_t stream) {
  const size_t nt_pad = (((size_t)a.S_pad >> a.tile_log2) + 3) & ~(size_t)3;
  const size_t smem = std::max(nt_pad * (sizeof(double) + sizeof(float)), (size_t)a.B * sizeof(float)) + 16;
  cudaError_t e = ensure_dynamic_smem(select_kernel, smem);
  if (e != cudaSuccess) return e;
  const int grid = (a.B + kSelTargets - 1) / kSelTargets;
  return launch_pdl(select_kernel, dim3(grid), dim3(kSelThreads), smem, stream, a);
}

cudaError_t launch_commit(Ctl* ctl, float* w, const Tiles& tiles, int S, int start, int n, const float* prio,
                          float alpha, int is_weight, cudaStream_t stream) {
  commit_kernel<<<1, 1024, 0, stream>>>(ctl, w, tiles, S, start, n, prio, alpha, is_weight);
  return cudaGetLastError();
}

cudaError_t launch_update(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids, int n,
                          const float* prio, float alpha, int is_weight, int wait_late, cudaStream_t stream) {
  const size_t smem = (size_t)n * 12;
  cudaError_t e = ensure_dynamic_smem(update_kernel<const float*>, smem);
  if (e != cudaSuccess) return e;
  const int grid = 1 + (n + 31) / 32;  // CTA 0: sum_ and weights; the others: one warp per element's tile
  return launch_pdl(update_kernel<const float*>, dim3(grid), dim3(1024), smem, stream, ctl, w, evicted, ids, n, prio,
                    alpha, is_weight, tiles, wait_late);
}

cudaError_t launch_update_inline(Ctl* ctl, float* w, const Tiles& tiles, const uint8_t* evicted, const int32_t* ids,
                                 int n, const float* prio_host, float alpha, int is_weight, int wait_late,
                                 cudaStream_t stream) {
  if (n > kInlinePrio) return cudaErrorInvalidValue;
  PrioInline pin;
  memcpy(pin.v, prio_host, sizeof(float) * (size_t)n);
  const size_t smem = (size_t)n * 12;
  cudaError_t e = ensure_dynamic_smem(update_kernel<PrioInline>, smem);
  if (e != cudaSuccess) return e;
  const int grid = 1 + (n + 31) / 32;
  return launch_pdl(update_kernel<PrioInline>, dim3(grid), dim3(1024), smem, stream, ctl, w, evicted, ids, n, pin,
                    alpha, is_weight, tiles, wait_late);
}

}  // namespace rr
