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
// is < 2^53 * u. Then every value this file computes (group sums, tile sums,
// prefixes, differences of prefixes) is an integer multiple of u below
// 2^53 * u, hence exact, and the three-level search below is bit-identical to
// the sequential walk. When the condition fails (weights spanning > 29 binades
// relative to the total) the kernel takes the ordered walk instead, which is
// the reference's loop verbatim, and counts it in Ctl::n_ordered.
//
// Dead slots hold weight 0 (zeroed on eviction; the reference never reads a
// popped slot's weight again: update() skips evicted ids and append overwrites
// the slot), so physical prefixes need no live-range mask.
#include <algorithm>
#include <cfloat>
#include <cmath>

#include "../../include/rela_b200.h"
#include "kernels.h"
#include "ptx.cuh"

namespace rr {

namespace {

constexpr int kSelThreads = 1024;
constexpr int kSelChunk = 512;  // targets staged through shared memory per round

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
// One CTA of 256 threads per 1024-slot tile: each thread owns 4 consecutive
// slots (one 16-byte load), 8 threads form a 32-slot group.
__global__ void __launch_bounds__(256) build_hierarchy_kernel(const float* __restrict__ w,
                                                              double* __restrict__ l1,
                                                              double* __restrict__ l2,
                                                              float* __restrict__ l2min) {
  const int tile = blockIdx.x;
  const int tid = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const float4 v = reinterpret_cast<const float4*>(w)[(size_t)tile * (kTile / 4) + tid];
  double s = ((double)v.x + (double)v.y) + ((double)v.z + (double)v.w);
  float mn = FLT_MAX;
  if (v.x != 0.0f) mn = fminf(mn, fabsf(v.x));
  if (v.y != 0.0f) mn = fminf(mn, fabsf(v.y));
  if (v.z != 0.0f) mn = fminf(mn, fabsf(v.z));
  if (v.w != 0.0f) mn = fminf(mn, fabsf(v.w));
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((tid & 7) == 0) l1[(size_t)tile * (kTile / kGroup) + (tid >> 3)] = s;
  // tile reduction: 4 groups per warp, 8 warps
  __shared__ double ws[8];
  __shared__ float wm[8];
  double t = s;
  if ((tid & 7) != 0) t = 0.0;
  t = warp_sum(t);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  if ((tid & 31) == 0) {
    ws[tid >> 5] = t;
    wm[tid >> 5] = mn;
  }
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
    float m = FLT_MAX;
    for (int k = 0; k < 8; ++k) {
      tot += ws[k];
      m = fminf(m, wm[k]);
    }
    l2[tile] = tot;
    l2min[tile] = m;
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
  const int tid = threadIdx.x;
  uint32_t v = 0;
  if (tid < 227) v = mt_mix(mt[tid], mt[tid + 1], mt[tid + 397]);
  __syncthreads();
  if (tid < 227) mt[tid] = v;
  __syncthreads();
  if (tid >= 227 && tid < 454) v = mt_mix(mt[tid], mt[tid + 1], mt[tid - 227]);
  __syncthreads();
  if (tid >= 227 && tid < 454) mt[tid] = v;
  __syncthreads();
  if (tid >= 454 && tid < 623) v = mt_mix(mt[tid], mt[tid + 1], mt[tid - 227]);
  __syncthreads();
  if (tid >= 454 && tid < 623) mt[tid] = v;
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

// Exact integer view of the prefix hierarchy ---------------------------------
// Under the exactness condition every weight, group sum, tile sum and prefix
// is an integer multiple of u = 2^ue below 2^52 * u. The search works on those
// integers ("q" values): no fp64 dependency chains, and comparisons against a
// target t become P >= ceil(t / u) (and "accSum > 0" becomes P >= 1).
__device__ __forceinline__ unsigned long long f2q(float w, int ue) {
  const uint32_t bits = __float_as_uint(w) & 0x7fffffffu;
  if (bits == 0u) return 0ull;
  const uint32_t e = bits >> 23;
  const unsigned long long m = e ? ((bits & 0x7fffffu) | 0x800000u) : (bits & 0x7fffffu);
  const int shift = (int)(e ? e : 1u) - 150 - ue;  // >= 0: ulp(w) >= u
  return shift < 40 ? (m << shift) : ~0ull >> 2;   // >= 2^62 can never be below a key
}

// x / 2^ue for a double that is a multiple of 2^ue; *inexact is raised when it
// is not, or when the quotient does not fit 2^62.
__device__ __forceinline__ unsigned long long d2q(double x, int ue, int* inexact) {
  const unsigned long long bits = (unsigned long long)__double_as_longlong(x) & 0x7fffffffffffffffull;
  if (bits == 0ull) return 0ull;
  const int e = (int)(bits >> 52);
  const unsigned long long m = (bits & 0xfffffffffffffull) | (e ? 0x10000000000000ull : 0ull);
  const int shift = (e ? e : 1) - 1075 - ue;
  if (shift >= 0) {
    if (shift > 9) {
      *inexact = 1;
      return ~0ull >> 2;
    }
    return m << shift;
  }
  if (shift <= -64) {
    *inexact = 1;
    return 0ull;
  }
  if (m & ((1ull << (-shift)) - 1ull)) *inexact = 1;
  return m >> (-shift);
}

}  // namespace

// -------------------------------------------------------------- K4/K5 select
// Single CTA. Dynamic shared memory: p2[nt2] doubles | targets[B] | wsel[B].
__global__ void __launch_bounds__(kSelThreads) select_kernel(const SelectArgs a) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  const int nt2 = a.S_pad / kTile;
  double* p2d = reinterpret_cast<double*>(sm_raw);
  unsigned long long* p2q = reinterpret_cast<unsigned long long*>(sm_raw);
  float* targets = reinterpret_cast<float*>(p2d + nt2);
  float* wsel = targets + a.B;
  int* s_tile = reinterpret_cast<int*>(wsel + a.B);
  double* stage_d = reinterpret_cast<double*>(s_tile + kSelChunk + ((2 * a.B) & 1));  // 8-byte aligned
  float* stage_f = reinterpret_cast<float*>(stage_d);
  __shared__ uint32_t mt[624];
  __shared__ unsigned long long red_q[32];
  __shared__ float red_f[32];
  __shared__ unsigned long long sh_total, sh_offset;
  __shared__ float sh_max;
  __shared__ int sh_exact, sh_fail;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  Ctl* ctl = a.ctl;

  pdl_launch_dependents();
  pdl_wait();
  if (tid == 0) ctl->probe[0] = clock64();
  // ---- 1. tile sums -> shared memory, min non-zero weight -> u = 2^ue,
  //         integer inclusive prefix of the tile sums (one pass, contiguous
  //         chunk per thread)
  const int per = (nt2 + kSelThreads - 1) / kSelThreads;
  const int c0 = min(tid * per, nt2), c1 = min(c0 + per, nt2);
  float mn = FLT_MAX;
  for (int i = c0; i < c1; ++i) {
    p2d[i] = a.l2[i];
    mn = fminf(mn, a.l2min[i]);
  }
  // values needed after the barriers: issue their loads now
  double head_l1 = 0.0;  // group sums / weights in front of head, for the offset O
  float head_w = 0.0f;
  if (warp == 0 && a.head > 0) {
    const int last = a.head - 1;
    const int t = last / kTile, g = (last % kTile) / kGroup, o = last % kGroup;
    if (lane < g) head_l1 = a.l1[(size_t)t * (kTile / kGroup) + lane];
    if (lane <= o) head_w = a.w[(size_t)t * kTile + (size_t)g * kGroup + lane];
  }
  const float sum_f_early = (float)ctl->sum;
  const int mti_early = ctl->mti;
  for (int i = tid; i < 624; i += kSelThreads) mt[i] = ctl->mt[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  if (lane == 0) red_f[warp] = mn;
  __syncthreads();
  {
    float m = red_f[lane];
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
  int inexact = 0;
  unsigned long long run = 0ull;
  for (int i = c0; i < c1; ++i) {
    run += d2q(p2d[i], ue, &inexact);
    p2q[i] = run;  // same 8 bytes as p2d[i]: each thread converts its own chunk in place
  }
  // block exclusive scan of the per-thread totals
  unsigned long long x = run;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) red_q[warp] = x;
  inexact = __syncthreads_or(inexact);
  if (warp == 0) {
    unsigned long long t = red_q[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned long long y = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= o) t += y;
    }
    red_q[lane] = t;
  }
  __syncthreads();
  {
    const unsigned long long off = (warp > 0 ? red_q[warp - 1] : 0ull) + (x - run);
    for (int i = c0; i < c1; ++i) p2q[i] += off;
  }
  __syncthreads();

  if (tid == 0) ctl->probe[1] = clock64();
  // ---- 2. exactness test, offset of head (physical inclusive prefix at head - 1)
  if (warp == 0) {
    const unsigned long long total = p2q[nt2 - 1];
    int exact = !inexact && total < (1ull << 52);
    if (a.force_ordered) exact = 0;
    unsigned long long part = 0ull;
    if (exact && a.head > 0) {
      const int last = a.head - 1;
      const int t = last / kTile;
      int dummy = 0;
      part = d2q(head_l1, ue, &dummy) + f2q(head_w, ue);
#pragma unroll
      for (int k = 16; k > 0; k >>= 1) part += __shfl_xor_sync(0xffffffffu, part, k);
      if (t > 0) part += p2q[t - 1];
    }
    if (lane == 0) {
      sh_total = total;
      sh_offset = part;
      sh_exact = exact;
      sh_fail = 0;
    }
  }
  __syncthreads();

  if (tid == 0) ctl->probe[2] = clock64();
  // ---- 3. targets: rand_i = U[0, segment) + i * segment, clamped (lines 265-280)
  const float sumf = sum_f_early;
  const float segment = __fdiv_rn(sumf, (float)a.B);
  {
    int pos = mti_early;
    int done = 0;
    while (done < a.B) {
      if (pos >= 624) {
        mt_twist(mt);
        pos = 0;
      }
      int m = min(a.B - done, 624 - pos);
      if (tid < m) {
        uint32_t y = mt_temper(mt[pos + tid]);
        float u = __fmul_rn(__uint2float_rn(y), 2.3283064365386963e-10f);
        if (u >= 1.0f) u = 0.99999994f;  // nextafterf(1, 0)
        const int i = done + tid;
        float r;
        if (a.contract_fma) {
          r = __fmaf_rn((float)i, segment, __fmaf_rn(u, segment, 0.0f));
        } else {
          r = __fadd_rn(__fadd_rn(__fmul_rn(u, segment), 0.0f), __fmul_rn((float)i, segment));
        }
        const float cap = __fsub_rn(sumf, 0.2f);
        targets[i] = (cap < r) ? cap : r;
      }
      pos += m;
      done += m;
      __syncthreads();
    }
    for (int i = tid; i < 624; i += kSelThreads) ctl->mt[i] = mt[i];
    if (tid == 0) ctl->mti = pos;
  }
  __syncthreads();

  if (tid == 0) ctl->probe[3] = clock64();
  // ---- 4. selection
  const int size = a.size;
  if (sh_exact) {
    // Both cases reduce to "first physical slot whose prefix reaches a key":
    //   unwrapped part [head, S):  P(s) - O >= t'   <=>  P(s) >= t' + O
    //   wrapped part   [0, head):  M1 + P(s) >= t'  <=>  P(s) >= t' - M1
    // with t' = max(ceil(target / u), 1) (the reference's accSum > 0 && accSum >= target),
    // O = prefix at head - 1 and M1 = total - O the mass of the unwrapped part.
    // The per-target group sums and weights are fetched by whole warps (one
    // coalesced 256 B / 128 B request per target) into padded shared-memory rows
    // and then walked by the target's own thread: scattered per-thread loads
    // would cost one L1 wavefront per lane per load on this single SM.
    const unsigned long long total = sh_total, off = sh_offset;
    const unsigned long long m1 = total - off;
    const double qscale = ldexp(1.0, -ue);  // 1 / u
    for (int i0 = 0; i0 < a.B; i0 += kSelChunk) {
      const int nT = min(kSelChunk, a.B - i0);
      int tile = -1, wrapped = 0;
      unsigned long long rem = 0ull;
      if (tid < nT) {
        const float tf = targets[i0 + tid];
        unsigned long long tq = 1ull;
        if (tf > 0.0f) {
          const double scaled = ldexp((double)tf, -ue);  // exact: power-of-two scaling
          tq = scaled >= 4.0e18 ? (1ull << 62) : (unsigned long long)__double2ll_ru(scaled);
          if (tq == 0ull) tq = 1ull;
        }
        wrapped = !(m1 >= tq);
        const unsigned long long key = wrapped ? tq - m1 : tq + off;
        int lo = 0, hi = nt2;  // first tile whose inclusive prefix reaches the key
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (p2q[mid] >= key) hi = mid; else lo = mid + 1;
        }
        if (lo < nt2) {
          tile = lo;
          rem = key - (tile > 0 ? p2q[tile - 1] : 0ull);  // 1 <= rem <= tile sum
        }
        s_tile[tid] = tile;
      }
      if (tid == 0) ctl->probe[8] = clock64();
      __syncthreads();
      if (tid == 0) ctl->probe[9] = clock64();
      {
        // all loads of the warp's targets are issued before the first store
        constexpr int kPerWarp = kSelChunk / (kSelThreads / 32);
        double v[kPerWarp];
#pragma unroll
        for (int k = 0; k < kPerWarp; ++k) {
          const int t = warp + k * (kSelThreads / 32);
          const int tl = t < nT ? s_tile[t] : -1;
          v[k] = tl >= 0 ? a.l1[(size_t)tl * (kTile / kGroup) + lane] : 0.0;
        }
#pragma unroll
        for (int k = 0; k < kPerWarp; ++k) {
          const int t = warp + k * (kSelThreads / 32);
          if (t < nT) stage_d[t * 33 + lane] = v[k];
        }
      }
      __syncthreads();
      if (tid == 0) ctl->probe[10] = clock64();
      int base = -1;
      // walks in fp64 on q-scaled values: every term is an integer below 2^52,
      // so the fused multiply-adds are exact and order-independent
      double remd = (double)rem;
      if (tid < nT && tile >= 0) {
        double acc = 0.0, before = 0.0;
        int grp = 0;
#pragma unroll 8
        for (int k = 0; k < kTile / kGroup; ++k) {
          acc = fma(stage_d[tid * 33 + k], qscale, acc);
          if (acc < remd) {
            ++grp;
            before = acc;
          }
        }
        if (grp >= kTile / kGroup) grp = kTile / kGroup - 1;
        remd -= before;
        base = tile * kTile + grp * kGroup;
      }
      if (tid < nT) s_tile[tid] = base;
      if (tid == 0) ctl->probe[11] = clock64();
      __syncthreads();
      {
        constexpr int kPerWarp = kSelChunk / (kSelThreads / 32);
        float v[kPerWarp];
#pragma unroll
        for (int k = 0; k < kPerWarp; ++k) {
          const int t = warp + k * (kSelThreads / 32);
          const int bs = t < nT ? s_tile[t] : -1;
          v[k] = bs >= 0 ? a.w[(size_t)bs + lane] : 0.0f;
        }
#pragma unroll
        for (int k = 0; k < kPerWarp; ++k) {
          const int t = warp + k * (kSelThreads / 32);
          if (t < nT) stage_f[t * 33 + lane] = v[k];
        }
      }
      __syncthreads();
      if (tid == 0) ctl->probe[12] = clock64();
      if (tid < nT) {
        int slot = -1;
        float wv = 0.0f;
        if (base >= 0) {
          double acc = 0.0;
          int j = 0;
#pragma unroll 8
          for (int k = 0; k < kGroup; ++k) {
            acc = fma((double)stage_f[tid * 33 + k], qscale, acc);
            j += (acc < remd);
          }
          if (j >= kGroup) j = kGroup - 1;
          slot = base + j;
          wv = stage_f[tid * 33 + j];
          if (wrapped && slot >= a.head) slot = -1;  // ran past the wrapped part
        }
        int logical = -1;
        if (slot >= 0) logical = slot >= a.head ? slot - a.head : slot + a.S - a.head;
        if (slot < 0 || slot >= a.S || logical >= size) {
          sh_fail = 1;  // reference: assert(false) at lines 297-302
          slot = a.head;
          wv = a.w[slot];
        }
        a.ids[i0 + tid] = slot;
        wsel[i0 + tid] = wv;
      }
      if (tid == 0) ctl->probe[13] = clock64();
      __syncthreads();
    }
  } else {
    // ordered walk: the reference's loop (lines 274-308), staged 1024 weights
    // at a time through shared memory and consumed by one thread in order
    __shared__ float chunk[kSelThreads];
    __shared__ int sh_next_i;
    __shared__ double sh_acc;
    __shared__ float sh_w;
    __shared__ int sh_id;
    if (tid == 0) {
      sh_next_i = 0;
      sh_acc = 0.0;
      sh_w = 0.0f;
      sh_id = 0;
    }
    __syncthreads();
    for (int base = 0; base <= size; base += kSelThreads) {
      int idx = base + tid;
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
            wsel[i] = w;
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
      sh_fail = 1;
      for (int i = sh_next_i; i < a.B; ++i) {
        a.ids[i] = a.head;
        wsel[i] = a.w[a.head];
      }
    }
  }
  __syncthreads();

  if (tid == 0) ctl->probe[4] = clock64();
  // ---- 5. mark sampled slots live (getElementAndMark, line 126)
  for (int i = tid; i < a.B; i += kSelThreads) a.evicted[a.ids[i]] = 0;
  __syncthreads();

  if (tid == 0) ctl->probe[5] = clock64();
  // ---- 6. eviction (blockPop, lines 84-103), after the marks like the reference
  if (a.pop_n > 0) {
    if (sh_exact) {
      unsigned long long part = 0ull;
      for (int j = tid; j < a.pop_n; j += kSelThreads) {
        int slot = a.head + j;
        if (slot >= a.S) slot -= a.S;
        part += f2q(a.w[slot], ue);
      }
#pragma unroll
      for (int k = 16; k > 0; k >>= 1) part += __shfl_xor_sync(0xffffffffu, part, k);
      if (lane == 0) red_q[warp] = part;
      __syncthreads();
      if (tid == 0) {
        unsigned long long tot = 0ull;
        for (int k = 0; k < kSelThreads / 32; ++k) tot += red_q[k];
        // diff = -(w_0 + w_1 + ...) exactly (< 2^52 u), then sum_ += diff (line 96)
        ctl->sum += -ldexp((double)tot, ue);
      }
    } else if (tid == 0) {
      double diff = 0.0;
      for (int j = 0; j < a.pop_n; ++j) {
        int slot = a.head + j;
        if (slot >= a.S) slot -= a.S;
        diff -= (double)a.w[slot];
      }
      ctl->sum += diff;
    }
    __syncthreads();
    for (int j = tid; j < a.pop_n; j += kSelThreads) {
      int slot = a.head + j;
      if (slot >= a.S) slot -= a.S;
      a.evicted[slot] = 1;
      a.w[slot] = 0.0f;
    }
  }

  if (tid == 0) ctl->probe[6] = clock64();
  // ---- 7. importance weights (lines 320-322)
  float mx = -INFINITY;
  for (int i = tid; i < a.B; i += kSelThreads) {
    float x = __fdiv_rn(wsel[i], sumf);
    x = __fmul_rn((float)a.size_is, x);
    x = (float)pow((double)x, (double)(-a.beta));
    wsel[i] = x;
    mx = fmaxf(mx, x);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  __syncthreads();
  if (lane == 0) red_f[warp] = mx;
  __syncthreads();
  if (warp == 0) {
    float m = red_f[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) sh_max = m;
  }
  __syncthreads();
  for (int i = tid; i < a.B; i += kSelThreads) a.is_weights[i] = __fdiv_rn(wsel[i], sh_max);

  if (tid == 0) {
    ctl->probe[7] = clock64();
    if (sh_exact) ctl->n_exact += 1; else ctl->n_ordered += 1;
    ctl->n_popped += (unsigned long long)a.pop_n;
    if (sh_fail) ctl->error = RR_ERR_OVERRUN;
  }
}

// ------------------------------------------------------------- add commit
// weights_[j] = priority^alpha for the block's slots, float block sum in index
// order, sum_ += (double) blockSum (rela/prioritized_replay.h:58-73,188).
__global__ void __launch_bounds__(1024) commit_kernel(Ctl* ctl, float* w, int S, int start, int n,
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
}

// ------------------------------------------------------ K7 priority update
// ConcurrentQueue::update (rela/prioritized_replay.h:105-119): in index order,
// skip evicted ids, diff += (float)(new - old), weights_[id] = new; a repeated
// id sees the value written by its previous occurrence. Sampled ids are
// non-decreasing in ring order (the reference's walk never rewinds), so
// repeats are always adjacent: occurrence i reads new[i-1], the last one writes.
// The double accumulation of the float differences is done in parallel when
// it is provably exact (sum |d| < 2^52 * ulp(min |d|)), else in order.
__global__ void __launch_bounds__(1024) update_kernel(Ctl* ctl, float* w, const uint8_t* __restrict__ evicted,
                                                      const int32_t* __restrict__ ids, int n,
                                                      const float* __restrict__ prio, float alpha,
                                                      int is_weight) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  int32_t* s_id = reinterpret_cast<int32_t*>(sm_raw);  // -1 for evicted
  float* s_new = reinterpret_cast<float*>(s_id + n);
  float* s_diff = s_new + n;
  __shared__ double r_sum[32], r_abs[32];
  __shared__ float r_min[32];
  __shared__ int sh_exact;
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  pdl_launch_dependents();
  pdl_wait();
  for (int i = tid; i < n; i += blockDim.x) {
    const int id = ids[i];
    const float p = prio[i];
    const bool ev = evicted[id] != 0;
    s_new[i] = is_weight ? p : pow_alpha(p, alpha);
    s_id[i] = ev ? -1 : id;
  }
  __syncthreads();
  double sum = 0.0, asum = 0.0;
  float mn = FLT_MAX;
  for (int i = tid; i < n; i += blockDim.x) {
    const int id = s_id[i];
    float d = 0.0f;
    if (id >= 0) {
      const bool has_prev = i > 0 && s_id[i - 1] == id;
      const float old = has_prev ? s_new[i - 1] : w[id];
      d = __fsub_rn(s_new[i], old);
      sum += (double)d;
      asum += fabs((double)d);
      if (d != 0.0f) mn = fminf(mn, fabsf(d));
    }
    s_diff[i] = d;
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

// ----------------------------------------------------------------- launches
cudaError_t launch_build_hierarchy(const SelectArgs& a, cudaStream_t stream) {
  return launch_pdl(build_hierarchy_kernel, dim3(a.S_pad / kTile), dim3(256), 0, stream, a.w, a.l1, a.l2, a.l2min);
}

cudaError_t launch_select(const SelectArgs& a, cudaStream_t stream) {
  const size_t smem = (size_t)(a.S_pad / kTile) * sizeof(double) + (size_t)a.B * 2 * sizeof(float) +
                      (size_t)(kSelChunk + 2) * sizeof(int) + (size_t)std::min(a.B, kSelChunk) * 33 * sizeof(double);
  static size_t configured = 0;
  if (smem > 40 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  return launch_pdl(select_kernel, dim3(1), dim3(kSelThreads), smem, stream, a);
}

cudaError_t launch_commit(Ctl* ctl, float* w, int S, int start, int n, const float* prio, float alpha,
                          int is_weight, cudaStream_t stream) {
  commit_kernel<<<1, 1024, 0, stream>>>(ctl, w, S, start, n, prio, alpha, is_weight);
  return cudaGetLastError();
}

cudaError_t launch_update(Ctl* ctl, float* w, const uint8_t* evicted, const int32_t* ids, int n,
                          const float* prio, float alpha, int is_weight, cudaStream_t stream) {
  const size_t smem = (size_t)n * 12;
  static size_t configured = 0;
  if (smem > 40 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  return launch_pdl(update_kernel, dim3(1), dim3(1024), smem, stream, ctl, w, evicted, ids, n, prio, alpha, is_weight);
}

}  // namespace rr
