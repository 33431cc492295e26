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
#include <cfloat>
#include <cmath>

#include "../../include/rela_b200.h"
#include "kernels.h"

namespace rr {

namespace {

constexpr int kSelThreads = 1024;

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

// Physical inclusive prefix helpers -------------------------------------
struct Hier {
  const float* w;
  const double* l1;
  const double* p2;  // inclusive prefix of tile sums (shared memory)
  int nt2;
};

// predicate of the reference walk on a logical prefix value
__device__ __forceinline__ bool hit(double v, double target) {
  return v > 0.0 && v >= target;
}

// smallest physical slot s in [0, S_pad) with hit(P(s) + delta, target), or -1.
__device__ int search_slot(const Hier& h, double delta, double target) {
  // level 2: binary search over tile prefixes
  int lo = 0, hi = h.nt2;  // first tile with hit(p2[tile] + delta)
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (hit(h.p2[mid] + delta, target)) hi = mid; else lo = mid + 1;
  }
  if (lo >= h.nt2) return -1;
  const int tile = lo;
  double acc = (tile > 0 ? h.p2[tile - 1] : 0.0);
  // level 1: groups of the tile, in order
  const double* g = h.l1 + (size_t)tile * (kTile / kGroup);
  int grp = 0;
  for (; grp < kTile / kGroup - 1; ++grp) {
    double nxt = acc + g[grp];
    if (hit(nxt + delta, target)) break;
    acc = nxt;
  }
  // level 0: slots of the group, in order
  const float* ws = h.w + (size_t)tile * kTile + (size_t)grp * kGroup;
  int j = 0;
  for (; j < kGroup - 1; ++j) {
    double nxt = acc + (double)ws[j];
    if (hit(nxt + delta, target)) break;
    acc = nxt;
  }
  return tile * kTile + grp * kGroup + j;
}

}  // namespace

// -------------------------------------------------------------- K4/K5 select
// Single CTA. Dynamic shared memory: p2[nt2] doubles | targets[B] | wsel[B].
__global__ void __launch_bounds__(kSelThreads) select_kernel(const SelectArgs a) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  const int nt2 = a.S_pad / kTile;
  double* p2 = reinterpret_cast<double*>(sm_raw);
  float* targets = reinterpret_cast<float*>(p2 + nt2);
  float* wsel = targets + a.B;
  __shared__ uint32_t mt[624];
  __shared__ double red_d[32];
  __shared__ float red_f[32];
  __shared__ double sh_total, sh_offset;
  __shared__ float sh_max;
  __shared__ int sh_exact, sh_fail;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  Ctl* ctl = a.ctl;

  if (tid == 0) ctl->probe[0] = clock64();
  // ---- 1. inclusive scan of tile sums into shared memory, min non-zero weight
  double carry = 0.0;
  float mn = FLT_MAX;
  for (int base = 0; base < nt2; base += kSelThreads) {
    int i = base + tid;
    double v = (i < nt2) ? a.l2[i] : 0.0;
    if (i < nt2) mn = fminf(mn, a.l2min[i]);
    // warp inclusive scan
    double x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      double y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) red_d[warp] = x;
    __syncthreads();
    if (warp == 0) {
      double t = red_d[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        double y = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += y;
      }
      red_d[lane] = t;  // inclusive over warps
    }
    __syncthreads();
    double woff = (warp > 0) ? red_d[warp - 1] : 0.0;
    if (i < nt2) p2[i] = carry + woff + x;
    carry += red_d[31];
    __syncthreads();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  if (lane == 0) red_f[warp] = mn;
  // mt19937 state
  for (int i = tid; i < 624; i += kSelThreads) mt[i] = ctl->mt[i];
  __syncthreads();

  if (tid == 0) ctl->probe[1] = clock64();
  // ---- 2. exactness test, offset of head, totals
  if (warp == 0) {
    float m = red_f[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
    // O = physical inclusive prefix at head - 1
    double part = 0.0;
    if (a.head > 0) {
      const int last = a.head - 1;
      const int t = last / kTile, g = (last % kTile) / kGroup, o = last % kGroup;
      if (lane < g) part += a.l1[(size_t)t * (kTile / kGroup) + lane];
      if (lane <= o) part += (double)a.w[(size_t)t * kTile + (size_t)g * kGroup + lane];
      part = warp_sum(part);
      if (t > 0) part += p2[t - 1];
    }
    if (lane == 0) {
      const double total = p2[nt2 - 1];
      sh_total = total;
      sh_offset = part;
      int exact = 1;
      if (m != FLT_MAX) {
        // ulp(m) * 2^53: every multiple of ulp(m) below it is a double
        int e;
        frexpf(m, &e);                           // m = f * 2^e, f in [0.5, 1)
        int ulp_exp = (e - 1) - 23;              // ulp of a normal float
        if (ulp_exp < -149) ulp_exp = -149;      // denormals
        exact = total < ldexp(1.0, ulp_exp + 52);  // 2^52, not 2^53: margin for an inexact `total`
      }
      if (a.force_ordered) exact = 0;
      sh_exact = exact;
      sh_fail = 0;
    }
  }
  __syncthreads();

  if (tid == 0) ctl->probe[2] = clock64();
  // ---- 3. targets: rand_i = U[0, segment) + i * segment, clamped (lines 265-280)
  const float sumf = (float)ctl->sum;
  const float segment = __fdiv_rn(sumf, (float)a.B);
  {
    int pos = ctl->mti;
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
    Hier h{a.w, a.l1, p2, nt2};
    const double total = sh_total, off = sh_offset;
    const double m1 = total - off;  // mass of [head, S)
    for (int i = tid; i < a.B; i += kSelThreads) {
      const double target = (double)targets[i];
      int slot;
      if (hit(m1, target)) {
        slot = search_slot(h, -off, target);
      } else {
        slot = search_slot(h, m1, target);
        if (slot >= a.head) slot = -1;  // ran past the wrapped part
      }
      int logical = -1;
      if (slot >= 0) logical = slot >= a.head ? slot - a.head : slot + a.S - a.head;
      if (slot < 0 || slot >= a.S || logical >= size) {
        sh_fail = 1;  // reference: assert(false) at lines 297-302
        slot = a.head;
      }
      a.ids[i] = slot;
      wsel[i] = a.w[slot];
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
    double part = 0.0;
    if (sh_exact) {
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
        for (int k = 0; k < kSelThreads / 32; ++k) tot += red_d[k];
        ctl->sum += -tot;
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
  build_hierarchy_kernel<<<a.S_pad / kTile, 256, 0, stream>>>(a.w, a.l1, a.l2, a.l2min);
  return cudaGetLastError();
}

cudaError_t launch_select(const SelectArgs& a, cudaStream_t stream) {
  const size_t smem = (size_t)(a.S_pad / kTile) * sizeof(double) + (size_t)a.B * 2 * sizeof(float);
  static size_t configured = 0;
  if (smem > 40 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  select_kernel<<<1, kSelThreads, smem, stream>>>(a);
  return cudaGetLastError();
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
  update_kernel<<<1, 1024, smem, stream>>>(ctl, w, evicted, ids, n, prio, alpha, is_weight);
  return cudaGetLastError();
}

}  // namespace rr
