// pl_filter.cu -- splat / blur / slice over a cached lattice (sm_100a).
//
// Value layout in HBM: pixel-major src/dst [n][C] (the reference's own layout,
// permutohedral.h:212-232) and vertex-major lattice values [M+1][Cc], where
// row M is a permanent zero row that absent blur neighbours point at
// (the reference's `zero` vector, permutohedral.h:494-495,518,522).  Channels
// are the contiguous dimension everywhere, so one pixel's / vertex's channel
// vector is a run of float4s: scatter (splat) issues vector reductions
// (red.global.add.v4.f32 -> REDG.E.ADD.F32x4), gathers (blur, slice) issue
// 128-bit loads.
//
// Very wide value matrices (the backward pass filters 2L(1+d) channels,
// crf/gaussian_matrix.py:450-458) are processed in channel chunks so the two
// ping-pong lattice buffers stay bounded; the lattice itself is reused.
#include <cooperative_groups.h>
#include <math.h>

#include "pl_common.cuh"

namespace pl {

namespace {

// ---- small vector helpers ---------------------------------------------------
template <int VEC>
struct Vec;
template <>
struct Vec<4> {
  float4 v;
  __device__ __forceinline__ static Vec load(const float* p) {
    Vec r;
    r.v = *reinterpret_cast<const float4*>(p);
    return r;
  }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
  __device__ __forceinline__ static Vec zero() {
    Vec r;
    r.v = make_float4(0.f, 0.f, 0.f, 0.f);
    return r;
  }
};
template <>
struct Vec<1> {
  float v;
  __device__ __forceinline__ static Vec load(const float* p) {
    Vec r;
    r.v = *p;
    return r;
  }
  __device__ __forceinline__ void store(float* p) const { *p = v; }
  __device__ __forceinline__ static Vec zero() {
    Vec r;
    r.v = 0.f;
    return r;
  }
};

__device__ __forceinline__ void red_add(float* addr, const Vec<4>& x) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x.v.x),
               "f"(x.v.y), "f"(x.v.z), "f"(x.v.w)
               : "memory");
}
__device__ __forceinline__ void red_add(float* addr, const Vec<1>& x) { atomicAdd(addr, x.v); }

template <typename F>
__device__ __forceinline__ Vec<4> map2(const Vec<4>& a, const Vec<4>& b, F f) {
  Vec<4> r;
  r.v.x = f(a.v.x, b.v.x);
  r.v.y = f(a.v.y, b.v.y);
  r.v.z = f(a.v.z, b.v.z);
  r.v.w = f(a.v.w, b.v.w);
  return r;
}
template <typename F>
__device__ __forceinline__ Vec<1> map2(const Vec<1>& a, const Vec<1>& b, F f) {
  Vec<1> r;
  r.v = f(a.v, b.v);
  return r;
}

// ---------------------------------------------------------------------------
// splat                                              permutohedral.h:452-455
//   val[vid[p][r]][c] += bary[p][r] * src[p][c]
// One thread per (pixel, channel vector); the d+1 (vertex, weight) pairs of a
// pixel are re-read by the threads sharing it (L1 broadcast).
// ---------------------------------------------------------------------------
template <int VEC, typename Idx>
__global__ void __launch_bounds__(256)
splat_kernel(const float* __restrict__ src, int64_t src_stride, int Cc,
             const int32_t* __restrict__ vid, const float* __restrict__ bary, int D1, Idx n,
             float* __restrict__ val) {
  const Idx V = (Idx)(Cc / VEC);
  const Idx total = n * V;
  for (Idx t = (Idx)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (Idx)gridDim.x * blockDim.x) {
    const Idx p = t / V;
    const int c = (int)(t - p * V) * VEC;
    const Vec<VEC> s = Vec<VEC>::load(src + (int64_t)p * src_stride + c);
    const int32_t* pv = vid + (int64_t)p * D1;
    const float* pw = bary + (int64_t)p * D1;
    for (int r = 0; r < D1; r++) {
      const float w = pw[r];
      Vec<VEC> ws;
      ws.v = s.v;
      Vec<VEC> wv;
      wv = map2(ws, ws, [w](float a, float) { return __fmul_rn(w, a); });
      red_add(val + (int64_t)pv[r] * Cc + c, wv);
    }
  }
}

// Warp-aggregated form for narrow inputs (V = Cc / VEC in {1, 2, 4} threads per pixel): the lanes
// of a warp hold consecutive pixels, and consecutive pixels mostly hit the same vertex for a given
// remainder.  Runs of equal vertex ids are summed with a segmented shuffle reduction (5 steps) and
// only the head of a run issues the atomic: ~5x fewer atomics on natural images.  The sums within a
// run are taken in a fixed order; runs still meet in memory in any order, as before.
__device__ __forceinline__ Vec<4> shfl_down_vec(const Vec<4>& a, int d) {
  Vec<4> r;
  r.v.x = __shfl_down_sync(0xffffffffu, a.v.x, d);
  r.v.y = __shfl_down_sync(0xffffffffu, a.v.y, d);
  r.v.z = __shfl_down_sync(0xffffffffu, a.v.z, d);
  r.v.w = __shfl_down_sync(0xffffffffu, a.v.w, d);
  return r;
}
__device__ __forceinline__ Vec<1> shfl_down_vec(const Vec<1>& a, int d) {
  Vec<1> r;
  r.v = __shfl_down_sync(0xffffffffu, a.v, d);
  return r;
}

template <int VEC, int V, typename Idx>
__global__ void __launch_bounds__(256)
splat_agg_kernel(const float* __restrict__ src, int64_t src_stride, int Cc,
                 const int32_t* __restrict__ vid, const float* __restrict__ bary, int D1, Idx n,
                 float* __restrict__ val) {
  const Idx total = n * (Idx)V;
  const int lane = threadIdx.x & 31;
  const Idx stride = (Idx)gridDim.x * blockDim.x;
  // warp-uniform trip count: every lane of a warp takes part in the shuffles
  for (Idx base = (Idx)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < total; base += stride) {
    const Idx t = base + lane;
    const bool ok = t < total;
    const Idx p = ok ? t / V : 0;
    const int c = ok ? (int)(t - p * V) * VEC : 0;
    const Vec<VEC> s = ok ? Vec<VEC>::load(src + (int64_t)p * src_stride + c) : Vec<VEC>::zero();
    for (int r = 0; r < D1; r++) {
      const int v = ok ? vid[(int64_t)p * D1 + r] : (-1 - lane);
      const float w = ok ? bary[(int64_t)p * D1 + r] : 0.f;
      Vec<VEC> acc = map2(s, s, [w](float a, float) { return __fmul_rn(w, a); });
      // the lane V places down holds the same channels of the previous pixel
      const int vprev = __shfl_up_sync(0xffffffffu, v, V);
      const bool head = (lane < V) || (vprev != v);
      const unsigned heads = __ballot_sync(0xffffffffu, head);
#pragma unroll
      for (int d = V; d < 32; d <<= 1) {
        const Vec<VEC> o = shfl_down_vec(acc, d);
        // lanes lane+V, lane+2V, ..., lane+d of my column: none of them may start a new run
        constexpr unsigned kCol =
            V == 1 ? 0xffffffffu : (V == 2 ? 0x55555555u : (V == 4 ? 0x11111111u : 0x01010101u));
        const unsigned between =
            (lane + d < 32) ? ((kCol << (lane % V)) & (((1u << d) - 1u) << (lane + 1))) : 0u;
        if (lane + d < 32 && (heads & between) == 0)
          acc = map2(acc, o, [](float a, float b) { return a + b; });
      }
      if (ok && head) red_add(val + (int64_t)v * Cc + c, acc);
    }
  }
}

// Odd narrow inputs (C = 3, 5, 6, 7: an RGB image, say): one thread per pixel, the channels in
// a loop, the same run aggregation per channel.
template <typename Idx>
__global__ void __launch_bounds__(256)
splat_pix_kernel(const float* __restrict__ src, int64_t src_stride, int Cc,
                 const int32_t* __restrict__ vid, const float* __restrict__ bary, int D1, Idx n,
                 float* __restrict__ val) {
  const int lane = threadIdx.x & 31;
  const Idx stride = (Idx)gridDim.x * blockDim.x;
  for (Idx base = (Idx)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride) {
    const Idx p = base + lane;
    const bool ok = p < n;
    float s[8];
#pragma unroll
    for (int c = 0; c < 8; c++) s[c] = (ok && c < Cc) ? src[(int64_t)p * src_stride + c] : 0.f;
    for (int r = 0; r < D1; r++) {
      const int v = ok ? vid[(int64_t)p * D1 + r] : (-1 - lane);
      const float w = ok ? bary[(int64_t)p * D1 + r] : 0.f;
      const int vprev = __shfl_up_sync(0xffffffffu, v, 1);
      const bool head = (lane == 0) || (vprev != v);
      const unsigned heads = __ballot_sync(0xffffffffu, head);
      float acc[8];
#pragma unroll
      for (int c = 0; c < 8; c++) acc[c] = __fmul_rn(w, s[c]);
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const bool take = (lane + d < 32) && ((heads & (((1u << d) - 1u) << (lane + 1))) == 0);
#pragma unroll
        for (int c = 0; c < 8; c++) {
          const float o = __shfl_down_sync(0xffffffffu, acc[c], d);
          if (take) acc[c] += o;
        }
      }
      if (ok && head) {
#pragma unroll
        for (int c = 0; c < 8; c++)
          if (c < Cc) atomicAdd(val + (int64_t)v * Cc + c, acc[c]);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// blur, one axis                                     permutohedral.h:502-527
//   out[v] = 2 * (0.25 in[v-] + 0.5 in[v] + 0.25 in[v+])
// evaluated as fma(0.5, in[v+], fma(0.5, in[v-], in[v])): the powers of two
// are exact, so each fma rounds exactly where the reference's adds round.
// ---------------------------------------------------------------------------
template <int VEC, typename Idx>
__global__ void __launch_bounds__(256)
blur_kernel(const float* __restrict__ in, float* __restrict__ out, const int2* __restrict__ nbr_j,
            Idx M, int Cc) {
  const Idx V = (Idx)(Cc / VEC);
  const Idx total = M * V;
  for (Idx t = (Idx)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (Idx)gridDim.x * blockDim.x) {
    const Idx v = t / V;
    const int c = (int)(t - v * V) * VEC;
    const int2 nb = nbr_j[v];
    const Vec<VEC> o = Vec<VEC>::load(in + (int64_t)v * Cc + c);
    const Vec<VEC> a = Vec<VEC>::load(in + (int64_t)nb.x * Cc + c);
    const Vec<VEC> b = Vec<VEC>::load(in + (int64_t)nb.y * Cc + c);
    Vec<VEC> r = map2(a, o, [](float x, float y) { return __fmaf_rn(0.5f, x, y); });
    r = map2(b, r, [](float x, float y) { return __fmaf_rn(0.5f, x, y); });
    r.store(out + (int64_t)v * Cc + c);
  }
}

// ---------------------------------------------------------------------------
// slice (+ fused epilogue)                           permutohedral.h:473-483
//   dst[p][c] = sum_r bary[p][r] * val[vid[p][r]][c] / (1 + 2^-d)  - sub + add
// EXACT: per-term multiply, divide, add, as the reference spells it.
// otherwise the constant is folded into the weight and accumulated with FMA.
// ---------------------------------------------------------------------------
template <int VEC, bool EXACT, typename Idx>
__global__ void __launch_bounds__(256)
slice_kernel(const float* __restrict__ val, int Cc, const int32_t* __restrict__ vid,
             const float* __restrict__ bary, int D1, Idx n, float denom, float* dst,
             int64_t dst_stride, const float* sub, int64_t sub_stride, const float* add,
             int64_t add_stride) {
  const Idx V = (Idx)(Cc / VEC);
  const Idx total = n * V;
  const float inv = 1.0f / denom;
  for (Idx t = (Idx)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (Idx)gridDim.x * blockDim.x) {
    const Idx p = t / V;
    const int c = (int)(t - p * V) * VEC;
    const int32_t* pv = vid + (int64_t)p * D1;
    const float* pw = bary + (int64_t)p * D1;
    Vec<VEC> acc = Vec<VEC>::zero();
    for (int r = 0; r < D1; r++) {
      const float w = pw[r];
      const Vec<VEC> x = Vec<VEC>::load(val + (int64_t)pv[r] * Cc + c);
      if (EXACT) {
        acc = map2(acc, x, [w, denom](float a, float b) {
          return __fadd_rn(a, __fdiv_rn(__fmul_rn(w, b), denom));
        });
      } else {
        const float wi = w * inv;
        acc = map2(acc, x, [wi](float a, float b) { return __fmaf_rn(wi, b, a); });
      }
    }
    // E = E0 + (filtered - src), crf_module.py:101 with gaussian_matrix.py:352
    if (sub != nullptr) {
      const Vec<VEC> s = Vec<VEC>::load(sub + (int64_t)p * sub_stride + c);
      acc = map2(acc, s, [](float a, float b) { return __fsub_rn(a, b); });
    }
    if (add != nullptr) {
      const Vec<VEC> s = Vec<VEC>::load(add + (int64_t)p * add_stride + c);
      acc = map2(acc, s, [](float a, float b) { return __fadd_rn(b, a); });
    }
    acc.store(dst + (int64_t)p * dst_stride + c);
  }
}

// All d+1 axes in ONE cooperative launch: a grid-wide barrier separates the Jacobi passes
// (the lattice of a stereo frame is a few MB, L2 resident; 6 launches of a few microseconds
// each were launch-bound).  Optionally zeroes the buffer that does NOT hold the result, which
// is the next mean-field iteration's splat target.
template <typename Idx, int BATCH>
__global__ void __launch_bounds__(256)
blur_all_kernel(float* __restrict__ buf0, float* __restrict__ buf1, const int2* __restrict__ nbr,
                Idx M, int Cc, int D1, int zero_other) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  const Idx V = (Idx)(Cc / 4);
  const Idx total = M * V;
  const Idx stride = (Idx)gridDim.x * blockDim.x;
  float* in = buf0;
  float* out = buf1;
  for (int j = 0; j < D1; j++) {
    const int2* nbr_j = nbr + (size_t)j * M;
    // BATCH independent items per trip: their neighbour ids, then all 3 * BATCH row loads, are in
    // flight together.  Lattices that spill out of L2 are DRAM-latency bound (ncu: issue-active 10 %,
    // half the warp slots idle) and want the deeper batch; L2-resident ones want BATCH = 1 and more CTAs.
    for (Idx t0 = (Idx)blockIdx.x * blockDim.x + threadIdx.x; t0 < total; t0 += stride * BATCH) {
      int64_t off[BATCH], na[BATCH], nc[BATCH];
      bool on[BATCH];
#pragma unroll
      for (int k = 0; k < BATCH; k++) {
        const Idx t = t0 + (Idx)k * stride;
        on[k] = t < total;
        const Idx v = on[k] ? t / V : 0;
        const int c = on[k] ? (int)(t - v * V) * 4 : 0;
        const int2 nb = nbr_j[v];
        off[k] = (int64_t)v * Cc + c;
        na[k] = (int64_t)nb.x * Cc + c;
        nc[k] = (int64_t)nb.y * Cc + c;
      }
      float4 o[BATCH], a[BATCH], b[BATCH];
#pragma unroll
      for (int k = 0; k < BATCH; k++) {
        o[k] = __ldcg(reinterpret_cast<const float4*>(in + off[k]));
        a[k] = __ldcg(reinterpret_cast<const float4*>(in + na[k]));
        b[k] = __ldcg(reinterpret_cast<const float4*>(in + nc[k]));
      }
#pragma unroll
      for (int k = 0; k < BATCH; k++) {
        if (!on[k]) continue;
        float4 r;
        r.x = __fmaf_rn(0.5f, b[k].x, __fmaf_rn(0.5f, a[k].x, o[k].x));
        r.y = __fmaf_rn(0.5f, b[k].y, __fmaf_rn(0.5f, a[k].y, o[k].y));
        r.z = __fmaf_rn(0.5f, b[k].z, __fmaf_rn(0.5f, a[k].z, o[k].z));
        r.w = __fmaf_rn(0.5f, b[k].w, __fmaf_rn(0.5f, a[k].w, o[k].w));
        *reinterpret_cast<float4*>(out + off[k]) = r;
      }
    }
    grid.sync();
    float* t2 = in;
    in = out;
    out = t2;
  }
  if (zero_other) {  // `in` holds the result; `out` is free
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (Idx t = (Idx)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride)
      *reinterpret_cast<float4*>(out + (int64_t)t * 4) = z;
  }
}

// (Two barrier-free alternatives for these small lattices were measured and dropped, 740 x 500 x 64, M = 8 750,
// against 27 us for this kernel: a W-channel slice of the whole lattice in ONE CTA's shared memory with
// __syncthreads() between the passes -- 47 us, 41 us with the neighbour ids of a pass requested together --
// and the 8-CTA cluster form over DSMEM below -- 25-39 us.  ncu puts 65 % of this kernel's stall samples in
// grid.sync(), but a single CTA's latency chain per pass is longer still.)
// Small lattices (a stereo frame with the module's default guide scales has M ~ 10^4
// vertices, a few MB of values): every thread owns ITEMS (vertex, float4) items for the whole
// blur.  Its 2(d+1) neighbour offsets are fetched once, its own value stays in registers
// across the axes, so a pass is two L2 gathers + one store + the grid barrier.  d = 5 only.
constexpr int kBlurRegThreads = 512;
template <int ITEMS>
__global__ void __launch_bounds__(kBlurRegThreads, 2)
blur_reg_kernel(float* __restrict__ buf0, float* __restrict__ buf1, const int2* __restrict__ nbr,
                uint32_t M, int Cc, int zero_other) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  constexpr int D1 = 6;
  const uint32_t V = (uint32_t)(Cc / 4);
  const uint32_t total = M * V;
  uint32_t off[ITEMS], na[ITEMS][D1], nb[ITEMS][D1];
  float4 val[ITEMS];
  bool live[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const uint32_t t = ((uint32_t)blockIdx.x * ITEMS + i) * kBlurRegThreads + threadIdx.x;
    live[i] = t < total;
    const uint32_t v = live[i] ? t / V : 0;
    const uint32_t c = live[i] ? (t - v * V) * 4 : 0;
    off[i] = v * (uint32_t)Cc + c;
#pragma unroll
    for (int j = 0; j < D1; j++) {
      const int2 p = nbr[(size_t)j * M + v];
      na[i][j] = (uint32_t)p.x * (uint32_t)Cc + c;
      nb[i][j] = (uint32_t)p.y * (uint32_t)Cc + c;
    }
    val[i] = live[i] ? __ldcg(reinterpret_cast<const float4*>(buf0 + off[i]))
                     : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float* in = buf0;
  float* out = buf1;
#pragma unroll
  for (int j = 0; j < D1; j++) {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      if (live[i]) {
        const float4 a = __ldcg(reinterpret_cast<const float4*>(in + na[i][j]));
        const float4 b = __ldcg(reinterpret_cast<const float4*>(in + nb[i][j]));
        float4 r;
        r.x = __fmaf_rn(0.5f, b.x, __fmaf_rn(0.5f, a.x, val[i].x));
        r.y = __fmaf_rn(0.5f, b.y, __fmaf_rn(0.5f, a.y, val[i].y));
        r.z = __fmaf_rn(0.5f, b.z, __fmaf_rn(0.5f, a.z, val[i].z));
        r.w = __fmaf_rn(0.5f, b.w, __fmaf_rn(0.5f, a.w, val[i].w));
        *reinterpret_cast<float4*>(out + off[i]) = r;
        val[i] = r;
      }
    }
    grid.sync();
    float* t2 = in;
    in = out;
    out = t2;
  }
  if (zero_other) {  // `in` holds the result (D1 is even: buf0); `out` is free
#pragma unroll
    for (int i = 0; i < ITEMS; i++)
      if (live[i]) *reinterpret_cast<float4*>(out + off[i]) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// (Also measured and dropped on these lattices: TWO axes per pass -- three grid barriers instead of six,
// 8 gathers per item and pass instead of 2 x 2, bit-identical -- 34.4 us against 26.5: a pass is not only
// the barrier, the uncached gathers of rows other SMs have just written cost as much again, and they
// scale with the gather count.)
// The same lattices WITHOUT any barrier between the passes (26.5 -> 14 us at M = 8 750 x 64 channels).
// Every pass writes its own scratch buffer (pass j -> flow[j], the last one -> the result buffer), all
// pre-filled with a sentinel bit pattern no arithmetic result can have: 0xFFFFFFFF, a NaN payload the
// FMA unit never produces (a NaN result is always 0x7FFFFFFF).  A thread of pass j + 1 re-reads its two
// neighbour values from flow[j] until none of the four components is the sentinel.  Each 32-bit component
// is written exactly once per call, so a non-sentinel component is final and the four need no joint
// atomicity; nothing is ordered against anything else, so there is no counter and no flag.  Per pass the
// chain is a store becoming visible plus the neighbour's load round trip, instead of fence + atomic
// arrival + flag poll + load.  (The fence behind the store is not needed for correctness; it makes the
// store visible sooner: 16.3 -> 14 us.  Backing off between polls costs 4 us.)
// The sentinel state is restored by the NEXT call: the scratch has two sets, a call uses set s and
// refills set 1 - s with fire-and-forget stores.  The launch is cooperative only for its co-residency
// guarantee: every producer must be running.  Row M (absent neighbour) is never read, the id is tested
// instead.  `buf0` is only read and cannot be zeroed here -- another thread may still be reading it in
// pass 0 when this one is done -- so the next splat target is a THIRD buffer, `zero_buf`, that nothing
// else touches while this kernel runs (the mean-field loop rotates three value buffers), or the caller
// clears buf0 behind the kernel.
// Same fp32 operations in the same order as blur_reg_kernel: bit-identical results
// (tests/test_blur_forms_gpu.py).  A poll that never succeeds traps after ~2 s.
constexpr uint32_t kFlowSentinel = 0xFFFFFFFFu;
__device__ __forceinline__ float4 ld_relaxed_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_f4(float* p, const float4& v) {
  asm volatile("st.relaxed.gpu.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ bool flow_ready(const float4& v) {
  return __float_as_uint(v.x) != kFlowSentinel && __float_as_uint(v.y) != kFlowSentinel &&
         __float_as_uint(v.z) != kFlowSentinel && __float_as_uint(v.w) != kFlowSentinel;
}
// ITEMS: adjacent float4 channel groups of ONE vertex per thread (they share the neighbour ids).
template <int ITEMS>
__global__ void __launch_bounds__(kBlurRegThreads, 2)
blur_flow_kernel(const float* __restrict__ buf0, float* __restrict__ result, float* __restrict__ flow_use,
                 float* __restrict__ flow_refill, size_t flow_stride, float* __restrict__ zero_buf,
                 const int2* __restrict__ nbr, uint32_t M, int Cc) {
  constexpr int D1 = 6;
  const uint32_t VT = (uint32_t)(Cc / (4 * ITEMS));     // threads per vertex
  const uint32_t total = M * VT;
  const uint32_t t = (uint32_t)blockIdx.x * kBlurRegThreads + threadIdx.x;
  if (t >= total) return;
  const uint32_t v = t / VT;
  const uint32_t c = (t - v * VT) * (4 * ITEMS);
  const uint32_t off = v * (uint32_t)Cc + c;
  int2 own[D1];
#pragma unroll
  for (int j = 0; j < D1; j++) own[j] = nbr[(size_t)j * M + v];
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  {
    const float sf = __uint_as_float(kFlowSentinel);
    const float4 s4 = make_float4(sf, sf, sf, sf);
#pragma unroll
    for (int j = 0; j < D1 - 1; j++)
#pragma unroll
      for (int i = 0; i < ITEMS; i++)
        *reinterpret_cast<float4*>(flow_refill + j * flow_stride + off + 4 * i) = s4;
  }
  // the neighbour table and the other scratch set are nobody else's business; the input and the spare
  // buffer belong to the kernel in front until it has completed (programmatic dependent launch)
  pdl_wait();
  pdl_launch_dependents();
  float4 val[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; i++) val[i] = __ldcg(reinterpret_cast<const float4*>(buf0 + off + 4 * i));
  if (zero_buf != nullptr) {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) *reinterpret_cast<float4*>(zero_buf + off + 4 * i) = zero;
  }
#pragma unroll
  for (int j = 0; j < D1; j++) {
    const uint32_t ia = (uint32_t)own[j].x, ib = (uint32_t)own[j].y;
    float4 a[ITEMS], b[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) a[i] = b[i] = zero;
    if (j == 0) {   // the input is complete before the launch
#pragma unroll
      for (int i = 0; i < ITEMS; i++) {
        if (ia < M) a[i] = __ldcg(reinterpret_cast<const float4*>(buf0 + ia * (uint32_t)Cc + c + 4 * i));
        if (ib < M) b[i] = __ldcg(reinterpret_cast<const float4*>(buf0 + ib * (uint32_t)Cc + c + 4 * i));
      }
    } else {
      const float* in = flow_use + (size_t)(j - 1) * flow_stride;
      const float* pa = in + ia * (uint32_t)Cc + c;
      const float* pb = in + ib * (uint32_t)Cc + c;
      bool ra = ia >= M, rb = ib >= M;
      uint32_t spins = 0;
      while (!(ra && rb)) {
        if (!ra) {
          bool ok = true;
#pragma unroll
          for (int i = 0; i < ITEMS; i++) {
            a[i] = ld_relaxed_f4(pa + 4 * i);
            ok = ok && flow_ready(a[i]);
          }
          ra = ok;
        }
        if (!rb) {
          bool ok = true;
#pragma unroll
          for (int i = 0; i < ITEMS; i++) {
            b[i] = ld_relaxed_f4(pb + 4 * i);
            ok = ok && flow_ready(b[i]);
          }
          rb = ok;
        }
        if (++spins > (1u << 22)) {
          printf("libpl_lattice: blur_flow_kernel: pass %d never saw its neighbours (block %d, thread %d)\n", j,
                 (int)blockIdx.x, (int)threadIdx.x);
          __trap();
        }
      }
    }
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      float4 r;
      r.x = __fmaf_rn(0.5f, b[i].x, __fmaf_rn(0.5f, a[i].x, val[i].x));
      r.y = __fmaf_rn(0.5f, b[i].y, __fmaf_rn(0.5f, a[i].y, val[i].y));
      r.z = __fmaf_rn(0.5f, b[i].z, __fmaf_rn(0.5f, a[i].z, val[i].z));
      r.w = __fmaf_rn(0.5f, b[i].w, __fmaf_rn(0.5f, a[i].w, val[i].w));
      val[i] = r;
      if (j < D1 - 1)
        st_relaxed_f4(flow_use + (size_t)j * flow_stride + off + 4 * i, r);
      else
        *reinterpret_cast<float4*>(result + off + 4 * i) = r;
    }
    if (j < D1 - 1) __threadfence();
  }
}

// ---------------------------------------------------------------------------
// blur inside thread-block clusters, no grid-wide barrier.
//
// The channels of the lattice are independent, so one CLUSTER of 8 CTAs takes a group of four
// channels (one float4 per vertex) of the WHOLE lattice into its distributed shared memory -- CTA r
// owns vertices [r * per, (r + 1) * per) in two ping-pong buffers -- runs all d+1 Jacobi passes there
// (own value from local shared memory, the two neighbours with ld.shared::cluster from whichever CTA
// owns them) and separates the passes with the hardware cluster barrier (~380 cycles) instead of
// grid.sync() (ncu on blur_reg_kernel: 65 % of the stall samples in the barrier, ~2.7 us per pass).
// Same arithmetic, so bit-identical results.  Up to 51 200 vertices (2 x 6 400 x 16 bytes per CTA); used for
// lattices too large for blur_reg_kernel, where it replaces blur_all_kernel.
// ---------------------------------------------------------------------------
constexpr int kCluster = 8;
constexpr int kClusterThreads = 1024;
constexpr uint32_t kClusterMaxPer = 6400;    // vertices per CTA: 2 * 6400 * 16 B = 204 800 B of shared memory

__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float4 ld_cluster_f4(uint32_t local_addr, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(remote)
               : "memory");
  return v;
}

__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kClusterThreads, 1)
blur_cluster_kernel(float* __restrict__ cur, float* __restrict__ other, const int2* __restrict__ nbr,
                    uint32_t M, int Cc, int D1, uint32_t per, int zero_other) {
  extern __shared__ __align__(16) unsigned char smem_blur[];
  float4* buf0 = reinterpret_cast<float4*>(smem_blur);
  float4* buf1 = buf0 + per;
  const uint32_t rank = blockIdx.x % kCluster;       // == %cluster_ctarank for a 1-D grid of 1-D clusters
  const uint32_t cgrp = blockIdx.x / kCluster;       // channel group (four channels)
  const uint32_t v0 = rank * per;
  const uint32_t nv = v0 < M ? min(per, M - v0) : 0u;
  const int tid = threadIdx.x;
  float* base = cur + 4 * (size_t)cgrp;

  for (uint32_t i = tid; i < nv; i += blockDim.x)
    buf0[i] = __ldcg(reinterpret_cast<const float4*>(base + (size_t)(v0 + i) * Cc));
  cluster_barrier();

  float4* in = buf0;
  float4* out = buf1;
  for (int j = 0; j < D1; j++) {
    const int2* nbr_j = nbr + (size_t)j * M + v0;
    const uint32_t in_addr = (uint32_t)__cvta_generic_to_shared(in);
    for (uint32_t i = tid; i < nv; i += blockDim.x) {
      const int2 nb = nbr_j[i];
      const float4 o = in[i];
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;     // absent neighbour (index M): zero row
      if ((uint32_t)nb.x < M) {
        const uint32_t ow = (uint32_t)nb.x / per;
        a = ld_cluster_f4(in_addr + ((uint32_t)nb.x - ow * per) * 16u, ow);
      }
      if ((uint32_t)nb.y < M) {
        const uint32_t ow = (uint32_t)nb.y / per;
        b = ld_cluster_f4(in_addr + ((uint32_t)nb.y - ow * per) * 16u, ow);
      }
      float4 r;
      r.x = __fmaf_rn(0.5f, b.x, __fmaf_rn(0.5f, a.x, o.x));
      r.y = __fmaf_rn(0.5f, b.y, __fmaf_rn(0.5f, a.y, o.y));
      r.z = __fmaf_rn(0.5f, b.z, __fmaf_rn(0.5f, a.z, o.z));
      r.w = __fmaf_rn(0.5f, b.w, __fmaf_rn(0.5f, a.w, o.w));
      out[i] = r;
    }
    cluster_barrier();      // every CTA's `out` is complete (and nobody reads `in` any more)
    float4* t = in;
    in = out;
    out = t;
  }
  // result back in place; the other global buffer is the next splat target
  float* obase = other + 4 * (size_t)cgrp;
  for (uint32_t i = tid; i < nv; i += blockDim.x) {
    *reinterpret_cast<float4*>(base + (size_t)(v0 + i) * Cc) = in[i];
    if (zero_other) *reinterpret_cast<float4*>(obase + (size_t)(v0 + i) * Cc) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15u) == 0; }

template <int VEC, typename Idx>
int apply_chunk(Lattice* L, const float* src, int64_t src_stride, float* dst, int64_t dst_stride,
                int Cc, const float* sub, int64_t sub_stride, const float* add, int64_t add_stride,
                uint32_t flags, cudaStream_t st) {
  const int D1 = L->d + 1;
  const int64_t M = L->M, n = L->n;
  const size_t row = sizeof(float) * (size_t)Cc;
  // zero the accumulation buffer (incl. row M) and the other buffer's row M
  float* cur = nullptr;
  float* nxt = nullptr;
  PL_TRY(lattice_splat_buffers(L, Cc, st, &cur, &nxt));

  const uint64_t pix_items = (uint64_t)n * (uint64_t)(Cc / VEC);
  const uint64_t vtx_items = (uint64_t)M * (uint64_t)(Cc / VEC);
  // the tile kernels give 16 lanes to a pixel row: up to 32 channels most of them idle and the
  // per-entry kernels with warp-aggregated atomics win (measured at 1 MP: C = 4..16 2-2.7x,
  // C = 32 1.3x; C = 64 the tile kernels are 1.8x ahead)
  const bool tiled = (VEC == 4) && !(flags & PL_APPLY_FLAT_KERNELS) && L->ntiles > 0 && Cc > 32;
  static const bool no_sorted = getenv("PL_SPLAT_NO_SORTED") != nullptr;
  if (tiled) {
    PL_TRY(splat_tiled(L, src, src_stride, Cc, cur, st));
  } else if ((Cc <= 3 || (VEC == 4 && Cc == 4)) && !no_sorted &&
             !(flags & PL_APPLY_FLAT_KERNELS) && L->ntiles > 0) {
    PL_TRY(splat_sorted(L, src, src_stride, Cc, cur, st));
  } else {
    const int V = Cc / VEC;
    if (V == 1) {
      PL_LAUNCH((splat_agg_kernel<VEC, 1, Idx>), grid_for(pix_items, 256, 8), 256, 0, st, src,
                src_stride, Cc, L->vid, L->bary, D1, (Idx)n, cur);
    } else if (V == 2) {
      PL_LAUNCH((splat_agg_kernel<VEC, 2, Idx>), grid_for(pix_items, 256, 8), 256, 0, st, src,
                src_stride, Cc, L->vid, L->bary, D1, (Idx)n, cur);
    } else if (V == 4) {
      PL_LAUNCH((splat_agg_kernel<VEC, 4, Idx>), grid_for(pix_items, 256, 8), 256, 0, st, src,
                src_stride, Cc, L->vid, L->bary, D1, (Idx)n, cur);
    } else if (V == 8) {
      PL_LAUNCH((splat_agg_kernel<VEC, 8, Idx>), grid_for(pix_items, 256, 8), 256, 0, st, src,
                src_stride, Cc, L->vid, L->bary, D1, (Idx)n, cur);
    } else if (VEC == 1 && Cc <= 8) {
      PL_LAUNCH((splat_pix_kernel<Idx>), grid_for((uint64_t)n, 256, 8), 256, 0, st, src, src_stride,
                Cc, L->vid, L->bary, D1, (Idx)n, cur);
    } else {
      PL_LAUNCH((splat_kernel<VEC, Idx>), grid_for(pix_items, 256, 8), 256, 0, st, src, src_stride,
                Cc, L->vid, L->bary, D1, (Idx)n, cur);
    }
  }
  if (VEC == 4) {
    float* res = nullptr;
    PL_TRY(lattice_blur(L, cur, nxt, Cc, st, &res));
    cur = res;
  } else {
    for (int j = 0; j < D1; j++) {
      PL_LAUNCH((blur_kernel<VEC, Idx>), grid_for(vtx_items, 256, 8), 256, 0, st, cur, nxt,
                L->nbr + (size_t)j * M, (Idx)M, Cc);
      float* t = cur;
      cur = nxt;
      nxt = t;
    }
  }
  const float denom = 1 + powf(2, -L->d);  // permutohedral.h:480
  if (tiled) {
    PL_TRY(slice_tiled(L, cur, Cc, denom, dst, dst_stride, sub, sub_stride, add, add_stride,
                       (flags & PL_APPLY_EXACT_SLICE) != 0, st));
  } else if (flags & PL_APPLY_EXACT_SLICE) {
    PL_LAUNCH((slice_kernel<VEC, true, Idx>), grid_for(pix_items, 256, 8), 256, 0, st, cur, Cc,
              L->vid, L->bary, D1, (Idx)n, denom, dst, dst_stride, sub, sub_stride, add,
              add_stride);
  } else {
    PL_LAUNCH((slice_kernel<VEC, false, Idx>), grid_for(pix_items, 256, 8), 256, 0, st, cur, Cc,
              L->vid, L->bary, D1, (Idx)n, denom, dst, dst_stride, sub, sub_stride, add,
              add_stride);
  }
  return PL_OK;
}

}  // namespace

// Grow-only ping-pong value buffers of (M+1) x Cc floats.
int lattice_workspace(Lattice* L, int Cc, cudaStream_t st, bool third) {
  const int64_t need = (L->M + 1) * (int64_t)Cc;
  if (L->ws_floats < need) {
    // val[1] and val[0] share one block (val[1] first): see lattice_splat_buffers
    if (L->val[1]) dev_free(L->val[1], st);
    if (L->val[2]) dev_free(L->val[2], st);
    L->val[0] = L->val[1] = L->val[2] = nullptr;
    L->ws_floats = 0;
    PL_TRY(dev_alloc((void**)&L->val[1], sizeof(float) * (size_t)need * 2, st));
    L->ws_floats = need;
  }
  if (third && L->val[2] == nullptr)
    PL_TRY(dev_alloc((void**)&L->val[2], sizeof(float) * (size_t)L->ws_floats, st));
  return PL_OK;
}

// The two value buffers for a filter with Cc channels, ready for a splat into *cur: *cur is all zero
// (rows 0..M) and row M of *nxt is zero (absent blur neighbours read it).  *cur is placed right behind
// *nxt's (M+1) x Cc floats, so that ONE memset covers row M of *nxt and all of *cur.
int lattice_splat_buffers(Lattice* L, int Cc, cudaStream_t st, float** cur, float** nxt) {
  const size_t rowf = (size_t)Cc, rows = (size_t)L->M + 1;
  L->val[0] = L->val[1] + rows * rowf;
  *nxt = L->val[1];
  *cur = L->val[0];
  PL_CUDA(cudaMemsetAsync(*nxt + (size_t)L->M * rowf, 0, sizeof(float) * rowf * (rows + 1), st));
  return PL_OK;
}

// items per thread of the barrier-free form for this shape; 0: not applicable
static int blur_flow_items(const Lattice* L, int Cc) {
  const char* form = getenv("PL_BLUR_FORM");
  if (form != nullptr && form[0] == 'r') return 0;
  if (L->d != 5 || Cc % 4 != 0 || L->M <= 0 || (uint64_t)(L->M + 1) * Cc >= 0x7fffffffull) return 0;
  const uint64_t vtx_items = (uint64_t)L->M * (uint64_t)(Cc / 4);
  const uint64_t threads = (uint64_t)PL_SM_COUNT * 2 * kBlurRegThreads;
  if (vtx_items <= threads) return 1;
  if (Cc % 8 == 0 && vtx_items <= 2 * threads) return 2;
  return 0;
}
bool lattice_blur_is_flow(const Lattice* L, int Cc) { return blur_flow_items(L, Cc) > 0; }

int lattice_flow_prepare(Lattice* L, int Cc, cudaStream_t st, void** fill_ptr, size_t* fill_bytes) {
  if (fill_bytes != nullptr) *fill_bytes = 0;
  const int64_t per = L->M * (int64_t)Cc;
  if (L->flow_floats >= per) return PL_OK;
  dev_free(L->flow, st);
  L->flow = nullptr;
  L->flow_floats = 0;
  PL_TRY(dev_alloc((void**)&L->flow, sizeof(float) * (size_t)per * 10, st));
  L->flow_floats = per;
  L->flow_set = 0;
  L->flow_dirty[0] = 0;
  L->flow_dirty[1] = per;     // uninitialised memory: to be refilled in full
  const size_t bytes = sizeof(float) * (size_t)per * 5;
  if (fill_ptr != nullptr && fill_bytes != nullptr) {
    *fill_ptr = L->flow;
    *fill_bytes = bytes;
  } else {
    PL_CUDA(cudaMemsetAsync(L->flow, 0xFF, bytes, st));
  }
  return PL_OK;
}

// d+1 Jacobi passes starting from `cur` (scratch `nxt`); *out = buffer holding the result.
// Both buffers must have a zero row M.  Cc % 4 == 0.
int lattice_blur(Lattice* L, float* cur, float* nxt, int Cc, cudaStream_t st, float** out,
                 bool zero_other, float* spare, float** zeroed) {
  if (zeroed != nullptr) *zeroed = nullptr;
  const int D1 = L->d + 1;
  const int64_t M = L->M;
  const uint64_t vtx_items = (uint64_t)M * (uint64_t)(Cc / 4);
  // one cooperative launch for all axes (all CTAs must be co-resident)
  static int coop_occ[64];             // per device: resident CTAs per SM (0: no cooperative launch)
  static PerDeviceOnce coop_once;
  const int devi = current_device_index();
  if (coop_once.need()) {
    int coop = 0, o1 = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, devi);
    if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, blur_all_kernel<uint32_t, 1>, 256, 0) ==
                    cudaSuccess)
      coop_occ[devi] = o1 > 8 ? 8 : o1;
    else
      coop_occ[devi] = 0;
  }
  const int coop_ctas_per_sm = coop_occ[devi];
  // cluster form: the lattice's four-channel slices fit the shared memory of 8 CTAs
  // (measured: 80 us against 100 us for blur_all_kernel at M = 44 561 x 64 channels, 68 against 96 at
  // M = 40 300; on lattices small enough for blur_reg_kernel below it is no faster -- 25-39 us against 27 --
  // the cluster barrier of eight 1024-thread CTAs costs more than its ~380-cycle small-CTA figure; a variant
  // with the remote addresses and the running value in registers was slower still and was dropped)
  static const bool no_cluster = getenv("PL_BLUR_NO_CLUSTER") != nullptr;   // A/B timing knob
  const bool reg_form_fits = coop_ctas_per_sm > 0 && D1 == 6 &&
                             vtx_items <= (uint64_t)PL_SM_COUNT * 2 * kBlurRegThreads * 2;
  if (!no_cluster && !reg_form_fits && M > 0 && (uint64_t)M <= (uint64_t)kCluster * kClusterMaxPer &&
      (uint64_t)(M + 1) * Cc < 0xffffffffull) {
    const uint32_t per = (((uint32_t)M + kCluster - 1) / kCluster + 7u) & ~7u;
    const size_t smem = (size_t)per * 2 * sizeof(float4);
    static PerDeviceOnce clu_attr;
    if (clu_attr.need())
      PL_CUDA(cudaFuncSetAttribute(blur_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(kClusterMaxPer * 2 * sizeof(float4))));
    PL_LAUNCH(blur_cluster_kernel, (unsigned)(kCluster * (Cc / 4)), kClusterThreads, smem, st, cur, nxt, L->nbr,
              (uint32_t)M, Cc, D1, per, zero_other ? 1 : 0);
    *out = cur;
    return PL_OK;
  }
  // barrier-free form: one or two items per thread on 2 x 148 co-resident CTAs
  const int flow_items = coop_ctas_per_sm > 0 ? blur_flow_items(L, Cc) : 0;
  if (flow_items > 0) {
    static int flow_occ[64];
    static PerDeviceOnce flow_once;
    if (flow_once.need()) {
      int o1 = 0, o2 = 0;
      const bool ok = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, blur_flow_kernel<1>, kBlurRegThreads, 0) ==
                          cudaSuccess &&
                      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, blur_flow_kernel<2>, kBlurRegThreads, 0) ==
                          cudaSuccess;
      flow_occ[devi] = ok ? (o1 < o2 ? o1 : o2) : 0;
    }
    if (flow_occ[devi] >= 2) {
      const int64_t per = M * (int64_t)Cc;
      const bool fresh = L->flow_floats < per;   // a memset sits in front of this launch
      PL_TRY(lattice_flow_prepare(L, Cc, st, nullptr, nullptr));
      const int s_use = L->flow_set, s_fill = 1 - s_use;
      const size_t stride = (size_t)L->flow_floats;
      float* use = L->flow + (size_t)s_use * 5 * stride;
      float* refill = L->flow + (size_t)s_fill * 5 * stride;
      // the kernel refills [0, per) of every pass buffer of the other set; a wider stretch left by a
      // call with more channels is refilled here
      if (L->flow_dirty[s_fill] > per) {
        for (int j = 0; j < 5; j++)
          PL_CUDA(cudaMemsetAsync(refill + j * stride + per, 0xFF,
                                  sizeof(float) * (size_t)(L->flow_dirty[s_fill] - per), st));
      }
      L->flow_dirty[s_fill] = 0;
      L->flow_dirty[s_use] = per;
      L->flow_set = s_fill;
      unsigned grid = ceil_div_u(vtx_items / flow_items, (uint64_t)kBlurRegThreads);
      uint32_t M32 = (uint32_t)M;
      const int2* nbr = L->nbr;
      const float* in0 = cur;
      // the next splat target: `spare` zeroed inside the kernel, else `cur` by a memset behind it
      float* zb = (zero_other && zeroed != nullptr && spare != nullptr && spare != cur && spare != nxt) ? spare
                                                                                                       : nullptr;
      size_t stride_arg = stride;
      void* args[] = {&in0, &nxt, &use, &refill, &stride_arg, &zb, &nbr, &M32, &Cc};
      if (g_prof_on) prof_begin("blur_flow_kernel", st);
      const void* fn = flow_items == 1 ? (const void*)blur_flow_kernel<1> : (const void*)blur_flow_kernel<2>;
      // programmatic dependent launch unless a memset sits in front (the scratch was just allocated)
      cudaError_t e = launch_ex(fn, dim3(grid), dim3(kBlurRegThreads), 0, st, args, true, !fresh);
      if (g_prof_on) prof_end(st);
      g_launches++;
      PL_CUDA(e);
      if (zb != nullptr)
        *zeroed = zb;
      else if (zero_other)
        PL_CUDA(cudaMemsetAsync(cur, 0, sizeof(float) * (size_t)M * Cc, st));
      *out = nxt;
      return PL_OK;
    }
  }
  // register-resident variant: d = 5 and at most 2 items per thread on 2 x 148 CTAs
  if (coop_ctas_per_sm > 0 && D1 == 6 && (uint64_t)(M + 1) * Cc < 0x7fffffffull &&
      vtx_items <= (uint64_t)PL_SM_COUNT * 2 * kBlurRegThreads * 2) {
    const int items = vtx_items <= (uint64_t)PL_SM_COUNT * 2 * kBlurRegThreads ? 1 : 2;
    unsigned grid = ceil_div_u(vtx_items, (uint64_t)kBlurRegThreads * items);
    uint32_t M32 = (uint32_t)M;
    int zo = zero_other ? 1 : 0;
    const int2* nbr = L->nbr;
    void* args[] = {&cur, &nxt, &nbr, &M32, &Cc, &zo};
    const void* fn = items == 1 ? (const void*)blur_reg_kernel<1> : (const void*)blur_reg_kernel<2>;
    if (g_prof_on) prof_begin("blur_reg_kernel", st);
    cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kBlurRegThreads), args, 0, st);
    if (g_prof_on) prof_end(st);
    g_launches++;
    PL_CUDA(e);
    *out = cur;  // six passes: back in the buffer it started from
    return PL_OK;
  }
  if (coop_ctas_per_sm > 0 && vtx_items < 0x7fffffffull) {
    // The passes are latency bound (ncu, profiles/r2b_summary.md: issue-active 10 %, half the warp
    // slots idle at 4 CTAs per SM).  Measured on lattices of 10 MB ... 450 MB per buffer: 6 CTAs per SM
    // while both buffers sit in L2, 8 beyond (1.69 ms against 2.16 ms with 4 at M = 438 000 x 256
    // channels = 0.49 of the HBM peak); four items per thread and trip was never better than that.
    const bool in_l2 = (uint64_t)(M + 1) * Cc * 8ull <= (64ull << 20);
    const int ctas = in_l2 ? (coop_ctas_per_sm < 6 ? coop_ctas_per_sm : 6) : coop_ctas_per_sm;
    unsigned grid = grid_for(vtx_items, 256, (unsigned)ctas);
    uint32_t M32 = (uint32_t)M;
    int zo = zero_other ? 1 : 0;
    int d1 = D1;
    const int2* nbr = L->nbr;
    void* args[] = {&cur, &nxt, &nbr, &M32, &Cc, &d1, &zo};
    const void* fn = (const void*)blur_all_kernel<uint32_t, 1>;
    if (g_prof_on) prof_begin("blur_all_kernel", st);
    cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(256), args, 0, st);
    if (g_prof_on) prof_end(st);
    g_launches++;
    PL_CUDA(e);
    *out = (D1 & 1) ? nxt : cur;
    return PL_OK;
  }
  float* a = cur;
  float* b = nxt;
  for (int j = 0; j < D1; j++) {
    if (vtx_items < 0x7fffffffull) {
      PL_LAUNCH((blur_kernel<4, uint32_t>), grid_for(vtx_items, 256, 8), 256, 0, st, a, b,
                L->nbr + (size_t)j * M, (uint32_t)M, Cc);
    } else {
      PL_LAUNCH((blur_kernel<4, uint64_t>), grid_for(vtx_items, 256, 8), 256, 0, st, a, b,
                L->nbr + (size_t)j * M, (uint64_t)M, Cc);
    }
    float* t = a;
    a = b;
    b = t;
  }
  if (zero_other) PL_CUDA(cudaMemsetAsync(b, 0, sizeof(float) * (size_t)M * Cc, st));
  *out = a;
  return PL_OK;
}

int lattice_apply_dot(Lattice* L, const float* src, int64_t src_stride, int C, const float* rowscale,
                      int64_t rs_stride, const float* other, int64_t other_stride, float* out,
                      int64_t out_stride, float coef, cudaStream_t st, int J) {
  const int64_t rows = L->M + 1;
  if (J < 1 || (J > 1 && (J != 5 || L->d != 5 || rowscale == nullptr || L->M >= (1 << 24)))) {
    set_last_error("pl_lattice_apply_dot: %d row factors at once is not supported here", J);
    return PL_ERR_INVALID;
  }
  const int Cs = C;   // channels of the source; the value rows carry J * C
  C = J * Cs;
  if (Cs <= 0 || Cs % 4 != 0 || src == nullptr || other == nullptr || out == nullptr || src_stride < Cs ||
      other_stride < Cs || src_stride % 4 != 0 || other_stride % 4 != 0 || !aligned16(src) ||
      !aligned16(other) || L->ntiles <= 0 || (uint64_t)rows * (uint64_t)C * sizeof(float) > (2ull << 30)) {
    set_last_error("pl_lattice_apply_dot: unsupported arguments (C=%d)", C);
    return PL_ERR_INVALID;
  }
  PL_TRY(lattice_workspace(L, C, st));
  float* cur = nullptr;
  float* nxt = nullptr;
  PL_TRY(lattice_splat_buffers(L, C, st, &cur, &nxt));
  PL_TRY(splat_tiled(L, src, src_stride, Cs, cur, st, rowscale, rs_stride, J));
  float* res = nullptr;
  PL_TRY(lattice_blur(L, cur, nxt, C, st, &res));
  const float denom = 1 + powf(2, -L->d);  // permutohedral.h:480
  return slice_dot_tiled(L, res, Cs, denom, other, other_stride, out, out_stride, coef, st, J);
}

int lattice_apply(Lattice* L, const float* src, int64_t src_stride, float* dst, int64_t dst_stride,
                  int C, const float* sub, int64_t sub_stride, const float* add, int64_t add_stride,
                  uint32_t flags, cudaStream_t st) {
  if (C <= 0 || src == nullptr || dst == nullptr || src_stride < C || dst_stride < C ||
      (sub != nullptr && sub_stride < C) || (add != nullptr && add_stride < C)) {
    set_last_error("pl_lattice_apply: bad arguments (C=%d)", C);
    return PL_ERR_INVALID;
  }
  const bool vec4 = (C % 4 == 0) && (src_stride % 4 == 0) && (dst_stride % 4 == 0) &&
                    aligned16(src) && aligned16(dst) &&
                    (sub == nullptr || (sub_stride % 4 == 0 && aligned16(sub))) &&
                    (add == nullptr || (add_stride % 4 == 0 && aligned16(add)));
  // channel chunk: both ping-pong buffers together stay under ~4 GiB
  const int64_t rows = L->M + 1;
  int64_t max_c = (int64_t)((2ull << 30) / (sizeof(float) * (uint64_t)rows));
  if (max_c < 4) max_c = 4;
  if (vec4) max_c &= ~(int64_t)3;
  const int chunk = (int)(C < max_c ? C : max_c);

  PL_TRY(lattice_workspace(L, chunk, st));

  for (int c0 = 0; c0 < C; c0 += chunk) {
    const int Cc = (C - c0) < chunk ? (C - c0) : chunk;
    const uint64_t items = (uint64_t)(L->n > L->M ? L->n : L->M) * (uint64_t)Cc;
    const bool small = items < 0x7fffffffull;
    const float* s = src + c0;
    float* o = dst + c0;
    const float* sb = sub ? sub + c0 : nullptr;
    const float* ad = add ? add + c0 : nullptr;
    int r;
    if (vec4 && small)
      r = apply_chunk<4, uint32_t>(L, s, src_stride, o, dst_stride, Cc, sb, sub_stride, ad,
                                   add_stride, flags, st);
    else if (vec4)
      r = apply_chunk<4, uint64_t>(L, s, src_stride, o, dst_stride, Cc, sb, sub_stride, ad,
                                   add_stride, flags, st);
    else if (small)
      r = apply_chunk<1, uint32_t>(L, s, src_stride, o, dst_stride, Cc, sb, sub_stride, ad,
                                   add_stride, flags, st);
    else
      r = apply_chunk<1, uint64_t>(L, s, src_stride, o, dst_stride, Cc, sb, sub_stride, ad,
                                   add_stride, flags, st);
    PL_TRY(r);
  }
  return PL_OK;
}

}  // namespace pl
