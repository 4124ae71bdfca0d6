// pl_readout.cu -- last slice of the mean-field loop fused with the depth readout (sm_100a).
//
// Inference wants a depth map, not the [n][L] energies: crf/mb_stereo_crf.py:62-66
// (`logits2average_depth`: sum_l softmax(logits)_l * label_l with logits = -E) and the
// notebooks' argmax read-out (Experiments/trainableDenseCRF.ipynb cell 8).  readout_tiled_kernel
// does the final  E = E0 + slice(blurred lattice) / (1 + 2^-d) - P  (crf_module.py:101,103) per
// pixel in registers and reduces it on the spot, so the last E (4 n L bytes written and read
// again by torch's softmax) never reaches HBM: reads 8 n L + lists, writes 4 n (+ 4 n).
//
// softargmax_rows_kernel is the same reduction over energies already in HBM (any L): the
// fall-back behind pl_meanfield_depth for label counts the tiled kernel does not take, and
// the pixel-major form of logits2average_depth.
#include <math.h>

#include "pl_common.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;
constexpr int kThreads = 256;
constexpr int kStageBytes = 64 * 1024;

struct Red {
  float m, s, t;   // running max of -e, sum of exp, sum of exp * label
  float best;      // smallest energy and its label index (first on ties)
  int arg;
};

__device__ __forceinline__ void red_init(Red& r) {
  r.m = -INFINITY; r.s = 0.f; r.t = 0.f; r.best = INFINITY; r.arg = 0x7fffffff;
}
// online softmax update with one element
__device__ __forceinline__ void red_push(Red& r, float e, float label, int idx) {
  const float x = -e;
  if (x > r.m) {
    const float c = __expf(r.m - x);
    r.s = r.s * c + 1.f;
    r.t = r.t * c + label;
    r.m = x;
  } else {
    const float p = __expf(x - r.m);
    r.s += p;
    r.t = fmaf(p, label, r.t);
  }
  if (e < r.best || (e == r.best && idx < r.arg)) { r.best = e; r.arg = idx; }
}
__device__ __forceinline__ void red_merge(Red& a, float m, float s, float t, float best, int arg) {
  const float M = fmaxf(a.m, m);
  const float ca = (a.m == -INFINITY) ? 0.f : __expf(a.m - M);
  const float cb = (m == -INFINITY) ? 0.f : __expf(m - M);
  a.s = a.s * ca + s * cb;
  a.t = a.t * ca + t * cb;
  a.m = M;
  if (best < a.best || (best == a.best && arg < a.arg)) { a.best = best; a.arg = arg; }
}
__device__ __forceinline__ void red_group(Red& r, int G) {
  for (int o = G >> 1; o > 0; o >>= 1) {
    const float m = __shfl_xor_sync(0xffffffffu, r.m, o, G);
    const float s = __shfl_xor_sync(0xffffffffu, r.s, o, G);
    const float t = __shfl_xor_sync(0xffffffffu, r.t, o, G);
    const float b = __shfl_xor_sync(0xffffffffu, r.best, o, G);
    const int a = __shfl_xor_sync(0xffffffffu, r.arg, o, G);
    red_merge(r, m, s, t, b, a);
  }
}

// One CTA per tile of 128 pixels; G lanes per pixel (power of two, <= 32), each lane owns the
// float4 columns lane, lane + G, ... of the pixel's L labels.
__global__ void __launch_bounds__(kThreads, 3)     // 80 registers: 3 CTAs per SM beat 4 with spills and 2 with 123
readout_tiled_kernel(const float* __restrict__ val, int L, int64_t n, int d1,
                     const uint16_t* __restrict__ loc, const float* __restrict__ bary,
                     const int32_t* __restrict__ vid, const int2* __restrict__ useg,
                     const int* __restrict__ tile_off, const int* __restrict__ tile_U,
                     float inv_denom, const float* __restrict__ p, const float* __restrict__ e0,
                     const float* __restrict__ labels, float* __restrict__ depth,
                     int* __restrict__ label_out, int G, int stage_bytes) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int TE = kTile * d1;
  float* s_w = reinterpret_cast<float*>(smem_raw);          // [TE]
  int* s_idx = reinterpret_cast<int*>(s_w + TE);            // [TE]
  float* s_lab = reinterpret_cast<float*>(s_idx + TE);      // [L]
  float* s_val = s_lab + ((L + 3) & ~3);                    // [U][L] when staged

  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);
  const int U = tile_U[tile];
  const int V = L >> 2;
  const bool staged = (size_t)U * L * 4 <= (size_t)stage_bytes;

  // the tile's records first: they belong to the lattice, not to the kernel in front of this one
  const int valid = rows * d1;
  for (int i = tid; i < valid; i += kThreads) {
    s_w[i] = bary[tile * TE + i] * inv_denom;
    s_idx[i] = staged ? (int)loc[tile * TE + i] : vid[tile * TE + i];
  }
  for (int i = tid; i < L; i += kThreads) s_lab[i] = labels ? labels[i] : (float)i;
  pdl_wait();      // the blurred lattice, P and E0 from here on (programmatic dependent launch)
  if (staged) {
    const int2* seg = useg + tile_off[tile];
    for (int i = tid; i < U * V; i += kThreads) {
      const int u = i / V, v = i - u * V;
      *reinterpret_cast<float4*>(s_val + u * L + 4 * v) =
          *reinterpret_cast<const float4*>(val + (int64_t)seg[u].x * L + 4 * v);
    }
  }
  __syncthreads();

  const int g = tid / G, l = tid % G, groups = kThreads / G;
  if (V <= 4 * G) {
    // A lane owns at most four float4 of its pixel's energies: keep them in registers and reduce in
    // two passes (group minimum, then one exp per element and two sums) -- the online form below costs
    // an exp and a branch per element plus five shuffles and two exps per merge step.
    for (int r = g; r < rows; r += groups) {
      const int64_t px = p0 + r;
      // P and E0 of this lane's columns first: their HBM latency runs under the gathers below
      float4 pp[4], ee[4];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int v = l + k * G;
        if (v < V) {
          pp[k] = *reinterpret_cast<const float4*>(p + px * L + 4 * v);
          ee[k] = *reinterpret_cast<const float4*>(e0 + px * L + 4 * v);
        }
      }
      // vertex-major: weight and row index of a vertex are read once for the lane's (up to) four columns;
      // per (pixel, label) the sum still runs over the d+1 vertices in entry order
      float4 acc[4];
#pragma unroll
      for (int k = 0; k < 4; k++) acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = 0; j < d1; j++) {
        const float w = s_w[r * d1 + j];
        const int ix = s_idx[r * d1 + j];
        const float* row = staged ? s_val + ix * L + 4 * l : val + (int64_t)ix * L + 4 * l;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          if (l + k * G < V) {
            const float4 x = *reinterpret_cast<const float4*>(row + 4 * k * G);
            acc[k].x = __fmaf_rn(w, x.x, acc[k].x);
            acc[k].y = __fmaf_rn(w, x.y, acc[k].y);
            acc[k].z = __fmaf_rn(w, x.z, acc[k].z);
            acc[k].w = __fmaf_rn(w, x.w, acc[k].w);
          }
        }
      }
      float en[16];
      float best = INFINITY;
      int arg = 0x7fffffff;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int v = l + k * G;
        if (v < V) {
          en[4 * k] = __fadd_rn(ee[k].x, __fsub_rn(acc[k].x, pp[k].x));
          en[4 * k + 1] = __fadd_rn(ee[k].y, __fsub_rn(acc[k].y, pp[k].y));
          en[4 * k + 2] = __fadd_rn(ee[k].z, __fsub_rn(acc[k].z, pp[k].z));
          en[4 * k + 3] = __fadd_rn(ee[k].w, __fsub_rn(acc[k].w, pp[k].w));
#pragma unroll
          for (int c = 0; c < 4; c++)     // ascending label index: the first minimum wins
            if (en[4 * k + c] < best) { best = en[4 * k + c]; arg = 4 * v + c; }
        } else {
          en[4 * k] = en[4 * k + 1] = en[4 * k + 2] = en[4 * k + 3] = INFINITY;
        }
      }
      for (int o = G >> 1; o > 0; o >>= 1) {
        const float b = __shfl_xor_sync(0xffffffffu, best, o, G);
        const int a = __shfl_xor_sync(0xffffffffu, arg, o, G);
        if (b < best || (b == best && a < arg)) { best = b; arg = a; }
      }
      float sum = 0.f, tsum = 0.f;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int v = l + k * G;
        if (v < V) {
          const float4 lb = *reinterpret_cast<const float4*>(s_lab + 4 * v);
          const float q0 = __expf(best - en[4 * k]), q1 = __expf(best - en[4 * k + 1]);
          const float q2 = __expf(best - en[4 * k + 2]), q3 = __expf(best - en[4 * k + 3]);
          sum += (q0 + q1) + (q2 + q3);
          tsum = fmaf(q0, lb.x, fmaf(q1, lb.y, fmaf(q2, lb.z, fmaf(q3, lb.w, tsum))));
        }
      }
      for (int o = G >> 1; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o, G);
        tsum += __shfl_xor_sync(0xffffffffu, tsum, o, G);
      }
      if (l == 0) {
        if (depth) depth[px] = tsum / sum;
        if (label_out) label_out[px] = arg;
      }
    }
    return;
  }
  for (int r = g; r < rows; r += groups) {
    const int64_t px = p0 + r;
    Red red;
    red_init(red);
    for (int v = l; v < V; v += G) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = 0; j < d1; j++) {
        const float w = s_w[r * d1 + j];
        const int ix = s_idx[r * d1 + j];
        const float4 x = staged ? *reinterpret_cast<const float4*>(s_val + ix * L + 4 * v)
                                : *reinterpret_cast<const float4*>(val + (int64_t)ix * L + 4 * v);
        acc.x = __fmaf_rn(w, x.x, acc.x);
        acc.y = __fmaf_rn(w, x.y, acc.y);
        acc.z = __fmaf_rn(w, x.z, acc.z);
        acc.w = __fmaf_rn(w, x.w, acc.w);
      }
      const float4 pp = *reinterpret_cast<const float4*>(p + px * L + 4 * v);
      const float4 ee = *reinterpret_cast<const float4*>(e0 + px * L + 4 * v);
      const float4 lb = *reinterpret_cast<const float4*>(s_lab + 4 * v);
      red_push(red, __fadd_rn(ee.x, __fsub_rn(acc.x, pp.x)), lb.x, 4 * v);
      red_push(red, __fadd_rn(ee.y, __fsub_rn(acc.y, pp.y)), lb.y, 4 * v + 1);
      red_push(red, __fadd_rn(ee.z, __fsub_rn(acc.z, pp.z)), lb.z, 4 * v + 2);
      red_push(red, __fadd_rn(ee.w, __fsub_rn(acc.w, pp.w)), lb.w, 4 * v + 3);
    }
    red_group(red, G);
    if (l == 0) {
      if (depth) depth[px] = red.t / red.s;
      if (label_out) label_out[px] = red.arg;
    }
  }
}

// energies already in HBM: one warp per row, any L
__global__ void __launch_bounds__(kThreads)
softargmax_rows_kernel(const float* __restrict__ e, int64_t n, int L,
                       const float* __restrict__ labels, float sign, float* __restrict__ depth,
                       int* __restrict__ label_out) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (kThreads / 32);
  for (int64_t row = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); row < n; row += warps) {
    Red red;
    red_init(red);
    for (int i = lane; i < L; i += 32)
      red_push(red, sign * e[row * L + i], labels ? labels[i] : (float)i, i);
    red_group(red, 32);
    if (lane == 0) {
      if (depth) depth[row] = red.t / red.s;
      if (label_out) label_out[row] = red.arg;
    }
  }
}

}  // namespace

bool readout_tiled_supported(const Lattice* L, int labels) {
  return L != nullptr && L->ntiles > 0 && L->loc != nullptr && (labels & 3) == 0 && labels >= 4;
}

// depth / label_out from the blurred lattice `val` of the last iteration, P and E0
int readout_tiled(Lattice* L, const float* val, int labels, float denom, const float* p,
                  const float* e0, const float* label_values, float* depth, int* label_out,
                  cudaStream_t st) {
  const int d1 = L->d + 1;
  const int TE = kTile * d1;
  // lanes per pixel: up to four float4 per lane, the most the register path holds (L = 64: 4 lanes, two
  // shuffle rounds per reduction, the entry weights / indices amortised over 16 labels; measured
  // 123 us with 16 lanes, 100 with 8, 89 with 4 at 740 x 500 x 64)
  int G = 1;
  while (G < 32 && G < labels / 16) G <<= 1;
  // staging sized by the lattice's largest tile (at most kStageBytes): the kernel is latency bound, and
  // a 740x500 default-scale lattice needs 32 KB, not 64 -- 5 CTAs per SM instead of 3
  if (L->max_U < 0 && L->max_U_ev != nullptr) {
    PL_CUDA(cudaEventSynchronize(L->max_U_ev));
    L->max_U = L->h_max_U[1];
    L->mean_U = L->ntiles > 0 ? (int)((uint32_t)L->h_max_U[0] / (uint64_t)L->ntiles) : 0;
  }
  size_t stage = kStageBytes;
  if (L->max_U > 0 && (size_t)L->max_U * labels * 4 < stage) stage = ((size_t)L->max_U * labels * 4 + 1023) & ~(size_t)1023;
  const size_t smem = sizeof(float) * (size_t)(2 * TE + ((labels + 3) & ~3)) + stage;
  static PerDeviceOnce attr_done;
  if (attr_done.need()) {
    PL_CUDA(cudaFuncSetAttribute(readout_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 100 * 1024));
  }
  if (smem > 100 * 1024) {
    set_last_error("readout_tiled: d + 1 = %d too large", d1);
    return PL_ERR_INVALID;
  }
  PL_LAUNCH_PDL(readout_tiled_kernel, (unsigned)L->ntiles, kThreads, smem, st, val, labels, L->n, d1,
            L->loc, L->bary, L->vid, L->useg, L->tile_off, L->tile_U, 1.0f / denom, p, e0,
            label_values, depth, label_out, G, (int)stage);
  return PL_OK;
}

// sign = +1: rows are energies E (soft-argmax of softmax(-E));  sign = -1: rows are logits
int softargmax_rows(const float* e, int64_t n, int labels, const float* label_values, float sign,
                    float* depth, int* label_out, cudaStream_t st) {
  if (n <= 0) return PL_OK;
  const unsigned grid = grid_for((uint64_t)n * 32, kThreads, 8);
  PL_LAUNCH(softargmax_rows_kernel, grid, kThreads, 0, st, e, n, labels, label_values, sign, depth,
            label_out);
  return PL_OK;
}

}  // namespace pl
