// pl_fused.cu -- one kernel per mean-field iteration (sm_100a), L <= 64 labels.
//
// crf/crf_module.py:100-102, pixel-major, per tile of PL_TILE pixels:
//     E  = E0 + slice(blurred lattice) - P_prev        (skipped when FIRST: E = E0)
//     Q  = softmax(-E)                                  (row reduction inside a 16-lane group)
//     P  = scale * Q Mu^T                               (128 x L x L contraction from shared memory)
//     P -> HBM (next iteration's "- P" term) and splat(P) into the zeroed lattice buffer
// so one iteration moves 3 n L floats through HBM (read E0, read P_prev, write P)
// instead of the 6 n L of the slice / softmax_compat / splat sequence, and E / Q
// never leave the SM.  Lattice rows the tile needs are staged in shared memory
// once (tile-local vertex list from pl_tiles.cu); the splat is the segment sum
// of pl_tiles.cu with one red.global.add.v4 per (tile, vertex, float4).
#include <math.h>

#include "pl_common.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;   // 128 pixels
constexpr int kThreads = 256;
constexpr int kLD = 64;          // shared-memory row pitch in floats (L <= 64)
constexpr int kGroup = 16;       // lanes per pixel row
constexpr int kGroups = kThreads / kGroup;
constexpr int kStageRows = 160;  // lattice rows staged per tile (>= kTile so P can alias it)
constexpr int kRowsPerGroup = kTile / kGroups;  // 8

__device__ __forceinline__ void red_add_v4(float* addr, float4 x) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x.x), "f"(x.y),
               "f"(x.z), "f"(x.w)
               : "memory");
}

__device__ __forceinline__ float group_max(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o, 16));
  return v;
}
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o, 16);
  return v;
}

template <bool FIRST>
__global__ void __launch_bounds__(kThreads, 2)
mf_fused_kernel(const float* __restrict__ val_in, float* __restrict__ val_out,
                const float* __restrict__ e0, const float* __restrict__ p_prev,
                float* __restrict__ p_next, const float* __restrict__ muT, float scale, int L,
                int64_t n, int d1, float inv_denom, const uint16_t* __restrict__ loc,
                const float* __restrict__ bary, const int32_t* __restrict__ vid,
                const uint8_t* __restrict__ ent_pix, const float* __restrict__ ent_w,
                const int2* __restrict__ useg, const int* __restrict__ tile_off,
                const int* __restrict__ tile_U) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_val = reinterpret_cast<float*>(smem_raw);       // [kStageRows][kLD]; later P [kTile][kLD]
  float* s_q = s_val + kStageRows * kLD;                   // [kTile][kLD]  E then Q
  float* s_mu = s_q + kTile * kLD;                         // [L][kLD]      Mu^T (k-major)
  float* s_w = s_mu + 64 * kLD;                            // [TE] slice weights (entry order)
  int* s_idx = reinterpret_cast<int*>(s_w + kTile * d1);   // [TE] slice row index (local or global)
  float* s_sw = reinterpret_cast<float*>(s_idx + kTile * d1);        // [TE] splat weights (sorted)
  uint8_t* s_spix = reinterpret_cast<uint8_t*>(s_sw + kTile * d1);   // [TE] splat pixel (sorted)

  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  const int V = L >> 2;  // float4 lanes per row, <= 16
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);
  const int valid = rows * d1;
  const int U = tile_U[tile];
  const int2* seg = useg + tile_off[tile];
  const bool staged = U <= kStageRows;
  const int g = tid / kGroup, l = tid % kGroup;

  // ---- stage: Mu^T, entry lists, lattice rows ------------------------------------------
  for (int i = tid; i < L * V; i += kThreads) {
    const int k = i / V, v = i - k * V;
    *reinterpret_cast<float4*>(s_mu + k * kLD + 4 * v) =
        *reinterpret_cast<const float4*>(muT + k * L + 4 * v);
  }
  for (int i = tid; i < valid; i += kThreads) {
    s_sw[i] = ent_w[tile * TE + i];
    s_spix[i] = ent_pix[tile * TE + i];
    if (!FIRST) {
      s_w[i] = bary[tile * TE + i] * inv_denom;
      s_idx[i] = staged ? (int)loc[tile * TE + i] : vid[tile * TE + i];
    }
  }
  if (!FIRST && staged) {
    for (int i = tid; i < U * V; i += kThreads) {
      const int u = i / V, v = i - u * V;
      *reinterpret_cast<float4*>(s_val + u * kLD + 4 * v) =
          *reinterpret_cast<const float4*>(val_in + (int64_t)seg[u].x * L + 4 * v);
    }
  }
  // prefetch this thread's E0 / P_prev vectors (8 rows) while the staging loads fly
  float4 e_reg[kRowsPerGroup], pp_reg[kRowsPerGroup];
#pragma unroll
  for (int i = 0; i < kRowsPerGroup; i++) {
    const int r = g + i * kGroups;
    e_reg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    pp_reg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < rows && l < V) {
      e_reg[i] = *reinterpret_cast<const float4*>(e0 + (p0 + r) * L + 4 * l);
      if (!FIRST) pp_reg[i] = *reinterpret_cast<const float4*>(p_prev + (p0 + r) * L + 4 * l);
    }
  }
  __syncthreads();

  // ---- slice + energy + softmax -> Q in shared memory ------------------------------------
#pragma unroll
  for (int i = 0; i < kRowsPerGroup; i++) {
    const int r = g + i * kGroups;
    float4 e = e_reg[i];
    if (!FIRST && r < rows && l < V) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = 0; j < d1; j++) {
        const float w = s_w[r * d1 + j];
        const int ix = s_idx[r * d1 + j];
        const float4 x = staged
                             ? *reinterpret_cast<const float4*>(s_val + ix * kLD + 4 * l)
                             : *reinterpret_cast<const float4*>(val_in + (int64_t)ix * L + 4 * l);
        acc.x = __fmaf_rn(w, x.x, acc.x);
        acc.y = __fmaf_rn(w, x.y, acc.y);
        acc.z = __fmaf_rn(w, x.z, acc.z);
        acc.w = __fmaf_rn(w, x.w, acc.w);
      }
      // E = E0 + (filtered - P)   (crf_module.py:101, gaussian_matrix.py:352)
      e.x = __fadd_rn(e.x, __fsub_rn(acc.x, pp_reg[i].x));
      e.y = __fadd_rn(e.y, __fsub_rn(acc.y, pp_reg[i].y));
      e.z = __fadd_rn(e.z, __fsub_rn(acc.z, pp_reg[i].z));
      e.w = __fadd_rn(e.w, __fsub_rn(acc.w, pp_reg[i].w));
    }
    // softmax(-E) over the row: all 16 lanes of the group take part in the shuffles
    const bool live = (r < rows) && (l < V);
    float m = live ? fmaxf(fmaxf(-e.x, -e.y), fmaxf(-e.z, -e.w)) : -INFINITY;
    m = group_max(m);
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) {
      x.x = expf(-e.x - m);
      x.y = expf(-e.y - m);
      x.z = expf(-e.z - m);
      x.w = expf(-e.w - m);
    }
    const float s = group_sum(x.x + x.y + x.z + x.w);
    const float inv = live ? 1.0f / s : 0.f;
    x.x *= inv; x.y *= inv; x.z *= inv; x.w *= inv;
    if (r < kTile) *reinterpret_cast<float4*>(s_q + r * kLD + 4 * l) = x;  // zeros beyond rows / L
  }
  __syncthreads();

  // ---- P = scale * Q Mu^T : thread owns rows ty*8..+7, columns 4*tx..+3 ------------------
  {
    const int ty = tid >> 4, tx = tid & 15;
    float4 acc[kRowsPerGroup];
#pragma unroll
    for (int i = 0; i < kRowsPerGroup; i++) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* qrow = s_q + (ty * kRowsPerGroup) * kLD;
    for (int k = 0; k < L; k += 4) {
      const float4 m0 = *reinterpret_cast<const float4*>(s_mu + (k + 0) * kLD + 4 * tx);
      const float4 m1 = *reinterpret_cast<const float4*>(s_mu + (k + 1) * kLD + 4 * tx);
      const float4 m2 = *reinterpret_cast<const float4*>(s_mu + (k + 2) * kLD + 4 * tx);
      const float4 m3 = *reinterpret_cast<const float4*>(s_mu + (k + 3) * kLD + 4 * tx);
#pragma unroll
      for (int i = 0; i < kRowsPerGroup; i++) {
        const float4 q = *reinterpret_cast<const float4*>(qrow + i * kLD + k);
        acc[i].x = fmaf(q.x, m0.x, acc[i].x); acc[i].y = fmaf(q.x, m0.y, acc[i].y);
        acc[i].z = fmaf(q.x, m0.z, acc[i].z); acc[i].w = fmaf(q.x, m0.w, acc[i].w);
        acc[i].x = fmaf(q.y, m1.x, acc[i].x); acc[i].y = fmaf(q.y, m1.y, acc[i].y);
        acc[i].z = fmaf(q.y, m1.z, acc[i].z); acc[i].w = fmaf(q.y, m1.w, acc[i].w);
        acc[i].x = fmaf(q.z, m2.x, acc[i].x); acc[i].y = fmaf(q.z, m2.y, acc[i].y);
        acc[i].z = fmaf(q.z, m2.z, acc[i].z); acc[i].w = fmaf(q.z, m2.w, acc[i].w);
        acc[i].x = fmaf(q.w, m3.x, acc[i].x); acc[i].y = fmaf(q.w, m3.y, acc[i].y);
        acc[i].z = fmaf(q.w, m3.z, acc[i].z); acc[i].w = fmaf(q.w, m3.w, acc[i].w);
      }
    }
    // s_val is dead (all slice reads finished before the barrier above): reuse it for P
#pragma unroll
    for (int i = 0; i < kRowsPerGroup; i++) {
      const int r = ty * kRowsPerGroup + i;
      float4 p = acc[i];
      p.x *= scale; p.y *= scale; p.z *= scale; p.w *= scale;
      *reinterpret_cast<float4*>(s_val + r * kLD + 4 * tx) = p;
      if (r < rows && tx < V) *reinterpret_cast<float4*>(p_next + (p0 + r) * L + 4 * tx) = p;
    }
  }
  __syncthreads();

  // ---- splat P (segment sums per distinct vertex of the tile) ----------------------------
  if (l < V) {
    for (int u = g; u < U; u += kGroups) {
      const int2 sg = seg[u];
      const int st = sg.y & 0xFFFF, cnt = sg.y >> 16;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = st; k < st + cnt; k++) {
        const float w = s_sw[k];
        const float4 x = *reinterpret_cast<const float4*>(s_val + (int)s_spix[k] * kLD + 4 * l);
        acc.x = __fmaf_rn(w, x.x, acc.x);
        acc.y = __fmaf_rn(w, x.y, acc.y);
        acc.z = __fmaf_rn(w, x.z, acc.z);
        acc.w = __fmaf_rn(w, x.w, acc.w);
      }
      red_add_v4(val_out + (int64_t)sg.x * L + 4 * l, acc);
    }
  }
}

size_t fused_smem(int d1) {
  return sizeof(float) * (size_t)(kStageRows * kLD + kTile * kLD + 64 * kLD) +
         (size_t)kTile * d1 * (4 + 4 + 4 + 1) + 16;
}

}  // namespace

bool mf_fused_supported(const Lattice* L, int labels) {
  return labels >= 4 && labels <= 64 && (labels % 4) == 0 && L->ntiles > 0 &&
         fused_smem(L->d + 1) <= 110 * 1024;
}

// One iteration.  first: E = E0 (no slice, p_prev unused).  val_out must be zeroed.
int mf_fused_iteration(Lattice* L, bool first, const float* val_in, float* val_out, const float* e0,
                       const float* p_prev, float* p_next, const float* muT, float scale,
                       int labels, cudaStream_t st) {
  const int d1 = L->d + 1;
  const size_t smem = fused_smem(d1);
  static PerDeviceOnce attr_done;
  if (attr_done.need()) {
    PL_CUDA(cudaFuncSetAttribute(mf_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 110 * 1024));
    PL_CUDA(cudaFuncSetAttribute(mf_fused_kernel<false>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
  }
  const float inv_denom = 1.0f / (1 + powf(2, -L->d));
  if (first) {
    PL_LAUNCH((mf_fused_kernel<true>), (unsigned)L->ntiles, kThreads, smem, st, val_in, val_out, e0,
              p_prev, p_next, muT, scale, labels, L->n, d1, inv_denom, L->loc, L->bary, L->vid,
              L->ent_pix, L->ent_w, L->useg, L->tile_off, L->tile_U);
  } else {
    PL_LAUNCH((mf_fused_kernel<false>), (unsigned)L->ntiles, kThreads, smem, st, val_in, val_out,
              e0, p_prev, p_next, muT, scale, labels, L->n, d1, inv_denom, L->loc, L->bary, L->vid,
              L->ent_pix, L->ent_w, L->useg, L->tile_off, L->tile_U);
  }
  return PL_OK;
}

}  // namespace pl
