// pl_backward.cu -- one backward step of the mean-field loop behind the trainable CRF (sm_100a).
//
// What autograd derives for one iteration of CRFasRNN.forward (crf/crf_module.py:99-102 with
// charb.forward :71-75), after the filter's own backward has produced dP = K dE - dE
// ("Matrix is symmetric", crf/gaussian_matrix.py:446), pixel-major, L = 64 labels:
//
//     dQ  = dP Mu                          [n x L] = [n x L] [L x L]
//     dMu += dP^T Q                        [L x L], summed over all pixels (unscaled)
//     dE  = -s Q (dQ - rowsum(dQ Q))       softmax(-E) backward with the compatibility scale s
//     dE0 += dE                            running gradient of the unaries
//
// One pass over HBM: a tile of 128 pixels of dP and Q is staged in shared memory once and feeds
// both products (fp32 FMAs: these are gradients checked against the reference's own, no tensor-core
// rounding in the way), dQ never leaves the chip, dE overwrites dP in place.  Persistent CTAs keep
// their 64 x 64 share of dMu in registers over all their tiles and add it to HBM once at the end.
#include <math.h>

#include "pl_common.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;   // 128 pixels
constexpr int kThreads = 256;
constexpr int kL = 64;
constexpr int kPI = kL + 4;      // row pitch in floats (272 bytes: float4-aligned)

__device__ __forceinline__ void fma4x(float4& acc, float a, const float4 b) {
  acc.x = fmaf(a, b.x, acc.x);
  acc.y = fmaf(a, b.y, acc.y);
  acc.z = fmaf(a, b.z, acc.z);
  acc.w = fmaf(a, b.w, acc.w);
}

__global__ void __launch_bounds__(kThreads, 2)
mf_backward_step_kernel(float* __restrict__ dp, const float* __restrict__ q, const float* __restrict__ mu,
                        float* __restrict__ de0, float* __restrict__ dmu, int64_t n, float scale) {
  extern __shared__ __align__(16) float s_bw[];
  float* s_dp = s_bw;                       // [128][kPI]  dP tile, later the dQ tile
  float* s_q = s_dp + kTile * kPI;          // [128][kPI]
  float* s_mu = s_q + kTile * kPI;          // [64][64]    Mu[j][k]

  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t ntiles = (n + kTile - 1) / kTile;
  for (int i = tid; i < kL * kL / 4; i += kThreads)
    reinterpret_cast<float4*>(s_mu)[i] = reinterpret_cast<const float4*>(mu)[i];

  // dQ: 4 rows x 8 columns per thread;  dMu: 8 x 4 per thread, the two halves of the CTA take half
  // of the tile's pixels each (both add their share to HBM at the end)
  const int tx = tid & 7, ty = tid >> 3;
  const int kb = tid & 15, jb = (tid >> 4) & 7, half = tid >> 7;
  float4 m_acc[8];
#pragma unroll
  for (int i = 0; i < 8; i++) m_acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  const int g = tid >> 4, l = tid & 15;     // row-wise phases: 16 lanes per row, 16 rows per pass

  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t p0 = tile * kTile;
    const int rows = (int)min((int64_t)kTile, n - p0);
    __syncthreads();   // the previous tile's row phase is done with both buffers (and Mu is staged)
#pragma unroll
    for (int i = 0; i < kTile / 16; i++) {
      const int r = g + i * 16;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
      if (r < rows) {
        a = *reinterpret_cast<const float4*>(dp + (p0 + r) * kL + 4 * l);
        b = *reinterpret_cast<const float4*>(q + (p0 + r) * kL + 4 * l);
      }
      *reinterpret_cast<float4*>(s_dp + r * kPI + 4 * l) = a;
      *reinterpret_cast<float4*>(s_q + r * kPI + 4 * l) = b;
    }
    __syncthreads();

    // ---- dQ = dP Mu ---------------------------------------------------------------------------
    float4 d_acc[4][2];
#pragma unroll
    for (int i = 0; i < 4; i++) d_acc[i][0] = d_acc[i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int j = 0; j < kL; j += 4) {
      float4 a[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = *reinterpret_cast<const float4*>(s_dp + (4 * ty + i) * kPI + j);
#pragma unroll
      for (int jj = 0; jj < 4; jj++) {
        // float4 columns tx and tx + 8: the 8 lanes of a quarter warp read 128 contiguous bytes
        const float4 b0 = *reinterpret_cast<const float4*>(s_mu + (j + jj) * kL + 4 * tx);
        const float4 b1 = *reinterpret_cast<const float4*>(s_mu + (j + jj) * kL + 4 * tx + 32);
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const float av = jj == 0 ? a[i].x : (jj == 1 ? a[i].y : (jj == 2 ? a[i].z : a[i].w));
          fma4x(d_acc[i][0], av, b0);
          fma4x(d_acc[i][1], av, b1);
        }
      }
    }
    // ---- dMu += dP^T Q ------------------------------------------------------------------------
#pragma unroll 4
    for (int p = half * (kTile / 2); p < (half + 1) * (kTile / 2); p++) {
      const float4 a0 = *reinterpret_cast<const float4*>(s_dp + p * kPI + 8 * jb);
      const float4 a1 = *reinterpret_cast<const float4*>(s_dp + p * kPI + 8 * jb + 4);
      const float4 b = *reinterpret_cast<const float4*>(s_q + p * kPI + 4 * kb);
      fma4x(m_acc[0], a0.x, b);
      fma4x(m_acc[1], a0.y, b);
      fma4x(m_acc[2], a0.z, b);
      fma4x(m_acc[3], a0.w, b);
      fma4x(m_acc[4], a1.x, b);
      fma4x(m_acc[5], a1.y, b);
      fma4x(m_acc[6], a1.z, b);
      fma4x(m_acc[7], a1.w, b);
    }
    __syncthreads();   // everyone is done reading the dP tile
#pragma unroll
    for (int i = 0; i < 4; i++) {
      *reinterpret_cast<float4*>(s_dp + (4 * ty + i) * kPI + 4 * tx) = d_acc[i][0];
      *reinterpret_cast<float4*>(s_dp + (4 * ty + i) * kPI + 4 * tx + 32) = d_acc[i][1];
    }
    __syncthreads();

    // ---- dE = -s Q (dQ - rowsum(dQ Q)), in place over dP;  dE0 += dE ----------------------------
#pragma unroll 2
    for (int i = 0; i < kTile / 16; i++) {
      const int r = g + i * 16;
      const float4 d = *reinterpret_cast<const float4*>(s_dp + r * kPI + 4 * l);
      const float4 b = *reinterpret_cast<const float4*>(s_q + r * kPI + 4 * l);
      float inner = d.x * b.x + d.y * b.y + d.z * b.z + d.w * b.w;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) inner += __shfl_xor_sync(0xffffffffu, inner, o, 16);
      if (r < rows) {
        float4 e;
        e.x = -scale * b.x * (d.x - inner);
        e.y = -scale * b.y * (d.y - inner);
        e.z = -scale * b.z * (d.z - inner);
        e.w = -scale * b.w * (d.w - inner);
        *reinterpret_cast<float4*>(dp + (p0 + r) * kL + 4 * l) = e;
        if (de0 != nullptr) {
          float4 t = *reinterpret_cast<const float4*>(de0 + (p0 + r) * kL + 4 * l);
          t.x += e.x; t.y += e.y; t.z += e.z; t.w += e.w;
          *reinterpret_cast<float4*>(de0 + (p0 + r) * kL + 4 * l) = t;
        }
      }
    }
  }

  if (dmu != nullptr) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      float* o = dmu + (8 * jb + i) * kL + 4 * kb;
      atomicAdd(o + 0, m_acc[i].x);
      atomicAdd(o + 1, m_acc[i].y);
      atomicAdd(o + 2, m_acc[i].z);
      atomicAdd(o + 3, m_acc[i].w);
    }
  }
  (void)lane;
}

}  // namespace

bool mf_backward_step_supported(int L) { return L == kL; }

int mf_backward_step(float* dp, const float* q, const float* mu, float* de0, float* dmu, int64_t n,
                     int L, float scale, cudaStream_t st) {
  if (L != kL) {
    set_last_error("mf_backward_step: L = %d not supported (L = 64 only)", L);
    return PL_ERR_INVALID;
  }
  if (n == 0) return PL_OK;
  const size_t smem = sizeof(float) * (2 * (size_t)kTile * kPI + kL * kL);
  static PerDeviceOnce attr_done;
  if (attr_done.need()) {
    PL_CUDA(cudaFuncSetAttribute(mf_backward_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
  }
  const int64_t ntiles = (n + kTile - 1) / kTile;
  const int64_t cap = 2 * PL_SM_COUNT;
  PL_LAUNCH(mf_backward_step_kernel, (unsigned)(ntiles < cap ? ntiles : cap), kThreads, smem, st, dp, q, mu,
            de0, dmu, n, scale);
  return PL_OK;
}

}  // namespace pl
