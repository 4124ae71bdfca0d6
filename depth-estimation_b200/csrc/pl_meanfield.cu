// pl_meanfield.cu -- per-pixel pieces of the mean-field loop (sm_100a).
//
// crf/crf_module.py:93-104 (CRFasRNN.forward) per iteration does
//     Q = softmax(-E, dim=1);  P = Mu(Q);  E = E0 + W(P, refs)
// with Mu = charb.forward (:71-75): a 1x1 conv2d with the L x L matrix
// Mu[l][k] = sqrt(gamma^2 + (label_l - label_k)^2) - gamma, times exp(s).
// Here everything is pixel-major [n][L] (labels contiguous), so the softmax is
// a row reduction and the compatibility transform is a row-times-matrix
// product; both are fused into one kernel so Q never goes to HBM unless asked.
// The E update is fused into the slice kernel (pl_filter.cu).
//
// softmax_compat_kernel: a CTA owns TP = 32 consecutive pixels.
//   phase 1: one warp per pixel row: row max, exp, sum -> Q in shared memory
//   phase 2: thread (lg, pg) owns label l = lg (+ k*LB) and PPT pixels;
//            per 4 k: 4 coalesced loads of muT[k][l] (L1/L2 resident, the
//            matrix is at most L*L*4 = 256 KB) and PPT broadcast LDS.128 of Q.
// This is the CUDA-core path, fp32-exact summation order (k ascending); it is
// HBM-bound up to L ~ 64.  (A tcgen05 path for L >= 128 is future work and
// would slot in behind the same entry point.)
#include <math.h>

#include "pl_common.cuh"

namespace pl {

namespace {

constexpr int kTP = 32;  // pixels per CTA

template <int PPT>
__global__ void __launch_bounds__(256)
softmax_compat_kernel(const float* __restrict__ e, const float* __restrict__ muT,
                      float* __restrict__ pout, float* __restrict__ qout, int64_t n, int L, int L4,
                      float scale) {
  extern __shared__ __align__(16) float qs[];  // [kTP][L4], zero padded to L4 = roundup(L, 4)
  constexpr int LB = 256 / (kTP / PPT);        // label lanes per pixel group
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t p0 = (int64_t)blockIdx.x * kTP;

  // phase 1: softmax(-E) per row (F.softmax(-E, dim=1), crf_module.py:99,102)
  for (int pp = wid; pp < kTP; pp += 8) {
    const int64_t p = p0 + pp;
    float* q = qs + pp * L4;
    if (p < n) {
      const float* row = e + p * L;
      float m = -INFINITY;
      for (int l = lane; l < L; l += 32) m = fmaxf(m, -row[l]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      float s = 0.f;
      for (int l = lane; l < L; l += 32) {
        const float x = expf(-row[l] - m);
        q[l] = x;
        s += x;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float inv = 1.0f / s;
      for (int l = lane; l < L4; l += 32) {
        const float x = l < L ? q[l] * inv : 0.f;
        q[l] = x;
        if (qout != nullptr && l < L) qout[p * L + l] = x;
      }
    } else {
      for (int l = lane; l < L4; l += 32) q[l] = 0.f;
    }
  }
  __syncthreads();

  // phase 2: P[p][l] = scale * sum_k muT[k][l] * Q[p][k]
  const int lg = tid % LB, pg = tid / LB;
  const float* qbase = qs + (pg * PPT) * L4;
  for (int l = lg; l < L; l += LB) {
    float acc[PPT];
#pragma unroll
    for (int i = 0; i < PPT; i++) acc[i] = 0.f;
    for (int k = 0; k < L4; k += 4) {
      const float m0 = muT[(int64_t)k * L + l];
      const float m1 = (k + 1 < L) ? muT[(int64_t)(k + 1) * L + l] : 0.f;
      const float m2 = (k + 2 < L) ? muT[(int64_t)(k + 2) * L + l] : 0.f;
      const float m3 = (k + 3 < L) ? muT[(int64_t)(k + 3) * L + l] : 0.f;
#pragma unroll
      for (int i = 0; i < PPT; i++) {
        const float4 q = *reinterpret_cast<const float4*>(qbase + i * L4 + k);
        acc[i] = fmaf(q.x, m0, acc[i]);
        acc[i] = fmaf(q.y, m1, acc[i]);
        acc[i] = fmaf(q.z, m2, acc[i]);
        acc[i] = fmaf(q.w, m3, acc[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < PPT; i++) {
      const int64_t p = p0 + pg * PPT + i;
      if (p < n) pout[p * L + l] = acc[i] * scale;
    }
  }
}

// dst[b][p][c] = alpha * src[b][c][p]: 32x32 tiles through padded shared memory
__global__ void __launch_bounds__(256)
transpose_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int64_t cols,
                 float alpha) {
  // src is [batch][rows][cols], dst is [batch][cols][rows]
  __shared__ float tile[32][33];
  const int64_t b = blockIdx.y;
  const float* s = src + b * (int64_t)rows * cols;
  float* d = dst + b * (int64_t)rows * cols;
  // flat tile index (rows can be millions of pixels: too many for grid.y)
  const unsigned tiles_c = (unsigned)((cols + 31) / 32);
  const int64_t c0 = (int64_t)(blockIdx.x % tiles_c) * 32;
  const int r0 = (int)(blockIdx.x / tiles_c) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int r = r0 + ty + i;
    const int64_t c = c0 + tx;
    if (r < rows && c < cols) tile[ty + i][tx] = s[(int64_t)r * cols + c];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int64_t c = c0 + ty + i;
    const int r = r0 + tx;
    if (r < rows && c < cols) d[c * rows + r] = alpha * tile[tx][ty + i];
  }
}

// Backward of Q = softmax(-E) chained with the scale of the compatibility transform, fused with the
// accumulation of the unary gradient (crf_module.py:99-102 differentiated):
//   dE[p][k] = -scale * q[p][k] * (dq[p][k] - sum_k' dq[p][k'] q[p][k']);   acc[p][k] += dE[p][k]
// LPR lanes per row, float4 per lane and step.  `de` may alias `dq`.
template <int LPR>
__global__ void __launch_bounds__(256)
softmax_backward_kernel(const float* dq, const float* __restrict__ q, float* de, float* acc,
                        int64_t n, int L, float scale) {
  constexpr int RPW = 32 / LPR;        // rows per warp
  constexpr int MAXC = 4;              // float4 steps per lane: L <= 16 * LPR
  const int lane = threadIdx.x & 31, l = lane % LPR;
  const int V = L >> 2;
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t r0 = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW; r0 < n;
       r0 += warps * RPW) {
    const int64_t row = r0 + lane / LPR;
    const bool ok = row < n;
    float4 a[MAXC], b[MAXC];
    float inner = 0.f;
#pragma unroll
    for (int c = 0; c < MAXC; c++) {
      const int v = l + c * LPR;
      a[c] = make_float4(0.f, 0.f, 0.f, 0.f);
      b[c] = a[c];
      if (ok && v < V) {
        a[c] = *reinterpret_cast<const float4*>(dq + row * L + 4 * v);
        b[c] = *reinterpret_cast<const float4*>(q + row * L + 4 * v);
      }
      inner += a[c].x * b[c].x + a[c].y * b[c].y + a[c].z * b[c].z + a[c].w * b[c].w;
    }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) inner += __shfl_xor_sync(0xffffffffu, inner, o, LPR);
#pragma unroll
    for (int c = 0; c < MAXC; c++) {
      const int v = l + c * LPR;
      if (ok && v < V) {
        float4 g;
        g.x = -scale * b[c].x * (a[c].x - inner);
        g.y = -scale * b[c].y * (a[c].y - inner);
        g.z = -scale * b[c].z * (a[c].z - inner);
        g.w = -scale * b[c].w * (a[c].w - inner);
        *reinterpret_cast<float4*>(de + row * L + 4 * v) = g;
        if (acc != nullptr) {
          float4 t = *reinterpret_cast<const float4*>(acc + row * L + 4 * v);
          t.x += g.x; t.y += g.y; t.z += g.z; t.w += g.w;
          *reinterpret_cast<float4*>(acc + row * L + 4 * v) = t;
        }
      }
    }
  }
}

// any L / alignment: one warp per row, scalars
__global__ void __launch_bounds__(256)
softmax_backward_scalar_kernel(const float* dq, const float* __restrict__ q, float* de, float* acc,
                               int64_t n, int L, float scale) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < n; row += warps) {
    float inner = 0.f;
    for (int k = lane; k < L; k += 32) inner += dq[row * L + k] * q[row * L + k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) inner += __shfl_xor_sync(0xffffffffu, inner, o);
    for (int k = lane; k < L; k += 32) {
      const float g = -scale * q[row * L + k] * (dq[row * L + k] - inner);
      de[row * L + k] = g;
      if (acc != nullptr) acc[row * L + k] += g;
    }
  }
}

}  // namespace

int softmax_compat(const float* e, const float* muT, float* p, float* q_out, int64_t n, int L,
                   float scale, cudaStream_t st) {
  if (e == nullptr || muT == nullptr || p == nullptr || n <= 0 || L <= 0) {
    set_last_error("pl_softmax_compat: bad arguments");
    return PL_ERR_INVALID;
  }
  const int L4 = (L + 3) & ~3;
  const size_t smem = sizeof(float) * (size_t)kTP * L4;
  if (smem > 200 * 1024) {
    set_last_error("pl_softmax_compat: L = %d too large for the shared-memory tile", L);
    return PL_ERR_INVALID;
  }
  const unsigned grid = ceil_div_u((uint64_t)n, kTP);
#define PL_SC(PPT)                                                                               \
  do {                                                                                           \
    if (smem > 48 * 1024)                                                                        \
      PL_CUDA(cudaFuncSetAttribute(softmax_compat_kernel<PPT>,                                   \
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
    PL_LAUNCH((softmax_compat_kernel<PPT>), grid, 256, smem, st, e, muT, p, q_out, n, L, L4,     \
              scale);                                                                            \
  } while (0)
  if (L <= 32) PL_SC(4);        // 32 label lanes x 8 pixel groups
  else if (L <= 64) PL_SC(8);   // 64 x 4
  else if (L <= 128) PL_SC(16); // 128 x 2
  else PL_SC(32);               // 256 x 1
#undef PL_SC
  return PL_OK;
}

static int transpose(const float* src, float* dst, int64_t batch, int64_t rows, int64_t cols,
                     float alpha, cudaStream_t st) {
  const uint64_t tiles = (uint64_t)((rows + 31) / 32) * (uint64_t)((cols + 31) / 32);
  if (src == nullptr || dst == nullptr || batch <= 0 || rows <= 0 || cols <= 0 || batch > 65535 ||
      rows > 0x7fffffff || tiles > 0x7fffffffull) {
    set_last_error("layout transpose: bad arguments");
    return PL_ERR_INVALID;
  }
  dim3 grid((unsigned)tiles, (unsigned)batch, 1);
  PL_LAUNCH(transpose_kernel, grid, 256, 0, st, src, dst, (int)rows, cols, alpha);
  return PL_OK;
}

int softmax_backward(const float* dq, const float* q, float* de, float* acc, int64_t n, int L, float scale,
                     cudaStream_t st) {
  if (n == 0) return PL_OK;
  const bool vec = (L % 4 == 0) && L <= 512 &&
                   ((((uintptr_t)dq | (uintptr_t)q | (uintptr_t)de | (uintptr_t)acc) & 15u) == 0);
  if (!vec) {
    PL_LAUNCH(softmax_backward_scalar_kernel, grid_for((uint64_t)n * 32, 256, 8), 256, 0, st, dq, q, de, acc,
              n, L, scale);
    return PL_OK;
  }
  const int V = L / 4;
  if (V <= 4) {
    PL_LAUNCH(softmax_backward_kernel<4>, grid_for((uint64_t)n * 4, 256, 8), 256, 0, st, dq, q, de, acc, n, L, scale);
  } else if (V <= 8) {
    PL_LAUNCH(softmax_backward_kernel<8>, grid_for((uint64_t)n * 8, 256, 8), 256, 0, st, dq, q, de, acc, n, L, scale);
  } else if (V <= 16) {
    PL_LAUNCH(softmax_backward_kernel<16>, grid_for((uint64_t)n * 16, 256, 8), 256, 0, st, dq, q, de, acc, n, L, scale);
  } else {
    PL_LAUNCH(softmax_backward_kernel<32>, grid_for((uint64_t)n * 32, 256, 8), 256, 0, st, dq, q, de, acc, n, L, scale);
  }
  return PL_OK;
}

int nchw_to_nlc(const float* src, float* dst, int64_t batch, int C, int64_t n, float alpha,
                cudaStream_t st) {
  return transpose(src, dst, batch, C, n, alpha, st);  // [b][C][n] -> [b][n][C]
}

int nlc_to_nchw(const float* src, float* dst, int64_t batch, int C, int64_t n, float alpha,
                cudaStream_t st) {
  return transpose(src, dst, batch, n, C, alpha, st);  // [b][n][C] -> [b][C][n]
}

}  // namespace pl
