// pl_cost.cu -- plane-sweep matching cost volume (the unaries the CRF reads), sm_100a.
//
// Replaces crf/depth.py:24-53 (`disparity_badness`): for disparity i in [0, L)
//     out[y][x][i] = sum_c crit(img1[y][x][c], img2[y][x - i][c])          (0 where x - i < 0:
//                    the right image is zero-padded on the left, depth.py:39)
//     cost[y][x][i] = sum over the ws x ws window of out, `reflect` boundary
//                    (scipy.ndimage.convolve default: d c b a | a b c d | d c b a)
// crit = AD |a - b|, SD (a - b)^2 or nprod -a*b (depth.py:18-23).  The reference computes in
// float64 on the host; here fp32 with every sum taken directly over its 9 terms (no running
// sums, so no cancellation): measured <= 2e-6 of the output scale against the float64 arithmetic.
//
// Output is pixel-major [h*w][L] -- the layout pl_meanfield consumes -- so the volume is born
// in HBM and the 4*n*L-byte host -> device copy of the unaries disappears.  HBM-bound on the
// output write (4 n L bytes); the two images (24 n bytes) stay in L2.
//
// CTA = 8 rows x 64 columns x 8 disparities: per-pixel costs of the haloed patch into shared
// memory, horizontal window sums, vertical window sums, one 32-byte sector per pixel written.
#include "pl_common.cuh"

namespace pl {

namespace {

constexpr int kTX = 64, kTY = 8, kLC = 8, kThreads = 256;
constexpr int kMaxR = 8;  // window radius supported (ws <= 17)

__device__ __forceinline__ int reflect(int i, int n) {
  // scipy 'reflect': (d c b a | a b c d | d c b a); windows are much smaller than the image,
  // but tiny images may need more than one fold
  while (i < 0 || i >= n) i = (i < 0) ? (-i - 1) : (2 * n - 1 - i);
  return i;
}

template <int CRIT>
__device__ __forceinline__ float crit(float a, float b) {
  if (CRIT == 0) return fabsf(a - b);
  if (CRIT == 1) return (a - b) * (a - b);
  return -a * b;
}

template <int CRIT>
__global__ void __launch_bounds__(kThreads)
cost_volume_kernel(const float* __restrict__ im1, const float* __restrict__ im2,
                   float* __restrict__ out, int h, int w, int ch, int L, int R) {
  extern __shared__ float sm[];
  const int PW = kTX + 2 * R, PH = kTY + 2 * R;
  float* s_d = sm;                       // [kLC][PH][PW]   per-pixel cost of the haloed patch
  float* s_h = sm + kLC * PH * PW;       // [kLC][PH][kTX]  horizontal window sums
  const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * kTY, l0 = blockIdx.z * kLC;
  const int tid = threadIdx.x;

  // per-pixel cost at reflected coordinates (the reflection acts on `out`, i.e. before the shift)
  for (int i = tid; i < PH * PW; i += kThreads) {
    const int py = i / PW, px = i - py * PW;
    const int y = reflect(y0 + py - R, h), x = reflect(x0 + px - R, w);
    const float* a = im1 + ((int64_t)y * w + x) * ch;
    float acc[kLC];
#pragma unroll
    for (int ll = 0; ll < kLC; ll++) acc[ll] = 0.f;
    for (int c = 0; c < ch; c++) {
      const float av = __ldg(a + c);
#pragma unroll
      for (int ll = 0; ll < kLC; ll++) {
        const int xs = x - (l0 + ll);
        const float bv = (xs >= 0) ? __ldg(im2 + ((int64_t)y * w + xs) * ch + c) : 0.f;
        acc[ll] += crit<CRIT>(av, bv);
      }
    }
#pragma unroll
    for (int ll = 0; ll < kLC; ll++) s_d[(ll * PH + py) * PW + px] = acc[ll];
  }
  __syncthreads();
  // horizontal sums: [kLC][PH][kTX]
  for (int i = tid; i < kLC * PH * kTX; i += kThreads) {
    const int x = i % kTX, rest = i / kTX;   // rest = ll * PH + py
    const float* p = s_d + rest * PW + x;
    float s = 0.f;
    for (int k = 0; k <= 2 * R; k++) s += p[k];
    s_h[i] = s;
  }
  __syncthreads();
  // vertical sums and store: thread <-> (pixel, all kLC disparities) so a pixel's 32 bytes go out together
  for (int i = tid; i < kTY * kTX; i += kThreads) {
    const int x = i % kTX, y = i / kTX;
    if (x0 + x >= w || y0 + y >= h) continue;
    float v[kLC];
#pragma unroll
    for (int ll = 0; ll < kLC; ll++) {
      const float* p = s_h + (ll * PH + y) * kTX + x;
      float s = 0.f;
      for (int k = 0; k <= 2 * R; k++) s += p[k * kTX];
      v[ll] = s;
    }
    float* o = out + ((int64_t)(y0 + y) * w + (x0 + x)) * L + l0;
    if (l0 + kLC <= L && (L & 3) == 0) {
      *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
    } else {
#pragma unroll
      for (int ll = 0; ll < kLC; ll++)
        if (l0 + ll < L) o[ll] = v[ll];
    }
  }
}

}  // namespace

int cost_volume(const float* im1, const float* im2, float* out, int h, int w, int ch, int L, int ws,
                int criterion, cudaStream_t st) {
  if (h <= 0 || w <= 0 || ch <= 0 || L <= 0 || ws <= 0 || (ws & 1) == 0 || ws > 2 * kMaxR + 1 ||
      criterion < 0 || criterion > 2) {
    set_last_error("pl_cost_volume: bad arguments (h=%d w=%d ch=%d L=%d ws=%d crit=%d)", h, w, ch, L,
                   ws, criterion);
    return PL_ERR_INVALID;
  }
  const int R = ws / 2;
  const int PW = kTX + 2 * R, PH = kTY + 2 * R;
  const size_t smem = sizeof(float) * (size_t)kLC * PH * (PW + kTX);
  dim3 grid((w + kTX - 1) / kTX, (h + kTY - 1) / kTY, (L + kLC - 1) / kLC);
  static PerDeviceOnce attr;
  if (attr.need()) {
    PL_CUDA(cudaFuncSetAttribute(cost_volume_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 128 * 1024));
    PL_CUDA(cudaFuncSetAttribute(cost_volume_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 128 * 1024));
    PL_CUDA(cudaFuncSetAttribute(cost_volume_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 128 * 1024));
  }
  if (criterion == 0) {
    PL_LAUNCH(cost_volume_kernel<0>, grid, kThreads, smem, st, im1, im2, out, h, w, ch, L, R);
  } else if (criterion == 1) {
    PL_LAUNCH(cost_volume_kernel<1>, grid, kThreads, smem, st, im1, im2, out, h, w, ch, L, R);
  } else {
    PL_LAUNCH(cost_volume_kernel<2>, grid, kThreads, smem, st, im1, im2, out, h, w, ch, L, R);
  }
  return PL_OK;
}

}  // namespace pl
