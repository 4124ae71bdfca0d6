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
#include <stdlib.h>

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

// ---------------------------------------------------------------------------------------------
// Fast path for the reference's default 9 x 9 window (R = 4) and float4-aligned label blocks.
// Same arithmetic, same summation order per output (so results are bit-identical to the generic
// kernel above), less redundant work:
//   * CTA = 16 rows x 64 columns x 8 disparities (halo overhead 1.69x instead of 2.25x);
//   * per-pixel costs: both images are first packed to one float4 per pixel (pack_pixels_kernel); a
//     thread takes 8 consecutive patch pixels, and the 15 right-image pixels they need for 8 disparities
//     are loaded once each with one LDG.128 (23 vector loads per 192 cost terms instead of 216 scalar
//     ones: the phase is bound by its load instructions);
//   * window sums: a thread produces 8 horizontal neighbours from 16 values (LDS.128), 4 vertical ones
//     from 12 -- 2-3 shared-memory reads per output instead of 9.
// ---------------------------------------------------------------------------------------------
// [h*w][ch] (ch <= 4) -> one float4 per pixel, missing channels zero: crit(0, 0) = 0 for all three
// criteria, so the padded channels add exact zeros.  One LDG.128 per pixel instead of three LDG.32:
// the cost phase is bound by its load instructions (measured: the kernel's time follows their count).
__global__ void __launch_bounds__(256)
pack_pixels_kernel(const float* __restrict__ im1, const float* __restrict__ im2, float4* __restrict__ out,
                   int64_t n, int ch) {
  // out[0 .. n) = left image, out[n .. 2n) = right image
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < 2 * n; i += (int64_t)gridDim.x * blockDim.x) {
    const float* im = i < n ? im1 + i * ch : im2 + (i - n) * ch;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < ch; c++) v[c] = im[c];
    out[i] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

template <int CRIT>
__device__ __forceinline__ float crit4(const float4 a, const float4 b, int ch) {
  // channels in ascending order, like the reference's sum over axis 2 (depth.py:49)
  float r = crit<CRIT>(a.x, b.x);
  if (ch > 1) r += crit<CRIT>(a.y, b.y);
  if (ch > 2) r += crit<CRIT>(a.z, b.z);
  if (ch > 3) r += crit<CRIT>(a.w, b.w);
  return r;
}

constexpr int kTY2 = 16, kR2 = 4;
constexpr int kPW2 = kTX + 2 * kR2;   // 72
constexpr int kPH2 = kTY2 + 2 * kR2;  // 24

template <int CRIT, int LC>
__global__ void __launch_bounds__(kThreads, LC == 8 ? 2 : 4)
cost_volume9_kernel(const float4* __restrict__ im1, const float4* __restrict__ im2,
                    float* __restrict__ out, int h, int w, int ch, int L) {
  extern __shared__ __align__(16) float sm[];
  float* s_d = sm;                          // [LC][kPH2][kPW2]
  float* s_h = sm + LC * kPH2 * kPW2;      // [LC][kPH2][kTX]
  const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * kTY2, l0 = blockIdx.z * LC;
  const int tid = threadIdx.x;

  // ---- per-pixel costs of the haloed patch: groups of 8 consecutive patch pixels ----------------
  // (staging the image patches in shared memory first was tried: 277 us against 235 us for these
  //  L1 / L2-served loads at 740 x 500 x 64 -- bank conflicts on the 3-channel stride and one more phase)
  for (int i = tid; i < kPH2 * (kPW2 / 8); i += kThreads) {
    const int py = i / (kPW2 / 8), pg = i - py * (kPW2 / 8);
    const int y = reflect(y0 + py - kR2, h);
    const int xa = x0 + pg * 8 - kR2;
    float acc[8][LC];
#pragma unroll
    for (int q = 0; q < 8; q++)
#pragma unroll
      for (int ll = 0; ll < LC; ll++) acc[q][ll] = 0.f;
    if (xa >= 0 && xa + 7 < w) {
      // interior: pixel q sits at x = xa + q; disparity l0 + ll reads the right image at x - l0 - ll =
      // pixel k = q - ll + LC - 1 of the run of 8 + LC - 1 pixels starting at xb.  Each right pixel is
      // loaded once (one LDG.128) and serves every (q, ll) pair on its anti-diagonal.
      const float4* a = im1 + (int64_t)y * w + xa;
      const float4* b = im2 + (int64_t)y * w;
      const int xb = xa - l0 - (LC - 1);
      float4 av[8];
#pragma unroll
      for (int q = 0; q < 8; q++) av[q] = __ldg(a + q);
#pragma unroll
      for (int k = 0; k < 8 + LC - 1; k++) {
        const float4 bv = (xb + k >= 0) ? __ldg(b + xb + k) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < 8; q++) {
          const int ll = q + LC - 1 - k;
          if (ll >= 0 && ll < LC) acc[q][ll] = crit4<CRIT>(av[q], bv, ch);
        }
      }
    } else {
      // image border: reflected coordinates, pixel by pixel (the reflection acts before the shift)
#pragma unroll 1
      for (int q = 0; q < 8; q++) {
        const int x = reflect(xa + q, w);
        const float4 av = __ldg(im1 + (int64_t)y * w + x);
        float t[LC];
#pragma unroll
        for (int ll = 0; ll < LC; ll++) {
          const int xs = x - (l0 + ll);
          const float4 bv = (xs >= 0) ? __ldg(im2 + (int64_t)y * w + xs) : make_float4(0.f, 0.f, 0.f, 0.f);
          t[ll] = crit4<CRIT>(av, bv, ch);
        }
#pragma unroll
        for (int ll = 0; ll < LC; ll++)
#pragma unroll
          for (int qq = 0; qq < 8; qq++)
            if (qq == q) acc[qq][ll] = t[ll];
      }
    }
#pragma unroll
    for (int ll = 0; ll < LC; ll++) {
      float* d = s_d + (ll * kPH2 + py) * kPW2 + pg * 8;
      *reinterpret_cast<float4*>(d) = make_float4(acc[0][ll], acc[1][ll], acc[2][ll], acc[3][ll]);
      *reinterpret_cast<float4*>(d + 4) = make_float4(acc[4][ll], acc[5][ll], acc[6][ll], acc[7][ll]);
    }
  }
  __syncthreads();
  // ---- horizontal 9-term sums, 8 neighbours per thread: [LC][kPH2][kTX] -------------------------
  for (int i = tid; i < LC * kPH2 * (kTX / 8); i += kThreads) {
    const int xg = i % (kTX / 8), rest = i / (kTX / 8);   // rest = ll * kPH2 + py
    const float4* p = reinterpret_cast<const float4*>(s_d + rest * kPW2 + xg * 8);
    float v[16];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const float4 t = p[k];
      v[4 * k] = t.x; v[4 * k + 1] = t.y; v[4 * k + 2] = t.z; v[4 * k + 3] = t.w;
    }
    float r[8];
#pragma unroll
    for (int q = 0; q < 8; q++) {
      float sacc = 0.f;
#pragma unroll
      for (int k = 0; k <= 2 * kR2; k++) sacc += v[q + k];
      r[q] = sacc;
    }
    float* o = s_h + rest * kTX + xg * 8;
    *reinterpret_cast<float4*>(o) = make_float4(r[0], r[1], r[2], r[3]);
    *reinterpret_cast<float4*>(o + 4) = make_float4(r[4], r[5], r[6], r[7]);
  }
  __syncthreads();
  // ---- vertical 9-term sums, 4 rows per thread; a pixel's 8 disparities leave as one 32-byte sector --
  for (int i = tid; i < (kTY2 / 4) * kTX; i += kThreads) {
    const int x = i % kTX, yg = i / kTX;
    if (x0 + x >= w) continue;
    float r[4][LC];
#pragma unroll
    for (int ll = 0; ll < LC; ll++) {
      const float* p = s_h + (ll * kPH2 + yg * 4) * kTX + x;
      float v[12];
#pragma unroll
      for (int k = 0; k < 12; k++) v[k] = p[k * kTX];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        float sacc = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * kR2; k++) sacc += v[q + k];
        r[q][ll] = sacc;
      }
    }
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int y = y0 + yg * 4 + q;
      if (y >= h) continue;
      float* o = out + ((int64_t)y * w + (x0 + x)) * L + l0;
#pragma unroll
      for (int c4 = 0; c4 < LC / 4; c4++)
        *reinterpret_cast<float4*>(o + 4 * c4) =
            make_float4(r[q][4 * c4], r[q][4 * c4 + 1], r[q][4 * c4 + 2], r[q][4 * c4 + 3]);
    }
  }
}

template <int CRIT, int LC>
int launch_cost9(const float* im1_raw, const float* im2_raw, float* out, int h, int w, int ch, int L,
                 cudaStream_t st) {
  // pack both images to float4 pixels (stream-ordered scratch, 32 n bytes)
  const int64_t n = (int64_t)h * w;
  float4* packed = nullptr;
  PL_TRY(dev_alloc((void**)&packed, sizeof(float4) * (size_t)n * 2, st));
  PL_LAUNCH(pack_pixels_kernel, grid_for((uint64_t)2 * n, 256, 8), 256, 0, st, im1_raw, im2_raw, packed, n, ch);
  const float4* im1 = packed;
  const float4* im2 = packed + n;
  const size_t smem = sizeof(float) * (size_t)LC * kPH2 * (kPW2 + kTX);   // 104 448 bytes
  static PerDeviceOnce attr;
  if (attr.need())
    PL_CUDA(cudaFuncSetAttribute((cost_volume9_kernel<CRIT, LC>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((w + kTX - 1) / kTX, (h + kTY2 - 1) / kTY2, L / LC);
  PL_LAUNCH((cost_volume9_kernel<CRIT, LC>), grid, kThreads, smem, st, im1, im2, out, h, w, ch, L);
  dev_free(packed, st);
  return PL_OK;
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
  if (R == kR2 && L % kLC == 0 && ch <= 4 && ((uintptr_t)out & 15u) == 0) {   // the reference's default window
    // (a 4-label chunk per CTA -- 64 registers, twice the CTAs per SM -- measured slower: 209 against 183 us)
    if (criterion == 0) return launch_cost9<0, 8>(im1, im2, out, h, w, ch, L, st);
    if (criterion == 1) return launch_cost9<1, 8>(im1, im2, out, h, w, ch, L, st);
    return launch_cost9<2, 8>(im1, im2, out, h, w, ch, L, st);
  }
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
