// pl_noalloc.cu -- pl_softmax_compat_noalloc: the softmax + compatibility transform of
// pl_softmax_compat for callers that cannot allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// pl_softmax_compat needs scratch memory: the compatibility matrix pre-split into TF32 hi / lo
// operand planes in the UMMA canonical layout (mf_tc_mu_bytes(L) bytes), which it takes from the
// stream-ordered pool.  Here that scratch is a file-scope device array of a fixed, modest size
// (L <= 64: 33 280 bytes) that every GPU loading the library holds until the process exits.
//
// The array's NAME is a host-side stand-in in host code, not a device address: its address is
// obtained once per device, with that device current, through cudaGetSymbolAddress, and that
// pointer is what the kernels receive.
//
// Every call shares the array and finds the previous call's contents in it.  Nothing in it has to
// start at zero: mu_split_kernel rewrites every byte of the planes the following kernel reads
// (pads included) before that kernel runs, in stream order.  Calls that OVERLAP on the GPU (two
// streams, two host threads) race on the array: serialise them.
#include "pl_common.cuh"

namespace pl {

namespace {

constexpr int kNoAllocMaxL = PL_NOALLOC_MAX_LABELS;                                  // 64
constexpr size_t kScratchBytes = 2 * (size_t)(kNoAllocMaxL / 4) * (kNoAllocMaxL * 16 + 16);   // 33 280

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// device address of g_noalloc_scratch on each device (nullptr: not looked up yet)
float* g_scratch_ptr[64] = {};

int scratch_for_current_device(float** out) {
  int dev = 0;
  PL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_last_error("pl_softmax_compat_noalloc: device index %d out of range", dev);
    return PL_ERR_INVALID;
  }
  if (g_scratch_ptr[dev] == nullptr) {
    int major = 0;
    PL_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) {
      set_last_error("device %d is sm_%dx; this library is built for sm_100a only", dev, major);
      return PL_ERR_NODEVICE;
    }
    void* p = nullptr;
    PL_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (p == nullptr || ((uintptr_t)p & 255u) != 0) {
      set_last_error("pl_softmax_compat_noalloc: bad scratch address %p", p);
      return PL_ERR_CUDA;
    }
    g_scratch_ptr[dev] = reinterpret_cast<float*>(p);
  }
  *out = g_scratch_ptr[dev];
  return PL_OK;
}

}  // namespace

}  // namespace pl

extern "C" int pl_softmax_compat_noalloc(const float* e_dev, const float* muT_dev, float* p_dev,
                                         float* q_out_dev, int64_t n, int L, float scale, void* stream) {
  if (e_dev == nullptr || muT_dev == nullptr || p_dev == nullptr || n <= 0) {
    pl::set_last_error("pl_softmax_compat_noalloc: null pointer or n <= 0");
    return PL_ERR_INVALID;
  }
  if (L > PL_NOALLOC_MAX_LABELS) {
    pl::set_last_error("pl_softmax_compat_noalloc: L = %d exceeds the static scratch (at most %d labels); "
                       "use pl_softmax_compat", L, PL_NOALLOC_MAX_LABELS);
    return PL_ERR_INVALID;
  }
  if (!pl::mf_tc_supported(nullptr, L)) {
    pl::set_last_error("pl_softmax_compat_noalloc: L = %d not supported (16, 32 or 64)", L);
    return PL_ERR_INVALID;
  }
  if ((((uintptr_t)e_dev | (uintptr_t)p_dev | (uintptr_t)q_out_dev | (uintptr_t)muT_dev) & 15u) != 0) {
    pl::set_last_error("pl_softmax_compat_noalloc: pointers must be 16-byte aligned");
    return PL_ERR_INVALID;
  }
  if (pl::mf_tc_mu_bytes(L) > pl::kScratchBytes) {   // cannot happen for L <= 64; guards the constant
    pl::set_last_error("pl_softmax_compat_noalloc: scratch too small for L = %d", L);
    return PL_ERR_INVALID;
  }
  float* scratch = nullptr;
  PL_TRY(pl::scratch_for_current_device(&scratch));
  cudaStream_t st = (cudaStream_t)stream;
  // split Mu into the scratch (every byte of the planes is rewritten), then the same kernel
  // pl_softmax_compat launches; no allocation anywhere on this path
  PL_TRY(pl::mf_tc_prepare_mu(muT_dev, L, scratch, st));
  return pl::softmax_compat_tc(e_dev, scratch, p_dev, q_out_dev, n, L, scale, st);
}
