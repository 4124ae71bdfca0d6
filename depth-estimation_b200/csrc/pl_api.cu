// pl_api.cu -- extern "C" entry points of libpl_lattice.so (see include/pl_lattice.h).
#include <math.h>
#include <stdlib.h>
#include <stdarg.h>
#include <string.h>

#include <new>
#include <string>
#include <deque>
#include <iterator>
#include <mutex>
#include <unordered_map>
#include <utility>
#include <vector>

#include "pl_common.cuh"

namespace pl {

unsigned long long g_launches = 0;
static thread_local char t_last_error[512] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_last_error, sizeof(t_last_error), fmt, ap);
  va_end(ap);
}

// ---- optional per-launch timing (pl_profile / pl_profile_report) -------------
bool g_prof_on = false;
struct ProfRec {
  const char* name;
  cudaEvent_t a, b;
};
static std::vector<ProfRec> g_prof;

void prof_begin(const char* name, cudaStream_t st) {
  ProfRec r;
  r.name = name;
  cudaEventCreate(&r.a);
  cudaEventCreate(&r.b);
  cudaEventRecord(r.a, st);
  g_prof.push_back(r);
}
void prof_end(cudaStream_t st) {
  if (!g_prof.empty()) cudaEventRecord(g_prof.back().b, st);
}

// Stream-ordered allocations from the device's default memory pool; the pool
// keeps freed blocks (release threshold = max) so repeated builds do not go
// back to the driver.
static bool g_pool_ready[64] = {false};

static int ensure_pool(int dev) {
  if (dev >= 0 && dev < 64 && !g_pool_ready[dev]) {
    cudaMemPool_t pool;
    PL_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
    uint64_t thr = UINT64_MAX;
    PL_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    g_pool_ready[dev] = true;
  }
  return PL_OK;
}

// A small exact-size cache in front of the pool.  A frame allocates and frees ~30 blocks of the same
// sizes in the same order (0.1 - 34 MB at 740 x 500).  A freed block is kept here with its size, device
// and the stream it was last used on, and handed back only to a request of exactly that size on that
// device AND that stream, which is safe in stream order without any event: in the steady state of a
// streaming loop no allocation reaches the driver (824 against 829 us per frame, and the pool's own
// split / re-join reuse cannot drift: with a growing pool cudaMallocAsync was seen to take 200 us for
// the 33.5 MB hash table, occasionally 10 - 150 ms with the device idle).
// Beyond PL_ALLOC_CACHE_MB (default 2048, 0 = off) the oldest blocks go back to the pool.
namespace {
struct CacheEnt {
  void* p;
  size_t bytes;
  cudaStream_t st;
  int dev;
};
std::mutex g_cache_mu;
std::unordered_map<void*, std::pair<size_t, int>> g_live;   // blocks handed out: size, device
std::deque<CacheEnt> g_cache;                                // freed blocks, oldest first
size_t g_cache_bytes = 0;

size_t cache_cap() {
  static const size_t cap = [] {
    const char* e = getenv("PL_ALLOC_CACHE_MB");
    const long long mb = e != nullptr ? atoll(e) : 2048;
    return (size_t)(mb > 0 ? mb : 0) << 20;
  }();
  return cap;
}

// back to the pool, in the order of the stream the block was last used on; if that stream is gone by
// now, the synchronising form
void cache_release(const CacheEnt& e) {
  int cur = 0;
  cudaGetDevice(&cur);
  if (cur != e.dev) cudaSetDevice(e.dev);
  if (cudaFreeAsync(e.p, e.st) != cudaSuccess) {
    (void)cudaGetLastError();
    cudaFree(e.p);
  }
  if (cur != e.dev) cudaSetDevice(cur);
}
}  // namespace

int dev_alloc(void** p, size_t bytes, cudaStream_t st) {
  *p = nullptr;
  if (bytes == 0) bytes = 16;
  int dev = 0;
  PL_CUDA(cudaGetDevice(&dev));
  if (cache_cap() > 0) {
    std::lock_guard<std::mutex> lock(g_cache_mu);
    for (auto it = g_cache.rbegin(); it != g_cache.rend(); ++it) {
      if (it->bytes == bytes && it->st == st && it->dev == dev) {
        *p = it->p;
        g_cache_bytes -= bytes;
        g_cache.erase(std::next(it).base());
        g_live[*p] = {bytes, dev};
        return PL_OK;
      }
    }
  }
  PL_CUDA(cudaMallocAsync(p, bytes, st));
  if (cache_cap() > 0) {
    std::lock_guard<std::mutex> lock(g_cache_mu);
    g_live[*p] = {bytes, dev};
  }
  return PL_OK;
}

void dev_free(void* p, cudaStream_t st) {
  if (p == nullptr) return;
  if (cache_cap() > 0) {
    std::vector<CacheEnt> evict;
    bool kept = false;
    {
      std::lock_guard<std::mutex> lock(g_cache_mu);
      auto it = g_live.find(p);
      if (it != g_live.end()) {
        const size_t bytes = it->second.first;
        const int dev = it->second.second;
        g_live.erase(it);
        if (bytes <= cache_cap()) {
          g_cache.push_back({p, bytes, st, dev});
          g_cache_bytes += bytes;
          kept = true;
          while (g_cache_bytes > cache_cap() && !g_cache.empty()) {
            evict.push_back(g_cache.front());
            g_cache_bytes -= g_cache.front().bytes;
            g_cache.pop_front();
          }
        }
      }
    }
    for (const CacheEnt& e : evict) cache_release(e);
    if (kept) return;
  }
  cudaFreeAsync(p, st);
}

static void lattice_release(Lattice* L, cudaStream_t st) {
  dev_free(L->vid, st);
  dev_free(L->bary, st);
  dev_free(L->greedy, st);
  dev_free(L->rank, st);
  dev_free(L->vkeys, st);
  dev_free(L->nbr, st);
  dev_free(L->ent_pix, st);
  dev_free(L->ent_w, st);
  dev_free(L->loc, st);
  dev_free(L->useg, st);
  dev_free(L->tile_off, st);
  dev_free(L->rec_slice, st);
  dev_free(L->rec_splat, st);
  dev_free(L->val[1], st);   // val[0] lives in the same block
  dev_free(L->val[2], st);
  dev_free(L->flow, st);
  if (L->max_U_ev != nullptr) cudaEventDestroy(L->max_U_ev);
  if (L->aux_ev != nullptr) cudaEventDestroy(L->aux_ev);
  delete L;
}

static int check_device() {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_last_error("no CUDA device: %s", cudaGetErrorString(e));
    return PL_ERR_NODEVICE;
  }
  static int checked[64] = {0};
  if (dev < 64 && !checked[dev]) {
    int major = 0;
    PL_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) {
      set_last_error("device %d is sm_%dx; this library is built for sm_100a only", dev, major);
      return PL_ERR_NODEVICE;
    }
    checked[dev] = 1;
  }
  return ensure_pool(dev);
}

// neighbour table as the tests want it: int32 [d+1][M][2], -1 = absent
__global__ void __launch_bounds__(256)
export_nbr_kernel(const int2* __restrict__ nbr, int64_t total, int M, int32_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int2 v = nbr[i];
    out[2 * i + 0] = v.x >= M ? -1 : v.x;
    out[2 * i + 1] = v.y >= M ? -1 : v.y;
  }
}

static int lattice_export(const Lattice* L, int what, void* dst, cudaStream_t st) {
  const size_t E = (size_t)L->n * (L->d + 1);
  const void* src = nullptr;
  size_t bytes = 0;
  switch (what) {
    case PL_EXPORT_GREEDY: src = L->greedy; bytes = sizeof(int16_t) * E; break;
    case PL_EXPORT_RANK: src = L->rank; bytes = sizeof(int8_t) * E; break;
    case PL_EXPORT_BARY: src = L->bary; bytes = sizeof(float) * E; break;
    case PL_EXPORT_VERTEX: src = L->vid; bytes = sizeof(int32_t) * E; break;
    case PL_EXPORT_KEYS: src = L->vkeys; bytes = sizeof(int16_t) * (size_t)L->M * L->d; break;
    case PL_EXPORT_NEIGHBOURS: {
      const int64_t total = L->M * (L->d + 1);
      PL_LAUNCH(export_nbr_kernel, grid_for((uint64_t)total, 256, 8), 256, 0, st, L->nbr, total,
                (int)L->M, (int32_t*)dst);
      return PL_OK;
    }
    default:
      set_last_error("pl_lattice_export: unknown artefact %d", what);
      return PL_ERR_INVALID;
  }
  PL_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st));
  return PL_OK;
}

}  // namespace pl

using pl::Lattice;

extern "C" {

int pl_lattice_create(const float* ref_dev, int64_t n, int d, uint32_t flags, void* stream,
                      pl_lattice** out) {
  if (out == nullptr) return PL_ERR_INVALID;
  *out = nullptr;
  if (ref_dev == nullptr || n <= 0 || d < 1 || d > PL_MAX_D) {
    pl::set_last_error("pl_lattice_create: bad arguments (n=%lld, d=%d)", (long long)n, d);
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  cudaStream_t st = (cudaStream_t)stream;
  Lattice* L = new (std::nothrow) Lattice();
  if (!L) return PL_ERR_INVALID;
  L->n = n;
  L->d = d;
  L->flags = flags;
  L->last_stream = st;
  cudaGetDevice(&L->device);
  int r = pl::lattice_build(L, ref_dev, st);
  if (r != PL_OK) {
    pl::lattice_release(L, st);
    return r;
  }
  *out = reinterpret_cast<pl_lattice*>(L);
  return PL_OK;
}

int pl_lattice_create_ijrgb(const float* rgb_planes_dev, int h, int w, float s_ij, float s_rgb,
                            uint32_t flags, void* stream, pl_lattice** out) {
  if (out == nullptr) return PL_ERR_INVALID;
  *out = nullptr;
  if (rgb_planes_dev == nullptr || h <= 0 || w <= 0 || !(s_ij != 0.f) || !(s_rgb != 0.f)) {
    pl::set_last_error("pl_lattice_create_ijrgb: bad arguments (h=%d, w=%d)", h, w);
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  cudaStream_t st = (cudaStream_t)stream;
  Lattice* L = new (std::nothrow) Lattice();
  if (!L) return PL_ERR_INVALID;
  L->n = (int64_t)h * w;
  L->d = 5;
  L->flags = flags;
  L->last_stream = st;
  cudaGetDevice(&L->device);
  pl::GuideSrc gs;
  gs.rgb = rgb_planes_dev; gs.h = h; gs.w = w; gs.s_ij = s_ij; gs.s_rgb = s_rgb;
  gs.diag = sqrt((double)((int64_t)h * h + (int64_t)w * w));   // np.sqrt(h**2 + w**2)
  int r = pl::lattice_build(L, nullptr, st, &gs);
  if (r != PL_OK) {
    pl::lattice_release(L, st);
    return r;
  }
  *out = reinterpret_cast<pl_lattice*>(L);
  return PL_OK;
}

void pl_lattice_destroy(pl_lattice* lat) {
  if (lat == nullptr) return;
  Lattice* L = reinterpret_cast<Lattice*>(lat);
  // stream-ordered release on the stream that used the lattice last (torch side streams are
  // non-blocking: freeing on the legacy stream would not be ordered against them)
  int cur = -1;
  cudaGetDevice(&cur);
  if (cur != L->device) cudaSetDevice(L->device);
  pl::lattice_release(L, L->last_stream);
  if (cur >= 0 && cur != L->device) cudaSetDevice(cur);
}

int64_t pl_lattice_num_points(const pl_lattice* lat) { return reinterpret_cast<const Lattice*>(lat)->n; }
int pl_lattice_dim(const pl_lattice* lat) { return reinterpret_cast<const Lattice*>(lat)->d; }
int64_t pl_lattice_num_vertices(const pl_lattice* lat) { return reinterpret_cast<const Lattice*>(lat)->M; }
int64_t pl_lattice_device_bytes(const pl_lattice* lat) {
  const Lattice* L = reinterpret_cast<const Lattice*>(lat);
  // the two value buffers, the rotating loop's third one and the barrier-free blur's 2 x 5 pass buffers
  return L->bytes + (int64_t)sizeof(float) * ((L->val[2] != nullptr ? 3 : 2) * L->ws_floats + 10 * L->flow_floats);
}

int pl_lattice_apply(pl_lattice* lat, const float* src_dev, int64_t src_stride, float* dst_dev,
                     int64_t dst_stride, int C, const float* sub_dev, int64_t sub_stride,
                     const float* add_dev, int64_t add_stride, uint32_t flags, void* stream) {
  if (lat == nullptr) return PL_ERR_INVALID;
  reinterpret_cast<Lattice*>(lat)->last_stream = (cudaStream_t)stream;
  return pl::lattice_apply(reinterpret_cast<Lattice*>(lat), src_dev, src_stride, dst_dev,
                           dst_stride, C, sub_dev, sub_stride, add_dev, add_stride, flags,
                           (cudaStream_t)stream);
}

int pl_lattice_apply_dot(pl_lattice* lat, const float* src_dev, int64_t src_stride, int C,
                         const float* rowscale_dev, int64_t rowscale_stride, const float* other_dev,
                         int64_t other_stride, float* out_dev, int64_t out_stride, float coef, void* stream) {
  if (lat == nullptr) return PL_ERR_INVALID;
  reinterpret_cast<Lattice*>(lat)->last_stream = (cudaStream_t)stream;
  return pl::lattice_apply_dot(reinterpret_cast<Lattice*>(lat), src_dev, src_stride, C, rowscale_dev,
                               rowscale_stride, other_dev, other_stride, out_dev, out_stride, coef,
                               (cudaStream_t)stream);
}

int pl_lattice_apply_dot_multi(pl_lattice* lat, const float* src_dev, int64_t src_stride, int C,
                               const float* rowscale_dev, int64_t rowscale_stride, int J,
                               const float* other_dev, int64_t other_stride, float* out_dev,
                               int64_t out_stride, float coef, void* stream) {
  if (lat == nullptr) return PL_ERR_INVALID;
  reinterpret_cast<Lattice*>(lat)->last_stream = (cudaStream_t)stream;
  return pl::lattice_apply_dot(reinterpret_cast<Lattice*>(lat), src_dev, src_stride, C, rowscale_dev,
                               rowscale_stride, other_dev, other_stride, out_dev, out_stride, coef,
                               (cudaStream_t)stream, J);
}

int pl_filter(const float* src_dev, const float* ref_dev, float* dst_dev, int64_t n, int C, int d,
              void* stream) {
  pl_lattice* lat = nullptr;
  // the reference's own call shape: reproduce its table-growth behaviour too
  PL_TRY(pl_lattice_create(ref_dev, n, d, PL_BUILD_REF_GROW_QUIRK, stream, &lat));
  int r = pl_lattice_apply(lat, src_dev, C, dst_dev, C, C, nullptr, 0, nullptr, 0,
                           PL_APPLY_DEFAULT, stream);
  pl::lattice_release(reinterpret_cast<Lattice*>(lat), (cudaStream_t)stream);
  return r;
}

int pl_filter_host(const float* src_host, const float* ref_host, float* dst_host, int64_t n, int C,
                   int d, void* stream) {
  if (src_host == nullptr || ref_host == nullptr || dst_host == nullptr || n <= 0 || C <= 0) {
    pl::set_last_error("pl_filter_host: bad arguments");
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  cudaStream_t st = (cudaStream_t)stream;
  float *src = nullptr, *ref = nullptr, *dst = nullptr;
  const size_t sb = sizeof(float) * (size_t)n * C, rb = sizeof(float) * (size_t)n * d;
  PL_TRY(pl::dev_alloc((void**)&src, sb, st));
  PL_TRY(pl::dev_alloc((void**)&ref, rb, st));
  PL_TRY(pl::dev_alloc((void**)&dst, sb, st));
  int r = PL_OK;
  cudaError_t e = cudaMemcpyAsync(ref, ref_host, rb, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(src, src_host, sb, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) {
    r = pl_filter(src, ref, dst, n, C, d, stream);
    if (r == PL_OK) e = cudaMemcpyAsync(dst_host, dst, sb, cudaMemcpyDeviceToHost, st);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  pl::dev_free(src, st);
  pl::dev_free(ref, st);
  pl::dev_free(dst, st);
  if (e != cudaSuccess) {
    pl::set_last_error("pl_filter_host: %s", cudaGetErrorString(e));
    return PL_ERR_CUDA;
  }
  return r;
}

static int compat_large(const float* e, int nblocks, const float* muT, float* p, float* q_out,
                        int64_t n, int L, int l0, int N, float scale, cudaStream_t st,
                        const float* const* block_ptrs = nullptr, float* const* out_block_ptrs = nullptr);

int pl_softmax_compat_ex(const float* e_dev, const float* muT_dev, float* p_dev, float* q_out_dev,
                         int64_t n, int L, float scale, uint32_t flags, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const bool aligned = (((uintptr_t)e_dev | (uintptr_t)p_dev | (uintptr_t)q_out_dev) & 15u) == 0;
  if (!(flags & PL_MEANFIELD_NO_TENSOR) && aligned && e_dev && muT_dev && p_dev && n > 0 &&
      pl::mf_tc_supported(nullptr, L)) {
    // tcgen05 path: Mu is split into tf32 hi / lo operand planes once, then one kernel
    PL_TRY(pl::check_device());
    float* mu_canon = nullptr;
    PL_TRY(pl::dev_alloc((void**)&mu_canon, pl::mf_tc_mu_bytes(L), st));
    int r = pl::mf_tc_prepare_mu(muT_dev, L, mu_canon, st);
    if (r == PL_OK) r = pl::softmax_compat_tc(e_dev, mu_canon, p_dev, q_out_dev, n, L, scale, st);
    pl::dev_free(mu_canon, st);
    return r;
  }
  if (!(flags & PL_MEANFIELD_NO_TENSOR) && aligned && e_dev && muT_dev && p_dev && n > 0 &&
      L > 64 && L <= 256 && pl::compat_tc_supported(L, L, L)) {
    PL_TRY(pl::check_device());
    return compat_large(e_dev, 1, muT_dev, p_dev, q_out_dev, n, L, 0, L, scale, st);
  }
  return pl::softmax_compat(e_dev, muT_dev, p_dev, q_out_dev, n, L, scale, st);
}

static int compat_large(const float* e, int nblocks, const float* muT, float* p, float* q_out,
                        int64_t n, int L, int l0, int N, float scale, cudaStream_t st,
                        const float* const* block_ptrs, float* const* out_block_ptrs) {
  uint8_t* rec = nullptr;
  PL_TRY(pl::dev_alloc((void**)&rec, pl::compat_tc_mu_bytes(L, N), st));
  int r = pl::compat_tc_prepare_mu(muT, L, l0, N, rec, st);
  if (r == PL_OK)
    r = pl::compat_tc_run(e, nblocks, L / nblocks, rec, p, q_out, n, L, N, scale, st, block_ptrs,
                          out_block_ptrs);
  pl::dev_free(rec, st);
  return r;
}

int pl_softmax_compat_blocked(const float* e_blocks_dev, int nblocks, const float* muT_dev,
                              float* p_dev, int64_t n, int L, int l0, int N, float scale,
                              void* stream) {
  if (e_blocks_dev == nullptr || muT_dev == nullptr || p_dev == nullptr || n <= 0 || nblocks < 1 ||
      L <= 0 || L % nblocks != 0 || l0 < 0 || N <= 0 || l0 + N > L ||
      (((uintptr_t)e_blocks_dev | (uintptr_t)p_dev) & 15u) != 0 ||
      !pl::compat_tc_supported(L, N, L / nblocks)) {
    pl::set_last_error("pl_softmax_compat_blocked: unsupported arguments (L=%d, N=%d, nblocks=%d)", L,
                       N, nblocks);
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  return compat_large(e_blocks_dev, nblocks, muT_dev, p_dev, nullptr, n, L, l0, N, scale,
                      (cudaStream_t)stream);
}

int pl_softmax_compat_peers(const float* const* e_block_ptrs, int nblocks, const float* muT_dev,
                            float* p_dev, int64_t n, int L, int l0, int N, float scale,
                            void* stream) {
  if (e_block_ptrs == nullptr || muT_dev == nullptr || p_dev == nullptr || n <= 0 || nblocks < 1 ||
      nblocks > 16 || L <= 0 || L % nblocks != 0 || l0 < 0 || N <= 0 || l0 + N > L ||
      ((uintptr_t)p_dev & 15u) != 0 || !pl::compat_tc_supported(L, N, L / nblocks)) {
    pl::set_last_error("pl_softmax_compat_peers: unsupported arguments (L=%d, N=%d, nblocks=%d)", L, N,
                       nblocks);
    return PL_ERR_INVALID;
  }
  for (int b = 0; b < nblocks; b++)
    if (e_block_ptrs[b] == nullptr || ((uintptr_t)e_block_ptrs[b] & 15u) != 0) {
      pl::set_last_error("pl_softmax_compat_peers: block %d is null or misaligned", b);
      return PL_ERR_INVALID;
    }
  PL_TRY(pl::check_device());
  return compat_large(e_block_ptrs[0], nblocks, muT_dev, p_dev, nullptr, n, L, l0, N, scale,
                      (cudaStream_t)stream, e_block_ptrs);
}

int pl_softmax_compat_shards(const float* const* e_block_ptrs, float* const* p_block_ptrs, int nblocks,
                             const float* muT_dev, int64_t row0, int64_t nrows, int L, float scale,
                             void* stream) {
  if (e_block_ptrs == nullptr || p_block_ptrs == nullptr || muT_dev == nullptr || nrows < 0 || row0 < 0 ||
      nblocks < 1 || nblocks > 16 || L <= 0 || L % nblocks != 0 ||
      !pl::compat_tc_supported(L, L, L / nblocks)) {
    pl::set_last_error("pl_softmax_compat_shards: unsupported arguments (L=%d, nblocks=%d)", L, nblocks);
    return PL_ERR_INVALID;
  }
  if (nrows == 0) return PL_OK;
  const int bw = L / nblocks;
  const float* eb[16];
  float* pb[16];
  for (int b = 0; b < nblocks; b++) {
    if (e_block_ptrs[b] == nullptr || p_block_ptrs[b] == nullptr ||
        (((uintptr_t)e_block_ptrs[b] | (uintptr_t)p_block_ptrs[b]) & 15u) != 0) {
      pl::set_last_error("pl_softmax_compat_shards: block %d is null or misaligned", b);
      return PL_ERR_INVALID;
    }
    eb[b] = e_block_ptrs[b] + row0 * bw;
    pb[b] = p_block_ptrs[b] + row0 * bw;
  }
  PL_TRY(pl::check_device());
  return compat_large(eb[0], nblocks, muT_dev, nullptr, nullptr, nrows, L, 0, L, scale,
                      (cudaStream_t)stream, eb, pb);
}

// ---- peer memory (CUDA IPC) for the label shards ------------------------------------------
int pl_peer_alloc(size_t bytes, void** dev_ptr_out, void* ipc_handle_out) {
  if (dev_ptr_out == nullptr || ipc_handle_out == nullptr || bytes == 0) return PL_ERR_INVALID;
  *dev_ptr_out = nullptr;
  PL_TRY(pl::check_device());
  PL_CUDA(cudaMalloc(dev_ptr_out, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, *dev_ptr_out);
  if (e != cudaSuccess) {
    cudaFree(*dev_ptr_out);
    *dev_ptr_out = nullptr;
    pl::set_last_error("cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
    return PL_ERR_CUDA;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == PL_IPC_HANDLE_BYTES, "handle size");
  memcpy(ipc_handle_out, &h, sizeof(h));
  return PL_OK;
}

int pl_peer_open(const void* ipc_handle, void** dev_ptr_out) {
  if (ipc_handle == nullptr || dev_ptr_out == nullptr) return PL_ERR_INVALID;
  *dev_ptr_out = nullptr;
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle, sizeof(h));
  PL_CUDA(cudaIpcOpenMemHandle(dev_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return PL_OK;
}

int pl_peer_close(void* dev_ptr) {
  if (dev_ptr == nullptr) return PL_OK;
  PL_CUDA(cudaIpcCloseMemHandle(dev_ptr));
  return PL_OK;
}

int pl_peer_free(void* dev_ptr) {
  if (dev_ptr == nullptr) return PL_OK;
  PL_CUDA(cudaFree(dev_ptr));
  return PL_OK;
}

int pl_softmax_compat(const float* e_dev, const float* muT_dev, float* p_dev, float* q_out_dev,
                      int64_t n, int L, float scale, void* stream) {
  return pl_softmax_compat_ex(e_dev, muT_dev, p_dev, q_out_dev, n, L, scale, 0u, stream);
}

// Read-out request of pl_meanfield_depth (all null: plain pl_meanfield, E goes to e_out_dev).
struct Readout {
  float* depth = nullptr;
  int32_t* label = nullptr;
  const float* labels = nullptr;
  bool on() const { return depth != nullptr || label != nullptr; }
};

static int meanfield_impl(pl_lattice* lat, const float* e0_dev, const float* muT_dev, float* e_out_dev,
                          float* work_dev, int L, int niters, float scale, uint32_t flags,
                          void* stream, const Readout& ro) {
  if (lat == nullptr || e0_dev == nullptr || muT_dev == nullptr || e_out_dev == nullptr ||
      work_dev == nullptr || L <= 0 || niters < 0 || e_out_dev == e0_dev) {
    pl::set_last_error("pl_meanfield: bad arguments");
    return PL_ERR_INVALID;
  }
  Lattice* Lt = reinterpret_cast<Lattice*>(lat);
  cudaStream_t st = (cudaStream_t)stream;
  Lt->last_stream = st;
  const int64_t n = Lt->n;
  if (niters == 0) {
    PL_CUDA(cudaMemcpyAsync(e_out_dev, e0_dev, sizeof(float) * (size_t)n * L,
                            cudaMemcpyDeviceToDevice, st));
    return PL_OK;
  }
  const bool aligned = (((uintptr_t)e0_dev | (uintptr_t)e_out_dev | (uintptr_t)work_dev |
                         (uintptr_t)muT_dev) & 15u) == 0;
  const bool fusable = aligned && !(flags & (PL_APPLY_FLAT_KERNELS | PL_APPLY_EXACT_SLICE |
                                             PL_MEANFIELD_UNFUSED));
  const bool use_tc = fusable && !(flags & PL_MEANFIELD_NO_TENSOR) && pl::mf_tc_supported(Lt, L);
  // warp-specialised form (pl_fused_ws.cu; L in {32, 64}, d = 5) unless the caller asks for the tile form
  static const bool env_tile = getenv("PL_MEANFIELD_TILE_KERNEL") != nullptr;   // A/B timing knob
  const bool use_ws = use_tc && !env_tile && !(flags & PL_MEANFIELD_TILE_KERNEL) &&
                      pl::mf_ws_supported(Lt, L);
  if (use_tc || (fusable && pl::mf_fused_supported(Lt, L))) {
    // One kernel per iteration (pl_fused_tc.cu on the tensor cores, pl_fused.cu on the CUDA
    // cores): P lives in work_dev and is updated in place (a tile reads its own rows of
    // P_prev into registers before it writes them).
    // small lattices: the blur has no grid barrier and cannot zero its own input, so three value
    // buffers rotate (result / zeroed next target / free) instead of two
    const bool rotate = pl::lattice_blur_is_flow(Lt, L);
    PL_TRY(pl::lattice_workspace(Lt, L, st, rotate));
    float* mu_canon = nullptr;
    if (use_tc) {
      PL_TRY(pl::dev_alloc((void**)&mu_canon, pl::mf_tc_mu_bytes(L), st));
      PL_TRY(pl::mf_tc_prepare_mu(muT_dev, L, mu_canon, st));
    }
    float* res = nullptr;           // buffer holding the blurred lattice of the previous iteration
    float* target = nullptr;        // the splat target: zeroed here (first iteration) or by the previous blur
    float* other = nullptr;         // where the blur may write
    PL_TRY(pl::lattice_splat_buffers(Lt, L, st, &target, &other));
    for (int it = 0; it < niters; it++) {
      if (use_ws) {
        // the first iteration's kernel also fills the blur's sentinel scratch (saves a memset per frame)
        void* fill = nullptr;
        size_t fill_bytes = 0;
        if (it == 0 && rotate) PL_TRY(pl::lattice_flow_prepare(Lt, L, st, &fill, &fill_bytes));
        PL_TRY(pl::mf_ws_iteration(Lt, it == 0, res, target, e0_dev, work_dev, work_dev, mu_canon,
                                   scale, L, st, fill, fill_bytes));
      } else if (use_tc) {
        PL_TRY(pl::mf_tc_iteration(Lt, it == 0, res, target, e0_dev, work_dev, work_dev, mu_canon,
                                   scale, L, st));
      } else {
        PL_TRY(pl::mf_fused_iteration(Lt, it == 0, res, target, e0_dev, work_dev, work_dev,
                                      muT_dev, scale, L, st));
      }
      // the iteration kernel is done with `res`: it is the buffer nothing touches during the blur
      float* spare = rotate ? (res != nullptr ? res : Lt->val[2]) : nullptr;
      float* zeroed = nullptr;
      float* out = nullptr;
      PL_TRY(pl::lattice_blur(Lt, target, other, L, st, &out, it + 1 < niters, spare, &zeroed));
      if (zeroed != nullptr) {       // rotation: result in `other`, next target = the zeroed spare
        float* input = target;
        target = zeroed;
        other = input;
      } else if (out == target) {    // in place: `other` was zeroed (or this was the last iteration)
        float* t = target;
        target = other;
        other = t;
      }                              // else: result in `other`, `target` was zeroed again
      res = out;
    }
    pl::dev_free(mu_canon, st);
    // E = E0 + (slice - P) of the last iteration (crf_module.py:101,103)
    const float denom = 1 + powf(2, -Lt->d);
    if (ro.on() && pl::readout_tiled_supported(Lt, L))
      return pl::readout_tiled(Lt, res, L, denom, work_dev, e0_dev, ro.labels, ro.depth, ro.label, st);
    PL_TRY(pl::slice_tiled(Lt, res, L, denom, e_out_dev, L, work_dev, L, e0_dev, L, false, st));
    if (ro.on())
      return pl::softargmax_rows(e_out_dev, n, L, ro.labels, 1.0f, ro.depth, ro.label, st);
    return PL_OK;
  }
  for (int it = 0; it < niters; it++) {
    // P = scale * Mu softmax(-E)          crf_module.py:99-100,102 with :71-75
    PL_TRY(pl_softmax_compat_ex(it == 0 ? e0_dev : e_out_dev, muT_dev, work_dev, nullptr, n, L,
                                scale, flags & PL_MEANFIELD_NO_TENSOR, stream));
    // E = E0 + (filter(P) - P)            crf_module.py:101 with gaussian_matrix.py:350-352
    PL_TRY(pl::lattice_apply(Lt, work_dev, L, e_out_dev, L, L, work_dev, L, e0_dev, L, flags, st));
  }
  if (ro.on()) return pl::softargmax_rows(e_out_dev, n, L, ro.labels, 1.0f, ro.depth, ro.label, st);
  return PL_OK;
}

int pl_meanfield(pl_lattice* lat, const float* e0_dev, const float* muT_dev, float* e_out_dev,
                 float* work_dev, int L, int niters, float scale, uint32_t flags, void* stream) {
  return meanfield_impl(lat, e0_dev, muT_dev, e_out_dev, work_dev, L, niters, scale, flags, stream,
                        Readout());
}

int pl_meanfield_depth(pl_lattice* lat, const float* e0_dev, const float* muT_dev,
                       float* depth_out_dev, int32_t* label_out_dev, float* work_dev,
                       const float* labels_dev, int L, int niters, float scale, uint32_t flags,
                       void* stream) {
  if (lat == nullptr || (depth_out_dev == nullptr && label_out_dev == nullptr) || L <= 0) {
    pl::set_last_error("pl_meanfield_depth: bad arguments");
    return PL_ERR_INVALID;
  }
  Lattice* Lt = reinterpret_cast<Lattice*>(lat);
  cudaStream_t st = (cudaStream_t)stream;
  Readout ro;
  ro.depth = depth_out_dev; ro.label = label_out_dev; ro.labels = labels_dev;
  // The fused read-out needs no E buffer; the other paths (niters == 0, label counts or flags
  // the fused iteration does not take) write E to a temporary first.
  const bool aligned = (((uintptr_t)e0_dev | (uintptr_t)work_dev | (uintptr_t)muT_dev) & 15u) == 0;
  const bool fused = niters > 0 && aligned &&
                     !(flags & (PL_APPLY_FLAT_KERNELS | PL_APPLY_EXACT_SLICE | PL_MEANFIELD_UNFUSED)) &&
                     ((!(flags & PL_MEANFIELD_NO_TENSOR) && pl::mf_tc_supported(Lt, L)) ||
                      pl::mf_fused_supported(Lt, L)) &&
                     pl::readout_tiled_supported(Lt, L);
  if (fused)   // e_out is never written on this path: any non-null pointer distinct from e0
    return meanfield_impl(lat, e0_dev, muT_dev, work_dev, work_dev, L, niters, scale, flags, stream, ro);
  float* e_tmp = nullptr;
  PL_TRY(pl::dev_alloc((void**)&e_tmp, sizeof(float) * (size_t)Lt->n * L, st));
  int r = meanfield_impl(lat, e0_dev, muT_dev, e_tmp, work_dev, L, niters, scale, flags, stream, ro);
  if (r == PL_OK && niters == 0)
    r = pl::softargmax_rows(e_tmp, Lt->n, L, labels_dev, 1.0f, depth_out_dev, label_out_dev, st);
  pl::dev_free(e_tmp, st);
  return r;
}

int pl_softargmax(const float* rows_dev, int64_t n, int L, const float* labels_dev, float sign,
                  float* depth_out_dev, int32_t* label_out_dev, void* stream) {
  if (rows_dev == nullptr || n < 0 || L <= 0 || (depth_out_dev == nullptr && label_out_dev == nullptr)) {
    pl::set_last_error("pl_softargmax: bad arguments");
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  return pl::softargmax_rows(rows_dev, n, L, labels_dev, sign, depth_out_dev, label_out_dev,
                             (cudaStream_t)stream);
}

int pl_cost_volume(const float* left_dev, const float* right_dev, float* out_dev, int h, int w,
                   int channels, int max_disp, int window, int criterion, void* stream) {
  if (left_dev == nullptr || right_dev == nullptr || out_dev == nullptr) {
    pl::set_last_error("pl_cost_volume: null pointer");
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  return pl::cost_volume(left_dev, right_dev, out_dev, h, w, channels, max_disp, window, criterion,
                         (cudaStream_t)stream);
}

int pl_meanfield_backward_step(float* dp_dev, const float* q_dev, const float* mu_dev, float* de0_dev,
                               float* dmu_dev, int64_t n, int L, float scale, void* stream) {
  if (dp_dev == nullptr || q_dev == nullptr || mu_dev == nullptr || n < 0 ||
      (((uintptr_t)dp_dev | (uintptr_t)q_dev | (uintptr_t)mu_dev | (uintptr_t)de0_dev) & 15u) != 0) {
    pl::set_last_error("pl_meanfield_backward_step: null or misaligned pointer");
    return PL_ERR_INVALID;
  }
  if (!pl::mf_backward_step_supported(L)) {
    pl::set_last_error("pl_meanfield_backward_step: L = %d not supported (use pl_softmax_backward)", L);
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  return pl::mf_backward_step(dp_dev, q_dev, mu_dev, de0_dev, dmu_dev, n, L, scale, (cudaStream_t)stream);
}

int pl_softmax_backward(const float* dq_dev, const float* q_dev, float* de_dev, float* acc_dev, int64_t n,
                        int L, float scale, void* stream) {
  if (dq_dev == nullptr || q_dev == nullptr || de_dev == nullptr || n < 0 || L <= 0) {
    pl::set_last_error("pl_softmax_backward: bad arguments");
    return PL_ERR_INVALID;
  }
  PL_TRY(pl::check_device());
  return pl::softmax_backward(dq_dev, q_dev, de_dev, acc_dev, n, L, scale, (cudaStream_t)stream);
}

int pl_nchw_to_nlc(const float* src_dev, float* dst_dev, int64_t batch, int C, int64_t n,
                   float alpha, void* stream) {
  return pl::nchw_to_nlc(src_dev, dst_dev, batch, C, n, alpha, (cudaStream_t)stream);
}
int pl_nlc_to_nchw(const float* src_dev, float* dst_dev, int64_t batch, int C, int64_t n,
                   float alpha, void* stream) {
  return pl::nlc_to_nchw(src_dev, dst_dev, batch, C, n, alpha, (cudaStream_t)stream);
}

int pl_lattice_export(const pl_lattice* lat, int what, void* dst_dev, void* stream) {
  if (lat == nullptr || dst_dev == nullptr) return PL_ERR_INVALID;
  const_cast<Lattice*>(reinterpret_cast<const Lattice*>(lat))->last_stream = (cudaStream_t)stream;
  return pl::lattice_export(reinterpret_cast<const Lattice*>(lat), what, dst_dev,
                            (cudaStream_t)stream);
}

void pl_profile(int enable) { pl::g_prof_on = enable != 0; }

int pl_profile_report(char* buf, int cap) {
  // one line per kernel: "<name>\t<launches>\t<total_ms>"; clears the records
  struct Agg { std::string name; int count; double ms; };
  std::vector<Agg> agg;
  for (auto& r : pl::g_prof) {
    cudaEventSynchronize(r.b);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.a, r.b);
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
    std::string nm(r.name);
    size_t i = 0;
    for (; i < agg.size(); i++) if (agg[i].name == nm) break;
    if (i == agg.size()) agg.push_back({nm, 0, 0.0});
    agg[i].count++;
    agg[i].ms += ms;
  }
  pl::g_prof.clear();
  std::string out;
  char line[256];
  for (auto& a : agg) {
    snprintf(line, sizeof(line), "%s\t%d\t%.6f\n", a.name.c_str(), a.count, a.ms);
    out += line;
  }
  if (buf != nullptr && cap > 0) {
    strncpy(buf, out.c_str(), (size_t)cap - 1);
    buf[cap - 1] = 0;
  }
  return (int)out.size();
}

uint64_t pl_launch_count(void) { return pl::g_launches; }
int pl_debug_tc_timeouts(void) { return pl::tc_timeouts() + pl::ctc_timeouts() + pl::ws_timeouts(); }
int pl_version(void) { return PL_VERSION; }

const char* pl_error_string(int code) {
  switch (code) {
    case PL_OK: return "ok";
    case PL_ERR_INVALID: return "invalid argument";
    case PL_ERR_CUDA: return "CUDA runtime error";
    case PL_ERR_OVERFLOW: return "lattice hash table overflow";
    case PL_ERR_NODEVICE: return "no sm_100 CUDA device";
    default: return "unknown error";
  }
}
const char* pl_last_error(void) { return pl::t_last_error; }

}  // extern "C"
