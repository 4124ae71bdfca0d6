// pl_common.cuh -- shared declarations for libpl_lattice (sm_100a only).
#pragma once
#include <atomic>
#include <vector>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/pl_lattice.h"

#define PL_TILE 128      // points per lattice tile (pl_tiles.cu)
#define PL_SM_COUNT 148  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

namespace pl {

// ---- error plumbing --------------------------------------------------------
void set_last_error(const char* fmt, ...);
extern unsigned long long g_launches;  // kernels launched by this library

#define PL_CUDA(expr)                                                                   \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      pl::set_last_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                  \
                         cudaGetErrorString(_e));                                       \
      return PL_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)

#define PL_TRY(expr)              \
  do {                            \
    int _r = (expr);              \
    if (_r != PL_OK) return _r;   \
  } while (0)

// Every kernel launch goes through this so pl_launch_count() is honest.  With
// pl_profile(1) each launch is additionally bracketed by CUDA events on its
// own stream (bench.py reads per-kernel durations for the roofline that way).
void prof_begin(const char* name, cudaStream_t st);
void prof_end(cudaStream_t st);
extern bool g_prof_on;

#define PL_LAUNCH(kernel, grid, block, smem, stream, ...)          \
  do {                                                             \
    if (pl::g_prof_on) pl::prof_begin(#kernel, (stream));          \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);    \
    if (pl::g_prof_on) pl::prof_end((stream));                     \
    pl::g_launches++;                                              \
    PL_CUDA(cudaGetLastError());                                   \
  } while (0)

// Programmatic dependent launch: a kernel launched with `pdl` may start while the kernel in front of
// it in the stream is still running; it must not read anything that kernel writes (nor write what it
// reads) before pdl_wait(), which returns once that kernel has completed and its writes are visible.
// pdl_launch_dependents() (after pdl_wait(), so completion stays transitive along the stream) lets
// the NEXT kernel begin its own prologue.  Both are no-ops in a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// PL_NO_PDL (environment, read once) launches everything the ordinary way: A/B timing knob.
inline bool pdl_enabled() {
  static const bool on = getenv("PL_NO_PDL") == nullptr;
  return on;
}

inline cudaError_t launch_ex(const void* fn, dim3 grid, dim3 block, size_t smem, cudaStream_t st, void** args,
                             bool cooperative, bool pdl) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  unsigned na = 0;
  if (cooperative) {
    at[na].id = cudaLaunchAttributeCooperative;
    at[na].val.cooperative = 1;
    na++;
  }
  if (pdl && pdl_enabled()) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    na++;
  }
  cfg.attrs = at;
  cfg.numAttrs = na;
  return cudaLaunchKernelExC(&cfg, fn, args);
}

// PL_LAUNCH for a kernel that calls pdl_wait(): launched so that it may start under the tail of the
// kernel in front of it.  Only where that kernel is one of ours (not a memset or a copy).
inline cudaLaunchConfig_t pdl_config(dim3 grid, dim3 block, size_t smem, cudaStream_t st, cudaLaunchAttribute* at) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cfg;
}
#define PL_LAUNCH_PDL(kernel, grid, block, smem, stream, ...)                                       \
  do {                                                                                              \
    if (pl::g_prof_on) pl::prof_begin(#kernel, (stream));                                           \
    cudaLaunchAttribute at_[1];                                                                     \
    cudaLaunchConfig_t cfg_ = pl::pdl_config(dim3(grid), dim3(block), (smem), (stream), at_);       \
    cudaError_t e_ = cudaLaunchKernelEx(&cfg_, kernel, __VA_ARGS__);                                \
    if (pl::g_prof_on) pl::prof_end((stream));                                                      \
    pl::g_launches++;                                                                               \
    PL_CUDA(e_);                                                                                    \
  } while (0)

// Opt-in attributes (cudaFuncSetAttribute) are per DEVICE: a process that drives several GPUs has
// to set them once on each.  `static PerDeviceOnce once; if (once.need()) { ... }`
// The flag is also per THREAD: the caller's block has no completion signal, so with a shared flag a
// second host thread could find it set and launch before the first thread had finished setting the
// attribute ("invalid argument" for a kernel that needs > 48 KB of shared memory: seen once with the
// two host threads of the pipelined end-to-end bench).  Setting an attribute twice is harmless.
struct PerDeviceOnce {
  uint32_t id;
  PerDeviceOnce() {
    static std::atomic<uint32_t> next{0};
    id = next.fetch_add(1);
  }
  bool need() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
    thread_local std::vector<uint8_t> seen;     // [id * 64 + device] for the calling thread
    const size_t k = (size_t)id * 64 + (size_t)dev;
    if (seen.size() <= k) seen.resize(k + 64, 0);
    if (seen[k]) return false;
    seen[k] = 1;
    return true;
  }
};
static inline int current_device_index() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  return dev;
}

static inline unsigned int ceil_div_u(uint64_t a, uint64_t b) { return (unsigned int)((a + b - 1) / b); }

// Grid for a grid-stride kernel: enough CTAs for `items` but capped at a
// multiple of the SM count (waves of 148 x ctas_per_sm).
static inline unsigned int grid_for(uint64_t items, unsigned int block, unsigned int ctas_per_sm) {
  uint64_t need = (items + block - 1) / block;
  uint64_t cap = (uint64_t)PL_SM_COUNT * ctas_per_sm;
  if (need < 1) need = 1;
  return (unsigned int)(need < cap ? need : cap);
}

// ---- the lattice handle ----------------------------------------------------
struct Lattice {
  int64_t n = 0;       // points
  int d = 0;           // feature dims
  int64_t M = 0;       // vertices
  uint32_t flags = 0;
  int device = 0;
  // stream of the most recent API call that used this lattice: pl_lattice_destroy releases the
  // buffers in ITS order (cudaFreeAsync), so kernels still queued there finish first
  cudaStream_t last_stream = nullptr;
  // persistent device arrays
  int32_t* vid = nullptr;    // [n][d+1] vertex id per (point, remainder)
  float* bary = nullptr;     // [n][d+1] barycentric weight per (point, remainder)
  int16_t* greedy = nullptr; // [n][d+1]
  int8_t* rank = nullptr;    // [n][d+1]
  int16_t* vkeys = nullptr;  // [M][d]
  int2* nbr = nullptr;       // [d+1][M] {minus, plus} neighbour ids; M = absent (zero row)
  // tile-local view (pl_tiles.cu): per tile of PL_TILE points, entries sorted by vertex
  int64_t ntiles = 0;
  uint8_t* ent_pix = nullptr;  // [ntiles*TE] point index within the tile, sorted order
  float* ent_w = nullptr;      // [ntiles*TE] barycentric weight, sorted order
  uint16_t* loc = nullptr;     // [ntiles*TE] per original entry: index into the tile's vertex list
  int2* useg = nullptr;        // per tile, per distinct vertex: {vertex id, start | count << 16}
  int* tile_off = nullptr;     // [ntiles] offset of the tile's list in useg
  int* tile_U = nullptr;       // [ntiles] distinct vertices in the tile (inside tile_off's block)
  // the same lists packed per tile for single bulk (TMA) copies, pads zeroed (pl_fused_tc.cu):
  //   rec_slice[tile] = { bary f32[TE] | loc u16[TE] }                      (entry order)
  //   rec_splat[tile] = { ent_w f32[TE] | vertex id i32[TE] | ent_pix u8[TE] }  (sorted order)
  uint8_t* rec_slice = nullptr;
  uint8_t* rec_splat = nullptr;
  // largest distinct-vertex count of a tile: reduced on the device after the tile build, copied
  // to a pinned slot, read (after the event) by the first tiled slice to size its staging buffer
  int max_U = -1;
  int mean_U = 0;              // mean distinct-vertex count per tile (same read-back)
  int* h_max_U = nullptr;
  cudaEvent_t max_U_ev = nullptr;
  cudaEvent_t aux_ev = nullptr;     // marks the point of the caller's stream a side-stream copy has to wait for
  // grow-only value workspace: two ping-pong buffers of (M+1) x ws_C floats, and a third one for the
  // mean-field loop's buffer rotation on small lattices (lattice_workspace(..., third = true))
  float* val[3] = {nullptr, nullptr, nullptr};
  int64_t ws_floats = 0;     // floats per buffer
  // scratch of the barrier-free small-lattice blur (pl_filter.cu blur_flow_kernel): two sets of d
  // pass buffers of M x flow_C floats; set `flow_set` is all-sentinel and is the one the next call uses
  float* flow = nullptr;
  int64_t flow_floats = 0;   // floats per pass buffer
  int flow_set = 0;
  int64_t flow_dirty[2] = {0, 0};   // floats from the start of each pass buffer of a set that may not be sentinel
  int64_t bytes = 0;         // persistent device bytes
};

// build (pl_build.cu).  With `guide` (d = 5) the features are evaluated inside the elevation
// kernel from the RGB planes (ijrgbGuide, crf_module.py:106-114) and `ref` is not read.
struct GuideSrc {
  const float* rgb = nullptr;  // [3][h][w] planes of one image
  int h = 0, w = 0;
  float s_ij = 1.f, s_rgb = 1.f;
  double diag = 1.0;           // sqrt(h^2 + w^2) in float64, as numpy computes it
};
int lattice_build(Lattice* L, const float* ref, cudaStream_t st, const GuideSrc* guide = nullptr);
// tile-local view + tiled kernels (pl_tiles.cu)
int lattice_build_tiles(Lattice* L, cudaStream_t st);
// Small device -> host results (M of the build, max U of the tile build) are copied on a per-device side
// stream behind an event of the caller's stream, so the copy engine works beside the next kernels instead
// of between them (each in-stream copy cost ~9 us of an 820 us frame: the copy plus two gaps).
// Records L->max_U_ev behind the copy; the host waits on that event when it needs the value.
int readback_small(Lattice* L, void* host_dst, const void* dev_src, size_t bytes, cudaStream_t st);
int* pinned_slot();   // pl_tiles.cu: two pinned ints for a small device -> host result, nullptr if unavailable
int splat_tiled(Lattice* L, const float* src, int64_t src_stride, int Cc, float* val,
                cudaStream_t st, const float* rowscale = nullptr, int64_t rs_stride = 0, int J = 1);
int splat_sorted(Lattice* L, const float* src, int64_t src_stride, int Cc, float* val, cudaStream_t st);
int slice_dot_tiled(Lattice* L, const float* val, int Cc, float denom, const float* other,
                    int64_t other_stride, float* out, int64_t out_stride, float coef, cudaStream_t st,
                    int J = 1);
// out[p] += coef * <filter(src * rowscale)[p][:], other[p][:]>  (pl_filter.cu; tiled kernels only)
// J > 1: J row factors rowscale[p][j] at once, out[p][j] (d = 5, J = 5 only)
int lattice_apply_dot(Lattice* L, const float* src, int64_t src_stride, int C, const float* rowscale,
                      int64_t rs_stride, const float* other, int64_t other_stride, float* out,
                      int64_t out_stride, float coef, cudaStream_t st, int J = 1);
int slice_tiled(Lattice* L, const float* val, int Cc, float denom, float* dst, int64_t dst_stride,
                const float* sub, int64_t sub_stride, const float* add, int64_t add_stride,
                bool exact, cudaStream_t st);
// value path (pl_filter.cu)
int lattice_apply(Lattice* L, const float* src, int64_t src_stride, float* dst, int64_t dst_stride,
                  int C, const float* sub, int64_t sub_stride, const float* add, int64_t add_stride,
                  uint32_t flags, cudaStream_t st);
int lattice_workspace(Lattice* L, int Cc, cudaStream_t st, bool third = false);
bool lattice_blur_is_flow(const Lattice* L, int Cc);
int lattice_splat_buffers(Lattice* L, int Cc, cudaStream_t st, float** cur, float** nxt);
// scratch of the barrier-free blur for Cc channels.  If it had to be (re)allocated, its first set must be
// filled with 0xFF bytes before the next blur: with fill_ptr == nullptr that is done here by a memset,
// else the region is handed to the caller (*fill_bytes == 0: nothing to fill).
int lattice_flow_prepare(Lattice* L, int Cc, cudaStream_t st, void** fill_ptr, size_t* fill_bytes);
// `zero_other`: leave a zeroed buffer behind for the next splat -- the one of cur / nxt that does not
// hold the result, unless `zeroed` is given and comes back non-null: then it names the buffer that was
// zeroed instead (`spare`, a third buffer of the same shape that nothing reads while the blur runs).
int lattice_blur(Lattice* L, float* cur, float* nxt, int Cc, cudaStream_t st, float** out,
                 bool zero_other = false, float* spare = nullptr, float** zeroed = nullptr);
// fused mean-field iteration (pl_fused.cu)
bool mf_fused_supported(const Lattice* L, int labels);
int mf_fused_iteration(Lattice* L, bool first, const float* val_in, float* val_out, const float* e0,
                       const float* p_prev, float* p_next, const float* muT, float scale,
                       int labels, cudaStream_t st);
// fused mean-field iteration on the tensor cores (pl_fused_tc.cu), L in {16, 32, 64}
bool mf_tc_supported(const Lattice* L, int labels);   // L == nullptr: stand-alone compat
size_t mf_tc_mu_bytes(int labels);
int mf_tc_prepare_mu(const float* muT, int labels, float* mu_canon, cudaStream_t st);
int mf_tc_iteration(Lattice* L, bool first, const float* val_in, float* val_out, const float* e0,
                    const float* p_prev, float* p_next, const float* mu_canon, float scale,
                    int labels, cudaStream_t st);
int softmax_compat_tc(const float* e, const float* mu_canon, float* p, float* q_out, int64_t n,
                      int labels, float scale, cudaStream_t st);
int tc_timeouts();
// the same iteration, warp-specialised: main group (slice / softmax / MMA / epilogue) and splat
// group hand the P tile over through shared memory (pl_fused_ws.cu), L in {32, 64}, d = 5
bool mf_ws_supported(const Lattice* L, int labels);
int mf_ws_iteration(Lattice* L, bool first, const float* val_in, float* val_out, const float* e0,
                    const float* p_prev, float* p_next, const float* mu_canon, float scale,
                    int labels, cudaStream_t st, void* fill = nullptr, size_t fill_bytes = 0);
int ws_timeouts();
// softmax + compatibility for large label counts / label shards on tcgen05 (pl_compat_tc.cu)
bool compat_tc_supported(int L, int N, int block_w);
size_t compat_tc_mu_bytes(int L, int N);
int compat_tc_prepare_mu(const float* muT, int L, int l0, int N, uint8_t* mu_rec, cudaStream_t st);
int compat_tc_run(const float* e, int nblocks, int block_w, const uint8_t* mu_rec, float* p,
                  float* q_out, int64_t n, int L, int N, float scale, cudaStream_t st,
                  const float* const* block_ptrs = nullptr, float* const* out_block_ptrs = nullptr);
int ctc_timeouts();
// mean field (pl_meanfield.cu)
int softmax_compat(const float* e, const float* muT, float* p, float* q_out, int64_t n, int L,
                   float scale, cudaStream_t st);
// one backward step of the mean-field loop, L = 64 (pl_backward.cu)
bool mf_backward_step_supported(int L);
int mf_backward_step(float* dp, const float* q, const float* mu, float* de0, float* dmu, int64_t n,
                     int L, float scale, cudaStream_t st);
int softmax_backward(const float* dq, const float* q, float* de, float* acc, int64_t n, int L, float scale,
                     cudaStream_t st);
int nchw_to_nlc(const float* src, float* dst, int64_t batch, int C, int64_t n, float alpha,
                cudaStream_t st);
int nlc_to_nchw(const float* src, float* dst, int64_t batch, int C, int64_t n, float alpha,
                cudaStream_t st);

// matching cost volume (pl_cost.cu) and fused depth read-out (pl_readout.cu)
int cost_volume(const float* im1, const float* im2, float* out, int h, int w, int ch, int L, int ws,
                int criterion, cudaStream_t st);
bool readout_tiled_supported(const Lattice* L, int labels);
int readout_tiled(Lattice* L, const float* val, int labels, float denom, const float* p,
                  const float* e0, const float* label_values, float* depth, int* label_out,
                  cudaStream_t st);
int softargmax_rows(const float* e, int64_t n, int labels, const float* label_values, float sign,
                    float* depth, int* label_out, cudaStream_t st);

// stream-ordered allocation helpers (pl_api.cu)
int dev_alloc(void** p, size_t bytes, cudaStream_t st);
void dev_free(void* p, cudaStream_t st);

}  // namespace pl
