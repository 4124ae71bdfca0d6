/*
 * pl_lattice.h -- C ABI of the B200-native permutohedral-lattice filter and
 * dense-CRF mean-field loop (libpl_lattice.so, built from
 * depth-estimation_b200/csrc/ for sm_100a).
 *
 * This is the drop-in boundary for the reference's one native export,
 *     at::Tensor filter(at::Tensor src, at::Tensor ref)
 *         crf/lattice/lite/lattice.cpp:6-16  (pybind module `lattice`,
 *         JIT-loaded by crf/gaussian_matrix.py:15-16 as `latticefilter`),
 * whose body is PermutohedralLattice::filter (crf/lattice/lite/permutohedral.h
 * :199-321): build lattice from ref[n,d] -> splat src[n,C] -> blur -> slice.
 *
 * Plain pointers and sizes only; no torch types.  All `*_dev` pointers are
 * device pointers on the current CUDA device, fp32, row-major; `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  Every call is
 * asynchronous on `stream` unless stated otherwise.  Functions return PL_OK or
 * a PL_ERR_* code; pl_error_string() / pl_last_error() describe it.  There is
 * no CPU fallback anywhere in this library.
 *
 * The reference rebuilds the lattice inside every filter() call
 * (permutohedral.h:208,236-238).  Here the lattice is a handle that is built
 * once per feature matrix and applied many times (every mean-field iteration
 * and the backward pass reuse it); pl_filter() is the one-shot spelling with
 * exactly the reference's call shape.
 */
#ifndef PL_LATTICE_H
#define PL_LATTICE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PL_VERSION 1

enum {
  PL_OK = 0,
  PL_ERR_INVALID = 1,   /* bad argument (null pointer, n<=0, d out of range, ...) */
  PL_ERR_CUDA = 2,      /* a CUDA runtime call failed; see pl_last_error() */
  PL_ERR_OVERFLOW = 3,  /* lattice hash table could not be sized (internal retry exhausted) */
  PL_ERR_NODEVICE = 4   /* no CUDA device / not an sm_100 device */
};

#define PL_MAX_D 16 /* feature dimensions supported (reference: char rank, d <= 127) */

typedef struct pl_lattice pl_lattice; /* opaque */

/* ---- lattice handle ------------------------------------------------------- */

/* pl_lattice_create flags */
#define PL_BUILD_DEFAULT 0u
/* Reproduce the reference's hash-table growth quirk (permutohedral.h:102-103
 * with :62: the lookup that triggers a table doubling probes from a slot
 * computed with the old capacity, which can create an extra duplicate vertex
 * per doubling; after the next doubling the duplicate is the copy lookups
 * find).  On: vertex count, vertex ids, keys and blur neighbours equal the
 * reference's bit for bit (replayed on the device from the clean build, see
 * csrc/pl_build.cu section 6).  Off: one vertex per distinct key (the
 * algorithm as published).  The two are identical whenever the reference
 * table never grows (M < 16383).  pl_filter / pl_filter_host always use it. */
#define PL_BUILD_REF_GROW_QUIRK 1u

/* Builds the lattice for feature vectors ref_dev[n][d] (what the reference's
 * splat() derives from `position`: permutohedral.h:376-450 minus the value
 * accumulation, plus the neighbour table blur() looks up at :503-522).
 * Synchronises `stream` once (the vertex count has to reach the host). */
int pl_lattice_create(const float* ref_dev, int64_t n, int d, uint32_t flags, void* stream,
                      pl_lattice** out);
/* Same lattice from an RGB image: the features of ijrgbGuide.forward (crf_module.py:106-114),
 *   (i / diag / s_ij, j / diag / s_ij, r / s_rgb, g / s_rgb, b / s_rgb),  diag = sqrt(h^2 + w^2)
 * are evaluated inside the elevation kernel with the reference's rounding (float64 mgrid / diag
 * cast to fp32, then fp32 divisions), so the [n][5] feature matrix is never materialised.
 * rgb_planes_dev: [3][h][w] fp32 (one image of the module's NCHW input). */
int pl_lattice_create_ijrgb(const float* rgb_planes_dev, int h, int w, float s_ij, float s_rgb,
                            uint32_t flags, void* stream, pl_lattice** out);
/* Releases the lattice's device memory in stream order on the stream of the most recent call
 * that used it (kernels still queued there finish first); safe right after an asynchronous call. */
void pl_lattice_destroy(pl_lattice* lat);

int64_t pl_lattice_num_points(const pl_lattice* lat);   /* n */
int pl_lattice_dim(const pl_lattice* lat);              /* d */
int64_t pl_lattice_num_vertices(const pl_lattice* lat); /* M == hashTable.size() (permutohedral.h:45) */
int64_t pl_lattice_device_bytes(const pl_lattice* lat); /* persistent device memory held */

/* pl_lattice_apply flags */
#define PL_APPLY_DEFAULT 0u
/* Evaluate slice exactly as permutohedral.h:480 does (per-term fp32 divide by
 * 1+2^-d, no FMA contraction) instead of folding the constant into the
 * barycentric weight. */
#define PL_APPLY_EXACT_SLICE 1u
/* Use the flat per-entry splat / slice kernels instead of the tile-local ones
 * (pl_tiles.cu); for A/B measurements and tests. */
#define PL_APPLY_FLAT_KERNELS 2u
/* pl_meanfield only: run the separate softmax_compat / splat / blur / slice
 * kernels instead of the fused per-iteration kernel (pl_fused.cu). */
#define PL_MEANFIELD_UNFUSED 4u
/* pl_meanfield / pl_softmax_compat: keep the compatibility transform on the CUDA cores
 * (fp32 FMA, k ascending) instead of the tcgen05 3xTF32 path used for L in {16, 32, 64}. */
#define PL_MEANFIELD_NO_TENSOR 8u
/* pl_meanfield: use the tile form of the tcgen05 iteration kernel (pl_fused_tc.cu, two CTAs per
 * SM) even where the warp-specialised form (pl_fused_ws.cu; L in {32, 64}, d = 5) exists. */
#define PL_MEANFIELD_TILE_KERNEL 16u

/* dst[p][c] = filter(src)[p][c] - (sub ? sub[p][c] : 0) + (add ? add[p][c] : 0)
 *
 * filter = splat (permutohedral.h:452-455) -> blur (:486-548) -> slice
 * (:473-483) on the cached lattice; identical to the reference's
 * filter(src, ref) when sub == add == NULL.  `sub = src` gives the adjacency
 * form W(x) = filter(x) - x of crf/gaussian_matrix.py:302,352; `add = E0`
 * fuses the mean-field update E = E0 + W(..) of crf/crf_module.py:101.
 * src/dst/sub/add are [n][C] with row strides given in floats (>= C);
 * dst may alias sub or add but not src. */
int pl_lattice_apply(pl_lattice* lat, const float* src_dev, int64_t src_stride, float* dst_dev,
                     int64_t dst_stride, int C, const float* sub_dev, int64_t sub_stride,
                     const float* add_dev, int64_t add_stride, uint32_t flags, void* stream);

/* The filter contracted on the spot, for the guide gradient of LatticeFilter.backward
 * (crf/gaussian_matrix.py:450-463), whose filtered outer products `filter(g (x) ref)` are only ever
 * used inside inner products with the rows of src / g:
 *   out[p * out_stride] += coef * sum_c filter(x)[p][c] * other[p][c],   x[p][c] = src[p][c] * rowscale[p * rowscale_stride]
 * (rowscale_dev NULL: x = src).  Neither x nor filter(x) is written to memory.  C % 4 == 0, 16-byte aligned
 * rows, lattice built with its tile view (the default); otherwise PL_ERR_INVALID. */
int pl_lattice_apply_dot(pl_lattice* lat, const float* src_dev, int64_t src_stride, int C,
                         const float* rowscale_dev, int64_t rowscale_stride, const float* other_dev,
                         int64_t other_stride, float* out_dev, int64_t out_stride, float coef, void* stream);
/* The same for J row factors at once (the d = 5 features of the guide, J = 5 only; d = 5 lattices):
 *   out[p * out_stride + j] += coef * <filter(src * rowscale[p * rowscale_stride + j])[p], other[p]>,  j < J.
 * The source rows are staged and gathered once for all J sources, the rows of `other` once for all J
 * contractions. */
int pl_lattice_apply_dot_multi(pl_lattice* lat, const float* src_dev, int64_t src_stride, int C,
                               const float* rowscale_dev, int64_t rowscale_stride, int J,
                               const float* other_dev, int64_t other_stride, float* out_dev,
                               int64_t out_stride, float coef, void* stream);

/* ---- the reference's call shape ------------------------------------------- */

/* lattice.cpp:6-10  filter(src[n,C], ref[n,d]) -> dst[n,C], device buffers. */
int pl_filter(const float* src_dev, const float* ref_dev, float* dst_dev, int64_t n, int C, int d,
              void* stream);

/* Same with HOST buffers (what the reference's CPU op takes and returns):
 * copies src/ref to the device, filters, copies dst back, synchronises.
 * Pinned host memory makes the copies asynchronous; pageable works too. */
int pl_filter_host(const float* src_host, const float* ref_host, float* dst_host, int64_t n, int C,
                   int d, void* stream);

/* ---- mean-field pieces (crf/crf_module.py:41-53, 66-75, 93-104) ----------- */

/* P[p][l] = scale * sum_k Mu[l][k] * softmax_k(-E[p][.])[k]
 * i.e. Q = F.softmax(-E, dim=1) followed by charb.forward's 1x1 conv2d with
 * the L x L compatibility matrix times exp(s) (crf_module.py:71-75), in
 * pixel-major layout.  muT_dev is Mu transposed, [L(k)][L(l)] row-major.
 * q_out_dev (optional, may be NULL) receives Q. */
int pl_softmax_compat(const float* e_dev, const float* muT_dev, float* p_dev, float* q_out_dev,
                      int64_t n, int L, float scale, void* stream);
/* Same with flags (PL_MEANFIELD_NO_TENSOR). */
int pl_softmax_compat_ex(const float* e_dev, const float* muT_dev, float* p_dev, float* q_out_dev,
                         int64_t n, int L, float scale, uint32_t flags, void* stream);

/* Label-sharded form of the same transform (label-channel sharding, one all-gather
 * per mean-field iteration).  The caller holds E for ALL L labels in the blocked layout an
 * all-gather over G label shards produces, e_blocks_dev[b][p][w] with b < nblocks,
 * w < L / nblocks, label = b * (L / nblocks) + w (nblocks = 1: plain [n][L]); the softmax runs
 * over all L labels and the rank computes its own N output labels l0 .. l0+N-1:
 *   p_dev[p][j] = scale * sum_k Mu[l0 + j][k] * softmax_k(-E[p][.])[k],  p_dev is [n][N].
 * muT_dev is the full transposed matrix [L(k)][L(l)].  tcgen05 path (3xTF32) when
 * L % 64 == 0 and N % 32 == 0, N <= 256; otherwise PL_ERR_INVALID. */
int pl_softmax_compat_blocked(const float* e_blocks_dev, int nblocks, const float* muT_dev,
                              float* p_dev, int64_t n, int L, int l0, int N, float scale,
                              void* stream);

/* pl_softmax_compat_blocked with the G label blocks at G separate addresses (host array of device
 * pointers, each [n][L/G]).  The pointers may be PEER-GPU memory opened with pl_peer_open: every
 * rank then reads the other ranks' E shards in place over NVLink inside the kernel that consumes
 * them -- no all-gathered copy is written to or re-read from HBM.  The caller orders the ranks
 * (the shards must be complete before the call: one tiny NCCL collective on the same stream). */
int pl_softmax_compat_peers(const float* const* e_block_ptrs, int nblocks, const float* muT_dev,
                            float* p_dev, int64_t n, int L, int l0, int N, float scale,
                            void* stream);

/* The pixel / label TRANSPOSE form of the sharded iteration (softmax and compatibility
 * couple labels but not pixels, the filter couples pixels but not labels): every rank holds label
 * block b of E and of P for ALL pixels, e_block_ptrs[b] / p_block_ptrs[b] = [n][L / nblocks], its own
 * and -- through pl_peer_open -- its peers'.  This call does the softmax + compatibility transform of
 * pixel rows row0 .. row0 + nrows - 1 (this rank's pixel shard) for all L labels, reading the rows
 * from every block in place and writing the P rows straight into the owners' blocks:
 *   P[b][p][j] = scale * sum_k Mu[b * bw + j][k] * softmax_k(-E[.][p][.])[k],   row0 <= p < row0 + nrows.
 * Per rank and iteration n/G * L floats cross NVLink in each direction instead of the n * L floats of
 * the all-gather form, and no rank repeats another's softmax.  L % 64 == 0, L <= 256, 16-byte aligned
 * blocks; the caller orders the ranks around the call (E complete before, P complete after). */
int pl_softmax_compat_shards(const float* const* e_block_ptrs, float* const* p_block_ptrs, int nblocks,
                             const float* muT_dev, int64_t row0, int64_t nrows, int L, float scale,
                             void* stream);

/* Device buffers other processes on this node can map (CUDA IPC; one process per GPU).
 * pl_peer_alloc: cudaMalloc + export handle (PL_IPC_HANDLE_BYTES opaque bytes to send to the peers).
 * pl_peer_open : map a peer's buffer into this process (not valid in the exporting process).
 * pl_peer_close / pl_peer_free: unmap / free. */
#define PL_IPC_HANDLE_BYTES 64
int pl_peer_alloc(size_t bytes, void** dev_ptr_out, void* ipc_handle_out);
int pl_peer_open(const void* ipc_handle, void** dev_ptr_out);
int pl_peer_close(void* dev_ptr);
int pl_peer_free(void* dev_ptr);

/* pl_softmax_compat for callers that cannot allocate: code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc.  (No counterpart in the reference.)
 * Same arguments, same result; L in {16, 32, 64} only.  The scratch the
 * transform needs (the compatibility matrix split into TF32 hi / lo operand planes, 33 280 bytes at
 * L = 64) lives in a file-scope `static __device__` array inside the library, which every GPU that
 * loads the library holds until the process exits; the call performs no cudaMalloc / cudaFree /
 * pool allocation of any kind.  L > PL_NOALLOC_MAX_LABELS is rejected with PL_ERR_INVALID.
 * All calls share that one array per device (it keeps the last call's contents; each call rewrites
 * all of it before use, in stream order): calls that OVERLAP on the GPU race -- serialise them. */
#define PL_NOALLOC_MAX_LABELS 64
int pl_softmax_compat_noalloc(const float* e_dev, const float* muT_dev, float* p_dev, float* q_out_dev,
                              int64_t n, int L, float scale, void* stream);

/* Whole inference loop of CRFasRNN.forward (crf_module.py:93-104) with
 * W = lattice adjacency (gaussian_matrix.py:341-352), pixel-major:
 *   E = E0; repeat niters: P = scale * Mu softmax(-E); E = E0 + filter(P) - P
 * e0_dev, e_out_dev: [n][L].  Returns E of the last iteration (the module
 * returns -E).  work_dev: caller-provided scratch of n*L floats. */
int pl_meanfield(pl_lattice* lat, const float* e0_dev, const float* muT_dev, float* e_out_dev,
                 float* work_dev, int L, int niters, float scale, uint32_t flags, void* stream);

/* pl_meanfield followed by the depth read-out of crf/mb_stereo_crf.py:62-66
 * (logits2average_depth: sum_l softmax(-E)_l * labels[l]) and / or the argmax label of the
 * notebooks (Experiments/trainableDenseCRF.ipynb cell 8), fused into the last slice so the
 * final E never reaches HBM.  depth_out_dev [n] f32 and label_out_dev [n] i32 may each be NULL;
 * labels_dev [L] f32 or NULL for labels 0..L-1 (mb_stereo_crf.py:64).  work_dev: n*L floats. */
int pl_meanfield_depth(pl_lattice* lat, const float* e0_dev, const float* muT_dev,
                       float* depth_out_dev, int32_t* label_out_dev, float* work_dev,
                       const float* labels_dev, int L, int niters, float scale, uint32_t flags,
                       void* stream);
/* The same read-out of rows already in HBM: rows_dev [n][L]; sign = +1 rows are energies E,
 * sign = -1 rows are logits (softmax(logits), mb_stereo_crf.py:63). */
int pl_softargmax(const float* rows_dev, int64_t n, int L, const float* labels_dev, float sign,
                  float* depth_out_dev, int32_t* label_out_dev, void* stream);

/* Plane-sweep matching cost volume, the unaries the CRF reads: crf/depth.py:34-53
 * (disparity_badness).  left_dev / right_dev: [h][w][channels] fp32 (the reference's numpy image
 * layout); out_dev: [h*w][max_disp] fp32, pixel-major -- the layout pl_meanfield consumes.
 *   out[y][x][i] = window-sum (window x window, scipy 'reflect' boundary) of
 *                  sum_c crit(left[y][x][c], right[y][x - i][c])   (0 where x - i < 0)
 * criterion: 0 = AD |a-b| (depth.py:20), 1 = SD (a-b)^2 (:18), 2 = nprod -a*b (:22).
 * window odd, <= 17.  The reference uses max_disp = w / 6 (depth.py:40). */
#define PL_COST_AD 0
#define PL_COST_SD 1
#define PL_COST_NPROD 2
int pl_cost_volume(const float* left_dev, const float* right_dev, float* out_dev, int h, int w,
                   int channels, int max_disp, int window, int criterion, void* stream);

/* One backward step of the mean-field loop in one pass (L = 64; other L: PL_ERR_INVALID, compose
 * pl_softmax_backward with two GEMMs): what autograd derives for crf_module.py:99-102 after the filter's
 * own backward (gaussian_matrix.py:446) has produced dP = K dE - dE:
 *   dQ = dP Mu;   dmu += dP^T Q (unscaled, atomically);   dE = -scale Q (dQ - rowsum(dQ Q));   de0 += dE
 * dp_dev [n][L]: dP in, dE out (in place); q_dev [n][L]: the iteration's softmax; mu_dev [L][L]: Mu[j][k]
 * row-major; de0_dev [n][L] or NULL; dmu_dev [L][L] or NULL. */
int pl_meanfield_backward_step(float* dp_dev, const float* q_dev, const float* mu_dev, float* de0_dev,
                               float* dmu_dev, int64_t n, int L, float scale, void* stream);

/* Backward of one mean-field iteration's softmax, what autograd derives for
 * `Q = F.softmax(-E, dim=1)` followed by the compatibility scale (crf_module.py:99-102, charb.forward :75):
 *   de[p][k] = -scale * q[p][k] * (dq[p][k] - sum_k' dq[p][k'] q[p][k']);   acc[p][k] += de[p][k]
 * dq_dev, q_dev, de_dev: [n][L] fp32 (de_dev may alias dq_dev); acc_dev: [n][L] running unary gradient
 * or NULL. */
int pl_softmax_backward(const float* dq_dev, const float* q_dev, float* de_dev, float* acc_dev, int64_t n,
                        int L, float scale, void* stream);

/* Layout helpers for the module surface (NCHW tensors <-> pixel-major):
 * dst[b][p][c] = alpha * src[b][c][p]   and back. */
int pl_nchw_to_nlc(const float* src_dev, float* dst_dev, int64_t batch, int C, int64_t n,
                   float alpha, void* stream);
int pl_nlc_to_nchw(const float* src_dev, float* dst_dev, int64_t batch, int C, int64_t n,
                   float alpha, void* stream);

/* ---- inspection (parity tests read the integer artefacts through these) --- */

enum {
  PL_EXPORT_GREEDY = 0,   /* int16 [n][d+1]  permutohedral.h:398-399,418,428 */
  PL_EXPORT_RANK = 1,     /* int8  [n][d+1]  permutohedral.h:407-433 */
  PL_EXPORT_BARY = 2,     /* f32   [n][d+1]  permutohedral.h:436-441 (replay weight :459) */
  PL_EXPORT_VERTEX = 3,   /* int32 [n][d+1]  replay offset / vd (:458), first-touch numbering */
  PL_EXPORT_KEYS = 4,     /* int16 [M][d]    hashTable.getKeys() (:48) */
  PL_EXPORT_NEIGHBOURS = 5/* int32 [d+1][M][2], -1 = absent (:516-522) */
};
/* Copies one artefact into dst_dev (device memory, sized as listed). */
int pl_lattice_export(const pl_lattice* lat, int what, void* dst_dev, void* stream);

/* ---- bookkeeping ---------------------------------------------------------- */

/* Per-launch timing for the roofline report: pl_profile(1) brackets every
 * kernel launch with CUDA events on its stream; pl_profile_report() waits for
 * them, writes one "<kernel>\t<launches>\t<total_ms>" line per kernel into
 * buf (NUL-terminated, truncated to cap), clears the records and returns the
 * untruncated length.  Not for use inside a timed region. */
void pl_profile(int enable);
int pl_profile_report(char* buf, int cap);

/* Number of kernels this library has launched in this process so far. */
uint64_t pl_launch_count(void);
/* Number of bounded mbarrier waits that timed out inside the tensor-core kernels since the
 * library was loaded (0 unless something is badly wrong; synchronises the device). */
int pl_debug_tc_timeouts(void);
int pl_version(void);
const char* pl_error_string(int code);
const char* pl_last_error(void); /* detail of the most recent failure on this thread */

#ifdef __cplusplus
}
#endif
#endif /* PL_LATTICE_H */
