/*
 * tsplat.h — C-ABI of libtsplat.so, the sm_100a differentiable Gaussian-splat rasterizer.
 *
 * This is the drop-in boundary for the render path of TranSplat
 * (/root/reference/src/model/decoder/cuda_splatting.py).  Every entry point below replaces
 * one binding of the external extension `diff_gaussian_rasterization`
 * (requirements.txt:17) that the reference reaches through
 *   - cuda_splatting.py:6-9      import GaussianRasterizationSettings, GaussianRasterizer
 *   - cuda_splatting.py:108-133  settings construction + rasterizer(...) call (render_cuda)
 *   - cuda_splatting.py:201-226  same call pattern (render_cuda_orthographic)
 *   - cuda_splatting.py:323-348  same call pattern (render_cuda_with_stats; reads radii)
 *   - cuda_splatting.py:404      render_depth_cuda -> a second render_cuda pass
 * and, one level up, the batched decoder call
 *   - decoder_splatting_cuda.py:41-97  DecoderSplattingCUDA.forward / render_depth.
 *
 * Conventions
 *   - plain C, no torch types; every pointer is a DEVICE pointer owned by the caller unless the
 *     parameter name ends in _host;
 *   - the library never allocates or frees device memory, never synchronises the device and
 *     only enqueues work on the stream it is given;
 *   - return value 0 = ok, otherwise a TSPLAT_E* code; tsplat_last_error() gives the text
 *     (thread-local);
 *   - the size of the tile/Gaussian pair list ("num_rendered", D) is not known on the host:
 *     the caller supplies a capacity d_cap; if D > d_cap the status block's `overflow` flag is
 *     set, pairs beyond the capacity are dropped, and the caller retries with a larger
 *     workspace (tsplat_status_read tells it how large).  A forward whose lists overflowed also
 *     POISONS its outputs on the device (NaN in the first colour value and in the fused loss), so a
 *     training loop cannot consume a truncated step unnoticed.
 */
#ifndef TSPLAT_H_
#define TSPLAT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TSPLAT_VERSION 200

/* error codes */
#define TSPLAT_OK 0
#define TSPLAT_EINVAL 1    /* bad shape / null pointer / unsupported option          */
#define TSPLAT_ENOSPC 2    /* workspace too small for (problem, d_cap)                */
#define TSPLAT_ECUDA 3     /* a CUDA runtime call or launch failed                    */

/* tile geometry is part of the parity contract (tiles_touched, keys, ranges depend on it) */
#define TSPLAT_TILE 16

/* layouts of the spherical-harmonics tensor */
#define TSPLAT_SH_GM3 0 /* [G, M, 3]  what the reference passes as `shs` (cuda_splatting.py:84,129) */
#define TSPLAT_SH_G3M 1 /* [G, 3, M]  Gaussians.harmonics as the encoder emits it (types.py:11)     */
/* layouts of the covariance tensor */
#define TSPLAT_COV_TRIU6 0 /* [G, 6] xx,xy,xz,yy,yz,zz (cuda_splatting.py:124,132)  */
#define TSPLAT_COV_FULL9 1 /* [G, 3, 3] Gaussians.covariances (types.py:10)          */

/* problem flags */
#define TSPLAT_FLAG_REZERO_GRAD_ACC 1 /* tsplat_backward: clear the screen-space accumulators first.  The
                                         forward pass leaves them zero, so the FIRST backward after a forward
                                         needs no clearing; set this for every further backward over the same
                                         forward state (retain_graph) */

#define TSPLAT_FLAG_SORT_ONESWEEP 2   /* tsplat_forward: bin with the global onesweep LSD radix sort over the full
                                         (view*T + tile | depth) key instead of the default bucket path (count per
                                         tile -> scan -> emit -> per-tile sort in shared memory).  Same results. */

#define TSPLAT_FLAG_DEBUG 4           /* the reference's debug=True (GaussianRasterizationSettings.debug, cuda_splatting.py:120):
                                         synchronise the stream after every stage and report the first failing
                                         one (TSPLAT_ECUDA + tsplat_last_error names it).  Blocks the host. */
#define TSPLAT_FLAG_BWD_RENDER_ONLY 8      /* tsplat_backward[_mse]: stop after the render-backward stage (the screen-space
                                              accumulators are complete; no per-Gaussian gradient is written yet) */
#define TSPLAT_FLAG_BWD_GAUSSIANS_ONLY 16  /* tsplat_backward[_mse]: only the per-Gaussian stage, for the Gaussians
                                              [g_begin, g_begin + g_count) - a view-sharded caller reduces the
                                              gradients of one range over the ranks while the next is computed
                                              (transplat_b200/parallel.py) */

#define TSPLAT_FLAG_BWD_MULTIMEM 32         /* tsplat_backward[_mse]: the four per-Gaussian gradient pointers are MULTICAST
                                              addresses of symmetric slabs replicated on every GPU of an NVSwitch domain
                                              (cuMulticast* / torch symmetric memory).  The per-Gaussian kernel then ADDS
                                              its values with multimem.red.add.f32: the switch reduces them into every
                                              replica, so after a barrier every rank holds the gradients summed over all
                                              ranks' views - the all-reduce of view sharding happens inside the
                                              kernel that produces the gradients, overlapped with it.  The caller zeroes
                                              its replica and synchronises the ranks before and after the call. */

/* fused depth channel (render_depth_cuda, cuda_splatting.py:375-417) */
#define TSPLAT_DEPTH_NONE 0
#define TSPLAT_DEPTH_DEPTH 1
#define TSPLAT_DEPTH_DISPARITY 2
#define TSPLAT_DEPTH_RELATIVE_DISPARITY 3
#define TSPLAT_DEPTH_LOG 4

/*
 * One camera = 64 floats (256 B).  Matrices use the reference's flat convention: element
 * (row r, col c) of the conventional column-vector matrix is m[4*c + r]
 * (cuda_splatting.py:94-96 passes transposes).
 */
typedef struct TsplatCamera {
  float view[16];     /* GaussianRasterizationSettings.viewmatrix   (cuda_splatting.py:95,115)  */
  float proj[16];     /* GaussianRasterizationSettings.projmatrix   (cuda_splatting.py:96,116)  */
  float campos[3];    /* GaussianRasterizationSettings.campos       (cuda_splatting.py:118)     */
  float tanfovx;      /* (cuda_splatting.py:111) */
  float tanfovy;      /* (cuda_splatting.py:112) */
  float bg[3];        /* GaussianRasterizationSettings.bg           (cuda_splatting.py:113)     */
  float mean_scale;   /* s = 1/near applied to means (cuda_splatting.py:74-77); 1 at level 1   */
  float cov_scale;    /* s*s applied to covariances (cuda_splatting.py:76); 1 at level 1       */
  float depth_row[4]; /* row 2 of inverse(extrinsics), unscaled (cuda_splatting.py:388-391)     */
  float near_plane;   /* unscaled near (relative_disparity / log, cuda_splatting.py:395-400)    */
  float far_plane;    /* unscaled far                                                           */
  float reserved[16];
} TsplatCamera;

/*
 * One call = B scenes x V target views per scene, all scenes with G Gaussians and all views HxW.
 * Level 1 (GaussianRasterizer.forward) uses B = V = 1, SH_GM3, COV_TRIU6, DEPTH_NONE.
 * Level 2 (DecoderSplattingCUDA.forward) passes the encoder's tensors untouched.
 */
typedef struct TsplatProblem {
  int32_t B;          /* scenes                                                                  */
  int32_t G;          /* Gaussians per scene                                                     */
  int32_t V;          /* views per scene                                                         */
  int32_t H, W;       /* image size (settings.image_height/width, cuda_splatting.py:109-110)     */
  int32_t sh_degree;  /* active degree 0..4 (settings.sh_degree, cuda_splatting.py:117)          */
  int32_t sh_coeffs;  /* M coefficients stored per channel; 0 => `sh` holds colors_precomp [G,3] */
  int32_t sh_layout;  /* TSPLAT_SH_*                                                             */
  int32_t cov_layout; /* TSPLAT_COV_*                                                            */
  int32_t depth_mode; /* TSPLAT_DEPTH_*; != NONE adds the fused 4th (depth) channel              */
  int32_t sh_l4;      /* 1: evaluate the l=4 band when sh_degree>=4, 0: ignore it   */
  int32_t flags;      /* TSPLAT_FLAG_* bits                                                      */
  const float* means;   /* [B, G, 3]                                                             */
  const float* cov;     /* [B, G, 6] or [B, G, 3, 3]                                             */
  const float* sh;      /* [B, G, M, 3] / [B, G, 3, M] / colors_precomp [B, G, 3]                */
  const float* opacity; /* [B, G]                                                                */
  const TsplatCamera* cams; /* [B*V]                                                             */
  int32_t g_begin;    /* tsplat_backward with TSPLAT_FLAG_BWD_GAUSSIANS_ONLY: the Gaussian range      */
  int32_t g_count;    /* [g_begin, g_begin + g_count) of every scene (g_begin a multiple of 64; 0 = all) */
} TsplatProblem;

/* Byte offsets of the inspectable workspace sections (all 256-B aligned).  N = B*V views. */
typedef struct TsplatLayout {
  uint64_t total_bytes;
  uint64_t status;        /* TsplatStatus                                                         */
  uint64_t depths;        /* float  [N, G]   camera-space z (the low key word)                    */
  uint64_t radii;         /* int32  [N, G]                                                        */
  /* the next three are INTERLEAVED in one record of rec_stride (48) bytes per (view, Gaussian):
   * x, y | conic.x, conic.y, conic.z, opacity | r, g, b, depth feature | id bits, cull threshold;
   * element i of each section lives at section offset + i * rec_stride                            */
  uint64_t xy;            /* float2 [N, G]   pixel-space centre                                   */
  uint64_t conic_opacity; /* float4 [N, G]                                                        */
  uint64_t rgbd;          /* float4 [N, G]   r, g, b, depth feature                               */
  uint64_t tiles_touched; /* uint32 [N, G]                                                        */
  uint64_t clamped;       /* uint8  [N, G]   bit c set => channel c was clamped at 0              */
  uint64_t keys;          /* uint64 [d_cap]  sorted: ((view*T + tile) << 32) | depth bits         */
  uint64_t vals;          /* uint32 [d_cap]  sorted Gaussian index within its scene               */
  uint64_t ranges;        /* uint2  [N*T]    [first, last) per (view, tile)                       */
  uint64_t final_T;       /* float  [N, H, W]                                                     */
  uint64_t n_contrib;     /* uint32 [N, H, W]                                                     */
  uint64_t grad_acc;      /* float  [N, G, 12] screen-space gradient accumulators (zeroed by the forward,
                             filled by the backward): moments of u = dL/dG * G over the pixel offsets
                             (dx, dy, dx^2, dx dy, dy^2) [conic, -1/2 and W/2, H/2 applied by the preprocess
                             backward], then dL/d opacity, r g b, depth feature, 2 pad               */
  int32_t tiles_x, tiles_y; /* grid                                                               */
  int32_t sort_bits;      /* number of key bits the sort covers                                   */
  int32_t sort_passes;
  int32_t rec_stride;     /* byte stride of the interleaved xy / conic_opacity / rgbd sections    */
  int32_t reserved0;
  /* per-8x4-block lists (what the render kernels stream; see TsplatStatus.num_block_entries) */
  uint64_t block_jobs;    /* uint4 [N*T*8]  one render job per (view, tile, block), tiles in heaviest-first order:
                             first reference, reference count, (view*T + tile)*8 + block, 0            */
  uint64_t block_refs;    /* uint2 [e_cap]: Gaussian id, 1-based position in the TILE list; a block's references
                             are contiguous and in list order                                          */
  uint64_t e_cap;         /* capacity of block_refs in entries (TSPLAT_BLOCK_ENTRY_RATIO * d_cap)     */
} TsplatLayout;

/* Device-side status block (first bytes of the workspace). */
typedef struct TsplatStatus {
  uint64_t num_rendered; /* D over all views of the call (may exceed d_cap when overflow != 0)    */
  uint32_t overflow;     /* 1 if D > d_cap or E > TSPLAT_BLOCK_ENTRY_RATIO * d_cap                 */
  uint32_t reserved;
  uint64_t num_block_entries; /* E: entries of the per-8x4-block lists the render kernels stream (each pair is
                                 listed once per pixel block of its tile it can touch: 1.1-1.3 D for encoder-sized
                                 Gaussians, up to 8 D for tile-covering ones); capacity = TSPLAT_BLOCK_ENTRY_RATIO * d_cap */
  uint64_t reserved1;
} TsplatStatus;
#define TSPLAT_BLOCK_ENTRY_RATIO 2

int tsplat_version(void);
const char* tsplat_last_error(void);

/* Replaces the resize lambdas + GeometryState/BinningState/ImageState::fromChunk carving of the
 * extension: one caller-owned workspace, size known up front. */
int tsplat_workspace_layout(const TsplatProblem* p, uint64_t d_cap, TsplatLayout* out);

/* Replaces _C.rasterize_gaussians.
 *   out_color [B*V, 3, H, W]; out_depth [B*V, H, W] or NULL (required iff depth_mode != NONE);
 *   out_radii [B*V, G] int32 or NULL. */
int tsplat_forward(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes,
                   uint64_t d_cap, float* out_color, float* out_depth, int32_t* out_radii,
                   void* cuda_stream);

/* tsplat_forward without a caller-owned workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments and results as tsplat_forward; the workspace is a static
 * device array of TSPLAT_NOALLOC_WORKSPACE_BYTES that every GPU loading the library holds until the process exits.
 *   - a problem whose tsplat_workspace_layout(p, d_cap).total_bytes exceeds that size is rejected with TSPLAT_ENOSPC;
 *   - the call makes no cudaMalloc / cudaFree / allocating library call; the caller's current device must be the one
 *     that owns the pointers (the array's device address is resolved per device);
 *   - all calls on a device share the array: calls that overlap on the GPU (other streams, concurrent graph replays)
 *     race and must be serialised by the caller; the array keeps the previous call's contents and the forward clears
 *     what must start at zero itself;
 *   - an overflow of d_cap poisons the outputs (NaN in the first colour value of every view), as in tsplat_forward;
 *     there is no backward over the static workspace. */
#define TSPLAT_NOALLOC_WORKSPACE_BYTES (32u << 20)
int tsplat_forward_noalloc(const TsplatProblem* p, uint64_t d_cap, float* out_color, float* out_depth,
                           int32_t* out_radii, void* cuda_stream);

/* Replaces _C.rasterize_gaussians_backward.  Must follow tsplat_forward on
 * the same (problem, workspace).  Gradients are summed over the V views of each scene and
 * WRITTEN (not accumulated):
 *   dL_dmeans [B,G,3]; dL_dcov [B,G,6] or [B,G,3,3] (upper triangle filled, lower zero);
 *   dL_dsh same layout as p->sh; dL_dopacity [B,G];
 *   dL_dmeans2D [B*V, G, 3] or NULL (gradient of the dummy screen-space tensor, level 1). */
int tsplat_backward(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes,
                    uint64_t d_cap, const float* dL_dcolor, const float* dL_ddepth,
                    float* dL_dmeans, float* dL_dcov, float* dL_dsh, float* dL_dopacity,
                    float* dL_dmeans2D, void* cuda_stream);

/* The step AFTER the render path fused into it: the MSE colour loss of the reference,
 *   delta = prediction.color - target; loss = weight * (delta ** 2).mean()        src/loss/loss_mse.py:23-31
 * tsplat_forward_mse = tsplat_forward + the sum of squared differences, accumulated in the render kernel's epilogue
 * (one partial per (view, tile) CTA, summed in a fixed order by the last CTA: deterministic).  tsplat_backward_mse =
 * tsplat_backward whose upstream colour gradient is formed in the render-backward prologue:
 *   dL/dcolor = dL_dcolor (may be NULL) + (*dL_dloss) * scale * (color - target),    scale = 2 * weight / numel,
 * so neither delta nor dL/dcolor is ever materialised and no elementwise kernel runs between forward and backward. */
typedef struct TsplatMse {
  const float* target;   /* [B*V, 3, H, W] ground-truth images (device)                                     */
  float* loss_sum;       /* forward: 1 float (device) <- loss_scale * sum over views, channels, pixels of delta^2 */
  const float* color;    /* backward: the out_color written by the forward                                    */
  const float* dL_dloss; /* backward: 1 float (device), upstream gradient of the loss VALUE                   */
  float scale;           /* backward: 2 * weight / numel (host)                                               */
  float loss_scale;      /* forward: weight / numel (host), so *loss_sum is the reference's loss value        */
  float* psnr;           /* forward, optional (NULL = skip): [B*V] per-view PSNR of the clipped images,
                            compute_psnr of src/evaluation/metrics.py:12-19 as logged by model_wrapper.py:149-153 */
} TsplatMse;
int tsplat_forward_mse(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                       float* out_color, float* out_depth, int32_t* out_radii, const TsplatMse* mse,
                       void* cuda_stream);
int tsplat_backward_mse(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                        const float* dL_dcolor, const float* dL_ddepth, float* dL_dmeans, float* dL_dcov,
                        float* dL_dsh, float* dL_dopacity, float* dL_dmeans2D, const TsplatMse* mse,
                        void* cuda_stream);

/* Replaces _C.mark_visible (GaussianRasterizer.markVisible).  present [G] uint8. */
int tsplat_mark_visible(int32_t G, const float* means, const float* viewmatrix,
                        const float* projmatrix, uint8_t* present, void* cuda_stream);

/* The scales / rotations input mode of GaussianRasterizer.forward ("exactly one of (scales, rotations) /
 * cov3D_precomp"): computeCov3D of the extension and its backward.  scales [G,3], rotations [G,4] as (r, x, y, z), used
 * as given (the extension does not normalise the quaternion); cov6 [G,6] = upper triangle of R diag(s^2) R^T with
 * s = scale_modifier * scale.  TranSplat always passes cov3D_precomp (cuda_splatting.py:124,132); at the reference's
 * call sites scale_modifier is 1.0 (cuda_splatting.py:114). */
int tsplat_cov3d_forward(int32_t G, const float* scales, const float* rotations, float scale_modifier, float* cov6,
                         void* cuda_stream);
int tsplat_cov3d_backward(int32_t G, const float* scales, const float* rotations, float scale_modifier,
                          const float* dL_dcov6, float* dL_dscales, float* dL_drotations, void* cuda_stream);

/* The one exchange on the path (view sharding): sums a gradient slab over the GPUs of one NVSwitch domain,
 * in place, through the switch's reduction unit (multimem.ld_reduce + multimem.st; NVLS).  Replaces the autograd sum
 * over the `repeat`-ed Gaussian copies (decoder_splatting_cuda.py:59-62) when those copies' views live on different
 * GPUs - with DDP the reference would all-reduce the encoder's parameter gradients instead.
 *   multicast_slab: the MULTICAST alias of a buffer replicated on every rank (symmetric memory; cuMulticast*), 16-byte
 *   aligned.  Pieces (at most 4, offsets and counts in floats, multiples of 4) are reduced; rank r of `world` handles the
 *   r-th part of every piece.  num_ctas <= 0 picks the default.  The caller orders the launch between two barriers
 *   over the ranks: every replica completely written before, every rank's launch finished before the sums are read. */
int tsplat_nvls_all_reduce(void* multicast_slab, const uint64_t* piece_offset_floats, const uint64_t* piece_count_floats,
                           int32_t num_pieces, int32_t rank, int32_t world, int32_t num_ctas, void* cuda_stream);

/* Device-side camera preparation for the batched decoder path: fills cams[N]
 * from what DecoderSplattingCUDA.forward receives, in one launch and with no host synchronisation.
 * Replaces the host arithmetic of render_cuda / render_depth_cuda:
 *   cuda_splatting.py:73-80 (scale invariance), :89-91 (get_fov, projection.py:233-247, tan),
 *   :93-96 (get_projection_matrix :26-53, extrinsics.inverse(), view @ projection),
 *   :111-112 (the two .item() host reads per view), :388-391 (unscaled inverse row for depth).
 *   extrinsics [N,4,4] camera-to-world (row-major), intrinsics [N,3,3] normalised, near/far [N],
 *   bg [N,3] with bg_stride = 3, or one [3] colour for every camera with bg_stride = 0. */
int tsplat_pack_cameras(int32_t N, const float* extrinsics, const float* intrinsics,
                        const float* near_plane, const float* far_plane, const float* bg,
                        int32_t bg_stride, int32_t scale_invariant, TsplatCamera* cams,
                        void* cuda_stream);

/* Copies the status block to host memory asynchronously on the stream.  status_host should be pinned
 * (cudaHostAlloc / cudaHostRegister, 16-byte aligned, 32 bytes): it is then written by a one-warp kernel through its
 * device alias, so the read never queues on a copy engine behind another stream's bulk transfer; pageable
 * memory falls back to cudaMemcpyAsync. */
int tsplat_status_read(const void* workspace, TsplatStatus* status_host, void* cuda_stream);

/* Optional per-stage device timing (cudaEvents recorded on the caller's stream around each stage;
 * process-global, off by default, a ring of the last 256 forward(+backward) calls).  Replaces
 * nothing in the reference, which only has wall-clock NVTX ranges around the whole decoder
 * (src/misc/benchmarker.py:44-145, model_wrapper.py:209).
 * tsplat_profile_read synchronises on the recorded events and fills ms_host[0..8] with the SUM of
 * each stage's durations over the recorded calls and ms_host[9] with the number of calls summed:
 * 0 zero-fill, 1 preprocess, 2 bin_emit, 3 sort, 4 ranges_gather (tile ranges on the onesweep path + the per-block
 * lists of both paths: block_lists_kernel), 5 render_forward,
 * 6 zero-fill (bwd), 7 render_backward, 8 preprocess_backward.  n must be >= 10. */
#define TSPLAT_NUM_STAGES 9
int tsplat_profile_enable(int on);
int tsplat_profile_read(float* ms_host, int n);

/* ------------------------------------------------------------------------------------------------------------
 * The step BEFORE the render path: the per-pixel Gaussian producer.  Replaces the torch op chain of
 *   GaussianAdapter.forward           src/model/encoder/common/gaussian_adapter.py:50-98
 *   build_covariance / quaternion_to_matrix   src/model/encoder/common/gaussians.py:7-43
 *   get_world_rays / unproject        src/geometry/projection.py:74-113
 *   rotate_sh_with_precomputed        src/misc/sh_rotation_optimized.py:56-101
 * with one kernel forward and one backward, as called from its only call site (encoder_trans.py:307-318): N = b*v
 * context views, n = rays*surfaces*gaussians_per_pixel Gaussians per view, every per-view quantity broadcast over n.
 * Outputs are written in the layouts the rasterizer reads (means [N,n,3], covariances [N,n,3,3], harmonics
 * [N,n,3,d_sh]), i.e. Gaussians{means, covariances, harmonics} after the encoder's rearrange "b v r srf spp -> b (v r
 * srf spp)" is a pure view.  Opacities pass through unchanged (not an argument).
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct TsplatAdapterProblem {
  int32_t N;          /* context views of the call (batch * views)                                        */
  int32_t n;          /* Gaussians per view                                                               */
  int32_t sh_degree;  /* cfg.sh_degree, 0..4; d_sh = (sh_degree + 1)^2                                    */
  int32_t H, W;       /* image_shape of the context views (pixel size of the scale multiplier, :66-68)    */
  float scale_min;    /* cfg.gaussian_scale_min                                                           */
  float scale_max;    /* cfg.gaussian_scale_max                                                           */
  float eps;          /* forward(..., eps=1e-8): quaternion normalisation (:72)                           */
  const float* extrinsics;  /* [N,4,4] camera-to-world, row-major                                         */
  const float* intrinsics;  /* [N,3,3] normalised                                                         */
  const float* coordinates; /* [N,n,2] ray coordinates in [0,1]^2                                         */
  const float* depths;      /* [N,n]                                                                      */
  const float* raw;         /* [N,n,7+3*d_sh] raw_gaussians: scales 3, rotation xyzw 4, sh [3][d_sh]      */
  const float* wigner;      /* [N, sum_l (2l+1)^2] per-degree Wigner-D blocks, row-major, concatenated
                               (precompute_wigner_d_matrices, sh_rotation_optimized.py:27-53)             */
  const float* sh_mask;     /* [d_sh] GaussianAdapter.sh_mask (:36-46)                                    */
} TsplatAdapterProblem;

/* means [N,n,3], covariances [N,n,3,3], harmonics [N,n,3,d_sh], scales [N,n,3], rotations [N,n,4] (xyzw, normalised) */
int tsplat_adapter_forward(const TsplatAdapterProblem* p, float* means, float* covariances, float* harmonics,
                           float* scales, float* rotations, void* cuda_stream);
/* gradients w.r.t. coordinates [N,n,2], depths [N,n], raw [N,n,7+3*d_sh] from those of means, covariances and
 * harmonics (scales / rotations are export-only outputs, gaussian_adapter.py:91-94: no gradient path). */
int tsplat_adapter_backward(const TsplatAdapterProblem* p, const float* dL_dmeans, const float* dL_dcovariances,
                            const float* dL_dharmonics, float* dL_dcoordinates, float* dL_ddepths, float* dL_draw,
                            void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* TSPLAT_H_ */
