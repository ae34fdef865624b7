/* pixelnerf_b200 -- C ABI of the B200 (sm_100a) render hot path.
 *
 * The reference (etiiiR/pixel-nerf) is pure Python/PyTorch and has no FFI of its
 * own; its boundary for this path is the Python API (make_model /
 * PixelNeRFNet.encode / forward, NeRFRenderer.from_conf / bind_parallel).  The
 * Python mirror of that API in pixelnerf_b200/{model,render} binds the entry
 * points below with ctypes (pixelnerf_b200/_lib.py); INTEGRATION.md shows the
 * stub.  Each entry point cites the reference code whose torch op sequence it
 * replaces (paths are relative to the reference root).
 *
 * Conventions: every pointer is a DEVICE pointer unless named host_*; tensors are
 * dense row-major fp32 unless stated; the library never allocates or frees device
 * memory (outputs and workspaces are caller-owned); kernels are enqueued on
 * `stream` and the call returns without synchronising; the return value is 0 on
 * success, otherwise a PN_ERR_* code and pn_last_error() (thread-local) holds a
 * message.  All functions are re-entrant and keep no per-process device state.
 */
#ifndef PIXELNERF_B200_H
#define PIXELNERF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PN_OK 0
#define PN_ERR_INVALID 1     /* bad argument / unsupported configuration */
#define PN_ERR_CUDA 2        /* CUDA runtime error */
#define PN_ERR_UNSUPPORTED 3 /* valid in the reference, deliberately not in this path */

#define PN_DTYPE_F32 0
#define PN_DTYPE_F16 1
#define PN_DTYPE_BF16 2

#define PN_D_HIDDEN 512  /* ResnetFC d_hidden of every shipped conf (conf/default_mv.conf:10,17) */
#define PN_D_LATENT 512  /* ResNet34 4-level latent (src/model/encoder.py:70) */
#define PN_D_CODE 42     /* 3 + 2*6*3 positional encoding + 3 view-dir (src/model/models.py:49-60) */
#define PN_N_BLOCKS 5
#define PN_COMBINE_LAYER 3

typedef void* pn_stream_t; /* cudaStream_t */

#if defined(__GNUC__)
#define PN_API __attribute__((visibility("default")))
#else
#define PN_API
#endif

PN_API int pn_version(void);
PN_API const char* pn_last_error(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
PN_API uint64_t pn_launch_count(void);

/* ---- (1) ray sampling ---------------------------------------------------------------- */

/* Device-resident seed word for the in-kernel draws (no counterpart in the reference, whose draws come from torch's
 * generator on the host: src/render/nerf.py:112,139,158).  The sampling launches made by the CALLING HOST THREAD after
 * this call (pn_sample_coarse, pn_sample_fine, pn_render_rays, pn_render_frames) use seed + *seed_dev; NULL (the
 * default) restores the host seed alone.  For CUDA graphs: a captured launch has its arguments baked in, so a replayed
 * training step would repeat one jitter pattern; with the word in device memory the owner advances it by a captured
 * device-side add and every replay draws anew.  The pointer must stay valid for as long as launches that saw it
 * may run. */
PN_API void pn_set_device_seed(const uint64_t* seed_dev);

/* NeRFRenderer.sample_coarse, src/render/nerf.py:98-118.  rays (R,8) =
 * [origin, dir, near, far]; host-made table lin = torch.linspace(0, 1-1/Kc, Kc);
 * u (R,Kc) uniform draws or NULL to draw in-kernel from (seed).  Every fp32
 * operation is rounded where torch rounds it (bit-exact with torch-CPU). */
PN_API int pn_sample_coarse(const float* rays, const float* u, const float* lin, float* z,
                     int64_t n_rays, int n_coarse, int lindisp, uint64_t seed, pn_stream_t stream);

/* NeRFRenderer.sample_fine + sample_fine_depth + cat + sort,
 * src/render/nerf.py:120-161, 285-295.  weights (R,Kc) and depth (R) are the coarse
 * composite outputs; u_bin/u_in (R,Kf) and noise (R,Kd) are injected draws or NULL.
 * z_sorted (R,Kc+Kf+Kd) is the merged ascending sample set; z_imp (R,Kf) and
 * z_depth (R,Kd) are optional outputs (NULL to skip).  Kf or Kd may be 0. */
PN_API int pn_sample_fine(const float* rays, const float* z_coarse, const float* weights, const float* depth,
                   const float* u_bin, const float* u_in, const float* noise,
                   float* z_sorted, float* z_imp, float* z_depth,
                   int64_t n_rays, int n_coarse, int n_imp, int n_depth, float depth_std,
                   int lindisp, uint64_t seed, pn_stream_t stream);

/* ---- (2) point features ---------------------------------------------------------------- */

/* NCHW fp32 encoder output (src/model/encoder.py:150-164) -> channels-last map
 * (V,Hl,Wl,C) in fp32 or fp16, so one bilinear tap is one contiguous C-vector. */
PN_API int pn_prepare_latent(const float* latent_nchw, void* latent_nhwc, int n_maps, int channels,
                      int hl, int wl, int dtype, pn_stream_t stream);

typedef struct pn_camera {
  const float* poses;   /* (SB*NS,3,4) world->camera, as cached by encode(), models.py:112-114 */
  const float* focal;   /* (SB*NS,2) [fx,-fy] already expanded per view, models.py:120-130 */
  const float* c;       /* (SB*NS,2) principal points, models.py:132-141 */
  float scale_x;        /* latent_scaling[0] / image_width,  encoder.py:98-99,161-163 */
  float scale_y;        /* latent_scaling[1] / image_height */
  int hl, wl;           /* latent map size */
  int n_objects;        /* SB */
  int n_views;          /* NS */
} pn_camera_t;

/* PixelNeRFNet.forward up to the MLP input, src/model/models.py:157-227 +
 * SpatialEncoder.index src/model/encoder.py:80-109 + PositionalEncoding
 * src/model/code.py:30-42.  Points come either from rays+z (renderer path:
 * x = o + z d, viewdir = d, src/render/nerf.py:185,203-208; n_samples = K, P = rays
 * per object * K) or from xyz/viewdirs (SB,P,3) when rays == NULL.
 * Rows are ordered (object, view, point).  feat_z (N,512) and feat_x (N,ldx) are
 * written in out_dtype; columns [42,ldx) of feat_x are zero. */
PN_API int pn_point_features(const pn_camera_t* cam, const void* latent_nhwc, int latent_dtype,
                      const float* rays, const float* z, int n_samples,
                      const float* xyz, const float* viewdirs,
                      int64_t points_per_object, void* feat_z, void* feat_x, int ldx, int out_dtype,
                      pn_stream_t stream);

/* Tensor-core path: the same projection / positional code as pn_point_features, but instead of
 * blending the latent it records WHERE the blend reads.  feat_x (N,64) fp16, 128 bytes per row:
 * columns [0,42) the positional code, [42,48) zero, and in bytes [96,128) the taps record = 8 32-bit
 * words [texel index of the 4 bilinear taps (view*Hl*Wl + y*Wl + x), their 4 fp32 weights]
 * (grid_sample(align_corners, border), encoder.py:102-108).
 * The chain kernel applies lin_z through per-texel tables (pn_prepare_latent_tables). */
PN_API int pn_point_taps(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                  const float* xyz, const float* viewdirs, int64_t points_per_object,
                  void* feat_x, pn_stream_t stream);

/* Backward of the latent gather (grid_sample backward w.r.t. its input, encoder.py:102-108):
 * g_latent_nhwc (V,Hl,Wl,512) fp32 += scatter of g_feat_z (N,512) with the bilinear weights.
 * Same point addressing as pn_point_features.  The caller zeroes g_latent_nhwc first and hands
 * its NCHW view to torch autograd for the ResNet encoder. */
PN_API int pn_point_features_bwd(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                          const float* xyz, int64_t points_per_object, const float* g_feat_z,
                          float* g_latent_nhwc, pn_stream_t stream);

/* Backward of the depth-guided sample positions (autograd of nerf.py:150-161 + :290-295; the only sample
 * positions the reference does not detach): g_depth (R) += sum over each ray's depth samples z_depth (R,Kd) that
 * lie strictly inside (near, far) of dL/dz at the sample's place in the sorted row z_sorted (R,K).  dL/dz =
 * g_z (R,K) from pn_composite_bwd (or NULL) + the path through the MLP inputs: code rows (g_h0 = dL/d lin_in
 * output, rows x 512, PN_DTYPE_F32 or PN_DTYPE_BF16, times lin_in_w (512,42), code.py:30-42) and latent rows
 * (g_feat_z rows x 512 fp32 times grid_sample's grid gradient, encoder.py:102-108, through models.py:162-212).
 * latent_nhwc is the fp32 channels-last latent; rows are ordered (object, view, ray, sample). */
PN_API int pn_depth_sample_bwd(const pn_camera_t* cam, const float* latent_nhwc, const float* rays,
                        const float* z_sorted, int n_samples, const float* z_depth, int n_depth,
                        const float* g_z, const void* g_h0, int g_h0_dtype, const float* g_feat_z,
                        const float* lin_in_w, int64_t rays_per_object, float* g_depth, pn_stream_t stream);

/* ---- (3) ResnetFC ---------------------------------------------------------------------- */

typedef struct pn_mlp_weights {
  const float* lin_in_w;  const float* lin_in_b;                 /* (512,42), (512)  resnetfc.py:88 */
  const float* fc0_w[PN_N_BLOCKS]; const float* fc0_b[PN_N_BLOCKS]; /* (512,512)     resnetfc.py:32 */
  const float* fc1_w[PN_N_BLOCKS]; const float* fc1_b[PN_N_BLOCKS]; /* (512,512)     resnetfc.py:33 */
  const float* lin_z_w[PN_COMBINE_LAYER]; const float* lin_z_b[PN_COMBINE_LAYER]; /* resnetfc.py:112-114 */
  const float* lin_out_w; const float* lin_out_b;                /* (4,512), (4)     resnetfc.py:92 */
} pn_mlp_weights_t;

/* Workspace for pn_mlp_forward_fp32 over n_rows = SB*NS*P feature rows. */
PN_API size_t pn_mlp_fp32_workspace_bytes(int64_t n_rows);

/* ResnetFC.forward (src/model/resnetfc.py:132-184) + output head
 * (src/model/models.py:257-266) in fp32 on CUDA cores: the 1e-4 parity path.
 * feat_z (N,512) fp32, feat_x (N,ldx>=42) fp32, rows ordered (object, view, point);
 * out (SB*P,4) = [sigmoid(rgb), relu(sigma)].  If pre_act is not NULL the raw
 * lin_out outputs (SB*P,4) are stored there too (needed by the backward). */
PN_API int pn_mlp_forward_fp32(const pn_mlp_weights_t* w, const float* feat_z, const float* feat_x, int ldx,
                        int n_objects, int n_views, int64_t points_per_object,
                        float* out, float* pre_act, void* workspace, size_t workspace_bytes,
                        pn_stream_t stream);

/* ---- (3b) ResnetFC training: forward keeping activations + hand-written backward ----------
 * (autograd of src/model/resnetfc.py:132-184 and the head src/model/models.py:257-266; driven by
 * loss.backward() in train/train.py:271-281).  fp32 on CUDA cores. */
typedef struct pn_mlp_grads {
  float* lin_in_w;  float* lin_in_b;
  float* fc0_w[PN_N_BLOCKS]; float* fc0_b[PN_N_BLOCKS];
  float* fc1_w[PN_N_BLOCKS]; float* fc1_b[PN_N_BLOCKS];
  float* lin_z_w[PN_COMBINE_LAYER]; float* lin_z_b[PN_COMBINE_LAYER];
  float* lin_out_w; float* lin_out_b;
} pn_mlp_grads_t;

PN_API size_t pn_mlp_train_acts_bytes(int64_t n_rows);
/* Same arithmetic as pn_mlp_forward_fp32; `acts` (pn_mlp_train_acts_bytes) receives the 11
 * activation tensors the backward needs. */
PN_API int pn_mlp_forward_fp32_train(const pn_mlp_weights_t* w, const float* feat_z, const float* feat_x, int ldx,
                              int n_objects, int n_views, int64_t points_per_object,
                              float* out, float* pre_act, void* acts, size_t acts_bytes, pn_stream_t stream);
PN_API size_t pn_mlp_backward_workspace_bytes(int64_t n_rows);
/* g_out (P,4) = dL/d[rgb, sigma] (the model outputs `out`).  Weight and bias gradients are
 * ACCUMULATED (+=) into `grads` (same shapes as the weights); wt holds the transposed fc_0 / fc_1 /
 * lin_z matrices ([in][out] row-major) for the data-gradient GEMMs.  If g_feat_z (N,512) is not NULL
 * it receives dL/d feat_z (the latent rows), for pn_point_features_bwd. */
/* byte offset in `workspace` where the call leaves dL/dh_0 (rows x 512 fp32), for pn_depth_sample_bwd */
PN_API size_t pn_mlp_backward_gh0_offset(int64_t n_rows, int n_views);
PN_API int pn_mlp_backward_fp32(const pn_mlp_weights_t* w, const pn_mlp_weights_t* wt, const float* feat_z,
                         const float* feat_x, int ldx, int n_objects, int n_views, int64_t points_per_object,
                         const void* acts, const float* out, const float* g_out, const pn_mlp_grads_t* grads,
                         float* g_feat_z, void* workspace, size_t workspace_bytes, pn_stream_t stream);

/* Tensor-core ResnetFC (tcgen05 / TMEM / TMA chain kernel, csrc/pn_mlp_tc.cu): fp16 GEMM
 * operands, fp32 accumulation and an fp32 residual stream held in tensor memory.
 * pn_prepare_weights converts one MLP's fp32 state_dict tensors once into the packed form
 * (15 zero-padded [512][512] fp16 layers + fp32 bias tables); `packed` is a caller-owned
 * 256-byte-aligned device buffer of pn_mlp_tc_packed_bytes(). */
PN_API size_t pn_mlp_tc_packed_bytes(void);

/* lin_z is linear and the latent row of a point is a convex blend of 4 texels, so
 * lin_z[b](blend) = blend(lin_z[b](texel)) (resnetfc.py:174-180 after encoder.py:102-108):
 * tables[b][texel][512] fp16 = lin_z[b](latent[texel]) + lin_z[b].bias + c_b, computed in fp32 once per
 * (encode, parameter version), with c_0 = lin_in.bias, c_b = blocks[b-1].fc_1.bias folded in (bilinear
 * weights sum to 1).  n_texels = SB*NS*Hl*Wl; latent_nhwc fp32; workspace n_texels*512 floats. */
PN_API size_t pn_latent_tables_bytes(int64_t n_texels);
PN_API int pn_prepare_latent_tables(const pn_mlp_weights_t* w, const float* latent_nhwc, int64_t n_texels,
                             void* tables, void* workspace, size_t workspace_bytes, pn_stream_t stream);
/* The same tables computed on the tensor cores, for training mode where they are rebuilt every step: latent_nhwc_f16
 * is the fp16 channels-last latent, lin_z weights enter as fp16 hi + lo pairs (two GEMMs per layer, fp32 sums).
 * workspace: pn_latent_tables_tc_workspace_bytes(n_texels), 256-byte aligned. */
PN_API size_t pn_latent_tables_tc_workspace_bytes(int64_t n_texels);
PN_API int pn_prepare_latent_tables_tc(const pn_mlp_weights_t* w, const void* latent_nhwc_f16, int64_t n_texels,
                                void* tables, void* workspace, size_t workspace_bytes, pn_stream_t stream);
PN_API int pn_prepare_weights(const pn_mlp_weights_t* w, void* packed, size_t packed_bytes, pn_stream_t stream);

typedef struct pn_mlp_tc_debug {
  float* h;   /* (rows, 512) fp32: residual stream h as read at `stage`, biases included */
  int stage;  /* 0..5: h before block `stage` (5 = input of lin_out) */
} pn_mlp_tc_debug_t;

PN_API size_t pn_mlp_tc_workspace_bytes(int64_t n_rows, int n_views);

/* ResnetFC.forward + head like pn_mlp_forward_fp32, on the tensor cores.  The latent rows enter as
 * `tables` (pn_prepare_latent_tables) + the taps record inside feat_x (N,64) fp16 (pn_point_taps), rows
 * ordered (object, view, point).  With n_views > 1 the per-view blocks 0-2 and the merged
 * blocks 3-4 run as two launches with the fp32 view mean in between (workspace).  dbg may
 * be NULL.  The first 4 bytes of workspace receive a nonzero code if a device-side wait
 * timed out (the launch then fails with a CUDA error instead of hanging). */
PN_API int pn_mlp_forward_tc(const void* packed, const void* tables, int64_t n_texels,
                      const void* feat_x, int n_objects, int n_views, int64_t points_per_object,
                      float* out, float* pre_act, void* workspace, size_t workspace_bytes,
                      const pn_mlp_tc_debug_t* dbg, pn_stream_t stream);

/* n_views > 1 in ONE launch, with the view mean (combine_interleaved, src/util/util.py:461-466, at block 3,
 * src/model/resnetfc.py:152-172) taken on chip.  pn_point_taps_grouped writes the rows of pn_point_taps in GROUPED
 * order: points are numbered object-major (SB * points_per_object of them); super-tile S holds points [128 S, 128 S + 128)
 * as one 128-row block per view -- row (S * n_views + v) * 128 + m = view v of point 128 S + m; rows of points past the
 * end are zero.  feat_x_grouped holds pn_point_taps_grouped_rows(n_points, n_views) rows of 128 bytes.
 * pn_mlp_forward_tc_grouped runs blocks 0-2 on the n_views blocks of a super-tile one after the other; row m of every
 * block is point m, so each epilogue thread keeps a running fp32 sum of its row's h over the views (in view order, like
 * torch.mean's sum; parked in the workspace between blocks, 128 KB per CTA, L2 resident), divides by n_views, rounds
 * the mean to fp16 ONCE (the two-launch form rounds every view's h to fp16 before the mean) and continues with blocks
 * 3-4 and the head on the same 128 rows, now one per point: no per-row HBM hand-off, no second launch.  out / pre_act (SB*P, 4) as pn_mlp_forward_tc.
 * workspace: pn_mlp_tc_grouped_workspace_bytes(), 256-byte aligned.  Inference only (the training forward keeps the
 * two-launch form whose saved operands the backward reads). */
PN_API int64_t pn_point_taps_grouped_rows(int64_t n_points, int n_views);
PN_API int pn_point_taps_grouped(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                          const float* xyz, const float* viewdirs, int64_t points_per_object,
                          void* feat_x_grouped, pn_stream_t stream);
PN_API size_t pn_mlp_tc_grouped_workspace_bytes(void);
PN_API int pn_mlp_forward_tc_grouped(const void* packed, const void* tables, int64_t n_texels,
                              const void* feat_x_grouped, int n_objects, int n_views, int64_t points_per_object,
                              float* out, float* pre_act, void* workspace, size_t workspace_bytes,
                              pn_stream_t stream);

/* The PRECISE tensor-core chain (mlp_precision = fp16x3): ResnetFC.forward + head (src/model/resnetfc.py:53-62, 132-184,
 * src/model/models.py:257-266) within the fp32 path's 1e-4 of the reference, on the tensor cores.  Every 512-wide
 * GEMM runs as three operand passes a_hi w_hi + a_hi w_lo + a_lo w_hi with a = a_hi + a_lo, w = w_hi + w_lo split into
 * fp16 pairs (22 significand bits each), fp32 accumulation and fp32 residual stream in tensor memory as in the fast
 * chain.  lin_in (K = 42) runs the same way on an fp16 PAIR of the positional code: pn_point_taps_f32 writes, next
 * to the usual taps rows feat_x (code rounded to fp16 = x_hi), a second (rows, 64) fp16 tensor feat_x_lo with the
 * remainder x_lo = fp16(x - x_hi) in columns [0,42) and zeros elsewhere (same row order), and the chain issues
 * x_hi w_hi + x_hi w_lo + x_lo w_hi.  lin_out + head are evaluated in fp32 on the CUDA cores, the lin_z tables stay
 * in fp32 (pn_prepare_latent_tables_f32: [3][n_texels][512] floats, same folded constants as the fp16 tables),
 * the view mean (n_views > 1: rows in GROUPED order, one launch) unrounded.  packed_x3: pn_prepare_weights_x3 into
 * pn_mlp_tc_packed_x3_bytes() bytes.  workspace: pn_mlp_tc_grouped_workspace_bytes().  Inference only. */
PN_API int pn_prepare_latent_tables_f32(const pn_mlp_weights_t* w, const float* latent_nhwc, int64_t n_texels,
                                 float* tables32, pn_stream_t stream);
PN_API int pn_point_taps_f32(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                      const float* xyz, const float* viewdirs, int64_t points_per_object,
                      void* feat_x, void* feat_x_lo, int grouped, pn_stream_t stream);
PN_API size_t pn_mlp_tc_packed_x3_bytes(void);
PN_API int pn_prepare_weights_x3(const pn_mlp_weights_t* w, void* packed, size_t packed_bytes, pn_stream_t stream);
PN_API int pn_mlp_forward_tc_x3(const void* packed_x3, const float* tables32, int64_t n_texels, const void* feat_x,
                         const void* feat_x_lo, int n_objects, int n_views, int64_t points_per_object,
                         float* out, float* pre_act, void* workspace, size_t workspace_bytes, pn_stream_t stream);

/* The TWO-PASS chain (mlp_precision = fp16x2): pn_mlp_forward_tc_x3 without its third operand pass -- every 512-wide
 * GEMM as a_hi w_hi + a_hi w_lo (weights as fp16 pairs, activations rounded to fp16 once), lin_z tables and head in
 * fp32 and lin_in on the fp16 pair of the code as in the precise chain.  Weight rounding and fp16 tables are the fast chain's two errors that are
 * coherent along a ray (the same perturbed network at every sample); activation rounding is independent per sample and
 * averages out in the composite.  This form stays within north_star's 2e-3 on every BASELINE config, DTU depth
 * included (measured <= 1.7e-3), for two thirds of the precise chain's tensor work.  Same arguments as
 * pn_mlp_forward_tc_x3.  Inference only. */
PN_API int pn_mlp_forward_tc_x2(const void* packed_x3, const float* tables32, int64_t n_texels, const void* feat_x,
                         const void* feat_x_lo, int n_objects, int n_views, int64_t points_per_object, float* out,
                         float* pre_act, void* workspace, size_t workspace_bytes, pn_stream_t stream);

/* pn_mlp_forward_tc for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and arithmetic as pn_mlp_forward_tc (bit-identical results), but the workspace is a
 * file-scope device array of PN_STATIC_SCRATCH_BYTES inside the library (csrc/pn_noalloc.cu) instead of a caller
 * buffer; the call performs no cudaMalloc / cudaFree and calls no allocating routine.  The caller still provides
 * every input and output buffer.
 *   - Limit: pn_mlp_tc_workspace_bytes(n_rows, n_views) must be <= PN_STATIC_SCRATCH_BYTES (32 MiB), i.e. any row
 *     count when n_views == 1 (256 bytes of scratch) and at most 32 767 (view, point) rows when n_views > 1 (1 KB
 *     per row); beyond that the call returns PN_ERR_INVALID with the sizes in pn_last_error().
 *   - Cost: every GPU that loads the library holds the 32 MiB until the process exits.
 *   - Sharing: all calls on a device use the same array, which keeps the previous call's contents; the call clears
 *     what must start at zero (the watchdog error word) itself.  Calls that overlap on the GPU (different streams
 *     or host threads on one device) RACE on the array: serialise them.  The launch uses the current device. */
#define PN_STATIC_SCRATCH_BYTES (32u << 20)
PN_API size_t pn_static_scratch_bytes(void);
PN_API int pn_mlp_forward_tc_static(const void* packed, const void* tables, int64_t n_texels,
                             const void* feat_x, int n_objects, int n_views, int64_t points_per_object,
                             float* out, float* pre_act, pn_stream_t stream);

/* ---- (3c) tensor-core GEMMs of the training path (dgrad / wgrad of the 512-wide linears) -- */

/* D[rows, n] = scale * A[rows, k] x B[k, n]  (both row-major 16-bit of the SAME format, 0 = fp16 or 1 = bf16; k % 64 == 0,
 * n % 256 == 0), fp32 accumulation on tcgen05.  Epilogue: optional relu-backward mask (mask_f16[rows, n], fp16 or bf16, is
 * the relu OUTPUT of the layer: D is zeroed where it is zero), optional `+= acc32` (accumulate != 0), then D is
 * written to acc32 (fp32) and/or out_bf16.  resid_bf16 (optional, [rows, n] bf16, may alias out_bf16) is added
 * to D before it is written; colsum0 / colsum1 (optional, [n] fp32, n <= 512) receive += the column sums of
 * the result (the bias gradient of the layer that produced A's consumer).  dgrad of a linear y = x W^T: A = dL/dy, B = W as stored ([out, in]).
 * err_word: optional device word receiving a watchdog code before a trap. */
PN_API int pn_gemm_tc_rows(const void* a, int a_fmt, const void* b, int b_fmt, int64_t n_rows, int k, int n,
                    const void* mask_f16, float* acc32, int accumulate, const void* resid_bf16, void* out_bf16,
                    float scale, float* colsum0, float* colsum1, unsigned int* err_word, pn_stream_t stream);

/* acc32[rows, n] (+)= scale * sum_s A_s[rows, 512] x B[s*512 .. s*512+512, n]: the same GEMM with its K = k dimension
 * spread over k/512 operands A_s stored one after the other, slab_rows rows apart (slab_rows % 256 == 0, >= n_rows;
 * rows past n_rows of a slab are never stored).  dL/dz = sum_b dL/dh_b W_z[b] of the MLP backward in one launch. */
PN_API int pn_gemm_tc_rows_slabs(const void* a_slabs, int64_t slab_rows, int fmt, const void* b, int64_t n_rows, int k,
                          int n, float* acc32, int accumulate, float scale, unsigned int* err_word,
                          pn_stream_t stream);

/* out[512, ldo] (fp32, first n_store columns) += G[rows, 512]^T x X[rows, ldx]  (ldx = 512, or 64 for lin_in's
 * padded code rows), split over row ranges, partial sums added with red.global.add: wgrad of y = x W^T with
 * G = dL/dy, X = the layer input.  The caller zeroes `out` first. */
PN_API int pn_gemm_tc_wgrad(const void* g, int g_fmt, int ldg, const void* x, int x_fmt, int ldx, int64_t n_rows,
                     int n_store, float* out, int ldo, unsigned int* err_word, pn_stream_t stream);

/* ---- (3d) tensor-core training path of the MLP (forward that keeps activations + backward) ------------- */

/* pn_mlp_forward_tc that also stores the bf16 relu'd operand of every 512-wide GEMM (11 slots of rows x 512:
 * slot 2b = relu(h_b), 2b+1 = relu(fc_0 out of block b), 10 = relu(h_5); slots 6.. hold n_objects*P rows when
 * n_views > 1; the slot pitch is `rows` rounded up to a multiple of 128).  `acts` needs
 * pn_mlp_tc_acts_bytes(rows) and 128-byte alignment. */
PN_API size_t pn_mlp_tc_acts_bytes(int64_t n_rows);
PN_API int pn_mlp_forward_tc_train(const void* packed, const void* tables, int64_t n_texels, const void* feat_x,
                            int n_objects, int n_views, int64_t points_per_object, float* out, float* pre_act,
                            void* workspace, size_t workspace_bytes, void* acts, size_t acts_bytes,
                            pn_stream_t stream);

/* bf16 row-major copies of fc_0[0..4], fc_1[0..4], lin_z[0..2] (the dgrad GEMMs' B operands). */
PN_API size_t pn_mlp_tc_w16_bytes(void);
PN_API int pn_prepare_weights_bf16(const pn_mlp_weights_t* w, void* w16, size_t w16_bytes, pn_stream_t stream);

/* Backward of ResnetFC + head (what autograd computes for resnetfc.py:132-184 / models.py:257-266) on the tensor
 * cores: g_out = dL/d out (n_objects*P, 4); every tensor of `grads` is accumulated into (+=); g_feat_z (optional,
 * rows x 512 fp32) receives dL/d(latent part of the MLP input) for pn_point_features_bwd.  feat_x is the forward's
 * (rows, 64) fp16 code rows with the embedded taps record; latent_nhwc the fp32 channels-last latent. */
PN_API size_t pn_mlp_backward_tc_workspace_bytes(int64_t n_rows, int n_views);
/* byte offset in `workspace` where the call leaves dL/dh_0 (rows x 512 bf16), for pn_depth_sample_bwd */
PN_API size_t pn_mlp_backward_tc_gh0_offset(int64_t n_rows, int n_views);
PN_API int pn_mlp_backward_tc(const pn_mlp_weights_t* w, const void* w16, const void* feat_x,
                       const float* latent_nhwc, int n_objects, int n_views, int64_t points_per_object,
                       const void* acts, const float* out, const float* g_out, const pn_mlp_grads_t* grads,
                       float* g_feat_z, void* workspace, size_t workspace_bytes, pn_stream_t stream);

/* ---- (5) either side of the render path --------------------------------------- */

/* Camera rays on the device: util.gen_rays (src/util/util.py:238-276) with unproj_map (:113-143) for
 * n_images camera-to-world poses (row-major, 4 floats per row, pose_stride floats per pose: 16 for 4x4,
 * 12 for 3x4).  rays (n_images, height, width, 8) = [origin, direction, z_near, z_far].  ndc=False only. */
PN_API int pn_gen_rays(const float* poses, int pose_stride, int n_images, int width, int height, float fx,
                float fy, float cx, float cy, float z_near, float z_far, float* rays, pn_stream_t stream);

/* (frames * 255).astype(np.uint8) of eval/gen_video.py:234-236 on the device: truncation, clamped to
 * [0,255].  n_values floats in, n_values bytes out. */
PN_API int pn_frame_to_u8(const float* rgb, unsigned char* out, int64_t n_values, pn_stream_t stream);

/* SpatialEncoder.forward's tail, src/model/encoder.py:150-163: every ResNet level map (n_maps, channels[l],
 * h[l], w[l]) fp32 NCHW is resized to level 0's size with F.interpolate(bilinear, align_corners=True), the
 * levels are concatenated over channels and the result is written channels-last (n_maps, h[0], w[0],
 * sum channels) in fp32 or fp16 -- the layout pn_prepare_latent would produce from the NCHW concat.
 * level_maps / level_channels / level_h / level_w are HOST arrays of n_levels (<= 4) entries; channel counts
 * must be multiples of 32 (ResNet34: 64, 64, 128, 256). */
PN_API int pn_encoder_latent(const float* const* level_maps, const int* level_channels, const int* level_h,
                      const int* level_w, int n_levels, int n_maps, void* latent_nhwc, int dtype,
                      pn_stream_t stream);

/* ---- (4) compositing ------------------------------------------------------------------- */

/* NeRFRenderer.composite after the model call, src/render/nerf.py:178-182, 228-249.
 * rgb_sigma (R,K,4), z (R,K), rays (R,8) -> weights (R,K), rgb (R,3), depth (R). */
PN_API int pn_composite_fwd(const float* rgb_sigma, const float* z, const float* rays,
                     float* weights, float* rgb, float* depth,
                     int64_t n_rays, int n_samples, int white_bkgd, pn_stream_t stream);

/* Hand-written backward of the above (autograd of nerf.py:228-249; formulas in
 * DESIGN.md): upstream g_rgb (R,3), g_depth (R) or NULL, g_weights (R,K) or NULL
 * -> g_rgb_sigma (R,K,4) and, if not NULL, g_z (R,K). */
PN_API int pn_composite_bwd(const float* rgb_sigma, const float* z, const float* rays,
                     const float* g_rgb, const float* g_depth, const float* g_weights,
                     float* g_rgb_sigma, float* g_z,
                     int64_t n_rays, int n_samples, int white_bkgd, pn_stream_t stream);

/* ---- (6) the whole render in one call ---------------------------------------------------- */

/* NeRFRenderer.forward without gradients, src/render/nerf.py:251-303, on the tensor-core path:
 * pn_sample_coarse -> pn_point_taps -> pn_mlp_forward_tc -> pn_composite_fwd -> pn_sample_fine ->
 * pn_point_taps -> pn_mlp_forward_tc -> pn_composite_fwd, enqueued back to back on `stream` with every
 * intermediate in `workspace` (pn_render_rays_workspace_bytes; 256-byte aligned).  n_imp = n_fine -
 * n_fine_depth importance samples and n_depth depth-guided ones; both 0 = coarse pass only. */
typedef struct pn_render_config {
  int n_coarse, n_imp, n_depth;
  float depth_std;   /* conf renderer.depth_std */
  int lindisp;       /* nerf.py:113-118 */
  int white_bkgd;    /* nerf.py:239-244 */
  uint64_t seed;     /* in-kernel draws when the matching pn_render_draws pointer is NULL */
} pn_render_config_t;

/* Injected draws in the order the reference consumes them (nerf.py:111,135,141,158); any may be NULL. */
typedef struct pn_render_draws {
  const float* u_coarse; /* (R,Kc) */
  const float* u_bin;    /* (R,n_imp) */
  const float* u_in;     /* (R,n_imp) */
  const float* noise;    /* (R,n_depth) */
} pn_render_draws_t;

/* One MLP as the chain kernel wants it: pn_prepare_weights + pn_prepare_latent_tables outputs. */
typedef struct pn_render_mlp {
  const void* packed;
  const void* tables;
} pn_render_mlp_t;

/* Outputs of one pass (nerf.py:305-316): rgb (R,3) and depth (R) are required; weights (R,K) and the
 * sample positions z (R,K) are optional (NULL = kept in the workspace only). */
typedef struct pn_render_out {
  float* rgb;
  float* depth;
  float* weights;
  float* z;
} pn_render_out_t;

/* Optional cudaEvent_t handles recorded on `stream` around the stages of pass 0 (coarse) and 1 (fine):
 * [pass][0] before, [pass][1] after.  NULL entries are skipped.  bench.py's per-kernel roofline times. */
typedef struct pn_render_events {
  void* taps[2][2];
  void* mlp[2][2];
  void* composite[2][2];
  void* sample[2][2];
} pn_render_events_t;

PN_API size_t pn_render_rays_workspace_bytes(int64_t n_rays, int n_views, const pn_render_config_t* cfg);

/* rays (R,8) ordered (object, ray) with R = cam->n_objects * rays-per-object; lin = device table
 * torch.linspace(0, 1-1/Kc, Kc) like pn_sample_coarse; mlp_fine may be NULL (the fine pass then runs
 * through mlp_coarse like models.py:242 when mlp_fine is None); out_fine is required iff n_imp+n_depth > 0.
 * The first 4 bytes of the workspace receive the chain kernel's watchdog code (see pn_mlp_forward_tc). */
PN_API int pn_render_rays(const pn_camera_t* cam, const pn_render_mlp_t* mlp_coarse, const pn_render_mlp_t* mlp_fine,
                   int64_t n_texels, const float* rays, int64_t n_rays, const float* lin,
                   const pn_render_config_t* cfg, const pn_render_draws_t* draws,
                   const pn_render_out_t* out_coarse, const pn_render_out_t* out_fine,
                   void* workspace, size_t workspace_bytes, const pn_render_events_t* events, pn_stream_t stream);

/* SURVEY 8f.1 -- ray generation fused into the renderer (util.gen_rays / unproj_map, src/util/util.py:113-143, 238-276,
 * feeding the loop of eval/gen_video.py:176-217): pn_render_rays with the rays of the call defined by FRAMES instead of
 * an (R, 8) tensor.  Ray r of the call is pixel first_ray + r of the concatenated (n_frames, height, width) grid; its
 * origin / unit direction are computed inside the kernels that need them from the frame's camera-to-world pose and the
 * intrinsics, with the arithmetic of pn_gen_rays (bit-identical results), near / far are the two constants.  A frame
 * costs 48-64 bytes of input instead of 32 bytes per ray.  One encoded object (cam->n_objects == 1).  Outputs are
 * indexed by r.  Workspace: pn_render_rays_workspace_bytes(n_rays, ...). */
typedef struct pn_frames {
  const float* poses;   /* DEVICE (n_frames, pose_stride) row-major camera-to-world, rows of 4 floats */
  int pose_stride;      /* 12 (3x4) or 16 (4x4) */
  int n_frames;
  int width, height;
  float fx, fy, cx, cy; /* focal lengths and principal point in pixels (util.unproj_map) */
  float z_near, z_far;
} pn_frames_t;
PN_API int pn_render_frames(const pn_camera_t* cam, const pn_render_mlp_t* mlp_coarse, const pn_render_mlp_t* mlp_fine,
                     int64_t n_texels, const pn_frames_t* frames, int64_t first_ray, int64_t n_rays,
                     const float* lin, const pn_render_config_t* cfg, const pn_render_draws_t* draws,
                     const pn_render_out_t* out_coarse, const pn_render_out_t* out_fine, void* workspace,
                     size_t workspace_bytes, const pn_render_events_t* events, pn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PIXELNERF_B200_H */
