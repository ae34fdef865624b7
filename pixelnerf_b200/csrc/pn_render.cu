// pn_render_rays: the whole no-grad render (reference src/render/nerf.py:251-303) enqueued from one C call.
// Host-side sequencing only -- every kernel lives in its own file and is reached through the same C entry
// points a caller could use one by one; this removes the per-stage Python/ctypes/allocator work from the
// frame (8 launches back to back on one stream, intermediates in a caller-owned workspace).
#include <nvtx3/nvToolsExt.h>

#include "pn_common.cuh"

namespace pn {
namespace render {

// NVTX ranges carrying the reference's profiler names (torch.autograd.profiler.record_function at
// src/render/nerf.py:175,264, src/model/models.py:156, src/model/encoder.py:90, src/model/resnetfc.py:139), so a
// timeline of this path reads like one of the reference.  Header-only NVTX: a no-op unless a profiler is attached.
struct Range {
  explicit Range(const char* name) { nvtxRangePushA(name); }
  ~Range() { nvtxRangePop(); }
};

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct Layout {
  size_t mlp, z_c, w_c, z_f, w_f, rgbs, fx, total;
};

static Layout layout(int64_t n_rays, int n_views, const pn_render_config_t& c) {
  const int k_f = c.n_imp + c.n_depth;
  const int k2 = c.n_coarse + k_f;
  const int k_max = k_f > 0 ? k2 : c.n_coarse;
  const int64_t r = n_rays > 0 ? n_rays : 0;
  Layout l;
  size_t off = 0;
  // one view: the chain's watchdog word; several views: the grouped chain's park buffer (view mean on chip)
  l.mlp = off;  off += align256(n_views > 1 ? pn_mlp_tc_grouped_workspace_bytes()
                                            : pn_mlp_tc_workspace_bytes(r * n_views * k_max, n_views));
  l.z_c = off;  off += align256((size_t)r * c.n_coarse * 4);
  l.w_c = off;  off += align256((size_t)r * c.n_coarse * 4);
  l.z_f = off;  off += align256((size_t)r * (k_f > 0 ? k2 : 0) * 4);
  l.w_f = off;  off += align256((size_t)r * (k_f > 0 ? k2 : 0) * 4);
  l.rgbs = off; off += align256((size_t)r * k_max * 16);
  l.fx = off;   off += align256((size_t)(n_views > 1 ? pn_point_taps_grouped_rows(r * k_max, n_views)
                                                       : r * n_views * k_max) * 128);
  l.total = off;
  return l;
}

static inline int record(void* ev, cudaStream_t st) {
  if (!ev) return PN_OK;
  PN_CUDA_OK(cudaEventRecord((cudaEvent_t)ev, st));
  return PN_OK;
}

#define PN_TRY(expr)            \
  do {                          \
    int _rc = (expr);           \
    if (_rc != PN_OK) return _rc; \
  } while (0)

// taps -> chain -> composite for one pass
static int pass(const pn_camera_t* cam, const pn_render_mlp_t* mlp, int64_t n_texels, const RaySrc& rays, int64_t n_rays,
                const float* z, int k, int white, float* rgbs, void* fx, void* mlp_ws, size_t mlp_ws_bytes,
                float* weights, float* rgb, float* depth, const pn_render_events_t* ev, int which, cudaStream_t st) {
  const int64_t ppo = n_rays / cam->n_objects * k;
  Range r_comp("renderer_composite");
  {
    Range r_model("model_inference");
    PN_TRY(record(ev ? ev->taps[which][0] : nullptr, st));
    const bool grouped = cam->n_views > 1;   // several views: grouped row order + on-chip view mean, one chain launch
    {
      Range r_index("encoder_index");     // projection + taps (+ positional_enc: same kernel)
      PN_TRY(point_taps_src(cam, rays, z, k, nullptr, nullptr, ppo, fx, grouped, st));
    }
    PN_TRY(record(ev ? ev->taps[which][1] : nullptr, st));
    PN_TRY(record(ev ? ev->mlp[which][0] : nullptr, st));
    {
      Range r_mlp("resnetfc_infer");
      if (grouped)
        PN_TRY(pn_mlp_forward_tc_grouped(mlp->packed, mlp->tables, n_texels, fx, cam->n_objects, cam->n_views, ppo, rgbs,
                                         nullptr, mlp_ws, mlp_ws_bytes, st));
      else
        PN_TRY(pn_mlp_forward_tc(mlp->packed, mlp->tables, n_texels, fx, cam->n_objects, cam->n_views, ppo, rgbs, nullptr,
                                 mlp_ws, mlp_ws_bytes, nullptr, st));
    }
    PN_TRY(record(ev ? ev->mlp[which][1] : nullptr, st));
  }
  PN_TRY(record(ev ? ev->composite[which][0] : nullptr, st));
  PN_TRY(composite_fwd_src(rgbs, z, rays, weights, rgb, depth, n_rays, k, white, st));
  PN_TRY(record(ev ? ev->composite[which][1] : nullptr, st));
  return PN_OK;
}

}  // namespace render
}  // namespace pn

extern "C" size_t pn_render_rays_workspace_bytes(int64_t n_rays, int n_views, const pn_render_config_t* cfg) {
  if (!cfg || n_views <= 0 || cfg->n_coarse <= 0 || cfg->n_imp < 0 || cfg->n_depth < 0) return 0;
  return pn::render::layout(n_rays, n_views, *cfg).total;
}

static int render_impl(const pn_camera_t* cam, const pn_render_mlp_t* mlp_coarse, const pn_render_mlp_t* mlp_fine,
                       int64_t n_texels, const pn::RaySrc& rays, int64_t n_rays, const float* lin,
                       const pn_render_config_t* cfg, const pn_render_draws_t* draws,
                       const pn_render_out_t* out_coarse, const pn_render_out_t* out_fine, void* workspace,
                       size_t workspace_bytes, const pn_render_events_t* events, pn_stream_t stream) {
  using namespace pn::render;
  using namespace pn;
  PN_CHECK_ARG(cam && mlp_coarse && cfg && out_coarse, "pn_render_rays: null pointer");
  PN_CHECK_ARG(cfg->n_coarse > 0 && cfg->n_imp >= 0 && cfg->n_depth >= 0, "pn_render_rays: bad sample counts");
  PN_CHECK_ARG(cam->n_objects > 0 && cam->n_views > 0, "pn_render_rays: bad camera sizes");
  PN_CHECK_ARG(n_rays >= 0 && n_rays % cam->n_objects == 0,
               "pn_render_rays: %lld rays do not split over %d objects", (long long)n_rays, cam->n_objects);
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers (nerf.py:23-27)
  const int k_f = cfg->n_imp + cfg->n_depth;
  const int k2 = cfg->n_coarse + k_f;
  PN_CHECK_ARG((rays.rays || rays.poses) && lin && workspace, "pn_render_rays: null pointer");
  PN_CHECK_ARG(mlp_coarse->packed && mlp_coarse->tables, "pn_render_rays: coarse MLP not prepared");
  PN_CHECK_ARG(!mlp_fine || (mlp_fine->packed && mlp_fine->tables), "pn_render_rays: fine MLP not prepared");
  PN_CHECK_ARG(out_coarse->rgb && out_coarse->depth, "pn_render_rays: coarse rgb/depth outputs are required");
  PN_CHECK_ARG(k_f == 0 || (out_fine && out_fine->rgb && out_fine->depth),
               "pn_render_rays: fine rgb/depth outputs are required when n_imp + n_depth > 0");
  PN_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "pn_render_rays: workspace must be 256-byte aligned");
  const Layout l = layout(n_rays, cam->n_views, *cfg);
  PN_CHECK_ARG(workspace_bytes >= l.total, "pn_render_rays: workspace too small (%zu < %zu)", workspace_bytes, l.total);
  cudaStream_t st = (cudaStream_t)stream;
  Range r_forward("renderer_forward");
  char* ws = (char*)workspace;
  const size_t mlp_bytes = l.z_c - l.mlp;
  float* z_c = out_coarse->z ? out_coarse->z : (float*)(ws + l.z_c);
  float* w_c = out_coarse->weights ? out_coarse->weights : (float*)(ws + l.w_c);
  float* rgbs = (float*)(ws + l.rgbs);
  void* fx = ws + l.fx;

  PN_TRY(record(events ? events->sample[0][0] : nullptr, st));
  PN_TRY(sample_coarse_src(rays, draws ? draws->u_coarse : nullptr, lin, z_c, n_rays, cfg->n_coarse, cfg->lindisp,
                           cfg->seed, st));
  PN_TRY(record(events ? events->sample[0][1] : nullptr, st));
  PN_TRY(pass(cam, mlp_coarse, n_texels, rays, n_rays, z_c, cfg->n_coarse, cfg->white_bkgd, rgbs, fx, ws + l.mlp,
              mlp_bytes, w_c, out_coarse->rgb, out_coarse->depth, events, 0, st));
  if (k_f == 0) return PN_OK;

  float* z_f = out_fine->z ? out_fine->z : (float*)(ws + l.z_f);
  float* w_f = out_fine->weights ? out_fine->weights : (float*)(ws + l.w_f);
  PN_TRY(record(events ? events->sample[1][0] : nullptr, st));
  PN_TRY(sample_fine_src(rays, z_c, w_c, out_coarse->depth, draws ? draws->u_bin : nullptr,
                         draws ? draws->u_in : nullptr, draws ? draws->noise : nullptr, z_f, nullptr, nullptr, n_rays,
                         cfg->n_coarse, cfg->n_imp, cfg->n_depth, cfg->depth_std, cfg->lindisp, cfg->seed, st));
  PN_TRY(record(events ? events->sample[1][1] : nullptr, st));
  PN_TRY(pass(cam, mlp_fine ? mlp_fine : mlp_coarse, n_texels, rays, n_rays, z_f, k2, cfg->white_bkgd, rgbs, fx,
              ws + l.mlp, mlp_bytes, w_f, out_fine->rgb, out_fine->depth, events, 1, st));
  return PN_OK;
}

extern "C" int pn_render_rays(const pn_camera_t* cam, const pn_render_mlp_t* mlp_coarse, const pn_render_mlp_t* mlp_fine,
                              int64_t n_texels, const float* rays, int64_t n_rays, const float* lin,
                              const pn_render_config_t* cfg, const pn_render_draws_t* draws,
                              const pn_render_out_t* out_coarse, const pn_render_out_t* out_fine, void* workspace,
                              size_t workspace_bytes, const pn_render_events_t* events, pn_stream_t stream) {
  PN_CHECK_ARG(rays || n_rays == 0, "pn_render_rays: null pointer");
  return render_impl(cam, mlp_coarse, mlp_fine, n_texels, pn::ray_src(rays), n_rays, lin, cfg, draws, out_coarse, out_fine,
                     workspace, workspace_bytes, events, stream);
}

// ray generation fused into the renderer.  The rays of the call are pixels [first_ray, first_ray + n_rays)
// of the concatenated (n_frames, H, W) grid of `frames`; every kernel that needs a ray computes it from the frame's pose
// and the intrinsics (pn_common.cuh: ray_load, the arithmetic of pn_gen_rays), so no (H, W, 8) tensor exists.
extern "C" int pn_render_frames(const pn_camera_t* cam, const pn_render_mlp_t* mlp_coarse, const pn_render_mlp_t* mlp_fine,
                                int64_t n_texels, const pn_frames_t* frames, int64_t first_ray, int64_t n_rays,
                                const float* lin, const pn_render_config_t* cfg, const pn_render_draws_t* draws,
                                const pn_render_out_t* out_coarse, const pn_render_out_t* out_fine, void* workspace,
                                size_t workspace_bytes, const pn_render_events_t* events, pn_stream_t stream) {
  PN_CHECK_ARG(frames && frames->poses, "pn_render_frames: null pointer");
  PN_CHECK_ARG(frames->pose_stride == 12 || frames->pose_stride == 16, "pn_render_frames: poses must be (n, 3|4, 4) row-major");
  PN_CHECK_ARG(frames->n_frames > 0 && frames->width > 0 && frames->height > 0, "pn_render_frames: bad frame sizes");
  PN_CHECK_ARG(frames->fx != 0.0f && frames->fy != 0.0f, "pn_render_frames: zero focal length");
  PN_CHECK_ARG(cam && cam->n_objects == 1, "pn_render_frames: one encoded object (a video of one scene)");
  const int64_t total = (int64_t)frames->n_frames * frames->width * frames->height;
  PN_CHECK_ARG(first_ray >= 0 && n_rays >= 0 && first_ray + n_rays <= total,
               "pn_render_frames: rays [%lld, %lld) outside the %lld pixels of the frames", (long long)first_ray,
               (long long)(first_ray + n_rays), (long long)total);
  pn::RaySrc src = {};
  src.poses = frames->poses; src.pose_stride = frames->pose_stride;
  src.width = frames->width; src.height = frames->height;
  src.fx = frames->fx; src.fy = frames->fy; src.cx = frames->cx; src.cy = frames->cy;
  src.z_near = frames->z_near; src.z_far = frames->z_far; src.first_ray = first_ray;
  return render_impl(cam, mlp_coarse, mlp_fine, n_texels, src, n_rays, lin, cfg, draws, out_coarse, out_fine, workspace,
                     workspace_bytes, events, stream);
}
