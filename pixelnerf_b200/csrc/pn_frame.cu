// Either side of the render path: camera rays generated on the device from a pose and the
// intrinsics (reference src/util/util.py:113-143 unproj_map, :238-276 gen_rays) so a frame needs ~100 bytes of
// input instead of 32 bytes per ray, and the rendered frame converted to 8-bit on the device before the
// device-to-host copy (reference eval/gen_video.py:234-236: (frames * 255).astype(np.uint8)).
#include "pn_common.cuh"

namespace pn {

__global__ void __launch_bounds__(256)
gen_rays_kernel(const float* __restrict__ poses, int pose_stride, int n_images, int width, int height,
                float fx, float fy, float cx, float cy, float z_near, float z_far, float4* __restrict__ rays) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t per = (int64_t)width * height;
  if (i >= per * n_images) return;
  // one definition of the camera model for the stand-alone kernel and the fused ray source (pn_common.cuh: ray_load)
  RaySrc src = {};
  src.poses = poses; src.pose_stride = pose_stride; src.width = width; src.height = height;
  src.fx = fx; src.fy = fy; src.cx = cx; src.cy = cy; src.z_near = z_near; src.z_far = z_far; src.first_ray = 0;
  float4 a4, b4;
  ray_load(src, i, a4, b4);
  rays[i * 2] = a4;
  rays[i * 2 + 1] = b4;
}

__global__ void __launch_bounds__(256)
frame_to_u8_kernel(const float* __restrict__ src, uint8_t* __restrict__ dst, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (i + 3 < n && (((uintptr_t)(src + i)) & 15) == 0) {
    const float4 t = *reinterpret_cast<const float4*>(src + i);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  } else {
    for (int j = 0; j < 4 && i + j < n; ++j) v[j] = src[i + j];
  }
  for (int j = 0; j < 4 && i + j < n; ++j) {
    // numpy's float -> uint8 cast truncates; values are clamped so that an out-of-range colour cannot wrap
    const float s = fminf(fmaxf(__fmul_rn(v[j], 255.0f), 0.0f), 255.0f);
    dst[i + j] = (uint8_t)(int)s;
  }
}

// Encoder hand-off: SpatialEncoder.forward's tail (reference src/model/encoder.py:150-163) --
// F.interpolate(level, size of level 0, bilinear, align_corners=True) for each ResNet level, torch.cat over
// channels -- fused with the NCHW -> channels-last conversion the gather kernels want (pn_prepare_latent).
// One block = one output row segment of 32 texels x 32 channels of one level: the four source taps are read
// with x fastest (neighbouring lanes read neighbouring or identical source texels), blended with ATen's
// index arithmetic (scale = (in-1)/(out-1), i0 = int(scale*dst), lambda = scale*dst - i0), transposed through
// shared memory and written with the channel fastest (128 contiguous bytes per warp row).
struct EncLevels {
  const float* map[4];
  int ch0[5];          // first output channel of each level (prefix sums), ch0[n_levels] = total
  int h[4], w[4];
  int n_levels;
};

template <typename OutT>
__global__ void __launch_bounds__(256)
encoder_latent_kernel(EncLevels lv, OutT* __restrict__ out, int n_ctiles) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int v = blockIdx.z / n_ctiles, ct = blockIdx.z - v * n_ctiles;
  const int c0 = ct * 32, y = blockIdx.y, x0 = blockIdx.x * 32;
  const int H = lv.h[0], W = lv.w[0], ctot = lv.ch0[lv.n_levels];
  int l = 0;
#pragma unroll
  for (int i = 1; i < 4; ++i)
    if (i < lv.n_levels && c0 >= lv.ch0[i]) l = i;
  const int hs = lv.h[l], ws = lv.w[l], cl = lv.ch0[l + 1] - lv.ch0[l];
  const float sh = H > 1 ? __fdiv_rn((float)(hs - 1), (float)(H - 1)) : 0.0f;
  const float sw = W > 1 ? __fdiv_rn((float)(ws - 1), (float)(W - 1)) : 0.0f;
  const float fy = __fmul_rn(sh, (float)y);
  const int y0 = (int)fy, y1 = y0 + (y0 < hs - 1 ? 1 : 0);
  const float ly1 = __fsub_rn(fy, (float)y0), ly0 = __fsub_rn(1.0f, ly1);
  const int x = x0 + tx;
  const float fx = __fmul_rn(sw, (float)x);
  const int xs0 = (int)fx, xs1 = xs0 + (xs0 < ws - 1 ? 1 : 0);
  const float lx1 = __fsub_rn(fx, (float)xs0), lx0 = __fsub_rn(1.0f, lx1);
  const float* src = lv.map[l] + ((size_t)v * cl + (c0 - lv.ch0[l])) * hs * ws;
  if (x < W) {
    for (int i = ty; i < 32; i += 8) {
      const float* m = src + (size_t)i * hs * ws;
      const float v00 = m[y0 * ws + xs0], v01 = m[y0 * ws + xs1], v10 = m[y1 * ws + xs0], v11 = m[y1 * ws + xs1];
      const float top = __fadd_rn(__fmul_rn(lx0, v00), __fmul_rn(lx1, v01));
      const float bot = __fadd_rn(__fmul_rn(lx0, v10), __fmul_rn(lx1, v11));
      tile[i][tx] = __fadd_rn(__fmul_rn(ly0, top), __fmul_rn(ly1, bot));
    }
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int xo = x0 + i;
    if (xo < W) out[(((size_t)v * H + y) * W + xo) * ctot + c0 + tx] = (OutT)tile[tx][i];
  }
}

}  // namespace pn

extern "C" int pn_encoder_latent(const float* const* level_maps, const int* level_channels, const int* level_h,
                                 const int* level_w, int n_levels, int n_maps, void* latent_nhwc, int dtype,
                                 pn_stream_t stream) {
  PN_CHECK_ARG(level_maps && level_channels && level_h && level_w && latent_nhwc, "pn_encoder_latent: null pointer");
  PN_CHECK_ARG(n_levels >= 1 && n_levels <= 4 && n_maps >= 0, "pn_encoder_latent: 1..4 levels");
  PN_CHECK_ARG(dtype == PN_DTYPE_F32 || dtype == PN_DTYPE_F16, "pn_encoder_latent: dtype must be f32 or f16");
  pn::EncLevels lv;
  lv.n_levels = n_levels;
  lv.ch0[0] = 0;
  for (int i = 0; i < 4; ++i) {
    const bool on = i < n_levels;
    lv.map[i] = on ? level_maps[i] : nullptr;
    lv.h[i] = on ? level_h[i] : 1;
    lv.w[i] = on ? level_w[i] : 1;
    lv.ch0[i + 1] = lv.ch0[i] + (on ? level_channels[i] : 0);
    PN_CHECK_ARG(!on || (level_maps[i] && level_channels[i] > 0 && level_channels[i] % 32 == 0 && level_h[i] > 0 &&
                         level_w[i] > 0),
                 "pn_encoder_latent: level %d needs a map, positive sizes and a channel count divisible by 32", i);
  }
  if (n_maps == 0) return PN_OK;
  const int n_ctiles = lv.ch0[n_levels] / 32;
  PN_CHECK_ARG((int64_t)n_maps * n_ctiles <= 65535 && lv.h[0] <= 65535, "pn_encoder_latent: grid too large");
  dim3 grid(pn::ceil_div(lv.w[0], 32), lv.h[0], n_maps * n_ctiles), block(32, 8);
  if (dtype == PN_DTYPE_F32)
    pn::encoder_latent_kernel<float><<<grid, block, 0, (cudaStream_t)stream>>>(lv, (float*)latent_nhwc, n_ctiles);
  else
    pn::encoder_latent_kernel<__half><<<grid, block, 0, (cudaStream_t)stream>>>(lv, (__half*)latent_nhwc, n_ctiles);
  PN_LAUNCH_OK("encoder_latent_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_gen_rays(const float* poses, int pose_stride, int n_images, int width, int height, float fx,
                           float fy, float cx, float cy, float z_near, float z_far, float* rays, pn_stream_t stream) {
  PN_CHECK_ARG(poses && rays, "pn_gen_rays: null pointer");
  PN_CHECK_ARG(n_images >= 0 && width > 0 && height > 0 && pose_stride >= 12, "pn_gen_rays: bad sizes");
  PN_CHECK_ARG(fx != 0.0f && fy != 0.0f, "pn_gen_rays: zero focal length");
  PN_CHECK_ARG(((uintptr_t)rays & 15) == 0, "pn_gen_rays: rays must be 16-byte aligned");
  const int64_t total = (int64_t)n_images * width * height;
  if (total == 0) return PN_OK;
  pn::gen_rays_kernel<<<pn::ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
      poses, pose_stride, n_images, width, height, fx, fy, cx, cy, z_near, z_far, reinterpret_cast<float4*>(rays));
  PN_LAUNCH_OK("gen_rays_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_frame_to_u8(const float* rgb, unsigned char* out, int64_t n_values, pn_stream_t stream) {
  if (n_values == 0) return PN_OK;
  PN_CHECK_ARG(rgb && out && n_values > 0, "pn_frame_to_u8: null pointer");
  pn::frame_to_u8_kernel<<<pn::ceil_div(pn::ceil_div(n_values, 4), 256), 256, 0, (cudaStream_t)stream>>>(rgb, out, n_values);
  PN_LAUNCH_OK("frame_to_u8_kernel");
  pn::count_launch();
  return PN_OK;
}
