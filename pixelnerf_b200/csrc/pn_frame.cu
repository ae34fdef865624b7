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
  const int img = (int)(i / per);
  const int pix = (int)(i - (int64_t)img * per);
  const int py = pix / width, px = pix - py * width;
  // unproj_map: ((x - cx) / fx, -(y - cy) / fy, -1), normalised
  const float X = __fdiv_rn((float)px - cx, fx);
  const float Y = -__fdiv_rn((float)py - cy, fy);
  const float Z = -1.0f;
  const float n = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(X, X), __fmul_rn(Y, Y)), 1.0f));
  const float ux = __fdiv_rn(X, n), uy = __fdiv_rn(Y, n), uz = __fdiv_rn(Z, n);
  const float* P = poses + (size_t)img * pose_stride;   // row-major camera-to-world, 4 floats per row
  const float dx = P[0] * ux + P[1] * uy + P[2] * uz;
  const float dy = P[4] * ux + P[5] * uy + P[6] * uz;
  const float dz = P[8] * ux + P[9] * uy + P[10] * uz;
  rays[i * 2] = make_float4(P[3], P[7], P[11], dx);
  rays[i * 2 + 1] = make_float4(dy, dz, z_near, z_far);
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

}  // namespace pn

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
