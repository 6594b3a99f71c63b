// taa.cu — TAA resolve: velocity reprojection, 4-neighbour history clamp, 0.1 blend,
// world-position rejection.  Replaces task "TAA" (reference src/taa.cpp:14-63,
// shaders/taa/resolve.comp:20-77).  target is RGBA16F (the image's real format, App. A.6 #1).
#include "common.cuh"

namespace vkr {

struct TaaK {
  Mat4 inverse_camera, prev_inverse_camera;
  CamK cam;
};

__device__ __forceinline__ f3 unpack_rgb16f(uint2 k) {
  float2 a = unpack_rg16f(k.x), b = unpack_rg16f(k.y);
  return mk3(a.x, a.y, b.x);
}

__device__ __forceinline__ f3 bilerp3(f3 a, f3 b, f3 c, f3 d, float wx, float wy) {
  return mk3(bilerp(a.x, b.x, c.x, d.x, wx, wy), bilerp(a.y, b.y, c.y, d.y, wx, wy), bilerp(a.z, b.z, c.z, d.z, wx, wy));
}

__device__ __forceinline__ int clampi(int v, int hi) { return min(max(v, 0), hi); }

// resolve.comp:72-77
__device__ __forceinline__ f3 world_pos(const Mip& depth, const Mat4& inv, float u, float v, const CamK& cam) {
  float d = sample_d24(depth, u, v);
  return mul_m4(inv, reconstruct_view_vec(u, v, d, cam), 1.0f);
}

__global__ void __launch_bounds__(256) taa_resolve_kernel(TaaK k, Mip history, Mip prev_depth, Mip depth, Mip velocity, Mip color, Mip out,
                                                          const float* __restrict__ srgb) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= out.w || py >= out.h) return;
  const float su = ((float)px + 0.5f) / (float)out.w, sv = ((float)py + 0.5f) / (float)out.h;

  float4 cc = sample_srgb8(color, su, sv, srgb);
  f3 current_color = mk3(cc.x, cc.y, cc.z);
  float2 vel = sample_rg16f(velocity, su, sv);
  float delta_len = sqrtf(FA(FM(vel.x, vel.x), FM(vel.y, vel.y)));
  bool reprojected = false;
  f3 out_color = mk3(0.f, 0.f, 0.f);
  float pu = su + vel.x, pv = sv + vel.y;

  if (pu >= 0 && pv >= 0 && pu <= 1 && pv <= 1) {
    // The five bilinear footprints (centre + 4 textureOffset neighbours, :41-45) share a
    // plus-shaped set of 12 texels; fetch each once.
    const int W = history.w, H = history.h;
    float x = __fsub_rn(__fmul_rn(pu, (float)W), 0.5f), y = __fsub_rn(__fmul_rn(pv, (float)H), 0.5f);
    float fx = floorf(x), fy = floorf(y);
    float wx = __fsub_rn(x, fx), wy = __fsub_rn(y, fy);
    int ix = (int)fx, iy = (int)fy;
    int xs[4] = {clampi(ix - 1, W - 1), clampi(ix, W - 1), clampi(ix + 1, W - 1), clampi(ix + 2, W - 1)};
    int ys[4] = {clampi(iy - 1, H - 1), clampi(iy, H - 1), clampi(iy + 1, H - 1), clampi(iy + 2, H - 1)};
    f3 t[4][4];
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
      for (int i = 0; i < 4; i++)
        if (!((i == 0 || i == 3) && (j == 0 || j == 3))) t[j][i] = unpack_rgb16f(ld64(history, xs[i], ys[j]));
    f3 hist = bilerp3(t[1][1], t[1][2], t[2][1], t[2][2], wx, wy);
    f3 c0 = bilerp3(t[1][2], t[1][3], t[2][2], t[2][3], wx, wy);   // offset (+1, 0)
    f3 c1 = bilerp3(t[2][1], t[2][2], t[3][1], t[3][2], wx, wy);   // offset (0, +1)
    f3 c2 = bilerp3(t[1][0], t[1][1], t[2][0], t[2][1], wx, wy);   // offset (-1, 0)
    f3 c3 = bilerp3(t[0][1], t[0][2], t[1][1], t[1][2], wx, wy);   // offset (0, -1)
    f3 cmin = mk3(gmin(c0.x, gmin(c1.x, gmin(c2.x, c3.x))), gmin(c0.y, gmin(c1.y, gmin(c2.y, c3.y))), gmin(c0.z, gmin(c1.z, gmin(c2.z, c3.z))));
    f3 cmax = mk3(gmax(c0.x, gmax(c1.x, gmax(c2.x, c3.x))), gmax(c0.y, gmax(c1.y, gmax(c2.y, c3.y))), gmax(c0.z, gmax(c1.z, gmax(c2.z, c3.z))));
    hist = mk3(gclamp(hist.x, cmin.x, cmax.x), gclamp(hist.y, cmin.y, cmax.y), gclamp(hist.z, cmin.z, cmax.z));
    out_color = mix3(hist, current_color, 0.1f);

    f3 v_world_cur = world_pos(depth, k.inverse_camera, su, sv, k.cam);
    f3 v_world_prev = world_pos(prev_depth, k.prev_inverse_camera, pu, pv, k.cam);
    f3 v_camera = mk3(k.inverse_camera.m[12], k.inverse_camera.m[13], k.inverse_camera.m[14]);
    float error = len3(v_world_cur - v_world_prev);
    float pixel_dist = len3(v_world_cur - v_camera);
    reprojected = (delta_len < 0.005f) || (error < gclamp(FM(FM(0.1f, pixel_dist), delta_len), 0.01f, 0.2f));
  }
  if (!reprojected) out_color = current_color;
  *((uint2*)(out.p + (size_t)py * out.pitch) + px) = make_uint2(pack_half2(out_color.x, out_color.y), pack_half2(out_color.z, 0.f));
}

}  // namespace vkr
using namespace vkr;

extern "C" int vkr_taa_resolve(const VkrTaaParams* p, const VkrImage* history, const VkrImage* prev_depth, const VkrImage* depth,
                               const VkrImage* velocity, const VkrImage* color, const VkrImage* out_target, void* stream) {
  if (!p || !image_ok(history, VKR_FORMAT_R16G16B16A16_SFLOAT) || !image_ok(prev_depth, VKR_FORMAT_D24_UNORM_S8_UINT) ||
      !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(velocity, VKR_FORMAT_R16G16_SFLOAT) ||
      !image_ok(color, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(out_target, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("taa_resolve: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (history->base == out_target->base) { set_error("taa_resolve: history and target alias"); return VKR_ERR_INVALID_ARGUMENT; }
  TaaK k;
  k.inverse_camera = make_mat(p->inverse_camera); k.prev_inverse_camera = make_mat(p->prev_inverse_camera);
  k.cam = make_cam(p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]);
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("taa_resolve: sRGB table upload failed"); return VKR_ERR_CUDA; }
  dim3 grid((out_target->width + 31) / 32, (out_target->height + 7) / 8);
  taa_resolve_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(k, make_mip(history, 0), make_mip(prev_depth, 0), make_mip(depth, 0),
                                                             make_mip(velocity, 0), make_mip(color, 0), make_mip(out_target, 0), srgb);
  return check_launch("taa_resolve_kernel");
}
