// synth.cu — deterministic synthetic G-buffer producer (see include/vkr_synth.h).
// Compiled with -fmad=false so the device kernel and the host twin evaluate the same
// IEEE fp32 expressions; the only double-precision steps (depth quantisation) are
// identical on both sides as well.
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>
#include "common.cuh"
#include "../../include/vkr_synth.h"

namespace vkr {

struct SynthFrameK {
  float inv_view[16], view[16], mvp[16], prev_mvp[16];
  float aspect, tg, zn, zf, jx, jy;
  int W, H;
};

struct SynthOut { uint32_t albedo, normal, material, velocity, depth; };

__host__ __device__ inline void xform(const float* M, float x, float y, float z, float w, float* o) {
  for (int r = 0; r < 4; r++) o[r] = M[r] * x + M[4 + r] * y + M[8 + r] * z + M[12 + r] * w;
}

__host__ __device__ inline uint32_t enc_unorm16(float v) {
  if (!(v > 0.f)) return 0u;
  if (v > 1.f) v = 1.f;
  return (uint32_t)floorf(v * 65535.0f + 0.5f);
}

__host__ __device__ inline uint32_t enc_normal(float nx, float ny, float nz) {  // gbuffer_encode.glsl:17-27
  float l1 = fabsf(nx) + fabsf(ny) + fabsf(nz);
  float inv = 1.0f / l1;
  float rx = nx * inv, ry = ny * inv;
  if (nz < 0.f) {
    float tx = (1.0f - fabsf(ry)) * (rx >= 0.f ? 1.f : -1.f);
    float ty = (1.0f - fabsf(rx)) * (ry >= 0.f ? 1.f : -1.f);
    rx = tx; ry = ty;
  }
  return enc_unorm16(0.5f * rx + 0.5f) | (enc_unorm16(0.5f * ry + 0.5f) << 16);
}

__host__ __device__ inline uint32_t enc_half2(float a, float b) {
  __half ha = __float2half_rn(a), hb = __float2half_rn(b);
  unsigned short ua, ub;
  memcpy(&ua, &ha, 2); memcpy(&ub, &hb, 2);
  return (uint32_t)ua | ((uint32_t)ub << 16);
}

__host__ __device__ inline SynthOut synth_pixel(const VkrSynthScene& sc, const SynthFrameK& f, int px, int py) {
  // un-jittered NDC of the surface point seen through this pixel centre (opaque_taa.vert:40)
  float ndx = (2.0f * ((float)px + 0.5f) / (float)f.W - 1.0f) - f.jx;
  float ndy = (2.0f * ((float)py + 0.5f) / (float)f.H - 1.0f) - f.jy;
  float dv[3] = {ndx * f.aspect * f.tg, ndy * f.tg, -1.0f};  // view-space ray direction (z = -1)
  float o4[4], d4[4];
  xform(f.inv_view, 0.f, 0.f, 0.f, 1.f, o4);
  xform(f.inv_view, dv[0], dv[1], dv[2], 0.f, d4);
  const float ox = o4[0], oy = o4[1], oz = o4[2], dx = d4[0], dy = d4[1], dz = d4[2];

  float best = 3.0e38f;
  float nx = 0.f, ny = 0.f, nz = 1.f;
  uint32_t alb = sc.sky_albedo, mat = 0xFF000000u;
  bool hit = false;

  if (dy < 0.f) {  // ground y = 0
    float t = -oy / dy;
    float gx = ox + t * dx, gz = oz + t * dz;
    if (t > 0.f && t < best && fabsf(gx) < 40.f && gz < sc.wall_z && gz > -40.f) {
      best = t; hit = true; nx = 0.f; ny = 1.f; nz = 0.f;
      int cx = (int)floorf(gx), cz = (int)floorf(gz);
      alb = sc.ground_albedo[(cx + cz) & 1]; mat = sc.ground_material;
    }
  }
  if (dz > 0.f) {  // back wall z = wall_z
    float t = (sc.wall_z - oz) / dz;
    float wx = ox + t * dx, wy = oy + t * dy;
    if (t > 0.f && t < best && fabsf(wx) <= sc.wall_half_width && wy >= 0.f && wy <= sc.wall_height) {
      best = t; hit = true; nx = 0.f; ny = 0.f; nz = -1.f; alb = sc.wall_albedo; mat = sc.wall_material;
    }
  }
  for (uint32_t i = 0; i < sc.object_count; i++) {
    const VkrSynthObject& ob = sc.objects[i];
    float cx = ob.center[0] - ox, cy = ob.center[1] - oy, cz = ob.center[2] - oz;
    if (ob.kind == 0u) {
      float r = ob.half_extent[0];
      float dd = dx * dx + dy * dy + dz * dz;
      float b = cx * dx + cy * dy + cz * dz;
      float c = cx * cx + cy * cy + cz * cz - r * r;
      float disc = b * b - dd * c;
      if (disc > 0.f) {
        float t = (b - sqrtf(disc)) / dd;
        if (t > 0.f && t < best) {
          best = t; hit = true;
          float ir = 1.0f / r;
          nx = (t * dx - cx) * ir; ny = (t * dy - cy) * ir; nz = (t * dz - cz) * ir;
          alb = ob.albedo_srgb8; mat = ob.material_srgb8;
        }
      }
    } else {
      float tmin = 0.f, tmax = 3.0e38f;
      int axis = 0; float sgn = 0.f;
      const float cc[3] = {cx, cy, cz}, dd3[3] = {dx, dy, dz};
      bool ok = true;
      for (int a = 0; a < 3; a++) {
        float e = ob.half_extent[a];
        if (dd3[a] != 0.f) {
          float inv = 1.0f / dd3[a];
          float t0 = (cc[a] - e) * inv, t1 = (cc[a] + e) * inv;
          float s = -1.f;
          if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; s = 1.f; }
          if (t0 > tmin) { tmin = t0; axis = a; sgn = s; }
          if (t1 < tmax) tmax = t1;
        } else if (fabsf(cc[a]) > e) ok = false;
      }
      if (ok && tmin < tmax && tmin > 0.f && tmin < best) {
        best = tmin; hit = true;
        nx = axis == 0 ? sgn : 0.f; ny = axis == 1 ? sgn : 0.f; nz = axis == 2 ? sgn : 0.f;
        alb = ob.albedo_srgb8; mat = ob.material_srgb8;
      }
    }
  }

  SynthOut out;
  out.albedo = alb; out.material = mat;
  if (!hit) {
    out.depth = 0xFFFFFFu; out.normal = enc_normal(0.f, 0.f, 1.f); out.velocity = 0u;
    return out;
  }
  // view-space z along this ray is exactly -t (direction has z = -1 in view space)
  double z = -(double)best;
  double zn = f.zn, zf = f.zf;
  double d = zf / (zf - zn) + zf * zn / (z * (zf - zn));  // gbuffer_encode.glsl:71-73
  if (d < 0.0) d = 0.0;
  if (d >= 1.0) { out.depth = 0xFFFFFFu; out.normal = enc_normal(0.f, 0.f, 1.f); out.velocity = 0u; out.albedo = sc.sky_albedo; out.material = 0xFF000000u; return out; }
  out.depth = (uint32_t)(d * 16777215.0 + 0.5);
  if (out.depth >= 0xFFFFFFu) out.depth = 0xFFFFFEu;  // keep geometry distinguishable from sky
  out.normal = enc_normal(nx, ny, nz);
  float wx = ox + best * dx, wy = oy + best * dy, wz = oz + best * dz;
  float pc[4];
  xform(f.prev_mvp, wx, wy, wz, 1.f, pc);
  float pnx = pc[0] / pc[3], pny = pc[1] / pc[3];
  out.velocity = enc_half2(0.5f * (pnx - ndx), 0.5f * (pny - ndy));  // opaque_taa.frag:45
  return out;
}

#ifndef VKR_HOST_ONLY   // device side: absent from oracle/libvkr_hostonly.so (see api.cu)
__global__ void __launch_bounds__(256) synth_kernel(const VkrSynthScene* __restrict__ scene_g, SynthFrameK f, Mip albedo, Mip normal, Mip material, Mip velocity, Mip depth) {
  __shared__ VkrSynthScene sc;
  for (int i = threadIdx.x; i < (int)(sizeof(VkrSynthScene) / 4); i += blockDim.x) ((uint32_t*)&sc)[i] = ((const uint32_t*)scene_g)[i];
  __syncthreads();
  int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= f.W || py >= f.H) return;
  SynthOut o = synth_pixel(sc, f, px, py);
  *((uint32_t*)(albedo.p + (size_t)py * albedo.pitch) + px) = o.albedo;
  *((uint32_t*)(normal.p + (size_t)py * normal.pitch) + px) = o.normal;
  *((uint32_t*)(material.p + (size_t)py * material.pitch) + px) = o.material;
  *((uint32_t*)(velocity.p + (size_t)py * velocity.pitch) + px) = o.velocity;
  *((uint32_t*)(depth.p + (size_t)py * depth.pitch) + px) = o.depth;
}

// Cube faces of a probe (stand-in for ProbeRenderer::render_cubemap, reference src/probe_renderer.cpp:78-168): face f sees
// direction s' X + t' Y - Z (Vulkan face order +x -x +y -y +z -z) through a 90-degree square camera; colour = the albedo code,
// distance = |view position| reconstructed from the D24 depth code the G-buffer producer would have written.
__global__ void __launch_bounds__(256) synth_cube_kernel(const VkrSynthScene* __restrict__ scene_g, float px, float py, float pz, int n, Mip color, Mip dist) {
  __shared__ VkrSynthScene sc;
  for (int i = threadIdx.x; i < (int)(sizeof(VkrSynthScene) / 4); i += blockDim.x) ((uint32_t*)&sc)[i] = ((const uint32_t*)scene_g)[i];
  __syncthreads();
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5), face = blockIdx.z;
  if (x >= n || y >= n) return;
  const float AX[6][9] = {{0, 0, -1, 0, -1, 0, -1, 0, 0}, {0, 0, 1, 0, -1, 0, 1, 0, 0}, {1, 0, 0, 0, 0, 1, 0, -1, 0},
                          {1, 0, 0, 0, 0, -1, 0, 1, 0},   {1, 0, 0, 0, -1, 0, 0, 0, -1}, {-1, 0, 0, 0, -1, 0, 0, 0, 1}};
  SynthFrameK f;
  for (int i = 0; i < 16; i++) { f.inv_view[i] = 0.f; f.view[i] = 0.f; f.mvp[i] = (i % 5 == 0) ? 1.f : 0.f; f.prev_mvp[i] = f.mvp[i]; }
  for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) f.inv_view[4 * c + r] = AX[face][3 * c + r];
  f.inv_view[12] = px; f.inv_view[13] = py; f.inv_view[14] = pz; f.inv_view[15] = 1.f;
  f.aspect = 1.f; f.tg = 1.f; f.zn = 0.05f; f.zf = 80.f; f.jx = f.jy = 0.f; f.W = n; f.H = n;
  const SynthOut o = synth_pixel(sc, f, x, y);
  const double d = (double)o.depth / 16777215.0, zn = 0.05, zf = 80.0;
  const double z = zn * zf / (d * (zf - zn) - zf);
  const float cx = 2.0f * ((float)x + 0.5f) / (float)n - 1.0f, cy = 2.0f * ((float)y + 0.5f) / (float)n - 1.0f;
  const float ray_len = sqrtf((1.0f + cx * cx) + cy * cy);
  const float distance = (float)(-z) * ray_len;
  *((uint32_t*)(color.p + (size_t)face * color.ls + (size_t)y * color.pitch) + x) = o.albedo;
  *((__half*)(dist.p + (size_t)face * dist.ls + (size_t)y * dist.pitch) + x) = __float2half_rn(distance);
}

#endif  // !VKR_HOST_ONLY

static int synth_check(const VkrSynthScene* scene, const VkrDrawTAAParams* p, const VkrImage* albedo, const VkrImage* normal,
                       const VkrImage* material, const VkrImage* velocity, const VkrImage* depth) {
  if (!scene || !p || scene->object_count > VKR_SYNTH_MAX_OBJECTS || !image_ok(albedo, VKR_FORMAT_R8G8B8A8_SRGB) ||
      !image_ok(normal, VKR_FORMAT_R16G16_UNORM) || !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) ||
      !image_ok(velocity, VKR_FORMAT_R16G16_SFLOAT) || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT)) {
    set_error("synth_gbuffer: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  uint32_t w = depth->width, h = depth->height;
  if (albedo->width != w || normal->width != w || material->width != w || velocity->width != w ||
      albedo->height != h || normal->height != h || material->height != h || velocity->height != h) {
    set_error("synth_gbuffer: G-buffer images have different sizes");
    return VKR_ERR_SIZE_MISMATCH;
  }
  return VKR_OK;
}

static SynthFrameK make_frame(const VkrDrawTAAParams* p, uint32_t w, uint32_t h) {
  SynthFrameK f;
  vkr_mat4_inverse(p->camera, f.inv_view);
  memcpy(f.view, p->camera, 64); memcpy(f.mvp, p->mvp, 64); memcpy(f.prev_mvp, p->prev_mvp, 64);
  f.aspect = p->fovy_aspect_znear_zfar[1];
  f.tg = std::tan(p->fovy_aspect_znear_zfar[0] / 2);
  f.zn = p->fovy_aspect_znear_zfar[2]; f.zf = p->fovy_aspect_znear_zfar[3];
  f.jx = p->jitter[0]; f.jy = p->jitter[1];
  f.W = (int)w; f.H = (int)h;
  return f;
}

}  // namespace vkr
using namespace vkr;

#ifndef VKR_HOST_ONLY
extern "C" int vkr_synth_gbuffer(const VkrSynthScene* scene, const VkrDrawTAAParams* p, const VkrImage* albedo, const VkrImage* normal,
                                 const VkrImage* material, const VkrImage* velocity, const VkrImage* depth, void* stream) {
  int rc = synth_check(scene, p, albedo, normal, material, velocity, depth);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  SynthFrameK f = make_frame(p, depth->width, depth->height);
  VkrSynthScene* dscene = nullptr;
  if (cudaMallocAsync(&dscene, sizeof(VkrSynthScene), st) != cudaSuccess) return check_launch("synth scene alloc");
  cudaMemcpyAsync(dscene, scene, sizeof(VkrSynthScene), cudaMemcpyHostToDevice, st);
  dim3 grid((f.W + 31) / 32, (f.H + 7) / 8);
  synth_kernel<<<grid, 256, 0, st>>>(dscene, f, make_mip(albedo, 0), make_mip(normal, 0), make_mip(material, 0), make_mip(velocity, 0), make_mip(depth, 0));
  rc = check_launch("synth_kernel");
  cudaFreeAsync(dscene, st);
  return rc;
}
#endif  // !VKR_HOST_ONLY

extern "C" int vkr_synth_gbuffer_host(const VkrSynthScene* scene, const VkrDrawTAAParams* p, const VkrImage* albedo, const VkrImage* normal,
                                      const VkrImage* material, const VkrImage* velocity, const VkrImage* depth, int threads) {
  int rc = synth_check(scene, p, albedo, normal, material, velocity, depth);
  if (rc) return rc;
  SynthFrameK f = make_frame(p, depth->width, depth->height);
  Mip ma = make_mip(albedo, 0), mn = make_mip(normal, 0), mm = make_mip(material, 0), mv = make_mip(velocity, 0), md = make_mip(depth, 0);
  int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  auto work = [&](int t) {
    for (int y = t; y < f.H; y += nt)
      for (int x = 0; x < f.W; x++) {
        SynthOut o = synth_pixel(*scene, f, x, y);
        *((uint32_t*)(ma.p + (size_t)y * ma.pitch) + x) = o.albedo;
        *((uint32_t*)(mn.p + (size_t)y * mn.pitch) + x) = o.normal;
        *((uint32_t*)(mm.p + (size_t)y * mm.pitch) + x) = o.material;
        *((uint32_t*)(mv.p + (size_t)y * mv.pitch) + x) = o.velocity;
        *((uint32_t*)(md.p + (size_t)y * md.pitch) + x) = o.depth;
      }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nt; t++) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  return VKR_OK;
}

#ifndef VKR_HOST_ONLY
extern "C" int vkr_synth_cubemap(const VkrSynthScene* scene, const float* position3, const VkrImage* cube_color, uint64_t color_face_stride,
                                 const VkrImage* cube_distance, uint64_t distance_face_stride, void* stream) {
  if (!scene || !position3 || scene->object_count > VKR_SYNTH_MAX_OBJECTS || !image_ok(cube_color, VKR_FORMAT_R8G8B8A8_SRGB) ||
      !image_ok(cube_distance, VKR_FORMAT_R16_SFLOAT) || !color_face_stride || !distance_face_stride) {
    set_error("synth_cubemap: null argument, unexpected format or missing face stride");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  const uint32_t n = cube_color->width;
  if (cube_color->height != n || cube_distance->width != n || cube_distance->height != n) {
    set_error("synth_cubemap: faces must be square and of one size");
    return VKR_ERR_SIZE_MISMATCH;
  }
  cudaStream_t st = (cudaStream_t)stream;
  VkrSynthScene* dscene = nullptr;
  if (cudaMallocAsync(&dscene, sizeof(VkrSynthScene), st) != cudaSuccess) return check_launch("synth scene alloc");
  cudaMemcpyAsync(dscene, scene, sizeof(VkrSynthScene), cudaMemcpyHostToDevice, st);
  Mip c = make_mip(cube_color, 0), d = make_mip(cube_distance, 0);
  c.ls = color_face_stride; d.ls = distance_face_stride;
  dim3 grid((n + 31) / 32, (n + 7) / 8, 6);
  synth_cube_kernel<<<grid, 256, 0, st>>>(dscene, position3[0], position3[1], position3[2], (int)n, c, d);
  int rc = check_launch("synth_cube_kernel");
  cudaFreeAsync(dscene, st);
  return rc;
}
#endif  // !VKR_HOST_ONLY
