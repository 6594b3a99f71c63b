// tex_fetch.cu — is the texture pipe worth it for the streaming passes?  
//
// Measures on the box, for a 3840x2160 RGBA16F pitch-linear image (the TAA history):
//   (a) the TAA footprint pattern — five bilinear samples at uv + {(0,0), (+-1,0), (0,+-1)} texels — through
//       cudaTextureObject_t (cudaResourceTypePitch2D, linear filter, clamp) against the manual form the TAA kernel uses
//       (12 shared texels as 64-bit loads, fp16 -> fp32, weight-form bilinear): time per launch and the result difference;
//   (b) point fetches with hardware format conversion (RGBA8 sRGB -> float4, RG16 UNORM -> float2) against the exact decode
//       of common.cuh: bit mismatches over every code.
// Build + run:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tex_fetch tools/ubench/tex_fetch.cu && /tmp/tex_fetch
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

static const int W = 3840, H = 2160;

__device__ __forceinline__ float4 ld_rgba16f(const unsigned char* p, size_t pitch, int x, int y) {
  const uint2 q = __ldg((const uint2*)(p + (size_t)y * pitch) + x);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&q.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&q.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ int clampi(int v, int hi) { return min(max(v, 0), hi); }
__device__ __forceinline__ float4 wsum4(float4 a, float4 b, float4 c, float4 d, float w00, float w10, float w01, float w11) {
  return make_float4(fmaf(d.x, w11, fmaf(c.x, w01, fmaf(b.x, w10, a.x * w00))), fmaf(d.y, w11, fmaf(c.y, w01, fmaf(b.y, w10, a.y * w00))),
                     fmaf(d.z, w11, fmaf(c.z, w01, fmaf(b.z, w10, a.z * w00))), fmaf(d.w, w11, fmaf(c.w, w01, fmaf(b.w, w10, a.w * w00))));
}
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

// synthetic reprojection: a smooth sub-texel offset field, like a slowly moving camera
__device__ __forceinline__ float2 prev_uv(int px, int py) {
  const float u = ((float)px + 0.5f) / W, v = ((float)py + 0.5f) / H;
  return make_float2(u + 0.0007f + 0.0003f * v, v - 0.0004f + 0.0002f * u);
}

__global__ void __launch_bounds__(256) manual_kernel(const unsigned char* __restrict__ img, size_t pitch, float4* __restrict__ out) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= W || py >= H) return;
  const float2 p = prev_uv(px, py);
  const float x = p.x * W - 0.5f, y = p.y * H - 0.5f;
  const float fx = floorf(x), fy = floorf(y), wx = x - fx, wy = y - fy;
  const int ix = (int)fx, iy = (int)fy;
  const float w11 = wx * wy, w10 = wx - w11, w01 = wy - w11, w00 = (1.0f - wx) - w01;
  // plus-shaped union of the five 2x2 footprints: 12 texels
  float4 t[4][4];
#pragma unroll
  for (int j = 0; j < 4; j++)
#pragma unroll
    for (int i = 0; i < 4; i++)
      if (!((i == 0 || i == 3) && (j == 0 || j == 3))) t[j][i] = ld_rgba16f(img, pitch, clampi(ix - 1 + i, W - 1), clampi(iy - 1 + j, H - 1));
  float4 s = wsum4(t[1][1], t[1][2], t[2][1], t[2][2], w00, w10, w01, w11);
  s = add4(s, wsum4(t[1][2], t[1][3], t[2][2], t[2][3], w00, w10, w01, w11));
  s = add4(s, wsum4(t[1][0], t[1][1], t[2][0], t[2][1], w00, w10, w01, w11));
  s = add4(s, wsum4(t[2][1], t[2][2], t[3][1], t[3][2], w00, w10, w01, w11));
  s = add4(s, wsum4(t[0][1], t[0][2], t[1][1], t[1][2], w00, w10, w01, w11));
  out[(size_t)py * W + px] = s;
}

__global__ void __launch_bounds__(256) tex_kernel(cudaTextureObject_t tex, float4* __restrict__ out) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= W || py >= H) return;
  const float2 p = prev_uv(px, py);
  const float x = p.x * W, y = p.y * H;        // unnormalised coordinates: texel centres at i + 0.5
  float4 s = tex2D<float4>(tex, x, y);
  s = add4(s, tex2D<float4>(tex, x + 1.0f, y));
  s = add4(s, tex2D<float4>(tex, x - 1.0f, y));
  s = add4(s, tex2D<float4>(tex, x, y + 1.0f));
  s = add4(s, tex2D<float4>(tex, x, y - 1.0f));
  out[(size_t)py * W + px] = s;
}

__global__ void srgb_point_kernel(cudaTextureObject_t tex, float4* out) {   // 256 x 1 RGBA8 sRGB image, texel i = (i, i, i, i)
  const int i = threadIdx.x;
  out[i] = tex2D<float4>(tex, i + 0.5f, 0.5f);
}
__global__ void rg16_point_kernel(cudaTextureObject_t tex, float2* out) {   // 65536 texels as 256 x 256, texel k = (k, 65535 - k)
  const int k = blockIdx.x * 256 + threadIdx.x;
  out[k] = tex2D<float2>(tex, (k & 255) + 0.5f, (k >> 8) + 0.5f);
}

static float time_ms(void (*launch)(void*), void* ctx, int reps) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; i++) launch(ctx);
  CK(cudaEventRecord(a));
  for (int i = 0; i < reps; i++) launch(ctx);
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  return ms / reps;
}

struct Ctx { unsigned char* img; size_t pitch; cudaTextureObject_t tex; float4* out; };
static void launch_manual(void* c) { Ctx* x = (Ctx*)c; manual_kernel<<<dim3((W + 31) / 32, (H + 7) / 8), 256>>>(x->img, x->pitch, x->out); }
static void launch_tex(void* c) { Ctx* x = (Ctx*)c; tex_kernel<<<dim3((W + 31) / 32, (H + 7) / 8), 256>>>(x->tex, x->out); }

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("device %s, texturePitchAlignment %zu, textureAlignment %zu\n", prop.name, prop.texturePitchAlignment, prop.textureAlignment);

  // ---- (a) RGBA16F bilinear footprints --------------------------------------------------------------------------
  Ctx c;
  c.pitch = ((size_t)W * 8 + 31) / 32 * 32;
  CK(cudaMalloc(&c.img, c.pitch * H));
  CK(cudaMalloc(&c.out, sizeof(float4) * W * H));
  std::vector<__half> host((c.pitch / 2) * H);
  uint32_t s = 12345u;
  for (int y = 0; y < H; y++)
    for (int x = 0; x < W; x++)
      for (int ch = 0; ch < 4; ch++) {
        s = s * 1664525u + 1013904223u;
        // smooth image + noise + hard vertical edges every 64 px: colour-like content in [0, 1.5]
        float v = 0.5f + 0.4f * sinf(0.01f * x + 0.02f * y + ch) + 0.05f * ((s >> 8) * (1.0f / 16777216.0f)) + ((x / 64) & 1 ? 0.5f : 0.0f);
        host[(size_t)y * (c.pitch / 2) + 4 * x + ch] = __float2half(v);
      }
  CK(cudaMemcpy(c.img, host.data(), c.pitch * H, cudaMemcpyHostToDevice));
  cudaResourceDesc rd = {};
  rd.resType = cudaResourceTypePitch2D;
  rd.res.pitch2D.devPtr = c.img; rd.res.pitch2D.width = W; rd.res.pitch2D.height = H; rd.res.pitch2D.pitchInBytes = c.pitch;
  rd.res.pitch2D.desc = cudaCreateChannelDescHalf4();
  cudaTextureDesc td = {};
  td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
  CK(cudaCreateTextureObject(&c.tex, &rd, &td, nullptr));

  std::vector<float4> ref((size_t)W * H), got((size_t)W * H);
  launch_manual(&c); CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(ref.data(), c.out, sizeof(float4) * W * H, cudaMemcpyDeviceToHost));
  launch_tex(&c); CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(got.data(), c.out, sizeof(float4) * W * H, cudaMemcpyDeviceToHost));
  double maxd = 0, sumsq = 0; size_t over4 = 0, over1 = 0;
  for (size_t i = 0; i < ref.size(); i++) {
    const float d[4] = {fabsf(ref[i].x - got[i].x), fabsf(ref[i].y - got[i].y), fabsf(ref[i].z - got[i].z), fabsf(ref[i].w - got[i].w)};
    for (int k = 0; k < 4; k++) { const double e = d[k] / 5.0; maxd = fmax(maxd, e); sumsq += e * e; over4 += e > 4e-3; over1 += e > 1e-3; }   // per sample (sum of 5)
  }
  const double n = 4.0 * ref.size();
  printf("bilinear RGBA16F, hardware vs fp32 weights (per sample): max |diff| %.3e, rms %.3e, frac > 1e-3: %.3e, frac > 4e-3: %.3e\n", maxd, sqrt(sumsq / n),
         over1 / n, over4 / n);
  const float tm = time_ms(launch_manual, &c, 50), tt = time_ms(launch_tex, &c, 50);
  printf("5 footprints per pixel at %dx%d: manual (12 x LDG.64 + cvt + weight-form) %.4f ms, tex2D x5 %.4f ms  (both also write 16 B per pixel)\n", W, H, tm, tt);

  // ---- (b) point fetch + hardware format conversion ------------------------------------------------------------
  {
    unsigned char* d8; size_t pitch8 = 1024;
    CK(cudaMalloc(&d8, pitch8));
    std::vector<unsigned char> h8(1024);
    for (int i = 0; i < 256; i++) for (int ch = 0; ch < 4; ch++) h8[4 * i + ch] = (unsigned char)i;
    CK(cudaMemcpy(d8, h8.data(), 1024, cudaMemcpyHostToDevice));
    cudaResourceDesc r8 = {};
    r8.resType = cudaResourceTypePitch2D;
    r8.res.pitch2D.devPtr = d8; r8.res.pitch2D.width = 256; r8.res.pitch2D.height = 1; r8.res.pitch2D.pitchInBytes = pitch8;
    r8.res.pitch2D.desc = cudaCreateChannelDesc<uchar4>();
    cudaTextureDesc t8 = {};
    t8.addressMode[0] = t8.addressMode[1] = cudaAddressModeClamp;
    t8.filterMode = cudaFilterModePoint; t8.readMode = cudaReadModeNormalizedFloat; t8.sRGB = 1;
    cudaTextureObject_t tx8;
    CK(cudaCreateTextureObject(&tx8, &r8, &t8, nullptr));
    float4* o8; CK(cudaMalloc(&o8, 256 * sizeof(float4)));
    srgb_point_kernel<<<1, 256>>>(tx8, o8);
    std::vector<float4> r(256);
    CK(cudaMemcpy(r.data(), o8, 256 * sizeof(float4), cudaMemcpyDeviceToHost));
    int mism_rgb = 0, mism_a = 0; double maxrel = 0;
    for (int i = 0; i < 256; i++) {
      const float cc = i / 255.0f;
      const float lin = cc <= 0.04045f ? cc / 12.92f : powf((cc + 0.055f) / 1.055f, 2.4f);     // the oracle's decode table entry
      mism_rgb += r[i].x != lin; mism_a += r[i].w != cc;
      if (lin > 0) maxrel = fmax(maxrel, fabs(r[i].x - lin) / lin);
    }
    printf("RGBA8 sRGB point fetch: %d of 256 rgb codes differ from the fp32 decode (max rel %.2e), %d of 256 alpha codes differ from k/255\n", mism_rgb, maxrel, mism_a);
  }
  {
    unsigned char* d16; size_t pitch16 = 256 * 4;
    CK(cudaMalloc(&d16, pitch16 * 256));
    std::vector<uint16_t> h16(2 * 65536);
    for (int k = 0; k < 65536; k++) { h16[2 * k] = (uint16_t)k; h16[2 * k + 1] = (uint16_t)(65535 - k); }
    CK(cudaMemcpy(d16, h16.data(), pitch16 * 256, cudaMemcpyHostToDevice));
    cudaResourceDesc r16 = {};
    r16.resType = cudaResourceTypePitch2D;
    r16.res.pitch2D.devPtr = d16; r16.res.pitch2D.width = 256; r16.res.pitch2D.height = 256; r16.res.pitch2D.pitchInBytes = pitch16;
    r16.res.pitch2D.desc = cudaCreateChannelDesc<ushort2>();
    cudaTextureDesc t16 = {};
    t16.addressMode[0] = t16.addressMode[1] = cudaAddressModeClamp;
    t16.filterMode = cudaFilterModePoint; t16.readMode = cudaReadModeNormalizedFloat;
    cudaTextureObject_t tx16;
    CK(cudaCreateTextureObject(&tx16, &r16, &t16, nullptr));
    float2* o16; CK(cudaMalloc(&o16, 65536 * sizeof(float2)));
    rg16_point_kernel<<<256, 256>>>(tx16, o16);
    std::vector<float2> r(65536);
    CK(cudaMemcpy(r.data(), o16, 65536 * sizeof(float2), cudaMemcpyDeviceToHost));
    int mism = 0;
    for (int k = 0; k < 65536; k++) mism += (r[k].x != (float)k / 65535.0f) + (r[k].y != (float)(65535 - k) / 65535.0f);
    printf("RG16 UNORM point fetch: %d of 131072 values differ from the correctly rounded k/65535\n", mism);
  }
  return 0;
}
