// lut.cu — init-time LUT pre-integration.  Replaces tasks "SSR_preintegrate" and
// "BRDF_preintegrate" (reference src/advanced_ssr.cpp:95-136;
// shaders/advanced_ssr/preintegrate.comp:45-84, preintegrate_ssr.comp:12-45).
// The sums are accumulated in the reference's order (i = 0..N-1, fp32) so the tables agree
// with the oracle to rounding of the individual terms.
#include "brdf.cuh"

namespace vkr {

__device__ __forceinline__ float G2_integrand(float t, float a, float b) {   // preintegrate.comp:77-84
  const float p = b - a;
  const float q = b + a;
  const float L = p * t + q;
  const float nom = (1 - t) * L;
  const float denom = (1 + t * t - 0.5f * L * L);
  return (L > 0) ? nom / (denom * denom) : 0.0f;
}

__global__ void __launch_bounds__(256) pdf_preintegrate_kernel(Mip out) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= out.w || y >= out.h) return;
  const int STEP_COUNT = 2000;
  const float a = 2 * ((float)x + 0.5f) / (float)out.w - 1;
  const float b = ((float)y + 0.5f) / (float)out.h;
  float sum = 0.0f;
  const float dt = 2.0f / STEP_COUNT;
  for (int i = 0; i < STEP_COUNT; i++) {
    float t = -1 + dt * ((float)i + 0.5f);
    sum += G2_integrand(t, a, b);
  }
  float g = 2.0f / STEP_COUNT * sum;
  *((float*)(out.p + (size_t)y * out.pitch) + x) = g;
}

__global__ void __launch_bounds__(256) brdf_preintegrate_kernel(const float4* __restrict__ samples, Mip out) {
  __shared__ float2 s_samples[128];
  if (threadIdx.x < 128) { float4 v = samples[threadIdx.x]; s_samples[threadIdx.x] = make_float2(v.x, v.y); }
  __syncthreads();
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= out.w || y >= out.h) return;
  const float roughness = ((float)x + 0.5f) / (float)out.w;
  const float NdotV = ((float)y + 0.5f) / (float)out.h;
  const float roughness2 = roughness * roughness;
  const f3 V = mk3(sqrtf(1 - NdotV * NdotV), 0, NdotV);
  float A = 0.0f, B = 0.0f;
  for (int i = 0; i < 128; i++) {
    f3 H = sampleGGXVNDF(V, roughness2, roughness2, s_samples[i].x, s_samples[i].y);
    f3 L = norm3(reflect3(-V, H));
    float NdotL = L.z;
    float alpha = powf(1 - dot3(V, H), 5.0f);
    float G1 = brdfG1(roughness2, NdotV);
    float G2 = brdfG2(NdotV, NdotL, roughness2);
    A += G2 / G1 * (1 - alpha);
    B += G2 / G1 * alpha;
  }
  A *= 1.0f / 128;
  B *= 1.0f / 128;
  *((uint32_t*)(out.p + (size_t)y * out.pitch) + x) = pack_half2(A, B);
}

}  // namespace vkr
using namespace vkr;

extern "C" int vkr_pdf_preintegrate(const VkrImage* out_pdf, void* stream) {
  if (!image_ok(out_pdf, VKR_FORMAT_R32_SFLOAT)) { set_error("pdf_preintegrate: expected an R32_SFLOAT image"); return VKR_ERR_INVALID_ARGUMENT; }
  dim3 grid((out_pdf->width + 31) / 32, (out_pdf->height + 7) / 8);
  pdf_preintegrate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(make_mip(out_pdf, 0));
  return check_launch("pdf_preintegrate_kernel");
}

extern "C" int vkr_brdf_preintegrate(const float* halton_vec4x128, const VkrImage* out_brdf, void* stream) {
  if (!halton_vec4x128 || !image_ok(out_brdf, VKR_FORMAT_R16G16_SFLOAT)) { set_error("brdf_preintegrate: bad arguments"); return VKR_ERR_INVALID_ARGUMENT; }
  dim3 grid((out_brdf->width + 31) / 32, (out_brdf->height + 7) / 8);
  brdf_preintegrate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const float4*)halton_vec4x128, make_mip(out_brdf, 0));
  return check_launch("brdf_preintegrate_kernel");
}
