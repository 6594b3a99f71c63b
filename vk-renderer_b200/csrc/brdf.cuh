// brdf.cuh — device restatement of reference src/shaders/include/brdf.glsl (GGX / Smith /
// Schlick / VNDF sampling / GGX-direction pdf LUT lookup), shared by the pass kernels.
#pragma once
#include "common.cuh"

namespace vkr {

__device__ __forceinline__ float DistributionGGX(f3 N, f3 H, float alpha) {            // brdf.glsl:31-38
  float NoH = dot3(N, H);
  float alpha2 = FM(alpha, alpha);
  float NoH2 = FM(NoH, NoH);
  float den = FA(FM(NoH2, alpha2), FS(1.f, NoH2));
  return FM((NoH2 > 0) ? 1.f : 0.f, alpha2) / FM(FM(VKR_PI, den), den);
}

__device__ __forceinline__ float brdfG1(float alpha2, float NdotV) {                   // brdf.glsl:43-47
  float NdotV2 = FM(NdotV, NdotV);
  float tgv2 = FS(1.f, NdotV2) / NdotV2;
  return 2.0f / FA(1.f, sqrtf(FA(1.f, FM(alpha2, tgv2))));
}

__device__ __forceinline__ float brdfG2(float NdotV, float NdotL, float alpha2) {      // brdf.glsl:49-56
  float NdotV2 = FM(NdotV, NdotV);
  float NdotL2 = FM(NdotL, NdotL);
  float L1 = sqrtf(FA(1.f, FM(alpha2, FS(1.f, NdotV2)) / NdotV2));
  float L2 = sqrtf(FA(1.f, FM(alpha2, FS(1.f, NdotL2)) / NdotL2));
  return 2.0f / FA(L1, L2);
}

// brdf.glsl:107-129 — pdf of the GGX reflection direction projected into the slice (LUT lookup)
__device__ __forceinline__ float sampleGGXdirPDF(const Mip& pdf_tex, f3 V, f3 N, f3 L, float alpha) {
  f3 Y = norm3(cross3(V, N));
  f3 X = norm3(cross3(Y, V));
  alpha = gclamp(alpha, 0.f, 0.9f);
  f3 Lproj = norm3(L - V * dot3(V, L));
  float cos_theta = dot3(X, Lproj);
  const float cos_phin = dot3(N, X);
  const float sin_phin = sqrtf(FS(1.f, FM(cos_phin, cos_phin)));
  const float alpha2 = FM(alpha, alpha);
  const float coef = sqrtf(FS(1.f, alpha2));
  const float a = FA(FM(FM(FM(0.5f, coef), cos_phin), cos_theta), 0.5f);
  const float b = FM(coef, sin_phin);
  return FM(alpha2 / FM(2 * VKR_PI, coef), sample_r32f(pdf_tex, a, b));
}

// brdf.glsl:135-155 — Heitz 2018 visible-normal sampling
__device__ __forceinline__ f3 sampleGGXVNDF(f3 Ve, float alpha_x, float alpha_y, float U1, float U2) {
  f3 Vh = norm3(mk3(FM(alpha_x, Ve.x), FM(alpha_y, Ve.y), Ve.z));
  float lensq = FA(FM(Vh.x, Vh.x), FM(Vh.y, Vh.y));
  f3 T1 = lensq > 0 ? mk3(-Vh.y, Vh.x, 0) * (1.0f / sqrtf(lensq)) : mk3(1, 0, 0);
  f3 T2 = cross3(Vh, T1);
  float r = sqrtf(U1);
  float phi = FM(2.0f * VKR_PI, U2);
  float t1 = FM(r, cosf(phi));
  float t2 = FM(r, sinf(phi));
  float s = FM(0.5f, FA(1.0f, Vh.z));
  t2 = FA(FM(FS(1.0f, s), sqrtf(FS(1.0f, FM(t1, t1)))), FM(s, t2));
  f3 Nh = t1 * T1 + t2 * T2 + sqrtf(gmax(0.0f, FS(FS(1.0f, FM(t1, t1)), FM(t2, t2)))) * Vh;
  return norm3(mk3(FM(alpha_x, Nh.x), FM(alpha_y, Nh.y), gmax(0.0f, Nh.z)));
}


// value-only restatement of sampleGGXVNDF (FMA contraction, single-MUFU rsqrt / sqrt; sin / cos stay the accurate forms:
// the direction they produce steers an up to 80-cell march)
__device__ __forceinline__ f3 sampleGGXVNDF_fast(f3 Ve, float alpha_x, float alpha_y, float U1, float U2) {
  f3 Vh = norm3f(mk3(alpha_x * Ve.x, alpha_y * Ve.y, Ve.z));
  float lensq = fmaf(Vh.x, Vh.x, Vh.y * Vh.y);
  float il = rsqrt_fast(lensq);
  f3 T1 = lensq > 0 ? mk3(-Vh.y * il, Vh.x * il, 0) : mk3(1, 0, 0);
  f3 T2 = mk3(Vh.y * T1.z - T1.y * Vh.z, Vh.z * T1.x - T1.z * Vh.x, Vh.x * T1.y - T1.x * Vh.y);
  float r = sqrt_fast(U1);
  float sp, cp;
  sincosf(FM(2.0f * VKR_PI, U2), &sp, &cp);
  float t1 = r * cp, t2 = r * sp;
  float s = 0.5f * (1.0f + Vh.z);
  t2 = fmaf(1.0f - s, sqrt_fast(fmaf(-t1, t1, 1.0f)), s * t2);
  float nz = sqrt_fast(fmaxf(0.0f, fmaf(-t2, t2, fmaf(-t1, t1, 1.0f))));
  f3 Nh = mk3(fmaf(t1, T1.x, fmaf(t2, T2.x, nz * Vh.x)), fmaf(t1, T1.y, fmaf(t2, T2.y, nz * Vh.y)), fmaf(t1, T1.z, fmaf(t2, T2.z, nz * Vh.z)));
  return norm3f(mk3(alpha_x * Nh.x, alpha_y * Nh.y, fmaxf(0.0f, Nh.z)));
}

// trace.comp:143-154 == gtao/main.comp:69-80
__device__ __forceinline__ f3 get_tangent(f3 n) {
  float max_xy = gmax(fabsf(n.x), fabsf(n.y));
  f3 t = (max_xy < 0.00001f) ? mk3(1, 0, 0) : mk3(n.y, -n.x, 0);
  return norm3(t);
}

}  // namespace vkr
