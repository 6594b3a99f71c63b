// f32x2.cuh — packed single precision (two IEEE fp32 lanes in one 64-bit register pair) for sm_100a.
//
// Blackwell issues fma/mul/add.rn.f32x2 (SASS FFMA2 / FMUL2 / FADD2) as ONE warp instruction doing the work of two; each
// lane rounds exactly like the scalar .rn form, so packing changes no result bit.  The chain's kernels are bound by
// issue slots (75-85 % issue-active with the FMA pipe at ~46 %, profiles/r01b_ncu_full_4k_summary.json), which is the
// case the packed forms help: tools/ubench/f32x2.cu measures 124 of 128 lane-FMA/clk/SM for FFMA2 against 104 for scalar
// FFMA, and 1.5x when ALU instructions compete for the same issue slots.
#pragma once
#include <cuda_runtime.h>

namespace vkr {

struct f2 { unsigned long long v; };
__device__ __forceinline__ f2 pk(float lo, float hi) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ f2 dup(float x) { return pk(x, x); }
__device__ __forceinline__ void unpk(f2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v)); }
__device__ __forceinline__ float hsum(f2 a) { float lo, hi; unpk(a, lo, hi); return lo + hi; }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { f2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }


__device__ __forceinline__ f2 sub2(f2 a, f2 b) { f2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }

// UNFUSED a * b + c.  ptxas contracts mul.rn.f32x2 followed by add.rn.f32x2 into FFMA2 — unlike the scalar .rn forms,
// and with -fmad=false too (checked with cuobjdump on CUDA 12.9) — which would change the bits of the paths that must
// round like the oracle's unfused fp32.  The sum is therefore written as fma(p, one, c) with a 1.0f the compiler cannot
// fold (it is read from constant memory at run time): p * 1 is exact, so the result is RN(RN(a * b) + c), and the
// instruction count is the same as mul + add.
static __constant__ float vkr_opaque_one = 1.0f;
__device__ __forceinline__ f2 one2() { return dup(vkr_opaque_one); }
__device__ __forceinline__ f2 add2_unfused(f2 p, f2 c, f2 one) { return fma2(p, one, c); }                       // p is a product
__device__ __forceinline__ f2 mul_add2_unfused(f2 a, f2 b, f2 c, f2 one) { return fma2(mul2(a, b), one, c); }
// GLSL mix(a, b, t) = a * (1 - t) + b * t on two value pairs at once, every operation rounded like common.cuh gmix()
__device__ __forceinline__ f2 gmix2(f2 a, f2 b, float one_minus_t, float t, f2 one) { return fma2(mul2(b, dup(t)), one, mul2(a, dup(one_minus_t))); }

}  // namespace vkr
