// Packed fp32x2 arithmetic of sm_100 (PTX add/sub/mul/fma.rn.f32x2 -> SASS FADD2 / FMUL2 / FFMA2): two IEEE
// round-to-nearest fp32 operations per instruction, bit-identical per lane to the scalar __f*_rn forms.
// The interpolation / tangent / scatter-weight math of the field kernels is 4 channels per lane (one
// float4 of a texel), i.e. two packed pairs: half the FMA-pipe instructions of the scalar code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nfi_f2 {

typedef unsigned long long f2;      // two fp32 in one 64-bit register pair (first element in the low half)

__device__ __forceinline__ f2 pack(float a, float b) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack(f2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f2 bcast(float a) { return pack(a, a); }      // ptxas folds this into a .F32 broadcast operand
__device__ __forceinline__ f2 fma(f2 a, f2 b, f2 c) { f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f2 add(f2 a, f2 b) { f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2 sub(f2 a, f2 b) { f2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2 mul(f2 a, f2 b) { f2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

struct f4 {                           // a float4 as two packed pairs: (x, y) and (z, w)
  f2 lo, hi;
};
__device__ __forceinline__ f4 from4(const float4& v) { f4 r; r.lo = pack(v.x, v.y); r.hi = pack(v.z, v.w); return r; }
__device__ __forceinline__ float4 to4(const f4& v) { float4 r; unpack(v.lo, r.x, r.y); unpack(v.hi, r.z, r.w); return r; }
__device__ __forceinline__ f4 zero4() { f4 r; r.lo = 0ull; r.hi = 0ull; return r; }
__device__ __forceinline__ f4 add(const f4& a, const f4& b) { f4 r; r.lo = add(a.lo, b.lo); r.hi = add(a.hi, b.hi); return r; }
__device__ __forceinline__ f4 sub(const f4& a, const f4& b) { f4 r; r.lo = sub(a.lo, b.lo); r.hi = sub(a.hi, b.hi); return r; }
__device__ __forceinline__ f4 mul(const f4& a, float s) { const f2 ss = bcast(s); f4 r; r.lo = mul(a.lo, ss); r.hi = mul(a.hi, ss); return r; }
// a * s + c
__device__ __forceinline__ f4 fma(const f4& a, float s, const f4& c) { const f2 ss = bcast(s); f4 r; r.lo = fma(a.lo, ss, c.lo); r.hi = fma(a.hi, ss, c.hi); return r; }

}  // namespace nfi_f2
