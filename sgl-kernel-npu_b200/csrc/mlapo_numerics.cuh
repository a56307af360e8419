// Numeric contract of mla_preprocess (activation types, the fp32->fp16->int8 saturating quantiser, vector pack/unpack) and the debug timeline marks.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/mla_preprocess_b200.h"
#include "sm100_ptx.cuh"

namespace mlapo {

// ============================================================================================ numerics
template <class T> struct Act;
template <> struct Act<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float f) { return __float2bfloat16_rn(f); }
  static __device__ __forceinline__ float lo(uint32_t w) { return __uint_as_float(w << 16); }
  static __device__ __forceinline__ float hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  static __device__ __forceinline__ float round(float f) { return __bfloat162float(__float2bfloat16_rn(f)); }
};
template <> struct Act<__half> {
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float f) { return __float2half_rn(f); }
  static __device__ __forceinline__ float lo(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w & 0xFFFF))); }
  static __device__ __forceinline__ float hi(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w >> 16))); }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  static __device__ __forceinline__ float round(float f) { return __half2float(__float2half_rn(f)); }
};

// fp32 -> fp16 (RNE) -> clamp [-128,127] -> round-half-even -> int8   (golden: test_mla:81-83)
// Written without the quarter-rate scalar conversions: the packed F2FP does the RNE rounding to fp16, and adding 1.5 * 2^23
// rounds the clamped value (|h| <= 128) to the nearest-even integer exactly, leaving it in the low mantissa bits.
// Returns a word whose LOW BYTE is the int8 result (two's complement); the upper bytes are the magic constant's.
__device__ __forceinline__ int sat_i8(float f) {
  const __half2 h2 = __floats2half2_rn(f, f);
  float h = __low2float(h2);
  h = fminf(fmaxf(h, -128.f), 127.f);
  return __float_as_int(__fadd_rn(h, 12582912.f));
}
// low bytes of four words -> one word (three byte permutes)
__device__ __forceinline__ uint32_t pack4_i8(int a, int b, int c, int d) {
  return __byte_perm(__byte_perm((uint32_t)a, (uint32_t)b, 0x0040), __byte_perm((uint32_t)c, (uint32_t)d, 0x0040), 0x5410);
}
template <class T>
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = Act<T>::lo(v.x); f[1] = Act<T>::hi(v.x); f[2] = Act<T>::lo(v.y); f[3] = Act<T>::hi(v.y);
  f[4] = Act<T>::lo(v.z); f[5] = Act<T>::hi(v.z); f[6] = Act<T>::lo(v.w); f[7] = Act<T>::hi(v.w);
}
__device__ __forceinline__ float load_descale(const void* p, int is_i64, int i) {
  if (is_i64) return __int_as_float((int)(reinterpret_cast<const long long*>(p)[i] & 0xFFFFFFFFll));
  return reinterpret_cast<const float*>(p)[i];
}

// Debug timeline (flag bit3): per-CTA clock64 marks, 16 slots per CTA, written only when the pointer is non-null — and only in
// the instrumented build (-DMLAPO_INSTRUMENTED, libmla_preprocess_b200_instr.so): the ~40 predicated marks of the product
// kernels cost 3-4 % of a decode call even with the pointer null (C2 33.44 -> 32.43 us, C1 17.63 -> 17.03 us, C5 24.2 -> 23.25 us
// without them, same box), so the shipped library compiles them out.
#ifdef MLAPO_INSTRUMENTED
#define MLAPO_TL_ON true
#else
#define MLAPO_TL_ON false
#endif
__device__ __forceinline__ void tl_mark(long long* tl, int slot) {
  if (MLAPO_TL_ON && tl) {
    const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    tl[cta * 16 + slot] = clock64();
  }
}
// Fine-grained trace of ONE thread of CTA 0 (rows 1024.. of the kernel's timeline table): entry idx <- clock64
__device__ __forceinline__ void tl_trace(long long* tl, bool who, int idx) {
  if (MLAPO_TL_ON && tl && who && blockIdx.x == 0 && idx < 2048) tl[1024 * 16 + idx] = clock64();
}
__device__ __forceinline__ void tl_mark_global(long long* tl, int slot) {
  if (MLAPO_TL_ON && tl) {
    const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    tl[cta * 16 + slot] = (long long)t;
  }
}


}  // namespace mlapo
