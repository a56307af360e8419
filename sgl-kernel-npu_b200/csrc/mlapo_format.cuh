// Load-time format tools around mla_preprocess: the formats the operator consumes or produces, as device
// kernels, so that Ascend-format checkpoints and caches can be prepared and inspected without leaving the GPU.
// The reference has these only as test helpers (tests/python/sgl_kernel_npu/test_mla_preprocess.py:38-54, 110-135).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mlapo {

// ND [rows, cols] -> FRACTAL_NZ [1, cols_pad/c0, rows_pad, c0] (rows padded to 16, cols to c0, zero fill); `eb` bytes per element.
// Element (r, c) lands at ((c / c0) * rows_pad + r) * c0 + c % c0   (transdata(), test_mla:38-54).
__global__ void transdata_nd_to_nz_kernel(const uint8_t* __restrict__ nd, uint8_t* __restrict__ nz, long long rows, long long cols,
                                          long long rows_pad, long long cols_pad, int c0, int eb) {
  const long long total = rows_pad * cols_pad;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long within = i % c0, r = (i / c0) % rows_pad, cb = i / (c0 * rows_pad);
    const long long c = cb * c0 + within;
    const bool in = r < rows && c < cols;
    for (int b = 0; b < eb; ++b) nz[i * eb + b] = in ? nd[(r * cols + c) * eb + b] : (uint8_t)0;
  }
}

// fp16-path descale carrier: int64 = (fp32 bits & 0xFFFFE000) | 1 << 46   (trans_descale_param(), test_mla:110-119)
__global__ void pack_descale_kernel(const uint32_t* __restrict__ f32_bits, unsigned long long* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = (unsigned long long)(f32_bits[i] & 0xFFFFE000u) | (1ull << 46);
}

// Rows of a paged NZ cache [blocks, width/c0, 128, c0] named by slot -> dense [n, width]; slot -1 or out of range -> zeros.
// Element c of slot s lives at block*128*width + (c/c0)*128*c0 + (s%128)*c0 + c%c0   (extract_from_nzcache(), test_mla:122-135;
// the writer is mla_preprocess_mix_bf16.hpp:2644-2675).
__global__ void gather_nz_rows_kernel(const uint8_t* __restrict__ cache, const int32_t* __restrict__ slots, uint8_t* __restrict__ out,
                                      long long n, int width, int c0, int eb, long long num_blocks) {
  const long long total = n * width;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / width;
    const int c = (int)(i % width);
    const long long s = slots[t];
    const bool ok = s >= 0 && s < num_blocks * 128;
    const long long src = ok ? (s / 128) * 128 * width + (long long)(c / c0) * 128 * c0 + (s % 128) * c0 + c % c0 : 0;
    for (int b = 0; b < eb; ++b) out[i * eb + b] = ok ? cache[src * eb + b] : (uint8_t)0;
  }
}

}  // namespace mlapo
