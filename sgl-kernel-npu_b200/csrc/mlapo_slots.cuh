// Slot producers: the integer ops that create and book-keep the `slot_mapping` mla_preprocess scatters by.
// Reference: csrc/alloc_extend/op_kernel/alloc_extend_kernel.cpp, csrc/cache_location_assign/op_kernel/cache_loc_assign_kernel.cpp,
// csrc/assign_cache_op/op_kernel/assign_cache_op.cpp; goldens in tests/python/sgl_kernel_npu/test_alloc_extend_slot.py:9-85,
// test_cache_assign.py:42-61, test_inplace_assign_cache.py:27-46.  All bit-exact integer work, one CTA per request; the exclusive
// prefix sums over the requests before it are recomputed by every CTA (a few thousand additions) so that no scratch is needed.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mlapo {

// sum over r < upto of f(r), by the whole CTA (blockDim.x a multiple of 32, <= 1024); every thread gets the result
template <class F>
__device__ __forceinline__ long long block_prefix(int upto, F f) {
  __shared__ long long s_part[32];
  long long v = 0;
  for (int r = threadIdx.x; r < upto; r += blockDim.x) v += f(r);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();  // s_part may still be read from a previous call
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
  __syncthreads();
  long long tot = 0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += s_part[w];
  return tot;
}

// alloc_extend: request i extends from pre_lens[i] to seq_lens[i] tokens.  Token at position p gets
//   p <  ceil(pre/ps)*ps : last_loc[i] + 1 + (p - pre)                         (the rest of the page the prefix ends in)
//   else                 : free_pages[P_i + p/ps - ceil(pre/ps)] * ps + p % ps  (new pages, taken in order)
// with P_i = new pages of the requests before i; out_indices is packed in request order; values[0] = total new pages.
template <class I, class O>
__global__ void alloc_extend_kernel(const I* __restrict__ pre_lens, const I* __restrict__ seq_lens, const I* __restrict__ last_loc,
                                    const I* __restrict__ free_pages, long long page_size, O* __restrict__ out_indices,
                                    I* __restrict__ values, int bs, long long out_capacity) {
  const int i = blockIdx.x;
  auto pages_of = [&](int r) -> long long {
    const long long pre = pre_lens[r], seq = seq_lens[r];
    return seq > pre ? (seq + page_size - 1) / page_size - (pre + page_size - 1) / page_size : 0;
  };
  const long long tok_off = block_prefix(i, [&](int r) { const long long e = (long long)seq_lens[r] - (long long)pre_lens[r]; return e > 0 ? e : 0; });
  const long long page_off = block_prefix(i, pages_of);
  const long long pre = pre_lens[i], seq = seq_lens[i];
  const long long first_new = (pre + page_size - 1) / page_size;  // index of the first page this request has to take
  for (long long p = pre + threadIdx.x; p < seq; p += blockDim.x) {
    const long long o = tok_off + (p - pre);
    if (o >= out_capacity) break;
    long long slot;
    if (p < first_new * page_size) slot = (long long)last_loc[i] + 1 + (p - pre);
    else slot = (long long)free_pages[page_off + p / page_size - first_new] * page_size + p % page_size;
    out_indices[o] = (O)slot;
  }
  if (i == bs - 1 && threadIdx.x == 0 && values) values[0] = (I)(page_off + pages_of(i));
}

// cache_loc_assign (update == 0): token_pool[req[i], start_i + j] = cache_loc[C_i + j];  cache_loc_update (update == 1): the
// reverse copy.  C_i = sum of (end - start) over the requests before i.
template <class R>
__global__ void cache_loc_kernel(const R* __restrict__ req, int32_t* __restrict__ token_pool, const long long* __restrict__ start,
                                 const long long* __restrict__ end, int32_t* __restrict__ cache_loc, long long row_size, int update) {
  const int i = blockIdx.x;
  const long long off = block_prefix(i, [&](int r) { return end[r] - start[r]; });
  int32_t* row = token_pool + (long long)req[i] * row_size + start[i];
  const long long n = end[i] - start[i];
  for (long long j = threadIdx.x; j < n; j += blockDim.x) {
    if (update) cache_loc[off + j] = row[j];
    else row[j] = cache_loc[off + j];
  }
}

// assign_cache_op: out[i, dst_start_i : dst_end_i] = src[src_start_i : src_end_i], elements of `eb` bytes.
template <class X>
__global__ void assign_cache_kernel(uint8_t* __restrict__ out, const uint8_t* __restrict__ src, const X* __restrict__ dst_start,
                                    const X* __restrict__ dst_end, const X* __restrict__ src_start, const X* __restrict__ src_end,
                                    long long row_elems, int eb) {
  const int i = blockIdx.x;
  long long n = (long long)dst_end[i] - (long long)dst_start[i];
  const long long ns = (long long)src_end[i] - (long long)src_start[i];
  if (ns < n) n = ns;
  uint8_t* d = out + ((long long)i * row_elems + (long long)dst_start[i]) * eb;
  const uint8_t* s = src + (long long)src_start[i] * eb;
  for (long long b = threadIdx.x; b < n * eb; b += blockDim.x) d[b] = s[b];
}

}  // namespace mlapo
