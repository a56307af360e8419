// Measurement helpers for bench.py and DESIGN.md (not part of the product library): built as
// tools/microbench/libmlapo_microbench.so by __graft_entry__.build(), called through ctypes.
//
//   mlapo_mb_i8_peak      dense int8 tensor-core peak of this GPU: every SM issues tcgen05.mma.kind::i8 (M128 N256 K32, int32
//                         accumulators in TMEM) on operand tiles that stay resident in shared memory, so nothing but the tensor
//                         pipe is exercised.  SURVEY.md §8(d) asks for this figure as the denominator of the prefill roofline.
//   mlapo_mb_late_copy    a stand-in for a producer kernel that uses programmatic dependent launch the aggressive way: it releases
//                         its dependents first and only then, after a delay, writes its output (tests/test_back_to_back_gpu.py
//                         puts it in front of mla_preprocess to check that the operator waits for its inputs by default).
//   mlapo_mb_dsmem        how fast a CTA can move a split-K partial tile into a peer CTA's shared memory inside a cluster of 8:
//                         (0) st.shared::cluster 16-byte stores, (1) cp.async.bulk shared::cta -> shared::cluster,
//                         (2) cp.reduce.async.bulk ... .add.s32 (the reduction done by the copy engine).  Decides whether MM1's
//                         split-K sums are better reduced on chip than through L2.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../sgl-kernel-npu_b200/csrc/sm100_ptx.cuh"

using namespace mlapo;

namespace {

constexpr int kATile = 128 * 128;   // [128 rows x 128 B] swizzled
constexpr int kBTile = 256 * 128;   // [256 rows x 128 B]

__global__ void __launch_bounds__(128, 1) i8_peak_kernel(int iters, unsigned seed) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  // operand bytes: any values do (the layout is what the descriptors describe); pseudo-random so that the data path toggles
  uint32_t x = seed * 2654435761u + threadIdx.x * 40503u + blockIdx.x * 9973u + 1u;
  for (int i = threadIdx.x; i < (kATile + kBTile) / 4; i += blockDim.x) {
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    reinterpret_cast<uint32_t*>(smem)[i] = x;
  }
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bar), 1);
    ptx::fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, 128, 256);
    const uint32_t a_base = ptx::smem_u32(smem), b_base = a_base + kATile;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = ptx::make_smem_desc(a_base + k * 32, 1024, 0);
        const uint64_t db = ptx::make_smem_desc(b_base + k * 32, 1024, 0);
        ptx::umma_i8(tmem + (it & 1) * 256, da, db, idesc, (it > 1 || k > 0) ? 1u : 0u);
      }
    }
    ptx::umma_commit(ptx::smem_u32(&bar));
    ptx::mbar_wait(ptx::smem_u32(&bar), 0);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------- MMA shapes of the op on CTA pairs
// The op's own instruction shapes, on resident tiles, every pair of SMs: M = 256 (cta_group::2), N and kind as given, K = 32
// bytes per instruction (int8: 32 elements, bf16: 16), K-major A [128 rows per CTA] and B [N/2 rows per CTA].  `chains` = how
// many accumulators the instruction stream alternates between (1: every instruction depends on the previous one's D).
template <bool kF16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) pair_shape_kernel(int iters, int n, int chains, unsigned seed) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint32_t x = seed * 2654435761u + threadIdx.x * 40503u + blockIdx.x * 9973u + 1u;
  for (int i = threadIdx.x; i < (kATile + kBTile / 2) / 4; i += blockDim.x) {
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    // bf16 operands: keep the exponent field away from inf / nan patterns (upper bits 0x3F80 | small mantissa)
    reinterpret_cast<uint32_t*>(smem)[i] = kF16 ? ((x & 0x007F007Fu) | 0x3F803F80u) : x;
  }
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bar), 1);
    ptx::fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    ptx::tmem_alloc_pair(ptx::smem_u32(&tmem_slot), 512);
    ptx::tmem_relinquish_pair();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_arrive();
  ptx::cluster_wait();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (ptx::cluster_ctarank() == 0 && threadIdx.x == 0) {
    const uint32_t idesc = kF16 ? ptx::make_idesc(ptx::kCFmtF32, ptx::kFmtBF16, ptx::kFmtBF16, 0, 0, 256, (uint32_t)n)
                                : ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, 256, (uint32_t)n);
    const uint32_t a_base = ptx::smem_u32(smem), b_base = a_base + kATile;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = ptx::make_smem_desc(a_base + k * 32, 1024, 0);
        const uint64_t db = ptx::make_smem_desc(b_base + k * 32, 1024, 0);
        const uint32_t d = tmem + (uint32_t)((it * 4 + k) % chains) * 256u;
        if (kF16) ptx::umma_f16_pair(d, da, db, idesc, it > 0 ? 1u : 0u);
        else ptx::umma_i8_pair(d, da, db, idesc, it > 0 ? 1u : 0u);
      }
    }
    ptx::umma_commit_pair(ptx::smem_u32(&bar));
  }
  if (threadIdx.x == 0) ptx::mbar_wait(ptx::smem_u32(&bar), 0);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_arrive();
  ptx::cluster_wait();
  if (threadIdx.x < 32) ptx::tmem_dealloc_pair(tmem, 512);
}

// The same on single CTAs (cta_group::1, M = 128).
template <bool kF16>
__global__ void __launch_bounds__(128, 1) single_shape_kernel(int iters, int m, int n, int chains, unsigned seed) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint32_t x = seed * 2654435761u + threadIdx.x * 40503u + blockIdx.x * 9973u + 1u;
  for (int i = threadIdx.x; i < (kATile + kBTile) / 4; i += blockDim.x) {
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    reinterpret_cast<uint32_t*>(smem)[i] = kF16 ? ((x & 0x007F007Fu) | 0x3F803F80u) : x;
  }
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bar), chains == 3 ? 2 : 1);
    ptx::fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  // chains == 4: the whole of warp 0 runs the issue loop (no divergent branch around it); the asm block elects the issuing lane
  if (chains == 4) {
    if (threadIdx.x < 32) {
      const uint32_t idesc = kF16 ? ptx::make_idesc(ptx::kCFmtF32, ptx::kFmtBF16, ptx::kFmtBF16, 0, 0, (uint32_t)m, (uint32_t)n)
                                  : ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, (uint32_t)m, (uint32_t)n);
      const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(smem), 1024, 0), db0 = ptx::make_smem_desc(ptx::smem_u32(smem) + kATile, 1024, 0);
      for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t da = da0 + 2 * k, db = db0 + 2 * k;   // +32 bytes = +2 in the descriptor's 16-byte address units
          const uint32_t acc = it > 0 ? 1u : 0u;
          if (kF16)
            asm volatile("{\n\t.reg .pred p, q;\n\telect.sync _|q, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
          else
            asm volatile("{\n\t.reg .pred p, q;\n\telect.sync _|q, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "@q tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
        }
      }
      if (ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&bar));
      __syncwarp();
    }
    if (threadIdx.x == 0) ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) ptx::tmem_dealloc(tmem, 512);
    return;
  }
  // chains == 3: TWO issuing threads (warps 0 and 1), each with its own accumulator and half of the instructions
  const bool two = chains == 3;
  if (threadIdx.x == 0 || (two && threadIdx.x == 32)) {
    const int who = threadIdx.x >> 5;
    const uint32_t idesc = kF16 ? ptx::make_idesc(ptx::kCFmtF32, ptx::kFmtBF16, ptx::kFmtBF16, 0, 0, (uint32_t)m, (uint32_t)n)
                                : ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, (uint32_t)m, (uint32_t)n);
    const uint32_t a_base = ptx::smem_u32(smem), b_base = a_base + kATile;
    const int my_iters = two ? iters / 2 : iters;
    for (int it = 0; it < my_iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = ptx::make_smem_desc(a_base + k * 32, 1024, 0);
        const uint64_t db = ptx::make_smem_desc(b_base + k * 32, 1024, 0);
        const uint32_t d = tmem + (two ? (uint32_t)who * 256u : (uint32_t)((it * 4 + k) % chains) * 256u);
        if (kF16) ptx::umma_f16(d, da, db, idesc, it > 0 ? 1u : 0u);
        else ptx::umma_i8(d, da, db, idesc, it > 0 ? 1u : 0u);
      }
    }
    ptx::umma_commit(ptx::smem_u32(&bar));
  }
  if (threadIdx.x == 0) ptx::mbar_wait(ptx::smem_u32(&bar), 0);
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------- DSMEM
constexpr int kDsChunk = 16384;        // bytes one CTA sends to one peer per round (a [32 x 128] int32 sub-tile)
constexpr int kDsThreads = 128;

__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// Every CTA of a cluster of 8 sends kDsChunk bytes to each of its 7 peers, `rounds` times; out[cluster] = cycles of rank 0.
__global__ void __cluster_dims__(8, 1, 1) __launch_bounds__(kDsThreads, 1) dsmem_kernel(int mode, int rounds, long long* out) {
  extern __shared__ __align__(1024) uint8_t sm[];
  uint8_t* src = sm;                       // 16 KB: what this CTA sends
  uint8_t* recv = sm + kDsChunk;           // 8 x 16 KB: slot r = what rank r sent (mode 2: everything is added into slot 0)
  __shared__ uint64_t bar;
  const uint32_t rank = ptx::cluster_ctarank();
  for (int i = threadIdx.x; i < kDsChunk / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(src)[i] = i + rank;
  for (int i = threadIdx.x; i < 8 * kDsChunk / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(recv)[i] = 0;
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bar), 1);
    ptx::fence_mbar_init();
  }
  ptx::fence_proxy_async_smem();
  __syncthreads();
  ptx::cluster_arrive();
  ptx::cluster_wait();
  const long long t0 = clock64();
  for (int r = 0; r < rounds; ++r) {
    if (mode == 0) {
      for (uint32_t p = 1; p < 8; ++p) {
        const uint32_t peer = (rank + p) & 7;
        const uint32_t dst = ptx::mapa(ptx::smem_u32(recv + rank * kDsChunk), peer);
        for (int i = threadIdx.x; i < kDsChunk / 16; i += blockDim.x) {
          const uint4 v = reinterpret_cast<const uint4*>(src)[i];
          st_cluster_v4(dst + i * 16, v.x, v.y, v.z, v.w);
        }
      }
      ptx::cluster_arrive();
      ptx::cluster_wait();
    } else {
      // the copy engine moves the chunk; its bytes are counted on the RECEIVER's barrier (7 peers x 16 KB per round)
      if (threadIdx.x == 0) {
        ptx::mbar_arrive_expect_tx(ptx::smem_u32(&bar), 7 * kDsChunk);
        for (uint32_t p = 1; p < 8; ++p) {
          const uint32_t peer = (rank + p) & 7;
          const uint32_t dst = ptx::mapa(ptx::smem_u32(recv + (mode == 2 ? 0 : rank * kDsChunk)), peer);
          const uint32_t rbar = ptx::mapa(ptx::smem_u32(&bar), peer);
          if (mode == 1)
            asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "r"(ptx::smem_u32(src)), "r"(kDsChunk), "r"(rbar) : "memory");
          else
            asm volatile("cp.reduce.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes.add.s32 [%0], [%1], %2, [%3];"
                         ::"r"(dst), "r"(ptx::smem_u32(src)), "r"(kDsChunk), "r"(rbar) : "memory");
        }
      }
      ptx::mbar_wait(ptx::smem_u32(&bar), r & 1);
      // nobody starts the next round (whose bytes would be counted on a barrier still in this round's phase) before every CTA
      // of the cluster has seen this round complete
      ptx::cluster_arrive();
      ptx::cluster_wait();
    }
  }
  const long long t1 = clock64();
  ptx::cluster_arrive();
  ptx::cluster_wait();
  if (threadIdx.x == 0 && rank == 0) out[blockIdx.x / 8] = t1 - t0;
}

__global__ void late_copy_kernel(uint4* dst, const uint4* src, long long n16, long long delay_cycles) {
  ptx::griddep_launch_dependents();   // dependents may start NOW, long before the stores below
  const long long t0 = clock64();
  while (clock64() - t0 < delay_cycles) {}
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += (long long)gridDim.x * blockDim.x) dst[i] = src[i];
}

}  // namespace

extern "C" {

// dst[0 .. n16) <- src (16-byte units) by a kernel that triggers its dependents at once and stores after `delay_cycles`.
int mlapo_mb_late_copy(void* dst, const void* src, long long n16, long long delay_cycles, void* stream) {
  late_copy_kernel<<<64, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<uint4*>(dst), static_cast<const uint4*>(src), n16, delay_cycles);
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// Runs the resident-tile MMA loop for about `target_ms` on every SM and returns the achieved dense int8 rate in TOP/s
// (2 * M * N * K per MMA); *ms_out = the measured duration.  Negative on CUDA errors.
double mlapo_mb_i8_peak(double target_ms, double* ms_out) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int smem = kATile + kBTile + 1024;
  if (cudaFuncSetAttribute(i8_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1.0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto run = [&](int iters) -> float {
    cudaEventRecord(e0);
    i8_peak_kernel<<<sms, 128, smem>>>(iters, 12345u);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) return -1.f;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
  };
  if (run(2000) < 0.f) return -2.0;                 // warm-up
  const float probe = run(20000);                   // ~5 ms at nominal rate
  if (probe <= 0.f) return -3.0;
  long long iters = (long long)(20000.0 * target_ms / probe);
  if (iters < 2000) iters = 2000;
  if (iters > 2000000000LL) iters = 2000000000LL;
  const float ms = run((int)iters);
  if (ms <= 0.f) return -4.0;
  if (ms_out) *ms_out = ms;
  const double ops = 2.0 * 128 * 256 * 32 * 4.0 * (double)iters * sms;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return ops / (ms * 1e-3) / 1e12;
}

// Cycles per instruction of the op's MMA shapes on CTA pairs (M = 256): kind 0 = int8 (K = 32), 1 = bf16 (K = 16); n = MMA N;
// chains = accumulators the stream alternates between.  Returns the achieved fraction of the nominal dense peak of that kind
// (4.5 POPS int8, 2.25 PFLOP/s bf16, at the SM clock measured during the run) and writes the average SM cycles per instruction.
double mlapo_mb_pair_shape(int kind, int n, int chains, double* cycles_per_mma);
// cta_group::1, M = 64 or 128 (same conventions as mlapo_mb_pair_shape)
double mlapo_mb_single_shape_m(int kind, int m, int n, int chains, double* cycles_per_mma);
double mlapo_mb_single_shape(int kind, int n, int chains, double* cycles_per_mma) { return mlapo_mb_single_shape_m(kind, 128, n, chains, cycles_per_mma); }
double mlapo_mb_single_shape_m(int kind, int m, int n, int chains, double* cycles_per_mma) {
  int dev = 0, sms = 0, khz = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  if (n < 32 || n > 256 || (n % 32) || chains < 1 || chains > 4 || (m != 64 && m != 128)) return -1.0;
  const int smem = kATile + kBTile;
  auto kern = kind ? single_shape_kernel<true> : single_shape_kernel<false>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -2.0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto run = [&](int iters) -> float {
    cudaEventRecord(e0);
    kern<<<sms, 128, smem>>>(iters, m, n, chains, 777u);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) return -1.f;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
  };
  if (run(2000) < 0.f) return -3.0;
  const int iters = 40000;
  const float ms = run(iters);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (ms <= 0.f) return -4.0;
  const double mmas = 4.0 * iters;
  if (cycles_per_mma) *cycles_per_mma = ms * 1e-3 * (khz * 1e3) / mmas;
  const double ops = 2.0 * m * n * (kind ? 16 : 32) * mmas * sms;
  return ops / (ms * 1e-3) / (kind ? 2.25e15 : 4.5e15);
}

double mlapo_mb_pair_shape(int kind, int n, int chains, double* cycles_per_mma) {
  int dev = 0, sms = 0, khz = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  if (n < 32 || n > 256 || (n % 32) || chains < 1 || chains > 2) return -1.0;
  const int smem = kATile + kBTile / 2;
  auto kern = kind ? pair_shape_kernel<true> : pair_shape_kernel<false>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -2.0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto run = [&](int iters) -> float {
    cudaEventRecord(e0);
    kern<<<sms & ~1, 128, smem>>>(iters, n, chains, 777u);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) return -1.f;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
  };
  if (run(2000) < 0.f) return -3.0;
  const int iters = 40000;
  const float ms = run(iters);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (ms <= 0.f) return -4.0;
  const double mmas = 4.0 * iters;                                  // per pair
  const double clk_hz = khz * 1e3;                                  // nominal max SM clock: short run, no power cap
  if (cycles_per_mma) *cycles_per_mma = ms * 1e-3 * clk_hz / mmas;
  const double ops = 2.0 * 256 * n * (kind ? 16 : 32) * mmas * ((sms & ~1) / 2);
  const double peak = kind ? 2.25e15 : 4.5e15;
  return ops / (ms * 1e-3) / peak;
}

// mode 0/1/2 as in the header comment; returns bytes per clock that ONE CTA sent (7 peers x 16 KB x rounds / cycles), median-ish
// (cluster 0); negative on errors.
double mlapo_mb_dsmem(int mode, int rounds) {
  const int smem = 9 * kDsChunk;
  if (cudaFuncSetAttribute(dsmem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1.0;
  long long* out = nullptr;
  if (cudaMalloc(&out, 64 * sizeof(long long)) != cudaSuccess) return -2.0;
  const int clusters = 16;
  dsmem_kernel<<<clusters * 8, kDsThreads, smem>>>(mode, 2, out);   // warm-up
  dsmem_kernel<<<clusters * 8, kDsThreads, smem>>>(mode, rounds, out);
  long long h[64];
  if (cudaMemcpy(h, out, clusters * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess) {
    fprintf(stderr, "dsmem microbench: %s\n", cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
    return -3.0;
  }
  cudaFree(out);
  long long worst = 0;
  for (int i = 0; i < clusters; ++i) worst = h[i] > worst ? h[i] : worst;
  return worst > 0 ? 7.0 * kDsChunk * rounds / (double)worst : -4.0;
}

}  // extern "C"
