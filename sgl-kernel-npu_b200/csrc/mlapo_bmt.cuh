// batch_matmul_transpose: C[m, b, :] = A[m, b, :] . B[b]  for every batch (= head) b, the W_UV half of MLA
// weight absorption after attention.  A [m, batch, k], B [batch, k, n] (ND) or [batch, n/16, k, 16] (NZ), C [m, batch, n], all
// fp16 or bf16, fp32 accumulation, one rounding to the output type.
// Reference: csrc/batch_matmul_transpose/op_host/batch_matmul_transpose.cpp:46-114 (host), op_kernel/batch_matmul_transpose_kernel.cpp.
//  - bmt_tc_kernel: tcgen05 path for ND B with k % 64 == 0 and n % 64 == 0.  One CTA per (n tile of 64/128, up to 256 rows, batch):
//    TMA walks the three tensors through rank-3 maps (no transposes are materialised: the "transpose" is the stride order),
//    A tiles K-major, B tiles as they lie in memory (N contiguous = MN-major operand), 4-stage ring over K in steps of 64.
//  - bmt_generic_kernel: any shape, ND or NZ B — one output element per thread.
#pragma once
#include "mlapo_numerics.cuh"
#include "mlapo_mm1.cuh"

namespace mlapo {

constexpr int kBmtThreads = 192;
constexpr int kBmtStages = 4;
constexpr int kBmtAtom = 8192;  // [64 k x 64 n] 16-bit atom of B
template <int BN> __host__ __device__ constexpr int bmt_stage_bytes() { return 2 * kTile16K + (BN / 64) * kBmtAtom; }
template <int BN> __host__ __device__ constexpr int bmt_smem_bytes() { return kBmtStages * bmt_stage_bytes<BN>() + kTile16K + 256 + 1024; }

template <class T, int BN>
__global__ void __launch_bounds__(kBmtThreads, 1)
bmt_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
              const __grid_constant__ CUtensorMap tm_c, int m_total, int k_total) {
  constexpr int STAGE = bmt_stage_bytes<BN>();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + kBmtStages * STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + kTile16K);
  uint64_t* full_bar = bars;       // [4]
  uint64_t* empty_bar = bars + 4;  // [4]
  uint64_t* done_bar = bars + 8;   // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * 2 * kBlockM, batch = blockIdx.z;
  const int mt = (m_total - m0 > kBlockM) ? 2 : 1;
  const int nk = k_total / 64;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_a);
    ptx::prefetch_tmap(&tm_b);
    ptx::prefetch_tmap(&tm_c);
    for (int i = 0; i < 9; ++i) ptx::mbar_init(ptx::smem_u32(bars + i), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 256);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t tx = (uint32_t)(mt * kTile16K + (BN / 64) * kBmtAtom);
      for (int i = 0; i < nk; ++i) {
        const int s = i & 3;
        if (i >= kBmtStages) ptx::mbar_wait(ptx::smem_u32(empty_bar + s), ((i >> 2) & 1) ^ 1);
        const uint32_t fb = ptx::smem_u32(full_bar + s);
        ptx::mbar_arrive_expect_tx(fb, tx);
        const uint32_t base = ptx::smem_u32(smem + s * STAGE);
        for (int j = 0; j < BN / 64; ++j)  // weights first: they are the bytes that come from DRAM
          ptx::tma_load_3d(base + 2 * kTile16K + j * kBmtAtom, &tm_b, fb, n0 + j * 64, i * 64, batch, ptx::kEvictFirst);
        for (int t = 0; t < mt; ++t) ptx::tma_load_3d(base + t * kTile16K, &tm_a, fb, i * 64, m0 + t * kBlockM, batch, ptx::kEvictFirst);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t kFmt = std::is_same<T, __nv_bfloat16>::value ? ptx::kFmtBF16 : ptx::kFmtF16;
    constexpr uint32_t idesc = ptx::make_idesc(ptx::kCFmtF32, kFmt, kFmt, 0, 1, kBlockM, BN);
    for (int i = 0; i < nk; ++i) {
      const int s = i & 3;
      ptx::mbar_wait(ptx::smem_u32(full_bar + s), (i >> 2) & 1);
      __syncwarp();
      ptx::tc_fence_after();
      {  // the converged warp issues (one elected lane inside the asm blocks): see sm100_ptx.cuh, umma_i8_elect
        const uint32_t a_base = ptx::smem_u32(smem + s * STAGE);
        const uint64_t da0 = ptx::make_smem_desc(a_base, 1024, 0);
        const uint64_t db0 = ptx::make_smem_desc(a_base + 2 * kTile16K, 1024, kBmtAtom);  // MN-major: SBO 8 k-rows, LBO next 64 columns
        for (int t = 0; t < mt; ++t) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // 4 x UMMA_K(16) = 64
            ptx::umma_f16_elect(tmem + t * 128, ptx::desc_advance(da0, t * kTile16K + kk * kUmmaKBytes), ptx::desc_advance(db0, kk * 2048), idesc,
                                (i > 0 || kk > 0) ? 1u : 0u);
        }
        ptx::umma_commit_elect(ptx::smem_u32(empty_bar + s));
      }
      __syncwarp();
    }
    ptx::umma_commit_elect(ptx::smem_u32(done_bar));
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const bool issuer = threadIdx.x == 64;
    const uint32_t st = ptx::smem_u32(staging);
    ptx::mbar_wait(ptx::smem_u32(done_bar), 0);
    __syncwarp();
    ptx::tc_fence_after();
    for (int t = 0; t < mt; ++t) {
      for (int half = 0; half < BN / 64; ++half) {
        uint32_t w[32];
        {
          uint32_t ra[32], rb[32];
          const uint32_t acc = tmem + ((uint32_t)(quarter * 32) << 16) + t * 128 + half * 64;
          ptx::tmem_ld_32x32b_x32(acc, ra);
          ptx::tmem_ld_32x32b_x32(acc + 32, rb);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            w[j] = Act<T>::pack(__uint_as_float(ra[2 * j]), __uint_as_float(ra[2 * j + 1]));
            w[16 + j] = Act<T>::pack(__uint_as_float(rb[2 * j]), __uint_as_float(rb[2 * j + 1]));
          }
        }
        if (issuer) ptx::tma_store_wait_read<0>();
        ptx::named_bar_sync(1, 128);
#pragma unroll
        for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(1, 128);
        if (issuer) {
          ptx::tma_store_3d(&tm_c, st, n0 + half * 64, m0 + t * kBlockM, batch);
          ptx::tma_store_commit();
        }
      }
    }
    if (issuer) ptx::tma_store_wait_read<0>();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, 256);
}

// Any shape; B either ND [batch, k, n] or NZ [batch, ceil(n/16), k, 16].  fp32 accumulation in ascending k.
template <class T>
__global__ void bmt_generic_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ c, long long m, long long batch,
                                   long long k, long long n, int b_is_nz) {
  const long long total = m * batch * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long ni = i % n, bi = (i / n) % batch, mi = i / (n * batch);
    const T* ar = a + (mi * batch + bi) * k;
    float acc = 0.f;
    if (b_is_nz) {
      const T* bp = b + ((bi * ((n + 15) / 16) + ni / 16) * k) * 16 + ni % 16;
      for (long long kk = 0; kk < k; ++kk) acc = fmaf(Act<T>::to_f(ar[kk]), Act<T>::to_f(bp[kk * 16]), acc);
    } else {
      const T* bp = b + bi * k * n + ni;
      for (long long kk = 0; kk < k; ++kk) acc = fmaf(Act<T>::to_f(ar[kk]), Act<T>::to_f(bp[kk * n]), acc);
    }
    c[i] = Act<T>::from_f(acc);
  }
}

}  // namespace mlapo
