// Microbenchmark: TMEM -> register bandwidth of tcgen05.ld for several shapes / warp counts (diagnostic, not product code).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_bw tmem_ld_bw.cu && ./tmem_ld_bw
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int X>
__device__ __forceinline__ void ld32x32(uint32_t taddr, uint32_t* r);
template <>
__device__ __forceinline__ void ld32x32<32>(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
template <>
__device__ __forceinline__ void ld32x32<16>(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void ld16x256_x8(uint32_t taddr, uint32_t* r) {  // 16 lanes x 256 bit x 8 = 64 columns of 16 lanes: 32 regs
  asm volatile(
      "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// mode 0: x32 + wait each; 1: two x32 back-to-back then wait; 2: four x32 then wait; 3: x16 + wait; 4: 16x256b.x8 pairs
template <int MODE>
__global__ void bench(int warps_active, int iters, long long* out, uint32_t* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  uint32_t acc = 0;
  long long t0 = 0, t1 = 0;
  if (warp < warps_active) {
    const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    __syncwarp();
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      // read all 512 columns of this warp's 32 lanes = 64 KB per warp per iteration
      if (MODE == 0) {
        for (int c = 0; c < 512; c += 32) { uint32_t r[32]; ld32x32<32>(base + c, r); wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) acc ^= r[j]; }
      } else if (MODE == 1) {
        for (int c = 0; c < 512; c += 64) { uint32_t r[64]; ld32x32<32>(base + c, r); ld32x32<32>(base + c + 32, r + 32); wait_ld();
#pragma unroll
          for (int j = 0; j < 64; ++j) acc ^= r[j]; }
      } else if (MODE == 2) {
        for (int c = 0; c < 512; c += 128) { uint32_t r[128]; ld32x32<32>(base + c, r); ld32x32<32>(base + c + 32, r + 32);
          ld32x32<32>(base + c + 64, r + 64); ld32x32<32>(base + c + 96, r + 96); wait_ld();
#pragma unroll
          for (int j = 0; j < 128; ++j) acc ^= r[j]; }
      } else if (MODE == 3) {
        for (int c = 0; c < 512; c += 16) { uint32_t r[16]; ld32x32<16>(base + c, r); wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j) acc ^= r[j]; }
      } else {
        for (int c = 0; c < 512; c += 64) { uint32_t r[64]; ld16x256_x8(base + c, r); ld16x256_x8(base + (16u << 16) + c, r + 32); wait_ld();
#pragma unroll
          for (int j = 0; j < 64; ++j) acc ^= r[j]; }
      }
    }
    t1 = clock64();
  }
  sink[threadIdx.x] = acc;
  if (lane == 0 && warp < warps_active) out[warp] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

template <int MODE>
void run(const char* name) {
  long long* out; uint32_t* sink;
  cudaMalloc(&out, 64 * 8); cudaMalloc(&sink, 1024 * 4);
  for (int w : {1, 4, 8, 16}) {
    const int iters = 50;
    bench<MODE><<<1, 512, 0>>>(w, iters, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    long long h[16]; cudaMemcpy(h, out, w * 8, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < w; ++i) mx = h[i] > mx ? h[i] : mx;
    const double bytes = (double)w * 32 * 512 * 4 * iters;
    printf("%-28s warps=%2d  cycles=%8lld  %.1f B/clk/SM  (%.0f cycles per 64KB warp-pass)\n", name, w, mx, bytes / mx, (double)mx / iters);
  }
}

int main() {
  run<0>("32x32b.x32 + wait");
  run<1>("2 x 32x32b.x32 then wait");
  run<2>("4 x 32x32b.x32 then wait");
  run<3>("32x32b.x16 + wait");
  run<4>("2 x 16x256b.x8 then wait");
  return 0;
}
