// tools/ubench_tmem.cu — two micro-benchmarks that decide how the thin (kx-stacked) layers are bounded:
//   (1) tcgen05.ld throughput (TMEM -> registers) per SM, by instruction shape and number of reading warps;
//   (2) tcgen05.mma issue cadence by N for the operand layouts conv_tc_kernel uses (no-swizzle interleaved A,
//       128-byte-swizzled B, K = 16 per instruction), M = 128, back-to-back dependent accumulation.
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench_tmem tools/ubench_tmem.cu
// Run on a B200; prints one line per configuration.  Not part of the library.
#include <cstdio>
#include <cstdlib>
#include "../image-segmentation_b200/csrc/ub_ptx.cuh"

using namespace ub;

template <int S> __device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t (&r)[64]);
template <int S> struct LdInfo;

template <> __device__ __forceinline__ void tmem_ld<0>(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr) : "memory");
}
template <> struct LdInfo<0> { static constexpr int regs = 16; static constexpr int lanes = 32;
  static constexpr const char* name = "32x32b.x16"; };

template <> __device__ __forceinline__ void tmem_ld<1>(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr) : "memory");
}
template <> struct LdInfo<1> { static constexpr int regs = 32; static constexpr int lanes = 32;
  static constexpr const char* name = "32x32b.x32"; };

template <> __device__ __forceinline__ void tmem_ld<2>(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63]) : "r"(taddr) : "memory");
}
template <> struct LdInfo<2> { static constexpr int regs = 64; static constexpr int lanes = 32;
  static constexpr const char* name = "32x32b.x64"; };

template <> __device__ __forceinline__ void tmem_ld<3>(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr) : "memory");
}
template <> struct LdInfo<3> { static constexpr int regs = 32; static constexpr int lanes = 16;
  static constexpr const char* name = "16x256b.x8"; };

template <> __device__ __forceinline__ void tmem_ld<4>(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile("tcgen05.ld.sync.aligned.16x128b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr) : "memory");
}
template <> struct LdInfo<4> { static constexpr int regs = 32; static constexpr int lanes = 16;
  static constexpr const char* name = "16x128b.x16"; };

template <> __device__ __forceinline__ void tmem_ld<5>(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile("tcgen05.ld.sync.aligned.16x64b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr) : "memory");
}
template <> struct LdInfo<5> { static constexpr int regs = 32; static constexpr int lanes = 16;
  static constexpr const char* name = "16x64b.x32"; };

// Every warp of the CTA reads its own 32-lane quadrant (warp % 4), `iters` times, two loads in flight.
template <int S>
__global__ void __launch_bounds__(512, 1) ld_bench(int iters, unsigned long long* out_cycles, uint32_t* sink) {
  __shared__ uint32_t s_tmem;
  __shared__ unsigned long long s_t0, s_t1;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { s_t0 = ~0ull; s_t1 = 0ull; }
  if (warp == 0) { tmem_alloc(&s_tmem, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = s_tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  uint32_t r[64], acc = 0;
  constexpr int cols = LdInfo<S>::regs * 32 / LdInfo<S>::lanes;  // TMEM columns one instruction covers (per lane)
  const unsigned long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    const uint32_t col = static_cast<uint32_t>((i * cols) & 255);
    tmem_ld<S>(base + col, r);
    if (LdInfo<S>::lanes == 16) tmem_ld<S>(base + (16u << 16) + col, r);   // the other half of the quadrant
    else tmem_ld<S>(base + ((col + cols) & 255u), r);
    tmem_ld_wait();
    acc ^= r[0] ^ r[LdInfo<S>::regs - 1];
  }
  const unsigned long long t1 = clock64();
  atomicMin(&s_t0, t0);
  atomicMax(&s_t1, t1);
  if (acc == 0x12345678u) sink[0] = acc;
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) out_cycles[blockIdx.x] = s_t1 - s_t0;
  if (warp == 0) { tc_fence_after(); tmem_dealloc(s_tmem, 512); }
}

template <int S>
void run_ld(unsigned long long* d_out, uint32_t* d_sink) {
  for (int warps : {4, 8, 16}) {
    const int iters = 2000;
    ld_bench<S><<<8, warps * 32>>>(iters, d_out, d_sink);
    ld_bench<S><<<8, warps * 32>>>(iters, d_out, d_sink);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[8];
    cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
    const double bytes = 2.0 * iters * warps * LdInfo<S>::regs * 32 * 4;
    printf("tcgen05.ld %-14s warps %2d : %8.1f B/cycle/SM  (%llu cycles, %s)\n", LdInfo<S>::name, warps, bytes / h[0], h[0],
           cudaGetErrorString(e));
  }
}

// MMA cadence: one thread issues n_mma MMAs (8 per unrolled loop body, descriptors precomputed: the loop is a handful of
// instructions per MMA, well under the tensor time), then one commit; cycles from the first issue to the commit's arrival.
// A: no-swizzle interleaved (LBO = plane stride, SBO = 8-pixel-group stride) with the conv kernel's tap offsets, or
// 128-byte swizzled; B: SW128 rows of 128 bytes.  ACC rotating accumulators (1 = every MMA depends on the previous one).
template <int N, int ACC>
__global__ void __launch_bounds__(128, 1) mma_bench(int n_mma, int lbo, int sbo, int a_sw128, unsigned long long* out_cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint32_t s_tmem;
  __shared__ uint64_t bar;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (64 * 1024 + N * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(&s_tmem, 512); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 32) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t da = a_sw128 ? umma_smem_desc<128>(smem_u32(smem)) : umma_smem_desc_interleaved(smem_u32(smem), lbo, sbo);
    const uint64_t db = umma_smem_desc<128>(smem_u32(smem + 64 * 1024));
    const uint32_t a_hi = static_cast<uint32_t>(da >> 32), b_hi = static_cast<uint32_t>(db >> 32);
    const uint32_t a_lo = static_cast<uint32_t>(da), b_lo = static_cast<uint32_t>(db);
    uint32_t al[8], bl[8], dd[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      al[j] = a_sw128 ? a_lo + 2u * (j & 3) : a_lo + static_cast<uint32_t>((j % 3) + (j / 3) * 10 + (j & 1) * (2 * lbo >> 4));
      bl[j] = b_lo + 2u * (j & 3);
      dd[j] = s_tmem + static_cast<uint32_t>((j % ACC) * (N <= 128 ? 128 : 256));
    }
#pragma unroll
    for (int j = 0; j < ACC; ++j) umma_bf16_lo<false>(dd[j], al[j], a_hi, bl[j], b_hi, idesc);
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    const unsigned long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n_mma; i += 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j) umma_bf16_lo<true>(dd[j], al[j], a_hi, bl[j], b_hi, idesc);
    }
    umma_commit(&bar);
    const unsigned long long t_issue = clock64();
    mbar_wait(&bar, 1);
    const unsigned long long t1 = clock64();
    out_cycles[2 * blockIdx.x] = t1 - t0;
    out_cycles[2 * blockIdx.x + 1] = t_issue - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(s_tmem, 512); }
}

template <int N, int ACC = 1>
void run_mma(unsigned long long* d_out, const char* layout, int lbo, int sbo, int a_sw128 = 0) {
  const int smem = 64 * 1024 + N * 128 + 2048;
  cudaFuncSetAttribute(mma_bench<N, ACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int n_mma = 2048;
  mma_bench<N, ACC><<<8, 128, smem>>>(n_mma, lbo, sbo, a_sw128, d_out);
  mma_bench<N, ACC><<<8, 128, smem>>>(n_mma, lbo, sbo, a_sw128, d_out);
  cudaError_t e = cudaDeviceSynchronize();
  unsigned long long h[2];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("tcgen05.mma M=128 N=%3d K=16 A=%-22s accumulators %d : %7.1f cycles/MMA (tensor floor %5.1f), issue loop %6.1f "
         "cycles/MMA (%s)\n", N, layout, ACC, static_cast<double>(h[0]) / n_mma, N / 2.0,
         static_cast<double>(h[1]) / n_mma, cudaGetErrorString(e));
}

int main() {
  unsigned long long* d_out;
  uint32_t* d_sink;
  cudaMalloc(&d_out, 64 * sizeof(unsigned long long));
  cudaMalloc(&d_sink, 64);
  run_ld<0>(d_out, d_sink);
  run_ld<1>(d_out, d_sink);
  run_ld<2>(d_out, d_sink);
  run_ld<3>(d_out, d_sink);
  run_ld<4>(d_out, d_sink);
  run_ld<5>(d_out, d_sink);

  // A layouts of conv_tc_kernel: standard tile (TW = 8: box 18 rows x 10 pixels, LBO = 2880, SBO = 160) and the
  // kx-stacked tile (TW = 32: box 6 rows x 32 pixels, LBO = 3072, SBO = 128)
  run_mma<32>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<64>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<128>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<256>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<32>(d_out, "kxs LBO3072 SBO128", 3072, 128);
  run_mma<64>(d_out, "kxs LBO3072 SBO128", 3072, 128);
  run_mma<96>(d_out, "kxs LBO3072 SBO128", 3072, 128);
  run_mma<128>(d_out, "kxs LBO3072 SBO128", 3072, 128);
  run_mma<192>(d_out, "kxs LBO3072 SBO128", 3072, 128);
  // is a floor above the tensor time a dependent-accumulation latency (rotating accumulators hide it) or an operand-fetch
  // cost of the no-swizzle layout (a 128-byte-swizzled A does not pay it)?
  run_mma<32, 2>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<64, 2>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<96, 2>(d_out, "kxs LBO3072 SBO128", 3072, 128);
  run_mma<128, 2>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<32, 4>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<64, 4>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<96, 4>(d_out, "kxs LBO3072 SBO128", 3072, 128);
  run_mma<128, 4>(d_out, "std LBO2880 SBO160", 2880, 160);
  run_mma<32>(d_out, "SW128", 0, 0, 1);
  run_mma<64>(d_out, "SW128", 0, 0, 1);
  run_mma<128>(d_out, "SW128", 0, 0, 1);
  run_mma<256>(d_out, "SW128", 0, 0, 1);
  run_mma<64, 2>(d_out, "SW128", 0, 0, 1);
  run_mma<64, 4>(d_out, "SW128", 0, 0, 1);
  return 0;
}
