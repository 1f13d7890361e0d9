// umma_probe.cu -- how long does one tcgen05.mma (M=128, K=16, bf16) occupy the issue path, as a function of
// N, of the number of independent TMEM accumulators the chain is spread over, and of how often the chain is
// committed?  Also: what does picking the issuing thread with `lane == 0` instead of elect.sync cost?
// Standalone: nvcc -gencode arch=compute_100a,code=sm_100a -o build/umma_probe tools/umma_probe.cu
// Operand contents are irrelevant (zeros); only timing is read.  Everything the loop needs is a compile-time
// constant so that the loop body is the MMA issue and nothing else (a runtime `%` costs ~100 cycles).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "../super_resolution_b200/csrc/tc_common.cuh"

using namespace hb::tc;

struct Res { long long issue, total; };

template <bool LEADER, int N, int NACC, int CE>      // CE: MMAs per commit (0 = one commit at the end)
__global__ void __launch_bounds__(128) probe(int groups, Res* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[2];
  __shared__ uint32_t slot;
  uint8_t* base = (uint8_t*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += 128) ((uint32_t*)base)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc(&slot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x < 32 && (LEADER ? elect_one() : threadIdx.x == 0)) {
    constexpr int G = CE ? CE : 8;
    const uint32_t idesc = make_idesc_bf16(128, N);
    const uint64_t ad = make_smem_desc(smem_u32(base), 128), bd = make_smem_desc(smem_u32(base + 24 * 1024), 128);
    const long long t0 = clock64();
    for (int g = 0; g < groups; ++g) {
#pragma unroll
      for (int k = 0; k < G; ++k)
        umma_bf16(tm + (uint32_t)((k % NACC) * N), ad + 2 * (k & 3), bd + 2 * (k & 3), idesc, (uint32_t)(g | (k >= NACC)));
      if (CE) umma_commit(&bar[1]);
    }
    const long long t1 = clock64();
    umma_commit(&bar[0]);
    mbar_wait(&bar[0], 0);
    const long long t2 = clock64();
    out[blockIdx.x].issue = t1 - t0;
    out[blockIdx.x].total = t2 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

static Res* d;
template <bool LEADER, int N, int NACC, int CE>
void run() {
  constexpr int G = CE ? CE : 8;
  const int groups = 512 / G;
  cudaFuncSetAttribute(probe<LEADER, N, NACC, CE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66 * 1024);
  probe<LEADER, N, NACC, CE><<<148, 128, 66 * 1024>>>(2, d);
  probe<LEADER, N, NACC, CE><<<148, 128, 66 * 1024>>>(groups, d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
  Res h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  double si = 0, st = 0;
  for (int i = 0; i < 148; ++i) { si += h[i].issue; st += h[i].total; }
  const double n = 148.0 * groups * G;
  printf("%5d %5d %5d %5d | %10.1f %10.1f %10.1f | %10.1f\n", (int)LEADER, N, NACC, CE, si / n, st / n, 128.0 * N / 256,
         CE ? st / n * CE : 0.0);
}

template <int N>
void sweep() {
  run<false, N, 1, 0>(); run<false, N, 1, 4>();
  run<true, N, 1, 0>(); run<true, N, 2, 0>();
  run<true, N, 1, 1>(); run<true, N, 1, 2>(); run<true, N, 1, 4>(); run<true, N, 1, 8>(); run<true, N, 1, 16>();
  run<true, N, 2, 4>(); run<true, N, 2, 8>();
}

int main() {
  cudaMalloc(&d, 148 * sizeof(Res));
  printf("%5s %5s %5s %5s | %10s %10s %10s | %10s\n", "elect", "N", "nacc", "cmt", "issue/mma", "total/mma", "ideal", "cyc/group");
  sweep<32>(); sweep<64>(); sweep<128>(); sweep<192>(); sweep<256>();
  return 0;
}
