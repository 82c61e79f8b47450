// Micro-benchmark: does a small kernel's CTA get onto an SM that already runs a big-shared-memory CTA, and how soon?
// BIG: 148 CTAs x 576 threads, large dynamic shared memory, spins ~40 us.  BIG2: the same, queued on another stream (cannot be
// placed while BIG runs).  SMALL: 32 CTAs x 128 threads on a third (high-priority) stream; CTA 0 stamps globaltimer when it starts.
// Prints start(SMALL) - start(BIG) for several configurations.  Build: nvcc -arch=sm_100a -O3 -o coresidency coresidency.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
__device__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__global__ void __launch_bounds__(576) big(unsigned long long* out, int slot, long long spin_ns, int regs_pad) {
  extern __shared__ float sm[];
  const unsigned long long t0 = gtimer();
  if (blockIdx.x == 0 && threadIdx.x == 0) out[slot] = t0;
  float acc = threadIdx.x;
  while ((long long)(gtimer() - t0) < spin_ns) { acc = acc * 1.0001f + 1.f; }
  if (acc == 12345.f) sm[threadIdx.x] = acc;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[slot + 1] = gtimer();
}
__global__ void __launch_bounds__(128) small(unsigned long long* out, int slot, long long spin_ns) {
  const unsigned long long t0 = gtimer();
  if (threadIdx.x == 0) atomicMin(&out[slot], t0);
  while ((long long)(gtimer() - t0) < spin_ns) {}
  if (threadIdx.x == 0) atomicMax(&out[slot + 1], gtimer());
}
int main(int argc, char** argv) {
  unsigned long long* d; cudaMalloc(&d, 64 * 8);
  int lo, hi; cudaDeviceGetStreamPriorityRange(&lo, &hi);
  cudaStream_t s1, s2, s3, s3lo;
  cudaStreamCreateWithPriority(&s1, cudaStreamNonBlocking, lo); cudaStreamCreateWithPriority(&s2, cudaStreamNonBlocking, lo);
  cudaStreamCreateWithPriority(&s3, cudaStreamNonBlocking, hi); cudaStreamCreateWithPriority(&s3lo, cudaStreamNonBlocking, lo);
  cudaFuncSetAttribute(big, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  struct Cfg { const char* name; int big_smem; bool big2; bool carve; bool hiprio; bool small_first; } cfgs[] = {
    {"big 204KB, no 2nd big, small default carveout, hi prio", 204 * 1024, false, false, true, false},
    {"big 204KB, no 2nd big, small MAX-SHARED carveout, hi prio", 204 * 1024, false, true, true, false},
    {"big 204KB, 2nd big queued, small MAX-SHARED, hi prio", 204 * 1024, true, true, true, false},
    {"big 204KB, 2nd big queued, small MAX-SHARED, LOW prio", 204 * 1024, true, true, false, false},
    {"big 204KB, 2nd big queued, small default carveout, hi prio", 204 * 1024, true, false, true, false},
    {"big 100KB, 2nd big queued (co-resident), small default, hi prio", 100 * 1024, true, false, true, false},
    {"big 40KB x2, small default, hi prio", 40 * 1024, true, false, true, false},
  };
  for (auto& c : cfgs) {
    cudaFuncSetAttribute(small, cudaFuncAttributePreferredSharedMemoryCarveout, c.carve ? 100 : -1);
    for (int rep = 0; rep < 3; ++rep) {
      unsigned long long init[8] = {0, 0, 0, 0, ~0ull, 0, 0, 0};
      cudaMemcpy(d, init, sizeof init, cudaMemcpyHostToDevice);
      cudaDeviceSynchronize();
      big<<<148, 576, c.big_smem, s1>>>(d, 0, 40000, 0);
      if (c.big2) big<<<148, 576, c.big_smem, s2>>>(d, 2, 40000, 0);
      small<<<32, 128, 0, c.hiprio ? s3 : s3lo>>>(d, 4, 5000);
      cudaDeviceSynchronize();
      unsigned long long h[8]; cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
      if (rep == 2) printf("%-70s small starts %+7.1f us after big start (big ran %.1f us, big2 start %+.1f us), err=%s\n", c.name,
                           ((double)h[4] - (double)h[0]) / 1e3, ((double)h[1] - (double)h[0]) / 1e3, c.big2 ? ((double)h[2] - (double)h[0]) / 1e3 : 0.0,
                           cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
