// Microbenchmarks that size the GEMM's promotion path on sm_100a:
//   (1) tcgen05.ld throughput (bytes/clk/SM) vs number of warps, load width and loads in flight;
//   (2) packed fp32 FMA (fma.rn.f32x2) throughput vs number of warps.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tmem_bw tmem_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

template <int X> struct Ld;
template <> struct Ld<32> {
  static __device__ __forceinline__ void ld(uint32_t a, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]),"=r"(v[1]),"=r"(v[2]),"=r"(v[3]),"=r"(v[4]),"=r"(v[5]),"=r"(v[6]),"=r"(v[7]),"=r"(v[8]),"=r"(v[9]),"=r"(v[10]),"=r"(v[11]),"=r"(v[12]),"=r"(v[13]),"=r"(v[14]),"=r"(v[15]),
        "=r"(v[16]),"=r"(v[17]),"=r"(v[18]),"=r"(v[19]),"=r"(v[20]),"=r"(v[21]),"=r"(v[22]),"=r"(v[23]),"=r"(v[24]),"=r"(v[25]),"=r"(v[26]),"=r"(v[27]),"=r"(v[28]),"=r"(v[29]),"=r"(v[30]),"=r"(v[31])
      : "r"(a));
  }
};
template <> struct Ld<16> {
  static __device__ __forceinline__ void ld(uint32_t a, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]),"=r"(v[1]),"=r"(v[2]),"=r"(v[3]),"=r"(v[4]),"=r"(v[5]),"=r"(v[6]),"=r"(v[7]),"=r"(v[8]),"=r"(v[9]),"=r"(v[10]),"=r"(v[11]),"=r"(v[12]),"=r"(v[13]),"=r"(v[14]),"=r"(v[15])
      : "r"(a));
  }
};
template <> struct Ld<64> {
  static __device__ __forceinline__ void ld(uint32_t a, uint32_t* v) { Ld<32>::ld(a, v); Ld<32>::ld(a + 32, v + 32); }
};

// X = columns per load instruction, INFLIGHT = loads issued before each wait
template <int X, int INFLIGHT>
__global__ void tmem_ld_kernel(int iters, long long* cycles, uint32_t* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t base = slot + (uint32_t((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  uint32_t v[INFLIGHT][X];
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < INFLIGHT; ++j) Ld<X>::ld(base + uint32_t(((it * INFLIGHT + j) * X) & 511 & ~(X - 1)), v[j]);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < INFLIGHT; ++j)
#pragma unroll
      for (int i = 0; i < X; i += 8) acc ^= v[j][i];
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) sink[0] = acc;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(slot), "r"(512));
}

template <int NACC>
__global__ void ffma2_kernel(int iters, long long* cycles, float* sink, float s) {
  float a[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) a[i] = float(threadIdx.x + i);
  float vv = float(threadIdx.x) * 1e-3f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; i += 2)
      asm volatile("{\n\t.reg .b64 acc, val, scl;\n\tmov.b64 acc, {%0, %1};\n\tmov.b64 val, {%2, %2};\n\tmov.b64 scl, {%3, %3};\n\t"
                   "fma.rn.f32x2 acc, val, scl, acc;\n\tmov.b64 {%0, %1}, acc;\n\t}" : "+f"(a[i]), "+f"(a[i + 1]) : "f"(vv), "f"(s));
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float r = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) r += a[i];
  if (r == 1.2345f) sink[0] = r;
}

template <int NACC>
__global__ void ffma_kernel(int iters, long long* cycles, float* sink, float s) {
  float a[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) a[i] = float(threadIdx.x + i);
  float vv = float(threadIdx.x) * 1e-3f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(vv), "f"(s));
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float r = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) r += a[i];
  if (r == 1.2345f) sink[0] = r;
}

static double avg(long long* h, int n) { double s = 0; for (int i = 0; i < n; ++i) s += double(h[i]); return s / n; }

int main() {
  const int nsm = 148, iters = 2000;
  long long* d_cyc; uint32_t* d_sink; float* d_fsink;
  CK(cudaMalloc(&d_cyc, nsm * sizeof(long long))); CK(cudaMalloc(&d_sink, 4)); CK(cudaMalloc(&d_fsink, 4));
  long long h[nsm];
#define RUN_LD(X, INF, WARPS) do { \
    tmem_ld_kernel<X, INF><<<nsm, WARPS * 32>>>(iters, d_cyc, d_sink); CK(cudaDeviceSynchronize()); \
    tmem_ld_kernel<X, INF><<<nsm, WARPS * 32>>>(iters, d_cyc, d_sink); CK(cudaDeviceSynchronize()); \
    CK(cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost)); \
    double c = avg(h, nsm); double bytes = double(iters) * INF * X * 4.0 * 32 * WARPS; \
    printf("tmem_ld x%-3d inflight=%d warps=%2d : %8.0f clk  %7.1f B/clk/SM  %6.1f clk per ld-instr per warp\n", X, INF, WARPS, c, bytes / c, c / (double(iters) * INF)); } while (0)
  RUN_LD(32, 1, 4); RUN_LD(32, 2, 4); RUN_LD(32, 4, 4);
  RUN_LD(32, 1, 8); RUN_LD(32, 2, 8); RUN_LD(32, 4, 8);
  RUN_LD(32, 1, 16); RUN_LD(32, 2, 16);
  RUN_LD(16, 1, 8); RUN_LD(16, 2, 8); RUN_LD(16, 4, 8); RUN_LD(16, 2, 16); RUN_LD(16, 4, 16);
  RUN_LD(64, 1, 4); RUN_LD(64, 1, 8); RUN_LD(64, 2, 4);
#define RUN_F(K, NACC, WARPS, PER) do { \
    K<NACC><<<nsm, WARPS * 32>>>(iters, d_cyc, d_fsink, 1.0001f); CK(cudaDeviceSynchronize()); \
    K<NACC><<<nsm, WARPS * 32>>>(iters, d_cyc, d_fsink, 1.0001f); CK(cudaDeviceSynchronize()); \
    CK(cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost)); \
    double c = avg(h, nsm); double fmas = double(iters) * NACC * 32 * WARPS; \
    printf(#K " nacc=%-3d warps=%2d : %8.0f clk  %6.1f fp32 FMA/clk/SM  (%.2f warp-instr/clk/SM)\n", NACC, WARPS, c, fmas / c, fmas / c / 32 / PER); } while (0)
  RUN_F(ffma2_kernel, 64, 4, 2); RUN_F(ffma2_kernel, 64, 8, 2); RUN_F(ffma2_kernel, 64, 16, 2); RUN_F(ffma2_kernel, 128, 8, 2);
  RUN_F(ffma_kernel, 64, 4, 1); RUN_F(ffma_kernel, 64, 8, 1); RUN_F(ffma_kernel, 64, 16, 1);
  return 0;
}
