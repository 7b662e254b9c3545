// grackle-b200 microbenchmark: how fast can one SM stage "one table row per cell" - the access
// pattern of the solve kernel's rate lookup (every cell at its own log-T bin, a 461 KB table that
// lives in L2) - with each of the three mechanisms sm_100a offers?
//
//   coop    what the product does (cell_io.cuh coop_fetch): 8 lanes read one cell's row as
//           consecutive 16-byte (value, difference) pairs with LDG.128, interpolate a + d t in
//           registers and store the VALUES column-major to shared memory; the owner reads 8 bytes per column
//   bulk    TMA 1-D bulk copy: every lane issues ONE cp.async.bulk (UBLKCP) of its cell's whole row
//           into the cell's own raw row in shared memory, completion on one mbarrier per warp; the
//           owner then reads the 16-byte pairs and interpolates
//   ldgsts  cp.async (LDGSTS.128) in the cooperative pattern: no register staging, raw pairs in
//           shared memory as for bulk
//
// bulk and ldgsts keep RAW pairs (16 bytes per column and cell), coop keeps interpolated values
// (8 bytes): at 512 threads per SM the product's 34 slot columns + per-cell state take 183 KB of the
// 227 KB, so raw rows of 32 columns (528 B x 512 = 270 KB) do not fit - the probe therefore runs
// 16 columns at 512 threads and 32 columns at 256 threads, every mechanism at both.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tests/bin/tma_row_probe tools/tma_row_probe.cu
//   tests/bin/tma_row_probe            (prints one line per mechanism and shape)
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int ROWS = 600;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t lcg(uint32_t &s)
{
  s = s * 1664525u + 1013904223u;
  return s;
}

__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}

// MODE 0 coop, 1 bulk, 2 ldgsts
template <int MODE, int NCOLS, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_probe(const double *table, double *sink, int iters)
{
  constexpr int ROWB = NCOLS * 16;
  constexpr int G = 8, CPP = 32 / G, PER = NCOLS / G;
  constexpr int CS = THREADS + 16 / G;           // column stride of the value slots (coop)
  constexpr int RAWB = ROWB + 16;                // padded raw row: conflict-free 16-byte reads
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t seed = (blockIdx.x * THREADS + tid) * 2654435761u + 12345u;
  double acc = 0.0;

  double *slots = reinterpret_cast<double *>(smem);                       // coop: [NCOLS][CS]
  unsigned char *raw = smem;                                              // bulk/ldgsts: [THREADS][RAWB]
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)THREADS * RAWB);   // one per warp
  if (MODE == 1) {
    if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + warp)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
  }
  const uint32_t bar = smem_u32(bars + warp);
  uint32_t parity = 0;

  for (int it = 0; it < iters; it++) {
    const uint32_t r = lcg(seed);
    const int bin = (int)(r % (ROWS - 1));
    const double t = (double)(r >> 8) * (1.0 / 16777216.0);
    const int off = bin * ROWB;

    if (MODE == 0) {
      const int g = lane & (G - 1), sub = lane / G;
      const char *src = reinterpret_cast<const char *>(table) + 16 * g;
      double *dst = slots + (size_t)(warp * 32) + g * CS + sub;
#pragma unroll
      for (int pass = 0; pass < G; pass++) {
        const int cell = pass * CPP + sub;
        const int o = __shfl_sync(0xffffffffu, off, cell);
        const double tt = __shfl_sync(0xffffffffu, t, cell);
#pragma unroll
        for (int k = 0; k < PER; k++) {
          const double2 p = __ldg(reinterpret_cast<const double2 *>(src + (unsigned)o + 16 * k * G));
          dst[pass * CPP + k * G * CS] = p.x + p.y * tt;
        }
      }
      __syncwarp();
      const double *mine = slots + tid;
#pragma unroll
      for (int c = 0; c < NCOLS; c++) acc += mine[c * CS];
      __syncwarp();
    } else {
      unsigned char *my_row = raw + (size_t)tid * RAWB;
      if (MODE == 1) {
        // the rows were read through the generic proxy in the previous iteration: order those reads
        // before the async-proxy writes of this one
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (lane == 0)
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(32 * ROWB) : "memory");
        __syncwarp();
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(my_row)), "l"(reinterpret_cast<const char *>(table) + off), "r"(ROWB), "r"(bar)
                     : "memory");
        while (!mbar_try_wait(bar, parity)) {}
        parity ^= 1u;
      } else {
        const int g = lane & (G - 1), sub = lane / G;
        const char *src = reinterpret_cast<const char *>(table) + 16 * g;
#pragma unroll
        for (int pass = 0; pass < G; pass++) {
          const int cell = pass * CPP + sub;
          const int o = __shfl_sync(0xffffffffu, off, cell);
          unsigned char *row = raw + (size_t)(warp * 32 + cell) * RAWB + 16 * g;
#pragma unroll
          for (int k = 0; k < PER; k++)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(row + 16 * k * G)),
                         "l"(src + (unsigned)o + 16 * k * G) : "memory");
        }
        asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
        __syncwarp();
      }
      const double2 *mine = reinterpret_cast<const double2 *>(my_row);
#pragma unroll
      for (int c = 0; c < NCOLS; c++) {
        const double2 p = mine[c];
        acc += p.x + p.y * t;
      }
      __syncwarp();
    }
  }
  if (acc == 1.2345e-300) sink[0] = acc;      // never true: keeps the work alive
}

template <int MODE, int NCOLS, int THREADS>
static void run(const char *name, const double *table, double *sink, int sms)
{
  constexpr int ROWB = NCOLS * 16;
  const size_t smem = MODE == 0 ? (size_t)NCOLS * (THREADS + 2) * 8 : (size_t)THREADS * (ROWB + 16) + 8 * (THREADS / 32);
  auto kern = k_probe<MODE, NCOLS, THREADS>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    std::printf("%-7s cols %2d threads %3d: %zu bytes of shared memory do not fit\n", name, NCOLS, THREADS, smem);
    cudaGetLastError();
    return;
  }
  const int iters = 4000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  kern<<<sms, THREADS, smem>>>(table, sink, 200);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0);
    kern<<<sms, THREADS, smem>>>(table, sink, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  const cudaError_t err = cudaGetLastError();
  const double rows = (double)sms * THREADS * iters;
  std::printf("%-7s cols %2d (%3d-byte rows) threads/SM %3d  smem %6.1f KB  %8.3f ms  %7.2f G rows/s  %6.2f TB/s of table bytes%s\n",
              name, NCOLS, ROWB, THREADS, smem / 1024.0, best, rows / best / 1e6, rows * ROWB / best / 1e9,
              err == cudaSuccess ? "" : "  CUDA ERROR");
  cudaEventDestroy(e0); cudaEventDestroy(e1);
}

int main()
{
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { std::printf("no CUDA device\n"); return 1; }
  const int sms = prop.multiProcessorCount;
  const size_t n = (size_t)ROWS * 48 * 2;       // the chemistry table of the product: 600 rows x 48 pairs
  std::vector<double> h(n);
  for (size_t i = 0; i < n; i++) h[i] = 1.0 + 1e-3 * (double)(i % 977);
  double *table, *sink;
  cudaMalloc(&table, n * sizeof(double));
  cudaMalloc(&sink, 64);
  cudaMemcpy(table, h.data(), n * sizeof(double), cudaMemcpyHostToDevice);
  std::printf("%s, %d SMs; table %zu KB (L2 resident); one row per cell per iteration at a pseudo-random bin\n", prop.name, sms,
              n * 8 / 1024);
  run<0, 16, 512>("coop", table, sink, sms);
  run<1, 16, 512>("bulk", table, sink, sms);
  run<2, 16, 512>("ldgsts", table, sink, sms);
  run<0, 32, 256>("coop", table, sink, sms);
  run<1, 32, 256>("bulk", table, sink, sms);
  run<2, 32, 256>("ldgsts", table, sink, sms);
  run<0, 32, 512>("coop", table, sink, sms);
  run<1, 32, 512>("bulk", table, sink, sms);
  return 0;
}
