// Device check of grackle_b200/csrc/fastmath.cuh: ddiv/drcp/dlog/dexp/dexp10 against the CUDA
// library versions over wide exponent ranges.  Prints "name max_rel_err count_bad".
// Built by tests/test_gpu_parity.py::test_fastmath_on_device (nvcc) and run on the GPU.
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
#include "../grackle_b200/csrc/fastmath.cuh"

__device__ double rel(double a, double b)
{
  if (a == b) return 0.0;
  if (isnan(a) && isnan(b)) return 0.0;
  return fabs(a - b) / fmax(fabs(b), 1e-300);
}

__global__ void k_check(double *worst, unsigned long long *bad)
{
  const int ea = (int)blockIdx.x - 300;            // exponent of a: 10^-300 .. 10^300
  const int eb = (int)threadIdx.x * 2 - 300;       // exponent of b
  const double a = 1.2345678901234567 * exp10((double)ea);
  const double b = 7.6543210987654321 * exp10((double)eb);
  const double lq = log10(fabs(a / b));
  if (lq > -290 && lq < 290) {                     // quotient representable
    const double e1 = rel(grb::ddiv(a, b), a / b);
    const double e2 = rel(grb::ddiv(-a, b), -a / b);
    const double e3 = rel(grb::ddiv(a, -b), a / -b);
    const double e = fmax(e1, fmax(e2, e3));
    if (e > 4e-16) atomicAdd(bad + 0, 1ULL);
    atomicMax((unsigned long long *)worst + 0, (unsigned long long)__double_as_longlong(e));
  }
  {
    const double e = rel(grb::drcp(b), 1.0 / b);
    if (e > 4e-16) atomicAdd(bad + 1, 1ULL);
    atomicMax((unsigned long long *)worst + 1, (unsigned long long)__double_as_longlong(e));
  }
  {
    const double e = fabs(grb::dlog(a) - log(a)) / fmax(fabs(log(a)), 1.0);
    if (e > 1e-15) atomicAdd(bad + 2, 1ULL);
    atomicMax((unsigned long long *)worst + 2, (unsigned long long)__double_as_longlong(e));
  }
  {
    const double x = (double)ea * 2.3 + (double)threadIdx.x * 0.001;      // -690 .. 690
    const double e = rel(grb::dexp(x), exp(x));
    if (e > 1e-15) atomicAdd(bad + 3, 1ULL);
    atomicMax((unsigned long long *)worst + 3, (unsigned long long)__double_as_longlong(e));
    const double y = (double)ea + (double)threadIdx.x * 0.003;           // -300 .. 301
    const double f = rel(grb::dexp10(y), exp10(y));
    if (f > 1e-15) atomicAdd(bad + 4, 1ULL);
    atomicMax((unsigned long long *)worst + 4, (unsigned long long)__double_as_longlong(f));
  }
}

int main()
{
  double *worst; unsigned long long *bad;
  cudaMallocManaged(&worst, 5 * sizeof(double));
  cudaMallocManaged(&bad, 5 * sizeof(unsigned long long));
  for (int i = 0; i < 5; i++) { worst[i] = 0; bad[i] = 0; }
  k_check<<<601, 301>>>(worst, bad);
  if (cudaDeviceSynchronize() != cudaSuccess) { std::printf("cuda error\n"); return 2; }
  const char *names[5] = {"div", "rcp", "log", "exp", "exp10"};
  int rc = 0;
  for (int i = 0; i < 5; i++) {
    std::printf("%s %.3e %llu\n", names[i], worst[i], bad[i]);
    if (bad[i]) rc = 1;
  }
  return rc;
}
