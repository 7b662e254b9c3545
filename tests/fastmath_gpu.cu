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

__global__ void k_check(double *worst, unsigned long long *bad, const double *tab)
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
    if (e > 4.5e-16) atomicAdd(bad + 0, 1ULL);
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
    // table-driven versions (the ones the kernels call)
    const double e6 = rel(grb::dexp(x, tab), exp(x));
    if (e6 > 1e-15) atomicAdd(bad + 6, 1ULL);
    atomicMax((unsigned long long *)worst + 6, (unsigned long long)__double_as_longlong(e6));
    const double e7 = rel(grb::dexp10(y, tab), exp10(y));
    if (e7 > 1e-15) atomicAdd(bad + 7, 1ULL);
    atomicMax((unsigned long long *)worst + 7, (unsigned long long)__double_as_longlong(e7));
  }
  {
    const double e = fabs(grb::dlog(a, tab) - log(a)) / fmax(fabs(log(a)), 1.0);
    if (e > 1e-15) atomicAdd(bad + 5, 1ULL);
    atomicMax((unsigned long long *)worst + 5, (unsigned long long)__double_as_longlong(e));
  }
}

int main()
{
  double *worst; unsigned long long *bad;
  double *tab;
  cudaMallocManaged(&worst, 8 * sizeof(double));
  cudaMallocManaged(&bad, 8 * sizeof(unsigned long long));
  cudaMallocManaged(&tab, GRB_FM_TABLE_DOUBLES * sizeof(double));
  grb::fm_fill_tables(tab);
  for (int i = 0; i < 8; i++) { worst[i] = 0; bad[i] = 0; }
  k_check<<<601, 301>>>(worst, bad, tab);
  if (cudaDeviceSynchronize() != cudaSuccess) { std::printf("cuda error\n"); return 2; }
  const char *names[8] = {"div", "rcp", "log", "exp", "exp10", "tlog", "texp", "texp10"};
  int rc = 0;
  for (int i = 0; i < 8; i++) {
    std::printf("%s %.3e %llu\n", names[i], worst[i], bad[i]);
    if (bad[i]) rc = 1;
  }
  return rc;
}
