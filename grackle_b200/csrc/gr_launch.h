// grackle-b200: launch entry points of kernels.cu, called from api.cu
#ifndef GRB200_LAUNCH_H
#define GRB200_LAUNCH_H
#include <cuda_runtime.h>
#include "gr_device.h"

#ifndef GRB_SOLVE_THREADS
#define GRB_SOLVE_THREADS 512    /* one CTA per SM: 16 warps, <=128 registers per thread */
#endif
#ifndef GRB_SOLVE_MINBLOCKS
#define GRB_SOLVE_MINBLOCKS 1     /* resident CTAs per SM the register allocator must allow */
#endif

void grb_launch_thermo(const GrParams &P, int mode, cudaStream_t st);     // 0 p, 1 T, 2 gamma
void grb_launch_fill(const GrParams &P, double value, cudaStream_t st);
void grb_launch_temperature_table(const GrParams &P, cudaStream_t st);
void grb_launch_cooling_time(const GrParams &P, cudaStream_t st);
void grb_launch_dust_temperature(const GrParams &P, cudaStream_t st);
void grb_launch_solve(const GrParams &P, int blocks, cudaStream_t st);
int grb_solve_variant_for(const GrParams &P);   // 0 general, 1 lean, 2 lean + H2-on-dust + UVB self-shielding
extern int grb_solve_variant;       // 0 = choose per call, 1 = always the general instantiation
int grb_solve_max_blocks_per_sm(int pc);
// cost-ordered work queue (order.cu)
size_t grb_cost_order_bytes(long long n);
const unsigned *grb_build_cost_order(const GrParams &P, void *scratch, size_t scratch_bytes, cudaStream_t st);
double grb_launch_fp64_peak(double *sink, int blocks, int iters, cudaStream_t st);
#endif
