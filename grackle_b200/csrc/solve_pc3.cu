// grackle-b200: the k_solve instantiations of primordial_chemistry = 3 (solve_kernel.cuh).
// One translation unit per network so that the five physics variants x two host-pipeline
// modes of each compile in parallel.
#include "solve_kernel.cuh"

void grb_launch_solve_pc3(const GrParams &P, int var, int blocks, cudaStream_t st)
{
  grb::launch_solve_pc<3>(P, var, blocks, st);
}

int grb_solve_max_blocks_pc3() { return grb::solve_max_blocks_pc<3>(); }
