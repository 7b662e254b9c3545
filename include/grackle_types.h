/* grackle-b200 public types.
 *
 * ABI contract: these three PODs must be layout-identical to the reference's
 * grackle_types.h:36-104 (grackle_field_data, code_units, grackle_version) so a
 * simulation code compiled against upstream headers can link this library.
 * Field arrays are SoA, Fortran order (i fastest): idx = i + ni*(j + nj*k)
 * (reference index_helper.h:50-56). Pointers may be HOST or DEVICE memory;
 * the library detects which per call ("boundary").
 */
#ifndef GRACKLE_B200_TYPES_H
#define GRACKLE_B200_TYPES_H

#include "grackle_float.h"

#ifdef GRACKLE_FLOAT_8
typedef double gr_float_t;
#define gr_float double
#else
#error "grackle-b200 is an FP64-only build (GRACKLE_FLOAT_8)"
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* one SoA view of the caller's grid; the library never owns these arrays */
typedef struct
{
  /* geometry: rank 1..3, per-axis extents, inclusive zero-based active window */
  int grid_rank;
  int *grid_dimension;
  int *grid_start;
  int *grid_end;
  gr_float grid_dx;

  /* mass densities in code units (every species, He carries mass 4, H2 mass 2) */
  gr_float *density;
  gr_float *HI_density;
  gr_float *HII_density;
  gr_float *HM_density;
  gr_float *HeI_density;
  gr_float *HeII_density;
  gr_float *HeIII_density;
  gr_float *H2I_density;
  gr_float *H2II_density;
  gr_float *DI_density;
  gr_float *DII_density;
  gr_float *HDI_density;
  gr_float *e_density;
  gr_float *metal_density;
  gr_float *dust_density;

  /* specific internal energy and (unused by the solve) velocities */
  gr_float *internal_energy;
  gr_float *x_velocity;
  gr_float *y_velocity;
  gr_float *z_velocity;

  /* optional per-cell source terms; only dereferenced when their flag is on */
  gr_float *volumetric_heating_rate;
  gr_float *specific_heating_rate;
  gr_float *RT_heating_rate;
  gr_float *RT_HI_ionization_rate;
  gr_float *RT_HeI_ionization_rate;
  gr_float *RT_HeII_ionization_rate;
  gr_float *RT_H2_dissociation_rate;
  gr_float *H2_self_shielding_length;
  gr_float *H2_custom_shielding_factor;
  gr_float *isrf_habing;
} grackle_field_data;

/* cgs conversion factors of the caller's unit system */
typedef struct
{
  int comoving_coordinates;
  double density_units;
  double length_units;
  double time_units;
  double velocity_units;
  double a_units;
  double a_value;
} code_units;

typedef struct
{
  const char *version;
  const char *branch;
  const char *revision;
} grackle_version;

#ifdef __cplusplus
}
#endif
#endif /* GRACKLE_B200_TYPES_H */
