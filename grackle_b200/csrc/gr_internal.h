// grackle-b200 internal interfaces between the host-side C++ (parameter handling,
// table generation, file loading, UVB scalars) and the CUDA side (device tables,
// kernels, API entry points).  Nothing here is exported from the shared library.
#ifndef GRB200_INTERNAL_H
#define GRB200_INTERNAL_H

#include "grackle.h"

#define GR_FLOAT_UNDEFINED (-99999.0)

// ---- rate_tables.cpp -------------------------------------------------------
int gr_build_rate_tables(chemistry_data *chem, chemistry_data_storage *rates, const code_units *units,
                         double co_length_units, double co_density_units);

// ---- table_io.cpp ----------------------------------------------------------
// group = "Primordial" or "Metals"; mirrors initialize_cloudy_data.c:29-316
int gr_load_cloudy_table(chemistry_data *chem, chemistry_data_storage *rates, cloudy_data *cloudy,
                         const char *group, const code_units *units, int read_data);
// mirrors initialize_UVbackground_data.c:30-341
int gr_load_uvb_table(chemistry_data *chem, chemistry_data_storage *rates);
void gr_free_cloudy_table(cloudy_data *cloudy, int primordial_chemistry, int uvb);

// ---- host.cpp --------------------------------------------------------------
// mirrors update_UVbackground_rates.c:25-289
int gr_update_uvb_rates(const chemistry_data *chem, const chemistry_data_storage *rates,
                        photo_rate_storage *out, const code_units *units);
void gr_comoving_units(const code_units *units, double *co_length, double *co_density);

// ---- device_tables.cu ------------------------------------------------------
int gr_device_attach(const chemistry_data *chem, const chemistry_data_storage *rates);
void gr_device_detach(const chemistry_data_storage *rates);

#endif
