/* ORACLE - TEST INFRASTRUCTURE ONLY (see fortran_restated.cpp header).
 *
 * Glue that lets the reference's own C sources link without libhdf5 and without
 * its generated files.  It provides:
 *   - the build-generated symbols auto_show_config / auto_show_flags /
 *     get_grackle_version (src/clib/Makefile:201-230 generates them upstream);
 *   - initialize_cloudy_data / initialize_UVbackground_data / _free_cloudy_data with
 *     the upstream signatures and the upstream post-processing
 *     (initialize_cloudy_data.c:62-305, initialize_UVbackground_data.c:30-341) but
 *     taking the RAW file contents from arrays injected by the test harness (which
 *     reads the HDF5 file with grackle_b200/h5lite.py) instead of calling libhdf5.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "grackle_macros.h"
#include "grackle_types.h"
#include "grackle_chemistry_data.h"

void auto_show_config(FILE *fp) { fprintf(fp, "oracle/_ref build: gcc -O2, no HDF5, Fortran restated in C++\n"); }
void auto_show_flags(FILE *fp) { fprintf(fp, "-O2 -ffp-contract=off -DLINUX -DGRACKLE_FLOAT_8\n"); }
grackle_version get_grackle_version(void)
{
  grackle_version v; v.version = "3.2.dev2"; v.branch = "oracle"; v.revision = "cdda20a6"; return v;
}

#define CLOUDY_MAX_DIMENSION 5
#define SMALL_LOG_VALUE -99.0

typedef struct {
  int present, old_style;
  long long rank, dims[CLOUDY_MAX_DIMENSION];
  double *params[CLOUDY_MAX_DIMENSION];   /* raw: Parameter1.., Temperature (linear K) */
  double *cooling, *heating, *mmw;        /* raw file values */
} injected_cloudy;

static injected_cloudy g_inj[2];          /* 0 = Primordial, 1 = Metals */
static struct { long long nz; double *v[15]; } g_uvb;   /* z,k24..k31,piHI,piHeI,piHeII,crsHI,crsHeI,crsHeII */

static double *dup(const double *p, long long n)
{
  if (!p) return NULL;
  double *q = malloc(sizeof(double) * (n > 0 ? n : 1));
  memcpy(q, p, sizeof(double) * n);
  return q;
}

/* harness API ------------------------------------------------------------ */
void oracle_inject_cloudy(int which, int old_style, long long rank, const long long *dims,
                          const double *const *params, const double *cooling,
                          const double *heating, const double *mmw)
{
  injected_cloudy *c = &g_inj[which];
  memset(c, 0, sizeof(*c));
  c->present = 1; c->old_style = old_style; c->rank = rank;
  long long size = 1;
  for (long long q = 0; q < rank; q++) { c->dims[q] = dims[q]; size *= dims[q]; c->params[q] = dup(params[q], dims[q]); }
  c->cooling = dup(cooling, size); c->heating = dup(heating, size); c->mmw = dup(mmw, size);
}

void oracle_inject_uvb(long long nz, const double *const *vecs)
{
  g_uvb.nz = nz;
  for (int i = 0; i < 15; i++) g_uvb.v[i] = dup(vecs[i], nz);
}

void oracle_clear_injected(void) { memset(g_inj, 0, sizeof(g_inj)); memset(&g_uvb, 0, sizeof(g_uvb)); }

/* upstream-signature loaders --------------------------------------------- */
int initialize_cloudy_data(chemistry_data *my_chemistry, chemistry_data_storage *my_rates,
                           cloudy_data *my_cloudy, char *group_name, code_units *my_units,
                           int read_data)
{
  long long q, w;
  my_cloudy->grid_parameters = malloc(CLOUDY_MAX_DIMENSION * sizeof(double *));
  my_cloudy->grid_dimension = malloc(CLOUDY_MAX_DIMENSION * sizeof(long long));
  for (q = 0; q < CLOUDY_MAX_DIMENSION; q++) { my_cloudy->grid_dimension[q] = 0; my_cloudy->grid_parameters[q] = NULL; }
  my_cloudy->cooling_data = my_cloudy->heating_data = my_cloudy->mmw_data = NULL;
  if (read_data == 0) { my_cloudy->grid_rank = 0; return SUCCESS; }

  const injected_cloudy *c = &g_inj[strcmp(group_name, "Primordial") == 0 ? 0 : 1];
  if (!c->present) { fprintf(stderr, "oracle: no injected Cloudy data for %s.\n", group_name); return FAIL; }

  double co_length_units, co_density_units;
  if (my_units->comoving_coordinates == TRUE) {
    co_length_units = my_units->length_units; co_density_units = my_units->density_units;
  } else {
    co_length_units = my_units->length_units * my_units->a_value * my_units->a_units;
    co_density_units = my_units->density_units / POW(my_units->a_value * my_units->a_units, 3);
  }
  double tbase1 = my_units->time_units;
  double xbase1 = co_length_units / (my_units->a_value * my_units->a_units);
  double dbase1 = co_density_units * POW(my_units->a_value * my_units->a_units, 3);
  double mh = 1.67e-24;
  double CoolUnit = (POW(my_units->a_units, 5) * POW(xbase1, 2) * POW(mh, 2)) / (POW(tbase1, 3) * dbase1);

  if (c->old_style) my_rates->cloudy_data_new = 0;
  my_cloudy->grid_rank = c->rank;
  my_cloudy->data_size = 1;
  for (q = 0; q < c->rank; q++) {
    my_cloudy->grid_dimension[q] = c->dims[q];
    my_cloudy->data_size *= c->dims[q];
    my_cloudy->grid_parameters[q] = malloc(c->dims[q] * sizeof(double));
    for (w = 0; w < c->dims[q]; w++)
      my_cloudy->grid_parameters[q][w] = (q < c->rank - 1) ? c->params[q][w] : log10(c->params[q][w]);
  }
  my_cloudy->cooling_data = malloc(my_cloudy->data_size * sizeof(double));
  for (q = 0; q < my_cloudy->data_size; q++) {
    my_cloudy->cooling_data[q] = c->cooling[q] > 0 ? log10(c->cooling[q]) : SMALL_LOG_VALUE;
    my_cloudy->cooling_data[q] -= log10(CoolUnit);
  }
  if (my_chemistry->UVbackground == 1) {
    if (!c->heating) { fprintf(stderr, "oracle: no Heating dataset injected.\n"); return FAIL; }
    my_cloudy->heating_data = malloc(my_cloudy->data_size * sizeof(double));
    for (q = 0; q < my_cloudy->data_size; q++) {
      my_cloudy->heating_data[q] = c->heating[q] > 0 ? log10(c->heating[q]) : SMALL_LOG_VALUE;
      my_cloudy->heating_data[q] -= log10(CoolUnit);
    }
  }
  if (my_chemistry->primordial_chemistry == 0 && strcmp(group_name, "Primordial") == 0) {
    if (!c->mmw) { fprintf(stderr, "oracle: no MMW dataset injected.\n"); return FAIL; }
    my_cloudy->mmw_data = dup(c->mmw, my_cloudy->data_size);
  }
  return SUCCESS;
}

int _free_cloudy_data(cloudy_data *my_cloudy, chemistry_data *my_chemistry, int primordial)
{
  (void)my_chemistry; (void)primordial;
  if (my_cloudy->grid_parameters)
    for (int i = 0; i < my_cloudy->grid_rank; i++) free(my_cloudy->grid_parameters[i]);
  free(my_cloudy->grid_parameters); free(my_cloudy->grid_dimension);
  free(my_cloudy->cooling_data); free(my_cloudy->heating_data); free(my_cloudy->mmw_data);
  memset(my_cloudy, 0, sizeof(*my_cloudy));
  return SUCCESS;
}

int initialize_UVbackground_data(chemistry_data *my_chemistry, chemistry_data_storage *my_rates)
{
  if (my_chemistry->UVbackground == 0 || my_chemistry->primordial_chemistry == 0) return SUCCESS;
  if (g_uvb.nz <= 0) { fprintf(stderr, "oracle: no injected UVB table.\n"); return FAIL; }
  UVBtable *t = &my_rates->UVbackground_table;
  long long Nz = g_uvb.nz;
  t->Nz = Nz;
  t->z = dup(g_uvb.v[0], Nz);
  t->k24 = dup(g_uvb.v[1], Nz); t->k25 = dup(g_uvb.v[2], Nz); t->k26 = dup(g_uvb.v[3], Nz);
  if (my_chemistry->primordial_chemistry > 1) {
    t->k27 = dup(g_uvb.v[4], Nz); t->k28 = dup(g_uvb.v[5], Nz); t->k29 = dup(g_uvb.v[6], Nz);
    t->k30 = dup(g_uvb.v[7], Nz); t->k31 = dup(g_uvb.v[8], Nz);
  }
  t->piHI = dup(g_uvb.v[9], Nz); t->piHeI = dup(g_uvb.v[10], Nz); t->piHeII = dup(g_uvb.v[11], Nz);
  if (my_chemistry->self_shielding_method > 0) {
    t->crsHI = dup(g_uvb.v[12], Nz); t->crsHeI = dup(g_uvb.v[13], Nz); t->crsHeII = dup(g_uvb.v[14], Nz);
  }
  t->zmin = t->z[0];
  t->zmax = t->z[Nz - 1];
  if (my_chemistry->UVbackground_redshift_on <= FLOAT_UNDEFINED) my_chemistry->UVbackground_redshift_on = t->zmax;
  if (my_chemistry->UVbackground_redshift_fullon <= FLOAT_UNDEFINED) my_chemistry->UVbackground_redshift_fullon = t->zmax;
  if (my_chemistry->UVbackground_redshift_drop <= FLOAT_UNDEFINED) my_chemistry->UVbackground_redshift_drop = t->zmin;
  if (my_chemistry->UVbackground_redshift_off <= FLOAT_UNDEFINED) my_chemistry->UVbackground_redshift_off = t->zmin;
  return SUCCESS;
}
