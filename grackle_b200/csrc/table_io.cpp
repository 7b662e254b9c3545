// grackle-b200: load the Cloudy cooling grids and the UV-background z-tables from a
// Grackle data file into chemistry_data_storage (host copies).  Semantics follow
// initialize_cloudy_data.c:29-316 and initialize_UVbackground_data.c:30-341 of the
// reference; the file is parsed by our own h5lite reader instead of libhdf5.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "gr_internal.h"
#include "h5lite.hpp"

namespace {
constexpr int kCloudyMaxDim = 5;          // CLOUDY_MAX_DIMENSION, initialize_cloudy_data.c:21
constexpr double kSmallLog = -99.0;       // SMALL_LOG_VALUE, :22

double *dup_array(const std::vector<double> &v)
{
  double *p = static_cast<double *>(std::malloc(sizeof(double) * (v.empty() ? 1 : v.size())));
  if (p && !v.empty()) std::memcpy(p, v.data(), sizeof(double) * v.size());
  return p;
}

bool open_file(const chemistry_data *chem, h5lite::File *f)
{
  std::string err;
  if (!chem->grackle_data_file || !f->open(chem->grackle_data_file, &err)) {
    std::fprintf(stderr, "grackle-b200: cannot read data file '%s': %s\n",
                 chem->grackle_data_file ? chem->grackle_data_file : "(null)", err.c_str());
    return false;
  }
  return true;
}
}  // namespace

int gr_load_cloudy_table(chemistry_data *chem, chemistry_data_storage *rates, cloudy_data *cl,
                         const char *group, const code_units *units, int read_data)
{
  cl->grid_parameters = static_cast<double **>(std::calloc(kCloudyMaxDim, sizeof(double *)));
  cl->grid_dimension = static_cast<long long *>(std::calloc(kCloudyMaxDim, sizeof(long long)));
  cl->grid_rank = 0;
  cl->data_size = 0;
  cl->cooling_data = cl->heating_data = cl->mmw_data = NULL;
  if (!read_data) return GR_SUCCESS;

  if (grackle_verbose)
    std::fprintf(stdout, "Initializing Cloudy cooling: %s.\ncloudy_table_file: %s.\n", group,
                 chem->grackle_data_file);

  // NB the reference converts the tables with mh = 1.67e-24 here
  double co_len, co_den;
  gr_comoving_units(units, &co_len, &co_den);
  const double aprod = units->a_value * units->a_units;
  const double tbase1 = units->time_units;
  const double xbase1 = co_len / aprod;
  const double dbase1 = co_den * std::pow(aprod, 3.0);
  const double mh = 1.67e-24;
  const double cool_unit = (std::pow(units->a_units, 5) * std::pow(xbase1, 2) * std::pow(mh, 2)) /
                           (std::pow(tbase1, 3) * dbase1);

  h5lite::File file;
  if (!open_file(chem, &file)) return GR_FAIL;

  for (const std::string &a : file.root().attr_names)
    if (a == "old_style") rates->cloudy_data_new = 0;

  const std::string base = std::string("/CoolingRates/") + group;
  const h5lite::Node *cool = file.root().find(base + "/Cooling");
  if (!cool || !cool->is_dataset || !cool->data.ok) {
    std::fprintf(stderr, "Can't open Cooling in %s.\n", chem->grackle_data_file);
    return GR_FAIL;
  }
  auto rank_it = cool->attrs.find("Rank");
  auto dim_it = cool->attrs.find("Dimension");
  if (rank_it == cool->attrs.end() || dim_it == cool->attrs.end() || !rank_it->second.is_int) {
    std::fprintf(stderr, "Failed to read Rank/Dimension attributes of Cooling dataset.\n");
    return GR_FAIL;
  }
  cl->grid_rank = rank_it->second.i64[0];
  if (cl->grid_rank > kCloudyMaxDim || cl->grid_rank < 1 ||
      (long long)dim_it->second.i64.size() < cl->grid_rank) {
    std::fprintf(stderr, "Error: rank of Cloudy cooling data must be between 1 and %d.\n", kCloudyMaxDim);
    return GR_FAIL;
  }
  cl->data_size = 1;
  for (long long q = 0; q < cl->grid_rank; ++q) {
    cl->grid_dimension[q] = dim_it->second.i64[q];
    cl->data_size *= cl->grid_dimension[q];
  }

  // axes: Parameter1..Parameter(rank-1) verbatim, last axis = log10(Temperature)
  for (long long q = 0; q < cl->grid_rank; ++q) {
    const bool is_temperature = (q == cl->grid_rank - 1);
    const std::string pname = is_temperature ? "Temperature" : "Parameter" + std::to_string(q + 1);
    auto it = cool->attrs.find(pname);
    if (it == cool->attrs.end() || (long long)it->second.f64.size() != cl->grid_dimension[q]) {
      std::fprintf(stderr, "Failed to read %s attribute in Cooling dataset.\n", pname.c_str());
      return GR_FAIL;
    }
    cl->grid_parameters[q] = dup_array(it->second.f64);
    if (is_temperature)
      for (long long w = 0; w < cl->grid_dimension[q]; ++w)
        cl->grid_parameters[q][w] = std::log10(cl->grid_parameters[q][w]);
    if (grackle_verbose)
      std::fprintf(stdout, "%s: %g to %g (%lld steps).\n", pname.c_str(), cl->grid_parameters[q][0],
                   cl->grid_parameters[q][cl->grid_dimension[q] - 1], cl->grid_dimension[q]);
  }

  // log10 of the rates, shifted to code units
  auto to_log_code = [&](const std::vector<double> &raw) {
    double *out = static_cast<double *>(std::malloc(sizeof(double) * cl->data_size));
    for (long long q = 0; q < cl->data_size; ++q) {
      out[q] = raw[q] > 0 ? std::log10(raw[q]) : kSmallLog;
      out[q] -= std::log10(cool_unit);
    }
    return out;
  };
  if ((long long)cool->data.f64.size() != cl->data_size) {
    std::fprintf(stderr, "Cooling dataset size does not match its Dimension attribute.\n");
    return GR_FAIL;
  }
  cl->cooling_data = to_log_code(cool->data.f64);

  if (chem->UVbackground == 1) {
    const h5lite::Node *heat = file.root().find(base + "/Heating");
    if (!heat || !heat->data.ok || (long long)heat->data.f64.size() != cl->data_size) {
      std::fprintf(stderr, "Can't open Heating in %s.\n", chem->grackle_data_file);
      return GR_FAIL;
    }
    cl->heating_data = to_log_code(heat->data.f64);
  }

  if (chem->primordial_chemistry == 0 && std::strcmp(group, "Primordial") == 0) {
    const h5lite::Node *mmw = file.root().find(base + "/MMW");
    if (!mmw || !mmw->data.ok || (long long)mmw->data.f64.size() != cl->data_size) {
      std::fprintf(stderr, "Can't open MMW in %s.\n", chem->grackle_data_file);
      return GR_FAIL;
    }
    cl->mmw_data = dup_array(mmw->data.f64);
  }
  return GR_SUCCESS;
}

void gr_free_cloudy_table(cloudy_data *cl, int, int)
{
  if (cl->grid_parameters)
    for (int q = 0; q < kCloudyMaxDim; ++q) std::free(cl->grid_parameters[q]);
  std::free(cl->grid_parameters);
  std::free(cl->grid_dimension);
  std::free(cl->cooling_data);
  std::free(cl->heating_data);
  std::free(cl->mmw_data);
  std::memset(cl, 0, sizeof(*cl));
}

int gr_load_uvb_table(chemistry_data *chem, chemistry_data_storage *rates)
{
  UVBtable &t = rates->UVbackground_table;
  std::memset(&t, 0, sizeof(t));
  if (chem->UVbackground == 0 || chem->primordial_chemistry == 0) return GR_SUCCESS;

  if (grackle_verbose)
    std::fprintf(stdout, "Initializing UV background.\nReading UV background data from %s.\n",
                 chem->grackle_data_file);
  h5lite::File file;
  if (!open_file(chem, &file)) return GR_FAIL;

  if (!file.root().find("/UVBRates/Info")) {
    std::fprintf(stderr, "Can't open 'UVBRates/Info' in %s.\n", chem->grackle_data_file);
    return GR_FAIL;
  }
  const h5lite::Node *zn = file.root().find("/UVBRates/z");
  if (!zn || !zn->data.ok || zn->data.f64.empty()) {
    std::fprintf(stderr, "Can't open redshift dataset ('UVBRates/z') in %s.\n", chem->grackle_data_file);
    return GR_FAIL;
  }
  const long long nz = (long long)zn->data.f64.size();
  t.Nz = nz;
  t.z = dup_array(zn->data.f64);

  bool ok = true;
  auto take = [&](const char *path, double **dst) {
    const h5lite::Node *n = file.root().find(path);
    if (!n || !n->data.ok || (long long)n->data.f64.size() != nz) {
      std::fprintf(stderr, "Error reading dataset '%s' in %s.\n", path, chem->grackle_data_file);
      ok = false;
      return;
    }
    *dst = dup_array(n->data.f64);
  };
  take("/UVBRates/Chemistry/k24", &t.k24);
  take("/UVBRates/Chemistry/k25", &t.k25);
  take("/UVBRates/Chemistry/k26", &t.k26);
  if (chem->primordial_chemistry > 1) {
    take("/UVBRates/Chemistry/k27", &t.k27);
    take("/UVBRates/Chemistry/k28", &t.k28);
    take("/UVBRates/Chemistry/k29", &t.k29);
    take("/UVBRates/Chemistry/k30", &t.k30);
    take("/UVBRates/Chemistry/k31", &t.k31);
  }
  take("/UVBRates/Photoheating/piHI", &t.piHI);
  take("/UVBRates/Photoheating/piHeII", &t.piHeII);
  take("/UVBRates/Photoheating/piHeI", &t.piHeI);
  if (chem->self_shielding_method > 0) {
    take("/UVBRates/CrossSections/hi_avg_crs", &t.crsHI);
    take("/UVBRates/CrossSections/heii_avg_crs", &t.crsHeII);
    take("/UVBRates/CrossSections/hei_avg_crs", &t.crsHeI);
  }
  if (!ok) return GR_FAIL;

  t.zmin = t.z[0];
  t.zmax = t.z[nz - 1];
  if (grackle_verbose)
    std::fprintf(stdout, "UV background information:\n  z_min = %6.3f\n  z_max = %6.3f\n", t.zmin, t.zmax);

  // unset on/off redshifts default to the table's range
  if (chem->UVbackground_redshift_on <= GR_FLOAT_UNDEFINED) chem->UVbackground_redshift_on = t.zmax;
  if (chem->UVbackground_redshift_fullon <= GR_FLOAT_UNDEFINED) chem->UVbackground_redshift_fullon = t.zmax;
  if (chem->UVbackground_redshift_drop <= GR_FLOAT_UNDEFINED) chem->UVbackground_redshift_drop = t.zmin;
  if (chem->UVbackground_redshift_off <= GR_FLOAT_UNDEFINED) chem->UVbackground_redshift_off = t.zmin;
  return GR_SUCCESS;
}
