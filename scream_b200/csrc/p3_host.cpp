// p3_host.cpp — lookup-table I/O + generation and runtime defaults (host, no CUDA).
#include "p3_host.hpp"

#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

namespace p3b200 {

void runtime_defaults(P3B200Runtime& r) {
  // p3_functions.hpp:114-133
  r.max_total_ni = 740.0e3;
  r.autoconversion_prefactor = 1350.0;
  r.autoconversion_qc_exponent = 2.47;
  r.autoconversion_nc_exponent = 1.79;
  r.autoconversion_radius = 25.0e-6;
  r.accretion_prefactor = 67.0;
  r.accretion_qc_exponent = 1.15;
  r.accretion_qr_exponent = 1.15;
  r.rain_selfcollection_prefactor = 5.78;
  r.rain_selfcollection_breakup_diameter = 0.00028;
  r.constant_mu_rain = 1.0;
  r.spa_ccn_to_nc_factor = 1.0;
  r.cldliq_to_ice_collection_factor = 0.5;
  r.rain_to_ice_collection_factor = 1.0;
  r.min_rime_rho = 50.0;
  r.max_rime_rho = 900.0;
  r.immersion_freezing_exponent = 0.65;
  r.deposition_nucleation_exponent = 0.304;
  r.ice_sedimentation_factor = 1.0;
  r.do_ice_production = 1;
  r.pad_ = 0;
}

// The text file is a whitespace-separated token stream: "VERSION x.y.z" then, for each of the
// 5 x 4 (density, rime-fraction) blocks, 50 ice rows (2 ints + 15 reals; reals 2..14 except 10 are
// kept -> 12 columns) followed by 50 x 30 collection rows (2 ints + 6 reals; reals 3 and 4 kept).
int tables_read_text(const char* path, const char* expect_version, P3B200Tables& out) {
  if (!path) return P3B200_EINVAL;
  std::ifstream in(path);
  if (!in) return P3B200_ETABLE;
  std::string tag, ver;
  in >> tag >> ver;
  if (tag != "VERSION") return P3B200_ETABLE;
  if (expect_version && ver != expect_version) return P3B200_ETABLE;
  for (int jj = 0; jj < P3B200_DENSIZE; ++jj)
    for (int ii = 0; ii < P3B200_RIMSIZE; ++ii) {
      for (int i = 0; i < P3B200_ISIZE; ++i) {
        int skip;
        in >> skip >> skip;
        int col = 0;
        for (int j = 0; j < 15; ++j) {
          double v;
          in >> v;
          if (j > 1 && j != 10) out.ice_table_vals[jj][ii][i][col++] = v;
        }
      }
      for (int i = 0; i < P3B200_ISIZE; ++i)
        for (int j = 0; j < P3B200_RCOLLSIZE; ++j) {
          int skip;
          in >> skip >> skip;
          int col = 0;
          for (int k = 0; k < 6; ++k) {
            double v;
            in >> v;
            if (k == 3 || k == 4) out.collect_table_vals[jj][ii][i][j][col++] = std::log10(v);
          }
        }
    }
  return in ? P3B200_OK : P3B200_ETABLE;
}

int tables_read_blob(const char* path, const char* expect_version, P3B200Tables& out) {
  if (!path) return P3B200_EINVAL;
  FILE* f = std::fopen(path, "rb");
  if (!f) return P3B200_ETABLE;
  char magic[4];
  uint32_t n = 0;
  int rc = P3B200_ETABLE;
  const size_t n_ice = sizeof(out.ice_table_vals) / sizeof(double);
  const size_t n_coll = sizeof(out.collect_table_vals) / sizeof(double);
  if (std::fread(magic, 1, 4, f) == 4 && std::memcmp(magic, "P3TB", 4) == 0 && std::fread(&n, 4, 1, f) == 1 && n < 64) {
    char ver[64] = {0};
    if (std::fread(ver, 1, n, f) == n && (!expect_version || std::string(ver) == expect_version)) {
      std::vector<double> coll(n_coll);
      if (std::fread(&out.ice_table_vals[0][0][0][0], sizeof(double), n_ice, f) == n_ice &&
          std::fread(coll.data(), sizeof(double), n_coll, f) == n_coll) {
        double* dst = &out.collect_table_vals[0][0][0][0][0];
        for (size_t i = 0; i < n_coll; ++i) dst[i] = std::log10(coll[i]);   // p3_table_ice_impl.hpp:63
        rc = P3B200_OK;
      }
    }
  }
  std::fclose(f);
  return rc;
}

// Rain fall-speed / ventilation tables by quadrature over the gamma PSD, statement order of the
// Fortran (micro_p3.F90:275-338; all literals are double there: components/eamxx/CMakeLists.txt:554-558).
// mu_r is the constant mu_r_constant = 1 (micro_p3_utils.F90:33) for every one of the 10 table columns, so
// one column is integrated and replicated.
void tables_generate_rain(P3B200Tables& out) {
  const double pi = 3.14159265358979323;
  const double thrd = 1.0 / 3.0, sxth = 1.0 / 6.0, piov6 = pi * sxth;   // micro_p3_utils.F90:105-109
  const double mu_r = 1.0;
  for (int i = 0; i < P3B200_MU_R_TABLE_DIM; ++i) out.mu_r_table_vals[i] = mu_r;
  const double dd = 2.0;
  for (int jj = 1; jj <= P3B200_VTABLE_DIM0; ++jj) {
    const double dm = (jj <= 20) ? (double(jj) * 10. - 5.) * 1.e-6 : (double(jj - 20) * 30. + 195.) * 1.e-6;
    const double lamr = (mu_r + 1.) / dm;
    double s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0;
    for (int kk = 1; kk <= 10000; ++kk) {
      const double dia = (double(kk) * dd - dd * 0.5) * 1.e-6;
      double amg = piov6 * 997. * (dia * dia * dia);
      amg = amg * 1000.;
      const double dmic = dia * 1.e+6;
      double vt;
      if (dmic <= 134.43)      vt = 4.5795e+3 * std::pow(amg, 2. * thrd);
      else if (dmic < 1511.64) vt = 4.962e+1 * std::pow(amg, thrd);
      else if (dmic < 3477.84) vt = 1.732e+1 * std::pow(amg, sxth);
      else                     vt = 9.17;
      const double ex = std::exp(-lamr * dia);
      const double l10 = std::log10(dia);
      s1 = s1 + vt * std::pow(10., mu_r * l10 + 4. * mu_r) * ex * dd * 1.e-6;
      s2 = s2 + std::pow(10., mu_r * l10 + 4. * mu_r) * ex * dd * 1.e-6;
      s3 = s3 + vt * std::pow(10., (mu_r + 3.) * l10 + 4. * mu_r) * ex * dd * 1.e-6;
      s4 = s4 + std::pow(10., (mu_r + 3.) * l10 + 4. * mu_r) * ex * dd * 1.e-6;
      s5 = s5 + std::pow(vt * dia, 0.5) * std::pow(10., (mu_r + 1.) * l10 + 3. * mu_r) * ex * dd * 1.e-6;
    }
    s2 = std::max(s2, 1.e-30);
    s4 = std::max(s4, 1.e-30);
    s5 = std::max(s5, 1.e-30);
    const double vn = s1 / s2, vm = s3 / s4;
    const double rv = std::pow(10., std::log10(s5) + (mu_r + 1.) * std::log10(lamr) - (3. * mu_r));
    for (int c = 0; c < P3B200_VTABLE_DIM1; ++c) {
      out.vn_table_vals[jj - 1][c] = vn;
      out.vm_table_vals[jj - 1][c] = vm;
      out.revap_table_vals[jj - 1][c] = rv;
    }
  }
  // p3_table3_impl.hpp:120-135
  static const double dnu[P3B200_DNUSIZE] = {0.000, -0.557, -0.430, -0.307, -0.186, -0.067, -0.050, -0.167,
                                            -0.282, -0.397, -0.512, -0.626, -0.739, -0.853, -0.966, -0.966};
  for (int i = 0; i < P3B200_DNUSIZE; ++i) out.dnu_table_vals[i] = dnu[i];
}

}  // namespace p3b200
