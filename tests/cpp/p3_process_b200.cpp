// p3_process_b200 — a C++ caller of the process-level face scream::P3Microphysics (include/p3_b200_process.hpp), the
// way the atmosphere driver uses the reference's P3Microphysics: construct from parameters, ask for the buffer size,
// hand over device memory (init_buffers), bind the fields (initialize_impl), step run_impl(dt) N times, finalize.
//
// usage: p3_process_b200 <table blob|text> <in.bin> <out.bin> ncol nlev dt nsteps predictNc prescribedCCN fluxes
// in.bin: the field-manager fields in the order of NAMES2D / NAMES1D below, raw doubles, (ncol, nlev) row-major;
// out.bin: the same fields after the steps, followed by the 15 property-check results of the last step
// (result, min, max per check).  tests/test_gpu_process.py compares with the oracle's preamble -> p3_main ->
// postamble sequence.
#include "p3_b200_process.hpp"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

using Real = double;
static const char* NAMES2D[] = {"p_mid", "p_dry_mid", "pseudo_density", "pseudo_density_dry", "cldfrac_tot", "nc_nuceat_tend",
                                "nccn", "ni_activated", "inv_qc_relvar", "T_mid", "qv", "qc", "qr", "qi", "qm", "nc", "nr", "ni",
                                "bm", "qv_prev_micro_step", "T_prev_micro_step", "eff_radius_qc", "eff_radius_qi",
                                "eff_radius_qr", "precip_total_tend", "nevapr", "rainfrac", "micro_liq_ice_exchange",
                                "micro_vap_liq_exchange", "micro_vap_ice_exchange"};
static const char* NAMES1D[] = {"precip_liq_surf_mass", "precip_ice_surf_mass", "vapor_flux", "water_flux", "ice_flux", "heat_flux"};
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 2; } } while (0)

int main(int argc, char** argv) {
  if (argc != 11) { std::fprintf(stderr, "usage: %s table in out ncol nlev dt nsteps predictNc prescribedCCN fluxes\n", argv[0]); return 1; }
  const int ncol = std::atoi(argv[4]), nlev = std::atoi(argv[5]), nsteps = std::atoi(argv[7]);
  const double dt = std::atof(argv[6]);
  try {
    scream::P3Microphysics::Params params;
    params.do_predict_nc = std::atoi(argv[8]) != 0;
    params.do_prescribed_ccn = std::atoi(argv[9]) != 0;
    params.enable_column_conservation_checks = std::atoi(argv[10]) != 0;
    params.p3_lookup_table = argv[1];
    scream::P3Microphysics p3(ncol, nlev, params);
    if (p3.name() != "p3" || p3.type() != scream::AtmosphereProcessType::Physics) return 4;

    std::map<std::string, Real*> dev;
    std::map<std::string, size_t> count;
    std::FILE* in = std::fopen(argv[2], "rb");
    if (!in) { std::perror(argv[2]); return 1; }
    auto load = [&](const char* name, size_t n) -> int {
      std::vector<Real> h(n);
      if (std::fread(h.data(), sizeof(Real), n, in) != n) { std::fprintf(stderr, "short read: %s\n", name); return 1; }
      Real* d = nullptr;
      CK(cudaMalloc(&d, n * sizeof(Real)));
      CK(cudaMemcpy(d, h.data(), n * sizeof(Real), cudaMemcpyHostToDevice));
      dev[name] = d; count[name] = n;
      return 0;
    };
    for (const char* n : NAMES2D) if (int rc = load(n, size_t(ncol) * nlev)) return rc;
    for (const char* n : NAMES1D) if (int rc = load(n, size_t(ncol))) return rc;
    std::fclose(in);

    void* buffer = nullptr;
    const size_t bytes = p3.requested_buffer_size_in_bytes();
    CK(cudaMalloc(&buffer, bytes));
    CK(cudaMemset(buffer, 0xff, bytes));     // the buffer manager hands out uninitialised memory
    p3.init_buffers(buffer, bytes);

    scream::P3Microphysics::Fields f;
    f.p_mid = dev["p_mid"]; f.p_dry_mid = dev["p_dry_mid"]; f.pseudo_density = dev["pseudo_density"];
    f.pseudo_density_dry = dev["pseudo_density_dry"]; f.cldfrac_tot = dev["cldfrac_tot"];
    f.nc_nuceat_tend = dev["nc_nuceat_tend"]; f.nccn = dev["nccn"]; f.ni_activated = dev["ni_activated"];
    f.inv_qc_relvar = dev["inv_qc_relvar"]; f.T_mid = dev["T_mid"]; f.qv = dev["qv"]; f.qc = dev["qc"]; f.qr = dev["qr"];
    f.qi = dev["qi"]; f.qm = dev["qm"]; f.nc = dev["nc"]; f.nr = dev["nr"]; f.ni = dev["ni"]; f.bm = dev["bm"];
    f.qv_prev_micro_step = dev["qv_prev_micro_step"]; f.T_prev_micro_step = dev["T_prev_micro_step"];
    f.precip_liq_surf_mass = dev["precip_liq_surf_mass"]; f.precip_ice_surf_mass = dev["precip_ice_surf_mass"];
    f.eff_radius_qc = dev["eff_radius_qc"]; f.eff_radius_qi = dev["eff_radius_qi"]; f.eff_radius_qr = dev["eff_radius_qr"];
    f.precip_total_tend = dev["precip_total_tend"]; f.nevapr = dev["nevapr"]; f.rainfrac = dev["rainfrac"];
    f.micro_liq_ice_exchange = dev["micro_liq_ice_exchange"]; f.micro_vap_liq_exchange = dev["micro_vap_liq_exchange"];
    f.micro_vap_ice_exchange = dev["micro_vap_ice_exchange"];
    if (params.enable_column_conservation_checks) {
      f.vapor_flux = dev["vapor_flux"]; f.water_flux = dev["water_flux"]; f.ice_flux = dev["ice_flux"]; f.heat_flux = dev["heat_flux"];
    }
    p3.initialize_impl(f, scream::RunType::Initial);
    p3.fail_on_check = false;
    long us = 0;
    for (int s = 0; s < nsteps; ++s) { p3.run_impl(dt); us += p3.last_p3_main_microseconds(); }

    std::FILE* out = std::fopen(argv[3], "wb");
    if (!out) { std::perror(argv[3]); return 1; }
    auto dump = [&](const char* name) -> int {
      std::vector<Real> h(count[name]);
      CK(cudaMemcpy(h.data(), dev[name], count[name] * sizeof(Real), cudaMemcpyDeviceToHost));
      std::fwrite(h.data(), sizeof(Real), count[name], out);
      return 0;
    };
    for (const char* n : NAMES2D) if (int rc = dump(n)) return rc;
    for (const char* n : NAMES1D) if (int rc = dump(n)) return rc;
    for (const auto& c : p3.property_check_results()) {
      const double rec[3] = {double(c.result), c.min_val, c.max_val};
      std::fwrite(rec, sizeof(double), 3, out);
    }
    std::fclose(out);
    std::printf("p3_process_b200: %d columns x %d levels, %d x run_impl, it = %d, %ld us in p3_main, qv repaired: %d\n", ncol, nlev,
                nsteps, p3.infrastructure.it, us, int(p3.qv_was_repaired()));
    p3.finalize_impl();
    for (auto& kv : dev) cudaFree(kv.second);
    cudaFree(buffer);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "p3_process_b200: %s\n", e.what());
    return 3;
  }
  return 0;
}
