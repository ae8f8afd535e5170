// p3_run_and_cmp_b200 — a complete C++ caller of p3::Functions<Real,B200Device>::p3_main, the analogue of
// the reference's p3/tests/p3_run_and_cmp.cpp:218-352 (initial condition -> N steps of p3_main -> dump of
// every output field).  It is written against the reference-shaped face (include/p3_b200_functions.hpp)
// exactly the way P3Microphysics::run_impl drives it (eamxx_p3_run.cpp:18-36): bump `it`, reset the
// workspace manager, zero the three history arrays, call p3_main.
//
// usage: p3_run_and_cmp_b200 <table blob|text> <in.bin> <out.bin> ncol nlev dt nsteps predictNc prescribedCCN
// in.bin / out.bin: the 37 fields of include/p3_b200.h in enum order, raw doubles, (ncol, levels) row-major.
// The comparison with the oracle is done by tests/test_gpu_cpp_driver.py.
#include "p3_b200_functions.hpp"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

using Real = double;
using P3F = scream::p3::Functions<Real, scream::B200Device>;

static int levels_of(int f, int nlev) {
  if (f >= P3B200_F_PRECIP_LIQ_SURF) return 1;
  return (f == P3B200_F_PRECIP_LIQ_FLUX || f == P3B200_F_PRECIP_ICE_FLUX) ? nlev + 1 : nlev;
}
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 2; } } while (0)

int main(int argc, char** argv) {
  if (argc != 10) { std::fprintf(stderr, "usage: %s table in out ncol nlev dt nsteps predictNc prescribedCCN\n", argv[0]); return 1; }
  const int ncol = std::atoi(argv[4]), nlev = std::atoi(argv[5]), nsteps = std::atoi(argv[7]);
  const double dt = std::atof(argv[6]);
  const bool predictNc = std::atoi(argv[8]) != 0, prescribedCCN = std::atoi(argv[9]) != 0;
  try {
    P3F::P3LookupTables lookup_tables = P3F::p3_init(argv[1], 0);
    std::vector<Real*> dev(P3B200_F_NUM, nullptr);
    std::vector<size_t> count(P3B200_F_NUM);
    std::FILE* in = std::fopen(argv[2], "rb");
    if (!in) { std::perror(argv[2]); return 1; }
    for (int f = 0; f < P3B200_F_NUM; ++f) {
      count[f] = size_t(ncol) * levels_of(f, nlev);
      std::vector<Real> h(count[f]);
      if (std::fread(h.data(), sizeof(Real), count[f], in) != count[f]) { std::fprintf(stderr, "short read, field %d\n", f); return 1; }
      CK(cudaMalloc(&dev[f], count[f] * sizeof(Real)));
      CK(cudaMemcpy(dev[f], h.data(), count[f] * sizeof(Real), cudaMemcpyHostToDevice));
    }
    std::fclose(in);

    auto v2 = [&](int f) { return P3F::view_2d<Real>(dev[f], ncol, levels_of(f, nlev)); };
    P3F::P3PrognosticState prog_state;
    prog_state.qc = v2(P3B200_F_QC); prog_state.nc = v2(P3B200_F_NC); prog_state.qr = v2(P3B200_F_QR);
    prog_state.nr = v2(P3B200_F_NR); prog_state.qi = v2(P3B200_F_QI); prog_state.qm = v2(P3B200_F_QM);
    prog_state.ni = v2(P3B200_F_NI); prog_state.bm = v2(P3B200_F_BM); prog_state.qv = v2(P3B200_F_QV);
    prog_state.th = v2(P3B200_F_TH);
    P3F::P3DiagnosticInputs diag_inputs;
    diag_inputs.nc_nuceat_tend = v2(P3B200_F_NC_NUCEAT_TEND); diag_inputs.nccn = v2(P3B200_F_NCCN);
    diag_inputs.ni_activated = v2(P3B200_F_NI_ACTIVATED); diag_inputs.inv_qc_relvar = v2(P3B200_F_INV_QC_RELVAR);
    diag_inputs.cld_frac_i = v2(P3B200_F_CLD_FRAC_I); diag_inputs.cld_frac_l = v2(P3B200_F_CLD_FRAC_L);
    diag_inputs.cld_frac_r = v2(P3B200_F_CLD_FRAC_R); diag_inputs.pres = v2(P3B200_F_PRES);
    diag_inputs.dz = v2(P3B200_F_DZ); diag_inputs.dpres = v2(P3B200_F_DPRES);
    diag_inputs.inv_exner = v2(P3B200_F_INV_EXNER); diag_inputs.qv_prev = v2(P3B200_F_QV_PREV);
    diag_inputs.t_prev = v2(P3B200_F_T_PREV);
    P3F::P3DiagnosticOutputs diag_outputs;
    diag_outputs.qv2qi_depos_tend = v2(P3B200_F_QV2QI_DEPOS_TEND);
    diag_outputs.precip_liq_surf = P3F::view_1d<Real>(dev[P3B200_F_PRECIP_LIQ_SURF], ncol);
    diag_outputs.precip_ice_surf = P3F::view_1d<Real>(dev[P3B200_F_PRECIP_ICE_SURF], ncol);
    diag_outputs.diag_eff_radius_qc = v2(P3B200_F_DIAG_EFF_RADIUS_QC);
    diag_outputs.diag_eff_radius_qi = v2(P3B200_F_DIAG_EFF_RADIUS_QI);
    diag_outputs.diag_eff_radius_qr = v2(P3B200_F_DIAG_EFF_RADIUS_QR);
    diag_outputs.rho_qi = v2(P3B200_F_RHO_QI);
    diag_outputs.precip_liq_flux = v2(P3B200_F_PRECIP_LIQ_FLUX); diag_outputs.precip_ice_flux = v2(P3B200_F_PRECIP_ICE_FLUX);
    diag_outputs.precip_total_tend = v2(P3B200_F_PRECIP_TOTAL_TEND); diag_outputs.nevapr = v2(P3B200_F_NEVAPR);
    P3F::P3HistoryOnly history_only;
    history_only.liq_ice_exchange = v2(P3B200_F_LIQ_ICE_EXCHANGE);
    history_only.vap_liq_exchange = v2(P3B200_F_VAP_LIQ_EXCHANGE);
    history_only.vap_ice_exchange = v2(P3B200_F_VAP_ICE_EXCHANGE);
    P3F::P3Infrastructure infrastructure;
    infrastructure.dt = dt; infrastructure.it = 0; infrastructure.its = 0; infrastructure.ite = ncol - 1;
    infrastructure.kts = 0; infrastructure.kte = nlev - 1;
    infrastructure.predictNc = predictNc; infrastructure.prescribedCCN = prescribedCCN;
    P3F::P3Runtime runtime_options;          // P3Runtime{740.0e3}: all defaults (p3_functions_f90.cpp:2061)
    P3F::WorkspaceManager workspace_mgr;
    P3F::check_columns = true;

    long total_us = 0;
    for (int s = 0; s < nsteps; ++s) {
      infrastructure.it++;
      workspace_mgr.reset_internals();
      // run_impl zero-fills the history arrays (eamxx_p3_run.cpp:27-29); the test bridge p3_main_f also zeroes
      // every other output on entry (p3_functions_f90.cpp:1941-1945), which makes precip_*_flux(nk) defined
      for (int f = P3B200_F_QV2QI_DEPOS_TEND; f < P3B200_F_NUM; ++f) CK(cudaMemset(dev[f], 0, count[f] * sizeof(Real)));
      total_us += P3F::p3_main(runtime_options, prog_state, diag_inputs, diag_outputs, infrastructure, history_only,
                               lookup_tables, workspace_mgr, ncol, nlev);
    }
    std::FILE* out = std::fopen(argv[3], "wb");
    if (!out) { std::perror(argv[3]); return 1; }
    for (int f = 0; f < P3B200_F_NUM; ++f) {
      std::vector<Real> h(count[f]);
      CK(cudaMemcpy(h.data(), dev[f], count[f] * sizeof(Real), cudaMemcpyDeviceToHost));
      std::fwrite(h.data(), sizeof(Real), count[f], out);
      cudaFree(dev[f]);
    }
    std::fclose(out);
    std::printf("p3_run_and_cmp_b200: %d columns x %d levels, %d steps, %ld us of device time\n", ncol, nlev, nsteps, total_us);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "p3_run_and_cmp_b200: %s\n", e.what());
    return 3;
  }
  return 0;
}
