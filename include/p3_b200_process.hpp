// p3_b200_process.hpp — the process-level face of libp3b200: a class with the members and the call sequence of
// EAMxx's P3Microphysics (components/eamxx/src/physics/p3/eamxx_p3_process_interface.hpp:22-60, 404-430;
// eamxx_p3_process_interface.cpp:220-425 initialize_impl; eamxx_p3_run.cpp:5-46 run_impl), on top of
// p3::Functions<Real, B200Device> (include/p3_b200_functions.hpp) and the C ABI (include/p3_b200.h).
//
// What is kept: name() / type(), the protected trio initialize_impl / run_impl / finalize_impl (public here: there is
// no AtmosphereProcess base class to befriend), requested_buffer_size_in_bytes() / init_buffers(), the member names
// (m_num_cols, m_num_levs, m_buffer, prog_state, diag_inputs, diag_outputs, history_only, lookup_tables,
// infrastructure, runtime_options, workspace_mgr), the Buffer members, the field names of the field manager, and the
// run_impl sequence: p3_preamble -> (zero the history fields) -> p3_main -> p3_postamble, followed by the property
// checks initialize_impl registers (:226-243) with the repair of qv.
// What is replaced: the FieldManager/GridsManager (set_grids) by a struct of device pointers named after the fields;
// Kokkos views by the non-owning PODs of p3_b200_functions.hpp; the ATMBufferManager by a caller-provided device
// allocation; the three Kokkos launches + 15 reductions of one run_impl by sm_100a kernels (p3b200_main_preamble =
// p3_preamble inside p3_main's streaming pass, with zero_history; p3b200_postamble_checked).
//
// Header-only, C++17, no Kokkos/EKAT.  Errors throw std::logic_error like EKAT_REQUIRE_MSG.
#ifndef P3_B200_PROCESS_HPP
#define P3_B200_PROCESS_HPP

#include "p3_b200_functions.hpp"

#include <array>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>

namespace scream {

enum class AtmosphereProcessType { Dynamics, Physics, Group };   // share/atm_process/atmosphere_process_utils.hpp
enum class RunType { Initial, Restarted };                       // share/scream_types.hpp:36-39

template <typename Real = double>
class P3MicrophysicsB200 {
 public:
  using P3F = p3::Functions<Real, B200Device>;
  using view_1d = typename P3F::template view_1d<Real>;
  using view_2d = typename P3F::template view_2d<Real>;

  // ---- what the reference reads from its ekat::ParameterList (eamxx_p3_process_interface.cpp:24-30, 222-223)
  struct Params {
    bool do_predict_nc = true;
    bool do_prescribed_ccn = true;
    bool enable_column_conservation_checks = false;   // has_column_conservation_check(): boundary fluxes are produced
    std::string p3_lookup_table;                      // text table or blob (p3::Functions::p3_init)
    int device = 0;
    typename P3F::P3Runtime runtime_options;          // load_runtime_options_from_file
  };

  // ---- the fields of the field manager (set_grids, eamxx_p3_process_interface.cpp:66-120): DEVICE pointers,
  // (ncol, nlev) level-contiguous with one common column stride; per-column fields have ncol entries.
  struct Fields {
    // required
    const Real *p_mid = nullptr, *p_dry_mid = nullptr, *pseudo_density = nullptr, *pseudo_density_dry = nullptr;
    const Real *cldfrac_tot = nullptr, *nc_nuceat_tend = nullptr, *nccn = nullptr, *ni_activated = nullptr;
    const Real *inv_qc_relvar = nullptr;
    // updated
    Real *T_mid = nullptr, *qv = nullptr, *qc = nullptr, *qr = nullptr, *qi = nullptr, *qm = nullptr;
    Real *nc = nullptr, *nr = nullptr, *ni = nullptr, *bm = nullptr;
    Real *qv_prev_micro_step = nullptr, *T_prev_micro_step = nullptr;
    Real *precip_liq_surf_mass = nullptr, *precip_ice_surf_mass = nullptr;     // [ncol], accumulated
    // computed
    Real *eff_radius_qc = nullptr, *eff_radius_qi = nullptr, *eff_radius_qr = nullptr;
    Real *precip_total_tend = nullptr, *nevapr = nullptr, *rainfrac = nullptr;
    Real *micro_liq_ice_exchange = nullptr, *micro_vap_liq_exchange = nullptr, *micro_vap_ice_exchange = nullptr;
    // only with enable_column_conservation_checks ([ncol])
    Real *vapor_flux = nullptr, *water_flux = nullptr, *ice_flux = nullptr, *heat_flux = nullptr;
    std::int64_t col_stride = 0;   // 0: nlev
  };

  // ---- local variables in memory provided by the caller (eamxx_p3_process_interface.hpp:358-396)
  struct Buffer {
    static constexpr int num_1d_scalar = 2;
    static constexpr int num_2d_vector = 8;     // SCREAM_P3_SMALL_KERNELS adds 48 temporaries the kernels never need
    static constexpr int num_2dp1_vector = 2;
    view_1d precip_liq_surf_flux, precip_ice_surf_flux;
    view_2d inv_exner, th_atm, cld_frac_l, cld_frac_i, dz, qv2qi_depos_tend, rho_qi, unused;
    view_2d precip_liq_flux, precip_ice_flux;   // nlev + 1 levels
  };

  P3MicrophysicsB200(int num_cols, int num_levs, const Params& params)
      : m_num_cols(num_cols), m_num_levs(num_levs), m_params(params) {
    if (num_cols < 0 || num_levs < 1 || num_levs > P3B200_MAX_NLEV) throw std::logic_error("P3Microphysics: bad grid size");
    infrastructure.it = 0;                     // eamxx_p3_process_interface.cpp:34
    infrastructure.predictNc = params.do_predict_nc;
    infrastructure.prescribedCCN = params.do_prescribed_ccn;
    runtime_options = params.runtime_options;
  }

  AtmosphereProcessType type() const { return AtmosphereProcessType::Physics; }
  std::string name() const { return "p3"; }

  // eamxx_p3_process_interface.cpp:124-143 (the WorkspaceManager's share is zero here: nothing lives in global memory)
  std::size_t requested_buffer_size_in_bytes() const {
    return align(Buffer::num_1d_scalar * std::size_t(m_num_cols) * sizeof(Real)) +
           Buffer::num_2d_vector * align(std::size_t(m_num_cols) * m_num_levs * sizeof(Real)) +
           Buffer::num_2dp1_vector * align(std::size_t(m_num_cols) * (m_num_levs + 1) * sizeof(Real));
  }

  // eamxx_p3_process_interface.cpp:147-218: carve the local views out of `device_mem`
  void init_buffers(void* device_mem, std::size_t bytes) {
    if (bytes < requested_buffer_size_in_bytes())
      throw std::logic_error("Error! Buffers size not sufficient.\n");      // :149
    char* p = static_cast<char*>(device_mem);
    auto take = [&](std::size_t n) { Real* r = reinterpret_cast<Real*>(p); p += align(n * sizeof(Real)); return r; };
    const int nc = m_num_cols, nk = m_num_levs;
    {
      Real* s = take(std::size_t(Buffer::num_1d_scalar) * nc);
      m_buffer.precip_liq_surf_flux = view_1d(s, nc);
      m_buffer.precip_ice_surf_flux = view_1d(s + nc, nc);
    }
    view_2d* mid[Buffer::num_2d_vector] = {&m_buffer.inv_exner, &m_buffer.th_atm, &m_buffer.cld_frac_l, &m_buffer.cld_frac_i,
                                           &m_buffer.dz, &m_buffer.qv2qi_depos_tend, &m_buffer.rho_qi, &m_buffer.unused};
    for (auto* v : mid) *v = view_2d(take(std::size_t(nc) * nk), nc, nk);
    m_buffer.precip_liq_flux = view_2d(take(std::size_t(nc) * (nk + 1)), nc, nk + 1);
    m_buffer.precip_ice_flux = view_2d(take(std::size_t(nc) * (nk + 1)), nc, nk + 1);
    m_buffers_set = true;
  }

  // eamxx_p3_process_interface.cpp:220-425: tables, argument structs, pre/post-processing bindings
  void initialize_impl(const Fields& f, RunType /*run_type*/ = RunType::Initial) {
    if (!m_buffers_set) throw std::logic_error("P3Microphysics::initialize_impl: init_buffers has not been called");
    require(f.p_mid && f.p_dry_mid && f.pseudo_density && f.pseudo_density_dry && f.cldfrac_tot && f.inv_qc_relvar,
            "a required input field is missing");
    require(f.T_mid && f.qv && f.qc && f.qr && f.qi && f.qm && f.nc && f.nr && f.ni && f.bm && f.qv_prev_micro_step &&
            f.T_prev_micro_step && f.precip_liq_surf_mass && f.precip_ice_surf_mass, "an updated field is missing");
    require(f.eff_radius_qc && f.eff_radius_qi && f.eff_radius_qr && f.precip_total_tend && f.nevapr && f.rainfrac &&
            f.micro_liq_ice_exchange && f.micro_vap_liq_exchange && f.micro_vap_ice_exchange, "a computed field is missing");
    require(!m_params.do_prescribed_ccn || f.nccn, "nccn is required with do_prescribed_ccn");
    require(m_params.do_prescribed_ccn || !m_params.do_predict_nc || (f.nc_nuceat_tend && f.ni_activated),
            "nc_nuceat_tend / ni_activated are required with do_predict_nc and without do_prescribed_ccn");
    if (m_params.enable_column_conservation_checks)
      require(f.vapor_flux && f.water_flux && f.ice_flux && f.heat_flux, "the four boundary flux fields are missing");
    m_fields = f;
    const int nc = m_num_cols, nk = m_num_levs;
    const std::int64_t cs = f.col_stride ? f.col_stride : nk;
    if (cs != nk) throw std::logic_error("P3Microphysics: the local buffers are dense; field column stride must equal nlev");

    lookup_tables = P3F::p3_init(m_params.p3_lookup_table, m_params.device);     // p3::p3_init + init_kokkos_*_tables

    auto v = [&](Real* p) { return view_2d(p, nc, nk, cs); };
    auto cv = [&](const Real* p) { return typename P3F::template view_2d<const Real>(p, nc, nk, cs); };
    prog_state.qc = v(f.qc); prog_state.nc = v(f.nc); prog_state.qr = v(f.qr); prog_state.nr = v(f.nr);
    prog_state.qi = v(f.qi); prog_state.qm = v(f.qm); prog_state.ni = v(f.ni); prog_state.bm = v(f.bm);
    prog_state.th = m_buffer.th_atm; prog_state.qv = v(f.qv);
    diag_inputs.nc_nuceat_tend = cv(f.nc_nuceat_tend);
    diag_inputs.nccn = m_params.do_prescribed_ccn ? cv(f.nccn) : cv(m_buffer.unused.data());   // :301-305
    diag_inputs.ni_activated = cv(f.ni_activated);
    diag_inputs.inv_qc_relvar = cv(f.inv_qc_relvar);
    diag_inputs.pres = cv(f.p_dry_mid);                    // :309 dry pressure
    diag_inputs.dpres = cv(f.pseudo_density_dry);          // :310 dry density
    diag_inputs.qv_prev = cv(f.qv_prev_micro_step);
    diag_inputs.t_prev = cv(f.T_prev_micro_step);
    diag_inputs.cld_frac_l = m_buffer.cld_frac_l; diag_inputs.cld_frac_i = m_buffer.cld_frac_i;
    diag_inputs.cld_frac_r = cv(f.rainfrac);
    diag_inputs.dz = m_buffer.dz; diag_inputs.inv_exner = m_buffer.inv_exner;
    diag_outputs.diag_eff_radius_qc = v(f.eff_radius_qc); diag_outputs.diag_eff_radius_qi = v(f.eff_radius_qi);
    diag_outputs.diag_eff_radius_qr = v(f.eff_radius_qr);
    diag_outputs.precip_total_tend = v(f.precip_total_tend); diag_outputs.nevapr = v(f.nevapr);
    diag_outputs.precip_liq_surf = m_buffer.precip_liq_surf_flux; diag_outputs.precip_ice_surf = m_buffer.precip_ice_surf_flux;
    diag_outputs.qv2qi_depos_tend = m_buffer.qv2qi_depos_tend; diag_outputs.rho_qi = m_buffer.rho_qi;
    diag_outputs.precip_liq_flux = m_buffer.precip_liq_flux; diag_outputs.precip_ice_flux = m_buffer.precip_ice_flux;
    history_only.liq_ice_exchange = v(f.micro_liq_ice_exchange);
    history_only.vap_liq_exchange = v(f.micro_vap_liq_exchange);
    history_only.vap_ice_exchange = v(f.micro_vap_ice_exchange);
    infrastructure.its = 0; infrastructure.ite = nc - 1; infrastructure.kts = 0; infrastructure.kte = nk - 1;   // :329-334

    // p3_preproc.set_variables (:286-291)
    pre_ = P3B200PreambleFields{};
    pre_.pmid = f.p_mid; pre_.pseudo_density = f.pseudo_density; pre_.pseudo_density_dry = f.pseudo_density_dry;
    pre_.T_atm = f.T_mid; pre_.cld_frac_t = f.cldfrac_tot;
    pre_.qv = f.qv; pre_.qc = f.qc; pre_.nc = f.nc; pre_.qr = f.qr; pre_.nr = f.nr; pre_.qi = f.qi; pre_.qm = f.qm;
    pre_.ni = f.ni; pre_.bm = f.bm; pre_.qv_prev = f.qv_prev_micro_step;
    pre_.inv_exner = m_buffer.inv_exner.data(); pre_.th_atm = m_buffer.th_atm.data();
    pre_.cld_frac_l = m_buffer.cld_frac_l.data(); pre_.cld_frac_i = m_buffer.cld_frac_i.data();
    pre_.cld_frac_r = f.rainfrac; pre_.dz = m_buffer.dz.data();
    pre_.col_stride = cs;
    // p3_postproc.set_variables (:345-353) and set_mass_and_energy_fluxes (:355-364)
    post_ = P3B200PostambleFields{};
    post_.pmid = f.p_mid; post_.pseudo_density = f.pseudo_density; post_.pseudo_density_dry = f.pseudo_density_dry;
    post_.th_atm = m_buffer.th_atm.data(); post_.T_atm = f.T_mid;
    post_.qv = f.qv; post_.qc = f.qc; post_.nc = f.nc; post_.qr = f.qr; post_.nr = f.nr; post_.qi = f.qi;
    post_.qm = f.qm; post_.ni = f.ni; post_.bm = f.bm;
    post_.diag_eff_radius_qc = f.eff_radius_qc; post_.diag_eff_radius_qi = f.eff_radius_qi;
    post_.diag_eff_radius_qr = f.eff_radius_qr;
    post_.T_prev = f.T_prev_micro_step; post_.qv_prev = f.qv_prev_micro_step;
    post_.precip_liq_surf_flux = m_buffer.precip_liq_surf_flux.data();
    post_.precip_ice_surf_flux = m_buffer.precip_ice_surf_flux.data();
    post_.precip_liq_surf_mass = f.precip_liq_surf_mass; post_.precip_ice_surf_mass = f.precip_ice_surf_mass;
    if (m_params.enable_column_conservation_checks) {
      post_.vapor_flux = f.vapor_flux; post_.water_flux = f.water_flux; post_.ice_flux = f.ice_flux; post_.heat_flux = f.heat_flux;
    }
    post_.col_stride = cs;
    m_initialized = true;
  }

  // eamxx_p3_run.cpp:5-46.  `stream`: optional cudaStream_t.  The call returns after the property checks, i.e. after
  // the work has completed (the reference fences after each of its three launches).
  void run_impl(const double dt, void* stream = nullptr) {
    if (!m_initialized) throw std::logic_error("P3Microphysics::run_impl before initialize_impl");
    p3b200_handle h = lookup_tables.handle.get();
    post_.dt = dt;                                                        // p3_postproc.m_dt = dt
    // p3_preproc: fused into p3_main's first kernel (p3b200_main_preamble) unless fuse_preamble is switched off
    if (!fuse_preamble) check(p3b200_preamble(h, &pre_, m_num_cols, m_num_levs, stream), "p3_preamble");
    infrastructure.dt = dt;
    infrastructure.it++;
    workspace_mgr.reset_internals();
    // get_field_out("micro_*_exchange").deep_copy(0.0) x3 + P3F::p3_main: the zero-fill is folded into the call
    m_last_p3_main_us = main_with_zeroed_history(h, stream);
    // p3_postproc, with the 15 property checks of initialize_impl folded into the pass
    check(p3b200_postamble_checked(h, &post_, m_num_cols, m_num_levs, m_checks.data(), stream), "p3_postamble");
    // AtmosphereProcess::run_property_check: repair what is repairable, fail on the rest
    m_repaired = false;
    for (int i = 0; i < P3B200_NUM_CHECKS; ++i) {
      if (m_checks[i].result == P3B200_CHECK_REPAIRABLE) {
        check(p3b200_repair_qv(h, m_fields.qv, post_.col_stride, m_num_cols, m_num_levs, stream), "repair qv");
        m_repaired = true;
      } else if (m_checks[i].result == P3B200_CHECK_FAIL && fail_on_check) {
        throw std::logic_error("P3Microphysics: property check " + std::to_string(i) + " failed: min " +
                               std::to_string(m_checks[i].min_val) + " max " + std::to_string(m_checks[i].max_val));
      }
    }
  }

  void finalize_impl() {
    lookup_tables = typename P3F::P3LookupTables{};
    m_initialized = false;
  }

  const std::array<P3B200CheckResult, P3B200_NUM_CHECKS>& property_check_results() const { return m_checks; }
  bool qv_was_repaired() const { return m_repaired; }
  Int last_p3_main_microseconds() const { return m_last_p3_main_us; }
  bool fail_on_check = true;
  bool fuse_preamble = true;

  // Keep track of field dimensions and the iteration count
  Int m_num_cols, m_num_levs;
  Buffer m_buffer;
  // the structures for each argument to p3_main
  typename P3F::P3PrognosticState prog_state;
  typename P3F::P3DiagnosticInputs diag_inputs;
  typename P3F::P3DiagnosticOutputs diag_outputs;
  typename P3F::P3HistoryOnly history_only;
  typename P3F::P3LookupTables lookup_tables;
  typename P3F::P3Infrastructure infrastructure;
  typename P3F::P3Runtime runtime_options;
  typename P3F::WorkspaceManager workspace_mgr;

 private:
  static std::size_t align(std::size_t n) { return (n + 255) / 256 * 256; }
  static void require(bool ok, const char* what) { if (!ok) throw std::logic_error(std::string("P3Microphysics: ") + what); }
  static void check(int rc, const char* what) {
    if (rc) throw std::logic_error(std::string("P3Microphysics: ") + what + ": " + p3b200_strerror(rc) + " " + p3b200_last_cuda_error());
  }
  Int main_with_zeroed_history(p3b200_handle, void* stream) {
    return P3F::p3_main(runtime_options, prog_state, diag_inputs, diag_outputs, infrastructure, history_only, lookup_tables,
                        workspace_mgr, m_num_cols, m_num_levs, stream, /*zero_history=*/true,
                        fuse_preamble ? &pre_ : nullptr);
  }

  Params m_params;
  Fields m_fields;
  P3B200PreambleFields pre_{};
  P3B200PostambleFields post_{};
  std::array<P3B200CheckResult, P3B200_NUM_CHECKS> m_checks{};
  bool m_buffers_set = false, m_initialized = false, m_repaired = false;
  Int m_last_p3_main_us = 0;
};

using P3Microphysics = P3MicrophysicsB200<double>;

}  // namespace scream
#endif  // P3_B200_PROCESS_HPP
