// p3_b200_functions.hpp — the host-side C++ face of libp3b200, in the shape of EAMxx's own
// p3::Functions<Scalar, Device> (components/eamxx/src/physics/p3/p3_functions.hpp).
//
// The reference binds P3Microphysics::run_impl (p3/eamxx_p3_run.cpp:5-46) to
//     P3F::p3_main(runtime_options, prog_state, diag_inputs, diag_outputs, infrastructure,
//                  history_only, lookup_tables, workspace_mgr, m_num_cols, m_num_levs);
// with P3F = p3::Functions<Real, DefaultDevice> (p3/eamxx_p3_process_interface.hpp:24).  This header
// provides the specialisation p3::Functions<Real, B200Device> with the SAME struct names, member names,
// argument order, return value (elapsed microseconds as Int) and error behaviour (host-side failures
// throw std::logic_error like EKAT_REQUIRE_MSG, p3_table_ice_impl.hpp:40-41), so that the call site above
// compiles unchanged with `using P3F = p3::Functions<Real, B200Device>`.
//
// Header-only, C++17, no Kokkos/EKAT: views are the 30-line non-owning PODs below (data, extents,
// column stride; level index contiguous = Kokkos LayoutRight with pack size 1).  Everything that
// computes is behind the C ABI of include/p3_b200.h (hand-written sm_100a kernels); there is no CPU path.
#ifndef P3_B200_FUNCTIONS_HPP
#define P3_B200_FUNCTIONS_HPP

#include "p3_b200.h"

#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>

namespace scream {

using Int = int;   // ekat/ekat.hpp:15

struct B200Device {};   // tag taking the place of DefaultDevice / HostDevice

namespace p3 {

// ---- non-owning views of DEVICE memory (replace KokkosTypes<Device>::view_1d / view_2d) ----------------
template <typename T>
struct view_1d {
  T* ptr = nullptr;
  Int n0 = 0;
  view_1d() = default;
  view_1d(T* p, Int e0) : ptr(p), n0(e0) {}
  template <typename U> view_1d(const view_1d<U>& o) : ptr(o.ptr), n0(o.n0) {}   // T -> const T
  T* data() const { return ptr; }
  Int extent(int) const { return n0; }
  Int extent_int(int) const { return n0; }
  size_t size() const { return size_t(n0); }
};

template <typename T>
struct view_2d {   // LayoutRight (nj, nk): level index fastest; a tracer sub-view has stride(0) > extent(1)
  T* ptr = nullptr;
  Int n0 = 0, n1 = 0;
  int64_t s0 = 0;
  view_2d() = default;
  view_2d(T* p, Int e0, Int e1, int64_t col_stride = -1) : ptr(p), n0(e0), n1(e1), s0(col_stride < 0 ? e1 : col_stride) {}
  template <typename U> view_2d(const view_2d<U>& o) : ptr(o.ptr), n0(o.n0), n1(o.n1), s0(o.s0) {}
  T* data() const { return ptr; }
  Int extent(int d) const { return d == 0 ? n0 : n1; }
  Int extent_int(int d) const { return extent(d); }
  int64_t stride(int d) const { return d == 0 ? s0 : 1; }
  size_t size() const { return size_t(n0) * size_t(n1); }
};

struct P3C {   // p3_functions.hpp:30-52
  enum { densize = P3B200_DENSIZE, rimsize = P3B200_RIMSIZE, isize = P3B200_ISIZE, ice_table_size = P3B200_ICE_TABLE_SIZE,
         rcollsize = P3B200_RCOLLSIZE, collect_table_size = P3B200_COLLECT_TABLE_SIZE,
         MU_R_TABLE_DIM = P3B200_MU_R_TABLE_DIM, VTABLE_DIM0 = P3B200_VTABLE_DIM0, VTABLE_DIM1 = P3B200_VTABLE_DIM1,
         dnusize = P3B200_DNUSIZE };
  static constexpr const char* p3_lookup_base = "p3_lookup_table_1.dat-v";
  static constexpr const char* p3_version = "4.1.1";
};

template <typename ScalarT, typename DeviceT>
struct Functions;   // only the B200Device specialisation exists in this library

template <typename ScalarT>
struct Functions<ScalarT, B200Device> {
  static_assert(sizeof(ScalarT) == 4 || sizeof(ScalarT) == 8, "Real is float or double");
  using Scalar = ScalarT;
  using Spack = ScalarT;        // pack size 1: the warp lanes take the place of SIMD packs
  using Device = B200Device;
  using P3C = p3::P3C;
  template <typename S> using view_1d = p3::view_1d<S>;
  template <typename S> using view_2d = p3::view_2d<S>;

  [[noreturn]] static void fail(int rc, const char* what) {
    throw std::logic_error(std::string(what) + ": " + p3b200_strerror(rc) + " " + p3b200_last_cuda_error());
  }

  // p3_functions.hpp:112-158
  struct P3Runtime {
    Scalar max_total_ni = 740.0e3;
    Scalar autoconversion_prefactor = 1350.0;
    Scalar autoconversion_qc_exponent = 2.47;
    Scalar autoconversion_nc_exponent = 1.79;
    Scalar autoconversion_radius = 25.0e-6;
    Scalar accretion_prefactor = 67.0;
    Scalar accretion_qc_exponent = 1.15;
    Scalar accretion_qr_exponent = 1.15;
    Scalar rain_selfcollection_prefactor = 5.78;
    Scalar rain_selfcollection_breakup_diameter = 0.00028;
    Scalar constant_mu_rain = 1.0;
    Scalar spa_ccn_to_nc_factor = 1.0;
    Scalar cldliq_to_ice_collection_factor = 0.5;
    Scalar rain_to_ice_collection_factor = 1.0;
    Scalar min_rime_rho = 50.0;
    Scalar max_rime_rho = 900.0;
    Scalar immersion_freezing_exponent = 0.65;
    Scalar deposition_nucleation_exponent = 0.304;
    Scalar ice_sedimentation_factor = 1.0;
    bool do_ice_production = true;

    // ParameterList is any type with `template<class T> T get(const std::string&, const T& default)`
    // (ekat::ParameterList has it); same keys as p3_functions.hpp:135-156.
    template <typename ParameterList>
    void load_runtime_options_from_file(ParameterList& params) {
#define P3B200_LOAD(name) name = Scalar(params.template get<double>(#name, double(name)))
      P3B200_LOAD(max_total_ni); P3B200_LOAD(autoconversion_prefactor); P3B200_LOAD(autoconversion_qc_exponent);
      P3B200_LOAD(autoconversion_nc_exponent); P3B200_LOAD(autoconversion_radius); P3B200_LOAD(accretion_prefactor);
      P3B200_LOAD(accretion_qc_exponent); P3B200_LOAD(accretion_qr_exponent); P3B200_LOAD(rain_selfcollection_prefactor);
      P3B200_LOAD(rain_selfcollection_breakup_diameter); P3B200_LOAD(constant_mu_rain); P3B200_LOAD(spa_ccn_to_nc_factor);
      P3B200_LOAD(cldliq_to_ice_collection_factor); P3B200_LOAD(rain_to_ice_collection_factor); P3B200_LOAD(min_rime_rho);
      P3B200_LOAD(max_rime_rho); P3B200_LOAD(immersion_freezing_exponent); P3B200_LOAD(deposition_nucleation_exponent);
      P3B200_LOAD(ice_sedimentation_factor);
#undef P3B200_LOAD
      do_ice_production = params.template get<bool>("do_ice_production", do_ice_production);
    }
  };

  // p3_functions.hpp:161-183
  struct P3PrognosticState {
    view_2d<Spack> qc, nc, qr, nr, qi, qm, ni, bm, qv, th;
  };
  // p3_functions.hpp:186-214
  struct P3DiagnosticInputs {
    view_2d<const Spack> nc_nuceat_tend, nccn, ni_activated, inv_qc_relvar, cld_frac_i, cld_frac_l, cld_frac_r,
                         pres, dz, dpres, inv_exner, qv_prev, t_prev;
  };
  // p3_functions.hpp:217-241
  struct P3DiagnosticOutputs {
    view_2d<Spack> qv2qi_depos_tend;
    view_1d<Scalar> precip_liq_surf, precip_ice_surf;
    view_2d<Spack> diag_eff_radius_qc, diag_eff_radius_qi, diag_eff_radius_qr, rho_qi,
                   precip_liq_flux, precip_ice_flux /* (nj, nk+1) */, precip_total_tend, nevapr;
  };
  // p3_functions.hpp:244-264
  struct P3Infrastructure {
    double dt = 0;
    Int it = 0, its = 0, ite = -1, kts = 0, kte = -1;
    bool predictNc = true, prescribedCCN = true;
    view_2d<const Scalar> col_location;   // debug only in the reference; unused here
  };
  // p3_functions.hpp:268-276
  struct P3HistoryOnly {
    view_2d<Spack> liq_ice_exchange, vap_liq_exchange, vap_ice_exchange;
  };

  // p3_functions.hpp:279-290.  In the reference these are Kokkos views of the tables; here the tables live
  // on the device inside the library handle, which this struct shares (ref-counted like a Kokkos view).
  struct P3LookupTables {
    std::shared_ptr<p3b200_ctx> handle;     // device tables + launch state (p3b200_create / p3b200_destroy)
    std::shared_ptr<P3B200Tables> host;     // host mirror: mu_r/vn/vm/revap/ice/collect/dnu _table_vals, as doubles
    int device = 0;
  };

  // ekat::WorkspaceManager stand-in (p3_functions.hpp:108-109).  The reference hands p3_main a pool of 52
  // per-team scratch arrays in global memory (eamxx_p3_process_interface.cpp:141,424); the B200 kernels
  // keep every per-column temporary in registers / shared memory, so nothing is pooled.
  struct WorkspaceManager {
    WorkspaceManager() = default;
    template <typename... Args> explicit WorkspaceManager(Args&&...) {}
    void reset_internals() const {}
    static size_t get_total_bytes_needed(int, int, int) { return 0; }
  };

  // Replaces init_kokkos_ice_lookup_tables (p3_table_ice_impl.hpp:18-78), init_kokkos_tables
  // (p3_table3_impl.hpp:87-148) and the Fortran p3_init_b generator behind them.  `table_path` is either
  // the reference's text table "<dir>/p3_lookup_table_1.dat-v4.1.1" or the binary blob shipped in
  // scream_b200/data (tools/make_table_blob.py); a wrong version throws like p3_table_ice_impl.hpp:40-41.
  static P3LookupTables p3_init(const std::string& table_path, int device = 0) {
    P3LookupTables t;
    t.host = std::make_shared<P3B200Tables>();
    t.device = device;
    const bool text = table_path.find(".dat") != std::string::npos;
    int rc = text ? p3b200_tables_read_text(table_path.c_str(), P3C::p3_version, t.host.get())
                  : p3b200_tables_read_blob(table_path.c_str(), P3C::p3_version, t.host.get());
    if (rc) fail(rc, "p3_init: ice lookup table");
    rc = p3b200_tables_generate_rain(t.host.get());
    if (rc) fail(rc, "p3_init: rain lookup tables");
    p3b200_handle h = nullptr;
    rc = p3b200_create(t.host.get(), device, (int)sizeof(Scalar), &h);
    if (rc) fail(rc, "p3_init: p3b200_create");
    t.handle = std::shared_ptr<p3b200_ctx>(h, [](p3b200_ctx* p) { p3b200_destroy(p); });
    return t;
  }

  // Static method p3_main (p3_functions.hpp:1356-1369; p3_main_impl.hpp:335-372): same argument list,
  // returns the device time of the call in microseconds.  Views hold DEVICE pointers.  `stream` is an
  // optional cudaStream_t; like the reference (Kokkos::fence, p3_main_impl.hpp:326) the call returns
  // after the work has completed.  `zero_history` (an extension): the call zero-fills the three history arrays
  // itself, replacing the three deep_copy(0.0) that precede p3_main in run_impl.
  static Int p3_main(const P3Runtime& runtime, const P3PrognosticState& prognostic_state,
                     const P3DiagnosticInputs& diagnostic_inputs, const P3DiagnosticOutputs& diagnostic_outputs,
                     const P3Infrastructure& infrastructure, const P3HistoryOnly& history_only,
                     const P3LookupTables& lookup_tables, const WorkspaceManager& /*workspace_mgr*/,
                     Int nj, Int nk, void* stream = nullptr, bool zero_history = false,
                     const P3B200PreambleFields* fused_preamble = nullptr) {
    if (!lookup_tables.handle) fail(P3B200_EINVAL, "p3_main: lookup tables not initialised (p3_init)");
    P3B200Fields f;
    for (int i = 0; i < P3B200_F_NUM; ++i) { f.ptr[i] = nullptr; f.col_stride[i] = 0; }
    auto set2 = [&](int id, const auto& v, Int levels, const char* name) {
      if (v.data() && (v.extent(0) < nj || v.extent(1) < levels))
        throw std::logic_error(std::string("p3_main: view too small: ") + name);
      f.ptr[id] = const_cast<void*>(static_cast<const void*>(v.data()));
      f.col_stride[id] = v.stride(0);
    };
    const auto& ps = prognostic_state; const auto& di = diagnostic_inputs; const auto& dg = diagnostic_outputs;
    set2(P3B200_F_QC, ps.qc, nk, "qc"); set2(P3B200_F_NC, ps.nc, nk, "nc"); set2(P3B200_F_QR, ps.qr, nk, "qr");
    set2(P3B200_F_NR, ps.nr, nk, "nr"); set2(P3B200_F_QI, ps.qi, nk, "qi"); set2(P3B200_F_QM, ps.qm, nk, "qm");
    set2(P3B200_F_NI, ps.ni, nk, "ni"); set2(P3B200_F_BM, ps.bm, nk, "bm"); set2(P3B200_F_QV, ps.qv, nk, "qv");
    set2(P3B200_F_TH, ps.th, nk, "th");
    set2(P3B200_F_NC_NUCEAT_TEND, di.nc_nuceat_tend, nk, "nc_nuceat_tend"); set2(P3B200_F_NCCN, di.nccn, nk, "nccn");
    set2(P3B200_F_NI_ACTIVATED, di.ni_activated, nk, "ni_activated");
    set2(P3B200_F_INV_QC_RELVAR, di.inv_qc_relvar, nk, "inv_qc_relvar");
    set2(P3B200_F_CLD_FRAC_I, di.cld_frac_i, nk, "cld_frac_i"); set2(P3B200_F_CLD_FRAC_L, di.cld_frac_l, nk, "cld_frac_l");
    set2(P3B200_F_CLD_FRAC_R, di.cld_frac_r, nk, "cld_frac_r"); set2(P3B200_F_PRES, di.pres, nk, "pres");
    set2(P3B200_F_DZ, di.dz, nk, "dz"); set2(P3B200_F_DPRES, di.dpres, nk, "dpres");
    set2(P3B200_F_INV_EXNER, di.inv_exner, nk, "inv_exner"); set2(P3B200_F_QV_PREV, di.qv_prev, nk, "qv_prev");
    set2(P3B200_F_T_PREV, di.t_prev, nk, "t_prev");
    set2(P3B200_F_QV2QI_DEPOS_TEND, dg.qv2qi_depos_tend, nk, "qv2qi_depos_tend");
    set2(P3B200_F_DIAG_EFF_RADIUS_QC, dg.diag_eff_radius_qc, nk, "diag_eff_radius_qc");
    set2(P3B200_F_DIAG_EFF_RADIUS_QI, dg.diag_eff_radius_qi, nk, "diag_eff_radius_qi");
    set2(P3B200_F_DIAG_EFF_RADIUS_QR, dg.diag_eff_radius_qr, nk, "diag_eff_radius_qr");
    set2(P3B200_F_RHO_QI, dg.rho_qi, nk, "rho_qi");
    set2(P3B200_F_PRECIP_LIQ_FLUX, dg.precip_liq_flux, nk + 1, "precip_liq_flux");
    set2(P3B200_F_PRECIP_ICE_FLUX, dg.precip_ice_flux, nk + 1, "precip_ice_flux");
    set2(P3B200_F_PRECIP_TOTAL_TEND, dg.precip_total_tend, nk, "precip_total_tend");
    set2(P3B200_F_NEVAPR, dg.nevapr, nk, "nevapr");
    set2(P3B200_F_LIQ_ICE_EXCHANGE, history_only.liq_ice_exchange, nk, "liq_ice_exchange");
    set2(P3B200_F_VAP_LIQ_EXCHANGE, history_only.vap_liq_exchange, nk, "vap_liq_exchange");
    set2(P3B200_F_VAP_ICE_EXCHANGE, history_only.vap_ice_exchange, nk, "vap_ice_exchange");
    f.ptr[P3B200_F_PRECIP_LIQ_SURF] = dg.precip_liq_surf.data(); f.col_stride[P3B200_F_PRECIP_LIQ_SURF] = 1;
    f.ptr[P3B200_F_PRECIP_ICE_SURF] = dg.precip_ice_surf.data(); f.col_stride[P3B200_F_PRECIP_ICE_SURF] = 1;

    P3B200Runtime r;
    r.max_total_ni = runtime.max_total_ni; r.autoconversion_prefactor = runtime.autoconversion_prefactor;
    r.autoconversion_qc_exponent = runtime.autoconversion_qc_exponent;
    r.autoconversion_nc_exponent = runtime.autoconversion_nc_exponent;
    r.autoconversion_radius = runtime.autoconversion_radius; r.accretion_prefactor = runtime.accretion_prefactor;
    r.accretion_qc_exponent = runtime.accretion_qc_exponent; r.accretion_qr_exponent = runtime.accretion_qr_exponent;
    r.rain_selfcollection_prefactor = runtime.rain_selfcollection_prefactor;
    r.rain_selfcollection_breakup_diameter = runtime.rain_selfcollection_breakup_diameter;
    r.constant_mu_rain = runtime.constant_mu_rain; r.spa_ccn_to_nc_factor = runtime.spa_ccn_to_nc_factor;
    r.cldliq_to_ice_collection_factor = runtime.cldliq_to_ice_collection_factor;
    r.rain_to_ice_collection_factor = runtime.rain_to_ice_collection_factor;
    r.min_rime_rho = runtime.min_rime_rho; r.max_rime_rho = runtime.max_rime_rho;
    r.immersion_freezing_exponent = runtime.immersion_freezing_exponent;
    r.deposition_nucleation_exponent = runtime.deposition_nucleation_exponent;
    r.ice_sedimentation_factor = runtime.ice_sedimentation_factor;
    r.do_ice_production = runtime.do_ice_production ? 1 : 0; r.pad_ = 0;

    P3B200Infra in;
    in.dt = infrastructure.dt; in.it = infrastructure.it; in.ncol = nj; in.nlev = nk;
    in.predictNc = infrastructure.predictNc ? 1 : 0; in.prescribedCCN = infrastructure.prescribedCCN ? 1 : 0;
    in.zero_flux_top = 0;   // reference p3_main with pack size 1 leaves precip_*_flux(nk) to the caller
    // the history arrays are left to run_impl's zero-fill (eamxx_p3_run.cpp:27-29) unless the caller folds it in
    in.zero_history = zero_history ? 1 : 0;
    int64_t us = 0;
    // fused_preamble (an extension): run_impl's p3_preamble of every level inside p3_main's first kernel
    const int rc = fused_preamble
        ? p3b200_main_preamble(lookup_tables.handle.get(), &r, &f, &in, fused_preamble, stream, &us, nullptr)
        : p3b200_main(lookup_tables.handle.get(), &r, &f, &in, stream, &us, nullptr);
    if (rc) fail(rc, "p3_main");
    int32_t bits = 0;
    if (check_columns && p3b200_error_summary(lookup_tables.handle.get(), &bits) == 0 && bits)
      // the reference aborts inside the kernel (EKAT_KERNEL_REQUIRE_MSG: T<=0 / NaN, physics_saturation_impl.hpp:29,33)
      throw std::logic_error("p3_main: a column raised a physics error flag (bits " + std::to_string(bits) + ")");
    return static_cast<Int>(us);
  }

  // Off by default like the reference's release build (asserts compiled out): set to true to turn the
  // per-column error bits (T<=0, NaN, Co_max<0) into an exception after each p3_main.
  static inline bool check_columns = false;
};

}  // namespace p3
}  // namespace scream
#endif  // P3_B200_FUNCTIONS_HPP
