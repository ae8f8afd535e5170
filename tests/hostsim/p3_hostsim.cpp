// tests/hostsim — DEBUG/TEST BUILD ONLY.  Compiles the product's column driver
// (scream_b200/csrc/p3_column.cuh, the exact source the CUDA kernel instantiates) for the host with a
// 1-lane "warp" so that control flow, indexing and shared-row reuse can be checked against the oracle
// in a container that has no GPU.  It is never loaded by the product package, bench.py's GPU arm or
// the C ABI: libp3b200.so has no CPU path.
#include "../../scream_b200/csrc/p3_column.cuh"
#include "../../scream_b200/csrc/p3_host.hpp"

#include <cstdlib>
#include <cstring>
#include <vector>

namespace {
struct WarpSim {
  static constexpr int size = 1;
  int lane() const { return 0; }
  bool any(bool b) const { return b; }
  template <class R> R max_nonneg(R v) const { return v; }
  int min_int(int v) const { return v; }
  int max_int(int v) const { return v; }
  int or_int(int v) const { return v; }
  void sync() const {}
  template <class R> static void atomic_add(R* p, R v) { *p = *p + v; }
  static void atomic_or(int* p, int v) { *p |= v; }
  template <class R> R shfl_up1(R v) const { return v; }
  template <class R> R shfl_from(R v, int) const { return v; }
  template <class R> R rot_up1(R v) const { return v; }
};

template <typename Real>
int run(const P3B200Tables* t, void* const* fields, const int64_t* strides, int ncol, int nlev, double dt,
        int predictNc, int prescribedCCN, int zero_flux_top, const P3B200Runtime* rt, int* diag) {
  using namespace p3b200;
  std::vector<Real> ice(12000), coll(60000), vn(3000), vm(3000), revap(3000), vmn(6000);
  for (int j = 0; j < 10; ++j)
    for (int i = 0; i < 300; ++i) {
      vmn[size_t(j * 300 + i) * 2 + 0] = Real(t->vm_table_vals[i][j]);
      vmn[size_t(j * 300 + i) * 2 + 1] = Real(t->vn_table_vals[i][j]);
    }
  for (int i = 0; i < 12000; ++i) ice[i] = Real((&t->ice_table_vals[0][0][0][0])[i]);
  for (int i = 0; i < 60000; ++i) coll[i] = Real((&t->collect_table_vals[0][0][0][0][0])[i]);
  for (int i = 0; i < 3000; ++i) {
    vn[i] = Real((&t->vn_table_vals[0][0])[i]); vm[i] = Real((&t->vm_table_vals[0][0])[i]);
    revap[i] = Real((&t->revap_table_vals[0][0])[i]);
  }
  MainArgs<Real> a;
  for (int f = 0; f < F_NUM; ++f) { a.f[f] = static_cast<Real*>(fields[f]); a.stride[f] = strides[f]; }
  a.ncol = ncol; a.nk = nlev; a.dt = Real(dt); a.predictNc = predictNc; a.prescribedCCN = prescribedCCN;
  a.zero_flux_top = zero_flux_top; a.ustride = 0; a.ustride_flux = 0;
  a.zero_history = 0; a.pre_pmid = a.pre_pd = a.pre_pdd = a.pre_T = a.pre_cft = nullptr; a.wl_idx = nullptr; a.wl_off = nullptr; a.lvlflags = nullptr;
  a.tb = Tables<Real>{ice.data(), coll.data(), vn.data(), vm.data(), revap.data(), vmn.data()};
  typename Phys<Real>::Loc3 tj;
  Phys<Real>::lookup3_from_dum1(Real(rt ? rt->constant_mu_rain : 1.0), Real(0), tj);
  const Real* rain_tab = (tj.rdumjj - Real(tj.dumjj) == 0) ? vmn.data() + size_t(tj.dumjj - 1) * 600 : nullptr;
  P3B200Runtime r;
  if (rt) r = *rt; else runtime_defaults(r);
  a.ro = make_runtime<Real>(r);
  std::vector<int> own_diag;
  if (!diag) { own_diag.assign(size_t(ncol) * D_NUM, 0); diag = own_diag.data(); }
  a.diag = diag; a.queue = nullptr;
  WarpSim w;
  const char* ice_env = std::getenv("P3B200_HOSTSIM_ICE");
  const bool ice_rolled = !(ice_env && std::strcmp(ice_env, "unrolled") == 0);
  std::vector<Real> ice_state;
  // the three phases in the order the kernels run them (p3_kernel.cu: launch_main)
  for (int i = 0; i < ncol; ++i) PhaseA<Real>::run_column_host(a, i);
  for (int i = 0; i < ncol; ++i) {
    const int* d = diag + size_t(i) * D_NUM;
    if (d[D_QC_KTOP] >= 0) Sediment<Real, WarpSim, SP_CLOUD, p3b200::MAX_NLEV>::run(w, a, i, d[D_QC_KTOP], d[D_QC_KBOT]);
  }
  for (int i = 0; i < ncol; ++i) {
    const int* d = diag + size_t(i) * D_NUM;
    if (d[D_QR_KTOP] >= 0) Sediment<Real, WarpSim, SP_RAIN, p3b200::MAX_NLEV>::run(w, a, i, d[D_QR_KTOP], d[D_QR_KBOT], rain_tab);
  }
  for (int i = 0; i < ncol; ++i) {
    const int* d = diag + size_t(i) * D_NUM;
    if (d[D_QI_KTOP] < 0) continue;
    // the device runs tall ice windows through Sediment::run_rolled (window in shared memory, slots as a rolled loop)
    // and short ones through run(): both here, selected by P3B200_HOSTSIM_ICE=rolled|unrolled (default rolled)
    using SI = Sediment<Real, WarpSim, SP_ICE, p3b200::MAX_NLEV>;
    if (ice_rolled) { ice_state.resize(SI::rolled_state_size); SI::run_rolled(w, a, i, d[D_QI_KTOP], d[D_QI_KBOT], ice_state.data()); }
    else SI::run(w, a, i, d[D_QI_KTOP], d[D_QI_KBOT]);
  }
  for (int i = 0; i < ncol; ++i) {
    const bool full = diag[size_t(i) * D_NUM + D_HYDRO_PART2] == 1;
    for (int k = 0; k < nlev; ++k) {
      if (full) PhaseC<Real>::template level<false>(a, i, k);
      else PhaseC<Real>::template defaults<false>(a, i, k);
    }
  }
  return 0;
}
}  // namespace

extern "C" int p3sim_main(const P3B200Tables* t, int real_bytes, void* const* fields, const int64_t* strides,
                          int ncol, int nlev, double dt, int predictNc, int prescribedCCN, int zero_flux_top,
                          const P3B200Runtime* rt, int* diag) {
  if (real_bytes == 8) return run<double>(t, fields, strides, ncol, nlev, dt, predictNc, prescribedCCN, zero_flux_top, rt, diag);
  if (real_bytes == 4) return run<float>(t, fields, strides, ncol, nlev, dt, predictNc, prescribedCCN, zero_flux_top, rt, diag);
  return -1;
}
extern "C" int p3sim_tables(const char* blob, P3B200Tables* out) {
  int rc = p3b200::tables_read_blob(blob, "4.1.1", *out);
  if (rc) return rc;
  p3b200::tables_generate_rain(*out);
  return 0;
}
