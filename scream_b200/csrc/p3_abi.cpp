// p3_abi.cpp — the extern "C" surface of libp3b200.so declared in include/p3_b200.h.
//
// Host-only C++ (CUDA runtime API, no device code): owns the handle (device tables, integer
// diagnostics, work-queue counters, events, the host<->device staging pipeline) and forwards to
// the kernel TU (p3_kernel.cu).  There is no CPU implementation of p3_main behind this file:
// every compute entry point fails with P3B200_ECUDA when no CUDA device is usable.
#include "p3_host.hpp"
#include "p3_kernel.cuh"
#include "p3_checks.cuh"
#include "p3_prepost.cuh"

#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cfloat>
#include <limits>
#include <mutex>
#include <string>
#include <vector>

using namespace p3b200;

namespace {

thread_local std::string g_last_cuda_error;

int cuda_fail(cudaError_t e, const char* what) {
  g_last_cuda_error = std::string(what) + ": " + cudaGetErrorString(e);
  return P3B200_ECUDA;
}
#define CU(call)                                            \
  do {                                                      \
    cudaError_t e__ = (call);                               \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call);   \
  } while (0)

constexpr int NQUEUE = 64;        // ring of work-queue counter groups so p3_main calls in flight on different streams never share one
constexpr int QSTRIDE = 32;       // counters per group: one per sedimentation launch (3 species x NUM_SED_CLASSES), one per part2 group, pad
constexpr int NSLOT = 4;          // host pipeline depth (H2D / kernel / D2H in flight, one spare)

// events around and between the five launches of one p3_main: m[0] = start, m[EV_END] = end
struct EventPair { cudaEvent_t m[NUM_PHASE_MARKS] = {nullptr}; };
constexpr int EV_END = PHASE_MARK_END;

struct Slot {                     // one in-flight column chunk of p3b200_main_host
  cudaStream_t stream = nullptr;
  void* dev = nullptr;            // chunk arrays, all 2-D fields padded to nlev+1
  size_t dev_bytes = 0;
  int* diag = nullptr;
  size_t diag_cols = 0;
};

}  // namespace

struct p3b200_ctx {
  int device = 0;
  int real_bytes = 8;
  void* d_tables = nullptr;       // ice | coll | vn | vm | revap in run precision
  Tables<double> tbd{};
  Tables<float> tbf{};
  int* d_diag = nullptr;
  size_t diag_cols = 0;
  unsigned int* d_queue = nullptr;
  int next_queue = 0;
  // phase-A scratch (work lists, per-level flag bytes: p3_kernel.cuh ScratchLayout), grown on demand.  One copy per
  // handle: a call's classification pass waits (scratch_free) until the previous call's column finish has run, so
  // calls on different streams (the host pipeline's slots) never overlap in it.
  void* d_scratch = nullptr;
  size_t scratch_bytes = 0;
  cudaEvent_t scratch_free = nullptr;
  bool scratch_pending = false;
  const P3B200PreambleFields* pre_for_next = nullptr;   // p3b200_main_preamble: consumed by the next enqueue
  std::vector<EventPair> events;  // one pair per kernel launch of the last call
  size_t events_used = 0;
  int last_ncol = 0;
  bool last_diag_in_handle = false;
  Slot slots[NSLOT];
  LaunchConfig cfg{};
  int cfg_nk = -1, cfg_ncol = -1;
  unsigned long long* d_checks = nullptr;   // 2*CHK_NUM min/max keys + 2 location slots (property checks)
  double* d_cons = nullptr;       // column conservation check: mass | energy | rel_mass | rel_energy [cons_cols] + 4 results
  size_t cons_cols = 0;
  double* d_budget = nullptr;     // [budget_cols][4] per-column values + BUDGET_PARTS*4 partials + 4 totals
  size_t budget_cols = 0;
  // helper streams/events that let the three sedimentation species run side by side (p3_kernel.cuh: SideStreams)
  cudaStream_t side_stream[NUM_SIDE_STREAMS] = {};
  cudaEvent_t side_event[NUM_SIDE_STREAMS + 1] = {};
  bool sed_concurrent = false;    // what the last p3b200_main did
};

namespace {

template <typename Real>
int upload_tables(p3b200_ctx* h, const P3B200Tables& t) {
  const size_t n_ice = 12000, n_coll = 60000, n_v = 3000;
  const size_t total = n_ice + n_coll + 3 * n_v + 2 * n_v;
  std::vector<Real> host(total);
  size_t o = 0;
  const double* src = &t.ice_table_vals[0][0][0][0];
  for (size_t i = 0; i < n_ice; ++i) host[o + i] = Real(src[i]);
  o += n_ice;
  src = &t.collect_table_vals[0][0][0][0][0];
  for (size_t i = 0; i < n_coll; ++i) host[o + i] = Real(src[i]);
  o += n_coll;
  const double* vs[3] = {&t.vn_table_vals[0][0], &t.vm_table_vals[0][0], &t.revap_table_vals[0][0]};
  for (int v = 0; v < 3; ++v) {
    for (size_t i = 0; i < n_v; ++i) host[o + i] = Real(vs[v][i]);
    o += n_v;
  }
  // (vm, vn) interleaved, one slice per mu_r column: [10][300][2]
  for (int j = 0; j < 10; ++j)
    for (int i = 0; i < 300; ++i) {
      host[o + (size_t)(j * 300 + i) * 2 + 0] = Real(t.vm_table_vals[i][j]);
      host[o + (size_t)(j * 300 + i) * 2 + 1] = Real(t.vn_table_vals[i][j]);
    }
  CU(cudaMalloc(&h->d_tables, total * sizeof(Real)));
  CU(cudaMemcpy(h->d_tables, host.data(), total * sizeof(Real), cudaMemcpyHostToDevice));
  const Real* base = static_cast<const Real*>(h->d_tables);
  Tables<Real> tb{base, base + n_ice, base + n_ice + n_coll, base + n_ice + n_coll + n_v, base + n_ice + n_coll + 2 * n_v,
                  base + n_ice + n_coll + 3 * n_v};
  if constexpr (sizeof(Real) == 8) h->tbd = tb; else h->tbf = tb;
  return P3B200_OK;
}

int ensure_events(p3b200_ctx* h, size_t n) {
  while (h->events.size() < n) {
    EventPair p;
    for (int i = 0; i < NUM_PHASE_MARKS; ++i) CU(cudaEventCreate(&p.m[i]));
    h->events.push_back(p);
  }
  return P3B200_OK;
}

int validate(const P3B200Fields* f, const P3B200Infra* in) {
  if (!f || !in) return P3B200_EINVAL;
  if (in->ncol < 0 || in->nlev < 1 || in->nlev > P3B200_MAX_NLEV || !(in->dt > 0)) return P3B200_EINVAL;
  for (int i = 0; i < P3B200_F_NUM; ++i) {
    const bool optional = (i == P3B200_F_INV_QC_RELVAR) ||
                          (i == P3B200_F_NCCN && !in->prescribedCCN) ||
                          ((i == P3B200_F_NC_NUCEAT_TEND || i == P3B200_F_NI_ACTIVATED) &&
                           (in->prescribedCCN || !in->predictNc));
    if (!f->ptr[i] && !optional) return P3B200_EINVAL;
    if (f->ptr[i] && i < P3B200_F_PRECIP_LIQ_SURF) {
      const int lev = (i == P3B200_F_PRECIP_LIQ_FLUX || i == P3B200_F_PRECIP_ICE_FLUX) ? in->nlev + 1 : in->nlev;
      if (f->col_stride[i] < lev) return P3B200_EINVAL;
    }
  }
  return P3B200_OK;
}

// Enqueue one kernel launch over `ncol` columns of device arrays.  ev (optional) brackets it.
template <typename Real>
int enqueue(p3b200_ctx* h, const P3B200Runtime* runtime, const P3B200Fields* f, const P3B200Infra* in,
            int ncol, cudaStream_t stream, int* diag, const EventPair* ev, const P3B200PostambleFields* post = nullptr,
            unsigned* post_mask = nullptr, bool allow_side = false) {
  MainArgs<Real> a;
  for (int i = 0; i < F_NUM; ++i) { a.f[i] = static_cast<Real*>(f->ptr[i]); a.stride[i] = f->col_stride[i]; }
  // common column stride of the 2-D fields (flux arrays apart): enables the one-multiply addressing of the level kernels
  a.ustride = 0; a.ustride_flux = 0;
  {
    int64_t u = -1;
    bool same = true;
    for (int i = 0; i < F_PRECIP_LIQ_SURF; ++i) {
      if (!f->ptr[i] || i == F_PRECIP_LIQ_FLUX || i == F_PRECIP_ICE_FLUX) continue;
      if (u < 0) u = f->col_stride[i];
      else if (u != f->col_stride[i]) same = false;
    }
    if (same && u > 0 && f->col_stride[F_PRECIP_LIQ_FLUX] == f->col_stride[F_PRECIP_ICE_FLUX]) {
      a.ustride = u; a.ustride_flux = f->col_stride[F_PRECIP_LIQ_FLUX];
    }
  }
  a.ncol = ncol; a.nk = in->nlev; a.dt = Real(in->dt);
  a.predictNc = in->predictNc; a.prescribedCCN = in->prescribedCCN; a.zero_flux_top = in->zero_flux_top;
  a.zero_history = in->zero_history ? 1 : 0;
  if constexpr (sizeof(Real) == 8) a.tb = h->tbd; else a.tb = h->tbf;
  P3B200Runtime r;
  if (runtime) r = *runtime; else runtime_defaults(r);
  a.ro = make_runtime<Real>(r);
  a.pre_pmid = a.pre_pd = a.pre_pdd = a.pre_T = a.pre_cft = nullptr;
  if (const P3B200PreambleFields* pf = h->pre_for_next) {
    a.pre_pmid = static_cast<const Real*>(pf->pmid); a.pre_pd = static_cast<const Real*>(pf->pseudo_density);
    a.pre_pdd = static_cast<const Real*>(pf->pseudo_density_dry); a.pre_T = static_cast<const Real*>(pf->T_atm);
    a.pre_cft = static_cast<const Real*>(pf->cld_frac_t);
    if (a.ustride == 0 || a.ustride != pf->col_stride) return P3B200_EINVAL;
  }
  a.diag = diag;
  a.queue = h->d_queue + h->next_queue * QSTRIDE;
  h->next_queue = (h->next_queue + 1) % NQUEUE;
  if (h->cfg_nk != in->nlev || h->cfg_ncol != ncol) {
    const int rc = plan_launch<Real>(in->nlev, ncol, &h->cfg);
    if (rc) return cuda_fail((cudaError_t)rc, "plan_launch");
    h->cfg_nk = in->nlev; h->cfg_ncol = ncol;
  }
  CU(cudaMemsetAsync(a.queue, 0, QSTRIDE * sizeof(unsigned int), stream));
  {
    const ScratchLayout sl = scratch_layout(ncol, in->nlev, h->cfg.a_tile_cols);
    if (sl.total > h->scratch_bytes) {
      if (h->d_scratch) { CU(cudaDeviceSynchronize()); CU(cudaFree(h->d_scratch)); h->d_scratch = nullptr; h->scratch_bytes = 0; }
      const size_t want = sl.total + sl.total / 8;
      CU(cudaMalloc(&h->d_scratch, want));
      h->scratch_bytes = want;
    }
    if (!h->scratch_free) CU(cudaEventCreateWithFlags(&h->scratch_free, cudaEventDisableTiming));
    if (h->scratch_pending) CU(cudaStreamWaitEvent(stream, h->scratch_free, 0));
    char* base = static_cast<char*>(h->d_scratch);
    a.wl_idx = reinterpret_cast<unsigned short*>(base + sl.wl_idx);
    a.wl_off = reinterpret_cast<int*>(base + sl.wl_off);
    a.lvlflags = reinterpret_cast<unsigned char*>(base + sl.lvlflags);
  }
  PostArgs<Real> pa;
  if (post) pa = make_post_args<Real>(post, ncol, in->nlev, post_mask);
  SideStreams side;
  for (int i = 0; i < NUM_SIDE_STREAMS; ++i) { side.stream[i] = h->side_stream[i]; side.join[i] = h->side_event[i + 1]; }
  side.fork = h->side_event[0];
  const bool use_side = allow_side && h->side_stream[0] && !std::getenv("P3B200_SED_SERIAL");
  h->sed_concurrent = use_side;
  const int rc = launch_main<Real>(a, h->cfg, stream, ev ? reinterpret_cast<void* const*>(ev->m) : nullptr,
                                   post ? &pa : nullptr, use_side ? &side : nullptr, h->scratch_free);
  if (rc) return cuda_fail((cudaError_t)rc, "p3_main_kernel launch");
  h->scratch_pending = true;
  return P3B200_OK;
}

int enqueue_any(p3b200_ctx* h, const P3B200Runtime* runtime, const P3B200Fields* f, const P3B200Infra* in,
                int ncol, cudaStream_t stream, int* diag, const EventPair* ev, const P3B200PostambleFields* post = nullptr,
                unsigned* post_mask = nullptr, bool allow_side = false) {
  return h->real_bytes == 8 ? enqueue<double>(h, runtime, f, in, ncol, stream, diag, ev, post, post_mask, allow_side)
                            : enqueue<float>(h, runtime, f, in, ncol, stream, diag, ev, post, post_mask, allow_side);
}

}  // namespace

extern "C" {

int p3b200_abi_version(void) { return P3B200_ABI_VERSION; }

const char* p3b200_strerror(int code) {
  switch (code) {
    case P3B200_OK: return "ok";
    case P3B200_EINVAL: return "invalid argument";
    case P3B200_ECUDA: return "CUDA runtime error (see p3b200_last_cuda_error)";
    case P3B200_ENOMEM: return "out of memory";
    case P3B200_ETABLE: return "lookup table malformed or wrong version";
    case P3B200_ECOLUMN: return "a column raised a physics error flag";
    case P3B200_EINTERNAL: return "two evaluations of a property check disagree";
    case P3B200_ETOOBIG: return "size beyond the fixed maximum of an allocation-free entry point (see p3b200_last_cuda_error)";
    default: return "unknown error";
  }
}

const char* p3b200_last_cuda_error(void) { return g_last_cuda_error.c_str(); }

int p3b200_device_count(void) {
  int n = 0;
  const cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { cuda_fail(e, "cudaGetDeviceCount"); return P3B200_ECUDA; }
  return n;
}

void p3b200_runtime_defaults(P3B200Runtime* out) { if (out) runtime_defaults(*out); }

int p3b200_tables_read_text(const char* path, const char* expect_version, P3B200Tables* out) {
  if (!out) return P3B200_EINVAL;
  return tables_read_text(path, expect_version, *out);
}
int p3b200_tables_read_blob(const char* path, const char* expect_version, P3B200Tables* out) {
  if (!out) return P3B200_EINVAL;
  return tables_read_blob(path, expect_version, *out);
}
int p3b200_tables_generate_rain(P3B200Tables* out) {
  if (!out) return P3B200_EINVAL;
  tables_generate_rain(*out);
  return P3B200_OK;
}

int p3b200_create(const P3B200Tables* host_tables, int device, int real_bytes, p3b200_handle* out) {
  if (!host_tables || !out || (real_bytes != 4 && real_bytes != 8)) return P3B200_EINVAL;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
  if (device < 0 || device >= n) { g_last_cuda_error = "no such CUDA device"; return P3B200_ECUDA; }
  CU(cudaSetDevice(device));
  p3b200_ctx* h = new (std::nothrow) p3b200_ctx;
  if (!h) return P3B200_ENOMEM;
  h->device = device;
  h->real_bytes = real_bytes;
  int rc = real_bytes == 8 ? upload_tables<double>(h, *host_tables) : upload_tables<float>(h, *host_tables);
  if (rc == P3B200_OK) {
    e = cudaMalloc(&h->d_queue, NQUEUE * QSTRIDE * sizeof(unsigned int));
    if (e != cudaSuccess) rc = cuda_fail(e, "cudaMalloc queue");
  }
  if (rc == P3B200_OK) {
    for (int i = 0; i < NUM_SIDE_STREAMS && rc == P3B200_OK; ++i) {
      e = cudaStreamCreateWithFlags(&h->side_stream[i], cudaStreamNonBlocking);
      if (e != cudaSuccess) rc = cuda_fail(e, "cudaStreamCreate side");
    }
    for (int i = 0; i < NUM_SIDE_STREAMS + 1 && rc == P3B200_OK; ++i) {
      e = cudaEventCreateWithFlags(&h->side_event[i], cudaEventDisableTiming);
      if (e != cudaSuccess) rc = cuda_fail(e, "cudaEventCreate side");
    }
  }
  if (rc != P3B200_OK) { p3b200_destroy(h); return rc; }
  *out = h;
  return P3B200_OK;
}

int p3b200_destroy(p3b200_handle h) {
  if (!h) return P3B200_EINVAL;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (auto& p : h->events) for (int i = 0; i < NUM_PHASE_MARKS; ++i) if (p.m[i]) cudaEventDestroy(p.m[i]);
  for (auto& s : h->slots) {
    if (s.dev) cudaFree(s.dev);
    if (s.diag) cudaFree(s.diag);
    if (s.stream) cudaStreamDestroy(s.stream);
  }
  for (auto& st : h->side_stream) if (st) cudaStreamDestroy(st);
  for (auto& evn : h->side_event) if (evn) cudaEventDestroy(evn);
  if (h->d_tables) cudaFree(h->d_tables);
  if (h->d_diag) cudaFree(h->d_diag);
  if (h->d_queue) cudaFree(h->d_queue);
  if (h->d_scratch) cudaFree(h->d_scratch);
  if (h->scratch_free) cudaEventDestroy(h->scratch_free);
  if (h->d_budget) cudaFree(h->d_budget);
  if (h->d_cons) cudaFree(h->d_cons);
  if (h->d_checks) cudaFree(h->d_checks);
  delete h;
  return P3B200_OK;
}

static int main_impl(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* fields,
                     const P3B200Infra* infra, void* cuda_stream, int64_t* elapsed_us, int32_t* diag_dev,
                     const P3B200PostambleFields* post, unsigned* post_mask) {
  if (!h) return P3B200_EINVAL;
  int rc = validate(fields, infra);
  if (rc) return rc;
  CU(cudaSetDevice(h->device));
  if (elapsed_us) *elapsed_us = 0;
  h->events_used = 0;
  h->last_ncol = infra->ncol;
  h->last_diag_in_handle = (diag_dev == nullptr);
  if (infra->ncol == 0) return P3B200_OK;
  int* diag = diag_dev;
  if (!diag) {
    if (h->diag_cols < (size_t)infra->ncol) {
      if (h->d_diag) { CU(cudaFree(h->d_diag)); h->d_diag = nullptr; h->diag_cols = 0; }
      CU(cudaMalloc(&h->d_diag, (size_t)infra->ncol * D_NUM * sizeof(int)));
      h->diag_cols = infra->ncol;
    }
    diag = h->d_diag;
  }
  rc = ensure_events(h, 1);
  if (rc) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  rc = enqueue_any(h, runtime, fields, infra, infra->ncol, stream, diag, &h->events[0], post, post_mask, true);
  if (rc) return rc;
  h->events_used = 1;
  if (elapsed_us) {
    CU(cudaEventSynchronize(h->events[0].m[EV_END]));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, h->events[0].m[0], h->events[0].m[EV_END]));
    *elapsed_us = (int64_t)(ms * 1000.0f + 0.5f);
  }
  return P3B200_OK;
}

int p3b200_main(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* fields,
                const P3B200Infra* infra, void* cuda_stream, int64_t* elapsed_us, int32_t* diag_dev) {
  return main_impl(h, runtime, fields, infra, cuda_stream, elapsed_us, diag_dev, nullptr, nullptr);
}

int p3b200_main_preamble(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* fields,
                         const P3B200Infra* infra, const P3B200PreambleFields* pre, void* cuda_stream,
                         int64_t* elapsed_us, int32_t* diag_dev) {
  if (!h || !fields || !infra || !pre) return P3B200_EINVAL;
  if (pre->col_stride < infra->nlev) return P3B200_EINVAL;
  // the preamble's outputs are p3_main's inputs: they must be the same arrays, with one column stride
  struct Alias { const void* p; int f; };
  const Alias same[] = {{pre->qv, P3B200_F_QV}, {pre->qc, P3B200_F_QC}, {pre->nc, P3B200_F_NC}, {pre->qr, P3B200_F_QR},
                        {pre->nr, P3B200_F_NR}, {pre->qi, P3B200_F_QI}, {pre->qm, P3B200_F_QM}, {pre->ni, P3B200_F_NI},
                        {pre->bm, P3B200_F_BM}, {pre->qv_prev, P3B200_F_QV_PREV}, {pre->inv_exner, P3B200_F_INV_EXNER},
                        {pre->th_atm, P3B200_F_TH}, {pre->cld_frac_l, P3B200_F_CLD_FRAC_L}, {pre->cld_frac_i, P3B200_F_CLD_FRAC_I},
                        {pre->cld_frac_r, P3B200_F_CLD_FRAC_R}, {pre->dz, P3B200_F_DZ}};
  for (const Alias& a : same)
    if (!a.p || a.p != fields->ptr[a.f] || fields->col_stride[a.f] != pre->col_stride) return P3B200_EINVAL;
  if (!pre->pmid || !pre->pseudo_density || !pre->pseudo_density_dry || !pre->T_atm || !pre->cld_frac_t) return P3B200_EINVAL;
  // cld_frac_l and cld_frac_i receive the same values but are two arrays of p3_main: they must not alias an input
  h->pre_for_next = pre;
  const int rc = main_impl(h, runtime, fields, infra, cuda_stream, elapsed_us, diag_dev, nullptr, nullptr);
  h->pre_for_next = nullptr;
  return rc;
}

int p3b200_get_diag(p3b200_handle h, int32_t* host_out, int32_t ncol) {
  if (!h || !host_out || ncol < 0) return P3B200_EINVAL;
  if (!h->last_diag_in_handle || (size_t)ncol > h->diag_cols || ncol > h->last_ncol) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(host_out, h->d_diag, (size_t)ncol * D_NUM * sizeof(int), cudaMemcpyDeviceToHost));
  return P3B200_OK;
}

int p3b200_error_summary(p3b200_handle h, int32_t* or_of_error_bits) {
  if (!h || !or_of_error_bits) return P3B200_EINVAL;
  *or_of_error_bits = 0;
  if (!h->last_diag_in_handle || h->last_ncol == 0) return P3B200_OK;
  std::vector<int32_t> d((size_t)h->last_ncol * D_NUM);
  const int rc = p3b200_get_diag(h, d.data(), h->last_ncol);
  if (rc) return rc;
  int32_t acc = 0;
  for (int i = 0; i < h->last_ncol; ++i) acc |= d[(size_t)i * D_NUM + D_ERROR];
  *or_of_error_bits = acc;
  return P3B200_OK;
}

int p3b200_last_kernel_ms(p3b200_handle h, double* ms, int64_t* launches) {
  if (!h || !ms) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  double total = 0;
  for (size_t i = 0; i < h->events_used; ++i) {
    CU(cudaEventSynchronize(h->events[i].m[EV_END]));
    float t = 0;
    CU(cudaEventElapsedTime(&t, h->events[i].m[0], h->events[i].m[EV_END]));
    total += t;
  }
  *ms = total;
  if (launches) *launches = (int64_t)h->events_used;
  return P3B200_OK;
}

}  // extern "C"

namespace {
template <typename Real>
int budget_impl(p3b200_ctx* h, const P3B200Fields* f, const P3B200Infra* in, cudaStream_t stream, double* totals_dev) {
  MainArgs<Real> a{};
  for (int i = 0; i < F_NUM; ++i) { a.f[i] = static_cast<Real*>(f->ptr[i]); a.stride[i] = f->col_stride[i]; }
  a.ncol = in->ncol; a.nk = in->nlev; a.dt = Real(in->dt);
  double* col_vals = h->d_budget;
  double* partials = h->d_budget + h->budget_cols * 4;
  const int rc = launch_budget<Real>(a, col_vals, partials, BUDGET_PARTS, totals_dev, stream);
  if (rc) return cuda_fail((cudaError_t)rc, "budget launches");
  return P3B200_OK;
}
}  // namespace

extern "C" {

int p3b200_water_energy_totals(p3b200_handle h, const P3B200Fields* fields, const P3B200Infra* infra,
                               void* cuda_stream, double* totals4_dev, double* totals4_host) {
  if (!h || !fields || !infra || infra->ncol < 0 || infra->nlev < 1) return P3B200_EINVAL;
  static const int need[] = {P3B200_F_QV, P3B200_F_QC, P3B200_F_QI, P3B200_F_QR, P3B200_F_TH, P3B200_F_INV_EXNER,
                             P3B200_F_DPRES, P3B200_F_PRECIP_LIQ_SURF, P3B200_F_PRECIP_ICE_SURF};
  for (int i : need) if (!fields->ptr[i]) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  if (h->budget_cols < (size_t)infra->ncol || !h->d_budget) {
    if (h->d_budget) { CU(cudaDeviceSynchronize()); CU(cudaFree(h->d_budget)); h->d_budget = nullptr; }
    const size_t cols = infra->ncol > 0 ? (size_t)infra->ncol : 1;
    CU(cudaMalloc(&h->d_budget, (cols * 4 + BUDGET_PARTS * 4 + 4) * sizeof(double)));
    h->budget_cols = cols;
  }
  double* own = h->d_budget + h->budget_cols * 4 + BUDGET_PARTS * 4;
  double* dst = totals4_dev ? totals4_dev : own;
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  const int rc = h->real_bytes == 8 ? budget_impl<double>(h, fields, infra, stream, dst)
                                    : budget_impl<float>(h, fields, infra, stream, dst);
  if (rc) return rc;
  if (totals4_host) {
    CU(cudaMemcpyAsync(totals4_host, dst, 4 * sizeof(double), cudaMemcpyDeviceToHost, stream));
    CU(cudaStreamSynchronize(stream));
  }
  return P3B200_OK;
}

// No allocation on this path: the scratch is the file-scope device array of p3_budget.cu (launch_budget_static), the
// result goes to the caller's device buffer, and nothing is copied back or synchronised.
int p3b200_water_energy_totals_static(p3b200_handle h, const P3B200Fields* fields, const P3B200Infra* infra,
                                      void* cuda_stream, double* totals4_dev) {
  if (!h || !fields || !infra || !totals4_dev || infra->ncol < 0 || infra->nlev < 1) return P3B200_EINVAL;
  if (infra->ncol > P3B200_STATIC_BUDGET_MAX_COLS) {
    g_last_cuda_error = "p3b200_water_energy_totals_static: ncol = " + std::to_string(infra->ncol) +
                        " exceeds P3B200_STATIC_BUDGET_MAX_COLS = " + std::to_string(P3B200_STATIC_BUDGET_MAX_COLS) +
                        " (use p3b200_water_energy_totals, or split the columns)";
    return P3B200_ETOOBIG;
  }
  static const int need[] = {P3B200_F_QV, P3B200_F_QC, P3B200_F_QI, P3B200_F_QR, P3B200_F_TH, P3B200_F_INV_EXNER,
                             P3B200_F_DPRES, P3B200_F_PRECIP_LIQ_SURF, P3B200_F_PRECIP_ICE_SURF};
  for (int i : need) if (!fields->ptr[i]) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  int rc;
  if (h->real_bytes == 8) {
    MainArgs<double> a{};
    for (int i = 0; i < F_NUM; ++i) { a.f[i] = static_cast<double*>(fields->ptr[i]); a.stride[i] = fields->col_stride[i]; }
    a.ncol = infra->ncol; a.nk = infra->nlev; a.dt = infra->dt;
    rc = launch_budget_static<double>(a, totals4_dev, cuda_stream);
  } else {
    MainArgs<float> a{};
    for (int i = 0; i < F_NUM; ++i) { a.f[i] = static_cast<float*>(fields->ptr[i]); a.stride[i] = fields->col_stride[i]; }
    a.ncol = infra->ncol; a.nk = infra->nlev; a.dt = float(infra->dt);
    rc = launch_budget_static<float>(a, totals4_dev, cuda_stream);
  }
  if (rc) return cuda_fail((cudaError_t)rc, "static budget launches");
  return P3B200_OK;
}

int p3b200_preamble(p3b200_handle h, const P3B200PreambleFields* f, int32_t ncol, int32_t nlev, void* cuda_stream) {
  if (!h || !f || ncol < 0 || nlev < 1 || f->col_stride < nlev) return P3B200_EINVAL;
  const void* need[] = {f->pmid, f->pseudo_density, f->pseudo_density_dry, f->T_atm, f->cld_frac_t, f->qv, f->qc, f->nc,
                        f->qr, f->nr, f->qi, f->qm, f->ni, f->bm, f->qv_prev, f->inv_exner, f->th_atm, f->cld_frac_l,
                        f->cld_frac_i, f->cld_frac_r, f->dz};
  for (const void* p : need) if (!p) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  const int rc = h->real_bytes == 8 ? preamble_from_abi<double>(f, ncol, nlev, cuda_stream)
                                    : preamble_from_abi<float>(f, ncol, nlev, cuda_stream);
  if (rc) return cuda_fail((cudaError_t)rc, "p3_preamble launch");
  return P3B200_OK;
}

static int postamble_impl(p3b200_handle h, const P3B200PostambleFields* f, int32_t ncol, int32_t nlev, void* cuda_stream,
                          unsigned* check_mask) {
  if (!h || !f || ncol < 0 || nlev < 1 || f->col_stride < nlev || !(f->dt > 0)) return P3B200_EINVAL;
  const void* need[] = {f->pmid, f->pseudo_density, f->pseudo_density_dry, f->th_atm, f->T_atm, f->qv, f->qc, f->nc,
                        f->qr, f->nr, f->qi, f->qm, f->ni, f->bm, f->diag_eff_radius_qc, f->diag_eff_radius_qi,
                        f->diag_eff_radius_qr, f->T_prev, f->qv_prev, f->precip_liq_surf_flux, f->precip_ice_surf_flux,
                        f->precip_liq_surf_mass, f->precip_ice_surf_mass};
  for (const void* p : need) if (!p) return P3B200_EINVAL;
  const int nflux = (f->vapor_flux != nullptr) + (f->water_flux != nullptr) + (f->ice_flux != nullptr) + (f->heat_flux != nullptr);
  if (nflux != 0 && nflux != 4) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  const int rc = h->real_bytes == 8 ? postamble_from_abi<double>(f, ncol, nlev, cuda_stream, check_mask)
                                    : postamble_from_abi<float>(f, ncol, nlev, cuda_stream, check_mask);
  if (rc) return cuda_fail((cudaError_t)rc, "p3_postamble launch");
  return P3B200_OK;
}

int p3b200_postamble(p3b200_handle h, const P3B200PostambleFields* f, int32_t ncol, int32_t nlev, void* cuda_stream) {
  return postamble_impl(h, f, ncol, nlev, cuda_stream, nullptr);
}

// ---- property checks (eamxx_p3_process_interface.cpp:226-243) ----
namespace {
static_assert((int)CHK_NUM == (int)P3B200_NUM_CHECKS, "check tables out of step");
static_assert(STATIC_BUDGET_MAX_COLS == P3B200_STATIC_BUDGET_MAX_COLS, "static budget scratch out of step with the header");
struct CheckSpec { double lb, ub; bool can_repair; bool column_field; };
inline CheckSpec check_spec(int i) { return {CHK_LB[i], CHK_UB[i], CHK_CAN_REPAIR[i], CHK_COLUMN_FIELD[i]}; }

int checks_prepare(p3b200_handle h, cudaStream_t stream) {
  if (!h->d_checks) CU(cudaMalloc(&h->d_checks, CHK_STATE_WORDS * sizeof(unsigned long long)));
  const int rc = checks_reset(h->d_checks, h->real_bytes, stream);
  if (rc) return cuda_fail((cudaError_t)rc, "p3_checks reset");
  return P3B200_OK;
}

// check_impl :202-226 on the reduced (min, max), then the locations of the failing ones
int checks_finish(p3b200_handle h, const void* const* field_ptrs, int64_t stride, int32_t ncol, int32_t nlev,
                  P3B200CheckResult* out, cudaStream_t stream) {
  unsigned long long keys[2 * P3B200_NUM_CHECKS];
  CU(cudaMemcpyAsync(keys, h->d_checks, sizeof(keys), cudaMemcpyDeviceToHost, stream));
  CU(cudaStreamSynchronize(stream));
  const double s_max = std::numeric_limits<double>::max();
  for (int i = 0; i < P3B200_NUM_CHECKS; ++i) {
    const CheckSpec c = check_spec(i);
    P3B200CheckResult& r = out[i];
    r.lb = c.lb; r.ub = c.ub;
    r.min_val = chk_decode(keys[2 * i]);
    r.max_val = chk_decode(keys[2 * i + 1]);
    r.min_loc = r.max_loc = -1;
    r.pass_lower = r.pass_upper = 1;
    r.reserved = 0;
    const double lb_rep = c.can_repair ? -s_max : c.lb, ub_rep = c.can_repair ? s_max : c.ub;
    if (r.min_val >= c.lb && r.max_val <= c.ub) {
      r.result = P3B200_CHECK_PASS;
      continue;
    } else if (r.min_val < lb_rep || r.max_val > ub_rep) {
      r.result = P3B200_CHECK_FAIL;
      if (r.min_val < lb_rep) r.pass_lower = 0;
      if (r.max_val > ub_rep) r.pass_upper = 0;
    } else {
      r.result = P3B200_CHECK_REPAIRABLE;
      if (r.min_val < c.lb) r.pass_lower = 0;
      if (r.max_val > c.ub) r.pass_upper = 0;
    }
    // locations (only printed by the reference when the check did not pass)
    unsigned long long* loc_dev = h->d_checks + CHK_STATE_LOC;
    CU(cudaMemsetAsync(loc_dev, 0xff, 2 * sizeof(unsigned long long), stream));
    const int nk = c.column_field ? 1 : nlev;
    const int64_t st = c.column_field ? 1 : stride;
    for (int side = 0; side < 2; ++side) {
      const double target = side ? r.max_val : r.min_val;
      const int rc = h->real_bytes == 8 ? find_loc<double>(field_ptrs[i], st, ncol, nk, target, loc_dev + side, stream)
                                        : find_loc<float>(field_ptrs[i], st, ncol, nk, target, loc_dev + side, stream);
      if (rc) return cuda_fail((cudaError_t)rc, "p3_find_loc launch");
    }
    unsigned long long locs[2];
    CU(cudaMemcpyAsync(locs, loc_dev, sizeof(locs), cudaMemcpyDeviceToHost, stream));
    CU(cudaStreamSynchronize(stream));
    r.min_loc = locs[0] == ~0ull ? -1 : (int64_t)locs[0];
    r.max_loc = locs[1] == ~0ull ? -1 : (int64_t)locs[1];
  }
  return P3B200_OK;
}
}  // namespace

static int property_checks_impl(p3b200_handle h, const P3B200CheckFields* f, int32_t ncol, int32_t nlev,
                                P3B200CheckResult* out, void* cuda_stream) {
  if (!h || !f || !out || ncol < 0 || nlev < 1 || f->col_stride < nlev) return P3B200_EINVAL;
  const void* ptrs[P3B200_NUM_CHECKS] = {f->T_mid, f->qv, f->qc, f->qi, f->qr, f->qm, f->nc, f->nr, f->ni, f->bm,
                                         f->precip_liq_surf_mass, f->precip_ice_surf_mass, f->eff_radius_qc,
                                         f->eff_radius_qi, f->eff_radius_qr};
  for (const void* p : ptrs) if (!p) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  int rc = checks_prepare(h, stream);
  if (rc) return rc;
  rc = h->real_bytes == 8 ? checks_from_abi<double>(f, ncol, nlev, cuda_stream, h->d_checks)
                          : checks_from_abi<float>(f, ncol, nlev, cuda_stream, h->d_checks);
  if (rc) return cuda_fail((cudaError_t)rc, "p3_checks launch");
  return checks_finish(h, ptrs, f->col_stride, ncol, nlev, out, stream);
}

int p3b200_property_checks(p3b200_handle h, const P3B200CheckFields* f, int32_t ncol, int32_t nlev,
                           P3B200CheckResult* out, void* cuda_stream) {
  return property_checks_impl(h, f, ncol, nlev, out, cuda_stream);
}

// Read the violation mask a fused postamble left behind and turn it into the 15 results: all Pass when it is zero,
// else the full reduction over the fields the postamble has just written (extreme values + locations).
static int checks_from_mask(p3b200_handle h, const P3B200PostambleFields* f, int32_t ncol, int32_t nlev,
                            P3B200CheckResult* out, void* cuda_stream) {
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  unsigned* mask_dev = reinterpret_cast<unsigned*>(h->d_checks + CHK_STATE_MASK);
  unsigned mask = 0;
  CU(cudaMemcpyAsync(&mask, mask_dev, sizeof(mask), cudaMemcpyDeviceToHost, stream));
  CU(cudaStreamSynchronize(stream));
  if (mask == 0) {
    // every check passed: the reference's Pass result carries no extreme values (check_impl :229-234)
    const double nan = std::numeric_limits<double>::quiet_NaN();
    for (int i = 0; i < P3B200_NUM_CHECKS; ++i)
      out[i] = P3B200CheckResult{CHK_LB[i], CHK_UB[i], nan, nan, -1, -1, P3B200_CHECK_PASS, 1, 1, 0};
    return P3B200_OK;
  }
  // some check did not pass: the failure message needs extreme values and locations -> the full reduction over
  // the fields the postamble has just written
  P3B200CheckFields c;
  c.T_mid = f->T_atm; c.qv = f->qv; c.qc = f->qc; c.qi = f->qi; c.qr = f->qr; c.qm = f->qm; c.nc = f->nc; c.nr = f->nr;
  c.ni = f->ni; c.bm = f->bm; c.precip_liq_surf_mass = f->precip_liq_surf_mass;
  c.precip_ice_surf_mass = f->precip_ice_surf_mass; c.eff_radius_qc = f->diag_eff_radius_qc;
  c.eff_radius_qi = f->diag_eff_radius_qi; c.eff_radius_qr = f->diag_eff_radius_qr; c.col_stride = f->col_stride;
  const int rc2 = property_checks_impl(h, &c, ncol, nlev, out, cuda_stream);
  if (rc2) return rc2;
  for (int i = 0; i < P3B200_NUM_CHECKS; ++i) {   // the two evaluations must agree on Pass / not-Pass
    const bool flagged = (mask >> (2 * i)) & 3u;
    if (flagged != (out[i].result != P3B200_CHECK_PASS)) return P3B200_EINTERNAL;
  }
  return P3B200_OK;
}

int p3b200_postamble_checked(p3b200_handle h, const P3B200PostambleFields* f, int32_t ncol, int32_t nlev,
                             P3B200CheckResult* out, void* cuda_stream) {
  if (!h || !f || !out) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  if (!h->d_checks) CU(cudaMalloc(&h->d_checks, CHK_STATE_WORDS * sizeof(unsigned long long)));
  unsigned* mask_dev = reinterpret_cast<unsigned*>(h->d_checks + CHK_STATE_MASK);
  CU(cudaMemsetAsync(mask_dev, 0, sizeof(unsigned), stream));
  const int rc = postamble_impl(h, f, ncol, nlev, cuda_stream, mask_dev);
  if (rc) return rc;
  return checks_from_mask(h, f, ncol, nlev, out, cuda_stream);
}

// p3_main with p3_postamble fused into the tail of phase C (and, with checks != NULL, the property-check mask).
int p3b200_main_postamble(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* fields,
                          const P3B200Infra* infra, const P3B200PostambleFields* post, P3B200CheckResult* checks,
                          void* cuda_stream, int64_t* elapsed_us, int32_t* diag_dev) {
  if (!h || !fields || !infra || !post) return P3B200_EINVAL;
  if (!(post->dt > 0) || post->col_stride < infra->nlev) return P3B200_EINVAL;
  // the tail works on the arrays p3_main has just written: they must be the same arrays, with one column stride
  struct Alias { const void* p; int f; };
  const Alias same[] = {{post->th_atm, P3B200_F_TH}, {post->qv, P3B200_F_QV}, {post->qc, P3B200_F_QC}, {post->nc, P3B200_F_NC},
                        {post->qr, P3B200_F_QR}, {post->nr, P3B200_F_NR}, {post->qi, P3B200_F_QI}, {post->qm, P3B200_F_QM},
                        {post->ni, P3B200_F_NI}, {post->bm, P3B200_F_BM},
                        {post->diag_eff_radius_qc, P3B200_F_DIAG_EFF_RADIUS_QC},
                        {post->diag_eff_radius_qi, P3B200_F_DIAG_EFF_RADIUS_QI},
                        {post->diag_eff_radius_qr, P3B200_F_DIAG_EFF_RADIUS_QR}};
  for (const Alias& a : same)
    if (!a.p || a.p != fields->ptr[a.f] || fields->col_stride[a.f] != post->col_stride) return P3B200_EINVAL;
  if (post->precip_liq_surf_flux != fields->ptr[P3B200_F_PRECIP_LIQ_SURF] ||
      post->precip_ice_surf_flux != fields->ptr[P3B200_F_PRECIP_ICE_SURF]) return P3B200_EINVAL;
  const void* need[] = {post->pmid, post->pseudo_density, post->pseudo_density_dry, post->T_atm, post->T_prev,
                        post->qv_prev, post->precip_liq_surf_mass, post->precip_ice_surf_mass};
  for (const void* p : need) if (!p) return P3B200_EINVAL;
  const int nflux = (post->vapor_flux != nullptr) + (post->water_flux != nullptr) + (post->ice_flux != nullptr) +
                    (post->heat_flux != nullptr);
  if (nflux != 0 && nflux != 4) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  unsigned* mask_dev = nullptr;
  if (checks) {
    if (!h->d_checks) CU(cudaMalloc(&h->d_checks, CHK_STATE_WORDS * sizeof(unsigned long long)));
    mask_dev = reinterpret_cast<unsigned*>(h->d_checks + CHK_STATE_MASK);
    CU(cudaMemsetAsync(mask_dev, 0, sizeof(unsigned), stream));
  }
  const int rc = main_impl(h, runtime, fields, infra, cuda_stream, elapsed_us, diag_dev, post, mask_dev);
  if (rc || !checks) return rc;
  return checks_from_mask(h, post, infra->ncol, infra->nlev, checks, cuda_stream);
}

int p3b200_repair_qv(p3b200_handle h, void* qv, int64_t col_stride, int32_t ncol, int32_t nlev, void* cuda_stream) {
  if (!h || !qv || ncol < 0 || nlev < 1 || col_stride < nlev) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  const CheckSpec c = check_spec(P3B200_CHECK_QV);
  const int rc = h->real_bytes == 8 ? clip_field<double>(qv, col_stride, ncol, nlev, c.lb, c.ub, cuda_stream)
                                    : clip_field<float>(qv, col_stride, ncol, nlev, c.lb, c.ub, cuda_stream);
  if (rc) return cuda_fail((cudaError_t)rc, "p3_clip launch");
  return P3B200_OK;
}

int p3b200_last_phase_ms(p3b200_handle h, double* ms6) {
  if (!h || !ms6) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  for (int p = 0; p < 6; ++p) ms6[p] = 0;
  // marks: 0 start, 1 phase A done, 2/3/4 cloud/rain/ice done, 5 joined, 6 end (p3_kernel.cuh)
  const bool conc = h->sed_concurrent;
  const int from[6] = {0, 1, conc ? 1 : 2, conc ? 1 : 3, 5, 1};
  const int to[6]   = {1, 2, 3, 4, 6, 5};
  for (size_t i = 0; i < h->events_used; ++i) {
    CU(cudaEventSynchronize(h->events[i].m[EV_END]));
    for (int p = 0; p < 6; ++p) {
      float t = 0;
      CU(cudaEventElapsedTime(&t, h->events[i].m[from[p]], h->events[i].m[to[p]]));
      ms6[p] += t;
    }
  }
  return P3B200_OK;
}

}  // extern "C"
namespace {
template <typename Real>
ConsArgs<Real> make_cons_args(const P3B200ConservationFields* f, int ncol, int nlev, double dt) {
  ConsArgs<Real> a;
  auto p = [](const void* v) { return static_cast<const Real*>(v); };
  a.pseudo_density = p(f->pseudo_density); a.T_mid = p(f->T_mid); a.horiz_winds = p(f->horiz_winds);
  a.qv = p(f->qv); a.qc = p(f->qc); a.qi = p(f->qi); a.qr = p(f->qr); a.ps = p(f->ps); a.phis = p(f->phis);
  a.vapor_flux = p(f->vapor_flux); a.water_flux = p(f->water_flux); a.ice_flux = p(f->ice_flux); a.heat_flux = p(f->heat_flux);
  a.col_stride = f->col_stride ? f->col_stride : nlev;
  a.wind_comp_stride = f->wind_comp_stride ? f->wind_comp_stride : nlev;
  a.wind_col_stride = f->wind_col_stride ? f->wind_col_stride : 2 * (int64_t)nlev;
  a.ncol = ncol; a.nk = nlev; a.dt = Real(dt);
  return a;
}
int cons_run(p3b200_ctx* h, const P3B200ConservationFields* f, int ncol, int nlev, double dt, bool check, void* stream) {
  if (!f || ncol < 0 || nlev < 1) return P3B200_EINVAL;
  if (!f->pseudo_density || !f->T_mid || !f->horiz_winds || !f->qv || !f->qc || !f->qi || !f->qr || !f->ps || !f->phis)
    return P3B200_EINVAL;
  if (check && (!f->vapor_flux || !f->water_flux || !f->ice_flux || !f->heat_flux)) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  if (check && (h->cons_cols < (size_t)ncol || !h->d_cons)) return P3B200_EINVAL;    // no matching _begin
  if (h->cons_cols < (size_t)ncol || !h->d_cons) {
    if (h->d_cons) { CU(cudaDeviceSynchronize()); CU(cudaFree(h->d_cons)); h->d_cons = nullptr; }
    CU(cudaMalloc(&h->d_cons, ((size_t)ncol * 4 + 4) * sizeof(double)));
    h->cons_cols = (size_t)ncol;
  }
  double* base = h->d_cons;
  const size_t n = h->cons_cols;
  const int rc = h->real_bytes == 8
      ? launch_conservation<double>(make_cons_args<double>(f, ncol, nlev, dt), check, base, base + n, base + 2 * n, base + 3 * n, base + 4 * n, stream)
      : launch_conservation<float>(make_cons_args<float>(f, ncol, nlev, dt), check, base, base + n, base + 2 * n, base + 3 * n, base + 4 * n, stream);
  if (rc) return cuda_fail((cudaError_t)rc, "p3_column_conservation launch");
  return P3B200_OK;
}
}  // namespace
extern "C" {

int p3b200_column_conservation_begin(p3b200_handle h, const P3B200ConservationFields* f, int32_t ncol, int32_t nlev,
                                     void* cuda_stream) {
  if (!h) return P3B200_EINVAL;
  return cons_run(h, f, ncol, nlev, 0.0, false, cuda_stream);
}

int p3b200_column_conservation_check(p3b200_handle h, const P3B200ConservationFields* f, int32_t ncol, int32_t nlev,
                                     double dt, double mass_tol, double energy_tol, P3B200ConservationResult* result,
                                     void* cuda_stream) {
  if (!h || !result || !(dt == dt)) return P3B200_EINVAL;    // the reference requires dt to be set (:118-119)
  const int rc = cons_run(h, f, ncol, nlev, dt, true, cuda_stream);
  if (rc) return rc;
  double out[4] = {0, -1, 0, -1};
  if (ncol > 0) {
    CU(cudaMemcpyAsync(out, h->d_cons + 4 * h->cons_cols, sizeof(out), cudaMemcpyDeviceToHost, (cudaStream_t)cuda_stream));
    CU(cudaStreamSynchronize((cudaStream_t)cuda_stream));
  } else {
    out[0] = out[2] = h->real_bytes == 4 ? -(double)FLT_MAX : -DBL_MAX;
  }
  result->mass_rel_err = out[0]; result->mass_col = (int64_t)out[1];
  result->energy_rel_err = out[2]; result->energy_col = (int64_t)out[3];
  result->mass_below_tol = out[0] < mass_tol; result->energy_below_tol = out[2] < energy_tol;     // :202-203
  result->pass = result->mass_below_tol && result->energy_below_tol;
  result->reserved = 0;
  return P3B200_OK;
}

int p3b200_column_conservation_get(p3b200_handle h, int32_t ncol, double* mass, double* energy, double* rel_mass,
                                   double* rel_energy) {
  if (!h || ncol < 0 || (size_t)ncol > h->cons_cols || !h->d_cons) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  double* dst[4] = {mass, energy, rel_mass, rel_energy};
  for (int i = 0; i < 4; ++i)
    if (dst[i]) CU(cudaMemcpy(dst[i], h->d_cons + (size_t)i * h->cons_cols, (size_t)ncol * sizeof(double), cudaMemcpyDeviceToHost));
  return P3B200_OK;
}

int p3b200_last_phase_a_ms(p3b200_handle h, double* ms3) {
  if (!h || !ms3) return P3B200_EINVAL;
  CU(cudaSetDevice(h->device));
  for (int p = 0; p < 3; ++p) ms3[p] = 0;
  // marks: 0 start, 7 classification pass done, 8 part2 launches done, 1 column finish done (p3_kernel.cuh)
  const int from[3] = {0, 7, 8};
  const int to[3]   = {7, 8, 1};
  for (size_t i = 0; i < h->events_used; ++i) {
    CU(cudaEventSynchronize(h->events[i].m[EV_END]));
    for (int p = 0; p < 3; ++p) {
      float t = 0;
      CU(cudaEventElapsedTime(&t, h->events[i].m[from[p]], h->events[i].m[to[p]]));
      ms3[p] += t;
    }
  }
  return P3B200_OK;
}

int p3b200_last_sed_concurrent(p3b200_handle h) { return (h && h->sed_concurrent) ? 1 : 0; }

}  // extern "C"

#include "p3_abi_host.inc"
