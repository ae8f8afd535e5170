/*
 * p3_b200.h — C ABI of the B200-native P3 microphysics column driver (libp3b200.so).
 *
 * This is the drop-in boundary for ONE hot path of EAMxx/SCREAM: p3::Functions<S,D>::p3_main
 * (reference: components/eamxx/src/physics/p3/p3_functions.hpp:1356-1369, implementation
 * p3/impl/p3_main_impl.hpp:76-372) and its callers P3Microphysics::run_impl
 * (p3/eamxx_p3_run.cpp:5-46) and the test bridge p3_main_f (p3/p3_functions_f90.cpp:1888-2103).
 *
 * Rules of the boundary
 *   - plain C: pointers, sizes, PODs; no C++/torch/Kokkos types; no exceptions cross it;
 *   - every entry point returns 0 on success or a negative P3B200_E* code; p3b200_strerror() names it;
 *   - p3b200_main takes DEVICE pointers only.  There is no CPU fallback: without a CUDA device
 *     p3b200_create fails with P3B200_ECUDA;
 *   - the caller owns all field memory; the handle owns the lookup tables, the per-column integer
 *     diagnostics and the work-queue counter;
 *   - field arrays are (ncol, nlev) with the LEVEL index contiguous (Kokkos LayoutRight views of the
 *     reference, p3_functions.hpp:161-276) and a per-field column stride in ELEMENTS (tracer
 *     sub-views of a bundled array have stride > nlev, share/field/field_impl.hpp:880-909);
 *     level 0 is the model top (p3_main_impl.hpp:97-99: kdir=-1, ktop=0, kbot=nk-1).
 */
#ifndef P3_B200_H
#define P3_B200_H

#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define P3B200_ABI_VERSION 2

/* error codes */
#define P3B200_OK        0
#define P3B200_EINVAL   -1   /* bad argument (null pointer, nlev out of range, real_bytes not 4/8 ...) */
#define P3B200_ECUDA    -2   /* CUDA runtime error (no device, launch failure ...); see p3b200_last_cuda_error */
#define P3B200_ENOMEM   -3   /* host or device allocation failed */
#define P3B200_ETABLE   -4   /* lookup-table blob/text malformed or wrong version (p3_table_ice_impl.hpp:40-41) */
#define P3B200_ECOLUMN  -5   /* a column raised a physics error flag (T<=0 / NaN / Co_max<0), see diag[P3B200_D_ERROR] */
#define P3B200_EINTERNAL -6  /* two evaluations of the same property check disagree (fields modified concurrently?) */
#define P3B200_ETOOBIG  -7   /* more columns than an allocation-free entry point's fixed scratch holds */

#define P3B200_MAX_NLEV 256  /* levels per column supported by the warp-per-column layout (8 per lane) */

/* Table extents: P3C, p3_functions.hpp:30-52 */
#define P3B200_DENSIZE 5
#define P3B200_RIMSIZE 4
#define P3B200_ISIZE 50
#define P3B200_ICE_TABLE_SIZE 12
#define P3B200_RCOLLSIZE 30
#define P3B200_COLLECT_TABLE_SIZE 2
#define P3B200_VTABLE_DIM0 300
#define P3B200_VTABLE_DIM1 10
#define P3B200_MU_R_TABLE_DIM 150
#define P3B200_DNUSIZE 16

/* P3LookupTables (p3_functions.hpp:279-290), host side, always double; converted to the run
 * precision on upload.  collect_table_vals holds log10 of the file values exactly as
 * init_kokkos_ice_lookup_tables stores them (p3_table_ice_impl.hpp:63). */
typedef struct P3B200Tables {
  double mu_r_table_vals[P3B200_MU_R_TABLE_DIM];
  double vn_table_vals[P3B200_VTABLE_DIM0][P3B200_VTABLE_DIM1];
  double vm_table_vals[P3B200_VTABLE_DIM0][P3B200_VTABLE_DIM1];
  double revap_table_vals[P3B200_VTABLE_DIM0][P3B200_VTABLE_DIM1];
  double ice_table_vals[P3B200_DENSIZE][P3B200_RIMSIZE][P3B200_ISIZE][P3B200_ICE_TABLE_SIZE];
  double collect_table_vals[P3B200_DENSIZE][P3B200_RIMSIZE][P3B200_ISIZE][P3B200_RCOLLSIZE][P3B200_COLLECT_TABLE_SIZE];
  double dnu_table_vals[P3B200_DNUSIZE];
} P3B200Tables;

/* P3Runtime (p3_functions.hpp:112-158): the 19 scalar knobs + do_ice_production, same names,
 * same defaults (p3b200_runtime_defaults). */
typedef struct P3B200Runtime {
  double max_total_ni;
  double autoconversion_prefactor;
  double autoconversion_qc_exponent;
  double autoconversion_nc_exponent;
  double autoconversion_radius;
  double accretion_prefactor;
  double accretion_qc_exponent;
  double accretion_qr_exponent;
  double rain_selfcollection_prefactor;
  double rain_selfcollection_breakup_diameter;
  double constant_mu_rain;
  double spa_ccn_to_nc_factor;
  double cldliq_to_ice_collection_factor;
  double rain_to_ice_collection_factor;
  double min_rime_rho;
  double max_rime_rho;
  double immersion_freezing_exponent;
  double deposition_nucleation_exponent;
  double ice_sedimentation_factor;
  int32_t do_ice_production;
  int32_t pad_;
} P3B200Runtime;

/* Flattened field order of P3PrognosticState (:161-183), P3DiagnosticInputs (:186-214),
 * P3DiagnosticOutputs (:217-241), P3HistoryOnly (:268-276). */
enum P3B200Field {
  P3B200_F_QC = 0, P3B200_F_NC, P3B200_F_QR, P3B200_F_NR, P3B200_F_QI, P3B200_F_QM, P3B200_F_NI,
  P3B200_F_BM, P3B200_F_QV, P3B200_F_TH,                                   /* prognostic, in/out */
  P3B200_F_NC_NUCEAT_TEND, P3B200_F_NCCN, P3B200_F_NI_ACTIVATED, P3B200_F_INV_QC_RELVAR,
  P3B200_F_CLD_FRAC_I, P3B200_F_CLD_FRAC_L, P3B200_F_CLD_FRAC_R, P3B200_F_PRES, P3B200_F_DZ,
  P3B200_F_DPRES, P3B200_F_INV_EXNER, P3B200_F_QV_PREV, P3B200_F_T_PREV,   /* diagnostic inputs */
  P3B200_F_QV2QI_DEPOS_TEND, P3B200_F_DIAG_EFF_RADIUS_QC, P3B200_F_DIAG_EFF_RADIUS_QI,
  P3B200_F_DIAG_EFF_RADIUS_QR, P3B200_F_RHO_QI, P3B200_F_PRECIP_LIQ_FLUX /* nlev+1 */,
  P3B200_F_PRECIP_ICE_FLUX /* nlev+1 */, P3B200_F_PRECIP_TOTAL_TEND, P3B200_F_NEVAPR, /* diag outputs */
  P3B200_F_LIQ_ICE_EXCHANGE, P3B200_F_VAP_LIQ_EXCHANGE, P3B200_F_VAP_ICE_EXCHANGE,    /* history */
  P3B200_F_PRECIP_LIQ_SURF /* [ncol] */, P3B200_F_PRECIP_ICE_SURF /* [ncol] */,
  P3B200_F_NUM
};

/* One array per field: device pointer + column stride in elements (ignored for the two [ncol] arrays).
 * P3B200_F_INV_QC_RELVAR, and the flag-dependent inputs that are dead for the chosen flags, may be NULL. */
typedef struct P3B200Fields {
  void*   ptr[P3B200_F_NUM];
  int64_t col_stride[P3B200_F_NUM];
} P3B200Fields;

/* P3Infrastructure (p3_functions.hpp:244-264).  its/ite/kts/kte are implied by ncol/nlev
 * (p3_main is always called with the whole local range, eamxx_p3_process_interface.cpp:329-334). */
typedef struct P3B200Infra {
  double  dt;
  int32_t it;
  int32_t ncol;            /* nj */
  int32_t nlev;            /* nk, 1..P3B200_MAX_NLEV */
  int32_t predictNc;
  int32_t prescribedCCN;
  int32_t zero_flux_top;   /* 1: also zero precip_*_flux[nlev] (what p3_main_f's pre-zeroing achieves,
                              p3_functions_f90.cpp:1941-1945); 0: reference p3_main pack-size-1 behaviour */
  int32_t zero_history;    /* 1: the call zero-fills liq_ice_exchange / vap_liq_exchange / vap_ice_exchange itself, i.e. it
                              replaces the caller's zero-fill that precedes every p3_main (eamxx_p3_run.cpp:27-29): the
                              streaming pass writes the zeros with its other stores and part2 stores non-zero values only.
                              0: reference p3_main behaviour (the arrays are written where part2 ran, untouched elsewhere) */
} P3B200Infra;

/* Per-column integer diagnostics (bit-exact parity targets), P3B200_D_NUM int32 per column. */
enum P3B200Diag {
  P3B200_D_NUC_POSSIBLE = 0, P3B200_D_HYDRO_PART1, P3B200_D_HYDRO_PART2,
  P3B200_D_QC_KTOP, P3B200_D_QC_KBOT, P3B200_D_QC_NSUB,
  P3B200_D_QR_KTOP, P3B200_D_QR_KBOT, P3B200_D_QR_NSUB,
  P3B200_D_QI_KTOP, P3B200_D_QI_KBOT, P3B200_D_QI_NSUB,
  P3B200_D_ERROR,          /* bit0 T<=0, bit1 T NaN (physics_saturation_impl.hpp:29,33), bit2 Co_max<0 (p3_upwind_impl.hpp:108) */
  P3B200_D_NUM = 16
};

typedef struct p3b200_ctx* p3b200_handle;

/* ---- library ---------------------------------------------------------------------------------- */
int         p3b200_abi_version(void);
const char* p3b200_strerror(int code);
const char* p3b200_last_cuda_error(void);
int         p3b200_device_count(void);                 /* >=0, or P3B200_ECUDA */

/* ---- tables: replaces init_kokkos_ice_lookup_tables (p3_table_ice_impl.hpp:18-78), init_kokkos_tables
 *      (p3_table3_impl.hpp:87-148) and the Fortran generator p3_init_b (micro_p3.F90:218-341). ------- */
void p3b200_runtime_defaults(P3B200Runtime* out);       /* P3Runtime defaults, p3_functions.hpp:114-133 */
/* Parse the reference's text table "p3_lookup_table_1.dat-v<version>" into ice/collect (log10 applied). */
int  p3b200_tables_read_text(const char* path, const char* expect_version, P3B200Tables* out);
/* Load the binary blob written by tools/make_table_blob.py (raw doubles; log10 applied on load). */
int  p3b200_tables_read_blob(const char* path, const char* expect_version, P3B200Tables* out);
/* Generate mu_r/vn/vm/revap by the p3_init_b quadrature and fill dnu (p3_table3_impl.hpp:120-135). */
int  p3b200_tables_generate_rain(P3B200Tables* out);

/* ---- handle ------------------------------------------------------------------------------------ */
/* real_bytes: 8 (SCREAM_DOUBLE_PRECISION) or 4.  device: CUDA ordinal.  Tables are copied to the device.
 * A handle is used by ONE host thread at a time (like the reference: one host thread per rank drives P3,
 * p3_main_impl.hpp:326); calls on different CUDA streams of one handle are allowed and are ordered where they share
 * the handle's scratch (the classification pass of a call waits for the previous call's column finish). */
int p3b200_create(const P3B200Tables* host_tables, int device, int real_bytes, p3b200_handle* out);
int p3b200_destroy(p3b200_handle h);

/* p3_main on DEVICE arrays.  Asynchronous on `cuda_stream` (a cudaStream_t, NULL = default stream)
 * unless elapsed_us != NULL, in which case it synchronises and returns the device time of the
 * five launches in microseconds like p3_main does (p3_main_impl.hpp:105,326-330).
 * diag_dev: optional device int32[ncol*P3B200_D_NUM]; NULL = keep them in the handle
 * (p3b200_get_diag). */
int p3b200_main(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* fields,
                const P3B200Infra* infra, void* cuda_stream, int64_t* elapsed_us, int32_t* diag_dev);

/* Copy the integer diagnostics of the last p3b200_main (diag_dev == NULL) to the host. */
int p3b200_get_diag(p3b200_handle h, int32_t* host_out, int32_t ncol);
/* OR of diag[P3B200_D_ERROR] over the columns of the last call (synchronises). 0 = clean. */
int p3b200_error_summary(p3b200_handle h, int32_t* or_of_error_bits);

/* Same as p3b200_main with HOST arrays (same layout): column-chunked H2D -> kernel -> D2H pipeline on internal
 * streams, copying straight from/to the caller's arrays (no staging copy).  Pass PINNED (page-locked) arrays
 * (cudaHostAlloc / cudaHostRegister): with pageable memory the copies are staged by the driver and the pipeline
 * degenerates to synchronous transfers.  This is the shape of P3Microphysics::run_impl for a host-resident
 * field manager and the e2e leg of bench.py.  Returns wall microseconds in elapsed_us. */
int p3b200_main_host(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* host_fields,
                     const P3B200Infra* infra, int64_t* elapsed_us, int32_t* diag_host);

/* Device kernel time (ms) of the last p3b200_main/p3b200_main_host measured with CUDA events on the
 * launching stream(s) (sum over chunks); synchronises. */
int p3b200_last_kernel_ms(p3b200_handle h, double* ms, int64_t* launches);
/* The same split by phase: ms6[0] phase A (init+part1+part2), ms6[1..3] cloud/rain/ice sedimentation, ms6[4] phase C
 * (homogeneous freezing + part3), ms6[5] the whole sedimentation stage (end of phase A to start of phase C); each
 * summed over the chunks of the last call.  p3b200_main runs the three species side by side on helper streams of the
 * handle (they share no data; environment P3B200_SED_SERIAL=1 restores stream order): ms6[1..3] are then each species'
 * time since the end of phase A, and p3b200_last_sed_concurrent returns 1. */
int p3b200_last_phase_ms(p3b200_handle h, double* ms6);
/* Split of phase A (ms6[0]) over its three kernels: ms3[0] classification pass (streaming: init + the deciding part of
 * part1 on every level), ms3[1] the part2 launches (one per process group), ms3[2] the column finish. */
int p3b200_last_phase_a_ms(p3b200_handle h, double* ms3);
int p3b200_last_sed_concurrent(p3b200_handle h);

/* ---- global water / energy budget (the one collective of the path: the caller all-reduces 4 doubles) ------
 * Domain totals over the columns of `fields` (device arrays), formed in double in a fixed order:
 *   totals4[0] = sum (qv+qc+qi+qr) * dpres / g            water mass            [kg m-2, summed over columns]
 *   totals4[1] = sum (T*Cpair + (LatVap+LatIce)*qv + LatIce*(qc+qr)) * dpres / g,  T = th/inv_exner   [J m-2]
 *   totals4[2] = sum RHO_H2O*(precip_liq_surf+precip_ice_surf)*dt     water that left through the surface
 *   totals4[3] = sum RHO_H2O*LatIce*precip_liq_surf*dt                energy that left through the surface
 * following the column formulas of share/property_checks/mass_and_energy_column_conservation_check.cpp:283-363
 * (kinetic and geopotential terms omitted: p3_main cannot change them).  Call it before p3b200_main (with
 * precip_*_surf zeroed) and after; conservation means  after[0]+after[2] == before[0]  and
 * after[1]+after[3] == before[1]  up to rounding.  Asynchronous on cuda_stream when totals4_host == NULL
 * (result in totals4_dev); with totals4_host != NULL it also copies the result back and synchronises. */
int p3b200_water_energy_totals(p3b200_handle h, const P3B200Fields* fields, const P3B200Infra* infra,
                               void* cuda_stream, double* totals4_dev, double* totals4_host);

/* ---- the same totals without any allocation ---------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * p3b200_water_energy_totals keeps its scratch (4 doubles per column + the partial sums) in a block the handle
 * allocates and grows on demand; this variant runs the same three kernels with that scratch in a file-scope
 * `static __device__` array of the library (p3_budget.cu), so the call itself performs no cudaMalloc / cudaFree and
 * calls no allocating library routine, copies nothing back and does not synchronise: the result (the same 4 doubles,
 * bit for bit) lands in totals4_dev, a DEVICE buffer of the caller's, in stream order on cuda_stream.
 *   limit   infra->ncol <= P3B200_STATIC_BUDGET_MAX_COLS; beyond it the call launches nothing and returns
 *           P3B200_ETOOBIG with the sizes in p3b200_last_cuda_error().  Shard larger domains, or use the variant above.
 *   cost    every GPU that loads the library holds the array, 1 MiB + 8 KiB, until the process exits.
 *   sharing all calls on a device (any handle, any stream) share the array, which keeps the last call's contents;
 *           nothing in it has to start at zero (every word read has been written earlier in the same call), but two
 *           calls that can overlap on the GPU race: order them on one stream or with events.
 *   set-up  the first call on each device looks the array's device address up (cudaGetSymbolAddress, not an
 *           allocation, but not something to do inside a graph capture): make one call per device beforehand. */
#define P3B200_STATIC_BUDGET_MAX_COLS 32768
int p3b200_water_energy_totals_static(p3b200_handle h, const P3B200Fields* fields, const P3B200Infra* infra,
                                      void* cuda_stream, double* totals4_dev);

/* ---- the column conservation check the atmosphere driver wraps around a process --------------------------
 * MassAndEnergyColumnConservationCheck (share/property_checks/mass_and_energy_column_conservation_check.cpp): before
 * the process runs the driver stores each column's total water mass and total energy (compute_current_mass / _energy,
 * :56-107); after it, check() (:110-215) forms the relative error of both against the stored value plus the boundary
 * fluxes the process reported (vapor/water/ice/heat flux, written by the postamble) times dt, and reduces each with
 * Kokkos::MaxLoc over the columns.  Pass when both maxima are below their tolerance.
 *   mass   = sum_k (qv+qc+qi+qr)*pseudo_density/g
 *   energy = sum_k (T_mid*Cpair + 0.5*(u^2+v^2) + (LatVap+LatIce)*qv + LatIce*(qc+qr))*pseudo_density/g + phis*ps/g
 * The within-column sums run in level order (the reference serialises them under ekatBFB), so the values have the bits
 * of the sequential sum.  Device pointers; 2-D fields (ncol, nlev) with `col_stride`; horiz_winds is (ncol, 2, nlev):
 * component c of column i starts at horiz_winds + i*wind_col_stride + c*wind_comp_stride.  Per-column arrays: ncol.
 * _begin stores the before-values in the handle; _check compares.  A NaN error never wins the MaxLoc (`err > max`),
 * equal errors report the lowest column; *_col is -1 when no column's error exceeded the reduction identity. */
typedef struct P3B200ConservationFields {
  const void *pseudo_density, *T_mid, *horiz_winds, *qv, *qc, *qi, *qr;
  const void *ps, *phis;                                       /* [ncol] */
  const void *vapor_flux, *water_flux, *ice_flux, *heat_flux;  /* [ncol]; only read by _check */
  int64_t col_stride, wind_col_stride, wind_comp_stride;
} P3B200ConservationFields;
typedef struct P3B200ConservationResult {
  double mass_rel_err, energy_rel_err;     /* the two maxima */
  int64_t mass_col, energy_col;            /* local column index of each */
  int32_t pass;                            /* 1: both below tolerance */
  int32_t mass_below_tol, energy_below_tol, reserved;
} P3B200ConservationResult;
int p3b200_column_conservation_begin(p3b200_handle h, const P3B200ConservationFields* f, int32_t ncol, int32_t nlev,
                                     void* cuda_stream);
int p3b200_column_conservation_check(p3b200_handle h, const P3B200ConservationFields* f, int32_t ncol, int32_t nlev,
                                     double dt, double mass_error_tolerance, double energy_error_tolerance,
                                     P3B200ConservationResult* result, void* cuda_stream);
/* the stored before-values and the per-column relative errors of the last _check (device -> host copies; NULL skips) */
int p3b200_column_conservation_get(p3b200_handle h, int32_t ncol, double* mass, double* energy, double* rel_mass,
                                   double* rel_energy);

/* ---- P3Microphysics::run_impl brackets p3_main with p3_preamble / p3_postamble ------------------------
 * (eamxx_p3_run.cpp:11-16, 39-44; functors eamxx_p3_process_interface.hpp:62-195, 200-262; the helper formulas
 * are share/util/scream_common_physics_functions_impl.hpp:88-97, 115-118, 162-165, 214-222, 327-331, 375-379,
 * 400-409).  Device pointers, (ncol, nlev) level-contiguous arrays that all share `col_stride`; the per-column
 * arrays have ncol entries.  Both are one streaming pass, one thread per (column, level); in-place fields
 * (wet <-> dry mixing ratios) are updated exactly where the reference updates them. */
typedef struct P3B200PreambleFields {
  /* in */
  const void *pmid, *pseudo_density, *pseudo_density_dry, *T_atm, *cld_frac_t;
  /* in/out: wet -> dry mixing ratio */
  void *qv, *qc, *nc, *qr, *nr, *qi, *qm, *ni, *bm, *qv_prev;
  /* out */
  void *inv_exner, *th_atm, *cld_frac_l, *cld_frac_i, *cld_frac_r, *dz;
  int64_t col_stride;
} P3B200PreambleFields;
int p3b200_preamble(p3b200_handle h, const P3B200PreambleFields* f, int32_t ncol, int32_t nlev, void* cuda_stream);

typedef struct P3B200PostambleFields {
  /* in */
  const void *pmid, *pseudo_density, *pseudo_density_dry, *th_atm;
  /* in/out */
  void *T_atm;                                                   /* T before p3 in, rescaled new T out */
  void *qv, *qc, *nc, *qr, *nr, *qi, *qm, *ni, *bm;              /* dry -> wet mixing ratio */
  void *diag_eff_radius_qc, *diag_eff_radius_qi, *diag_eff_radius_qr;   /* m -> micron */
  /* out */
  void *T_prev, *qv_prev;
  /* per column: fluxes from p3_main in, accumulated masses in/out, boundary fluxes out (all four NULL when
   * compute_mass_and_energy_fluxes is off) */
  const void *precip_liq_surf_flux, *precip_ice_surf_flux;
  void *precip_liq_surf_mass, *precip_ice_surf_mass;
  void *vapor_flux, *water_flux, *ice_flux, *heat_flux;
  int64_t col_stride;
  double dt;
} P3B200PostambleFields;
int p3b200_postamble(p3b200_handle h, const P3B200PostambleFields* f, int32_t ncol, int32_t nlev, void* cuda_stream);

/* ---- the property checks of the P3 process, as one fused reduction ---------------------------------
 * P3Microphysics::initialize_impl registers 15 checks (eamxx_p3_process_interface.cpp:226-243): two invariant
 * interval checks (T_mid, qv; qv repairable), eleven post-condition interval checks and two lower-bound checks.
 * The atmosphere driver evaluates each one after run_impl as its own MinMaxLoc reduction over the field
 * (share/property_checks/field_within_interval_check.cpp:80-226).  Here all 15 (min, max) pairs come out of ONE
 * pass: either folded into the postamble (p3b200_postamble_checked — the values are in registers there, the
 * checks cost no extra HBM traffic) or as a stand-alone pass over the 15 fields (p3b200_property_checks).
 *
 * out[i] follows the order of the add_*_check calls (P3B200_CHECK_* below).  `result` restates
 * check_impl :202-226 with the bounds of :226-243: PASS when lb <= min and max <= ub; otherwise REPAIRABLE
 * for a check registered with can_repair (only qv: its repairable bounds default to +-max), FAIL for the rest.
 * min_loc/max_loc are flat (col * nlev + lev; col for the two per-column fields) indices of the LOWEST index holding
 * the extreme value and are only computed when the check did not pass (-1 otherwise): the reference prints them
 * only in its failure message, and the order in which Kokkos joins equal values is unspecified.  Values that are
 * NaN are skipped, as the reference's `v < min` / `v > max` comparisons skip them.  Both calls synchronise the
 * stream (the reference's reductions return to the host as well).
 *
 * p3b200_postamble_checked carries only a 30-bit violation mask through the postamble (bit 2i: some value of check
 * i below lb, bit 2i+1: above ub — exactly "not Pass").  When the mask is zero all 15 results are PASS and
 * min_val/max_val are NaN (the reference's Pass result carries no values either, check_impl :229-234); otherwise
 * the call continues with the full reduction of p3b200_property_checks over the fields it has just written. */
enum { P3B200_CHECK_T_MID = 0, P3B200_CHECK_QV, P3B200_CHECK_QC, P3B200_CHECK_QI, P3B200_CHECK_QR, P3B200_CHECK_QM,
       P3B200_CHECK_NC, P3B200_CHECK_NR, P3B200_CHECK_NI, P3B200_CHECK_BM, P3B200_CHECK_PRECIP_LIQ_SURF_MASS,
       P3B200_CHECK_PRECIP_ICE_SURF_MASS, P3B200_CHECK_EFF_RADIUS_QC, P3B200_CHECK_EFF_RADIUS_QI,
       P3B200_CHECK_EFF_RADIUS_QR, P3B200_NUM_CHECKS };
enum { P3B200_CHECK_PASS = 0, P3B200_CHECK_REPAIRABLE = 1, P3B200_CHECK_FAIL = 2 };
typedef struct P3B200CheckResult {
  double lb, ub;                 /* the registered bounds (ub = DBL_MAX for the lower-bound checks) */
  double min_val, max_val;
  int64_t min_loc, max_loc;
  int32_t result;                /* P3B200_CHECK_PASS / _REPAIRABLE / _FAIL */
  int32_t pass_lower, pass_upper;
  int32_t reserved;
} P3B200CheckResult;
typedef struct P3B200CheckFields {
  const void *T_mid, *qv, *qc, *qi, *qr, *qm, *nc, *nr, *ni, *bm;
  const void *precip_liq_surf_mass, *precip_ice_surf_mass;      /* [ncol] */
  const void *eff_radius_qc, *eff_radius_qi, *eff_radius_qr;    /* microns, as the postamble leaves them */
  int64_t col_stride;
} P3B200CheckFields;
int p3b200_property_checks(p3b200_handle h, const P3B200CheckFields* f, int32_t ncol, int32_t nlev,
                           P3B200CheckResult out[P3B200_NUM_CHECKS], void* cuda_stream);
int p3b200_postamble_checked(p3b200_handle h, const P3B200PostambleFields* f, int32_t ncol, int32_t nlev,
                             P3B200CheckResult out[P3B200_NUM_CHECKS], void* cuda_stream);
/* p3_main followed by p3_postamble in one call: what P3Microphysics::run_impl does after the preamble
 * (eamxx_p3_run.cpp:26-44).  The postamble (T from theta with the dry/wet rescaling, dry -> wet, radii to microns,
 * T_prev / qv_prev, the per-column masses at level 0) is enqueued behind phase C as its own streaming launch.  (Round 1
 * ran it as the tail of phase C; that was measured slower than the two launches and has been removed.)  Results are
 * those of p3b200_main followed by p3b200_postamble.
 * `post` must name the SAME device arrays as `fields` for th_atm/qv/qc/nc/qr/nr/qi/qm/ni/bm, the three radii and
 * the two surface fluxes (P3B200_EINVAL otherwise), with post->col_stride equal to their column stride.
 * checks: NULL, or P3B200_NUM_CHECKS results filled as by p3b200_postamble_checked (the call then synchronises). */
int p3b200_main_postamble(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* fields,
                          const P3B200Infra* infra, const P3B200PostambleFields* post, P3B200CheckResult* checks,
                          void* cuda_stream, int64_t* elapsed_us, int32_t* diag_dev);
/* p3_main with the p3_preamble fused into its first kernel (the streaming classification pass): what run_impl does
 * before and with P3F::p3_main (eamxx_p3_run.cpp:11-36) in one call.  The pass reads the field manager's wet mixing
 * ratios, T_mid, p_mid, the two pseudo densities and cldfrac_tot directly, runs the level's p3_preamble
 * (eamxx_p3_process_interface.hpp:62-129) in registers and stores its results together with its own: the preamble's
 * separate pass over 31 arrays is gone.  Results are bit-identical to p3b200_preamble followed by p3b200_main (same
 * operations in the same order per level).  `pre` must name the SAME device arrays as `fields` for every array the
 * preamble writes (qv qc nc qr nr qi qm ni bm, qv_prev, inv_exner, th_atm, the three cloud fractions, dz), with
 * pre->col_stride equal to their common column stride (P3B200_EINVAL otherwise). */
int p3b200_main_preamble(p3b200_handle h, const P3B200Runtime* runtime, const P3B200Fields* fields,
                         const P3B200Infra* infra, const P3B200PreambleFields* pre, void* cuda_stream,
                         int64_t* elapsed_us, int32_t* diag_dev);
/* FieldWithinIntervalCheck::repair_impl (field_within_interval_check.cpp:338-372) for the one repairable check:
 * clip qv into its interval [1e-13, 0.2], in place. */
int p3b200_repair_qv(p3b200_handle h, void* qv, int64_t col_stride, int32_t ncol, int32_t nlev, void* cuda_stream);

/* ---- the reference's own C bridge, re-implemented on the GPU ---------------------------------------
 * Drop-in for p3_main_f (p3_functions_f90.hpp:945-952; p3_functions_f90.cpp:1888-2103): HOST pointers,
 * Fortran (column-major) layout arr[i + ncol*k], 1-based its/kts (must be 1), outputs zeroed on entry.
 * Real is double (p3_main_f) or float (p3_main_f_sp).  Uses a process-wide default handle created on
 * first use from p3b200_set_default_tables().  Returns elapsed microseconds (>=0) or a negative error. */
int p3b200_set_default_tables(const P3B200Tables* host_tables, int device);
int p3b200_p3_main_f(
  double* qc, double* nc, double* qr, double* nr, double* th_atm, double* qv, double dt,
  double* qi, double* qm, double* ni, double* bm, double* pres, double* dz,
  double* nc_nuceat_tend, double* nccn_prescribed, double* ni_activated, double* inv_qc_relvar, int it,
  double* precip_liq_surf, double* precip_ice_surf, int its, int ite, int kts, int kte,
  double* diag_eff_radius_qc, double* diag_eff_radius_qi, double* diag_eff_radius_qr, double* rho_qi,
  bool do_predict_nc, bool do_prescribed_CCN, double* dpres, double* inv_exner,
  double* qv2qi_depos_tend, double* precip_liq_flux, double* precip_ice_flux,
  double* cld_frac_r, double* cld_frac_l, double* cld_frac_i,
  double* liq_ice_exchange, double* vap_liq_exchange, double* vap_ice_exchange,
  double* qv_prev, double* t_prev);

#ifdef __cplusplus
}
#endif
#endif /* P3_B200_H */
