/* libsphb200 — C ABI of the B200-native SPH particle time-step.
 *
 * The reference (dmuley/sph-code) has no FFI layer: its boundary for this path is the Python module
 * namespace `navier_stokes_cleaned` (imported as `nsc`, sph/code_running.py:12).  Each entry point
 * below is what a binding for one function of that namespace calls; the Python module
 * `sph-code_b200/navier_stokes_cleaned.py` is that binding (ctypes), see INTEGRATION.md.
 *
 * Conventions
 *   - every array pointer is a DEVICE pointer (cudaMalloc / torch tensor .data_ptr()) unless its
 *     name ends in `_host`; arrays are C-contiguous float64 unless stated;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); calls are
 *     asynchronous on that stream except where a host result is returned;
 *   - return value 0 = success, non-zero = failure with the message in sphb200_last_error();
 *   - no exceptions cross the ABI; one context per GPU; a context is not thread-safe.
 */
#ifndef SPHB200_H
#define SPHB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sphb200_ctx sphb200_ctx;

/* ---- context ------------------------------------------------------------------------------ */
int sphb200_create(sphb200_ctx** out, int device);
int sphb200_destroy(sphb200_ctx* ctx);
const char* sphb200_last_error(void);
int sphb200_version(void);
/* physical constants the reference takes from scipy.constants (navier_stokes_cleaned.py:21-36):
 * Boltzmann k, atomic mass constant amu, hydrogen mass m_h, reference mass m_0 */
int sphb200_set_constants(sphb200_ctx* ctx, double k_boltzmann, double amu, double m_h, double m_0);
/* tunables: "cell_occ" (mean particles per cell), "cell_edge" (fixed cell edge, 0 = automatic),
 * "knn_group" (lanes per target: 8, 16 or 32) */
int sphb200_set_option(sphb200_ctx* ctx, const char* name, double value);
/* counters: "launches", "knn_overflow", "knn_rescan", "knn_widen", "bad_index"; reading a flag
 * synchronises the stream and resets it */
int sphb200_get_counter(sphb200_ctx* ctx, const char* name, int64_t* value_host, void* stream);

/* measurement aid (bench.py roofline): DFMA thread-instructions per second of a register-resident chain on
 * this GPU, timed with CUDA events (synchronises) */
int sphb200_fp64_peak(sphb200_ctx* ctx, double* dfma_per_second_host, void* stream);

/* ---- A1: neighbors(points, dist, N_NEIGH)            navier_stokes_cleaned.py:541-552 ---------
 * points (N,3).  Outputs: idx (N,k) int64 ascending distance, self first, missing = N;
 * dists (N,k) (missing = 0); nontriv (N) int64; new_sizes (N) = row max of dists.
 * Exact k-NN (the reference's eps=0.1 approximate search is replaced by eps=0). */
int sphb200_neighbors(sphb200_ctx* ctx, const double* points, int64_t n, int k, double dist, int64_t* idx,
                      double* dists, int64_t* nontriv, double* new_sizes, void* stream);

/* ---- A2-A6: hydro_update(...)                         navier_stokes_cleaned.py:556-671 ---------
 * neighbor (N,k) int64; points, velocities (N,3); mass, sizes, particle_type, T, mu_array,
 * gamma_array (N); f_un (N,S) or NULL.  Outputs: hydro_accel, visc_accel (N,3); visc_heat, density,
 * num_density, dust_density (N); f_un_neighbor (S,N) or NULL.
 * Fails (IndexError in the binding) if a neighbour index is outside [-N, N). */
int sphb200_hydro_update(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                         const double* mass, const double* sizes, const double* f_un, int n_species,
                         const double* particle_type, const double* T, const double* mu_array,
                         const double* gamma_array, const double* velocities, double* hydro_accel,
                         double* visc_accel, double* visc_heat, double* density, double* num_density,
                         double* f_un_neighbor, double* dust_density, void* stream);

/* ---- A9: crossing_time(neighbor, velocities, sizes, particle_type)         nsc:776-786 ---------
 * result (seconds) is written to *result_host after synchronising the stream. */
int sphb200_crossing_time(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* velocities,
                          const double* sizes, const double* particle_type, double dt_0, double* result_host,
                          void* stream);

/* ---- A7-A8: loop-generation sums                                          nsc:673-816 ---------
 * `d` is the module-global length scale the reference reads (nsc.d).  neighbor entries < 0 are
 * padding (ragged rows).  mode: 0 density (nsc:693), 1 dust_density (nsc:704), 2 num_dens (nsc:744). */
int sphb200_loop_density(sphb200_ctx* ctx, int mode, int64_t n, int k, const int64_t* neighbor, const double* points,
                         const double* mass, const double* particle_type, const double* sizes,
                         const double* mu_array, double d, double* out, void* stream);
int sphb200_del_pressure(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                         const double* mass, const double* particle_type, const double* E_internal,
                         const double* gamma_array, double d, double* grad_pressure, void* stream);
int sphb200_artificial_viscosity(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                                 const double* particle_type, const double* sizes, const double* mass,
                                 const double* densities, const double* velocities, const double* T,
                                 const double* gamma_array, const double* mu_array, double d, double* visc_accel,
                                 double* visc_heat, void* stream);

/* ---- F1: net_impulse(points,mass,sizes,velocities,particle_type,neighbor,f_un)   nsc:719-742 ----
 * Dust-gas drag: accel_onto (N,3) = sum over each row's dust neighbours with positive Weigh2_dust;
 * accel_reaction (N,3) = the opposite push scattered onto those neighbours (self excluded).
 * f_un is (N, n_species) row-major on the device; grain_mass_host / sigma_effective_host are the
 * n_species-long HOST tables of nsc:69-79 (grain_mass(), sigma_effective()), n_species <= 64. */
int sphb200_net_impulse(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                        const double* mass, const double* sizes, const double* velocities,
                        const double* particle_type, const double* f_un, int n_species, const double* grain_mass_host,
                        const double* sigma_effective_host, double* accel_onto, double* accel_reaction, void* stream);

/* ---- F2: grav_force_calculation_new(mass, points, sizes)                 nsc:252-415 --------------
 * The softened monopole law the reference applies bin by bin, a_i = G sum_j m_j (x_j - x_i) /
 * (|x_j - x_i|^2 + softening^2)^(3/2), summed directly over all particles (O(N^2), fp64); the binding passes
 * softening = median(sizes) as nsc:359 does.  accel (N,3). */
int sphb200_gravity(sphb200_ctx* ctx, int64_t n, const double* mass, const double* points, double G, double softening,
                    double* accel, void* stream);

/* ---- crossing time without any device allocation ---------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * sphb200_crossing_time (nsc:776-786), but the int32 neighbour list and the minimum slot live in a fixed file-scope
 * device array of 16 MiB + 256 B that every GPU loading the library holds until exit; the call needs no context and
 * performs no cudaMalloc / cudaFree.  LIMIT: n * k <= 4 194 304 neighbour entries (k = 40: n <= 104 857); larger
 * inputs are rejected with an error.  All pointers are DEVICE pointers, including the outputs: result_dev (1 double,
 * seconds) and bad_index_dev (1 int, may be NULL: entries of `neighbor` outside [-n, n), which are skipped); nothing
 * is synchronised.  The first call on a device looks up the array's address (cudaGetSymbolAddress); make it before a
 * stream capture starts.  Every call on a device shares the array: calls that overlap on the GPU (different streams,
 * concurrently replayed graphs) race; calls ordered on one stream are independent. */
int sphb200_crossing_time_noalloc(int device, int64_t n, int k, const int64_t* neighbor, const double* velocities,
                                  const double* sizes, const double* particle_type, double dt_0, double* result_dev,
                                  int* bad_index_dev, void* stream);

/* ---- A10-A13: device-resident particle state and the whole time step ---------------------------
 * The reference keeps its particle state as module-level numpy arrays in the driver
 * (sph/code_running.py:121-167) and advances it inline (:217-540).  A session holds the same
 * struct-of-arrays in HBM and advances it without leaving the GPU.
 * Field names are the driver's variable names: "points", "velocities", "total_accel" (N,3);
 * "mass", "sizes", "T", "E_internal", "mu_array", "gamma_array", "particle_type", "densities_0" (N).
 * Read-only results: "densities", "num_densities", "dust_densities", "visc_heat" (N); "hydro_accel",
 * "visc_accel", "pressure_accel" (N,3); "neighbor" (N,k) int64; "num_nontrivial" (N) int64.
 * upload/download take DEVICE pointers in the caller's particle order.  upload also accepts "neighbor"
 * ((N,k) int64 rows in the caller's indices, N = missing): the list a checkpoint saved, so that the next
 * step's crossing time (code_running.py:222) sees the same rows as the uninterrupted run. */
int sphb200_session_create(sphb200_ctx* ctx, int64_t n, int k, double dist);
/* A context holds at most one session: create fails while one is live.  The function-level calls above
 * (A1-A9, F1) share the context's scratch buffers with it and are refused on a context with a live
 * session - give them a context of their own.  destroy ends the session (sphb200_destroy does it too). */
int sphb200_session_destroy(sphb200_ctx* ctx);
/* same run, new particle set of n particles (the first n_owned are owned, the rest ghosts; -1 = all owned):
 * keeps k, dist, run parameters and device scalars; fill the arrays with sphb200_session_upload_rows */
int sphb200_session_resize(sphb200_ctx* ctx, int64_t n, int64_t n_owned, void* stream);
int sphb200_session_upload(sphb200_ctx* ctx, const char* field, const void* src, int64_t count, void* stream);
int sphb200_session_download(sphb200_ctx* ctx, const char* field, void* dst, int64_t count, void* stream);
/* Whole particle rows, (n_rows, ncol >= 18) doubles in the caller's order: columns 0-16 are the eleven
 * state fields in the order listed above, column 17 the k-NN trial radius (previous k-th distance);
 * further columns belong to the caller.  download writes rows [0, n_rows); upload needs n_rows == N.
 * Also accepted by upload/download by name: "densities", "visc_pi", "knn_h" (owner-computed values a
 * slab decomposition overwrites for ghost particles) and, download only, "dist_sq" (|x|^2 of the slow
 * owned particles, +inf for the rest).
 * Extra phases: "kick", "length", "smooth" (the three parts of "integrate"), "rho0"; "sort_keys" = "sort"
 * that carries only points, velocities and particle_type into the new order: a host-resident step uploads
 * the other fields afterwards (uploads land in the current order), overlapping the copy with the search. */
/* one field at `count` caller indices (device int64): out/in is (count, width) doubles */
int sphb200_session_gather(sphb200_ctx* ctx, const char* field, const int64_t* idx, int64_t count, double* out, void* stream);
int sphb200_session_scatter(sphb200_ctx* ctx, const char* field, const int64_t* idx, int64_t count, const double* in, void* stream);
int sphb200_session_download_rows(sphb200_ctx* ctx, double* rows, int64_t n_rows, int ncol, void* stream);
int sphb200_session_upload_rows(sphb200_ctx* ctx, const double* rows, int64_t n_rows, int ncol, void* stream);
/* pieces of the same: download with column 18 filled from `tag` (device, caller order: the particle id a slab
 * decomposition sends along with a row); upload of the rows of the caller indices [first, first + n_rows) */
int sphb200_session_download_rows_tagged(sphb200_ctx* ctx, double* rows, int64_t n_rows, int ncol, const double* tag,
                                         void* stream);
int sphb200_session_upload_rows_at(sphb200_ctx* ctx, const double* rows, int64_t first, int64_t n_rows, int ncol, void* stream);
/* slab decomposition, after the search: writes [1 if some owned particle's ball (x, h) leaves [lo, hi) along x,
 * largest h of the owned particles] into the float64 buffer "cover" (sphb200_session_buffer), to be
 * all_reduced (MAX) in place */
int sphb200_session_cover_check(sphb200_ctx* ctx, double lo, double hi, void* stream);
/* SURVEY row F4, variable-N state (code_running.py:361-381 grows every array when a supernova injects
 * particles): append n_new rows (same layout) as caller indices N .. N+n_new-1, keeping the existing state
 * and scalars.  Run the phases "sort" and "knn" afterwards (the driver's nsc.neighbors call at :386); the
 * next step then takes dv = a dt (:484-485) and new_smoothing = 1 (:534-535), as the reference does while its
 * old arrays are shorter.  Drag tables are dropped (call sphb200_session_set_drag again). */
int sphb200_session_append_rows(sphb200_ctx* ctx, const double* rows, int64_t n_new, int ncol, void* stream);
/* run parameters: "dt_0", "MAX_AGE", "N_INT_PER_PARTICLE", "raise_factor", "AU", "dist", "age", "d", "dt", "ct",
 * "n_owned" (caller indices >= n_owned are ghosts of a slab decomposition: excluded from the crossing-time
 * minimum and from the percentile behind d) */
int sphb200_session_set_scalar(sphb200_ctx* ctx, const char* name, double value);
/* F1 inside the step: with f_un (device, (N, n_species) row-major, caller order) the force phase adds
 * the dust-gas drag of code_running.py:455,462-463,473 to visc_accel (tables as in sphb200_net_impulse);
 * f_un == NULL switches it off (the default).  Not available on a slab session with ghosts. */
int sphb200_session_set_drag(sphb200_ctx* ctx, const void* f_un, int n_species, const double* grain_mass_host,
                             const double* sigma_effective_host, void* stream);
/* device scalars of the last step: "ct", "dt", "d", "age", "max_dist", "n_slow" (synchronises) */
int sphb200_session_scalar(sphb200_ctx* ctx, const char* name, double* value_host, void* stream);
/* code_running.py:171-179: first neighbour query + hydro_update, densities_0 <- densities */
int sphb200_session_startup(sphb200_ctx* ctx, void* stream);
/* one pass of the while-body, code_running.py:217-540 (gas/dust only, gravity and chemistry zero; drag as
 * set by sphb200_session_set_drag, off by default) */
int sphb200_session_step(sphb200_ctx* ctx, void* stream);
/* ---- slab decomposition on the device (the reference has only independent replicas,
 * sph/code_script_3.sl:19).  domain.py drives the collectives; these entry points keep everything N-sized on
 * the GPU.  Rows are NCOL >= 19 doubles: the 17 state columns, the k-NN trial radius, the global particle id.
 *   enable_gid      the session carries a global id per particle (column 18 of uploaded / downloaded rows);
 *   migrate_pack    phase 0: classify against the slab faces bounds[0..world] (device), counts_out (device
 *                   int32[world + 1]) = movers per destination rank, then the number of particles that stay;
 *                   phase 1: drop the ghosts, write the movers to rows_out grouped by destination, compact the
 *                   rest (caller index = compacted position); then call set_count(kept, kept);
 *   receive_rows    append rows from other ranks as owned particles (movers) or as ghosts;
 *   halo_pack       phase 0: count the owned particles within w of every other rank's slab (counts_out int32[world]);
 *                   phase 1: pack them as rows grouped by destination, with their caller indices in idx_out (the
 *                   send lists later ghost refreshes gather at with sphb200_session_gather). */
int sphb200_session_enable_gid(sphb200_ctx* ctx);
int sphb200_session_migrate_pack(sphb200_ctx* ctx, int phase, const double* bounds, int world, int rank, double* rows_out,
                                 int64_t cap_rows, int ncol, int* counts_out, void* stream);
int sphb200_session_set_count(sphb200_ctx* ctx, int64_t n, int64_t n_owned);
int sphb200_session_receive_rows(sphb200_ctx* ctx, const double* rows, int64_t n_new, int ncol, int as_ghosts, void* stream);
int sphb200_session_halo_pack(sphb200_ctx* ctx, int phase, const double* bounds, int world, int rank, double w, double* rows_out,
                              int64_t* idx_out, int64_t cap_rows, int ncol, int* counts_out, void* stream);
/* F2 inside the step: softening > 0 switches self-gravity on (phase "gravity" after the search, as
 * code_running.py:448; "integrate" adds it, :477), <= 0 off.  Result field "grav_accel" (N,3). */
int sphb200_session_set_gravity(sphb200_ctx* ctx, double G, double softening);
/* Synthetic motion for the measured workloads and the moving-particle parity tests (the reference has no
 * counterpart: its own step cannot displace particles by a sizeable fraction of h and stay finite, DESIGN.md
 * section 2): points[i] += amplitude * u(id_i, seed), u uniform in [-1,1)^3, a pure function of the particle's
 * caller index and the seed (host mirror: sph-code_b200/ics.py perturbation()).  Ghosts are left alone. */
int sphb200_session_perturb(sphb200_ctx* ctx, double amplitude, uint64_t seed, void* stream);
/* device pointer + element count of an internal array a caller-side collective works on in place:
 * "cover": float64 [bad, hmax] of sphb200_session_cover_check;
 * "length_hist": int64 histogram of the percentile's radix select (code_running.py:515-518), valid between
 * the phases "length_hist" and "length_pick"; a slab decomposition all_reduces (SUM) it there, five times
 * per step ("length_begin", then 5 x ("length_hist", all_reduce, "length_pick")). */
int sphb200_session_buffer(sphb200_ctx* ctx, const char* name, void** ptr, int64_t* count);
/* one phase of the step, for per-kernel timing: "crossing", "dt_clamp", "sort", "knn", "pack",
 * "density", "pi", "force", "integrate" */
int sphb200_session_phase(sphb200_ctx* ctx, const char* phase, void* stream);

#ifdef __cplusplus
}
#endif
#endif
