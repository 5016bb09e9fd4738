// Internal launch interfaces between the translation units of libsphb200.
#pragma once
#include "common.cuh"

// cells.cu: cell list on the device; guess = previous k-th distances in SOURCE order (nullptr: none); sync_host
// copies the chosen grid back to ctx->grid (function-level API only: one host synchronisation)
int build_cells(sphb200_ctx* ctx, const double* pts, int64_t n, double edge_hint, const double* guess, int knn_k,
                bool sync_host, cudaStream_t st);

int scan_exclusive_u32(sphb200_ctx* ctx, uint32_t* a, int m, cudaStream_t st);

// session.cu
void session_free(sphb200_ctx* ctx);

// knn.cu (general kernel: warp per target, any block width)
int knn_cap_for(int k);
int launch_knn(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double guess_factor, int32_t* list, double* dists,
               double* h_out, int32_t* nontriv, cudaStream_t st);
int launch_knn_worklist(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double guess_factor,
                        const int* worklist, const int* nwork, int32_t* list, double* dists, double* h_out, int32_t* nontriv,
                        int by_slot, cudaStream_t st);
// knn_group.cu (fast path: one CTA per 2x2x2 cell group, candidates staged by bulk copies, fp32 filter + fp64 exact)
int knn_group_max_k();
int launch_knn_group(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double margin, uint16_t* list16,
                     int32_t* list, double* dists, double* h_out, int32_t* nontriv, cudaStream_t st);
// knn_lane.cu (session fast path: one target per lane, single trial radius, misses go to the general kernel)
int launch_knn_lane(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double margin, uint16_t* list16,
                    int32_t* list, double* h_out, int32_t* nontriv, cudaStream_t st);
int launch_knn_unsort(sphb200_ctx* ctx, const int32_t* list, const double* dists, const double* h, const int32_t* nontriv,
                      int n, int k, int64_t* idx_out, double* dist_out, int64_t* nontriv_out, double* h_out,
                      cudaStream_t st);

// hydro.cu  (vectorised generation, nsc:556-671)
struct PackIn {
  const double* points;     // (N,3) or nullptr when r0.xyz is already filled (sorted session path)
  const double* vel;        // (N,3)
  const double* mass;       // (N)
  const double* h;          // (N)
  const double* ptype;      // (N) 0 gas, 1 star, 2 dust (float64, as the reference stores it)
  const double* T;          // (N)
  const double* mu;         // (N)
  const double* gamma;      // (N)
};
// Neighbour rows of a pass.  list16 != nullptr: (N,k) uint16 tile offsets (tile.cuh) for the cell-tiled kernels, and
// `list` holds the int32 rows of the escaped targets targets[0 .. *ntargets) by slot.  list16 == nullptr: `list`
// holds (N,k) int32 rows of every particle.
struct RowsRef {
  const uint16_t* list16;
  const int32_t* list;
  const int* targets;
  const int* ntargets;
};
int launch_pack(sphb200_ctx* ctx, int n, const PackIn& in, cudaStream_t st);
int launch_convert_list(sphb200_ctx* ctx, const int64_t* nb, int n, int k, int32_t* list, cudaStream_t st);
int launch_density(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, double* rho, double* dust, double* num, cudaStream_t st);
int launch_pi(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, cudaStream_t st);
int launch_repack(sphb200_ctx* ctx, int n, int what, const double* rho, cudaStream_t st);
int launch_force(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, const double* mass, double* hydro_accel,
                 double* visc_accel, double* visc_heat, cudaStream_t st);
int launch_species(sphb200_ctx* ctx, int n, int k, int S, const int32_t* list, const double* f_un, double* out,
                   cudaStream_t st);
int launch_crossing_time(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, const double* vel, const double* sizes,
                         const double* ptype, double dt_0, double* out_dev, const uint32_t* id, uint32_t n_owned, cudaStream_t st);
int launch_expand_list(sphb200_ctx* ctx, int n, int k, const uint16_t* list16, const int32_t* glist, int32_t* out, cudaStream_t st);

// gravity.cu (F2: direct Plummer-softened sum, nsc:252-415)
int launch_gravity_direct(sphb200_ctx* ctx, int n, const double* pos, const double* mass, double G, double eps, double* out,
                          cudaStream_t st);

// loopgen.cu  (loop generation, nsc:673-816)
struct LoopIn {
  const double* points; const double* vel; const double* mass; const double* sizes; const double* ptype;
  const double* T; const double* mu; const double* gamma; const double* E; const double* dens;
  double d;
};
int launch_loop_density(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, int mode, double* out,
                        cudaStream_t st);
int launch_loop_del_pressure(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, double* out,
                             cudaStream_t st);
int launch_loop_viscosity(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, double* accel,
                          double* heat, cudaStream_t st);
int launch_drag(sphb200_ctx* ctx, int n, int k, const int32_t* list, const uint32_t* id, const double* mg,
                const double* mc, const double* rho, const double* rho_dust, const double* ptype, double* onto,
                double* reaction, double* visc_accel, cudaStream_t st);
int launch_grain_means(sphb200_ctx* ctx, int n, int ns, const double* f_un, const double* meff, const double* seff,
                       double* mg, double* mc, cudaStream_t st);
int launch_loop_net_impulse(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, const double* mg,
                            const double* mc, double* onto, double* reaction, cudaStream_t st);
