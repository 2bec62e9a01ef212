/* d4est_b200.h -- C ABI of the B200 (sm_100a) matrix-free SIPG/DG Laplacian path.
 *
 * Plain C: pointers and sizes only, no torch / C++ types.  This is the drop-in
 * boundary behind d4est's operator API.  Every entry point names the reference
 * interface (file:line under the d4est tree) it replaces.  The d4est-facing
 * glue that calls these from d4est's own structs (p4est_t, d4est_mesh_data_t,
 * d4est_elliptic_data_t ...) is include/d4est_b200_glue.h; see INTEGRATION.md.
 *
 * Conventions:
 *   - 3-D only.  N = deg+1 GLL nodes per direction, Nq = deg_quad+1 quadrature
 *     points per direction.  Node id = i_r + N*(i_s + N*i_t) (r fastest).
 *   - nodal vectors are element-concatenated in p4est local order; element e
 *     starts at node_off[e] (d4est: ed->nodal_stride, Mesh/d4est_mesh.c:2441).
 *   - all arithmetic FP64.
 *   - return value 0 = success; nonzero = D4B_E* error code, message via
 *     d4b_last_error().  The reference itself has no error returns (it aborts,
 *     Utilities/d4est_util.h:171); the glue layer turns nonzero into D4EST_ABORT.
 */
#ifndef D4EST_B200_H
#define D4EST_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define D4B_MAX_DEG 11      /* N  <= 12 */
#define D4B_MAX_DEGQ 15     /* Nq <= 16; the completeness kernel additionally needs its scratch of one element to fit
                               227 KB of shared memory (d4b_mesh_create reports it: e.g. deg 11 with deg_quad >= 14
                               does not).  Template instances exist for deg 1..7, deg_quad = deg..deg+3. */

enum { D4B_OK = 0, D4B_EINVAL = 1, D4B_ECUDA = 2, D4B_ENOMEM = 3, D4B_ENCCL = 4, D4B_ENODEV = 5 };

enum { D4B_QUAD_GAUSS_LEGENDRE = 0, D4B_QUAD_GAUSS_LOBATTO = 1 };

/* ------------------------------------------------------------------------
 * Half-mortar record: one (local element, face, sub-face) side of an interface
 * or a physical-boundary face.  The face iterator of the reference
 * (Mesh/d4est_mortars_with_opt.c:612-806) visits every interface once per side
 * ("s_m" loop :659); one record is emitted per local "minus" element per
 * mortar sub-face, so each element owns all records that write into its Au and
 * no two records of different elements write the same output (no atomics).
 * ---------------------------------------------------------------------- */
typedef struct d4b_hm {
  int32_t nbr;      /* neighbour element: local id, n_local+ghost id, or -1 = boundary   */
  int8_t f_m;       /* my p4est face 0..5 (Appendix A.2)                                */
  int8_t f_p;       /* neighbour's face                                                 */
  int8_t tcode;     /* bit0: axes aligned, bit1: flip fast axis, bit2: flip slow axis:
                       m-order point (a,b) -> p-order point, restating
                       d4est_operators_reorient_face_data (dGMath/d4est_operators.c:1993) */
  int8_t child_m;   /* -1: my whole face is the mortar; 0..3: mortar = this sub-face of
                       my face (I am the coarse side), bits (fast, slow)                */
  int8_t child_p;   /* same for the neighbour, child index in the neighbour's own frame */
  int8_t degq;      /* mortar quadrature degree = max(deg_quad_m, deg_quad_p)
                       (dGMath/d4est_laplacian_with_opt_flux.c:302-316)                 */
  int8_t face_slot; /* 0..5 = f_m (kept for alignment / debugging)                      */
  int8_t pad;       /* boundary records: 0 = Dirichlet, 1 = Robin (fac rows 0-2 zero, row 3 = sj*alpha,
                       bc_g = sj*beta at the mortar points; ..._flux_sipg.c:377-476)                  */
  int32_t pad0;     /* explicit: keeps the 8-byte fields aligned, sizeof == 40        */
  int64_t fac_off;  /* offset in doubles into fac[]: [7][Nqm^2] = a_m[3], a_p[3], sj*sigma
                       (interior) or [4][Nqm^2] = a_m[3], sj*sigma (boundary)            */
  int64_t g_off;    /* boundary only: offset into bc_g[] of Nqm^2 Dirichlet values, or -1 */
  int64_t pad2;
} d4b_hm_t;         /* 40 bytes */

/* Flat, pointer-and-size description of one mesh level on one rank.  Built by the
 * glue from d4est structs (d4est_mesh_data_t Mesh/d4est_mesh.h:117-162,
 * d4est_element_data_t Mesh/d4est_element_data.h:13-46) or synthetically
 * (d4b_brick_desc_*).  All arrays are host pointers owned by the caller and are
 * copied to HBM by d4b_mesh_create. */
typedef struct d4b_mesh_desc {
  int32_t n_local;          /* local elements                                          */
  int32_t n_ghost;          /* ghost elements (d4est_ghost_t Mesh/d4est_ghost.h:7)     */
  int32_t quad_type;        /* D4B_QUAD_*  ([quadrature] name = legendre | lobatto)    */
  int32_t reserved;
  const int32_t *deg;       /* [n_local+n_ghost]                                       */
  const int64_t *node_off;  /* [n_local+n_ghost+1]; ghosts continue after local_nodes  */
  const int32_t *degq;      /* [n_local]  deg_quad                                     */
  const int64_t *quad_off;  /* [n_local+1] volume quadrature node offsets              */
  const double *G;          /* [6*quad_off[n_local]] element-blocked symmetric metric:
                               G[6*quad_off[e] + c*Nq^3 + q], c = (00,01,02,11,12,22),
                               G_ll' = w J sum_k dr_l/dx_k dr_l'/dx_k
                               (Quadrature/d4est_quadrature.c:339-375)                  */
  const int32_t *hm_begin;  /* [n_local+1] CSR of half-mortar records per element      */
  const d4b_hm_t *hm;       /* [hm_begin[n_local]]                                     */
  const double *fac;        /* mortar factor pool                                      */
  int64_t n_fac;
  const double *bc_g;       /* boundary-value pool (may be NULL when homogeneous)      */
  int64_t n_bc_g;
  /* halo (d4est_ghost_data_exchange Mesh/d4est_ghost_data.c:143): ghost elements are
   * ordered by owner rank as in p4est's ghost layer                                    */
  int32_t n_peers;
  int32_t reserved2;
  const int32_t *peer_rank;   /* [n_peers]                                             */
  const int32_t *send_begin;  /* [n_peers+1] into send_elem                            */
  const int32_t *send_elem;   /* local element ids (mirrors), per peer                 */
  const int32_t *recv_begin;  /* [n_peers+1] ghost element index ranges per peer       */
} d4b_mesh_desc_t;

typedef struct d4b_mesh d4b_mesh_t; /* opaque: device-resident level */

const char *d4b_last_error(void);
int d4b_device_count(int *count);
int d4b_set_device(int dev);   /* one device per process: a second, different device is refused */
const char *d4b_version(void);

/* ---- mesh residency (hook: d4est_mesh_update, Mesh/d4est_mesh.c:2790) ------------ */
int d4b_mesh_create(const d4b_mesh_desc_t *desc, d4b_mesh_t **out);
void d4b_mesh_destroy(d4b_mesh_t *m);
int64_t d4b_mesh_local_nodes(const d4b_mesh_t *m);
int64_t d4b_mesh_ghost_nodes(const d4b_mesh_t *m);
int64_t d4b_mesh_bytes(const d4b_mesh_t *m);          /* HBM footprint of the level    */
int64_t d4b_mesh_algorithmic_bytes(const d4b_mesh_t *m); /* SURVEY 8(d) bytes per apply */

/* Synthetic uniform brick (BASELINE config 2: "Poisson on a uniform brick mesh, single degree"),
 * generated DIRECTLY into the resident tables by CUDA kernels -- same tables the glue produces
 * from the reference's [geometry] name = brick (Geometry/d4est_geometry_brick.c:121,152) with
 * face_h_type = FACE_H_EQ_TREE_H (Mesh/d4est_mesh.c:717-726), penalty maxp_sqr_over_minh
 * (dGMath/d4est_laplacian_with_opt_flux_sipg.c:1314) and homogeneous Dirichlet data.  Elements are
 * in Morton order (p4est's order inside one tree when nx = ny = nz = 2^level). */
typedef struct d4b_brick_params {
  int32_t nx, ny, nz;       /* elements per direction                                   */
  int32_t deg, degq;        /* uniform degree and quadrature degree                     */
  int32_t quad_type;        /* D4B_QUAD_*                                               */
  double x0, x1, y0, y1, z0, z1;
  double penalty_prefactor; /* [flux] sipg_penalty_prefactor                            */
  double face_h;            /* h in the penalty; <= 0 means 1/nx (= dq/ROOT_LEN)        */
} d4b_brick_params_t;
int d4b_brick_create(const d4b_brick_params_t *p, d4b_mesh_t **out);
/* One z-slab per rank of the brick nx x ny x (nz*nranks) over the same box (the p4est partition of a
 * brick forest cut along z): ghost planes towards rank-1 / rank+1 and the mirror lists that
 * d4b_halo_exchange uses (d4est_ghost_t / d4est_ghost_data_exchange, Mesh/d4est_ghost_data.c:143). */
int d4b_brick_create_slab(const d4b_brick_params_t *p, int rank, int nranks, d4b_mesh_t **out);

/* ---- operator apply: replaces d4est_laplacian_with_opt_apply_aij
 *      (dGMath/d4est_laplacian_with_opt.c:289) and its twin d4est_laplacian_apply_aij
 *      (dGMath/d4est_laplacian.c:317).  Au is overwritten.
 *      d4b_apply_host takes any host memory; buffers from d4b_host_alloc_pinned (or cudaHostAlloc /
 *      cudaHostRegister with the mapped flag) let uploads, kernels and downloads overlap fully.      */
int d4b_apply_host(d4b_mesh_t *m, const double *u_host, double *Au_host);
int d4b_apply_dev(d4b_mesh_t *m, const double *u_dev, double *Au_dev, void *cuda_stream);
/* ghost values supplied by the caller (single-process tests of the N>1 layout) */
int d4b_set_ghost_host(d4b_mesh_t *m, const double *ughost_host);
int64_t d4b_kernel_launches(void);   /* number of d4b kernels launched so far          */
/* how the local elements of a level are split over the element kernels: counts[0] register-blocked /
 * line-task kernels (regular elements), [1] templated hanging / mixed-p kernel, [2] completeness kernel */
int d4b_mesh_kernel_split(const d4b_mesh_t *m, int64_t counts[3]);

/* ---- rank boundary (one p4est rank = one GPU): replaces d4est_ghost_data_exchange
 *      (Mesh/d4est_ghost_data.c:143-255) and sc_allreduce (Solver/d4est_solver_cg.c:139,165) by
 *      ncclSend/ncclRecv + ncclAllReduce.  The 128-byte id is created on rank 0
 *      (d4b_comm_unique_id) and broadcast by the host program (MPI_Bcast in d4est).  With a
 *      communicator, d4b_apply_dev exchanges the halo of u before the element kernels.        */
int d4b_comm_unique_id(char *out128);
int d4b_comm_init(int rank, int nranks, const char *id128);
int d4b_comm_finalize(void);
int d4b_comm_size(void);
int d4b_halo_exchange(d4b_mesh_t *m, const double *u_dev, void *cuda_stream);

/* ---- device vectors / BLAS-1 (LinearAlgebra/d4est_linalg.c:179-235) -------------- */
int d4b_dev_alloc(size_t bytes, void **dptr);
int d4b_dev_free(void *dptr);
int d4b_h2d(void *dst_dev, const void *src_host, size_t bytes, void *cuda_stream);
int d4b_d2h(void *dst_host, const void *src_dev, size_t bytes, void *cuda_stream);
int d4b_sync(void *cuda_stream);
int d4b_host_alloc_pinned(size_t bytes, void **hptr);
int d4b_host_free_pinned(void *hptr);
int d4b_vec_axpy(int64_t n, double a, const double *x, double *y, void *cuda_stream);      /* y += a x      */
int d4b_vec_xpby(int64_t n, const double *x, double b, double *y, void *cuda_stream);      /* y = x + b y   */
int d4b_vec_axpyeqz(int64_t n, double a, const double *x, const double *y, double *z, void *cuda_stream);
int d4b_vec_scale(int64_t n, double a, double *x, void *cuda_stream);
int d4b_vec_copy(int64_t n, const double *x, double *y, void *cuda_stream);
int d4b_vec_zero(int64_t n, double *x, void *cuda_stream);
int d4b_vec_dot(int64_t n, const double *x, const double *y, double *result_host, void *cuda_stream);
/* Allocation-free dot product.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * result_dev[0] = sum_i x[i] y[i] over THIS rank's n entries (no NCCL call: the caller reduces over ranks), written
 * on the device, in stream order; x, y, result_dev are device pointers.  Calls no cudaMalloc / cudaFree / pinned
 * allocation: the block partials live in a static device array of 75 776 bytes that every GPU loading the library
 * holds until exit.  n > D4B_NOALLOC_DOT_MAX_N is rejected with D4B_EINVAL.  All calls share that array: calls that
 * overlap on the GPU (different streams, no ordering) race; calls on one stream are safe.  The first call on a
 * device resolves the array's address (cudaGetSymbolAddress); make it outside a graph capture. */
#define D4B_NOALLOC_DOT_MAX_N ((int64_t)1 << 28)
int d4b_vec_dot_noalloc(int64_t n, const double *x, const double *y, double *result_dev, void *cuda_stream);

/* ---- Chebyshev smoother + CG spectral bound (device vectors; `work` = caller scratch)
 *      d4est_solver_multigrid_smoother_cheby_iterate_aux  Solver/d4est_solver_multigrid_smoother_cheby.c:81
 *      cg_eigs                                            Solver/d4est_solver_cg_eigs.c:115            */
int d4b_cheby_iterate(d4b_mesh_t *m, double *u, const double *rhs, double *r, int iter, double lmin,
                      double lmax, int residual_at_end, double *work_2n, void *cuda_stream);
int d4b_cg_eigs(d4b_mesh_t *m, double *u, const double *rhs, int imax, int use_new, double *spectral_bound,
                double *work_3n, void *cuda_stream);
/* ---- Krylov drivers: d4est_solver_cg_solve Solver/d4est_solver_cg.c:76,
 *      d4est_solver_fcg_solve Solver/d4est_solver_fcg_improved.c:97 (pc = d4est_krylov_pc_t.pc_apply) */
typedef int (*d4b_pc_fcn_t)(void *ctx, const double *r_dev, double *z_dev, void *cuda_stream);
int d4b_cg_solve(d4b_mesh_t *m, double *u, const double *rhs, int imax, double atol, double rtol,
                 int *iters, double *res2, double *work_3n, void *cuda_stream);
int d4b_fcg_solve(d4b_mesh_t *m, double *u, const double *rhs, int imax, double atol, double rtol,
                  d4b_pc_fcn_t pc, void *pc_ctx, int *iters, double *final_res, double *work_6n,
                  void *cuda_stream);
/* options: "fcg_textbook_update" (default 0 = the arithmetic d4est_solver_fcg_solve really performs, see
 * d4b_solver.cu; 1 = d_k = v_k - gamma/rho d_{k-1});  test switches of the d4b_apply_host pipeline, whose defaults
 * come from the environment variables of the same name in capitals with a D4B_ prefix, read once:
 * "no_pipeline", "no_late_pass", "pipe_trace", "pipe_chunks" */
int d4b_set_option(const char *name, int value);

/* ---- multigrid on cached levels.  Transfer table between two levels (one entry per COARSE
 *      element; d4est_solver_multigrid_refine_data_t{hrefine,degh[8],degH} Solver/d4est_solver_multigrid.h:12-18):
 *      kind 0 = same cell (p-prolongation, 1 child), kind 1 = 8 children (hp-prolongation, child c =
 *      bits (r,s,t)); offsets are into the coarse / fine nodal vectors.                              */
typedef struct d4b_mg_transfer d4b_mg_transfer_t;
typedef struct d4b_mg d4b_mg_t;
typedef struct d4b_mg_params {  /* [multigrid], [mg_smoother_cheby], [mg_bottom_solver_*] input keys */
  int32_t vcycle_imax; double vcycle_rtol, vcycle_atol;
  int32_t cheby_imax, cheby_eigs_cg_imax; double cheby_eigs_lmax_lmin_ratio, cheby_eigs_max_multiplier;
  int32_t eigs_reuse_fromdownvcycle, eigs_reuse_fromlastvcycle, use_new_cg_eigs;
  int32_t bottom_kind;          /* 0 = mg_bottom_solver_cheby, 1 = mg_bottom_solver_cg */
  int32_t bottom_imax, bottom_eigs_cg_imax; double bottom_eigs_lmax_lmin_ratio, bottom_eigs_max_multiplier;
  double bottom_atol, bottom_rtol;
} d4b_mg_params_t;
int d4b_mg_transfer_create(int n_coarse, const int32_t *kind, const int32_t *degH, const int64_t *coarse_off,
                           const int32_t *child_deg /*[8 n]*/, const int64_t *child_off /*[8 n]*/,
                           d4b_mg_transfer_t **out);
void d4b_mg_transfer_destroy(d4b_mg_transfer_t *T);
int d4b_mg_restrict(d4b_mg_transfer_t *T, const double *fine_dev, double *coarse_dev, void *cuda_stream);
int d4b_mg_prolong(d4b_mg_transfer_t *T, const double *coarse_dev, double *fine_dev, void *cuda_stream);
/* ---- AMR field transfer (old grid -> refined grid -> 2:1-balanced grid), driven by the logs of d4est_amr_step:
 *      d4est_amr_interpolate_field              hpAMR/d4est_amr.c:406-485
 *      d4est_amr_interpolate_field_on_element   hpAMR/d4est_amr.c:338-403 (code < 0: 8 children of the same degree by
 *                                               hp-prolongation; code >= deg: p-prolongation; coarsening aborts there,
 *                                               D4B_EINVAL here)
 *      d4b_amr_transfer_create builds the transfer table of ONE sweep (element i of degree |deg_in[i]| with code[i]);
 *      apply it with d4b_mg_prolong.  The _dev form runs both sweeps on device vectors (aux_dev: n_aux doubles from
 *      _sizes); the _host form stages host vectors (the call the glue makes for d4est's double** field). */
int d4b_amr_transfer_create(int n, const int32_t *deg_in, const int32_t *code, d4b_mg_transfer_t **out,
                            int64_t *nodes_in, int64_t *nodes_out);
int d4b_amr_interpolate_field_sizes(int n_init, const int32_t *initial_log, const int32_t *refinement_log, int n_bal,
                                    const int32_t *balance_log, int64_t *n_in, int64_t *n_aux, int64_t *n_out);
int d4b_amr_interpolate_field_dev(int n_init, const int32_t *initial_log, const int32_t *refinement_log, int n_bal,
                                  const int32_t *balance_log, const double *in_dev, double *aux_dev, double *out_dev,
                                  void *cuda_stream);
int d4b_amr_interpolate_field_host(int n_init, const int32_t *initial_log, const int32_t *refinement_log, int n_bal,
                                   const int32_t *balance_log, const double *field_in, int64_t n_in, double *field_out,
                                   int64_t n_out);
/* levels[0] = coarsest ... levels[n-1] = finest; transfers[l] joins l (coarse) and l+1 (fine) */
int d4b_mg_create(int n_levels, d4b_mesh_t **levels, d4b_mg_transfer_t **transfers, const d4b_mg_params_t *par,
                  d4b_mg_t **out);
void d4b_mg_destroy(d4b_mg_t *g);
int d4b_mg_get_eigs(d4b_mg_t *g, double *eigs_per_level);
int d4b_mg_vcycle(d4b_mg_t *g, double *u_dev, const double *rhs_dev, double *r2_local, void *cuda_stream);
int d4b_mg_solve(d4b_mg_t *g, double *u_dev, const double *rhs_dev, int *vcycles, double *r2_final, void *cuda_stream);
int d4b_mg_pc_apply(void *mg, const double *r_dev, double *z_dev, void *cuda_stream);
int d4b_mg_fcg_solve(d4b_mg_t *g, double *u_dev, const double *rhs_dev, int imax, double atol, double rtol,
                     int *iters, double *final_res, double *work_6n, void *cuda_stream);
int d4b_mg_num_levels(const d4b_mg_t *g);

/* ---- HOST-vector forms at the granularity d4est's solver stack calls them: one upload / download around MANY
 *      applies (a smoother call is 15 + 15 + 1 of them).  These are what the d4est-facing glue
 *      (include/d4est_b200_glue.h) installs behind d4est's own interfaces:
 *      build_residual callbacks (EllipticSystem/d4est_elliptic_eqns.h:32; e.g. poisson_sinx_fcns.h:133-148),
 *      d4est_solver_multigrid_smoother_t.smooth / _bottom_solver_t.solve (Solver/d4est_solver_multigrid.h:129-144;
 *      bodies smoother_cheby.c:264-360, bottom_solver_cheby.c:59-113),
 *      d4est_krylov_pc_t.pc_apply (Solver/d4est_krylov_pc.h:13-33; body d4est_krylov_pc_multigrid.c:41-75),
 *      d4est_solver_fcg_solve (Solver/d4est_solver_fcg_improved.c:97). */
int d4b_residual_dev(d4b_mesh_t *m, const double *u_dev, const double *rhs_dev, double *r_dev, void *cuda_stream);
int d4b_residual_host(d4b_mesh_t *m, const double *u_host, const double *rhs_host, double *r_host);
int d4b_cheby_smooth_host(d4b_mesh_t *m, double *u_host, const double *rhs_host, double *r_host, int cheby_imax,
                          int eigs_cg_imax, double lmax_lmin_ratio, double max_multiplier, int use_new_cg_eigs,
                          int compute_eig, double *eig);
/* mass-type terms of apply_jac / build_residual: out (+)= I^T diag(w c) I v with c = J f(x, u0) at the volume
 * quadrature points (d4est_quadrature_apply_fofufofvlilj, Quadrature/d4est_quadrature.c:593; f = 1:
 * d4est_quadrature_apply_mass_matrix :776).  The coefficient is resident until it is set again. */
int d4b_mass_set_coef(d4b_mesh_t *m, const double *coef_quad_host);
int d4b_mass_apply_dev(d4b_mesh_t *m, const double *v_dev, double *out_dev, int accumulate, void *cuda_stream);
int d4b_mass_apply_host(d4b_mesh_t *m, const double *v_host, double *out_host, int accumulate);
int d4b_apply_jac_dev(d4b_mesh_t *m, const double *u_dev, double *Au_dev, void *cuda_stream);   /* A u + M_c u */
int d4b_mg_pc_apply_host(d4b_mg_t *g, const double *r_host, double *z_host);
int d4b_mg_fcg_solve_host(d4b_mg_t *g, double *u_host, const double *rhs_host, int imax, double atol, double rtol,
                          int *iters, double *final_res);

#ifdef __cplusplus
}
#endif
#endif
