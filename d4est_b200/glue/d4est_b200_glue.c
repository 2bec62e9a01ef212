/* d4est_b200_glue.c -- host C glue between d4est's structs and the flat C ABI.
 * Compiled against the d4est headers (in the d4est tree, or here by
 * oracle/ref_build when /root/reference is mounted).  See include/d4est_b200_glue.h.
 */
#include "d4est_b200_glue.h"
#include "../csrc/d4b_ops.h"
#include <d4est_mortars_with_opt.h>
#include <d4est_laplacian_flux.h>
#include <d4est_laplacian_flux_sipg.h>
#include <d4est_xyz_functions.h>
#include <d4est_laplacian_with_opt_flux_sipg.h>
#include <d4est_lgl.h>
#include <d4est_reference.h>
#include <d4est_util.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

/* ------------------------------------------------------------------ recorder */
typedef struct {
  int32_t elem;     /* owning local element */
  d4b_hm_t hm;
  int nfac;         /* doubles in this record's factor block */
} rec_t;

typedef struct {
  p4est_t *p4est;
  d4est_ghost_t *ghost;
  /* the three things the recording callbacks need from a flux-data struct; filled from
   * d4est_laplacian_with_opt_flux_data_t or from the baseline twin's d4est_laplacian_flux_data_t */
  d4est_laplacian_with_opt_flux_sipg_params_t *sp;   /* common leading fields of both sipg parameter structs */
  d4est_laplacian_bc_t bc_type; void *bc_data;
  int baseline;     /* flux data of the baseline twin: its Robin callbacks take ITS boundary-data struct */
  int n_local;
  rec_t *rec; size_t nrec, caprec;
  double *fac; size_t nfac, capfac;
  double *g; size_t ng, capg;
  int quad_type;
} recorder_t;

static void push_rec(recorder_t *R, const rec_t *r)
{
  if (R->nrec == R->caprec) {
    R->caprec = R->caprec ? 2 * R->caprec : 1024;
    R->rec = (rec_t *)realloc(R->rec, R->caprec * sizeof(rec_t));
  }
  R->rec[R->nrec++] = *r;
}
static double *grow(double **buf, size_t *n, size_t *cap, size_t add)
{
  if (*n + add > *cap) {
    *cap = (*cap ? 2 * *cap : 4096); while (*n + add > *cap) *cap *= 2;
    *buf = (double *)realloc(*buf, *cap * sizeof(double));
  }
  double *p = *buf + *n; *n += add; return p;
}

static int elem_ext_index(recorder_t *R, d4est_element_data_t *e)
{ /* local id, or n_local + ghost id (d4est_ghost_init sets ged->id = gid, Mesh/d4est_ghost.c:58) */
  if (e->mpirank == R->p4est->mpirank) return e->id;
  return R->n_local + (int)(e - R->ghost->ghost_elements);
}

/* m-order point -> p-order point code; restates the decisions of
 * d4est_operators_reorient_face_data (dGMath/d4est_operators.c:2029-2075) */
static int orientation_tcode(int f_m, int f_p, int o)
{
  int ft[9];
  p4est_expand_face_transform((f_m <= f_p) ? f_m : f_p,
                              (P4EST_FACES)*o + ((f_m <= f_p) ? f_p : f_m), ft);
  int aligned = ((ft[1] - ft[0]) * (ft[4] - ft[3]) > 0);
  return (aligned ? 1 : 0) | (ft[6] ? 2 : 0) | (ft[7] ? 4 : 0);
}

/* source (p-order) index of m-order point (a,b) on an n x n face grid */
static int tcode_src(int tcode, int n, int a, int b)
{
  int x = (tcode & 1) ? a : b, y = (tcode & 1) ? b : a;
  if (tcode & 2) x = n - 1 - x;
  if (tcode & 4) y = n - 1 - y;
  return x + n * y;
}

/* verify tcode against the reference's own reorientation on an index ramp */
static void check_tcode(d4est_operators_t *ops, int f_m, int f_p, int o, int tcode)
{
  enum { dg = 2, n = 3 };
  double in[n * n], out[n * n];
  for (int i = 0; i < n * n; i++) in[i] = i;
  d4est_operators_reorient_face_data(ops, in, 2, dg, o, f_m, f_p, out);
  for (int b = 0; b < n; b++)
    for (int a = 0; a < n; a++)
      if ((int)out[a + n * b] != tcode_src(tcode, n, a, b))
        D4EST_ABORT("d4est_b200: orientation code does not reproduce reorient_face_data");
}

static double penalty_sigma(recorder_t *R, d4est_element_data_t *em, d4est_element_data_t *ep,
                            int deg_m, double hm, int deg_p, double hp)
{ /* dGMath/d4est_laplacian_with_opt_flux_sipg.c:620-639 (interface), :75-86 (boundary) */
  d4est_laplacian_with_opt_flux_sipg_params_t *sp = R->sp;
  double mult = d4est_laplacian_with_opt_flux_sipg_get_multiplier(sp, em->region, ep->region);
  return mult * sp->sipg_penalty_fcn(deg_m, hm, deg_p, hp, sp->sipg_penalty_prefactor);
}

/* Recording twin of d4est_laplacian_with_opt_flux_interface
 * (dGMath/d4est_laplacian_with_opt_flux.c:212): called by the reference iterator once
 * per side; emits one half-mortar per LOCAL minus element per mortar sub-face. */
static void record_interface(p4est_t *p4est, d4est_element_data_t **e_m, int faces_m, int f_m,
                             int mortar_side_id_m, int *e_m_is_ghost,
                             d4est_element_data_t **e_p, int faces_p, int f_p,
                             int mortar_side_id_p, int *e_p_is_ghost, int orientation,
                             d4est_operators_t *ops, d4est_geometry_t *geom,
                             d4est_quadrature_t *quad, d4est_mesh_data_t *factors, void *ctx)
{
  recorder_t *R = (recorder_t *)ctx;
  int faces_mortar = (faces_m > faces_p) ? faces_m : faces_p;
  d4est_element_data_t *e_p_or[(P4EST_HALF)];
  d4est_element_data_reorient_f_p_elements_to_f_m_order(e_p, (P4EST_DIM)-1, f_m, f_p, orientation,
                                                        faces_p, e_p_or);
  int degq_m[(P4EST_HALF)], nq2[(P4EST_HALF)], nq2_p[(P4EST_HALF)], sub_p[(P4EST_HALF)];
  int total = 0;
  for (int i = 0; i < faces_m; i++)
    for (int j = 0; j < faces_p; j++) {
      degq_m[i + j] = d4est_util_max_int(e_m[i]->deg_quad, e_p_or[j]->deg_quad); /* :302-316 */
      nq2[i + j] = d4est_lgl_get_nodes(2, degq_m[i + j]);
      total += nq2[i + j];
    }
  for (int i = 0; i < faces_mortar; i++) {
    sub_p[i] = (faces_mortar == (P4EST_HALF))
                   ? d4est_reference_reorient_face_order(2, f_m, f_p, orientation, i) : i;
    nq2_p[sub_p[i]] = nq2[i];
  }
  int use_id = -1; /* slot owner: first local minus element (:445-456) */
  for (int i = 0; i < faces_m; i++)
    if (e_m[i]->mpirank == p4est->mpirank) { use_id = e_m[i]->id; break; }
  if (use_id < 0) return;
  size_t s0 = (size_t)factors->local_strides[use_id].mortar_quad_stride[f_m];
  const double *sj = &factors->sj_m_mortar_quad[s0];
  const double *hm = &factors->hm_mortar_quad[s0];
  const double *hp = &factors->hp_mortar_quad[s0];
  const double *nn = &factors->n_m_mortar_quad[3 * s0];               /* [d*total + k]          */
  const double *dm = &factors->drst_dxyz_m_mortar_quad[9 * s0];       /* [(l+3d)*total + k]     */
  const double *dp = &factors->drst_dxyz_p_mortar_quad_porder[9 * s0];
  int tcode = orientation_tcode(f_m, f_p, orientation);
  check_tcode(ops, f_m, f_p, orientation, tcode);

  int off = 0;
  for (int f = 0; f < faces_mortar; f++) {
    int im = (faces_m == faces_mortar) ? f : 0;
    int ip = (faces_p == faces_mortar) ? f : 0;
    d4est_element_data_t *em = e_m[im], *ep = e_p_or[ip];
    int n2 = nq2[f], nq = degq_m[f] + 1;
    if (em->mpirank == p4est->mpirank) {
      int off_p = 0; /* p-order offset of this sub-face (:771-796) */
      for (int b = 0; b < sub_p[f]; b++) off_p += nq2_p[b];
      rec_t r; memset(&r, 0, sizeof r);
      r.elem = em->id;
      r.hm.nbr = elem_ext_index(R, ep);
      r.hm.f_m = (int8_t)f_m; r.hm.f_p = (int8_t)f_p; r.hm.tcode = (int8_t)tcode;
      r.hm.child_m = (int8_t)((faces_m < faces_mortar) ? f : -1);
      r.hm.child_p = (int8_t)((faces_p < faces_mortar) ? sub_p[f] : -1);
      r.hm.degq = (int8_t)degq_m[f]; r.hm.face_slot = (int8_t)f_m;
      r.hm.fac_off = (int64_t)R->nfac; r.hm.g_off = -1;
      r.nfac = 7 * n2;
      double *F = grow(&R->fac, &R->nfac, &R->capfac, (size_t)7 * n2);
      double half_m = (faces_m < faces_mortar) ? 0.5 : 1.0; /* :805-808, sipg.c:945 */
      double half_p = (faces_p < faces_mortar) ? 0.5 : 1.0; /* :810-813 */
      for (int b = 0; b < nq; b++)
        for (int a = 0; a < nq; a++) {
          int k = a + nq * b, ks = off + k;
          int kp = off_p + tcode_src(tcode, nq, a, b);
          for (int l = 0; l < 3; l++) {
            double am = 0., ap = 0.;
            for (int d = 0; d < 3; d++) {
              double sn = sj[ks] * nn[(size_t)d * total + ks];
              am += sn * dm[(size_t)(l + 3 * d) * total + ks];
              ap += sn * dp[(size_t)(l + 3 * d) * total + kp];
            }
            F[(size_t)l * n2 + k] = half_m * am;
            F[(size_t)(3 + l) * n2 + k] = half_p * ap;
          }
          F[(size_t)6 * n2 + k] =
              sj[ks] * penalty_sigma(R, e_m[0], e_p[0], em->deg, hm[ks], ep->deg, hp[ks]);
        }
      push_rec(R, &r);
    }
    off += n2;
  }
  (void)mortar_side_id_m; (void)mortar_side_id_p; (void)e_m_is_ghost; (void)e_p_is_ghost;
  (void)geom; (void)quad;
}

/* Robin boundary face: Au += Lift(E^T w sj (alpha u - beta)), no gradient terms
 * (d4est_laplacian_with_opt_flux_sipg_robin_aux, dGMath/d4est_laplacian_with_opt_flux_sipg.c:377-476; twin
 * d4est_laplacian_flux_sipg_robin_aux).  alpha = robin_coeff and beta = robin_rhs are host callbacks of the mortar
 * point coordinates; they are evaluated ONCE here.  The callbacks also receive the reference's boundary-data struct:
 * it is handed over with its geometric members filled (coordinates, normals, sj, h, metric) and the members that
 * depend on u left NULL -- a coefficient that reads u would make the operator nonlinear and cannot be tabulated. */
static void record_boundary_robin(recorder_t *R, d4est_element_data_t *e_m, int f_m, int mortar_side_id_m,
                                  d4est_mesh_data_t *factors)
{
  d4est_laplacian_with_opt_robin_bc_t *bc = (d4est_laplacian_with_opt_robin_bc_t *)R->bc_data;
  int degq = e_m->deg_quad, nq = degq + 1, n2 = nq * nq;
  size_t s0 = (size_t)factors->local_strides[e_m->id].mortar_quad_stride[f_m];
  size_t b0 = (size_t)3 * factors->local_strides[e_m->id].boundary_quad_stride[f_m];
  double *sj = &factors->sj_m_mortar_quad[s0];
  d4est_laplacian_with_opt_flux_boundary_data_t bd; memset(&bd, 0, sizeof bd);
  d4est_quadrature_mortar_t fo;
  fo.dq = e_m->dq; fo.tree = e_m->tree; fo.face = f_m; fo.mortar_side_id = mortar_side_id_m;
  fo.mortar_subface_id = 0; fo.q[0] = e_m->q[0]; fo.q[1] = e_m->q[1]; fo.q[2] = e_m->q[2];
  bd.face_object = &fo; bd.deg_mortar_quad = degq; bd.face_nodes_m_quad = n2;
  bd.face_nodes_m_lobatto = d4est_lgl_get_nodes(2, e_m->deg);
  bd.sj_on_f_m_quad = sj; bd.h_quad = &factors->hm_mortar_quad[s0];
  for (int d = 0; d < 3; d++) {
    bd.n_on_f_m_quad[d] = &factors->n_m_mortar_quad[3 * s0 + (size_t)d * n2];
    bd.xyz_on_f_m_quad[d] = &factors->xyz_m_mortar_quad[b0 + (size_t)d * n2];
    bd.xyz_on_f_m_lobatto[d] = &factors->xyz_m_mortar_lobatto[b0 + (size_t)d * n2];
    for (int d2 = 0; d2 < 3; d2++) bd.drst_dxyz_quad[d][d2] = &factors->drst_dxyz_m_mortar_quad[9 * s0 + (size_t)(d + 3 * d2) * n2];
  }
  rec_t r; memset(&r, 0, sizeof r);
  r.elem = e_m->id;
  r.hm.nbr = -1; r.hm.f_m = (int8_t)f_m; r.hm.f_p = -1; r.hm.tcode = 1; r.hm.pad = 1;
  r.hm.child_m = -1; r.hm.child_p = -1; r.hm.degq = (int8_t)degq; r.hm.face_slot = (int8_t)f_m;
  r.hm.fac_off = (int64_t)R->nfac; r.hm.g_off = (int64_t)R->ng; r.nfac = 4 * n2;
  double *F = grow(&R->fac, &R->nfac, &R->capfac, (size_t)4 * n2);
  double *g = grow(&R->g, &R->ng, &R->capg, (size_t)n2);
  for (int k = 0; k < n2; k++) {
    double x = bd.xyz_on_f_m_quad[0][k], y = bd.xyz_on_f_m_quad[1][k], z = bd.xyz_on_f_m_quad[2][k];
    F[k] = F[n2 + k] = F[2 * n2 + k] = 0.;
    /* the twin's callbacks expect d4est_laplacian_flux_boundary_data_t, a different layout: they get NULL */
    F[3 * n2 + k] = sj[k] * bc->robin_coeff(x, y, z, bc->user, R->baseline ? NULL : &bd, k);
    g[k] = sj[k] * bc->robin_rhs(x, y, z, bc->user, R->baseline ? NULL : &bd, k);
  }
  push_rec(R, &r);
}

/* Recording twin of d4est_laplacian_with_opt_flux_boundary (…_with_opt_flux.c:25) +
 * d4est_laplacian_with_opt_flux_sipg_dirichlet_aux (…_flux_sipg.c:15-191). */
static void record_boundary(p4est_t *p4est, d4est_element_data_t *e_m, int f_m,
                            int mortar_side_id_m, d4est_operators_t *ops,
                            d4est_geometry_t *geom, d4est_quadrature_t *quad,
                            d4est_mesh_data_t *factors, void *ctx)
{
  recorder_t *R = (recorder_t *)ctx;
  if (R->bc_type == BC_ROBIN) { record_boundary_robin(R, e_m, f_m, mortar_side_id_m, factors); return; }
  if (R->bc_type != BC_DIRICHLET)
    D4EST_ABORT("d4est_b200: boundary condition type not implemented on the device path");
  d4est_laplacian_with_opt_dirichlet_bc_t *bc = (d4est_laplacian_with_opt_dirichlet_bc_t *)R->bc_data;
  int degq = e_m->deg_quad, nq = degq + 1, n2 = nq * nq;
  int nl2 = d4est_lgl_get_nodes(2, e_m->deg);
  size_t s0 = (size_t)factors->local_strides[e_m->id].mortar_quad_stride[f_m];
  size_t b0 = (size_t)3 * factors->local_strides[e_m->id].boundary_quad_stride[f_m];
  const double *sj = &factors->sj_m_mortar_quad[s0];
  const double *hq = &factors->hm_mortar_quad[s0];
  const double *nn = &factors->n_m_mortar_quad[3 * s0];
  const double *dm = &factors->drst_dxyz_m_mortar_quad[9 * s0];
  rec_t r; memset(&r, 0, sizeof r);
  r.elem = e_m->id;
  r.hm.nbr = -1; r.hm.f_m = (int8_t)f_m; r.hm.f_p = -1; r.hm.tcode = 1;
  r.hm.child_m = -1; r.hm.child_p = -1; r.hm.degq = (int8_t)degq; r.hm.face_slot = (int8_t)f_m;
  r.hm.fac_off = (int64_t)R->nfac; r.hm.g_off = -1; r.nfac = 4 * n2;
  double *F = grow(&R->fac, &R->nfac, &R->capfac, (size_t)4 * n2);
  for (int k = 0; k < n2; k++) {
    for (int l = 0; l < 3; l++) {
      double am = 0.;
      for (int d = 0; d < 3; d++) am += sj[k] * nn[(size_t)d * n2 + k] * dm[(size_t)(l + 3 * d) * n2 + k];
      F[(size_t)l * n2 + k] = am;
    }
    F[(size_t)3 * n2 + k] = sj[k] * penalty_sigma(R, e_m, e_m, e_m->deg, hq[k], e_m->deg, hq[k]);
  }
  if (bc->dirichlet_fcn != zero_fcn) { /* inhomogeneous data, evaluated once on the host */
    r.hm.g_off = (int64_t)R->ng;
    double *g = grow(&R->g, &R->ng, &R->capg, (size_t)n2);
    if (bc->eval_method == EVAL_BNDRY_FCN_ON_LOBATTO) { /* sipg.c:88-124 */
      const double *xl = &factors->xyz_m_mortar_lobatto[b0];
      double gl[D4B_OPS_MAXN * D4B_OPS_MAXN];
      for (int i = 0; i < nl2; i++)
        gl[i] = bc->dirichlet_fcn(xl[i], xl[n2 + i], xl[2 * n2 + i], bc->user);
      d4est_quadrature_mortar_t fo;
      fo.dq = e_m->dq; fo.tree = e_m->tree; fo.face = f_m; fo.mortar_side_id = mortar_side_id_m;
      fo.mortar_subface_id = 0; fo.q[0] = e_m->q[0]; fo.q[1] = e_m->q[1]; fo.q[2] = e_m->q[2];
      d4est_quadrature_interpolate(ops, quad, geom, &fo, QUAD_OBJECT_MORTAR, QUAD_INTEGRAND_UNKNOWN,
                                   gl, e_m->deg, g, degq);
    } else { /* EVAL_BNDRY_FCN_ON_QUAD sipg.c:125-138 */
      const double *xq = &factors->xyz_m_mortar_quad[b0];
      for (int k = 0; k < n2; k++)
        g[k] = bc->dirichlet_fcn(xq[k], xq[n2 + k], xq[2 * n2 + k], bc->user);
    }
  }
  push_rec(R, &r);
  (void)p4est;
}

/* -------------------------------------------------------------------- flatten */
static d4b_mesh_desc_t *flatten_impl(p4est_t *p4est, d4est_ghost_t *ghost, void *sipg_params,
                                     d4est_laplacian_bc_t bc_type, void *bc_data, int baseline, d4est_operators_t *ops,
                                     d4est_geometry_t *geom, d4est_quadrature_t *quad, d4est_mesh_data_t *factors);

d4b_mesh_desc_t *d4est_b200_flatten(p4est_t *p4est, d4est_ghost_t *ghost,
                                    d4est_laplacian_with_opt_flux_data_t *flux_data,
                                    d4est_operators_t *ops, d4est_geometry_t *geom,
                                    d4est_quadrature_t *quad, d4est_mesh_data_t *factors)
{
  if (flux_data->flux_type != FLUX_SIPG) D4EST_ABORT("d4est_b200: only the SIPG flux is implemented");
  return flatten_impl(p4est, ghost, flux_data->flux_data, flux_data->bc_type, flux_data->bc_data, 0, ops, geom, quad, factors);
}

/* Same tables from the baseline twin's flux data (d4est_laplacian_flux_data_t, dGMath/d4est_laplacian_flux.h):
 * its sipg parameter struct has the same leading fields (penalty prefactor, region multipliers, penalty function,
 * dGMath/d4est_laplacian_flux_sipg.h:18-40 vs ..._with_opt_flux_sipg.h:18-40) and its Dirichlet struct the same
 * three members.  The twin evaluates every interface twice and uses max(deg) for the mortar Lobatto space; the two
 * variants agree to 1e-13 (Tests/Unit/d4est_test_laplacian_speedup.c:485), so one device path serves both. */
d4b_mesh_desc_t *d4est_b200_flatten_baseline(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_flux_data_t *fd,
                                             d4est_operators_t *ops, d4est_geometry_t *geom,
                                             d4est_quadrature_t *quad, d4est_mesh_data_t *factors)
{
  if (fd->flux_type != FLUX_SIPG) D4EST_ABORT("d4est_b200: only the SIPG flux is implemented");
  return flatten_impl(p4est, ghost, fd->flux_data, fd->bc_type, fd->bc_data, 1, ops, geom, quad, factors);
}

static d4b_mesh_desc_t *flatten_impl(p4est_t *p4est, d4est_ghost_t *ghost, void *sipg_params,
                                     d4est_laplacian_bc_t bc_type, void *bc_data, int baseline, d4est_operators_t *ops,
                                     d4est_geometry_t *geom, d4est_quadrature_t *quad, d4est_mesh_data_t *factors)
{
  int nl = (int)p4est->local_num_quadrants;
  int ng = (int)ghost->ghost->ghosts.elem_count;
  d4b_mesh_desc_t *D = (d4b_mesh_desc_t *)calloc(1, sizeof *D);
  if (quad->quad_type == QUAD_TYPE_GAUSS_LEGENDRE) D->quad_type = D4B_QUAD_GAUSS_LEGENDRE;
  else if (quad->quad_type == QUAD_TYPE_GAUSS_LEGENDRE_LOBATTO) D->quad_type = D4B_QUAD_GAUSS_LOBATTO;
  else D4EST_ABORT("d4est_b200: unsupported quadrature type");
  D->n_local = nl; D->n_ghost = ng;
  int32_t *deg = (int32_t *)malloc((size_t)(nl + ng) * sizeof(int32_t));
  int64_t *node_off = (int64_t *)malloc((size_t)(nl + ng + 1) * sizeof(int64_t));
  int32_t *degq = (int32_t *)malloc((size_t)(nl ? nl : 1) * sizeof(int32_t));
  int64_t *quad_off = (int64_t *)malloc((size_t)(nl + 1) * sizeof(int64_t));
  d4est_element_data_t **eds = (d4est_element_data_t **)malloc((size_t)(nl ? nl : 1) * sizeof(void *));

  int64_t nodes = 0, qn = 0; int e = 0;
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    p4est_tree_t *tree = p4est_tree_array_index(p4est->trees, tt);
    sc_array_t *tq = &tree->quadrants;
    for (size_t q = 0; q < tq->elem_count; ++q, ++e) {
      d4est_element_data_t *ed = (d4est_element_data_t *)p4est_quadrant_array_index(tq, q)->p.user_data;
      if (ed->id != e || ed->nodal_stride != nodes || ed->quad_stride != qn)
        D4EST_ABORT("d4est_b200: element ids/strides are not in p4est local order");
      if (ed->deg > D4B_MAX_DEG || ed->deg_quad > D4B_MAX_DEGQ) D4EST_ABORT("d4est_b200: degree too high");
      eds[e] = ed; deg[e] = ed->deg; degq[e] = ed->deg_quad;
      node_off[e] = nodes; quad_off[e] = qn;
      nodes += d4est_lgl_get_nodes(3, ed->deg); qn += d4est_lgl_get_nodes(3, ed->deg_quad);
    }
  }
  node_off[nl] = nodes; quad_off[nl] = qn;
  for (int g = 0; g < ng; g++) { /* ghost u follows local u (Mesh/d4est_ghost.c:54-70) */
    d4est_element_data_t *ged = &ghost->ghost_elements[g];
    deg[nl + g] = ged->deg;
    node_off[nl + g + 1] = node_off[nl + g] + d4est_lgl_get_nodes(3, ged->deg);
  }

  /* volume metric G = w J (dr/dx)(dr/dx)^T  (Quadrature/d4est_quadrature.c:339-375) */
  double *G = (double *)malloc((size_t)(6 * qn ? 6 * qn : 1) * sizeof(double));
  for (e = 0; e < nl; e++) {
    d4est_element_data_t *ed = eds[e];
    int nq = ed->deg_quad + 1, nq3 = nq * nq * nq;
    double xq[D4B_OPS_MAXN], wq[D4B_OPS_MAXN];
    d4b_ops_quad(ed->deg_quad, D->quad_type, xq, wq);
    double *J = d4est_mesh_get_jacobian_on_quadrature_points(factors, ed);
    d4est_mesh_data_on_element_t md = d4est_mesh_data_on_element(factors, ed);
    double *Ge = &G[6 * quad_off[e]];
    for (int k = 0; k < nq3; k++) {
      double w = wq[k % nq] * wq[(k / nq) % nq] * wq[k / (nq * nq)] * J[k];
      int c = 0;
      for (int l = 0; l < 3; l++)
        for (int lp = l; lp < 3; lp++, c++) {
          double s = 0.;
          for (int d = 0; d < 3; d++) s += md.rst_xyz_quad[l][d][k] * md.rst_xyz_quad[lp][d][k];
          Ge[(size_t)c * nq3 + k] = w * s;
        }
    }
  }

  /* faces: one walk of the reference iterator with recording callbacks */
  recorder_t R; memset(&R, 0, sizeof R);
  R.p4est = p4est; R.ghost = ghost; R.n_local = nl; R.quad_type = D->quad_type;
  R.sp = (d4est_laplacian_with_opt_flux_sipg_params_t *)sipg_params; R.bc_type = bc_type; R.bc_data = bc_data; R.baseline = baseline;
  d4est_mortars_with_opt_fcn_ptrs_t fp;
  fp.flux_interface_fcn = record_interface; fp.flux_boundary_fcn = record_boundary; fp.user_ctx = &R;
  d4est_mortars_with_opt_compute_flux_on_local_elements(p4est, ghost, ops, geom, quad, factors, &fp);

  /* stable counting sort by owning element; repack factor / g pools in that order */
  int32_t *hb = (int32_t *)calloc((size_t)nl + 2, sizeof(int32_t));
  for (size_t i = 0; i < R.nrec; i++) hb[R.rec[i].elem + 1]++;
  for (e = 0; e < nl; e++) hb[e + 1] += hb[e];
  int32_t *cur = (int32_t *)malloc((size_t)(nl + 1) * sizeof(int32_t));
  memcpy(cur, hb, (size_t)(nl + 1) * sizeof(int32_t));
  size_t *order = (size_t *)malloc((R.nrec ? R.nrec : 1) * sizeof(size_t));
  for (size_t i = 0; i < R.nrec; i++) order[cur[R.rec[i].elem]++] = i;
  d4b_hm_t *hm = (d4b_hm_t *)malloc((R.nrec ? R.nrec : 1) * sizeof(d4b_hm_t));
  double *fac = (double *)malloc((R.nfac ? R.nfac : 1) * sizeof(double));
  double *gq = R.ng ? (double *)malloc(R.ng * sizeof(double)) : NULL;
  size_t fo = 0, go = 0;
  for (size_t j = 0; j < R.nrec; j++) {
    rec_t *r = &R.rec[order[j]];
    hm[j] = r->hm;
    memcpy(&fac[fo], &R.fac[r->hm.fac_off], (size_t)r->nfac * sizeof(double));
    hm[j].fac_off = (int64_t)fo; fo += (size_t)r->nfac;
    if (r->hm.g_off >= 0) {
      int n2 = (r->hm.degq + 1) * (r->hm.degq + 1);
      memcpy(&gq[go], &R.g[r->hm.g_off], (size_t)n2 * sizeof(double));
      hm[j].g_off = (int64_t)go; go += (size_t)n2;
    }
  }
  free(R.rec); free(R.fac); free(R.g); free(cur); free(order); free(eds);
  D->deg = deg; D->node_off = node_off; D->degq = degq; D->quad_off = quad_off; D->G = G;
  D->hm_begin = hb; D->hm = hm; D->fac = fac; D->n_fac = (int64_t)fo; D->bc_g = gq; D->n_bc_g = (int64_t)go;

  /* halo lists from p4est's ghost layer (send = mirrors per rank, recv = ghosts per rank);
   * reference exchange: Mesh/d4est_ghost_data.c:143-255 */
  p4est_ghost_t *gl = ghost->ghost;
  int np = 0;
  for (int p = 0; p < p4est->mpisize; p++)
    if (gl->proc_offsets[p + 1] > gl->proc_offsets[p] ||
        gl->mirror_proc_offsets[p + 1] > gl->mirror_proc_offsets[p]) np++;
  int32_t *peer = (int32_t *)malloc((size_t)(np ? np : 1) * sizeof(int32_t));
  int32_t *sb = (int32_t *)calloc((size_t)np + 1, sizeof(int32_t));
  int32_t *rb = (int32_t *)calloc((size_t)np + 1, sizeof(int32_t));
  int nsend = (int)gl->mirror_proc_offsets[p4est->mpisize];
  int32_t *se = (int32_t *)malloc((size_t)(nsend ? nsend : 1) * sizeof(int32_t));
  int ip = 0, is = 0;
  for (int p = 0; p < p4est->mpisize; p++) {
    int nr = (int)(gl->proc_offsets[p + 1] - gl->proc_offsets[p]);
    int ns = (int)(gl->mirror_proc_offsets[p + 1] - gl->mirror_proc_offsets[p]);
    if (!nr && !ns) continue;
    peer[ip] = p; rb[ip] = (int32_t)gl->proc_offsets[p]; rb[ip + 1] = (int32_t)gl->proc_offsets[p + 1];
    sb[ip] = is;
    for (int j = 0; j < ns; j++) {
      p4est_locidx_t mi = gl->mirror_proc_mirrors[gl->mirror_proc_offsets[p] + j];
      p4est_quadrant_t *mq = p4est_quadrant_array_index(&gl->mirrors, (size_t)mi);
      se[is++] = (int32_t)mq->p.piggy3.local_num;
    }
    sb[ip + 1] = is; ip++;
  }
  D->n_peers = np; D->peer_rank = peer; D->send_begin = sb; D->send_elem = se; D->recv_begin = rb;
  return D;
}

void d4est_b200_flatten_free(d4b_mesh_desc_t *D)
{
  if (!D) return;
  free((void *)D->deg); free((void *)D->node_off); free((void *)D->degq); free((void *)D->quad_off);
  free((void *)D->G); free((void *)D->hm_begin); free((void *)D->hm); free((void *)D->fac);
  free((void *)D->bc_g); free((void *)D->peer_rank); free((void *)D->send_begin);
  free((void *)D->send_elem); free((void *)D->recv_begin); free(D);
}

/* ------------------------------------------------------------- level residency */
#define D4B_GLUE_MAX_LEVELS 64
static struct { d4est_mesh_data_t *key; d4b_mesh_t *mesh; } g_levels[D4B_GLUE_MAX_LEVELS];

static d4b_mesh_t *level_lookup(d4est_mesh_data_t *factors, int *free_slot)
{
  *free_slot = -1;
  for (int i = 0; i < D4B_GLUE_MAX_LEVELS; i++) {
    if (g_levels[i].key == factors) return g_levels[i].mesh;
    if (*free_slot < 0 && !g_levels[i].key) *free_slot = i;
  }
  if (*free_slot < 0) D4EST_ABORT("d4est_b200: too many resident levels");
  return NULL;
}
static d4b_mesh_t *level_store(int slot, d4est_mesh_data_t *factors, d4b_mesh_desc_t *D)
{
  d4b_mesh_t *m = NULL;
  if (d4b_mesh_create(D, &m) != D4B_OK) {
    fprintf(stderr, "d4est_b200: %s\n", d4b_last_error());
    D4EST_ABORT("d4est_b200: d4b_mesh_create failed (no CPU fallback)");
  }
  d4est_b200_flatten_free(D);
  g_levels[slot].key = factors; g_levels[slot].mesh = m;
  return m;
}

d4b_mesh_t *d4est_b200_level_get(p4est_t *p4est, d4est_ghost_t *ghost,
                                 d4est_laplacian_with_opt_flux_data_t *flux_data,
                                 d4est_operators_t *ops, d4est_geometry_t *geom,
                                 d4est_quadrature_t *quad, d4est_mesh_data_t *factors)
{
  int slot; d4b_mesh_t *m = level_lookup(factors, &slot);
  if (m) return m;
  return level_store(slot, factors, d4est_b200_flatten(p4est, ghost, flux_data, ops, geom, quad, factors));
}

d4b_mesh_t *d4est_b200_level_get_baseline(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_flux_data_t *fd,
                                          d4est_operators_t *ops, d4est_geometry_t *geom,
                                          d4est_quadrature_t *quad, d4est_mesh_data_t *factors)
{
  int slot; d4b_mesh_t *m = level_lookup(factors, &slot);
  if (m) return m;
  return level_store(slot, factors, d4est_b200_flatten_baseline(p4est, ghost, fd, ops, geom, quad, factors));
}

void d4est_b200_level_invalidate(d4est_mesh_data_t *factors)
{
  for (int i = 0; i < D4B_GLUE_MAX_LEVELS; i++)
    if (g_levels[i].key == factors) { d4b_mesh_destroy(g_levels[i].mesh); g_levels[i].key = NULL; g_levels[i].mesh = NULL; }
}
void d4est_b200_level_invalidate_all(void)
{
  for (int i = 0; i < D4B_GLUE_MAX_LEVELS; i++)
    if (g_levels[i].key) { d4b_mesh_destroy(g_levels[i].mesh); g_levels[i].key = NULL; g_levels[i].mesh = NULL; }
}

static void check_rc(int rc, const char *what)
{
  if (rc == D4B_OK) return;
  fprintf(stderr, "d4est_b200: %s: %s\n", what, d4b_last_error());
  D4EST_ABORT("d4est_b200: device call failed (no CPU fallback)");
}

/* ---------------------------------------------------------------- operator entry points */
void d4est_b200_laplacian_apply_aij(p4est_t *p4est, d4est_ghost_t *ghost,
                                    d4est_ghost_data_t *ghost_data, d4est_elliptic_data_t *vecs,
                                    d4est_laplacian_with_opt_flux_data_t *flux_data,
                                    d4est_operators_t *ops, d4est_geometry_t *geom,
                                    d4est_quadrature_t *quad, d4est_mesh_data_t *factors,
                                    int which_field)
{
  d4b_mesh_t *m = d4est_b200_level_get(p4est, ghost, flux_data, ops, geom, quad, factors);
  size_t off = (size_t)which_field * (size_t)vecs->local_nodes; /* …_with_opt.c:199 */
  check_rc(d4b_apply_host(m, vecs->u + off, vecs->Au + off), "apply");
  (void)ghost_data;
}

/* behind the baseline twin d4est_laplacian_apply_aij (dGMath/d4est_laplacian.c:317), same argument list */
void d4est_b200_laplacian_baseline_apply_aij(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                                             d4est_elliptic_data_t *vecs, d4est_laplacian_flux_data_t *flux_fcn_data,
                                             d4est_operators_t *ops, d4est_geometry_t *geom, d4est_quadrature_t *quad,
                                             d4est_mesh_data_t *factors, int which_field)
{
  d4b_mesh_t *m = d4est_b200_level_get_baseline(p4est, ghost, flux_fcn_data, ops, geom, quad, factors);
  size_t off = (size_t)which_field * (size_t)vecs->local_nodes;
  check_rc(d4b_apply_host(m, vecs->u + off, vecs->Au + off), "apply (baseline twin)");
  (void)ghost_data;
}

/* d4est_apply_operator_fcn_t (EllipticSystem/d4est_elliptic_eqns.h:12-24): install as prob_fcns.apply_lhs with
 * prob_fcns.user = the d4est_laplacian_with_opt_flux_data_t* of the problem */
void d4est_b200_apply_lhs(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                          d4est_elliptic_data_t *vecs, d4est_operators_t *ops, d4est_geometry_t *geom,
                          d4est_quadrature_t *quad, d4est_mesh_data_t *factors, void *user)
{
  d4est_b200_laplacian_apply_aij(p4est, ghost, ghost_data, vecs, (d4est_laplacian_with_opt_flux_data_t *)user, ops,
                                 geom, quad, factors, 0);
}

/* d4est_apply_operator_fcn_t for prob_fcns.build_residual (:32): Au = rhs - A u, the body of the Laplacian
 * problems' build_residual callbacks (Problems/Poisson/poisson_sinx_fcns.h:133-148), one upload of u and rhs */
void d4est_b200_build_residual(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                               d4est_elliptic_data_t *vecs, d4est_operators_t *ops, d4est_geometry_t *geom,
                               d4est_quadrature_t *quad, d4est_mesh_data_t *factors, void *user)
{
  d4b_mesh_t *m = d4est_b200_level_get(p4est, ghost, (d4est_laplacian_with_opt_flux_data_t *)user, ops, geom, quad, factors);
  check_rc(d4b_residual_host(m, vecs->u, vecs->rhs, vecs->Au), "build_residual");
  (void)ghost_data;
}

/* d4est_laplacian_build_rhs_with_strong_bc (dGMath/d4est_laplacian.c:16-134), same argument list:
 * rhs = M f - A(u = 0), the second term carrying the inhomogeneous Dirichlet data.  A(0) runs on the device; the
 * mass-matrix product with the source term stays the reference's own host code (setup, once per solve). */
void d4est_b200_laplacian_build_rhs_with_strong_bc(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                                                   d4est_operators_t *ops, d4est_geometry_t *geom,
                                                   d4est_quadrature_t *quad, d4est_mesh_data_t *factors,
                                                   d4est_elliptic_data_t *prob_vecs,
                                                   d4est_laplacian_flux_data_t *flux_fcn_data_for_build_rhs,
                                                   double *rhs, d4est_xyz_fcn_t problem_rhs_fcn,
                                                   d4est_mesh_init_field_option_t init_option, void *ctx, int which_field)
{
  int n = prob_vecs->local_nodes;
  double *u0 = P4EST_ALLOC_ZERO(double, n), *Au0 = P4EST_ALLOC_ZERO(double, n);
  /* the boundary data belongs to THIS flux struct: flatten it separately from the (homogeneous) operator level */
  d4b_mesh_desc_t *D = d4est_b200_flatten_baseline(p4est, ghost, flux_fcn_data_for_build_rhs, ops, geom, quad, factors);
  d4b_mesh_t *m = NULL;
  check_rc(d4b_mesh_create(D, &m), "build_rhs level");
  d4est_b200_flatten_free(D);
  check_rc(d4b_apply_host(m, u0, Au0), "build_rhs apply");
  d4b_mesh_destroy(m);
  double *f = P4EST_ALLOC(double, init_option == INIT_FIELD_ON_LOBATTO ? n : d4est_mesh_get_local_quad_nodes(p4est));
  d4est_mesh_init_field(p4est, f, problem_rhs_fcn, ops, geom, factors, init_option, ctx);
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    sc_array_t *tq = &p4est_tree_array_index(p4est->trees, tt)->quadrants;
    for (size_t q = 0; q < tq->elem_count; ++q) {
      d4est_element_data_t *ed = (d4est_element_data_t *)p4est_quadrant_array_index(tq, q)->p.user_data;
      d4est_quadrature_volume_t mo;
      mo.dq = ed->dq; mo.tree = ed->tree; mo.q[0] = ed->q[0]; mo.q[1] = ed->q[1]; mo.q[2] = ed->q[2]; mo.element_id = ed->id;
      double *J = d4est_mesh_get_jacobian_on_quadrature_points(factors, ed);
      if (init_option == INIT_FIELD_ON_LOBATTO)
        d4est_quadrature_apply_mass_matrix(ops, geom, quad, &mo, QUAD_OBJECT_VOLUME, QUAD_INTEGRAND_UNKNOWN,
                                           &f[ed->nodal_stride], ed->deg, J, ed->deg_quad, &rhs[ed->nodal_stride]);
      else
        d4est_quadrature_apply_galerkin_integral(ops, geom, quad, &mo, QUAD_OBJECT_VOLUME, QUAD_INTEGRAND_UNKNOWN,
                                                 &f[ed->quad_stride], ed->deg, J, ed->deg_quad, &rhs[ed->nodal_stride]);
    }
  }
  for (int i = 0; i < n; i++) rhs[i] -= Au0[i];
  P4EST_FREE(u0); P4EST_FREE(Au0); P4EST_FREE(f);
  (void)ghost_data; (void)which_field;
}

/* ---------------------------------------------------------------- mass-type terms
 * c = J f(x, u0) at the volume quadrature points, tabulated once per linearisation point with the user's own
 * callback exactly as d4est_quadrature_apply_fofufofvlilj does per element (Quadrature/d4est_quadrature.c:629-683,
 * interpolate_f = 0): u0 is interpolated to the quadrature points, f is evaluated there.  The device then applies
 * I^T diag(w c) I (the d4est_quadrature_apply_mass_matrix the reference ends in, :735-750) for every element at once. */
void d4est_b200_mass_set_fofu(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_with_opt_flux_data_t *flux_data,
                              d4est_operators_t *ops, d4est_geometry_t *geom, d4est_quadrature_t *quad,
                              d4est_mesh_data_t *factors, double *u0, d4est_xyzu_fcn_t fofu_fcn, void *fofu_ctx)
{
  d4b_mesh_t *m = d4est_b200_level_get(p4est, ghost, flux_data, ops, geom, quad, factors);
  size_t nq_total = 0;
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    sc_array_t *tq = &p4est_tree_array_index(p4est->trees, tt)->quadrants;
    for (size_t q = 0; q < tq->elem_count; ++q)
      nq_total += (size_t)d4est_lgl_get_nodes(3, ((d4est_element_data_t *)p4est_quadrant_array_index(tq, q)->p.user_data)->deg_quad);
  }
  double *coef = P4EST_ALLOC(double, nq_total ? nq_total : 1);
  double uq[D4B_OPS_MAXN * D4B_OPS_MAXN * D4B_OPS_MAXN];
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    sc_array_t *tq = &p4est_tree_array_index(p4est->trees, tt)->quadrants;
    for (size_t q = 0; q < tq->elem_count; ++q) {
      d4est_element_data_t *ed = (d4est_element_data_t *)p4est_quadrant_array_index(tq, q)->p.user_data;
      int n3 = d4est_lgl_get_nodes(3, ed->deg_quad);
      d4est_quadrature_volume_t mo;
      mo.dq = ed->dq; mo.tree = ed->tree; mo.q[0] = ed->q[0]; mo.q[1] = ed->q[1]; mo.q[2] = ed->q[2]; mo.element_id = ed->id;
      d4est_mesh_data_on_element_t md = d4est_mesh_data_on_element(factors, ed);
      double *J = d4est_mesh_get_jacobian_on_quadrature_points(factors, ed);
      if (u0)
        d4est_quadrature_interpolate(ops, quad, geom, &mo, QUAD_OBJECT_VOLUME, QUAD_INTEGRAND_UNKNOWN,
                                     &u0[ed->nodal_stride], ed->deg, uq, ed->deg_quad);
      double *c = &coef[ed->quad_stride];
      for (int i = 0; i < n3; i++)
        c[i] = J[i] * (fofu_fcn ? fofu_fcn(md.xyz_quad[0][i], md.xyz_quad[1][i], md.xyz_quad[2][i], u0 ? uq[i] : 0., fofu_ctx) : 1.);
    }
  }
  check_rc(d4b_mass_set_coef(m, coef), "mass coefficient");
  P4EST_FREE(coef);
}

/* vecs->Au += M_c vecs->u: the loop over elements of e.g. constant_density_star_fcns.h:540-603 in one device call */
void d4est_b200_mass_apply_add(p4est_t *p4est, d4est_ghost_t *ghost, d4est_laplacian_with_opt_flux_data_t *flux_data,
                               d4est_elliptic_data_t *vecs, d4est_operators_t *ops, d4est_geometry_t *geom,
                               d4est_quadrature_t *quad, d4est_mesh_data_t *factors)
{
  d4b_mesh_t *m = d4est_b200_level_get(p4est, ghost, flux_data, ops, geom, quad, factors);
  check_rc(d4b_mass_apply_host(m, vecs->u, vecs->Au, 1), "mass apply");
}
