/* d4b_mg.cu -- multigrid on cached device levels: element-local restriction / prolongation and the
 * V-cycle driver.  The reference defines its hierarchy by MUTATING the live p4est every cycle
 * (p4est_coarsen_ext / p4est_refine_ext, Solver/d4est_solver_multigrid.c:984,1199) and transfers in
 * the callbacks (Solver/d4est_solver_multigrid_callbacks.h:109 restriction r_H = P^T r_h, :236
 * prolongation e_h = P e_H, coarse degree = min over children :50-75).  Here the host walks that
 * hierarchy ONCE, emits one transfer table per level pair, and every level stays resident. */
#include "d4b_internal.h"
#include "d4b_ops.h"
#include <vector>
#include <cmath>
#include <cstdlib>
#include <cstring>

/* 1-D prolongation pool: P[degH][degh][variant] (Nh x NH), variant 0 = same cell (p-prolong,
 * dGMath/d4est_operators.c:995), 1/2 = lower/upper child (hp-prolong :944-993) */
struct mg_pool { int off[D4B_NDEG][D4B_NDEG][3]; const double *pool; };
__constant__ mg_pool c_mg;
static bool g_mg_ready = false;

static int mg_pool_init()
{
  if (g_mg_ready) return D4B_OK;
  static mg_pool h;
  std::vector<double> pool;
  double buf[D4B_OPS_MAXN * D4B_OPS_MAXN];
  for (int dH = 0; dH < D4B_NDEG; dH++)
    for (int dh = 0; dh < D4B_NDEG; dh++)
      for (int v = 0; v < 3; v++) {
        h.off[dH][dh][v] = (int)pool.size();
        if (dH < 1 || dh < dH) continue;
        d4b_ops_prolong(dH, dh, v - 1, buf);
        pool.insert(pool.end(), buf, buf + (dh + 1) * (dH + 1));
      }
  double *d = nullptr;
  D4B_CUDA(cudaMalloc((void **)&d, pool.size() * sizeof(double)));
  D4B_CUDA(cudaMemcpy(d, pool.data(), pool.size() * sizeof(double), cudaMemcpyHostToDevice));
  h.pool = d;
  D4B_CUDA(cudaMemcpyToSymbol(c_mg, &h, sizeof h));
  g_mg_ready = true;
  return D4B_OK;
}

struct d4b_mg_transfer {
  int n_coarse;
  int32_t *kind, *degH, *child_deg;     /* device: [nH], [nH], [8 nH] */
  int64_t *coarse_off, *child_off;      /* device: [nH], [8 nH]       */
  int max_n;
};

/* y(dims with axis ax -> rows) = M applied along ax; M(r,k) = M[r*rs + k*cs] */
__device__ __forceinline__ void mg_contract(const double *__restrict__ M, int rs, int cs, int rows, int cols,
                                            const double *in, double *out, int n0, int n1, int n2, int ax, bool acc)
{
  const int o0 = ax == 0 ? rows : n0, o1 = ax == 1 ? rows : n1, o2 = ax == 2 ? rows : n2;
  const int total = o0 * o1 * o2, stride = ax == 0 ? 1 : (ax == 1 ? n0 : n0 * n1);
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    int i0 = idx % o0, i1 = (idx / o0) % o1, i2 = idx / (o0 * o1);
    int r = ax == 0 ? i0 : (ax == 1 ? i1 : i2);
    int base = (ax == 0 ? 0 : i0) + n0 * ((ax == 1 ? 0 : i1) + n1 * (ax == 2 ? 0 : i2));
    double s = 0.;
    for (int k = 0; k < cols; k++) s = fma(__ldg(&M[r * rs + k * cs]), in[base + k * stride], s);
    out[idx] = acc ? out[idx] + s : s;
  }
  __syncthreads();
}

/* one CTA per coarse element.  dir = 0: restriction xH = sum_c P_c^T xh_c;  dir = 1: prolongation */
__global__ void __launch_bounds__(128)
k_mg_transfer(d4b_mg_transfer T, int dir, const double *__restrict__ src, double *__restrict__ dst, int NX)
{
  extern __shared__ double sm[];
  const int e = blockIdx.x, NX3 = NX * NX * NX;
  double *a = sm, *b = a + NX3, *c = b + NX3;
  const int dH = T.degH[e], NH = dH + 1, NH3 = NH * NH * NH;
  const int nchild = T.kind[e] == 1 ? 8 : 1;
  if (dir == 0) for (int i = threadIdx.x; i < NH3; i += blockDim.x) c[i] = 0.;
  else for (int i = threadIdx.x; i < NH3; i += blockDim.x) c[i] = src[T.coarse_off[e] + i];
  __syncthreads();
  for (int ch = 0; ch < nchild; ch++) {
    const int dh = T.child_deg[8 * e + ch], Nh = dh + 1, Nh3 = Nh * Nh * Nh;
    const int64_t fo = T.child_off[8 * e + ch];
    const int vx = nchild == 8 ? 1 + (ch & 1) : 0, vy = nchild == 8 ? 1 + ((ch >> 1) & 1) : 0,
              vz = nchild == 8 ? 1 + ((ch >> 2) & 1) : 0;
    const double *Px = c_mg.pool + c_mg.off[dH][dh][vx], *Py = c_mg.pool + c_mg.off[dH][dh][vy],
                 *Pz = c_mg.pool + c_mg.off[dH][dh][vz];
    if (dir == 0) { /* restriction: transposes (Nh -> NH per axis), accumulate over children */
      for (int i = threadIdx.x; i < Nh3; i += blockDim.x) a[i] = src[fo + i];
      __syncthreads();
      mg_contract(Px, 1, NH, NH, Nh, a, b, Nh, Nh, Nh, 0, false);
      mg_contract(Py, 1, NH, NH, Nh, b, a, NH, Nh, Nh, 1, false);
      mg_contract(Pz, 1, NH, NH, Nh, a, c, NH, NH, Nh, 2, true);
    } else {        /* prolongation */
      mg_contract(Px, NH, 1, Nh, NH, c, a, NH, NH, NH, 0, false);
      mg_contract(Py, NH, 1, Nh, NH, a, b, Nh, NH, NH, 1, false);
      mg_contract(Pz, NH, 1, Nh, NH, b, a, Nh, Nh, NH, 2, false);
      for (int i = threadIdx.x; i < Nh3; i += blockDim.x) dst[fo + i] = a[i];
      __syncthreads();
    }
  }
  if (dir == 0) for (int i = threadIdx.x; i < NH3; i += blockDim.x) dst[T.coarse_off[e] + i] = c[i];
}

template <class T> static int up(T **dst, const T *src, size_t n)
{
  D4B_CUDA(cudaMalloc((void **)dst, (n ? n : 1) * sizeof(T)));
  if (n) D4B_CUDA(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
  return D4B_OK;
}

struct d4b_mg {
  int n_levels;
  d4b_mesh **lvl;            /* 0 = coarsest ... n_levels-1 = finest */
  d4b_mg_transfer **tr;      /* tr[l]: between level l (coarse) and l+1 (fine) */
  d4b_mg_params_t par;
  double *eigs;              /* DEVICE: per level spectral bound (cheby->eigs[level]) */
  double *u, *rhs, *r;       /* per-level stacked work vectors for levels < finest */
  int64_t *off;              /* offset of level l inside the stacks */
  double *work;              /* 3 * max nodes */
  double *hbuf;              /* 8 * finest nodes: staging + FCG work of the host-vector entry points (lazy) */
  int vcycles_done;
  /* CUDA graph of one preconditioner application (d4b_mg_pc_apply): captured on the second call with the same
   * (r, z, stream), replayed afterwards */
  cudaGraphExec_t gexec; const double *g_r; double *g_z; cudaStream_t g_st; int graph_state, pc_calls;
};

extern "C" {

int d4b_mg_transfer_create(int n_coarse, const int32_t *kind, const int32_t *degH, const int64_t *coarse_off,
                           const int32_t *child_deg, const int64_t *child_off, d4b_mg_transfer_t **out)
{
  int rc = mg_pool_init(); if (rc) return rc;
  d4b_mg_transfer *T = (d4b_mg_transfer *)calloc(1, sizeof *T);
  T->n_coarse = n_coarse;
  for (int e = 0; e < n_coarse; e++) {
    int nc = kind[e] == 1 ? 8 : 1;
    if (degH[e] < 1 || degH[e] > D4B_MAX_DEG) { d4b_set_error("mg: degH out of range"); free(T); return D4B_EINVAL; }
    if (degH[e] + 1 > T->max_n) T->max_n = degH[e] + 1;
    for (int c = 0; c < nc; c++) {
      int dh = child_deg[8 * e + c];
      if (dh < degH[e] || dh > D4B_MAX_DEG) { d4b_set_error("mg: child degree below coarse degree"); free(T); return D4B_EINVAL; }
      if (dh + 1 > T->max_n) T->max_n = dh + 1;
    }
  }
  if ((rc = up(&T->kind, kind, n_coarse)) || (rc = up(&T->degH, degH, n_coarse)) ||
      (rc = up(&T->coarse_off, coarse_off, n_coarse)) || (rc = up(&T->child_deg, child_deg, 8 * (size_t)n_coarse)) ||
      (rc = up(&T->child_off, child_off, 8 * (size_t)n_coarse))) return rc;
  *out = T;
  return D4B_OK;
}
void d4b_mg_transfer_destroy(d4b_mg_transfer_t *T)
{
  if (!T) return;
  cudaFree(T->kind); cudaFree(T->degH); cudaFree(T->coarse_off); cudaFree(T->child_deg); cudaFree(T->child_off); free(T);
}
static int transfer(d4b_mg_transfer *T, int dir, const double *src, double *dst, cudaStream_t st)
{
  if (T->n_coarse == 0) return D4B_OK;
  size_t smem = 3 * (size_t)T->max_n * T->max_n * T->max_n * sizeof(double);
  static size_t attr = 0;
  if (smem > 48 * 1024 && smem > attr) {
    D4B_CUDA(cudaFuncSetAttribute(k_mg_transfer, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = smem;
  }
  k_mg_transfer<<<T->n_coarse, 128, smem, st>>>(*T, dir, src, dst, T->max_n);
  D4B_CUDA(cudaGetLastError()); d4b_count_launch(1);
  return D4B_OK;
}
/* r_H = P^T r_h  (multigrid_callbacks.h:109-175) */
int d4b_mg_restrict(d4b_mg_transfer_t *T, const double *fine, double *coarse, void *stream)
{ return transfer(T, 0, fine, coarse, (cudaStream_t)stream); }
/* e_h = P e_H    (multigrid_callbacks.h:236-300) */
int d4b_mg_prolong(d4b_mg_transfer_t *T, const double *coarse, double *fine, void *stream)
{ return transfer(T, 1, coarse, fine, (cudaStream_t)stream); }

/* ---- AMR field transfer (second half): d4est_amr_interpolate_field (hpAMR/d4est_amr.c:406-485) carries
 * a nodal field from the old grid to the new one in two sweeps driven by the logs of d4est_amr_step: initial grid ->
 * refined grid (element i of degree initial_log[i]: refinement_log[i] = -deg -> 8 children of the same degree by
 * hp-prolongation, = deg' >= deg -> p-prolongation, :338-403) and refined grid -> 2:1-balanced grid (balance_log[i] =
 * -deg -> 8 children, +deg -> copy).  Both sweeps are the prolongation of a transfer table (k_mg_transfer, one CTA
 * per old element), elements and children in log order, which is the order of the nodal vectors. */
int d4b_amr_transfer_create(int n, const int32_t *deg_in, const int32_t *code, d4b_mg_transfer_t **out,
                            int64_t *nodes_in, int64_t *nodes_out)
{
  *out = nullptr;
  std::vector<int32_t> kind(n), degH(n), cdeg(8 * (size_t)n, 0);
  std::vector<int64_t> coff(n), choff(8 * (size_t)n, 0);
  int64_t ni = 0, no = 0;
  for (int e = 0; e < n; e++) {
    const int d = deg_in[e] < 0 ? -deg_in[e] : deg_in[e], c = code[e];
    if (d < 1 || d > D4B_MAX_DEG) { d4b_set_error("amr transfer: element %d has degree %d", e, d); return D4B_EINVAL; }
    const int64_t nin = (int64_t)(d + 1) * (d + 1) * (d + 1);
    degH[e] = d; coff[e] = ni; ni += nin;
    if (c < 0) {   /* h-refinement keeps the degree (the reference asserts it, d4est_amr.c:352) */
      if (-c != d) { d4b_set_error("amr transfer: element %d: h-refinement code %d with degree %d", e, c, d); return D4B_EINVAL; }
      kind[e] = 1;
      for (int k = 0; k < 8; k++) { cdeg[8 * (size_t)e + k] = d; choff[8 * (size_t)e + k] = no; no += nin; }
    } else {       /* p-refinement; coarsening is not supported by the reference either (it aborts, :395) */
      if (c < d || c > D4B_MAX_DEG) { d4b_set_error("amr transfer: element %d: code %d with degree %d (coarsening is not supported)", e, c, d); return D4B_EINVAL; }
      kind[e] = 0; cdeg[8 * (size_t)e] = c; choff[8 * (size_t)e] = no; no += (int64_t)(c + 1) * (c + 1) * (c + 1);
    }
  }
  if (nodes_in) *nodes_in = ni;
  if (nodes_out) *nodes_out = no;
  return d4b_mg_transfer_create(n, kind.data(), degH.data(), coff.data(), cdeg.data(), choff.data(), out);
}

/* field_out (n_out nodes, balanced grid) from field_in (n_in nodes, initial grid); device pointers; aux_dev holds
 * the refined, not yet balanced field (sum of the first sweep's output sizes, returned by the _sizes call) */
int d4b_amr_interpolate_field_sizes(int n_init, const int32_t *initial_log, const int32_t *refinement_log, int n_bal,
                                    const int32_t *balance_log, int64_t *n_in, int64_t *n_aux, int64_t *n_out)
{
  int64_t a = 0, b = 0, c = 0;
  for (int i = 0; i < n_init; i++) {
    const int d = initial_log[i], r = refinement_log[i], q = r < 0 ? -r : r;
    a += (int64_t)(d + 1) * (d + 1) * (d + 1);
    b += (int64_t)(q + 1) * (q + 1) * (q + 1) * (r < 0 ? 8 : 1);
  }
  int64_t b2 = 0;
  for (int i = 0; i < n_bal; i++) {
    const int r = balance_log[i], q = r < 0 ? -r : r;
    b2 += (int64_t)(q + 1) * (q + 1) * (q + 1);
    c += (int64_t)(q + 1) * (q + 1) * (q + 1) * (r < 0 ? 8 : 1);
  }
  if (b != b2) { d4b_set_error("amr transfer: the refinement log produces %lld nodes, the balance log consumes %lld", (long long)b, (long long)b2); return D4B_EINVAL; }
  if (n_in) *n_in = a;
  if (n_aux) *n_aux = b;
  if (n_out) *n_out = c;
  return D4B_OK;
}
int d4b_amr_interpolate_field_dev(int n_init, const int32_t *initial_log, const int32_t *refinement_log, int n_bal,
                                  const int32_t *balance_log, const double *in_dev, double *aux_dev, double *out_dev,
                                  void *stream)
{
  int rc;
  if ((rc = d4b_amr_interpolate_field_sizes(n_init, initial_log, refinement_log, n_bal, balance_log, nullptr, nullptr, nullptr))) return rc;
  d4b_mg_transfer_t *T1 = nullptr, *T2 = nullptr;
  if ((rc = d4b_amr_transfer_create(n_init, initial_log, refinement_log, &T1, nullptr, nullptr))) return rc;
  if ((rc = d4b_amr_transfer_create(n_bal, balance_log, balance_log, &T2, nullptr, nullptr))) { d4b_mg_transfer_destroy(T1); return rc; }
  rc = d4b_mg_prolong(T1, in_dev, aux_dev, stream);
  if (!rc) rc = d4b_mg_prolong(T2, aux_dev, out_dev, stream);
  cudaError_t ce = cudaStreamSynchronize((cudaStream_t)stream);   /* the tables are freed below */
  d4b_mg_transfer_destroy(T1); d4b_mg_transfer_destroy(T2);
  if (!rc && ce != cudaSuccess) { d4b_set_error("amr transfer: %s", cudaGetErrorString(ce)); rc = D4B_ECUDA; }
  return rc;
}
int d4b_amr_interpolate_field_host(int n_init, const int32_t *initial_log, const int32_t *refinement_log, int n_bal,
                                   const int32_t *balance_log, const double *field_in, int64_t n_in, double *field_out,
                                   int64_t n_out)
{
  int64_t a, b, c; int rc;
  if ((rc = d4b_amr_interpolate_field_sizes(n_init, initial_log, refinement_log, n_bal, balance_log, &a, &b, &c))) return rc;
  if (a != n_in || c != n_out) {
    d4b_set_error("amr transfer: the logs describe %lld -> %lld nodes, the caller passed %lld -> %lld", (long long)a, (long long)c, (long long)n_in, (long long)n_out);
    return D4B_EINVAL;
  }
  double *d = nullptr;
  D4B_CUDA(cudaMalloc((void **)&d, (size_t)(a + b + c + 1) * sizeof(double)));
  cudaError_t ce = cudaMemcpy(d, field_in, (size_t)a * sizeof(double), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) {
    rc = d4b_amr_interpolate_field_dev(n_init, initial_log, refinement_log, n_bal, balance_log, d, d + a, d + a + b, nullptr);
    if (!rc) ce = cudaMemcpy(field_out, d + a + b, (size_t)c * sizeof(double), cudaMemcpyDeviceToHost);
  }
  cudaFree(d);
  if (!rc && ce != cudaSuccess) { d4b_set_error("amr transfer: %s", cudaGetErrorString(ce)); rc = D4B_ECUDA; }
  return rc;
}

int d4b_mg_create(int n_levels, d4b_mesh_t **levels, d4b_mg_transfer_t **transfers, const d4b_mg_params_t *par,
                  d4b_mg_t **out)
{
  if (n_levels < 2) { d4b_set_error("multigrid needs >= 2 levels (multigrid.c:1434)"); return D4B_EINVAL; }
  d4b_mg *g = (d4b_mg *)calloc(1, sizeof *g);
  g->n_levels = n_levels; g->par = *par;
  g->lvl = (d4b_mesh **)malloc(n_levels * sizeof(void *)); g->tr = (d4b_mg_transfer **)malloc(n_levels * sizeof(void *));
  g->off = (int64_t *)calloc(n_levels + 1, sizeof(int64_t)); D4B_CUDA(cudaMalloc((void **)&g->eigs, n_levels * sizeof(double))); D4B_CUDA(cudaMemset(g->eigs, 0, n_levels * sizeof(double)));
  int64_t tot = 0, mx = 0;
  for (int l = 0; l < n_levels; l++) {
    g->lvl[l] = levels[l]; g->tr[l] = l < n_levels - 1 ? transfers[l] : nullptr;
    g->off[l] = tot; if (l < n_levels - 1) tot += levels[l]->local_nodes;
    if (levels[l]->local_nodes > mx) mx = levels[l]->local_nodes;
  }
  size_t nb = (size_t)(tot ? tot : 1) * sizeof(double);
  D4B_CUDA(cudaMalloc((void **)&g->u, nb)); D4B_CUDA(cudaMalloc((void **)&g->rhs, nb));
  D4B_CUDA(cudaMalloc((void **)&g->r, nb + levels[n_levels - 1]->local_nodes * sizeof(double)));
  D4B_CUDA(cudaMalloc((void **)&g->work, 3 * (size_t)mx * sizeof(double)));
  *out = g;
  return D4B_OK;
}
void d4b_mg_destroy(d4b_mg_t *g)
{
  if (!g) return;
  if (g->gexec) cudaGraphExecDestroy(g->gexec);
  cudaFree(g->u); cudaFree(g->rhs); cudaFree(g->r); cudaFree(g->work); cudaFree(g->hbuf);
  free(g->lvl); free(g->tr); free(g->off); cudaFree(g->eigs); free(g);
}
int d4b_mg_get_eigs(d4b_mg_t *g, double *eigs)
{ D4B_CUDA(cudaMemcpy(eigs, g->eigs, g->n_levels * sizeof(double), cudaMemcpyDeviceToHost)); return D4B_OK; }

} /* extern C */

/* smoother = [cg_eigs] + Chebyshev (smoother_cheby.c:264-360); `down` selects the reuse policy of
 * d4est_solver_multigrid_smoother_cheby_update (:221-262) */
static int mg_smooth(d4b_mg *g, int l, double *u, const double *rhs, double *r, int down, void *stream)
{
  const d4b_mg_params_t &p = g->par;
  int compute = down ? !(p.eigs_reuse_fromlastvcycle && g->vcycles_done != 0)
                     : !(p.eigs_reuse_fromdownvcycle || (p.eigs_reuse_fromlastvcycle && g->vcycles_done != 0));
  int rc;
  cudaStream_t st = (cudaStream_t)stream;
  /* nothing comes back to the host: the bound goes from the level's scalar block to g->eigs[l] (device) inside the
   * coefficient kernel of the Chebyshev sweep */
  if (compute) {
    if ((rc = d4b_cg_eigs_dev(g->lvl[l], u, rhs, p.cheby_eigs_cg_imax, p.use_new_cg_eigs, g->work, st))) return rc;
    return d4b_cheby_dev(g->lvl[l], u, rhs, r, p.cheby_imax, d4b_scalar_tmax(g->lvl[l]), p.cheby_eigs_max_multiplier,
                         g->eigs + l, p.cheby_eigs_lmax_lmin_ratio, g->work, st);
  }
  return d4b_cheby_dev(g->lvl[l], u, rhs, r, p.cheby_imax, g->eigs + l, 1.0, g->eigs + l, p.cheby_eigs_lmax_lmin_ratio, g->work, st);
}

extern "C" {

/* One V-cycle on the cached levels: restates d4est_solver_multigrid_vcycle
 * (Solver/d4est_solver_multigrid.c:751-1330).  u, rhs: finest-level device vectors. */
int d4b_mg_vcycle(d4b_mg_t *g, double *u, const double *rhs, double *r2_local, void *stream)
{
  const int top = g->n_levels - 1;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  double *rtop = g->r + g->off[top]; /* finest-level residual lives after the coarse stack */
  /* down sweep (:847-1110) */
  for (int l = top; l > 0; --l) {
    double *ul = l == top ? u : g->u + g->off[l];
    const double *fl = l == top ? rhs : g->rhs + g->off[l];
    double *rl = l == top ? rtop : g->r + g->off[l];
    if (l != top) D4B_CUDA(cudaMemsetAsync(ul, 0, g->lvl[l]->local_nodes * sizeof(double), st));
    if ((rc = mg_smooth(g, l, ul, fl, rl, 1, stream))) return rc;
    if ((rc = d4b_mg_restrict(g->tr[l - 1], rl, g->rhs + g->off[l - 1], stream))) return rc;
  }
  /* bottom solve (:1112-1160) */
  {
    d4b_mesh *m0 = g->lvl[0];
    double *u0 = g->u + g->off[0], *f0 = g->rhs + g->off[0], *r0 = g->r + g->off[0];
    const d4b_mg_params_t &p = g->par;
    D4B_CUDA(cudaMemsetAsync(u0, 0, m0->local_nodes * sizeof(double), st));
    if (p.bottom_kind == 1) { /* mg_bottom_solver_cg (bottom_solver_cg.c:59-190) */
      if ((rc = d4b_cg_solve(m0, u0, f0, p.bottom_imax, p.bottom_atol, p.bottom_rtol, nullptr, nullptr, g->work, stream))) return rc;
    } else {                  /* mg_bottom_solver_cheby (bottom_solver_cheby.c:59-113) */
      if ((rc = d4b_cg_eigs_dev(m0, u0, f0, p.bottom_eigs_cg_imax, p.use_new_cg_eigs, g->work, st))) return rc;
      if ((rc = d4b_cheby_dev(m0, u0, f0, r0, p.bottom_imax, d4b_scalar_tmax(m0), p.bottom_eigs_max_multiplier, g->eigs,
                              p.bottom_eigs_lmax_lmin_ratio, g->work, st))) return rc;
    }
  }
  /* up sweep (:1168-1300): prolong the correction, add, post-smooth */
  for (int l = 0; l < top; ++l) {
    double *uf = (l + 1 == top) ? u : g->u + g->off[l + 1];
    const double *ff = (l + 1 == top) ? rhs : g->rhs + g->off[l + 1];
    double *rf = (l + 1 == top) ? rtop : g->r + g->off[l + 1];
    if ((rc = d4b_mg_prolong(g->tr[l], g->u + g->off[l], rf, stream))) return rc;
    if ((rc = d4b_vec_axpy(g->lvl[l + 1]->local_nodes, 1.0, rf, uf, stream))) return rc;
    if ((rc = mg_smooth(g, l + 1, uf, ff, rf, 0, stream))) return rc;
  }
  if (r2_local) {   /* the only host read of a V-cycle, and only when the caller wants the residual */
    if ((rc = d4b_norm2_dev(g->lvl[top], rtop, st))) return rc;
    return d4b_read_r2(g->lvl[top], r2_local, st);
  }
  return D4B_OK;
}

/* V-cycles until the residual criterion: restates d4est_solver_multigrid_solve (:1420-1500) */
int d4b_mg_solve(d4b_mg_t *g, double *u, const double *rhs, int *vcycles, double *r2_final, void *stream)
{
  const int top = g->n_levels - 1; const int64_t n = g->lvl[top]->local_nodes;
  double *Au = g->work, *r = g->work + n, r2, r2_last;
  int rc;
  if ((rc = d4b_apply_dev(g->lvl[top], u, Au, stream))) return rc;
  if ((rc = d4b_vec_axpyeqz(n, -1., Au, rhs, r, stream))) return rc;
  if ((rc = d4b_vec_dot(n, r, r, &r2, stream))) return rc;
  r2_last = r2; g->vcycles_done = 0;
  const double stoptol = g->par.vcycle_rtol * g->par.vcycle_rtol * r2 + g->par.vcycle_atol * g->par.vcycle_atol;
  while (g->vcycles_done < g->par.vcycle_imax && r2 > stoptol) {
    /* the residual after a V-cycle is the one its last smoother call left behind (multigrid.c:1315,
     * d4est_solver_multigrid_compute_residual outside the START state): no extra apply */
    if ((rc = d4b_mg_vcycle(g, u, rhs, &r2, stream))) return rc;
    g->vcycles_done++;
    if (sqrt(r2 / r2_last) >= .99) break;
    r2_last = r2;
  }
  if (vcycles) *vcycles = g->vcycles_done;
  if (r2_final) *r2_final = r2;
  return D4B_OK;
}

int g_mg_graph = 0;   /* d4b_set_option("mg_graph", 1) */

/* preconditioner apply z = MG(r) from a zero guess: restates d4est_krylov_pc_multigrid_apply
 * (Solver/d4est_krylov_pc_multigrid.c:41-75); signature = d4b_pc_fcn_t with ctx = d4b_mg_t* */
int d4b_mg_pc_apply(void *ctx, const double *r_dev, double *z_dev, void *stream)
{
  d4b_mg *g = (d4b_mg *)ctx;
  cudaStream_t st = (cudaStream_t)stream;
  const bool one_cycle = g->par.vcycle_imax == 1 && g->par.vcycle_rtol < 1. && g->par.vcycle_atol == 0.;
  /* The single-V-cycle application enqueues the same few thousand launches, event records and waits every time and
   * reads nothing back: on small levels (MG coarse levels, N-rank runs) the host cannot issue them as fast as the GPU
   * retires them.  From the second application with the same vectors and stream it is therefore captured once into
   * a CUDA graph (the fork/join side streams, the communication stream and the NCCL calls join the capture through
   * their events) and replayed.  The first application runs uncaptured: it does every lazy set-up (per-bucket
   * metadata, function attributes, constant tables).
   * OPT-IN (D4B_MG_GRAPH=1 / option "mg_graph"): measured on the 50 245-element curved hp hierarchy on one GPU the
   * replay issues 9.7 k instead of 487 k host launches per solve but takes the same time (6.21 s against 6.23 s) --
   * the coarse levels are bound by the dependent chain of small kernels on the GPU, not by the host.  Never used
   * with neighbour ranks: capturing the NCCL exchange hung a 2-rank run.  A bottom solver with a host-side stopping
   * test (mg_bottom_solver_cg) cannot be captured either. */
  static const bool env_graph = getenv("D4B_MG_GRAPH") != nullptr;
  if (one_cycle && (env_graph || g_mg_graph) && d4b_comm_size() == 1 && g->par.bottom_kind != 1 && g->graph_state >= 0 &&
      g->pc_calls >= 1 && st != nullptr) {
    if (g->graph_state == 1 && (g->g_r != r_dev || g->g_z != z_dev || g->g_st != st)) {
      cudaGraphExecDestroy(g->gexec); g->gexec = nullptr; g->graph_state = 0;
    }
    if (g->graph_state == 0) {
      cudaGraph_t graph = nullptr;
      int rc = D4B_OK;
      if (cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) == cudaSuccess) {
        if (cudaMemsetAsync(z_dev, 0, g->lvl[g->n_levels - 1]->local_nodes * sizeof(double), st) != cudaSuccess) rc = D4B_ECUDA;
        g->vcycles_done = 0;
        if (rc == D4B_OK) rc = d4b_mg_vcycle(g, z_dev, r_dev, nullptr, stream);
        const cudaError_t ce = cudaStreamEndCapture(st, &graph);
        if (rc == D4B_OK && ce == cudaSuccess && graph &&
            cudaGraphInstantiate(&g->gexec, graph, 0) == cudaSuccess) {
          g->graph_state = 1; g->g_r = r_dev; g->g_z = z_dev; g->g_st = st;
        } else {
          g->graph_state = -1; g->gexec = nullptr;   /* this hierarchy stays on the plain path */
          cudaGetLastError();
        }
        if (graph) cudaGraphDestroy(graph);
      } else {
        g->graph_state = -1; cudaGetLastError();
      }
    }
    if (g->graph_state == 1) {
      D4B_CUDA(cudaGraphLaunch(g->gexec, st));
      d4b_count_launch(1);
      g->vcycles_done = 1; g->pc_calls++;
      return D4B_OK;
    }
  }
  g->pc_calls++;
  D4B_CUDA(cudaMemsetAsync(z_dev, 0, g->lvl[g->n_levels - 1]->local_nodes * sizeof(double), st));
  if (one_cycle) {
    /* d4est_solver_multigrid_solve from a zero guess with vcycle_imax = 1: the initial residual is r itself, the
     * loop test r2 > rtol^2 r2 + atol^2 holds for every nonzero r, so exactly one V-cycle runs (multigrid.c:1463);
     * neither the initial residual nor the final one is needed: nothing is read back, the whole application
     * enqueues without a host synchronisation */
    g->vcycles_done = 0;
    int rc = d4b_mg_vcycle(g, z_dev, r_dev, nullptr, stream);
    g->vcycles_done = 1;
    return rc;
  }
  return d4b_mg_solve(g, z_dev, r_dev, nullptr, nullptr, stream);
}

/* MG-preconditioned flexible CG in one call (BASELINE config 4) */
int d4b_mg_fcg_solve(d4b_mg_t *g, double *u, const double *rhs, int imax, double atol, double rtol, int *iters,
                     double *final_res, double *work_6n, void *stream)
{
  return d4b_fcg_solve(g->lvl[g->n_levels - 1], u, rhs, imax, atol, rtol, d4b_mg_pc_apply, g, iters, final_res, work_6n, stream);
}


/* Host-vector forms: what d4est_krylov_pc_t.pc_apply (Solver/d4est_krylov_pc.h:13-33) and
 * d4est_solver_fcg_solve (Solver/d4est_solver_fcg_improved.c:97) hand over are host arrays; one upload and one
 * download bracket a whole V-cycle / a whole solve. */
static int mg_hbuf(d4b_mg *g)
{
  if (g->hbuf) return D4B_OK;
  D4B_CUDA(cudaMalloc((void **)&g->hbuf, (size_t)(8 * (g->lvl[g->n_levels - 1]->local_nodes + 2)) * sizeof(double)));
  return D4B_OK;
}
int d4b_mg_pc_apply_host(d4b_mg_t *g, const double *r_host, double *z_host)
{
  int rc = mg_hbuf(g); if (rc) return rc;
  d4b_mesh *top = g->lvl[g->n_levels - 1];
  const int64_t n = top->local_nodes; const size_t nb = (size_t)n * sizeof(double);
  double *r = g->hbuf, *z = r + n;
  D4B_CUDA(cudaMemcpyAsync(r, r_host, nb, cudaMemcpyHostToDevice, top->stream));
  if ((rc = d4b_mg_pc_apply(g, r, z, top->stream)) != D4B_OK) return rc;
  D4B_CUDA(cudaMemcpyAsync(z_host, z, nb, cudaMemcpyDeviceToHost, top->stream));
  D4B_CUDA(cudaStreamSynchronize(top->stream));
  return D4B_OK;
}
int d4b_mg_fcg_solve_host(d4b_mg_t *g, double *u_host, const double *rhs_host, int imax, double atol, double rtol,
                          int *iters, double *final_res)
{
  int rc = mg_hbuf(g); if (rc) return rc;
  d4b_mesh *top = g->lvl[g->n_levels - 1];
  const int64_t n = top->local_nodes; const size_t nb = (size_t)n * sizeof(double);
  double *u = g->hbuf, *rhs = u + n, *work = rhs + n;
  D4B_CUDA(cudaMemcpyAsync(u, u_host, nb, cudaMemcpyHostToDevice, top->stream));
  D4B_CUDA(cudaMemcpyAsync(rhs, rhs_host, nb, cudaMemcpyHostToDevice, top->stream));
  if ((rc = d4b_mg_fcg_solve(g, u, rhs, imax, atol, rtol, iters, final_res, work, top->stream)) != D4B_OK) return rc;
  D4B_CUDA(cudaMemcpyAsync(u_host, u, nb, cudaMemcpyDeviceToHost, top->stream));
  D4B_CUDA(cudaStreamSynchronize(top->stream));
  return D4B_OK;
}
int d4b_mg_num_levels(const d4b_mg_t *g) { return g->n_levels; }

} /* extern C */
