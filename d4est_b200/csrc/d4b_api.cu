/* d4b_api.cu -- C-ABI entry points (include/d4est_b200.h): residency, apply, copies. */
#include "d4b_internal.h"
#include "d4b_ops.h"
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <atomic>
#include <algorithm>

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void d4b_set_error(const char *fmt, ...)
{ va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap); }
void d4b_count_launch(int n) { g_launches += n; }

int d4b_launch_generic(const d4b_mesh *m, const double *u, double *Au, const int32_t *elem_list,
                       int n_list, cudaStream_t st);
size_t d4b_generic_smem_bytes(int NX, int NQX);
int d4b_launch_fast(const d4b_mesh *m, const double *u, double *Au, cudaStream_t st, int *handled);
int d4b_launch_phase(const d4b_mesh *m, const double *u, double *Au, cudaStream_t st, int phase);
int d4b_hp_build(d4b_mesh *m, const d4b_mesh_desc_t *D, const std::vector<int32_t> &cand, std::vector<int32_t> &rest);

/* run-time switches of the host pipeline: defaults from the environment (read ONCE), changed by d4b_set_option */
static struct { bool init; int no_late_pass, no_pipeline, pipe_trace, pipe_chunks, no_overlap, overlap_min, tp_min; } g_opt;
static void opts_init()
{
  if (g_opt.init) return;
  g_opt.no_late_pass = getenv("D4B_NO_LATE_PASS") != nullptr; g_opt.no_pipeline = getenv("D4B_NO_PIPELINE") != nullptr;
  g_opt.pipe_trace = getenv("D4B_PIPE_TRACE") != nullptr;
  g_opt.pipe_chunks = getenv("D4B_PIPE_CHUNKS") ? atoi(getenv("D4B_PIPE_CHUNKS")) : 0;
  g_opt.no_overlap = getenv("D4B_NO_OVERLAP") != nullptr;
  g_opt.overlap_min = getenv("D4B_OVERLAP_MIN") ? atoi(getenv("D4B_OVERLAP_MIN")) : 16384;
  g_opt.tp_min = getenv("D4B_NO_TP") ? 0x7fffffff : (getenv("D4B_TP_MIN") ? atoi(getenv("D4B_TP_MIN")) : -1);
  g_opt.init = true;
}
extern int g_fcg_textbook;   /* d4b_solver.cu */
extern int g_mg_graph;       /* d4b_mg.cu */
int d4b_opt_tp_min() { opts_init(); return g_opt.tp_min; }   /* d4b_apply_hp.cu */

extern "C" {

int d4b_set_option(const char *name, int value)
{
  opts_init();
  if (!name) { d4b_set_error("option name is NULL"); return D4B_EINVAL; }
  if (!strcmp(name, "fcg_textbook_update")) g_fcg_textbook = value;
  else if (!strcmp(name, "no_late_pass")) g_opt.no_late_pass = value;
  else if (!strcmp(name, "no_pipeline")) g_opt.no_pipeline = value;
  else if (!strcmp(name, "pipe_trace")) g_opt.pipe_trace = value;
  else if (!strcmp(name, "pipe_chunks")) g_opt.pipe_chunks = value;
  else if (!strcmp(name, "no_overlap")) g_opt.no_overlap = value;
  else if (!strcmp(name, "mg_graph")) g_mg_graph = value;
  else if (!strcmp(name, "two_pass_min_elements")) g_opt.tp_min = value;         /* read when a level is created */
  else if (!strcmp(name, "overlap_min_elements")) g_opt.overlap_min = value;   /* read when a level first exchanges */
  else { d4b_set_error("unknown option %s", name); return D4B_EINVAL; }
  return D4B_OK;
}

const char *d4b_last_error(void) { return g_err; }
const char *d4b_version(void) { return "d4est_b200 0.1 (sm_100a)"; }
int64_t d4b_kernel_launches(void) { return (int64_t)g_launches.load(); }

int d4b_device_count(int *count)
{
  int n = 0; cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { n = 0; cudaGetLastError(); }
  *count = n; return D4B_OK;
}
/* One device per process (one p4est rank = one GPU): the 1-D table pools, the per-instance __constant__ tables and
 * the function attributes are cached per process, so a second device in the same process is refused. */
static int g_device = -1;
int d4b_set_device(int dev)
{
  if (g_device >= 0 && dev != g_device) {
    d4b_set_error("d4b_set_device(%d): this process is already bound to device %d (one rank = one GPU)", dev, g_device);
    return D4B_EINVAL;
  }
  D4B_CUDA(cudaSetDevice(dev));
  g_device = dev;
  return D4B_OK;
}

int d4b_dev_alloc(size_t bytes, void **p) { D4B_CUDA(cudaMalloc(p, bytes ? bytes : 8)); return D4B_OK; }
int d4b_dev_free(void *p) { D4B_CUDA(cudaFree(p)); return D4B_OK; }
int d4b_h2d(void *d, const void *s, size_t n, void *st)
{ D4B_CUDA(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, (cudaStream_t)st)); return D4B_OK; }
int d4b_d2h(void *d, const void *s, size_t n, void *st)
{ D4B_CUDA(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, (cudaStream_t)st)); return D4B_OK; }
int d4b_sync(void *st) { D4B_CUDA(cudaStreamSynchronize((cudaStream_t)st)); return D4B_OK; }
int d4b_host_alloc_pinned(size_t n, void **p) { D4B_CUDA(cudaMallocHost(p, n ? n : 8)); return D4B_OK; }
int d4b_host_free_pinned(void *p) { D4B_CUDA(cudaFreeHost(p)); return D4B_OK; }

} /* extern C */

/* ------------------------------------------------------------ 1-D operator pool */
static d4b_ops_dev g_ops[2];
static bool g_ops_ready[2] = {false, false};

const d4b_ops_dev *d4b_ops_device(int qt)
{
  if (qt < 0 || qt > 1) { d4b_set_error("bad quad_type %d", qt); return nullptr; }
  if (g_ops_ready[qt]) return &g_ops[qt];
  std::vector<double> pool;
  d4b_ops_index &ix = g_ops[qt].idx;
  double buf[D4B_OPS_MAXN * D4B_OPS_MAXN], x[D4B_OPS_MAXN], w[D4B_OPS_MAXN];
  std::vector<std::vector<double>> Dm(D4B_NDEG);
  for (int dg = 0; dg < D4B_NDEG; dg++) {
    int N = dg + 1;
    if (dg == 0) { buf[0] = 0.; } else d4b_ops_diff(dg, buf);
    Dm[dg].assign(buf, buf + N * N);
    ix.D[dg] = (int)pool.size(); pool.insert(pool.end(), buf, buf + N * N);
    ix.Dt[dg] = (int)pool.size();
    for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) pool.push_back(buf[j * N + i]);
  }
  for (int dq = 0; dq < D4B_NDEGQ; dq++) {
    if (qt == 1 && dq == 0) { x[0] = 0.; w[0] = 2.; } else d4b_ops_quad(dq, qt, x, w);
    ix.w[dq] = (int)pool.size(); pool.insert(pool.end(), w, w + dq + 1);
  }
  for (int dg = 0; dg < D4B_NDEG; dg++)
    for (int dq = 0; dq < D4B_NDEGQ; dq++)
      for (int c = 0; c < 3; c++) {
        ix.I[dg][dq][c] = (int)pool.size();
        const int N = dg + 1, Nq = dq + 1;
        if (dg == 0 || (qt == 1 && dq == 0)) { for (int i = 0; i < N * Nq; i++) buf[i] = 1.0; }
        else d4b_ops_interp(dg, dq, qt, c - 1, buf);
        pool.insert(pool.end(), buf, buf + N * Nq);
        ix.ID[dg][dq][c] = (int)pool.size();   /* (I D)[q][j] = sum_k I[q][k] D[k][j] */
        for (int q = 0; q < Nq; q++)
          for (int j = 0; j < N; j++) {
            double sacc = 0.;
            for (int k = 0; k < N; k++) sacc += buf[q * N + k] * Dm[dg][k * N + j];
            pool.push_back(sacc);
          }
      }
  double *d = nullptr;
  if (cudaMalloc(&d, pool.size() * sizeof(double)) != cudaSuccess ||
      cudaMemcpy(d, pool.data(), pool.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) {
    d4b_set_error("ops pool upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  g_ops[qt].pool = d; g_ops_ready[qt] = true;
  return &g_ops[qt];
}

int d4b_mesh_make_streams(d4b_mesh *m)
{
  for (int k = 0; k < D4B_NSIDE; k++) {
    D4B_CUDA(cudaStreamCreateWithFlags(&m->side[k], cudaStreamNonBlocking));
    D4B_CUDA(cudaEventCreateWithFlags(&m->ev_join[k], cudaEventDisableTiming));
  }
  D4B_CUDA(cudaEventCreateWithFlags(&m->ev_fork, cudaEventDisableTiming));
  return D4B_OK;
}

template <class T> static int upload(T **dst, const T *src, size_t n, int64_t *bytes)
{
  *dst = nullptr;
  D4B_CUDA(cudaMalloc((void **)dst, (n ? n : 1) * sizeof(T)));
  if (n && src) D4B_CUDA(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice)); /* src NULL: filled on device */
  *bytes += (int64_t)(n * sizeof(T));
  return D4B_OK;
}

extern "C" {

int d4b_mesh_create(const d4b_mesh_desc_t *D, d4b_mesh_t **out)
{
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    d4b_set_error("no CUDA device: d4est_b200 has no CPU fallback"); return D4B_ENODEV;
  }
  if (!D || D->n_local < 0 || D->n_ghost < 0) { d4b_set_error("bad descriptor"); return D4B_EINVAL; }
  const int nl = D->n_local, ne = nl + D->n_ghost;
  d4b_mesh *m = (d4b_mesh *)calloc(1, sizeof(d4b_mesh));
  m->n_local = nl; m->n_ghost = D->n_ghost; m->quad_type = D->quad_type;
  m->local_nodes = D->node_off[nl]; m->ghost_nodes = D->node_off[ne] - D->node_off[nl];
  m->total_quad = D->quad_off[nl]; m->n_hm = D->hm_begin[nl]; m->n_fac = D->n_fac; m->n_bc_g = D->n_bc_g;
  int64_t alg = 0;
  for (int e = 0; e < ne; e++) {
    if (D->deg[e] < 1 || D->deg[e] > D4B_MAX_DEG) { d4b_set_error("deg[%d]=%d out of range", e, D->deg[e]); free(m); return D4B_EINVAL; }
    if (D->deg[e] > m->max_deg_ext) m->max_deg_ext = D->deg[e];
    int64_t N = D->deg[e] + 1;
    if (D->node_off[e + 1] - D->node_off[e] != N * N * N) { d4b_set_error("node_off inconsistent at %d", e); free(m); return D4B_EINVAL; }
  }
  for (int e = 0; e < nl; e++) {
    if (D->degq[e] < 0 || D->degq[e] > D4B_MAX_DEGQ) { d4b_set_error("degq[%d] out of range", e); free(m); return D4B_EINVAL; }
    if (D->deg[e] > m->max_deg) m->max_deg = D->deg[e];
    if (D->degq[e] > m->max_degq) m->max_degq = D->degq[e];
    if (D->degq[e] > m->max_degq_own) m->max_degq_own = D->degq[e];
    int64_t N = D->deg[e] + 1, Nq = D->degq[e] + 1;
    alg += 16 * N * N * N + 48 * Nq * Nq * Nq;
  }
  for (int64_t r = 0; D->hm && r < m->n_hm; r++) {
    const d4b_hm_t &h = D->hm[r];
    if (h.degq > D4B_MAX_DEGQ || h.nbr >= ne || h.f_m < 0 || h.f_m > 5) { d4b_set_error("bad half-mortar %lld", (long long)r); free(m); return D4B_EINVAL; }
    if (h.degq > m->max_degq) m->max_degq = h.degq;
    int64_t n2 = (int64_t)(h.degq + 1) * (h.degq + 1);
    alg += (h.nbr >= 0) ? 28 * n2 : 32 * n2; /* 7 doubles per unique mortar point, 4 on the boundary */
  }
  m->alg_bytes = alg;
  m->generic_smem = d4b_generic_smem_bytes(m->max_deg_ext + 1, m->max_degq + 1);
  if (m->generic_smem > 227 * 1024) {
    d4b_set_error("degrees too high for the generic kernel: needs %zu B shared memory", m->generic_smem);
    free(m); return D4B_EINVAL;
  }
  int rc;
#define UP(field, src, n) if ((rc = upload(&m->field, src, (size_t)(n), &m->bytes)) != D4B_OK) { d4b_mesh_destroy(m); return rc; }
  UP(deg, D->deg, ne) UP(node_off, D->node_off, ne + 1) UP(degq, D->degq, nl) UP(quad_off, D->quad_off, nl + 1)
  UP(G, D->G, 6 * m->total_quad) UP(hm_begin, D->hm_begin, nl + 1) UP(hm, D->hm, m->n_hm)
  UP(fac, D->fac, m->n_fac) UP(bc_g, D->bc_g, m->n_bc_g)
#undef UP
#define D4B_CUDA_M(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { d4b_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); d4b_mesh_destroy(m); return D4B_ECUDA; } } while (0)
  D4B_CUDA_M(cudaMalloc((void **)&m->u_ghost, (m->ghost_nodes ? m->ghost_nodes : 1) * sizeof(double)));
  D4B_CUDA_M(cudaMemset(m->u_ghost, 0, (m->ghost_nodes ? m->ghost_nodes : 1) * sizeof(double)));
  D4B_CUDA_M(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
  /* halo lists */
  m->n_peers = D->n_peers;
  if (D->n_peers > 0) {
    int np = D->n_peers;
    m->peer_rank_h = (int32_t *)malloc(np * sizeof(int32_t)); memcpy(m->peer_rank_h, D->peer_rank, np * sizeof(int32_t));
    m->send_begin_h = (int32_t *)malloc((np + 1) * sizeof(int32_t)); memcpy(m->send_begin_h, D->send_begin, (np + 1) * sizeof(int32_t));
    m->recv_begin_h = (int32_t *)malloc((np + 1) * sizeof(int32_t)); memcpy(m->recv_begin_h, D->recv_begin, (np + 1) * sizeof(int32_t));
    m->send_node_begin_h = (int64_t *)calloc(np + 1, sizeof(int64_t));
    m->recv_node_begin_h = (int64_t *)calloc(np + 1, sizeof(int64_t));
    int64_t sn = 0;
    std::vector<int64_t> soff((size_t)D->send_begin[np] + 1);
    for (int p = 0; p < np; p++) {
      m->send_node_begin_h[p] = sn;
      for (int j = D->send_begin[p]; j < D->send_begin[p + 1]; j++) {
        int e = D->send_elem[j]; soff[j] = sn; sn += D->node_off[e + 1] - D->node_off[e];
      }
      m->recv_node_begin_h[p] = D->node_off[nl + D->recv_begin[p]] - D->node_off[nl];
    }
    m->send_node_begin_h[np] = sn;
    m->recv_node_begin_h[np] = D->node_off[nl + D->recv_begin[np]] - D->node_off[nl];
    m->send_nodes = sn;
    int64_t dummy = 0;
    if ((rc = upload(&m->send_elem, D->send_elem, (size_t)D->send_begin[np], &dummy)) != D4B_OK) { d4b_mesh_destroy(m); return rc; }
    if ((rc = upload(&m->send_off, soff.data(), (size_t)D->send_begin[np], &dummy)) != D4B_OK) { d4b_mesh_destroy(m); return rc; }
    D4B_CUDA_M(cudaMalloc((void **)&m->send_buf, (sn ? sn : 1) * sizeof(double)));
  }
  if ((rc = d4b_fast_classify(m, D)) != D4B_OK) { d4b_mesh_destroy(m); return rc; }
  if ((rc = d4b_mesh_make_streams(m)) != D4B_OK) { d4b_mesh_destroy(m); return rc; }
  *out = m;
  return D4B_OK;
}

void d4b_mesh_destroy(d4b_mesh_t *m)
{
  if (!m) return;
  cudaFree(m->deg); cudaFree(m->node_off); cudaFree(m->degq); cudaFree(m->quad_off); cudaFree(m->G);
  cudaFree(m->hm_begin); cudaFree(m->hm); cudaFree(m->fac); cudaFree(m->bc_g); cudaFree(m->u_ghost); cudaFree(m->mass_coef);
  cudaFree(m->u_stage); cudaFree(m->Au_stage); cudaFree(m->host_work); cudaFree(m->sc); cudaFree(m->dot_part); if (m->sc_h) cudaFreeHost(m->sc_h); cudaFree(m->send_elem); cudaFree(m->send_buf); cudaFree(m->send_off);
  for (int i = 0; i < m->n_fast; i++) { cudaFree(m->fast[i].elem_list); cudaFree(m->fast[i].frec); cudaFree(m->fast[i].nrec); cudaFree(m->fast[i].v3_slots); }
  for (int i = 0; i < m->n_hp; i++) { cudaFree(m->hp[i].slots); cudaFree(m->hp[i].recs); }
  for (int i = 0; i < m->n_tp1; i++) { cudaFree(m->tp1[i].slots); cudaFree(m->tp1[i].recs); }
  for (int i = 0; i < m->n_tp2; i++) { cudaFree(m->tp2[i].slots); cudaFree(m->tp2[i].recs); }
  cudaFree(m->tpW);
  cudaFree(m->generic_list);
  free(m->peer_rank_h); free(m->send_begin_h); free(m->recv_begin_h);
  free(m->send_node_begin_h); free(m->recv_node_begin_h);
  for (int k = 0; k < D4B_NSIDE; k++) { if (m->side[k]) cudaStreamDestroy(m->side[k]); if (m->ev_join[k]) cudaEventDestroy(m->ev_join[k]); }
  if (m->ev_fork) cudaEventDestroy(m->ev_fork);
  if (m->split_built)
    for (int i = 0; i < m->n_fast; i++) {
      d4b_fast_bucket *two[2] = {&m->fast_early[i], &m->fast_late[i]};
      for (d4b_fast_bucket *b : two) { cudaFree(b->elem_list); cudaFree(b->frec); cudaFree(b->nrec); cudaFree(b->v3_slots); }
    }
  if (m->comm_stream) { cudaStreamDestroy(m->comm_stream); cudaEventDestroy(m->ev_cfork); cudaEventDestroy(m->ev_halo); }
  if (m->stream) cudaStreamDestroy(m->stream);
  if (m->copy_stream) { cudaStreamDestroy(m->copy_stream); cudaStreamDestroy(m->in_stream); for (int k = 0; k < D4B_PIPE_MAX; k++) { cudaEventDestroy(m->ev_out[k]); cudaEventDestroy(m->ev_in[k]); } cudaEventDestroy(m->ev_done);
    cudaFree(m->pipe_late.elem_list); cudaFree(m->pipe_late.frec); cudaFree(m->pipe_late.nrec); cudaFree(m->pipe_late.v3_slots); }
  free(m);
}

int64_t d4b_mesh_local_nodes(const d4b_mesh_t *m) { return m->local_nodes; }
int64_t d4b_mesh_ghost_nodes(const d4b_mesh_t *m) { return m->ghost_nodes; }
int64_t d4b_mesh_bytes(const d4b_mesh_t *m) { return m->bytes; }
int64_t d4b_mesh_algorithmic_bytes(const d4b_mesh_t *m) { return m->alg_bytes; }

int d4b_mesh_kernel_split(const d4b_mesh_t *m, int64_t counts[3])
{
  counts[0] = counts[1] = 0; counts[2] = m->n_generic;
  for (int i = 0; i < m->n_fast; i++) counts[0] += m->fast[i].count;
  for (int i = 0; i < m->n_hp; i++) counts[1] += m->hp[i].count;
  for (int i = 0; i < m->n_tp2; i++) counts[1] += m->tp2[i].count;   /* two-pass hp elements */
  return D4B_OK;
}

int d4b_set_ghost_host(d4b_mesh_t *m, const double *ug)
{
  if (m->ghost_nodes) D4B_CUDA(cudaMemcpy(m->u_ghost, ug, m->ghost_nodes * sizeof(double), cudaMemcpyHostToDevice));
  m->ghost_supplied = 1;
  return D4B_OK;
}

/* The reference always exchanges ghosts inside the apply (dGMath/d4est_laplacian_with_opt.c:327).  A level with
 * neighbour ranks therefore needs the NCCL communicator; without it the apply fails loudly instead of using a
 * stale halo -- unless the caller supplied the ghost values itself (d4b_set_ghost_host). */
static int d4b_halo_for_apply(d4b_mesh_t *m, const double *u, void *stream)
{
  if (m->n_peers == 0) return D4B_OK;
  if (d4b_comm_size() > 1) return d4b_halo_exchange(m, u, stream);
  if (m->ghost_supplied) return D4B_OK;
  d4b_set_error("level has %d neighbour ranks but neither d4b_comm_init nor d4b_set_ghost_host was called: the halo of u is undefined", m->n_peers);
  return D4B_ENCCL;
}

/* Split every regular bucket of a level with neighbour ranks into ghost-free batches and the rest (once). */
static int d4b_build_overlap_split(d4b_mesh_t *m, cudaStream_t st)
{
  int rc, any = 0;
  for (int i = 0; i < m->n_fast; i++) {
    const d4b_fast_bucket &b = m->fast[i];
    const int group = (b.nrec && d4b_v3_supported(b.N, b.Nq)) ? 32 / b.Nq : 1;
    if ((rc = d4b_v3_split_ghost(m, b, group, &m->fast_early[i], &m->fast_late[i], st)) != D4B_OK) return rc;
    any += m->fast_early[i].count;
  }
  if (!m->comm_stream) {
    /* highest priority: the block scheduler then places the ghost-dependent batches as soon as the halo is there,
     * ahead of the not yet dispatched CTAs of the ghost-free launch (at default priority they ran as a 47 us tail
     * after it, measured) */
    int prio_lo = 0, prio_hi = 0;
    D4B_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    D4B_CUDA(cudaStreamCreateWithPriority(&m->comm_stream, cudaStreamNonBlocking, prio_hi));
    D4B_CUDA(cudaEventCreateWithFlags(&m->ev_cfork, cudaEventDisableTiming));
    D4B_CUDA(cudaEventCreateWithFlags(&m->ev_halo, cudaEventDisableTiming));
  }
  /* below ~16k ghost-free elements (option overlap_min_elements) an apply is launch-bound (MG coarse levels, hp-dominated levels): the extra
   * launches of the split cost what the hidden exchange saves (MG-FCG on 2 GPUs: 6.53 s with, 6.50 s without) */
  m->split_built = (any > 0 && any >= g_opt.overlap_min) ? 1 : -1;
  return D4B_OK;
}

/* The reference exchanges the ghost data and then runs the element loop (dGMath/d4est_laplacian_with_opt.c:327).
 * Here the exchange (pack kernel + ncclSend/ncclRecv) runs on the level's communication stream while the batches
 * that touch no ghost element are computed on the caller's stream; the ghost-dependent rest of the level follows
 * the exchange in the communication stream.  Both are ordered after everything already queued on the caller's
 * stream, and the caller's stream ends up ordered after both, so the call keeps its in-stream semantics.
 * d4b_set_option("no_overlap", 1) (or D4B_NO_OVERLAP) restores exchange-then-apply. */
int d4b_apply_dev(d4b_mesh_t *m, const double *u, double *Au, void *stream)
{
  cudaStream_t st = (cudaStream_t)stream;
  int handled = 0, rc;
  opts_init();
  if (m->n_peers > 0 && d4b_comm_size() > 1 && !g_opt.no_overlap && m->split_built >= 0) {
    if (!m->split_built && (rc = d4b_build_overlap_split(m, st)) != D4B_OK) return rc;
    if (m->split_built == 1) {
      static const bool trace = getenv("D4B_OVERLAP_TRACE") != nullptr;
      cudaEvent_t tv[4];
      D4B_CUDA(cudaEventRecord(m->ev_cfork, st));
      D4B_CUDA(cudaStreamWaitEvent(m->comm_stream, m->ev_cfork, 0));
      if (trace) { for (int i = 0; i < 4; i++) cudaEventCreate(&tv[i]); cudaEventRecord(tv[0], m->comm_stream); }
      if ((rc = d4b_halo_exchange(m, u, m->comm_stream)) != D4B_OK) return rc;
      if (trace) cudaEventRecord(tv[1], m->comm_stream);
      if ((rc = d4b_launch_phase(m, u, Au, st, 1)) != D4B_OK) return rc;
      if (trace) cudaEventRecord(tv[2], st);
      if ((rc = d4b_launch_phase(m, u, Au, m->comm_stream, 2)) != D4B_OK) return rc;
      if (trace) cudaEventRecord(tv[3], m->comm_stream);
      D4B_CUDA(cudaEventRecord(m->ev_halo, m->comm_stream));
      D4B_CUDA(cudaStreamWaitEvent(st, m->ev_halo, 0));
      if (trace) {
        cudaEventSynchronize(tv[2]); cudaEventSynchronize(tv[3]);
        float a, b, c;
        cudaEventElapsedTime(&a, tv[0], tv[1]); cudaEventElapsedTime(&b, tv[0], tv[2]); cudaEventElapsedTime(&c, tv[0], tv[3]);
        int ne = 0, nl = 0;
        for (int i = 0; i < m->n_fast; i++) { ne += m->fast_early[i].count; nl += m->fast_late[i].count; }
        fprintf(stderr, "[d4b overlap] halo done %.3f ms, ghost-free batches (%d elements) done %.3f, ghost batches (%d) done %.3f\n", a, ne, b, nl, c);
        for (int i = 0; i < 4; i++) cudaEventDestroy(tv[i]);
      }
      return D4B_OK;
    }
  }
  if ((rc = d4b_halo_for_apply(m, u, stream)) != D4B_OK) return rc;
  rc = d4b_launch_fast(m, u, Au, st, &handled);
  if (rc != D4B_OK) return rc;
  return rc;
}

int d4b_apply_host(d4b_mesh_t *m, const double *u_host, double *Au_host)
{
  size_t nb = (size_t)m->local_nodes * sizeof(double);
  if (!m->u_stage) {
    D4B_CUDA(cudaMalloc((void **)&m->u_stage, nb ? nb : 8));
    D4B_CUDA(cudaMalloc((void **)&m->Au_stage, nb ? nb : 8));
  }
  /* One degree bucket that covers the whole level in element order (uniform mesh, no rank boundary) and a
   * range-capable kernel: the apply runs as a 3-stream pipeline over equal element ranges.  u arrives in chunks;
   * the kernel of range k starts when the last chunk it reads (own elements + face neighbours, computed once per
   * level) has arrived; the device->host copy of a finished range overlaps the following kernels and, PCIe
   * being full duplex, the remaining uploads.  Every element writes only its own nodes of Au. */
  /* can the kernel store straight into the caller's buffer (pinned, device-mapped host memory)? */
  double *Au_map = nullptr;
  opts_init();
  const bool env_no_late = g_opt.no_late_pass, env_no_pipe = g_opt.no_pipeline, env_trace = g_opt.pipe_trace;
  const int env_chunks = g_opt.pipe_chunks;
  if (!env_no_late) {
    cudaPointerAttributes pa;
    if (cudaPointerGetAttributes(&pa, Au_host) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer)
      Au_map = static_cast<double *>(pa.devicePointer);
    else
      cudaGetLastError();
  }
  /* with neighbour ranks the pipeline needs the late launch: the elements next to a ghost can only be computed
   * after the halo exchange, which follows the last upload */
  if (m->n_fast == 1 && m->n_generic == 0 && m->fast[0].count == m->n_local && m->fast[0].nrec &&
      (m->n_peers == 0 || Au_map) &&
      d4b_v3_supported(m->fast[0].N, m->fast[0].Nq) && m->n_local >= 8192 && !env_no_pipe) {
    const d4b_fast_bucket &b = m->fast[0];
    int NCH = 16;
    if (env_chunks >= 2 && env_chunks <= D4B_PIPE_MAX) NCH = env_chunks;
    const int64_t N3 = (int64_t)b.N * b.N * b.N;
    int rc0;
    if (!m->copy_stream) {
      D4B_CUDA(cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking));
      D4B_CUDA(cudaStreamCreateWithFlags(&m->in_stream, cudaStreamNonBlocking));
      for (int k = 0; k < D4B_PIPE_MAX; k++) {
        D4B_CUDA(cudaEventCreateWithFlags(&m->ev_out[k], cudaEventDisableTiming));
        D4B_CUDA(cudaEventCreateWithFlags(&m->ev_in[k], cudaEventDisableTiming));
      }
      D4B_CUDA(cudaEventCreateWithFlags(&m->ev_done, cudaEventDisableTiming));
      /* range boundaries: equal ranges, multiples of the elements per CTA (measured on the bench brick: 16
       * ranges beat 8, 24 and 32, and ranges tapering towards the end gained nothing -- the download stream
       * runs at the duplex PCIe rate and simply ends one to two chunks after the upload) */
      const int per = d4b_v3_elems_per_cta(b.N, b.Nq);
      int nr0 = 0;
      m->pipe_b[0] = 0;
      for (int k = 0; k < NCH; k++) {
        int e = k == NCH - 1 ? m->n_local : (int)((int64_t)(k + 1) * m->n_local / NCH / per * per);
        if (e > m->pipe_b[nr0]) m->pipe_b[++nr0] = e;
      }
      m->pipe_ranges = nr0;
      if ((rc0 = d4b_v3_range_need(m, b, m->pipe_ranges, m->pipe_b, m->pipe_need, m->stream)) != D4B_OK) return rc0;
      if ((rc0 = d4b_v3_build_late(m, b, m->pipe_ranges, m->pipe_b, 1, &m->pipe_late, m->stream)) != D4B_OK) return rc0;
    }
    const int nr = m->pipe_ranges;
    /* Elements that read a neighbour more than one chunk ahead (seams of the space-filling curve, a few per
     * cent) would hold their whole range back until most of u has arrived.  If the kernel can store straight
     * into the caller's buffer (pinned, device-mapped host memory), they are recomputed in a last launch after
     * the final upload and the final download, and every range runs as soon as the NEXT chunk is there (what
     * the early launch computes for them from partial data is overwritten).  Otherwise a range waits for the
     * last chunk it reads. */
    if (m->pipe_late.count == 0) Au_map = nullptr;
    const bool defer = Au_map != nullptr || m->pipe_late.count == 0;
    const bool trace = env_trace;
    cudaEvent_t tv[5];
    if (trace) { for (int i = 0; i < 5; i++) cudaEventCreate(&tv[i]); cudaEventRecord(tv[0], m->in_stream); }
    for (int k = 0; k < nr; k++) {
      int s0 = m->pipe_b[k], s1 = m->pipe_b[k + 1];
      D4B_CUDA(cudaMemcpyAsync(m->u_stage + s0 * N3, u_host + s0 * N3, (size_t)(s1 - s0) * N3 * sizeof(double),
                               cudaMemcpyHostToDevice, m->in_stream));
      D4B_CUDA(cudaEventRecord(m->ev_in[k], m->in_stream));
    }
    if (trace) cudaEventRecord(tv[1], m->in_stream);
    int waited = -1;
    for (int k = 0; k < nr; k++) {
      int s0 = m->pipe_b[k], s1 = m->pipe_b[k + 1];
      int need = defer ? k + 1 : m->pipe_need[k];
      need = need < k ? k : (need >= nr ? nr - 1 : need);
      if (need > waited) { D4B_CUDA(cudaStreamWaitEvent(m->stream, m->ev_in[need], 0)); waited = need; }
      if ((rc0 = d4b_launch_v3_range(m, b, m->u_stage, m->Au_stage, m->stream, s0, s1)) != D4B_OK) return rc0;
      D4B_CUDA(cudaEventRecord(m->ev_out[k], m->stream));
      D4B_CUDA(cudaStreamWaitEvent(m->copy_stream, m->ev_out[k], 0));
      D4B_CUDA(cudaMemcpyAsync(Au_host + s0 * N3, m->Au_stage + s0 * N3, (size_t)(s1 - s0) * N3 * sizeof(double),
                               cudaMemcpyDeviceToHost, m->copy_stream));
    }
    if (trace) { cudaEventRecord(tv[2], m->stream); cudaEventRecord(tv[3], m->copy_stream); }
    const bool dist = m->n_peers > 0 && d4b_comm_size() > 1;
    if (m->n_peers > 0 && !dist && !m->ghost_supplied) {
      d4b_set_error("level has %d neighbour ranks but neither d4b_comm_init nor d4b_set_ghost_host was called", m->n_peers);
      return D4B_ENCCL;
    }
    if (Au_map || dist) {
      if (waited < nr - 1) D4B_CUDA(cudaStreamWaitEvent(m->stream, m->ev_in[nr - 1], 0));
      if (dist && (rc0 = d4b_halo_exchange(m, m->u_stage, m->stream)) != D4B_OK) return rc0;
      D4B_CUDA(cudaEventRecord(m->ev_done, m->copy_stream));
      D4B_CUDA(cudaStreamWaitEvent(m->stream, m->ev_done, 0));
      if (Au_map && (rc0 = d4b_launch_v3(m, m->pipe_late, m->u_stage, Au_map, m->stream)) != D4B_OK) return rc0;
    }
    D4B_CUDA(cudaStreamSynchronize(m->copy_stream));
    D4B_CUDA(cudaStreamSynchronize(m->stream));
    D4B_CUDA(cudaStreamSynchronize(m->in_stream));
    if (trace) {
      cudaEventRecord(tv[4], m->stream); cudaEventSynchronize(tv[4]);
      float a, b2, c, d;
      cudaEventElapsedTime(&a, tv[0], tv[1]); cudaEventElapsedTime(&b2, tv[0], tv[2]);
      cudaEventElapsedTime(&c, tv[0], tv[3]); cudaEventElapsedTime(&d, tv[0], tv[4]);
      fprintf(stderr, "[d4b pipe] h2d done %.2f ms, early kernels done %.2f, d2h done %.2f, all done %.2f (late elements %d)\n",
              a, b2, c, d, m->pipe_late.count);
      for (int i = 0; i < 5; i++) cudaEventDestroy(tv[i]);
    }
    return D4B_OK;
  }
  D4B_CUDA(cudaMemcpyAsync(m->u_stage, u_host, nb, cudaMemcpyHostToDevice, m->stream));
  int rc = d4b_apply_dev(m, m->u_stage, m->Au_stage, m->stream);
  if (rc != D4B_OK) return rc;
  D4B_CUDA(cudaMemcpyAsync(Au_host, m->Au_stage, nb, cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return D4B_OK;
}

} /* extern C */

/* Split the local elements into degree buckets of "regular" elements (every face conforming with
 * an equal-degree neighbour, or a boundary) and the rest (generic kernel). */
int d4b_fast_classify(d4b_mesh *m, const d4b_mesh_desc_t *D)
{
  const int nl = m->n_local;
  std::vector<std::vector<int32_t>> el(D4B_NDEG * D4B_NDEGQ), fr(D4B_NDEG * D4B_NDEGQ), nr(D4B_NDEG * D4B_NDEGQ);
  /* (element, face) -> its single conforming record, for the neighbour-record table of v3 */
  std::vector<int32_t> rec_of;
  if (D->hm) {
    rec_of.assign((size_t)nl * 6, -1);
    for (int e = 0; e < nl; e++)
      for (int r = D->hm_begin[e]; r < D->hm_begin[e + 1]; r++) {
        const d4b_hm_t &h = D->hm[r];
        if (h.f_m < 0 || h.f_m > 5) continue;
        int32_t &slot = rec_of[(size_t)e * 6 + h.f_m];
        slot = (slot == -1 && h.child_m < 0 && h.child_p < 0) ? r : -3; /* -3: hanging / several records */
      }
  }
  std::vector<int32_t> gen, cand;
  const bool off = getenv("D4B_NO_FAST") != nullptr || getenv("D4B_ALL_HP") != nullptr;
  const bool no_hp = getenv("D4B_NO_HP") != nullptr || (getenv("D4B_NO_FAST") != nullptr && !getenv("D4B_ALL_HP"));
  for (int e = 0; e < nl; e++) {
    int dg = D->deg[e], dq = D->degq[e];
    bool ok = !off && d4b_fast_supported(dg + 1, dq + 1);
    int32_t rec[6] = {-1, -1, -1, -1, -1, -1};
    if (ok && D->hm) {
      if (D->hm_begin[e + 1] - D->hm_begin[e] != 6) ok = false;
      for (int r = D->hm_begin[e]; ok && r < D->hm_begin[e + 1]; r++) {
        const d4b_hm_t &h = D->hm[r];
        if (h.child_m >= 0 || h.child_p >= 0 || h.degq != dq || rec[h.f_m] >= 0 || (h.nbr < 0 && h.pad == 1)) { ok = false; break; }
        if (h.nbr >= 0 && D->deg[h.nbr] != dg) { ok = false; break; }
        rec[h.f_m] = r;
      }
    } else if (ok) {
      for (int f = 0; f < 6; f++) rec[f] = 6 * e + f; /* device-built brick: records in face order */
    }
    if (!ok) { (no_hp ? gen : cand).push_back(e); continue; }
    int k = dg * D4B_NDEGQ + dq;
    int32_t nrc[6];
    for (int f = 0; f < 6; f++) {
      if (!D->hm) { nrc[f] = -2; continue; } /* device-built brick: filled by d4b_brick_create */
      const d4b_hm_t &h = D->hm[rec[f]];
      nrc[f] = h.nbr < 0 ? -1 : -2;
      if (h.nbr >= 0 && h.nbr < nl && h.f_p >= 0 && h.f_p < 6) {
        int r2 = rec_of[(size_t)h.nbr * 6 + h.f_p];
        if (r2 >= 0 && D->hm[r2].nbr == e && D->hm[r2].degq == h.degq && D->hm[r2].f_p == h.f_m) nrc[f] = r2;
      }
    }
    el[k].push_back(e); fr[k].insert(fr[k].end(), rec, rec + 6); nr[k].insert(nr[k].end(), nrc, nrc + 6);
  }
  m->n_fast = 0;
  int64_t dummy = 0; int rc;
  for (int k = 0; k < D4B_NDEG * D4B_NDEGQ; k++) {
    if (el[k].empty()) continue;
    if (m->n_fast == D4B_MAX_BUCKETS) { d4b_set_error("more than %d (deg, deg_quad) pairs on one level", D4B_MAX_BUCKETS); return D4B_EINVAL; }
    d4b_fast_bucket &b = m->fast[m->n_fast++];
    b.N = k / D4B_NDEGQ + 1; b.Nq = k % D4B_NDEGQ + 1; b.count = (int)el[k].size();
    if ((rc = upload(&b.elem_list, el[k].data(), el[k].size(), &dummy)) != D4B_OK) return rc;
    if ((rc = upload(&b.frec, fr[k].data(), fr[k].size(), &dummy)) != D4B_OK) return rc;
    b.nrec = nullptr; b.v3_slots = nullptr;
    if ((rc = upload(&b.nrec, nr[k].data(), nr[k].size(), &dummy)) != D4B_OK) return rc;
  }
  if ((rc = d4b_hp_build(m, D, cand, gen)) != D4B_OK) return rc;
  m->n_generic = (int)gen.size();
  std::sort(gen.begin(), gen.end());
  if ((rc = upload(&m->generic_list, gen.data(), gen.size(), &dummy)) != D4B_OK) return rc;
  return D4B_OK;
}
