/* d4b_halo.cu -- rank-boundary traffic: forward halo of u for mirror elements and the scalar
 * reductions of the Krylov drivers.  Replaces d4est_ghost_data_exchange
 * (Mesh/d4est_ghost_data.c:143-255: per-element memcpy pack + sc_MPI_Isend/Irecv/Waitall) and
 * sc_allreduce (e.g. Solver/d4est_solver_cg.c:139,165) with a pack kernel + grouped
 * ncclSend/ncclRecv and ncclAllReduce over NVLink.  One p4est rank = one GPU. */
#include "d4b_internal.h"
#include <nccl.h>
#include <cstring>

static ncclComm_t g_comm = nullptr;
static int g_rank = 0, g_nranks = 1;

#define D4B_NCCL(call)                                                                  \
  do { ncclResult_t r_ = (call);                                                        \
    if (r_ != ncclSuccess) { d4b_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call,  \
                                           ncclGetErrorString(r_)); return D4B_ENCCL; } } while (0)

__global__ void k_pack(int n_send, const int32_t *__restrict__ send_elem, const int64_t *__restrict__ node_off,
                       const int64_t *__restrict__ send_off, const double *__restrict__ u, double *__restrict__ buf)
{ /* one CTA per mirror element: contiguous copy of its nodal values */
  int j = blockIdx.x; if (j >= n_send) return;
  int e = send_elem[j];
  const double *src = u + node_off[e]; double *dst = buf + send_off[j];
  int n = (int)(node_off[e + 1] - node_off[e]);
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

extern "C" {

int d4b_comm_unique_id(char *out128)
{ ncclUniqueId id; D4B_NCCL(ncclGetUniqueId(&id)); memcpy(out128, &id, sizeof id); return D4B_OK; }

int d4b_comm_init(int rank, int nranks, const char *id128)
{
  ncclUniqueId id; memcpy(&id, id128, sizeof id);
  D4B_NCCL(ncclCommInitRank(&g_comm, nranks, id, rank));
  g_rank = rank; g_nranks = nranks;
  return D4B_OK;
}
int d4b_comm_finalize(void)
{ if (g_comm) { ncclCommDestroy(g_comm); g_comm = nullptr; g_nranks = 1; g_rank = 0; } return D4B_OK; }
int d4b_comm_size(void) { return g_nranks; }

/* forward halo: after this m->u_ghost holds the neighbours' u for every ghost element */
int d4b_halo_exchange(d4b_mesh_t *m, const double *u_dev, void *stream)
{
  cudaStream_t st = (cudaStream_t)stream;
  if (m->n_peers == 0) return D4B_OK;
  if (!g_comm) { d4b_set_error("mesh has %d peer ranks but d4b_comm_init was not called", m->n_peers); return D4B_ENCCL; }
  int ns = m->send_begin_h[m->n_peers];
  if (ns > 0) {
    k_pack<<<ns, 128, 0, st>>>(ns, m->send_elem, m->node_off, m->send_off, u_dev, m->send_buf);
    D4B_CUDA(cudaGetLastError()); d4b_count_launch(1);
  }
  D4B_NCCL(ncclGroupStart());
  for (int p = 0; p < m->n_peers; p++) {
    int64_t s0 = m->send_node_begin_h[p], s1 = m->send_node_begin_h[p + 1];
    int64_t r0 = m->recv_node_begin_h[p], r1 = m->recv_node_begin_h[p + 1];
    if (s1 > s0) D4B_NCCL(ncclSend(m->send_buf + s0, (size_t)(s1 - s0), ncclDouble, m->peer_rank_h[p], g_comm, st));
    if (r1 > r0) D4B_NCCL(ncclRecv(m->u_ghost + r0, (size_t)(r1 - r0), ncclDouble, m->peer_rank_h[p], g_comm, st));
  }
  D4B_NCCL(ncclGroupEnd());
  return D4B_OK;
}

} /* extern C */

/* sum over ranks of k device scalars, in place and in stream: the reductions of CG / FCG / cg_eigs / the V-cycle
 * residual (sc_allreduce in the reference, e.g. Solver/d4est_solver_cg.c:139,165, fcg_improved.c:245) never leave
 * the device */
int d4b_allreduce_sum_dev(double *dev, int k, cudaStream_t st)
{
  if (g_nranks == 1 || !g_comm) return D4B_OK;
  D4B_NCCL(ncclAllReduce(dev, dev, (size_t)k, ncclDouble, ncclSum, g_comm, st));
  return D4B_OK;
}
