/* d4b_internal.h -- device-side data model shared by the .cu files (not part of the ABI). */
#ifndef D4B_INTERNAL_H
#define D4B_INTERNAL_H
#include "../../include/d4est_b200.h"
#include <cuda_runtime.h>

#define D4B_NDEG 12   /* deg 0..11  */
#define D4B_NDEGQ 16  /* degq 0..15 */

/* offsets (in doubles) into the 1-D operator pool of one quadrature type */
struct d4b_ops_index {
  int D[D4B_NDEG];                 /* N x N,   D[i*N+j] = l_j'(x_i)                      */
  int Dt[D4B_NDEG];                /* transpose                                          */
  int w[D4B_NDEGQ];                /* Nq quadrature weights                              */
  int I[D4B_NDEG][D4B_NDEGQ][3];   /* Nq x N interpolation GLL->quad; [0] full face,
                                      [1] child 0 (lower half), [2] child 1 (upper half) */
  int ID[D4B_NDEG][D4B_NDEGQ][3];  /* I D: reference derivative evaluated at the same points     */
};

struct d4b_ops_dev {
  d4b_ops_index idx;
  const double *pool;              /* device pointer */
};

/* one degree bucket of "regular" elements handled by a templated kernel (d4b_apply_fast.cu) */
struct d4b_fast_bucket {
  int N, Nq, count;
  int32_t *elem_list;   /* device: element ids                                          */
  int32_t *frec;        /* device: [count][6] half-mortar record index of face f        */
  void *v3_slots;       /* device: resolved per-slot metadata of d4b_apply_v3.cu (lazy)        */
  int32_t *nrec;        /* device: [count][6] record of the NEIGHBOUR for the same face (its a- is my
                           -a+), -1 boundary, -2 none (ghost / not mirrored: use my own a+)   */
};
#define D4B_MAX_BUCKETS 48
/* one (N, Nq) bucket of the templated hanging / mixed-p kernel (d4b_apply_hp.cu) */
struct d4b_hp_bucket {
  int N, Nq, count, n_rec, NMX, MX;   /* NMX: largest mortar Nm, MX: largest of N and the neighbours' N */
  void *slots, *recs;                 /* device: hp_slot[count], hp_rec[n_rec]                          */
};

/* one (N, Nq) bucket of the two-pass (twin-trace) hp kernels (d4b_apply_hp.cu): pass 1 or pass 2 */
struct d4b_tp_bucket {
  int N, Nq, count, n_rec, NMX;
  void *slots, *recs;                 /* device: tp_slot[count], tp_rec[n_rec] */
};

/* device-resident mesh level */
struct d4b_mesh {
  int n_local, n_ghost, quad_type, max_deg, max_degq, max_deg_ext, max_degq_own;
  int64_t local_nodes, ghost_nodes, total_quad, n_hm, n_fac, n_bc_g;
  int32_t *deg; int64_t *node_off; int32_t *degq; int64_t *quad_off;
  double *G; int32_t *hm_begin; d4b_hm_t *hm; double *fac; double *bc_g;
  double *u_ghost;                 /* [ghost_nodes] received halo                        */
  double *mass_coef;               /* [total_quad] coefficient of the mass-type term (d4b_mass_set_coef), or NULL */
  int ghost_supplied;              /* d4b_set_ghost_host was called (single-process tests of the N > 1 layout) */
  double *u_stage, *Au_stage;      /* staging for d4b_apply_host                         */
  double *sc, *dot_part, *sc_h;    /* device scalar block, reduction partials, pinned host mirror (d4b_solver.cu) */
  double *host_work;               /* 6 n doubles: staging + work of the host-vector solver entry points */
  int64_t bytes, alg_bytes;
  size_t generic_smem;
  /* halo lists */
  int n_peers; int32_t *peer_rank_h, *send_begin_h, *recv_begin_h; int32_t *send_elem;
  int64_t *send_node_begin_h, *recv_node_begin_h; double *send_buf; int64_t send_nodes; int64_t *send_off; /* device: offset of each send slot in send_buf */
  #define D4B_PIPE_MAX 32
  cudaStream_t stream, copy_stream, in_stream; cudaEvent_t ev_out[D4B_PIPE_MAX], ev_in[D4B_PIPE_MAX], ev_done;   /* d4b_apply_host pipeline */
  int pipe_ranges, pipe_b[D4B_PIPE_MAX + 1], pipe_need[D4B_PIPE_MAX];   /* range k = slots [pipe_b[k], pipe_b[k+1]) */
  d4b_fast_bucket pipe_late;       /* elements with a face neighbour more than one chunk ahead: computed after the last upload */
  d4b_fast_bucket fast[D4B_MAX_BUCKETS]; int n_fast;
  d4b_hp_bucket hp[D4B_MAX_BUCKETS]; int n_hp;
  #define D4B_NSIDE 6
  cudaStream_t side[D4B_NSIDE]; cudaEvent_t ev_fork, ev_join[D4B_NSIDE];   /* concurrent bucket launches of one apply */
  int32_t *generic_list; int n_generic;   /* elements left to the generic kernel */
  /* two-pass hp path: trace buckets (pass 1), flux buckets (pass 2), trace buffer W (2 Nm^2 doubles per record) */
  d4b_tp_bucket tp1[D4B_MAX_BUCKETS], tp2[D4B_MAX_BUCKETS]; int n_tp1, n_tp2; double *tpW; int64_t tpW_doubles;
  /* halo / compute overlap of a distributed apply (d4b_apply_dev): every regular bucket is split once into the
   * batches that touch no ghost element (launched while the exchange runs on comm_stream) and the rest */
  int split_built;                 /* 0 not yet, 1 built, -1 nothing to overlap (no ghost-free regular batch) */
  d4b_fast_bucket fast_early[D4B_MAX_BUCKETS], fast_late[D4B_MAX_BUCKETS];
  cudaStream_t comm_stream; cudaEvent_t ev_cfork, ev_halo;
};

/* Per-slot metadata, resolved once per bucket (v3_build_slots): one coalesced read per element
 * instead of the frec -> hm -> node_off pointer chase. */
struct v3_face {
  int64_t fac_off;   /* my record: a-[3][Q2] at +0, a+[3][Q2] at +3 Q2, sj*sigma at +6 Q2 (interior) / +3 Q2 (boundary) */
  int64_t g_off;     /* Dirichlet data offset or -1                                               */
  int64_t nbr_u;     /* >= 0: offset into u;  < 0: -(offset into u_ghost) - 1                      */
  int64_t spare;
  int32_t code;      /* bits 0-2 tcode, 3-5 f_p, 6-7 mode: 0 boundary, else interior face;
                        bit 8 INTERNAL: the neighbour sits in MY batch, faces me with f^1 and identity
                        orientation, so its trace comes by warp shuffle from the threads that own it;
                        bits 9-13: its index in the batch;
                        bits 14-18 (external faces): staging slot that holds the neighbour trace     */
  int32_t pad;
};
struct v3_slot { int64_t node_off, quad_off; v3_face f[6]; };   /* 256 bytes */

int d4b_v3_prepare_slots(const d4b_mesh *m, const d4b_fast_bucket &b, cudaStream_t st); /* any (N,Nq) */

void d4b_set_error(const char *fmt, ...);
const d4b_ops_dev *d4b_ops_device(int quad_type); /* lazily built + uploaded */
void d4b_count_launch(int n);
bool d4b_fast_supported(int N, int Nq);
bool d4b_hp_supported(int N, int Nq);
int d4b_launch_hp_bucket(const d4b_mesh *m, int i, const double *u, double *Au, cudaStream_t st);
int d4b_launch_tp_bucket(const d4b_mesh *m, int pass, int i, const double *u, double *Au, cudaStream_t st);
bool d4b_v3_supported(int N, int Nq);
int d4b_launch_v3(const d4b_mesh *m, const d4b_fast_bucket &b, const double *u, double *Au, cudaStream_t st);
int d4b_launch_v3_range(const d4b_mesh *m, const d4b_fast_bucket &b, const double *u, double *Au, cudaStream_t st,
                        int slot_begin, int slot_end);
int d4b_v3_elems_per_cta(int N, int Nq);
int d4b_v3_build_late(const d4b_mesh *m, const d4b_fast_bucket &b, int n_ranges, const int *bounds, int lookahead, d4b_fast_bucket *late, cudaStream_t st);
int d4b_v3_split_ghost(const d4b_mesh *m, const d4b_fast_bucket &b, int group, d4b_fast_bucket *early, d4b_fast_bucket *late, cudaStream_t st);
int d4b_v3_range_need(const d4b_mesh *m, const d4b_fast_bucket &b, int n_ranges, const int *bounds, int *need_host, cudaStream_t st);
/* split elements into fast buckets + generic list; D->hm == NULL means the uniform device-built brick */
int d4b_fast_classify(d4b_mesh *m, const d4b_mesh_desc_t *D);
/* sync-free solver pieces (d4b_solver.cu) used by the multigrid */
int d4b_cg_eigs_dev(d4b_mesh *m, double *u, const double *rhs, int imax, int use_new, double *work_3n, cudaStream_t st);
int d4b_cheby_dev(d4b_mesh *m, double *u, const double *rhs, double *r, int iter, const double *eig_src, double mult,
                  double *eig_dst, double ratio, double *work_2n, cudaStream_t st);
double *d4b_scalar_tmax(d4b_mesh *m);
int d4b_norm2_dev(d4b_mesh *m, const double *r, cudaStream_t st);
int d4b_read_r2(d4b_mesh *m, double *r2, cudaStream_t st);
int d4b_allreduce_sum_dev(double *dev, int k, cudaStream_t st);
int d4b_mesh_make_streams(d4b_mesh *m);
int d4b_launch_mass(const d4b_mesh *m, const double *coef, const double *v, double *out, int accumulate, cudaStream_t st);

#define D4B_CUDA(call)                                                                   \
  do { cudaError_t e_ = (call);                                                          \
    if (e_ != cudaSuccess) { d4b_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call,   \
                                           cudaGetErrorString(e_)); return D4B_ECUDA; } } while (0)
#endif
