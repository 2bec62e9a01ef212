/* d4b_brick.cu -- synthetic uniform brick generated straight into HBM (d4b_brick_create).
 * Affine map x = x0 + (x1-x0) * (i + (r+1)/2)/nx  =>  dr/dx = 2 nx/(x1-x0), J = hx hy hz / 8,
 * surface Jacobian of face dir d: product of the two tangential half-widths, n = +-e_d
 * (reference: Geometry/d4est_geometry_brick.c:121-180, Mesh/d4est_mortars_with_opt.c:191-203).
 * The factors are stored in full (per quadrature point), exactly as for a curved mesh. */
#include "d4b_internal.h"
#include "d4b_ops.h"
#include <vector>
#include <algorithm>
#include <cstdlib>

struct brick_dev {
  int nx, ny, nz, Nq, n_elem;
  int ghost_lo, ghost_hi;   /* first ghost id of the plane below / above the slab, or -1 (physical boundary) */
  double hx, hy, hz, sig;   /* element widths, sigma = prefactor p^2 / h */
  const int32_t *emap;      /* grid cell -> element id */
  const int32_t *ecoord;    /* element -> packed ix | iy<<10 | iz<<20 */
};

__global__ void brick_fill_G(brick_dev B, const double *__restrict__ wq, double *__restrict__ G)
{
  const int Nq = B.Nq, Nq3 = Nq * Nq * Nq;
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)B.n_elem * Nq3) return;
  int e = (int)(t / Nq3), q = (int)(t % Nq3);
  double w = wq[q % Nq] * wq[(q / Nq) % Nq] * wq[q / (Nq * Nq)] * (B.hx * B.hy * B.hz / 8.0);
  double *Ge = G + (int64_t)6 * Nq3 * e;
  double rx = 2.0 / B.hx, ry = 2.0 / B.hy, rz = 2.0 / B.hz;
  Ge[q] = w * rx * rx; Ge[Nq3 + q] = 0.; Ge[2 * Nq3 + q] = 0.;
  Ge[3 * Nq3 + q] = w * ry * ry; Ge[4 * Nq3 + q] = 0.; Ge[5 * Nq3 + q] = w * rz * rz;
}

__global__ void brick_fill_faces(brick_dev B, d4b_hm_t *__restrict__ hm, double *__restrict__ fac)
{
  const int n2 = B.Nq * B.Nq;
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)B.n_elem * 6) return;
  int e = (int)(t / 6), f = (int)(t % 6), d = f >> 1, s = f & 1;
  int c = B.ecoord[e], ix = c & 1023, iy = (c >> 10) & 1023, iz = (c >> 20) & 1023;
  int jx = ix + (d == 0 ? (s ? 1 : -1) : 0), jy = iy + (d == 1 ? (s ? 1 : -1) : 0),
      jz = iz + (d == 2 ? (s ? 1 : -1) : 0);
  bool bnd = jx < 0 || jy < 0 || jz < 0 || jx >= B.nx || jy >= B.ny || jz >= B.nz;
  int ghost = -1; /* neighbour in the slab of the rank below / above: ghost planes are ordered ix + nx iy */
  if (jz < 0 && B.ghost_lo >= 0) ghost = B.ghost_lo + jx + B.nx * jy;
  if (jz >= B.nz && B.ghost_hi >= 0) ghost = B.ghost_hi + jx + B.nx * jy;
  if (ghost >= 0) bnd = false;
  d4b_hm_t h;
  h.nbr = bnd ? -1 : (ghost >= 0 ? B.n_elem + ghost : B.emap[jx + B.nx * (jy + B.ny * jz)]);
  h.f_m = (int8_t)f; h.f_p = (int8_t)(bnd ? -1 : (f ^ 1)); h.tcode = 1; h.child_m = -1; h.child_p = -1;
  h.degq = (int8_t)(B.Nq - 1); h.face_slot = (int8_t)f; h.pad = 0; h.pad0 = 0;
  h.fac_off = t * 7 * n2; h.g_off = -1; h.pad2 = 0;
  hm[t] = h;
  double hd = d == 0 ? B.hx : (d == 1 ? B.hy : B.hz);
  double sj = (B.hx * B.hy * B.hz / hd) / 4.0;
  double an = sj * (s ? 1.0 : -1.0) * 2.0 / hd; /* sj n_d dr_d/dx_d */
  double *F = fac + h.fac_off;
  for (int q = 0; q < n2; q++) {
    for (int l = 0; l < 3; l++) F[l * n2 + q] = (l == d) ? an : 0.;
    if (bnd) F[3 * n2 + q] = sj * B.sig;
    else {
      for (int l = 0; l < 3; l++) F[(3 + l) * n2 + q] = (l == d) ? an : 0.;
      F[6 * n2 + q] = sj * B.sig;
    }
  }
}

/* neighbour-record table of the register-blocked kernel: records are in face order (6 e + f) and
 * the single bucket lists the elements in order, so the neighbour's record of my face f is 6 nbr + (f^1) */
__global__ void brick_fill_nrec(int n6, const d4b_hm_t *__restrict__ hm, int32_t *__restrict__ nrec)
{
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n6) return;
  int nb = hm[t].nbr;
  nrec[t] = nb < 0 ? -1 : (6 * nb >= n6 ? -2 /* ghost: no record here, fold with my own a+ */ : 6 * nb + ((t % 6) ^ 1));
}

static inline uint64_t morton3(uint32_t x, uint32_t y, uint32_t z)
{
  uint64_t k = 0;
  for (int b = 0; b < 10; b++)
    k |= ((uint64_t)((x >> b) & 1) << (3 * b)) | ((uint64_t)((y >> b) & 1) << (3 * b + 1)) |
         ((uint64_t)((z >> b) & 1) << (3 * b + 2));
  return k;
}

static int brick_create_slab(const d4b_brick_params_t *p, int rank, int nranks, d4b_mesh_t **out);
extern "C" int d4b_brick_create(const d4b_brick_params_t *p, d4b_mesh_t **out) { return brick_create_slab(p, 0, 1, out); }
/* z-slab `rank` of the brick nx x ny x (nz * nranks) over the same box: one slab per rank (= GPU), ghost
 * planes towards rank-1 / rank+1, mirror lists for d4b_halo_exchange (weak scaling: nz per rank is fixed) */
extern "C" int d4b_brick_create_slab(const d4b_brick_params_t *p, int rank, int nranks, d4b_mesh_t **out)
{
  if (nranks < 1 || rank < 0 || rank >= nranks) { d4b_set_error("d4b_brick_create_slab: bad rank"); return D4B_EINVAL; }
  return brick_create_slab(p, rank, nranks, out);
}

static int brick_create_slab(const d4b_brick_params_t *p, int rank, int nranks, d4b_mesh_t **out)
{
  *out = nullptr;
  if (!p || p->nx < 1 || p->ny < 1 || p->nz < 1 || p->nx > 1023 || p->ny > 1023 || p->nz > 1023 ||
      p->deg < 1 || p->deg > D4B_MAX_DEG || p->degq < 0 || p->degq > D4B_MAX_DEGQ) {
    d4b_set_error("d4b_brick_create: bad parameters"); return D4B_EINVAL;
  }
  const int64_t ne = (int64_t)p->nx * p->ny * p->nz;
  if (ne > 0x7fffffff / 8) { d4b_set_error("brick too large"); return D4B_EINVAL; }
  const int N = p->deg + 1, Nq = p->degq + 1;
  std::vector<std::pair<uint64_t, int32_t>> keys((size_t)ne);
  size_t k = 0;
  for (int iz = 0; iz < p->nz; iz++) for (int iy = 0; iy < p->ny; iy++) for (int ix = 0; ix < p->nx; ix++)
    keys[k++] = { morton3(ix, iy, iz), ix | (iy << 10) | (iz << 20) };
  std::sort(keys.begin(), keys.end());
  const bool has_lo = rank > 0, has_hi = rank < nranks - 1;
  const int64_t plane = (int64_t)p->nx * p->ny, ng = plane * ((has_lo ? 1 : 0) + (has_hi ? 1 : 0));
  std::vector<int32_t> ecoord((size_t)ne), emap((size_t)ne), deg((size_t)(ne + ng), p->deg), degq((size_t)ne, p->degq), hb((size_t)ne + 1);
  std::vector<int64_t> node_off((size_t)(ne + ng) + 1), quad_off((size_t)ne + 1);
  for (int64_t e = 0; e < ne; e++) {
    int c = keys[e].second; ecoord[e] = c;
    emap[(c & 1023) + p->nx * (((c >> 10) & 1023) + (int64_t)p->ny * ((c >> 20) & 1023))] = (int32_t)e;
  }
  for (int64_t e = 0; e <= ne; e++) { quad_off[e] = e * Nq * Nq * Nq; hb[e] = (int32_t)(6 * e); }
  for (int64_t e = 0; e <= ne + ng; e++) node_off[e] = e * N * N * N;
  /* halo lists: peers in rank order (below, above); mirrors = my bottom / top plane in ix + nx iy order */
  std::vector<int32_t> peer_rank, send_begin(1, 0), send_elem, recv_begin(1, 0);
  for (int side = 0; side < 2; side++) {
    if (!(side == 0 ? has_lo : has_hi)) continue;
    peer_rank.push_back(side == 0 ? rank - 1 : rank + 1);
    const int iz = side == 0 ? 0 : p->nz - 1;
    for (int iy = 0; iy < p->ny; iy++) for (int ix = 0; ix < p->nx; ix++)
      send_elem.push_back(emap[ix + p->nx * (iy + (int64_t)p->ny * iz)]);
    send_begin.push_back((int32_t)send_elem.size());
    recv_begin.push_back((int32_t)(recv_begin.back() + plane));
  }
  d4b_mesh_desc_t D; memset(&D, 0, sizeof D);
  int32_t zero1[1] = {0};
  D.n_ghost = (int32_t)ng; D.n_peers = (int32_t)peer_rank.size(); D.peer_rank = peer_rank.data();
  D.send_elem = send_elem.data();
  D.n_local = (int32_t)ne; D.quad_type = p->quad_type; D.deg = deg.data(); D.node_off = node_off.data();
  D.degq = degq.data(); D.quad_off = quad_off.data(); D.hm_begin = hb.data();
  D.n_fac = 6 * ne * 7 * Nq * Nq; D.send_begin = send_begin.data(); D.recv_begin = recv_begin.data();
  (void)zero1;
  d4b_mesh *m = nullptr;
  int rc = d4b_mesh_create(&D, &m);
  if (rc != D4B_OK) return rc;
  brick_dev B; B.nx = p->nx; B.ny = p->ny; B.nz = p->nz; B.Nq = Nq; B.n_elem = (int)ne;
  B.hx = (p->x1 - p->x0) / p->nx; B.hy = (p->y1 - p->y0) / p->ny; B.hz = (p->z1 - p->z0) / ((double)p->nz * nranks);
  B.ghost_lo = has_lo ? 0 : -1; B.ghost_hi = has_hi ? (int)(has_lo ? plane : 0) : -1;
  double h = p->face_h > 0 ? p->face_h : 1.0 / p->nx;
  B.sig = p->penalty_prefactor * (double)p->deg * (double)p->deg / h;
  int32_t *d_emap = nullptr, *d_ecoord = nullptr; double *d_w = nullptr;
  double xq[D4B_OPS_MAXN], wq[D4B_OPS_MAXN];
  d4b_ops_quad(p->degq, p->quad_type, xq, wq);
  D4B_CUDA(cudaMalloc((void **)&d_emap, ne * sizeof(int32_t)));
  D4B_CUDA(cudaMalloc((void **)&d_ecoord, ne * sizeof(int32_t)));
  D4B_CUDA(cudaMalloc((void **)&d_w, Nq * sizeof(double)));
  D4B_CUDA(cudaMemcpy(d_emap, emap.data(), ne * sizeof(int32_t), cudaMemcpyHostToDevice));
  D4B_CUDA(cudaMemcpy(d_ecoord, ecoord.data(), ne * sizeof(int32_t), cudaMemcpyHostToDevice));
  D4B_CUDA(cudaMemcpy(d_w, wq, Nq * sizeof(double), cudaMemcpyHostToDevice));
  B.emap = d_emap; B.ecoord = d_ecoord;
  int64_t nq = ne * Nq * Nq * Nq;
  brick_fill_G<<<(unsigned)((nq + 255) / 256), 256>>>(B, d_w, m->G);
  brick_fill_faces<<<(unsigned)((ne * 6 + 127) / 128), 128>>>(B, m->hm, m->fac);
  if (m->n_fast == 1 && m->fast[0].count == ne && m->fast[0].nrec)
    brick_fill_nrec<<<(unsigned)((ne * 6 + 255) / 256), 256>>>((int)(ne * 6), m->hm, m->fast[0].nrec);
  D4B_CUDA(cudaGetLastError());
  D4B_CUDA(cudaDeviceSynchronize());
  cudaFree(d_emap); cudaFree(d_ecoord); cudaFree(d_w);
  /* algorithmic bytes: u + Au + 6-component metric + 7 doubles per UNIQUE interior
   * mortar point (each appears in two half-mortars) + 4 per boundary point */
  int64_t nb = 2 * ((int64_t)p->ny * p->nz + (int64_t)p->nx * p->nz) + plane * ((has_lo ? 0 : 1) + (has_hi ? 0 : 1));
  int64_t ni = (6 * ne - nb) / 2;
  m->alg_bytes = ne * (16 * (int64_t)N * N * N + 48 * (int64_t)Nq * Nq * Nq) + (ni * 56 + nb * 32) * Nq * Nq;
  m->max_degq = p->degq;
  *out = m;
  return D4B_OK;
}
