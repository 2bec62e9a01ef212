/* d4b_apply_generic.cu -- element-centric SIPG apply for ARBITRARY (deg, deg_quad,
 * neighbour degree, hanging, orientation) combinations.  One CTA per element, all
 * sizes runtime.  This is the completeness kernel: every element that has no
 * templated fast kernel (d4b_apply_fast.cu) goes through here.  Both write disjoint
 * elements of Au, so there are no atomics anywhere on the path.
 *
 * Math (reference dGMath/d4est_laplacian_with_opt.c:289-368):
 *   Au_e = sum_l D_l^T W_l + A_face,
 *   W_l  = I^T [ sum_l' G_ll' (I D_l' u) ]  +  Lift_f( E^T w H_l )      (term 2)
 *   A_face = Lift_f( E^T w (term1 + term3) )
 * with, per mortar quadrature point (…_flux_sipg.c:645-677):
 *   term1 = -1/2 (q- + q+),  q = sum_l a_l du/dr_l,   H_l = -1/2 a-_l [u],  term3 = sj sigma [u].
 */
#include "d4b_internal.h"

__constant__ d4b_ops_dev c_ops[2];

struct d4b_gen_args {
  const int32_t *deg; const int64_t *node_off; const int32_t *degq; const int64_t *quad_off;
  const double *G; const int32_t *hm_begin; const d4b_hm_t *hm; const double *fac; const double *bc_g;
  const double *u, *u_ghost; double *Au;
  const int32_t *elem_list; int n_list; int n_local; int64_t local_nodes;
  int qt, NX, NQX, MX;
};

/* idx / d for 0 <= idx < 2^15, 1 <= d <= 64, with inv = (2^20 + d - 1) / d computed once per stage */
__device__ __forceinline__ int fdiv_small(int idx, unsigned inv) { return (int)(((unsigned long long)(unsigned)idx * inv) >> 20); }

template <int COLS>
__device__ __forceinline__ double dot_fixed(const double *__restrict__ Mr, int cs, const double *inb, int stride)
{
  double m[COLS], x[COLS];
#pragma unroll
  for (int k = 0; k < COLS; k++) { m[k] = __ldg(&Mr[k * cs]); x[k] = inb[k * stride]; }   /* all loads in flight */
  double s = 0.;
#pragma unroll
  for (int k = 0; k < COLS; k++) s = fma(m[k], x[k], s);
  return s;
}

/* out = (M applied along axis ax) in;  M(r,k) = M[r*rs + k*cs]; in dims (n0,n1,n2), n[ax]=cols */
__device__ __forceinline__ void contract(const double *__restrict__ M, int rs, int cs, int rows, int cols,
                                         const double *in, double *out, int n0, int n1, int n2, int ax,
                                         bool accumulate)
{
  const int o0 = ax == 0 ? rows : n0, o1 = ax == 1 ? rows : n1, o2 = ax == 2 ? rows : n2;
  const int total = o0 * o1 * o2;
  const int stride = ax == 0 ? 1 : (ax == 1 ? n0 : n0 * n1);
  const unsigned inv0 = ((1u << 20) + o0 - 1) / o0, inv1 = ((1u << 20) + o1 - 1) / o1;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int q0 = fdiv_small(idx, inv0), i0 = idx - q0 * o0;
    const int i2 = fdiv_small(q0, inv1), i1 = q0 - i2 * o1;
    int r = ax == 0 ? i0 : (ax == 1 ? i1 : i2);
    int base = (ax == 0 ? 0 : i0) + n0 * ((ax == 1 ? 0 : i1) + n1 * (ax == 2 ? 0 : i2));
    const double *Mr = M + r * rs, *inb = in + base;
    double s;
    switch (cols) { /* 1-D sizes are 2..16: fixed trip counts keep every load of the dot product in flight */
      case 2: s = dot_fixed<2>(Mr, cs, inb, stride); break;
      case 3: s = dot_fixed<3>(Mr, cs, inb, stride); break;
      case 4: s = dot_fixed<4>(Mr, cs, inb, stride); break;
      case 5: s = dot_fixed<5>(Mr, cs, inb, stride); break;
      case 6: s = dot_fixed<6>(Mr, cs, inb, stride); break;
      case 7: s = dot_fixed<7>(Mr, cs, inb, stride); break;
      case 8: s = dot_fixed<8>(Mr, cs, inb, stride); break;
      default:
        s = 0.;
        for (int k = 0; k < cols; k++) s = fma(__ldg(&Mr[k * cs]), inb[k * stride], s);
    }
    out[idx] = accumulate ? out[idx] + s : s;
  }
  __syncthreads();
}

__device__ __forceinline__ int face_vol_index(int N, int f, int a, int b)
{ /* Appendix A.2: face-local axes = remaining axes in increasing order, lower fastest */
  int layer = (f & 1) ? N - 1 : 0, dir = f >> 1;
  return dir == 0 ? layer + N * (a + N * b) : (dir == 1 ? a + N * (layer + N * b) : a + N * (b + N * layer));
}

/* out[0][a,b] = u on face f; out[1+l][a,b] = du/dr_l on face f  (slicer of D_l u,
 * reference dGMath/d4est_operators.c:1521 + :1385) */
__device__ __forceinline__ void face_traces(const double *uel, int N, int f, const double *__restrict__ D,
                                            double *out)
{
  const int N2 = N * N;
  for (int t = threadIdx.x; t < 4 * N2; t += blockDim.x) {
    int c = t / N2, ab = t % N2, a = ab % N, b = ab / N;
    int v = face_vol_index(N, f, a, b);
    if (c == 0) { out[ab] = uel[v]; continue; }
    int l = c - 1, st = l == 0 ? 1 : (l == 1 ? N : N2);
    int il = (v / st) % N, base = v - il * st;
    double s = 0.;
    for (int k = 0; k < N; k++) s = fma(__ldg(&D[il * N + k]), uel[base + k * st], s);
    out[c * N2 + ab] = s;
  }
  __syncthreads();
}

__device__ __forceinline__ int tcode_src(int tcode, int n, int a, int b)
{ /* m-order point -> p-order point (d4b_hm_t.tcode) */
  int x = (tcode & 1) ? a : b, y = (tcode & 1) ? b : a;
  if (tcode & 2) x = n - 1 - x;
  if (tcode & 4) y = n - 1 - y;
  return x + n * y;
}

extern "C" __global__ void __launch_bounds__(256)
d4b_apply_generic_kernel(const d4b_gen_args A)
{
  extern __shared__ double sm[];
  const int e = A.elem_list ? A.elem_list[blockIdx.x] : (int)blockIdx.x;
  if (e >= A.n_local) return;
  const d4b_ops_dev &ops = c_ops[A.qt];
  const double *pool = ops.pool;
  const int dg = A.deg[e], dq = A.degq[e];
  const int N = dg + 1, Nq = dq + 1, N2 = N * N, N3 = N2 * N, Nq3 = Nq * Nq * Nq;
  const int NX = A.NX, NQX = A.NQX;
  const int NX3 = NX * NX * NX, NQX3 = NQX * NQX * NQX, MX3 = A.MX * A.MX * A.MX;
  double *su = sm, *W = su + NX3, *acc = W + 3 * NX3, *f = acc + NX3;
  double *t1 = f + 3 * NQX3, *t2 = t1 + MX3, *fs = t2 + MX3; /* fs: face scratch */
  const double *D = pool + ops.idx.D[dg];
  const double *I = pool + ops.idx.I[dg][dq][0];
  const double *ue = A.u + A.node_off[e];

  { /* L2 prefetch of what this element reads once, later: its metric block and the factor records of its
     * half-mortars (8 threads per record) */
    const char *g = reinterpret_cast<const char *>(A.G + 6 * A.quad_off[e]);
    for (int o = threadIdx.x * 128; o < 6 * Nq3 * 8; o += blockDim.x * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + o));
    const int r1 = A.hm_begin[e + 1];
    for (int r = A.hm_begin[e] + (threadIdx.x >> 3); r < r1; r += blockDim.x >> 3) {
      const d4b_hm_t *h = A.hm + r;
      const int n2 = (h->degq + 1) * (h->degq + 1);
      const char *p = reinterpret_cast<const char *>(A.fac + h->fac_off);
      const int bytes = (h->nbr >= 0 ? 7 : 4) * n2 * 8;
      for (int o = (threadIdx.x & 7) * 128; o < bytes; o += 8 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
    }
  }
  for (int i = threadIdx.x; i < N3; i += blockDim.x) { su[i] = ue[i]; acc[i] = 0.; }
  __syncthreads();

  /* ---- volume: f_l = I (D_l u) at quadrature points (Quadrature/d4est_quadrature.c:339-375) */
  for (int l = 0; l < 3; l++) {
    contract(D, N, 1, N, N, su, t2, N, N, N, l, false);
    contract(I, N, 1, Nq, N, t2, t1, N, N, N, 0, false);
    contract(I, N, 1, Nq, N, t1, t2, Nq, N, N, 1, false);
    contract(I, N, 1, Nq, N, t2, f + l * NQX3, Nq, Nq, N, 2, false);
  }
  {
    const double *Ge = A.G + 6 * A.quad_off[e];
    for (int q = threadIdx.x; q < Nq3; q += blockDim.x) {
      double g0 = f[q], g1 = f[NQX3 + q], g2 = f[2 * NQX3 + q];
      double G00 = Ge[q], G01 = Ge[Nq3 + q], G02 = Ge[2 * Nq3 + q];
      double G11 = Ge[3 * Nq3 + q], G12 = Ge[4 * Nq3 + q], G22 = Ge[5 * Nq3 + q];
      f[q] = G00 * g0 + G01 * g1 + G02 * g2;
      f[NQX3 + q] = G01 * g0 + G11 * g1 + G12 * g2;
      f[2 * NQX3 + q] = G02 * g0 + G12 * g1 + G22 * g2;
    }
    __syncthreads();
  }
  for (int l = 0; l < 3; l++) { /* W_l = I^T f_l */
    contract(I, 1, N, N, Nq, f + l * NQX3, t1, Nq, Nq, Nq, 2, false);
    contract(I, 1, N, N, Nq, t1, t2, Nq, Nq, N, 1, false);
    contract(I, 1, N, N, Nq, t2, W + l * NX3, Nq, N, N, 0, false);
  }

  /* ---- faces: one half-mortar at a time */
  const int FX = NX * NX, QX = NQX * NQX, TX = NQX * (NX > NQX ? NX : NQX);
  double *ftm = fs, *ftp = ftm + 4 * FX, *mq = ftp + 4 * FX, *pq = mq + 4 * QX, *FH = pq + 4 * QX,
         *tmp = FH + 4 * QX;
  (void)TX;
  for (int r = A.hm_begin[e]; r < A.hm_begin[e + 1]; r++) {
    const d4b_hm_t h = A.hm[r];
    const int fm = h.f_m, mdq = h.degq, Nm = mdq + 1, Nm2 = Nm * Nm;
    const double *Ea = pool + ops.idx.I[dg][mdq][h.child_m < 0 ? 0 : 1 + (h.child_m & 1)];
    const double *Eb = pool + ops.idx.I[dg][mdq][h.child_m < 0 ? 0 : 1 + ((h.child_m >> 1) & 1)];
    const double *wq = pool + ops.idx.w[mdq];
    const double *F = A.fac + h.fac_off;
    face_traces(su, N, fm, D, ftm);
    contract(Ea, N, 1, Nm, N, ftm, tmp, N, N, 4, 0, false);
    contract(Eb, N, 1, Nm, N, tmp, mq, Nm, N, 4, 1, false);
    if (h.nbr >= 0) {
      const int dp = A.deg[h.nbr], Np = dp + 1;
      const double *un = (h.nbr < A.n_local) ? A.u + A.node_off[h.nbr]
                                             : A.u_ghost + (A.node_off[h.nbr] - A.local_nodes);
      const double *Dp = pool + ops.idx.D[dp];
      const double *Pa = pool + ops.idx.I[dp][mdq][h.child_p < 0 ? 0 : 1 + (h.child_p & 1)];
      const double *Pb = pool + ops.idx.I[dp][mdq][h.child_p < 0 ? 0 : 1 + ((h.child_p >> 1) & 1)];
      face_traces(un, Np, h.f_p, Dp, ftp);
      contract(Pa, Np, 1, Nm, Np, ftp, tmp, Np, Np, 4, 0, false);
      contract(Pb, Np, 1, Nm, Np, tmp, pq, Nm, Np, 4, 1, false);
    }
    for (int q = threadIdx.x; q < Nm2; q += blockDim.x) {
      int a = q % Nm, b = q / Nm;
      double w = wq[a] * wq[b];
      double am0 = F[q], am1 = F[Nm2 + q], am2 = F[2 * Nm2 + q];
      double um = mq[q];
      double qm = am0 * mq[Nm2 + q] + am1 * mq[2 * Nm2 + q] + am2 * mq[3 * Nm2 + q];
      double jump, t1v, sjs, hs;
      if (h.nbr >= 0) {
        int s = tcode_src(h.tcode, Nm, a, b);
        double qp = F[3 * Nm2 + q] * pq[Nm2 + s] + F[4 * Nm2 + q] * pq[2 * Nm2 + s] +
                    F[5 * Nm2 + q] * pq[3 * Nm2 + s];
        jump = um - pq[s]; t1v = -0.5 * (qm + qp); sjs = F[6 * Nm2 + q]; hs = -0.5;
      } else {
        double g = (h.g_off >= 0) ? A.bc_g[h.g_off + q] : 0.;
        jump = um - g; t1v = -qm; sjs = F[3 * Nm2 + q]; hs = -1.0;
        if (h.pad == 1) { /* Robin: sj (alpha u - beta), fac row 3 = sj alpha, bc_g = sj beta, rows 0-2 zero */
          FH[q] = w * (sjs * um - g); FH[Nm2 + q] = 0.; FH[2 * Nm2 + q] = 0.; FH[3 * Nm2 + q] = 0.;
          continue;
        }
      }
      FH[q] = w * (t1v + sjs * jump);
      FH[Nm2 + q] = w * hs * am0 * jump;
      FH[2 * Nm2 + q] = w * hs * am1 * jump;
      FH[3 * Nm2 + q] = w * hs * am2 * jump;
    }
    __syncthreads();
    /* Galerkin integral + projection onto my face space: E^T (…_flux_sipg.c:682-730, 833-910) */
    contract(Eb, 1, N, N, Nm, FH, tmp, Nm, Nm, 4, 1, false);
    contract(Ea, 1, N, N, Nm, tmp, ftm, Nm, N, 4, 0, false);
    for (int t = threadIdx.x; t < N2; t += blockDim.x) { /* lift (…_flux_sipg.c:918-958) */
      int v = face_vol_index(N, fm, t % N, t / N);
      acc[v] += ftm[t];
      W[v] += ftm[N2 + t];
      W[NX3 + v] += ftm[2 * N2 + t];
      W[2 * NX3 + v] += ftm[3 * N2 + t];
    }
    __syncthreads();
  }

  /* ---- Au = acc + sum_l D_l^T W_l */
  for (int l = 0; l < 3; l++) contract(D, 1, N, N, N, W + l * NX3, acc, N, N, N, l, true);
  double *Aue = A.Au + A.node_off[e];
  for (int i = threadIdx.x; i < N3; i += blockDim.x) Aue[i] = acc[i];
}

/* ---- mass-type term of apply_jac / build_residual: out (+)= I^T diag(w (x) w (x) w * c) I v per element, c given at the
 * volume quadrature points (c = J f(x, u0): d4est_quadrature_apply_fofufofvlilj, Quadrature/d4est_quadrature.c:593-760,
 * which ends in d4est_quadrature_apply_mass_matrix; with f = 1 the plain mass matrix of
 * d4est_laplacian_build_rhs_with_strong_bc).  Element-local, no faces: one CTA per element, run-time sizes. */
struct d4b_mass_args {
  const int32_t *deg; const int64_t *node_off; const int32_t *degq; const int64_t *quad_off;
  const double *coef, *v; double *out; int n_local, qt, NX, NQX, accumulate;
};
extern "C" __global__ void __launch_bounds__(128) d4b_mass_kernel(const d4b_mass_args A)
{
  extern __shared__ double sm[];
  const int e = blockIdx.x;
  if (e >= A.n_local) return;
  const d4b_ops_dev &ops = c_ops[A.qt];
  const double *pool = ops.pool;
  const int dg = A.deg[e], dq = A.degq[e], N = dg + 1, Nq = dq + 1, N3 = N * N * N, Nq3 = Nq * Nq * Nq;
  const int MX3 = (A.NX > A.NQX ? A.NX : A.NQX); 
  double *t1 = sm, *t2 = sm + MX3 * MX3 * MX3;
  const double *I = pool + ops.idx.I[dg][dq][0], *wq = pool + ops.idx.w[dq];
  const double *ve = A.v + A.node_off[e], *ce = A.coef + A.quad_off[e];
  for (int i = threadIdx.x; i < N3; i += blockDim.x) t1[i] = ve[i];
  __syncthreads();
  contract(I, N, 1, Nq, N, t1, t2, N, N, N, 0, false);
  contract(I, N, 1, Nq, N, t2, t1, Nq, N, N, 1, false);
  contract(I, N, 1, Nq, N, t1, t2, Nq, Nq, N, 2, false);
  for (int q = threadIdx.x; q < Nq3; q += blockDim.x) {
    const int a = q % Nq, b = (q / Nq) % Nq, c = q / (Nq * Nq);
    t2[q] *= wq[a] * wq[b] * wq[c] * ce[q];
  }
  __syncthreads();
  contract(I, 1, N, N, Nq, t2, t1, Nq, Nq, Nq, 2, false);
  contract(I, 1, N, N, Nq, t1, t2, Nq, Nq, N, 1, false);
  contract(I, 1, N, N, Nq, t2, t1, Nq, N, N, 0, false);
  double *oe = A.out + A.node_off[e];
  for (int i = threadIdx.x; i < N3; i += blockDim.x) oe[i] = A.accumulate ? oe[i] + t1[i] : t1[i];
}

size_t d4b_generic_smem_bytes(int NX, int NQX)
{
  int MX = NX > NQX ? NX : NQX;
  size_t n = (size_t)5 * NX * NX * NX + (size_t)3 * NQX * NQX * NQX + (size_t)2 * MX * MX * MX;
  n += (size_t)4 * (2 * NX * NX + 3 * NQX * NQX + NQX * MX);
  return n * sizeof(double);
}

static int generic_bind_ops(int qt);

int d4b_launch_mass(const d4b_mesh *m, const double *coef, const double *v, double *out, int accumulate, cudaStream_t st)
{
  int rc = generic_bind_ops(m->quad_type); if (rc) return rc;
  if (m->n_local == 0) return D4B_OK;
  d4b_mass_args A;
  A.deg = m->deg; A.node_off = m->node_off; A.degq = m->degq; A.quad_off = m->quad_off; A.coef = coef; A.v = v; A.out = out;
  A.n_local = m->n_local; A.qt = m->quad_type; A.NX = m->max_deg + 1; A.NQX = m->max_degq_own + 1; A.accumulate = accumulate;
  const int MX = A.NX > A.NQX ? A.NX : A.NQX;
  const size_t smem = 2 * (size_t)MX * MX * MX * sizeof(double);
  static size_t attr = 0;
  if (smem > 48 * 1024 && smem > attr) {
    D4B_CUDA(cudaFuncSetAttribute(d4b_mass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = smem;
  }
  d4b_mass_kernel<<<m->n_local, 128, smem, st>>>(A);
  D4B_CUDA(cudaGetLastError());
  d4b_count_launch(1);
  return D4B_OK;
}

static int generic_bind_ops(int qt)
{
  static bool bound[2] = {false, false};
  if (bound[qt]) return D4B_OK;
  const d4b_ops_dev *ops = d4b_ops_device(qt);
  if (!ops) return D4B_ECUDA;
  D4B_CUDA(cudaMemcpyToSymbol(c_ops, ops, sizeof(d4b_ops_dev), qt * sizeof(d4b_ops_dev), cudaMemcpyHostToDevice));
  D4B_CUDA(cudaDeviceSynchronize());
  bound[qt] = true;
  return D4B_OK;
}

int d4b_launch_generic(const d4b_mesh *m, const double *u, double *Au, const int32_t *elem_list,
                       int n_list, cudaStream_t st)
{
  static bool attr_set = false;
  if (!attr_set) {
    D4B_CUDA(cudaFuncSetAttribute(d4b_apply_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  227 * 1024));
    attr_set = true;
  }
  { int rcb = generic_bind_ops(m->quad_type); if (rcb) return rcb; }
  d4b_gen_args A;
  A.deg = m->deg; A.node_off = m->node_off; A.degq = m->degq; A.quad_off = m->quad_off; A.G = m->G;
  A.hm_begin = m->hm_begin; A.hm = m->hm; A.fac = m->fac; A.bc_g = m->bc_g;
  A.u = u; A.u_ghost = m->u_ghost; A.Au = Au; A.elem_list = elem_list;
  A.n_list = elem_list ? n_list : m->n_local; A.n_local = m->n_local; A.local_nodes = m->local_nodes;
  A.qt = m->quad_type; A.NX = m->max_deg_ext + 1; A.NQX = m->max_degq + 1;
  A.MX = A.NX > A.NQX ? A.NX : A.NQX;
  if (A.n_list == 0) return D4B_OK;
  d4b_apply_generic_kernel<<<A.n_list, 128, m->generic_smem, st>>>(A);
  D4B_CUDA(cudaGetLastError());
  d4b_count_launch(1);
  return D4B_OK;
}
