/* d4b_apply_fast.cu -- templated element-centric SIPG apply for "regular" elements:
 * every face is either a physical boundary or a conforming interface (no hanging side) whose
 * neighbour has the same (deg, deg_quad) -- any inter-tree orientation.  One template instance
 * per (N, Nq); elements are batched by degree bucket (lists built in d4b_fast_classify).  All
 * other elements go through d4b_apply_generic.cu.  Same math as the generic kernel (see its
 * header); restructured as sum-factorised LINE tasks: a thread pulls one line of N (Nq) values
 * from shared memory into registers, applies one or two 1-D matrices whose entries are
 * compile-time-indexed __constant__ operands of the DFMAs, and stores the result line.
 *
 * 1-D matrices:  B = I (GLL->quad interpolation, Nq x N),  C = I D (derivative evaluated at
 * quad points),  D (N x N) only for the face-normal derivative / normal lift.
 */
#include "d4b_internal.h"
#include "d4b_ops.h"
#include <vector>
#include <cstdlib>
#include <cstring>

#define FAST_MAXN 8
#define FAST_MAXQ 10
struct fast_tab { double B[FAST_MAXQ * FAST_MAXN], C[FAST_MAXQ * FAST_MAXN], D[FAST_MAXN * FAST_MAXN], w[FAST_MAXQ]; };
__constant__ fast_tab c_ft;

struct fast_args {
  const int64_t *node_off; const int64_t *quad_off; const double *G; const int32_t *hm_begin;
  const d4b_hm_t *hm; const double *fac; const double *bc_g;
  const double *u, *u_ghost; double *Au; const int32_t *elem_list; int n_list; int n_local;
  int64_t local_nodes;
};

/* out_k[o] = sum_i M_k[o*NI+i] in[i]   (TR: M_k[i*NO+o], i.e. transposed use of an NI x NO table) */
template <int NI, int NO, bool TR>
__device__ __forceinline__ void mat1(const double *M, const double (&x)[NI], double (&y)[NO])
{
#pragma unroll
  for (int o = 0; o < NO; o++) {
    double s = 0.;
#pragma unroll
    for (int i = 0; i < NI; i++) s = fma(TR ? M[i * NO + o] : M[o * NI + i], x[i], s);
    y[o] = s;
  }
}
template <int NI, int NO, bool TR>
__device__ __forceinline__ void mat1_acc(const double *M, const double (&x)[NI], double (&y)[NO])
{
#pragma unroll
  for (int o = 0; o < NO; o++) {
    double s = y[o];
#pragma unroll
    for (int i = 0; i < NI; i++) s = fma(TR ? M[i * NO + o] : M[o * NI + i], x[i], s);
    y[o] = s;
  }
}
template <int n> __device__ __forceinline__ void ld(const double *p, int st, double (&x)[n])
{
#pragma unroll
  for (int i = 0; i < n; i++) x[i] = p[i * st];
}
template <int n> __device__ __forceinline__ void st(double *p, int s, const double (&x)[n])
{
#pragma unroll
  for (int i = 0; i < n; i++) p[i * s] = x[i];
}

template <int N, int Nq> struct FL {
  static constexpr int M = N > Nq ? N : Nq;
  static constexpr int N2 = N * N, N3 = N * N * N, Q2 = Nq * Nq, Q3 = Nq * Nq * Nq, NQ = N * Nq;
  static constexpr int VB = M * M * M;
  static constexpr int TR = 4 * 2 * N2, S1 = 4 * 3 * NQ, QV = 4 * 4 * Q2, FH = 2 * 4 * Q2, R1 = 2 * 3 * NQ, R2 = 2 * 2 * N2;
  static constexpr int RA = TR > QV ? TR : QV, RB = S1 > FH ? S1 : FH;
  static constexpr int FACE = RA + RB + R1 + R2;
  static constexpr int SCR = 6 * VB > FACE ? 6 * VB : FACE;
  static constexpr int PER = 2 * N3 + SCR;          /* doubles of shared memory per element */
  static constexpr int TPE = N2;                     /* threads per element */
  static constexpr int EPB = (128 / TPE) > 0 ? (128 / TPE) : 1; /* elements per CTA */
  static constexpr int THREADS = TPE * EPB;
};

__device__ __forceinline__ int f_tsrc(int tcode, int n, int a, int b)
{
  int x = (tcode & 1) ? a : b, y = (tcode & 1) ? b : a;
  if (tcode & 2) x = n - 1 - x;
  if (tcode & 4) y = n - 1 - y;
  return x + n * y;
}

template <int N, int Nq>
__global__ void __launch_bounds__(FL<N, Nq>::THREADS)
d4b_apply_fast_kernel(const fast_args A, const int32_t *__restrict__ frec)
{
  using L = FL<N, Nq>;
  constexpr int N2 = L::N2, N3 = L::N3, Q2 = L::Q2, Q3 = L::Q3, NQ = L::NQ, VB = L::VB, TPE = L::TPE;
  extern __shared__ double sm[];
  const int le = threadIdx.x / TPE, lt = threadIdx.x % TPE;
  const int slot = blockIdx.x * L::EPB + le;
  const bool active = slot < A.n_list;
  const int e = active ? A.elem_list[slot] : 0;
  double *U = sm + le * L::PER, *ACC = U + N3, *S = ACC + N3;
  double *V0 = S, *V1 = S + VB, *V2 = S + 2 * VB, *V3 = S + 3 * VB, *V4 = S + 4 * VB, *V5 = S + 5 * VB;
  const double *B = c_ft.B, *C = c_ft.C, *D = c_ft.D;
  const double *ue = A.u + A.node_off[e];

  if (active)
#pragma unroll
    for (int m = 0; m < N3 / TPE; m++) U[lt + TPE * m] = ue[lt + TPE * m];
  __syncthreads();

  /* ---------------- volume (Quadrature/d4est_quadrature.c:263-375, sum-factorised) */
  if (active) { /* along t: lines (i,j) */
    for (int t = lt; t < N2; t += TPE) {
      double x[N], y[Nq];
      ld<N>(U + t, N2, x);
      mat1<N, Nq, false>(B, x, y); st<Nq>(V0 + t, N2, y);
      mat1<N, Nq, false>(C, x, y); st<Nq>(V1 + t, N2, y);
    }
  }
  __syncthreads();
  if (active) { /* along s: lines (i,c) */
    for (int t = lt; t < NQ; t += TPE) {
      int i = t % N, c = t / N;
      double x[N], y[Nq];
      ld<N>(V0 + i + N2 * c, N, x);
      mat1<N, Nq, false>(B, x, y); st<Nq>(V2 + i + NQ * c, N, y);
      mat1<N, Nq, false>(C, x, y); st<Nq>(V3 + i + NQ * c, N, y);
      ld<N>(V1 + i + N2 * c, N, x);
      mat1<N, Nq, false>(B, x, y); st<Nq>(V4 + i + NQ * c, N, y);
    }
  }
  __syncthreads();
  if (active) { /* along r: lines (b,c); metric; transposed r-contraction, all in registers */
    const double *Ge = A.G + 6 * A.quad_off[e];
    for (int t = lt; t < Q2; t += TPE) {
      double x[N], gr[Nq], gs[Nq], gt[Nq];
      ld<N>(V2 + N * t, 1, x); mat1<N, Nq, false>(C, x, gr);
      ld<N>(V3 + N * t, 1, x); mat1<N, Nq, false>(B, x, gs);
      ld<N>(V4 + N * t, 1, x); mat1<N, Nq, false>(B, x, gt);
      const double *Gq = Ge + Nq * t;
#pragma unroll
      for (int a = 0; a < Nq; a++) {
        double G00 = Gq[a], G01 = Gq[Q3 + a], G02 = Gq[2 * Q3 + a], G11 = Gq[3 * Q3 + a],
               G12 = Gq[4 * Q3 + a], G22 = Gq[5 * Q3 + a];
        double g0 = gr[a], g1 = gs[a], g2 = gt[a];
        gr[a] = G00 * g0 + G01 * g1 + G02 * g2;
        gs[a] = G01 * g0 + G11 * g1 + G12 * g2;
        gt[a] = G02 * g0 + G12 * g1 + G22 * g2;
      }
      mat1<Nq, N, true>(C, gr, x); st<N>(V0 + N * t, 1, x);
      mat1<Nq, N, true>(B, gs, x); st<N>(V1 + N * t, 1, x);
      mat1<Nq, N, true>(B, gt, x); st<N>(V5 + N * t, 1, x);
    }
  }
  __syncthreads();
  if (active) { /* transposed s-contraction: lines (i,c) */
    for (int t = lt; t < NQ; t += TPE) {
      int i = t % N, c = t / N;
      double x[Nq], y[N];
      ld<Nq>(V0 + i + NQ * c, N, x); mat1<Nq, N, true>(B, x, y);
      ld<Nq>(V1 + i + NQ * c, N, x); mat1_acc<Nq, N, true>(C, x, y);
      st<N>(V2 + i + N2 * c, N, y);
      ld<Nq>(V5 + i + NQ * c, N, x); mat1<Nq, N, true>(B, x, y);
      st<N>(V3 + i + N2 * c, N, y);
    }
  }
  __syncthreads();
  if (active) { /* transposed t-contraction: lines (i,j) -> ACC */
    for (int t = lt; t < N2; t += TPE) {
      double x[Nq], y[N];
      ld<Nq>(V2 + t, N2, x); mat1<Nq, N, true>(B, x, y);
      ld<Nq>(V3 + t, N2, x); mat1_acc<Nq, N, true>(C, x, y);
      st<N>(ACC + t, N2, y);
    }
  }
  __syncthreads();

  /* ---------------- faces, one direction (2 own faces + their 2 neighbours) at a time */
  double *RA = S, *RB = S + L::RA, *RC = RB + L::RB, *RD = RC + L::R1;
  const int32_t *fr = frec + 6 * (active ? slot : 0);
#pragma unroll 1
  for (int d = 0; d < 3; d++) {
    const int sa = d == 0 ? N : 1, sb = d == 2 ? N : N2, sn = d == 0 ? 1 : (d == 1 ? N : N2);
    const int la = d == 0 ? 1 : 0, lb = d == 2 ? 1 : 2;
    /* T: traces u and du/dn on the face nodes: [side][u,dn][a + N b]; sides: own0 own1 nbr0 nbr1 */
    if (active) {
      for (int t = lt; t < 3 * N2; t += TPE) {
        int which = t / N2, ab = t % N2, a = ab % N, b = ab / N;
        double x[N];
        if (which == 0) {
          ld<N>(U + a * sa + b * sb, sn, x);
          double d0 = 0., d1 = 0.;
#pragma unroll
          for (int m = 0; m < N; m++) { d0 = fma(D[m], x[m], d0); d1 = fma(D[(N - 1) * N + m], x[m], d1); }
          RA[ab] = x[0]; RA[N2 + ab] = d0; RA[2 * N2 + ab] = x[N - 1]; RA[3 * N2 + ab] = d1;
        } else {
          const d4b_hm_t *h = A.hm + fr[2 * d + which - 1];
          int nb = h->nbr;
          if (nb >= 0) {
            const double *un = (nb < A.n_local) ? A.u + A.node_off[nb] : A.u_ghost + (A.node_off[nb] - A.local_nodes);
            int fp = h->f_p, dp = fp >> 1, lay = (fp & 1) ? N - 1 : 0;
            int pa = dp == 0 ? N : 1, pb = dp == 2 ? N : N2, pn = dp == 0 ? 1 : (dp == 1 ? N : N2);
            const double *p = un + a * pa + b * pb;
            double dn = 0.;
#pragma unroll
            for (int m = 0; m < N; m++) { x[m] = __ldg(p + m * pn); dn = fma(D[lay * N + m], x[m], dn); }
            double *T = RA + (which + 1) * 2 * N2;
            T[ab] = (fp & 1) ? x[N - 1] : x[0]; T[N2 + ab] = dn;
          }
        }
      }
    }
    __syncthreads();
    /* S1: contraction along face axis a: [side][uB,uC,dnB][qa + Nq b] */
    if (active) {
      for (int t = lt; t < 8 * N; t += TPE) {
        int b = t % N, fld = (t / N) & 1, side = t / (2 * N);
        if (side >= 2 && A.hm[fr[2 * d + side - 2]].nbr < 0) continue;
        double x[N], y[Nq];
        ld<N>(RA + (side * 2 + fld) * N2 + N * b, 1, x);
        double *o = RB + side * 3 * NQ + Nq * b;
        if (fld == 0) {
          mat1<N, Nq, false>(B, x, y); st<Nq>(o, 1, y);
          mat1<N, Nq, false>(C, x, y); st<Nq>(o + NQ, 1, y);
        } else {
          mat1<N, Nq, false>(B, x, y); st<Nq>(o + 2 * NQ, 1, y);
        }
      }
    }
    __syncthreads();
    /* QV: contraction along face axis b: [side][u, du/da, du/db, du/dn][qa + Nq qb] */
    if (active) {
      for (int t = lt; t < 12 * Nq; t += TPE) {
        int qa = t % Nq, fld = (t / Nq) % 3, side = t / (3 * Nq);
        if (side >= 2 && A.hm[fr[2 * d + side - 2]].nbr < 0) continue;
        double x[N], y[Nq];
        ld<N>(RB + (side * 3 + fld) * NQ + qa, Nq, x);
        double *o = RA + side * 4 * Q2 + qa;
        if (fld == 0) {
          mat1<N, Nq, false>(B, x, y); st<Nq>(o, Nq, y);
          mat1<N, Nq, false>(C, x, y); st<Nq>(o + 2 * Q2, Nq, y);
        } else {
          mat1<N, Nq, false>(B, x, y); st<Nq>(o + (fld == 1 ? 1 : 3) * Q2, Nq, y);
        }
      }
    }
    __syncthreads();
    /* FLUX at mortar quadrature points (…_flux_sipg.c:645-677 interior, :139-191 Dirichlet):
     * FH[face][F, Ha, Hb, Hn][q] */
    if (active) {
      for (int t = lt; t < 2 * Q2; t += TPE) {
        int s = t / Q2, q = t % Q2, qa = q % Nq, qb = q / Nq;
        const d4b_hm_t *h = A.hm + fr[2 * d + s];
        const double *F = A.fac + h->fac_off;
        const double *o = RA + s * 4 * Q2;
        double w = c_ft.w[qa] * c_ft.w[qb];
        double am[3] = { F[q], F[Q2 + q], F[2 * Q2 + q] };
        double um = o[q];
        double qm = am[la] * o[Q2 + q] + am[lb] * o[2 * Q2 + q] + am[d] * o[3 * Q2 + q];
        double jump, t1, sjs, hs;
        int nb = h->nbr;
        if (nb >= 0) {
          int src = f_tsrc(h->tcode, Nq, qa, qb), dp = h->f_p >> 1;
          int pla = dp == 0 ? 1 : 0, plb = dp == 2 ? 1 : 2;
          const double *p = RA + (2 + s) * 4 * Q2 + src;
          double qp = F[(3 + pla) * Q2 + q] * p[Q2] + F[(3 + plb) * Q2 + q] * p[2 * Q2] + F[(3 + dp) * Q2 + q] * p[3 * Q2];
          jump = um - p[0]; t1 = -0.5 * (qm + qp); sjs = F[6 * Q2 + q]; hs = -0.5;
        } else {
          double g = (h->g_off >= 0) ? A.bc_g[h->g_off + q] : 0.;
          jump = um - g; t1 = -qm; sjs = F[3 * Q2 + q]; hs = -1.0;
        }
        double *fh = RB + s * 4 * Q2 + q;
        double wj = w * hs * jump;
        fh[0] = w * (t1 + sjs * jump); fh[Q2] = wj * am[la]; fh[2 * Q2] = wj * am[lb]; fh[3 * Q2] = wj * am[d];
      }
    }
    __syncthreads();
    /* BACK1: transposed contraction along b: R1[face][X1, Ha1, Hn1][qa + Nq b'] */
    if (active) {
      for (int t = lt; t < 6 * Nq; t += TPE) {
        int qa = t % Nq, k = (t / Nq) % 3, s = t / (3 * Nq);
        const double *fh = RB + s * 4 * Q2 + qa;
        double x[Nq], y[N];
        if (k == 0) {
          ld<Nq>(fh, Nq, x); mat1<Nq, N, true>(B, x, y);
          ld<Nq>(fh + 2 * Q2, Nq, x); mat1_acc<Nq, N, true>(C, x, y);
        } else {
          ld<Nq>(fh + (k == 1 ? 1 : 3) * Q2, Nq, x); mat1<Nq, N, true>(B, x, y);
        }
        st<N>(RC + (s * 3 + k) * NQ + qa, Nq, y);
      }
    }
    __syncthreads();
    /* BACK2: transposed contraction along a: R2[face][Y, Z][a' + N b'] */
    if (active) {
      for (int t = lt; t < 4 * N; t += TPE) {
        int b = t % N, k = (t / N) & 1, s = t / (2 * N);
        const double *r1 = RC + s * 3 * NQ + Nq * b;
        double x[Nq], y[N];
        if (k == 0) {
          ld<Nq>(r1, 1, x); mat1<Nq, N, true>(B, x, y);
          ld<Nq>(r1 + NQ, 1, x); mat1_acc<Nq, N, true>(C, x, y);
        } else {
          ld<Nq>(r1 + 2 * NQ, 1, x); mat1<Nq, N, true>(B, x, y);
        }
        st<N>(RD + (s * 2 + k) * N2 + N * b, 1, y);
      }
    }
    __syncthreads();
    /* LIFT (…_flux_sipg.c:918-958): face layer += Y, normal line += D[layer,:] Z */
    if (active) {
      for (int t = lt; t < N2; t += TPE) {
        int a = t % N, b = t / N;
        double *acc = ACC + a * sa + b * sb;
        double y0 = RD[t], z0 = RD[N2 + t], y1 = RD[2 * N2 + t], z1 = RD[3 * N2 + t];
#pragma unroll
        for (int m = 0; m < N; m++) {
          double v = acc[m * sn] + D[m] * z0 + D[(N - 1) * N + m] * z1;
          if (m == 0) v += y0;
          if (m == N - 1) v += y1;
          acc[m * sn] = v;
        }
      }
    }
    __syncthreads();
  }

  if (active) {
    double *Aue = A.Au + A.node_off[e];
#pragma unroll
    for (int m = 0; m < N3 / TPE; m++) Aue[lt + TPE * m] = ACC[lt + TPE * m];
  }
}

/* ------------------------------------------------------------------ host side */
int d4b_launch_generic(const d4b_mesh *m, const double *u, double *Au, const int32_t *elem_list,
                       int n_list, cudaStream_t st);

template <int N, int Nq>
static int launch_bucket(const d4b_mesh *m, const d4b_fast_bucket &b, const double *u, double *Au, cudaStream_t st)
{
  using L = FL<N, Nq>;
  static bool attr = false;
  size_t smem = (size_t)L::PER * L::EPB * sizeof(double);
  if (!attr) {
    D4B_CUDA(cudaFuncSetAttribute(d4b_apply_fast_kernel<N, Nq>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = true;
  }
  fast_args A;
  A.node_off = m->node_off; A.quad_off = m->quad_off; A.G = m->G; A.hm_begin = m->hm_begin; A.hm = m->hm;
  A.fac = m->fac; A.bc_g = m->bc_g; A.u = u; A.u_ghost = m->u_ghost; A.Au = Au;
  A.elem_list = b.elem_list; A.n_list = b.count; A.n_local = m->n_local; A.local_nodes = m->local_nodes;
  int grid = (b.count + L::EPB - 1) / L::EPB;
  d4b_apply_fast_kernel<N, Nq><<<grid, L::THREADS, smem, st>>>(A, b.frec);
  D4B_CUDA(cudaGetLastError());
  d4b_count_launch(1);
  return D4B_OK;
}

#define FAST_PAIRS(X) X(2,2) X(2,3) X(2,4) X(3,3) X(3,4) X(3,5) X(4,4) X(4,5) X(4,6) X(5,5) X(5,6) X(5,7) \
                      X(6,6) X(6,7) X(6,8) X(7,7) X(7,8) X(8,8) X(8,9)

bool d4b_fast_supported(int N, int Nq)
{
#define X(a, b) if (N == a && Nq == b) return true;
  FAST_PAIRS(X)
#undef X
  return false;
}

static int upload_tab(int qt, int N, int Nq, cudaStream_t st)
{
  static int cur_qt = -1, cur_N = -1, cur_Nq = -1;
  static fast_tab *h = nullptr;
  if (cur_qt == qt && cur_N == N && cur_Nq == Nq) return D4B_OK;
  if (!h) D4B_CUDA(cudaMallocHost((void **)&h, 4 * sizeof(fast_tab))); /* small ring: async copies stay valid */
  static int ring = 0;
  fast_tab *t = h + (ring++ & 3);
  memset(t, 0, sizeof *t);
  double I[D4B_OPS_MAXN * D4B_OPS_MAXN], Dm[D4B_OPS_MAXN * D4B_OPS_MAXN], x[D4B_OPS_MAXN], w[D4B_OPS_MAXN];
  d4b_ops_interp(N - 1, Nq - 1, qt, -1, I); d4b_ops_diff(N - 1, Dm); d4b_ops_quad(Nq - 1, qt, x, w);
  for (int q = 0; q < Nq; q++) {
    t->w[q] = w[q];
    for (int j = 0; j < N; j++) {
      t->B[q * N + j] = I[q * N + j];
      double s = 0.;
      for (int k = 0; k < N; k++) s += I[q * N + k] * Dm[k * N + j];
      t->C[q * N + j] = s;
    }
  }
  for (int i = 0; i < N * N; i++) t->D[i] = Dm[i];
  D4B_CUDA(cudaMemcpyToSymbolAsync(c_ft, t, sizeof(fast_tab), 0, cudaMemcpyHostToDevice, st));
  cur_qt = qt; cur_N = N; cur_Nq = Nq;
  return D4B_OK;
}

int d4b_launch_fast(const d4b_mesh *m, const double *u, double *Au, cudaStream_t st, int *handled)
{
  *handled = 1;
  for (int i = 0; i < m->n_fast; i++) {
    const d4b_fast_bucket &b = m->fast[i];
    int rc = upload_tab(m->quad_type, b.N, b.Nq, st);
    if (rc != D4B_OK) return rc;
    rc = D4B_EINVAL;
#define X(a, c) if (b.N == a && b.Nq == c) rc = launch_bucket<a, c>(m, b, u, Au, st);
    FAST_PAIRS(X)
#undef X
    if (rc != D4B_OK) return rc;
  }
  if (m->n_generic > 0) return d4b_launch_generic(m, u, Au, m->generic_list, m->n_generic, st);
  return D4B_OK;
}
