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
#define FAST_PAIRS(X) X(2,2) X(2,3) X(2,4) X(3,3) X(3,4) X(3,5) X(4,4) X(4,5) X(4,6) X(5,5) X(5,6) X(5,7) \
                      X(6,6) X(6,7) X(6,8) X(7,7) X(7,8) X(8,8) X(8,9)
/* one immutable table per template instance (written once per quadrature type): launches of different
 * instances on different streams never race on it */
enum {
#define X(a, b) FTI_##a##_##b,
  FAST_PAIRS(X)
#undef X
  FAST_NINST
};
template <int N, int Nq> struct fast_index;
#define X(a, b) template <> struct fast_index<a, b> { static constexpr int value = FTI_##a##_##b; };
FAST_PAIRS(X)
#undef X
__constant__ fast_tab c_fts[FAST_NINST];

struct fast_args {
  const int64_t *node_off; const int64_t *quad_off; const double *G; const int32_t *hm_begin;
  const d4b_hm_t *hm; const double *fac; const double *bc_g;
  const double *u, *u_ghost; double *Au; const int32_t *elem_list; int n_list; int n_local;
  int64_t local_nodes;
  const v3_slot *slots;   /* per-slot metadata resolved once per bucket (d4b_v3_prepare_slots): one read instead of
                             the frec -> hm -> node_off chase */
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

/* Shared-memory paddings of one instance.  The defaults are the hand-chosen round-1 strides; the specialisations
 * in d4b_fast_pads.inc come from tools/fast_bank_sim.py, a bank-conflict model of every shared access of the
 * kernel (threads of N*N-wide elements straddle half-warps, so the best strides are not the obvious odd ones).
 * Measured at p = 4 (ncu, 32^3 brick): shared wavefronts 52.1 M -> 41.4 M, 323.6 -> 305.9 us. */
template <int N, int Nq> struct fast_pad {
  static constexpr int SJ = N + 1, AS = N * N + (N == 4 ? 4 : 0), BS = N * Nq + 1, TB = N * N + 1, QP = Nq * Nq, PP = 0;
};
#define FAST_PAD(n, q, sj, as, bs, tb, qp, pp)                                                      \
  template <> struct fast_pad<n, q> { static constexpr int SJ = sj, AS = as, BS = bs, TB = tb, QP = qp, PP = pp; };
#include "d4b_fast_pads.inc"
#undef FAST_PAD

template <int N, int Nq> struct FL {
  using P = fast_pad<N, Nq>;
  static constexpr int N2 = N * N, N3 = N * N * N, Q2 = Nq * Nq, Q3 = Nq * Nq * Nq, NQ = N * Nq;
  static constexpr int SJ = P::SJ, SK = N * SJ, UA = N * SK;              /* U, ACC: (i,j,k) -> i + SJ j + SK k */
  static constexpr int AS = P::AS, AB = Nq * AS;                          /* A bufs: (i,j,c) -> i + N j + AS c  */
  static constexpr int BS = P::BS, BB = Nq * BS;                          /* B bufs: (i,b,c) -> i + BS b + N c  */
  static constexpr int VOL = 2 * AB + 3 * BB;
  static constexpr int TB = P::TB;                                        /* trace block stride */
  static constexpr int QP = P::QP;                                        /* stride of a Q2 block of the face scratch */
  static constexpr int TR = 8 * TB, QV = 4 * 2 * QP, R1 = 2 * 3 * NQ;
  static constexpr int S1 = 4 * 3 * NQ, FH = 2 * 4 * QP, R2 = 2 * 2 * N2;
  static constexpr int RA = (TR > QV ? (TR > R1 ? TR : R1) : (QV > R1 ? QV : R1));
  static constexpr int RB = (S1 > FH ? (S1 > R2 ? S1 : R2) : (FH > R2 ? FH : R2));
  static constexpr int FACE = RA + RB;
  static constexpr int SCR = VOL > FACE ? VOL : FACE;
  static constexpr int PER = 2 * UA + SCR + P::PP;   /* doubles of shared memory per element */
  static constexpr int TPE = N2;                      /* threads per element */
  static constexpr int EPB = (128 / TPE) > 0 ? (128 / TPE) : 1; /* elements per CTA */
  static constexpr int THREADS = TPE * EPB;
};

/* face-local axis strides for direction d in an (i,j,k) array with strides (1, sj, sk) */
__device__ __forceinline__ void face_strides(int d, int sj, int sk, int &sa, int &sb, int &sn)
{
  sa = d == 0 ? sj : 1; sb = d == 2 ? sj : sk; sn = d == 0 ? 1 : (d == 1 ? sj : sk);
}

template <int N, int Nq>
__global__ void __launch_bounds__(FL<N, Nq>::THREADS)
d4b_apply_fast_kernel(const fast_args A, const int32_t *__restrict__ frec)
{
  using L = FL<N, Nq>;
  constexpr int N2 = L::N2, N3 = L::N3, Q2 = L::Q2, Q3 = L::Q3, NQ = L::NQ, TPE = L::TPE;
  constexpr int SJ = L::SJ, SK = L::SK, AS = L::AS, BS = L::BS, TB = L::TB, QP = L::QP;
  extern __shared__ double sm[];
  const int le = threadIdx.x / TPE, lt = threadIdx.x % TPE;
  const int slot = blockIdx.x * L::EPB + le;
  const bool active = slot < A.n_list;
  const v3_slot *sl = A.slots + (active ? slot : 0);
  double *U = sm + le * L::PER, *ACC = U + L::UA, *S = ACC + L::UA;
  double *V0 = S, *V1 = S + L::AB, *V2 = S + 2 * L::AB, *V3 = V2 + L::BB, *V4 = V3 + L::BB;
  const fast_tab &c_ft = c_fts[fast_index<N, Nq>::value];
  const double *B = c_ft.B, *C = c_ft.C, *D = c_ft.D;
  const double *ue = A.u + __ldg(&sl->node_off);
  const int fa = lt % N, fb = lt / N; /* my face node (T stage) */

  /* neighbour normal lines of direction d, prefetched into registers one direction ahead */
  double xn[2][N];
  int nlay[2];
  auto prefetch = [&](int d) {
#pragma unroll
    for (int s = 0; s < 2; s++) {
      const v3_face *h = &sl->f[2 * d + s];
      const int code = __ldg(&h->code); nlay[s] = -1;
      if ((code >> 6) & 3) {
        const int64_t nu = __ldg(&h->nbr_u);
        const double *un = nu >= 0 ? A.u + nu : A.u_ghost + (-nu - 1);
        int fp = (code >> 3) & 7, dp = fp >> 1; nlay[s] = (fp & 1) ? N - 1 : 0;
        int pa, pb, pn; face_strides(dp, N, N2, pa, pb, pn);
        const double *p = un + fa * pa + fb * pb;
#pragma unroll
        for (int m = 0; m < N; m++) xn[s][m] = __ldg(p + m * pn);
      }
    }
  };
  if (active) {
    /* L2 prefetch of the metric block (read after the three interpolation stages) and of the six
     * face-factor records (read after the whole volume phase) */
    {
      const char *g = reinterpret_cast<const char *>(A.G + 6 * __ldg(&sl->quad_off));
      for (int o = lt * 128; o < 6 * Q3 * 8; o += TPE * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + o));
    }
    if (lt < 6) {
      const v3_face *h = &sl->f[lt];
      const char *p = reinterpret_cast<const char *>(A.fac + __ldg(&h->fac_off));
      const int bytes = (((__ldg(&h->code) >> 6) & 3) ? 7 : 4) * Q2 * 8;
      for (int o = 0; o < bytes; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
    }
    prefetch(0);
#pragma unroll
    for (int m = 0; m < N3 / TPE; m++) { /* g = lt + TPE m = (i,j) + N2 k */
      U[(lt % N) + SJ * (lt / N) + SK * m] = ue[lt + TPE * m];
    }
  }
  __syncthreads();

  /* ---------------- volume (Quadrature/d4est_quadrature.c:263-375, sum-factorised) */
  if (active) { /* along t: line (i,j) */
    const int i = lt % N, j = lt / N;
    double x[N], y[Nq];
    ld<N>(U + i + SJ * j, SK, x);
    mat1<N, Nq, false>(B, x, y); st<Nq>(V0 + lt, AS, y);
    mat1<N, Nq, false>(C, x, y); st<Nq>(V1 + lt, AS, y);
  }
  __syncthreads();
  if (active) { /* along s: lines (i,c) */
    for (int t = lt; t < NQ; t += TPE) {
      int i = t % N, c = t / N;
      double x[N], y[Nq];
      ld<N>(V0 + i + AS * c, N, x);
      mat1<N, Nq, false>(B, x, y); st<Nq>(V2 + i + N * c, BS, y);
      mat1<N, Nq, false>(C, x, y); st<Nq>(V3 + i + N * c, BS, y);
      ld<N>(V1 + i + AS * c, N, x);
      mat1<N, Nq, false>(B, x, y); st<Nq>(V4 + i + N * c, BS, y);
    }
  }
  __syncthreads();
  if (active) { /* along r: lines (b,c); metric; transposed r-contraction; in place, in registers */
    const double *Ge = A.G + 6 * __ldg(&sl->quad_off);
    for (int t = lt; t < Q2; t += TPE) {
      int b = t % Nq, c = t / Nq;
      double Gv[6][Nq];
#pragma unroll
      for (int k = 0; k < 6; k++)
#pragma unroll
        for (int a = 0; a < Nq; a++) Gv[k][a] = __ldg(Ge + k * Q3 + Nq * t + a);
      double x[N], gr[Nq], gs[Nq], gt[Nq];
      const int o = BS * b + N * c;
      ld<N>(V2 + o, 1, x); mat1<N, Nq, false>(C, x, gr);
      ld<N>(V3 + o, 1, x); mat1<N, Nq, false>(B, x, gs);
      ld<N>(V4 + o, 1, x); mat1<N, Nq, false>(B, x, gt);
#pragma unroll
      for (int a = 0; a < Nq; a++) {
        double g0 = gr[a], g1 = gs[a], g2 = gt[a];
        gr[a] = Gv[0][a] * g0 + Gv[1][a] * g1 + Gv[2][a] * g2;
        gs[a] = Gv[1][a] * g0 + Gv[3][a] * g1 + Gv[4][a] * g2;
        gt[a] = Gv[2][a] * g0 + Gv[4][a] * g1 + Gv[5][a] * g2;
      }
      mat1<Nq, N, true>(C, gr, x); st<N>(V2 + o, 1, x);
      mat1<Nq, N, true>(B, gs, x); st<N>(V3 + o, 1, x);
      mat1<Nq, N, true>(B, gt, x); st<N>(V4 + o, 1, x);
    }
  }
  __syncthreads();
  if (active) { /* transposed s-contraction: lines (i,c) */
    for (int t = lt; t < NQ; t += TPE) {
      int i = t % N, c = t / N;
      double x[Nq], y[N];
      ld<Nq>(V2 + i + N * c, BS, x); mat1<Nq, N, true>(B, x, y);
      ld<Nq>(V3 + i + N * c, BS, x); mat1_acc<Nq, N, true>(C, x, y);
      st<N>(V0 + i + AS * c, N, y);
      ld<Nq>(V4 + i + N * c, BS, x); mat1<Nq, N, true>(B, x, y);
      st<N>(V1 + i + AS * c, N, y);
    }
  }
  __syncthreads();
  if (active) { /* transposed t-contraction: line (i,j) -> ACC */
    const int i = lt % N, j = lt / N;
    double x[Nq], y[N];
    ld<Nq>(V0 + lt, AS, x); mat1<Nq, N, true>(B, x, y);
    ld<Nq>(V1 + lt, AS, x); mat1_acc<Nq, N, true>(C, x, y);
    st<N>(ACC + i + SJ * j, SK, y);
  }
  __syncthreads();

  /* ---------------- faces, one direction (2 own faces + their 2 neighbours) at a time */
  double *RA = S, *RB = S + L::RA;
#pragma unroll 1
  for (int d = 0; d < 3; d++) {
    int sa, sb, sn; face_strides(d, SJ, SK, sa, sb, sn);
    /* T: traces u and du/dn at my face node: RA[(side*2 + fld)*TB + a + N b]; sides own0 own1 nbr0 nbr1 */
    if (active) {
      double x[N];
      ld<N>(U + fa * sa + fb * sb, sn, x);
      double d0 = 0., d1 = 0.;
#pragma unroll
      for (int m = 0; m < N; m++) { d0 = fma(D[m], x[m], d0); d1 = fma(D[(N - 1) * N + m], x[m], d1); }
      RA[lt] = x[0]; RA[TB + lt] = d0; RA[2 * TB + lt] = x[N - 1]; RA[3 * TB + lt] = d1;
#pragma unroll
      for (int s = 0; s < 2; s++) {
        if (nlay[s] >= 0) {
          double dn = 0.;
          const double *Dr = D + nlay[s] * N;
#pragma unroll
          for (int m = 0; m < N; m++) dn = fma(Dr[m], xn[s][m], dn);
          RA[(4 + 2 * s) * TB + lt] = nlay[s] ? xn[s][N - 1] : xn[s][0];
          RA[(5 + 2 * s) * TB + lt] = dn;
        }
      }
    }
    const bool nbnd0 = active && nlay[0] < 0, nbnd1 = active && nlay[1] < 0; /* boundary flags of this direction */
    if (active && d < 2) prefetch(d + 1);
    __syncthreads();
    /* S1: contraction along face axis a: RB[(side*3 + {uB,uC,dnB})*NQ + qa + Nq b] */
    if (active) {
      for (int t = lt; t < 8 * N; t += TPE) {
        int b = t % N, fld = (t / N) & 1, side = t / (2 * N);
        if ((side == 2 && nbnd0) || (side == 3 && nbnd1)) continue;
        double x[N], y[Nq];
        ld<N>(RA + (side * 2 + fld) * TB + N * b, 1, x);
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
    /* QV: contraction along face axis b, then fold the gradient with the face factors:
     * RA[(side*2 + 0)*QP + q] = u,  RA[(side*2 + 1)*QP + q] = sum_l a_l du/dr_l  (q in MY point order) */
    if (active) {
      for (int t = lt; t < 4 * Nq; t += TPE) {
        int qa = t % Nq, side = t / Nq, s = side & 1;
        if ((side == 2 && nbnd0) || (side == 3 && nbnd1)) continue;
        const v3_face *h = &sl->f[2 * d + s];
        const double *F = A.fac + __ldg(&h->fac_off);
        const double *in = RB + side * 3 * NQ + qa;
        double x[N], u[Nq], ua[Nq], ub[Nq], un[Nq];
        ld<N>(in, Nq, x); mat1<N, Nq, false>(B, x, u); mat1<N, Nq, false>(C, x, ub);
        ld<N>(in + NQ, Nq, x); mat1<N, Nq, false>(B, x, ua);
        ld<N>(in + 2 * NQ, Nq, x); mat1<N, Nq, false>(B, x, un);
        double *o = RA + side * 2 * QP;
        if (side < 2) {
          const int ia = d == 0 ? 1 : 0, ib = d == 2 ? 1 : 2;
#pragma unroll
          for (int qb = 0; qb < Nq; qb++) {
            int q = qa + Nq * qb;
            o[q] = u[qb];
            o[QP + q] = F[ia * Q2 + q] * ua[qb] + F[ib * Q2 + q] * ub[qb] + F[d * Q2 + q] * un[qb];
          }
        } else {
          const int hc = __ldg(&h->code);
          const int tc = hc & 7, dp = ((hc >> 3) & 7) >> 1;
          const int ia = dp == 0 ? 1 : 0, ib = dp == 2 ? 1 : 2;
#pragma unroll
          for (int qb = 0; qb < Nq; qb++) { /* (qa,qb) is the neighbour's point; find mine */
            int xx = (tc & 2) ? Nq - 1 - qa : qa, yy = (tc & 4) ? Nq - 1 - qb : qb;
            int q = (tc & 1) ? xx + Nq * yy : yy + Nq * xx;
            o[q] = u[qb];
            o[QP + q] = F[(3 + ia) * Q2 + q] * ua[qb] + F[(3 + ib) * Q2 + q] * ub[qb] + F[(3 + dp) * Q2 + q] * un[qb];
          }
        }
      }
    }
    __syncthreads();
    /* FLUX at mortar quadrature points (…_flux_sipg.c:645-677 interior, :139-191 Dirichlet):
     * RB[(face*4 + {F,Ha,Hb,Hn})*QP + q] */
    if (active) {
      for (int t = lt; t < 2 * Q2; t += TPE) {
        int s = t / Q2, q = t % Q2, qa = q % Nq, qb = q / Nq;
        const v3_face *h = &sl->f[2 * d + s];
        const double *F = A.fac + __ldg(&h->fac_off);
        const bool bnd = s ? nbnd1 : nbnd0;
        double am0 = F[q], am1 = F[Q2 + q], am2 = F[2 * Q2 + q], sjs = F[(bnd ? 3 : 6) * Q2 + q];
        double w = c_ft.w[qa] * c_ft.w[qb];
        double um = RA[s * 2 * QP + q], qm = RA[(s * 2 + 1) * QP + q];
        double jump, t1, hs;
        if (!bnd) {
          jump = um - RA[(4 + 2 * s) * QP + q]; t1 = -0.5 * (qm + RA[(5 + 2 * s) * QP + q]); hs = -0.5;
        } else {
          const int64_t go = __ldg(&h->g_off);
          double g = (go >= 0) ? A.bc_g[go + q] : 0.;
          jump = um - g; t1 = -qm; hs = -1.0;
        }
        double *fh = RB + s * 4 * QP + q;
        double wj = w * hs * jump;
        fh[0] = w * (t1 + sjs * jump);
        fh[QP] = wj * (d == 0 ? am1 : am0);
        fh[2 * QP] = wj * (d == 2 ? am1 : am2);
        fh[3 * QP] = wj * (d == 0 ? am0 : (d == 1 ? am1 : am2));
      }
    }
    __syncthreads();
    /* BACK1: transposed contraction along b: RA[(face*3 + {X1,Ha1,Hn1})*NQ + qa + Nq b'] */
    if (active) {
      for (int t = lt; t < 6 * Nq; t += TPE) {
        int qa = t % Nq, k = (t / Nq) % 3, s = t / (3 * Nq);
        const double *fh = RB + s * 4 * QP + qa;
        double x[Nq], y[N];
        if (k == 0) {
          ld<Nq>(fh, Nq, x); mat1<Nq, N, true>(B, x, y);
          ld<Nq>(fh + 2 * QP, Nq, x); mat1_acc<Nq, N, true>(C, x, y);
        } else {
          ld<Nq>(fh + (k == 1 ? 1 : 3) * QP, Nq, x); mat1<Nq, N, true>(B, x, y);
        }
        st<N>(RA + (s * 3 + k) * NQ + qa, Nq, y);
      }
    }
    __syncthreads();
    /* BACK2: transposed contraction along a: RB[(face*2 + {Y,Z})*N2 + a' + N b'] */
    if (active) {
      for (int t = lt; t < 4 * N; t += TPE) {
        int b = t % N, k = (t / N) & 1, s = t / (2 * N);
        const double *r1 = RA + s * 3 * NQ + Nq * b;
        double x[Nq], y[N];
        if (k == 0) {
          ld<Nq>(r1, 1, x); mat1<Nq, N, true>(B, x, y);
          ld<Nq>(r1 + NQ, 1, x); mat1_acc<Nq, N, true>(C, x, y);
        } else {
          ld<Nq>(r1 + 2 * NQ, 1, x); mat1<Nq, N, true>(B, x, y);
        }
        st<N>(RB + (s * 2 + k) * N2 + N * b, 1, y);
      }
    }
    __syncthreads();
    /* LIFT (…_flux_sipg.c:918-958): face layer += Y, normal line += D[layer,:] Z */
    if (active) {
      double *acc = ACC + fa * sa + fb * sb;
      double y0 = RB[lt], z0 = RB[N2 + lt], y1 = RB[2 * N2 + lt], z1 = RB[3 * N2 + lt];
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
  if (active) {
    double *Aue = A.Au + __ldg(&sl->node_off);
#pragma unroll
    for (int m = 0; m < N3 / TPE; m++) Aue[lt + TPE * m] = ACC[(lt % N) + SJ * (lt / N) + SK * m];
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
  { int rcs = d4b_v3_prepare_slots(m, b, st); if (rcs != D4B_OK) return rcs; }
  fast_args A;
  A.slots = (const v3_slot *)b.v3_slots;
  A.node_off = m->node_off; A.quad_off = m->quad_off; A.G = m->G; A.hm_begin = m->hm_begin; A.hm = m->hm;
  A.fac = m->fac; A.bc_g = m->bc_g; A.u = u; A.u_ghost = m->u_ghost; A.Au = Au;
  A.elem_list = b.elem_list; A.n_list = b.count; A.n_local = m->n_local; A.local_nodes = m->local_nodes;
  int grid = (b.count + L::EPB - 1) / L::EPB;
  d4b_apply_fast_kernel<N, Nq><<<grid, L::THREADS, smem, st>>>(A, b.frec);
  D4B_CUDA(cudaGetLastError());
  d4b_count_launch(1);
  return D4B_OK;
}

bool d4b_fast_supported(int N, int Nq)
{
#define X(a, b) if (N == a && Nq == b) return true;
  FAST_PAIRS(X)
#undef X
  return false;
}

static int upload_tab(int qt, int N, int Nq, cudaStream_t st)
{
  static int have_qt[FAST_NINST];
  static bool init = false;
  if (!init) { for (int i = 0; i < FAST_NINST; i++) have_qt[i] = -1; init = true; }
  int idx = -1;
#define X(a, c) if (N == a && Nq == c) idx = FTI_##a##_##c;
  FAST_PAIRS(X)
#undef X
  if (idx < 0) { d4b_set_error("no line-task instance for N=%d Nq=%d", N, Nq); return D4B_EINVAL; }
  if (have_qt[idx] == qt) return D4B_OK;
  if (have_qt[idx] != -1) D4B_CUDA(cudaDeviceSynchronize());   /* other quadrature type still in flight */
  fast_tab tt, *t = &tt;
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
  D4B_CUDA(cudaMemcpyToSymbol(c_fts, t, sizeof(fast_tab), (size_t)idx * sizeof(fast_tab), cudaMemcpyHostToDevice));
  D4B_CUDA(cudaDeviceSynchronize());   /* the staged copy has landed before any stream launches */
  (void)st;
  have_qt[idx] = qt;
  return D4B_OK;
}

static int launch_fast_bucket(const d4b_mesh *m, const d4b_fast_bucket &b, const double *u, double *Au, cudaStream_t st)
{
  if (b.nrec && d4b_v3_supported(b.N, b.Nq)) return d4b_launch_v3(m, b, u, Au, st);
  int rc = upload_tab(m->quad_type, b.N, b.Nq, st);
  if (rc != D4B_OK) return rc;
  rc = D4B_EINVAL;
#define X(a, c) if (b.N == a && b.Nq == c) rc = launch_bucket<a, c>(m, b, u, Au, st);
  FAST_PAIRS(X)
#undef X
  return rc;
}

/* One apply = one launch per degree bucket (regular-element kernels, hp kernel) + the completeness kernel.
 * Buckets write disjoint elements of Au and only read u, so with more than one launch they run CONCURRENTLY:
 * fork from the caller's stream to the level's side streams, largest bucket first, join back.  A random-hp
 * level has 8-12 launches of very different size; run back to back the small ones cost a launch latency each
 * and the tail of every large one leaves SMs idle. */
/* phase 0: the whole level.  A distributed apply that overlaps its halo exchange (d4b_apply_dev) calls phase 1
 * -- the batches of the regular buckets that touch no ghost element, while the exchange is in flight -- and
 * phase 2: the remaining batches, the hp buckets and the completeness kernel, queued behind the exchange. */
int d4b_launch_phase(const d4b_mesh *m, const double *u, double *Au, cudaStream_t st, int phase)
{
  struct item { int kind, idx; double w; };
  static const bool serial = getenv("D4B_NO_CONCURRENT") != nullptr;
  const d4b_fast_bucket *fb = phase == 0 ? m->fast : (phase == 1 ? m->fast_early : m->fast_late);
  /* round 0: everything but pass 2 of the two-pass hp path; round 1: pass 2 (it reads the traces W that ALL pass-1
   * launches of round 0 write, so it starts after the join of round 0) */
  for (int round = 0; round < 2; round++) {
    item it[4 * D4B_MAX_BUCKETS + 1];
    int n = 0;
    if (round == 0) {
      for (int i = 0; i < m->n_fast; i++)
        if (fb[i].count > 0) it[n++] = {0, i, (double)fb[i].count * fb[i].Nq * fb[i].Nq * fb[i].Nq};
      if (phase != 1) {
        for (int i = 0; i < m->n_hp; i++) it[n++] = {1, i, 4.0 * m->hp[i].n_rec / 6.0 * m->hp[i].Nq * m->hp[i].Nq * m->hp[i].Nq};
        if (m->n_generic > 0) it[n++] = {2, 0, 40.0 * m->n_generic * 64};
        /* the trace launches first: round 1 waits for them */
        for (int i = 0; i < m->n_tp1; i++) it[n++] = {3, i, 1e15 + 2.0 * m->tp1[i].n_rec / 6.0 * m->tp1[i].Nq * m->tp1[i].Nq * m->tp1[i].Nq};
      }
    } else {
      if (phase == 1) break;
      for (int i = 0; i < m->n_tp2; i++) it[n++] = {4, i, 2.0 * m->tp2[i].n_rec / 6.0 * m->tp2[i].Nq * m->tp2[i].Nq * m->tp2[i].Nq};
    }
    if (n == 0) continue;
    for (int i = 1; i < n; i++) { item x = it[i]; int j = i; while (j > 0 && it[j - 1].w < x.w) { it[j] = it[j - 1]; j--; } it[j] = x; }
    const bool fork = n > 1 && !serial && m->ev_fork;
    int used = 0;
    if (fork) D4B_CUDA(cudaEventRecord(m->ev_fork, st));
    for (int i = 0; i < n; i++) {
      cudaStream_t s = st;
      if (fork && i > 0) {
        const int k = (i - 1) % D4B_NSIDE;
        s = m->side[k];
        if (k >= used) { D4B_CUDA(cudaStreamWaitEvent(s, m->ev_fork, 0)); used = k + 1; }
      }
      int rc;
      if (it[i].kind == 0) rc = launch_fast_bucket(m, fb[it[i].idx], u, Au, s);
      else if (it[i].kind == 1) rc = d4b_launch_hp_bucket(m, it[i].idx, u, Au, s);
      else if (it[i].kind == 2) rc = d4b_launch_generic(m, u, Au, m->generic_list, m->n_generic, s);
      else rc = d4b_launch_tp_bucket(m, it[i].kind == 3 ? 1 : 2, it[i].idx, u, Au, s);
      if (rc != D4B_OK) return rc;
    }
    for (int k = 0; k < used; k++) {
      D4B_CUDA(cudaEventRecord(m->ev_join[k], m->side[k]));
      D4B_CUDA(cudaStreamWaitEvent(st, m->ev_join[k], 0));
    }
  }
  return D4B_OK;
}

int d4b_launch_fast(const d4b_mesh *m, const double *u, double *Au, cudaStream_t st, int *handled)
{
  *handled = 1;
  return d4b_launch_phase(m, u, Au, st, 0);
}
