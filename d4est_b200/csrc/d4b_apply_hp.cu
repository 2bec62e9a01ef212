/* d4b_apply_hp.cu -- degree-templated SIPG apply for the elements of an hp-adaptive mesh that are NOT
 * "regular": hanging faces in both roles (I am the coarse side: 4 half-mortars on a face; I am a fine
 * side: the neighbour's trace is evaluated on my quarter of its face), mixed-p neighbours, mortar
 * quadrature degree above my own, any inter-tree orientation, local or ghost neighbours, Dirichlet
 * boundary faces.  One template instance per (N, Nq) of the element itself; the neighbour's degree, the
 * mortar degree and the child indices are per-record run-time data.
 *
 * Reference semantics (same operator, one side at a time, see d4b_apply_generic.cu):
 *   dGMath/d4est_laplacian_with_opt_flux.c:212-813   slice, reorient, project to mortar, interpolate, 1/2
 *   dGMath/d4est_laplacian_with_opt_flux_sipg.c:529-958  SIPG flux, Galerkin integral, back-projection, lift
 *   Mesh/d4est_mortars_with_opt.c:521,561            p-/hp-prolongation to the mortar space and its adjoint
 *   Quadrature/d4est_quadrature.c:263-375            volume stiffness term
 *
 * Mapping: ONE WARP PER ELEMENT, no block barrier anywhere (__syncwarp only), a CTA is just a few
 * independent warps.  Shared memory per warp is sized for the (N, Nq) bucket, not for the largest
 * element of the level.
 *   volume: sum-factorised line tasks as in d4b_apply_fast.cu (1-D tables B = I, C = I D of the element's
 *     own (N, Nq) are compile-time-indexed __constant__ operands of the DFMAs).
 *   faces: one half-mortar record at a time, all 32 lanes: T (face traces of u and du/dn, own side from shared
 *     memory, neighbour straight from global memory) -> S1 (contraction along face axis a with E_a, E_a D) ->
 *     QV (along b: u and sum_l a_l du/dr_l at the mortar points, both sides, in my point order) -> FLUX ->
 *     BACK1/BACK2 (adjoint of QV/S1: Galerkin integral + back-projection) -> per-face accumulators Y (lift) and
 *     Z (normal-derivative lift).  The contractions are output-stationary: a lane owns one table row (own side
 *     in registers, N being a template parameter) and sweeps the lines, which arrive as shared-memory
 *     broadcasts.  E = (child) interpolation to the mortar Gauss points composed with the p-/hp-prolongation,
 *     tables I[deg][degq_mortar][child] and I D from the global 1-D pool (L1-resident).
 *   lift: Au = volume + sum_faces lift(Y) + D_n^T lift(Z), one direction at a time (v2's LIFT stage).
 * Records are processed in their stored order, so results are bitwise reproducible run to run.
 */
#include "d4b_internal.h"
#include "d4b_ops.h"
#include <vector>
#include <cstdlib>
#include <cstring>
#include <algorithm>

struct hp_rec {       /* 32 bytes: one half-mortar, resolved once per level */
  int64_t fac_off;    /* my factor block: a-[3], (a+[3],) sj*sigma, Nm^2 each                           */
  int64_t g_off;      /* boundary: Dirichlet data or -1                                                */
  int64_t nbr_u;      /* >= 0 offset into u; < 0: -(offset into u_ghost) - 1                           */
  int32_t code;       /* bits 0-2 f_m, 3-5 f_p, 6-8 tcode, 9-11 child_m+1, 12-14 child_p+1, 15 boundary, 16 Robin */
  int32_t dims;       /* Nm | Np << 8                                                                  */
};
struct hp_slot { int64_t node_off, quad_off; int32_t rec_begin, nrec; int64_t pad; };

__constant__ d4b_ops_dev c_hp_ops[2];

template <int N, int Nq> struct hp_tab { double B[Nq * N], C[Nq * N], D[N * N]; };

/* (N, Nq, spare) */
#define HP_PAIRS(X) X(2, 2, 4) X(2, 3, 4) X(2, 4, 4) X(2, 5, 4) X(3, 3, 4) X(3, 4, 4) X(3, 5, 4) X(3, 6, 4) X(4, 4, 4) X(4, 5, 4) X(4, 6, 4) X(4, 7, 4) \
                    X(5, 5, 2) X(5, 6, 2) X(5, 7, 2) X(5, 8, 2) X(6, 6, 2) X(6, 7, 2) X(6, 8, 2) X(7, 7, 1) X(7, 8, 1) X(7, 9, 1) \
                    X(8, 8, 1) X(8, 9, 1) X(8, 10, 1)
#define HP_MAXNP 8     /* neighbour N handled by the templated line tasks */
#define HP_MAXNM 12

/* one immutable __constant__ table per template instance (written once per quadrature type) */
template <int N, int Nq> struct hp_const;
#define X(a, b, g)                                                                                          \
  __constant__ hp_tab<a, b> c_hp_##a##_##b;                                                                 \
  template <> struct hp_const<a, b> {                                                                       \
    static __device__ __forceinline__ const hp_tab<a, b> &get() { return c_hp_##a##_##b; }                  \
    static cudaError_t put(const hp_tab<a, b> *h) { return cudaMemcpyToSymbol(c_hp_##a##_##b, h, sizeof(*h)); } \
  };
HP_PAIRS(X)
#undef X

struct hp_args {
  const hp_slot *slots; const hp_rec *recs; const double *G, *fac, *bc_g, *u, *u_ghost; double *Au;
  int n_list, qt, NMX, MX, per_warp;
};

template <int N, int Nq> struct HL {
  static constexpr int N2 = N * N, N3 = N2 * N, Q2 = Nq * Nq, Q3 = Q2 * Nq, NQ = N * Nq;
  static constexpr int SJ = N + 1, SK = N * (N + 1), UA = N * SK;   /* U, ACC: (i,j,k) -> i + SJ j + SK k */
  static constexpr int AS = N2 + (N == 4 ? 4 : 0), AB = Nq * AS;     /* A bufs: (i,j,c) -> i + N j + AS c  */
  static constexpr int BS = NQ + 1, BB = Nq * BS;                    /* B bufs: (i,c,b) -> i + N c + BS b  */
  static constexpr int VOL = 2 * AB + 3 * BB;
  static constexpr int FIXED = 2 * UA + 12 * N2;
};
/* face scratch of one warp: RA = max(T: 4 M^2, QV: 4 Nm^2, BACK1: 3 Nm N), RB = max(S1: 6 Nm M, FLUX: 4 Nm^2) */
__host__ __device__ static inline int hp_ra_doubles(int NMX, int MX)
{ int a = 4 * MX * MX, b = 4 * NMX * NMX; return a > b ? a : b; }
__host__ __device__ static inline int hp_group_doubles(int NMX, int MX)
{ int rb = 6 * NMX * MX > 4 * NMX * NMX ? 6 * NMX * MX : 4 * NMX * NMX; return hp_ra_doubles(NMX, MX) + rb; }

template <int n> __device__ __forceinline__ void hld(const double *p, int st, double (&x)[n])
{
#pragma unroll
  for (int i = 0; i < n; i++) x[i] = p[i * st];
}
template <int n> __device__ __forceinline__ void hst(double *p, int s, const double (&x)[n])
{
#pragma unroll
  for (int i = 0; i < n; i++) p[i * s] = x[i];
}
/* y[o] = sum_i M(o,i) x[i], M compile-time-indexed (constant-bank operands) */
template <int NI, int NO, bool TR>
__device__ __forceinline__ void hmat(const double *M, const double (&x)[NI], double (&y)[NO])
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
__device__ __forceinline__ void hmat_acc(const double *M, const double (&x)[NI], double (&y)[NO])
{
#pragma unroll
  for (int o = 0; o < NO; o++) {
    double s = y[o];
#pragma unroll
    for (int i = 0; i < NI; i++) s = fma(TR ? M[i * NO + o] : M[o * NI + i], x[i], s);
    y[o] = s;
  }
}
__device__ __forceinline__ void hp_face_strides(int d, int sj, int sk, int &sa, int &sb, int &sn)
{
  sa = d == 0 ? sj : 1; sb = d == 2 ? sj : sk; sn = d == 0 ? 1 : (d == 1 ? sj : sk);
}
__device__ __forceinline__ int hp_tcode_src(int tcode, int n, int a, int b)
{ /* m-order point -> p-order point (d4b_hm_t.tcode) */
  int x = (tcode & 1) ? a : b, y = (tcode & 1) ? b : a;
  if (tcode & 2) x = n - 1 - x;
  if (tcode & 4) y = n - 1 - y;
  return x + n * y;
}

template <int N, int Nq>
__global__ void __launch_bounds__(128, (N <= 4 ? 5 : 4)) d4b_apply_hp_kernel(const hp_args A)
{
  using L = HL<N, Nq>;
  constexpr int N2 = L::N2, N3 = L::N3, Q2 = L::Q2, Q3 = L::Q3, NQ = L::NQ;
  constexpr int SJ = L::SJ, SK = L::SK, AS = L::AS, BS = L::BS;
  extern __shared__ double sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = blockIdx.x * (blockDim.x >> 5) + warp;
  if (slot >= A.n_list) return;                         /* warps are independent: no block barrier below */
  const hp_tab<N, Nq> &T = hp_const<N, Nq>::get();
  const double *B = T.B, *C = T.C, *D = T.D;
  double *U = sm + (size_t)warp * A.per_warp, *ACC = U + L::UA, *YZ = ACC + L::UA, *S = YZ + 12 * N2;
  const hp_slot sl = A.slots[slot];
  const double *ue = A.u + sl.node_off;

  { /* L2 prefetch of what is read once, later: the metric block and the factor blocks of the records */
    const char *g = reinterpret_cast<const char *>(A.G + 6 * sl.quad_off);
    for (int o = lane * 128; o < 6 * Q3 * 8; o += 32 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + o));
    for (int r = lane >> 2; r < sl.nrec; r += 8) {
      const hp_rec *h = A.recs + sl.rec_begin + r;
      const int nm = __ldg(&h->dims) & 255, bytes = ((__ldg(&h->code) >> 15) & 1 ? 4 : 7) * nm * nm * 8;
      const char *p = reinterpret_cast<const char *>(A.fac + __ldg(&h->fac_off));
      for (int o = (lane & 3) * 128; o < bytes; o += 4 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
    }
  }
  for (int i = lane; i < N3; i += 32) U[(i % N) + SJ * ((i / N) % N) + SK * (i / N2)] = ue[i];
  for (int i = lane; i < 12 * N2; i += 32) YZ[i] = 0.;
  __syncwarp();

  /* ---------------- volume (Quadrature/d4est_quadrature.c:263-375, sum-factorised) */
  {
    double *V0 = S, *V1 = S + L::AB, *V2 = S + 2 * L::AB, *V3 = V2 + L::BB, *V4 = V3 + L::BB;
    for (int t = lane; t < N2; t += 32) { /* along t: line (i,j) */
      const int i = t % N, j = t / N;
      double x[N], y[Nq];
      hld<N>(U + i + SJ * j, SK, x);
      hmat<N, Nq, false>(B, x, y); hst<Nq>(V0 + t, AS, y);
      hmat<N, Nq, false>(C, x, y); hst<Nq>(V1 + t, AS, y);
    }
    __syncwarp();
    for (int t = lane; t < NQ; t += 32) { /* along s: lines (i,c) */
      const int i = t % N, c = t / N;
      double x[N], y[Nq];
      hld<N>(V0 + i + AS * c, N, x);
      hmat<N, Nq, false>(B, x, y); hst<Nq>(V2 + i + N * c, BS, y);
      hmat<N, Nq, false>(C, x, y); hst<Nq>(V3 + i + N * c, BS, y);
      hld<N>(V1 + i + AS * c, N, x);
      hmat<N, Nq, false>(B, x, y); hst<Nq>(V4 + i + N * c, BS, y);
    }
    __syncwarp();
    const double *Ge = A.G + 6 * sl.quad_off;
    for (int t = lane; t < Q2; t += 32) { /* along r: lines (b,c); metric; transposed r-contraction in registers */
      const int b = t % Nq, c = t / Nq;
      double x[N], gr[Nq], gs[Nq], gt[Nq];
      const int o = BS * b + N * c;
      const double *Gt = Ge + Nq * t;
      hld<N>(V2 + o, 1, x); hmat<N, Nq, false>(C, x, gr);
      hld<N>(V3 + o, 1, x); hmat<N, Nq, false>(B, x, gs);
      hld<N>(V4 + o, 1, x); hmat<N, Nq, false>(B, x, gt);
      /* the metric is loaded point by point (six values live at a time instead of 6 Nq): the register budget of
       * the instance, and with it the number of resident warps, is set by this loop */
#pragma unroll
      for (int a = 0; a < Nq; a++) {
        const double G00 = __ldg(Gt + a), G01 = __ldg(Gt + Q3 + a), G02 = __ldg(Gt + 2 * Q3 + a),
                     G11 = __ldg(Gt + 3 * Q3 + a), G12 = __ldg(Gt + 4 * Q3 + a), G22 = __ldg(Gt + 5 * Q3 + a);
        const double g0 = gr[a], g1 = gs[a], g2 = gt[a];
        gr[a] = G00 * g0 + G01 * g1 + G02 * g2;
        gs[a] = G01 * g0 + G11 * g1 + G12 * g2;
        gt[a] = G02 * g0 + G12 * g1 + G22 * g2;
      }
      hmat<Nq, N, true>(C, gr, x); hst<N>(V2 + o, 1, x);
      hmat<Nq, N, true>(B, gs, x); hst<N>(V3 + o, 1, x);
      hmat<Nq, N, true>(B, gt, x); hst<N>(V4 + o, 1, x);
    }
    __syncwarp();
    for (int t = lane; t < NQ; t += 32) { /* transposed s-contraction: lines (i,c) */
      const int i = t % N, c = t / N;
      double x[Nq], y[N];
      hld<Nq>(V2 + i + N * c, BS, x); hmat<Nq, N, true>(B, x, y);
      hld<Nq>(V3 + i + N * c, BS, x); hmat_acc<Nq, N, true>(C, x, y);
      hst<N>(V0 + i + AS * c, N, y);
      hld<Nq>(V4 + i + N * c, BS, x); hmat<Nq, N, true>(B, x, y);
      hst<N>(V1 + i + AS * c, N, y);
    }
    __syncwarp();
    for (int t = lane; t < N2; t += 32) { /* transposed t-contraction: line (i,j) -> ACC */
      const int i = t % N, j = t / N;
      double x[Nq], y[N];
      hld<Nq>(V0 + t, AS, x); hmat<Nq, N, true>(B, x, y);
      hld<Nq>(V1 + t, AS, x); hmat_acc<Nq, N, true>(C, x, y);
      hst<N>(ACC + i + SJ * j, SK, y);
    }
    __syncwarp();
  }

  /* ---------------- faces: one half-mortar record at a time, all 32 lanes.  Contractions are OUTPUT-stationary:
   * a lane owns one row of the 1-D table (own side: in registers, N is a template parameter; neighbour side and
   * the adjoint stages: run-time length, table entries from the L1-resident pool) and sweeps the lines, whose
   * values come from shared memory as broadcasts. */
  const int MX = A.MX, M2 = MX * MX;
  double *RA = S, *RB = S + hp_ra_doubles(A.NMX, MX);
  const d4b_ops_dev &ops = c_hp_ops[A.qt];
  const double *pool = ops.pool;
#pragma unroll 1
  for (int r = 0; r < sl.nrec; r++) {
    const hp_rec R = A.recs[sl.rec_begin + r];
    const int fm = R.code & 7, fp = (R.code >> 3) & 7, tc = (R.code >> 6) & 7;
    const int cm = ((R.code >> 9) & 7) - 1, cp = ((R.code >> 12) & 7) - 1;
    const bool bnd = (R.code >> 15) & 1;
    const int Nm = R.dims & 255, Np = (R.dims >> 8) & 255, Nm2 = Nm * Nm, mdq = Nm - 1;
    const int d = fm >> 1, ia = d == 0 ? 1 : 0, ib = d == 2 ? 1 : 2;
    const int dp = fp >> 1, iap = dp == 0 ? 1 : 0, ibp = dp == 2 ? 1 : 2;
    const int ca = cm < 0 ? 0 : 1 + (cm & 1), cb = cm < 0 ? 0 : 1 + ((cm >> 1) & 1);
    const int cpa = cp < 0 ? 0 : 1 + (cp & 1), cpb = cp < 0 ? 0 : 1 + ((cp >> 1) & 1);
    const double *Ea = pool + ops.idx.I[N - 1][mdq][ca], *EaD = pool + ops.idx.ID[N - 1][mdq][ca];
    const double *Eb = pool + ops.idx.I[N - 1][mdq][cb], *EbD = pool + ops.idx.ID[N - 1][mdq][cb];
    const double *Pa = pool + ops.idx.I[Np - 1][mdq][cpa], *PaD = pool + ops.idx.ID[Np - 1][mdq][cpa];
    const double *Pb = pool + ops.idx.I[Np - 1][mdq][cpb], *PbD = pool + ops.idx.ID[Np - 1][mdq][cpb];
    const double *F = A.fac + R.fac_off;
    const int pl = Nm * MX;                       /* plane stride of the S1 buffers */
    const int qrow = lane % Nm, qsub = lane / Nm; /* lane = (table row, line subset) for the Nm-row stages */
    const int SQ = 32 / Nm;                       /* line subsets (lanes >= SQ * Nm idle in those stages)  */

    /* T: traces at face nodes.  RA[t] = u-, RA[M2 + t] = du-/dn, RA[2 M2 + t'] = u+, RA[3 M2 + t'] = du+/dn */
    {
      int sa, sb, sn; hp_face_strides(d, SJ, SK, sa, sb, sn);
      const int lay = (fm & 1) ? N - 1 : 0;
      const int Np2 = Np * Np, ntask = N2 + (bnd ? 0 : Np2);
      const double *un = bnd ? nullptr : (R.nbr_u >= 0 ? A.u + R.nbr_u : A.u_ghost + (-R.nbr_u - 1));
      const int pa = dp == 0 ? Np : 1, pb = dp == 2 ? Np : Np2, pn = dp == 0 ? 1 : (dp == 1 ? Np : Np2);
      const int layp = (fp & 1) ? Np - 1 : 0;
      const double *Dp = pool + ops.idx.D[Np - 1] + layp * Np;
      for (int t = lane; t < ntask; t += 32) {
        if (t < N2) {
          const int a = t % N, b = t / N;
          double x[N]; hld<N>(U + a * sa + b * sb, sn, x);
          double dn = 0.;
#pragma unroll
          for (int m = 0; m < N; m++) dn = fma(D[lay * N + m], x[m], dn);
          RA[t] = (fm & 1) ? x[N - 1] : x[0];
          RA[M2 + t] = dn;
        } else {
          const int tt = t - N2, b = tt / Np, a = tt - b * Np;
          const double *p = un + a * pa + b * pb;
          double dn = 0., ul = 0.;
#pragma unroll 4
          for (int m = 0; m < Np; m++) {
            const double v = __ldg(p + m * pn);
            dn = fma(__ldg(Dp + m), v, dn);
            if (m == layp) ul = v;
          }
          RA[2 * M2 + tt] = ul; RA[3 * M2 + tt] = dn;
        }
      }
    }
    __syncwarp();
    /* S1: along face axis a.  RB[(3 side + k) pl + q + Nm b], k = {E u, E D u, E dn}; lane = (q, subset of b) */
    if (qsub < SQ) {
      {
        double me[N], md[N];
#pragma unroll
        for (int i = 0; i < N; i++) { me[i] = __ldg(Ea + qrow * N + i); md[i] = __ldg(EaD + qrow * N + i); }
        for (int b = qsub; b < N; b += SQ) {
          double su = 0., sc = 0., sd = 0.;
#pragma unroll
          for (int i = 0; i < N; i++) {
            const double xu = RA[N * b + i], xn = RA[M2 + N * b + i];
            su = fma(me[i], xu, su); sc = fma(md[i], xu, sc); sd = fma(me[i], xn, sd);
          }
          double *o = RB + qrow + Nm * b;
          o[0] = su; o[pl] = sc; o[2 * pl] = sd;
        }
      }
      if (!bnd) {
        const double *pe = Pa + qrow * Np, *pd = PaD + qrow * Np;
        for (int b = qsub; b < Np; b += SQ) {
          const double *xu = RA + 2 * M2 + Np * b, *xn = RA + 3 * M2 + Np * b;
          double su = 0., sc = 0., sd = 0.;
#pragma unroll 4
          for (int i = 0; i < Np; i++) {
            const double e = __ldg(pe + i), ed = __ldg(pd + i), u0 = xu[i], n0 = xn[i];
            su = fma(e, u0, su); sc = fma(ed, u0, sc); sd = fma(e, n0, sd);
          }
          double *o = RB + 3 * pl + qrow + Nm * b;
          o[0] = su; o[pl] = sc; o[2 * pl] = sd;
        }
      }
    }
    __syncwarp();
    /* QV: along face axis b, then the fold with the face factors, in MY point order:
     * RA[q] = u-, RA[Nm2 + q] = sum_l a-_l du-/dr_l, RA[2 Nm2 + q] = u+, RA[3 Nm2 + q] = sum_l a+_l du+/dr_l.
     * lane = (qb, subset of qa) */
    if (qsub < SQ) {
      {
        double me[N], md[N];
#pragma unroll
        for (int i = 0; i < N; i++) { me[i] = __ldg(Eb + qrow * N + i); md[i] = __ldg(EbD + qrow * N + i); }
        for (int qa = qsub; qa < Nm; qa += SQ) {
          double su = 0., sa = 0., sb = 0., sn = 0.;
#pragma unroll
          for (int i = 0; i < N; i++) {
            const double x1 = RB[qa + Nm * i], x2 = RB[pl + qa + Nm * i], x3 = RB[2 * pl + qa + Nm * i];
            su = fma(me[i], x1, su); sa = fma(me[i], x2, sa); sb = fma(md[i], x1, sb); sn = fma(me[i], x3, sn);
          }
          const int q = qa + Nm * qrow;
          RA[q] = su;
          RA[Nm2 + q] = __ldg(F + ia * Nm2 + q) * sa + __ldg(F + ib * Nm2 + q) * sb + __ldg(F + d * Nm2 + q) * sn;
        }
      }
      if (!bnd) {
        const double *pe = Pb + qrow * Np, *pd = PbD + qrow * Np;
        for (int qa = qsub; qa < Nm; qa += SQ) {
          const double *x1 = RB + 3 * pl + qa, *x2 = x1 + pl, *x3 = x2 + pl;
          double su = 0., sa = 0., sb = 0., sn = 0.;
#pragma unroll 4
          for (int i = 0; i < Np; i++) {
            const double e = __ldg(pe + i), ed = __ldg(pd + i), v1 = x1[Nm * i], v2 = x2[Nm * i], v3 = x3[Nm * i];
            su = fma(e, v1, su); sa = fma(e, v2, sa); sb = fma(ed, v1, sb); sn = fma(e, v3, sn);
          }
          /* (qa, qrow) is the neighbour's point; mine is */
          const int xx = (tc & 2) ? Nm - 1 - qa : qa, yy = (tc & 4) ? Nm - 1 - qrow : qrow;
          const int q = (tc & 1) ? xx + Nm * yy : yy + Nm * xx;
          RA[2 * Nm2 + q] = su;
          RA[3 * Nm2 + q] = __ldg(F + (3 + iap) * Nm2 + q) * sa + __ldg(F + (3 + ibp) * Nm2 + q) * sb +
                            __ldg(F + (3 + dp) * Nm2 + q) * sn;
        }
      }
    }
    __syncwarp();
    /* FLUX at the mortar quadrature points (..._flux_sipg.c:645-677 interior, :139-191 Dirichlet):
     * RB[k Nm2 + q], k = 0: tested with v; 1, 2: tested with dv/d(axis a), dv/d(axis b); 3: with dv/dn */
    {
      const double *wq = pool + ops.idx.w[mdq];
      for (int q = lane; q < Nm2; q += 32) {
        const int b = q / Nm, a = q - b * Nm;
        const double w = __ldg(wq + a) * __ldg(wq + b);
        const double am0 = __ldg(F + q), am1 = __ldg(F + Nm2 + q), am2 = __ldg(F + 2 * Nm2 + q);
        const double um = RA[q], qm = RA[Nm2 + q];
        double jump, t1, sjs, hs;
        if (!bnd) {
          jump = um - RA[2 * Nm2 + q]; t1 = -0.5 * (qm + RA[3 * Nm2 + q]); sjs = __ldg(F + 6 * Nm2 + q); hs = -0.5;
        } else {
          const double gv = R.g_off >= 0 ? __ldg(A.bc_g + R.g_off + q) : 0.;
          jump = um - gv; t1 = -qm; sjs = __ldg(F + 3 * Nm2 + q); hs = -1.0;
          if ((R.code >> 16) & 1) { /* Robin: sj (alpha u - beta); rows 0-2 of the block are zero */
            RB[q] = w * (sjs * um - gv); RB[Nm2 + q] = 0.; RB[2 * Nm2 + q] = 0.; RB[3 * Nm2 + q] = 0.;
            continue;
          }
        }
        const double wj = w * hs * jump;
        RB[q] = w * (t1 + sjs * jump);
        RB[Nm2 + q] = wj * (d == 0 ? am1 : am0);
        RB[2 * Nm2 + q] = wj * (d == 2 ? am1 : am2);
        RB[3 * Nm2 + q] = wj * (d == 0 ? am0 : (d == 1 ? am1 : am2));
      }
    }
    __syncwarp();
    /* BACK1: adjoint of QV along b.  RA[k NmN + qa + Nm j], k = 0: Eb^T F + (Eb D)^T Hb, 1: Eb^T Ha, 2: Eb^T Hn;
     * lane = (j, subset of qa) */
    const int NmN = Nm * N;
    constexpr int SN = 32 / N;
    const int jrow = lane % N, jsub = lane / N;
    if (jsub < SN) {
      for (int qa = jsub; qa < Nm; qa += SN) {
        double x1 = 0., ha = 0., hn = 0.;
        const double *f0 = RB + qa;
#pragma unroll 4
        for (int qb = 0; qb < Nm; qb++) {
          const double e = __ldg(Eb + qb * N + jrow), ed = __ldg(EbD + qb * N + jrow);
          const double v0 = f0[Nm * qb], v1 = f0[Nm2 + Nm * qb], v2 = f0[2 * Nm2 + Nm * qb], v3 = f0[3 * Nm2 + Nm * qb];
          x1 = fma(e, v0, x1); x1 = fma(ed, v2, x1); ha = fma(e, v1, ha); hn = fma(e, v3, hn);
        }
        double *o = RA + qa + Nm * jrow;
        o[0] = x1; o[NmN] = ha; o[2 * NmN] = hn;
      }
    }
    __syncwarp();
    /* BACK2: adjoint of S1 along a, accumulated straight into the per-face Y (lift onto the face layer) and
     * Z (normal-derivative lift); lane = (a', subset of b') */
    if (jsub < SN) {
      double *yz = YZ + fm * 2 * N2;
      for (int bp = jsub; bp < N; bp += SN) {
        double y = 0., z = 0.;
        const double *x1 = RA + Nm * bp;
#pragma unroll 4
        for (int qa = 0; qa < Nm; qa++) {
          const double e = __ldg(Ea + qa * N + jrow), ed = __ldg(EaD + qa * N + jrow);
          y = fma(e, x1[qa], y); y = fma(ed, x1[NmN + qa], y); z = fma(e, x1[2 * NmN + qa], z);
        }
        yz[jrow + N * bp] += y; yz[N2 + jrow + N * bp] += z;
      }
    }
    __syncwarp();
  }


  /* ---------------- LIFT (..._flux_sipg.c:918-958): face layer += Y, normal line += D[layer,:] Z; then store */
#pragma unroll 1
  for (int d = 0; d < 3; d++) {
    int sa, sb, sn; hp_face_strides(d, SJ, SK, sa, sb, sn);
    for (int t = lane; t < N2; t += 32) {
      double *acc = ACC + (t % N) * sa + (t / N) * sb;
      const double y0 = YZ[(4 * d) * N2 + t], z0 = YZ[(4 * d + 1) * N2 + t];
      const double y1 = YZ[(4 * d + 2) * N2 + t], z1 = YZ[(4 * d + 3) * N2 + t];
#pragma unroll
      for (int m = 0; m < N; m++) {
        double v = acc[m * sn] + D[m] * z0 + D[(N - 1) * N + m] * z1;
        if (m == 0) v += y0;
        if (m == N - 1) v += y1;
        acc[m * sn] = v;
      }
    }
    __syncwarp();
  }
  double *Aue = A.Au + sl.node_off;
  for (int i = lane; i < N3; i += 32) Aue[i] = ACC[(i % N) + SJ * ((i / N) % N) + SK * (i / N2)];
}

/* ====================================================================================================
 * Two-pass ("twin trace") form of the same operator for hp elements whose neighbours are all local.
 * In the one-side-at-a-time formulation every half-mortar record evaluates BOTH sides of its mortar: its own
 * trace (u-, sum_l a-_l du-/dr_l at the mortar points) and the neighbour's -- which is exactly the own-side trace
 * of the neighbour's TWIN record (same mortar, roles swapped), with a+ = -a-(twin) at the same physical points
 * (checked on every curved hp fixture: 2e-15 of the block maximum, hanging and mixed-p faces included).  So:
 *   pass 1 (d4b_tp_trace_kernel): per element, volume term -> Au, and per record T/S1/QV of the OWN side only ->
 *           W[record] = (u-, q-) at the mortar points (2 Nm^2 doubles, own point order);
 *   pass 2 (d4b_tp_flux_kernel): per element and record, u+ = W[twin].u (reordered by tcode), q+ = -W[twin].q ->
 *           FLUX -> BACK1/BACK2 -> Y, Z; LIFT onto the volume result; Au.
 * The neighbour-side stages (run-time Np loops, the neighbour's whole element read from global memory per record)
 * disappear; what is added is 2 Nm^2 doubles written and 4 Nm^2 read per record.  Elements with a ghost neighbour
 * keep the single-pass kernel above (the twin's trace lives on another rank).  Pass 1 also runs, trace only, over
 * the local neighbours of two-pass elements that are handled by other kernels. */
#ifndef TP_TRACE_MINB
#define TP_TRACE_MINB(N) ((N) <= 4 ? 8 : ((N) <= 6 ? 5 : 4))
#endif
#ifndef TP_FLUX_MINB
#define TP_FLUX_MINB(N) ((N) <= 4 ? 8 : ((N) <= 6 ? 6 : 4))
#endif
struct tp_rec {        /* 48 bytes */
  int64_t fac_off, g_off, w_own, w_twin;   /* w_twin < 0: boundary */
  int32_t code;        /* bits 0-2 f_m, 6-8 tcode, 9-11 child_m+1, 15 boundary, 16 Robin */
  int32_t dims;        /* Nm */
  int64_t pad;
};
struct tp_slot { int64_t node_off, quad_off; int32_t rec_begin, nrec, flags, pad; };   /* flags bit 0: two-pass element */
struct tp_args {
  const tp_slot *slots; const tp_rec *recs; const double *G, *fac, *bc_g, *u; double *Au, *W;
  int n_list, qt, NMX, per_warp;
};
__host__ __device__ static inline int tp1_scratch(int N, int NMX) { return 2 * N * N + 3 * NMX * N; }
__host__ __device__ static inline int tp2_scratch(int N, int NMX) { return 3 * NMX * N + 4 * NMX * NMX; }

/* own-side traces of one record; NMT = number of mortar points per direction when known at compile time (0: run time) */
template <int N, int Nq, int NMT>
__device__ __forceinline__ void tp_trace_record(const tp_args &A, const tp_rec &R, const double *U, double *RA, double *RB,
                                                const double *D, int lane)
{
  using L = HL<N, Nq>;
  constexpr int N2 = L::N2, SJ = L::SJ, SK = L::SK;
  const d4b_ops_dev &ops = c_hp_ops[A.qt];
  const double *pool = ops.pool;
  const int fm = R.code & 7, cm = ((R.code >> 9) & 7) - 1;
  const int Nm = NMT ? NMT : R.dims, Nm2 = Nm * Nm, mdq = Nm - 1;
  const int d = fm >> 1, ia = d == 0 ? 1 : 0, ib = d == 2 ? 1 : 2;
  const int ca = cm < 0 ? 0 : 1 + (cm & 1), cb = cm < 0 ? 0 : 1 + ((cm >> 1) & 1);
  const double *Ea = pool + ops.idx.I[N - 1][mdq][ca], *EaD = pool + ops.idx.ID[N - 1][mdq][ca];
  const double *Eb = pool + ops.idx.I[N - 1][mdq][cb], *EbD = pool + ops.idx.ID[N - 1][mdq][cb];
  const double *F = A.fac + R.fac_off;
  const int pl = Nm * N;
  const int qrow = lane % Nm, qsub = lane / Nm, SQ = 32 / Nm;
  {
    int sa, sb, sn; hp_face_strides(d, SJ, SK, sa, sb, sn);
    const int lay = (fm & 1) ? N - 1 : 0;
    for (int t = lane; t < N2; t += 32) {
      const int a = t % N, b = t / N;
      double x[N]; hld<N>(U + a * sa + b * sb, sn, x);
      double dn = 0.;
#pragma unroll
      for (int m = 0; m < N; m++) dn = fma(D[lay * N + m], x[m], dn);
      RA[t] = (fm & 1) ? x[N - 1] : x[0];
      RA[N2 + t] = dn;
    }
  }
  __syncwarp();
  if (qsub < SQ) {
    double me[N], md[N];
#pragma unroll
    for (int i = 0; i < N; i++) { me[i] = __ldg(Ea + qrow * N + i); md[i] = __ldg(EaD + qrow * N + i); }
    for (int b = qsub; b < N; b += SQ) {
      double su = 0., sc = 0., sd = 0.;
#pragma unroll
      for (int i = 0; i < N; i++) {
        const double xu = RA[N * b + i], xn = RA[N2 + N * b + i];
        su = fma(me[i], xu, su); sc = fma(md[i], xu, sc); sd = fma(me[i], xn, sd);
      }
      double *o = RB + qrow + Nm * b;
      o[0] = su; o[pl] = sc; o[2 * pl] = sd;
    }
  }
  __syncwarp();
  if (qsub < SQ) {
    double me[N], md[N];
#pragma unroll
    for (int i = 0; i < N; i++) { me[i] = __ldg(Eb + qrow * N + i); md[i] = __ldg(EbD + qrow * N + i); }
    double *Wr = A.W + R.w_own;
    for (int qa = qsub; qa < Nm; qa += SQ) {
      const int q = qa + Nm * qrow;
      const double fa = __ldg(F + ia * Nm2 + q), fb = __ldg(F + ib * Nm2 + q), fn = __ldg(F + d * Nm2 + q);
      double su = 0., sa = 0., sb = 0., sn = 0.;
#pragma unroll
      for (int i = 0; i < N; i++) {
        const double x1 = RB[qa + Nm * i], x2 = RB[pl + qa + Nm * i], x3 = RB[2 * pl + qa + Nm * i];
        su = fma(me[i], x1, su); sa = fma(me[i], x2, sa); sb = fma(md[i], x1, sb); sn = fma(me[i], x3, sn);
      }
      Wr[q] = su;
      Wr[Nm2 + q] = fa * sa + fb * sb + fn * sn;
    }
  }
  __syncwarp();
}

template <int N, int Nq>
__global__ void __launch_bounds__(128, TP_TRACE_MINB(N)) d4b_tp_trace_kernel(const tp_args A)
{
  using L = HL<N, Nq>;
  constexpr int N2 = L::N2, N3 = L::N3, Q2 = L::Q2, Q3 = L::Q3, NQ = L::NQ;
  constexpr int SJ = L::SJ, SK = L::SK, AS = L::AS, BS = L::BS;
  extern __shared__ double sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = blockIdx.x * (blockDim.x >> 5) + warp;
  if (slot >= A.n_list) return;
  const hp_tab<N, Nq> &T = hp_const<N, Nq>::get();
  const double *B = T.B, *C = T.C, *D = T.D;
  double *U = sm + (size_t)warp * A.per_warp, *S = U + L::UA;
  const tp_slot sl = A.slots[slot];
  const double *ue = A.u + sl.node_off;
  if (sl.flags & 1) {
    const char *g = reinterpret_cast<const char *>(A.G + 6 * sl.quad_off);
    for (int o = lane * 128; o < 6 * Q3 * 8; o += 32 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + o));
  }
  for (int r = lane >> 2; r < sl.nrec; r += 8) {
    const tp_rec *h = A.recs + sl.rec_begin + r;
    const int nm = __ldg(&h->dims), bytes = 3 * nm * nm * 8;
    const char *p = reinterpret_cast<const char *>(A.fac + __ldg(&h->fac_off));
    for (int o = (lane & 3) * 128; o < bytes; o += 4 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
  }
  for (int i = lane; i < N3; i += 32) U[(i % N) + SJ * ((i / N) % N) + SK * (i / N2)] = ue[i];
  __syncwarp();
  /* ---------------- volume term of two-pass elements, straight to Au (same code as the single-pass kernel) */
  if (sl.flags & 1) {
    double *V0 = S, *V1 = S + L::AB, *V2 = S + 2 * L::AB, *V3 = V2 + L::BB, *V4 = V3 + L::BB;
    for (int t = lane; t < N2; t += 32) {
      const int i = t % N, j = t / N;
      double x[N], y[Nq];
      hld<N>(U + i + SJ * j, SK, x);
      hmat<N, Nq, false>(B, x, y); hst<Nq>(V0 + t, AS, y);
      hmat<N, Nq, false>(C, x, y); hst<Nq>(V1 + t, AS, y);
    }
    __syncwarp();
    for (int t = lane; t < NQ; t += 32) {
      const int i = t % N, c = t / N;
      double x[N], y[Nq];
      hld<N>(V0 + i + AS * c, N, x);
      hmat<N, Nq, false>(B, x, y); hst<Nq>(V2 + i + N * c, BS, y);
      hmat<N, Nq, false>(C, x, y); hst<Nq>(V3 + i + N * c, BS, y);
      hld<N>(V1 + i + AS * c, N, x);
      hmat<N, Nq, false>(B, x, y); hst<Nq>(V4 + i + N * c, BS, y);
    }
    __syncwarp();
    const double *Ge = A.G + 6 * sl.quad_off;
    for (int t = lane; t < Q2; t += 32) {
      const int b = t % Nq, c = t / Nq;
      double x[N], gr[Nq], gs[Nq], gt[Nq];
      const int o = BS * b + N * c;
      const double *Gt = Ge + Nq * t;
      hld<N>(V2 + o, 1, x); hmat<N, Nq, false>(C, x, gr);
      hld<N>(V3 + o, 1, x); hmat<N, Nq, false>(B, x, gs);
      hld<N>(V4 + o, 1, x); hmat<N, Nq, false>(B, x, gt);
#pragma unroll
      for (int a = 0; a < Nq; a++) {
        const double G00 = __ldg(Gt + a), G01 = __ldg(Gt + Q3 + a), G02 = __ldg(Gt + 2 * Q3 + a),
                     G11 = __ldg(Gt + 3 * Q3 + a), G12 = __ldg(Gt + 4 * Q3 + a), G22 = __ldg(Gt + 5 * Q3 + a);
        const double g0 = gr[a], g1 = gs[a], g2 = gt[a];
        gr[a] = G00 * g0 + G01 * g1 + G02 * g2;
        gs[a] = G01 * g0 + G11 * g1 + G12 * g2;
        gt[a] = G02 * g0 + G12 * g1 + G22 * g2;
      }
      hmat<Nq, N, true>(C, gr, x); hst<N>(V2 + o, 1, x);
      hmat<Nq, N, true>(B, gs, x); hst<N>(V3 + o, 1, x);
      hmat<Nq, N, true>(B, gt, x); hst<N>(V4 + o, 1, x);
    }
    __syncwarp();
    for (int t = lane; t < NQ; t += 32) {
      const int i = t % N, c = t / N;
      double x[Nq], y[N];
      hld<Nq>(V2 + i + N * c, BS, x); hmat<Nq, N, true>(B, x, y);
      hld<Nq>(V3 + i + N * c, BS, x); hmat_acc<Nq, N, true>(C, x, y);
      hst<N>(V0 + i + AS * c, N, y);
      hld<Nq>(V4 + i + N * c, BS, x); hmat<Nq, N, true>(B, x, y);
      hst<N>(V1 + i + AS * c, N, y);
    }
    __syncwarp();
    double *Aue = A.Au + sl.node_off;
    for (int t = lane; t < N2; t += 32) {   /* transposed t-contraction: line (i,j) -> Au[t + N2 k] */
      double x[Nq], y[N];
      hld<Nq>(V0 + t, AS, x); hmat<Nq, N, true>(B, x, y);
      hld<Nq>(V1 + t, AS, x); hmat_acc<Nq, N, true>(C, x, y);
      hst<N>(Aue + t, N2, y);
    }
    __syncwarp();
  }

  /* ---------------- own-side traces of the records: T -> S1 -> QV -> W */
  double *RA = S, *RB = S + 2 * N2;
  tp_rec R = A.recs[sl.rec_begin];
#pragma unroll 1
  for (int r = 0; r < sl.nrec; r++) {
    tp_rec Rn = R;   /* the next record is requested now and used in the next iteration */
    if (r + 1 < sl.nrec) Rn = A.recs[sl.rec_begin + r + 1];
    switch (R.dims - Nq) {
      case 0: tp_trace_record<N, Nq, Nq>(A, R, U, RA, RB, D, lane); break;
      case 1: tp_trace_record<N, Nq, Nq + 1>(A, R, U, RA, RB, D, lane); break;
      case 2: tp_trace_record<N, Nq, Nq + 2>(A, R, U, RA, RB, D, lane); break;
      default: tp_trace_record<N, Nq, 0>(A, R, U, RA, RB, D, lane); break;
    }
    R = Rn;
  }
}

/* flux, Galerkin integral and back-projection of one record into the per-face accumulators Y, Z */
template <int N, int Nq, int NMT>
__device__ __forceinline__ void tp_flux_record(const tp_args &A, const tp_rec &R, double *YZ, double *RA, double *RB, int lane)
{
  using L = HL<N, Nq>;
  constexpr int N2 = L::N2;
  const d4b_ops_dev &ops = c_hp_ops[A.qt];
  const double *pool = ops.pool;
  const int fm = R.code & 7, tc = (R.code >> 6) & 7, cm = ((R.code >> 9) & 7) - 1;
  const bool bnd = (R.code >> 15) & 1;
  const int Nm = NMT ? NMT : R.dims, Nm2 = Nm * Nm, mdq = Nm - 1;
  const int d = fm >> 1;
  const int ca = cm < 0 ? 0 : 1 + (cm & 1), cb = cm < 0 ? 0 : 1 + ((cm >> 1) & 1);
  const double *Ea = pool + ops.idx.I[N - 1][mdq][ca], *EaD = pool + ops.idx.ID[N - 1][mdq][ca];
  const double *Eb = pool + ops.idx.I[N - 1][mdq][cb], *EbD = pool + ops.idx.ID[N - 1][mdq][cb];
  const double *F = A.fac + R.fac_off;
  const double *Wo = A.W + R.w_own, *Wt = A.W + (bnd ? R.w_own : R.w_twin);
  /* FLUX at the mortar quadrature points (..._flux_sipg.c:645-677 interior, :139-191 Dirichlet, :377-476 Robin) */
  {
    const double *wq = pool + ops.idx.w[mdq];
    for (int q = lane; q < Nm2; q += 32) {
      const int b = q / Nm, a = q - b * Nm;
      const double w = __ldg(wq + a) * __ldg(wq + b);
      const double am0 = __ldg(F + q), am1 = __ldg(F + Nm2 + q), am2 = __ldg(F + 2 * Nm2 + q);
      const double um = __ldg(Wo + q), qm = __ldg(Wo + Nm2 + q);
      double jump, t1, sjs, hs;
      if (!bnd) {
        const int qp = hp_tcode_src(tc, Nm, a, b);   /* my point in the twin's point order */
        jump = um - __ldg(Wt + qp); t1 = -0.5 * (qm - __ldg(Wt + Nm2 + qp)); sjs = __ldg(F + 6 * Nm2 + q); hs = -0.5;
      } else {
        const double gv = R.g_off >= 0 ? __ldg(A.bc_g + R.g_off + q) : 0.;
        jump = um - gv; t1 = -qm; sjs = __ldg(F + 3 * Nm2 + q); hs = -1.0;
        if ((R.code >> 16) & 1) {
          RB[q] = w * (sjs * um - gv); RB[Nm2 + q] = 0.; RB[2 * Nm2 + q] = 0.; RB[3 * Nm2 + q] = 0.;
          continue;
        }
      }
      const double wj = w * hs * jump;
      RB[q] = w * (t1 + sjs * jump);
      RB[Nm2 + q] = wj * (d == 0 ? am1 : am0);
      RB[2 * Nm2 + q] = wj * (d == 2 ? am1 : am2);
      RB[3 * Nm2 + q] = wj * (d == 0 ? am0 : (d == 1 ? am1 : am2));
    }
  }
  __syncwarp();
  const int NmN = Nm * N;
  constexpr int SN = 32 / N;
  const int jrow = lane % N, jsub = lane / N;
  if (jsub < SN) {
    for (int qa = jsub; qa < Nm; qa += SN) {
      double x1 = 0., ha = 0., hn = 0.;
      const double *f0 = RB + qa;
#pragma unroll (NMT ? NMT : 4)
      for (int qb = 0; qb < Nm; qb++) {
        const double e = __ldg(Eb + qb * N + jrow), ed = __ldg(EbD + qb * N + jrow);
        const double v0 = f0[Nm * qb], v1 = f0[Nm2 + Nm * qb], v2 = f0[2 * Nm2 + Nm * qb], v3 = f0[3 * Nm2 + Nm * qb];
        x1 = fma(e, v0, x1); x1 = fma(ed, v2, x1); ha = fma(e, v1, ha); hn = fma(e, v3, hn);
      }
      double *o = RA + qa + Nm * jrow;
      o[0] = x1; o[NmN] = ha; o[2 * NmN] = hn;
    }
  }
  __syncwarp();
  if (jsub < SN) {
    double *yz = YZ + fm * 2 * N2;
    for (int bp = jsub; bp < N; bp += SN) {
      double y = 0., z = 0.;
      const double *x1 = RA + Nm * bp;
#pragma unroll (NMT ? NMT : 4)
      for (int qa = 0; qa < Nm; qa++) {
        const double e = __ldg(Ea + qa * N + jrow), ed = __ldg(EaD + qa * N + jrow);
        y = fma(e, x1[qa], y); y = fma(ed, x1[NmN + qa], y); z = fma(e, x1[2 * NmN + qa], z);
      }
      yz[jrow + N * bp] += y; yz[N2 + jrow + N * bp] += z;
    }
  }
  __syncwarp();
}

template <int N, int Nq>
__global__ void __launch_bounds__(128, TP_FLUX_MINB(N)) d4b_tp_flux_kernel(const tp_args A)
{
  using L = HL<N, Nq>;
  constexpr int N2 = L::N2, N3 = L::N3, SJ = L::SJ, SK = L::SK;
  extern __shared__ double sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = blockIdx.x * (blockDim.x >> 5) + warp;
  if (slot >= A.n_list) return;
  const hp_tab<N, Nq> &T = hp_const<N, Nq>::get();
  const double *D = T.D;
  double *ACC = sm + (size_t)warp * A.per_warp, *YZ = ACC + L::UA, *RA = YZ + 12 * N2, *RB = RA + 3 * A.NMX * N;
  const tp_slot sl = A.slots[slot];
  for (int i = lane; i < 12 * N2; i += 32) YZ[i] = 0.;
  /* the volume result that pass 1 left in Au, requested now, needed at the LIFT */
  double *Aue = A.Au + sl.node_off;
  for (int i = lane; i < N3; i += 32) {   /* asynchronous: nothing waits for it before the LIFT */
    const unsigned sa = (unsigned)__cvta_generic_to_shared(ACC + (i % N) + SJ * ((i / N) % N) + SK * (i / N2));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(Aue + i) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  tp_rec R = A.recs[sl.rec_begin];
#pragma unroll 1
  for (int r = 0; r < sl.nrec; r++) {
    tp_rec Rn = R;   /* the next record is requested now and used in the next iteration */
    if (r + 1 < sl.nrec) Rn = A.recs[sl.rec_begin + r + 1];
    switch (R.dims - Nq) {
      case 0: tp_flux_record<N, Nq, Nq>(A, R, YZ, RA, RB, lane); break;
      case 1: tp_flux_record<N, Nq, Nq + 1>(A, R, YZ, RA, RB, lane); break;
      case 2: tp_flux_record<N, Nq, Nq + 2>(A, R, YZ, RA, RB, lane); break;
      default: tp_flux_record<N, Nq, 0>(A, R, YZ, RA, RB, lane); break;
    }
    R = Rn;
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncwarp();
  /* LIFT onto the volume result that pass 1 left in Au */
#pragma unroll 1
  for (int d = 0; d < 3; d++) {
    int sa, sb, sn; hp_face_strides(d, SJ, SK, sa, sb, sn);
    for (int t = lane; t < N2; t += 32) {
      double *acc = ACC + (t % N) * sa + (t / N) * sb;
      const double y0 = YZ[(4 * d) * N2 + t], z0 = YZ[(4 * d + 1) * N2 + t];
      const double y1 = YZ[(4 * d + 2) * N2 + t], z1 = YZ[(4 * d + 3) * N2 + t];
#pragma unroll
      for (int m = 0; m < N; m++) {
        double v = acc[m * sn] + D[m] * z0 + D[(N - 1) * N + m] * z1;
        if (m == 0) v += y0;
        if (m == N - 1) v += y1;
        acc[m * sn] = v;
      }
    }
    __syncwarp();
  }
  for (int i = lane; i < N3; i += 32) Aue[i] = ACC[(i % N) + SJ * ((i / N) % N) + SK * (i / N2)];
}



/* ------------------------------------------------------------------ host side */
struct hp_bucket_host { std::vector<hp_slot> slots; std::vector<hp_rec> recs; int NMX = 0, MX = 0; };

bool d4b_hp_supported(int N, int Nq)
{
#define X(a, b, g) if (N == a && Nq == b) return true;
  HP_PAIRS(X)
#undef X
  return false;
}

template <int N, int Nq> static size_t hp_warp_doubles(int NMX, int MX)
{
  using L = HL<N, Nq>;
  int scr = hp_group_doubles(NMX, MX);
  if (scr < L::VOL) scr = L::VOL;
  return (size_t)L::FIXED + scr;
}
static size_t hp_warp_doubles_rt(int N, int Nq, int NMX, int MX)
{
#define X(a, b, g) if (N == a && Nq == b) return hp_warp_doubles<a, b>(NMX, MX);
  HP_PAIRS(X)
#undef X
  return (size_t)1 << 40;
}

int d4b_opt_tp_min();
struct tp_bucket_host { std::vector<tp_slot> slots; std::vector<tp_rec> recs; int NMX = 0; };

static int tp_upload(tp_bucket_host *hb, d4b_tp_bucket *out, int *n_out)
{
  *n_out = 0;
  for (int k = 0; k < D4B_NDEG * D4B_NDEGQ; k++) {
    tp_bucket_host &b = hb[k];
    if (b.slots.empty()) continue;
    if (*n_out == D4B_MAX_BUCKETS) { d4b_set_error("more than %d (deg, deg_quad) pairs on one level", D4B_MAX_BUCKETS); return D4B_EINVAL; }
    d4b_tp_bucket &o = out[(*n_out)++];
    o.N = k / D4B_NDEGQ + 1; o.Nq = k % D4B_NDEGQ + 1; o.count = (int)b.slots.size(); o.NMX = b.NMX; o.n_rec = (int)b.recs.size();
    D4B_CUDA(cudaMalloc(&o.slots, b.slots.size() * sizeof(tp_slot)));
    D4B_CUDA(cudaMemcpy(o.slots, b.slots.data(), b.slots.size() * sizeof(tp_slot), cudaMemcpyHostToDevice));
    D4B_CUDA(cudaMalloc(&o.recs, (b.recs.size() ? b.recs.size() : 1) * sizeof(tp_rec)));
    if (!b.recs.empty()) D4B_CUDA(cudaMemcpy(o.recs, b.recs.data(), b.recs.size() * sizeof(tp_rec), cudaMemcpyHostToDevice));
  }
  return D4B_OK;
}

/* the record of the neighbour that shares the mortar of record r of element e (roles swapped), or -1 */
static int tp_find_twin(const d4b_mesh_desc_t *D, int e, int r)
{
  const d4b_hm_t &h = D->hm[r];
  const int p = h.nbr;
  for (int q = D->hm_begin[p]; q < D->hm_begin[p + 1]; q++) {
    const d4b_hm_t &g = D->hm[q];
    if (g.nbr == e && g.f_m == h.f_p && g.f_p == h.f_m && g.child_m == h.child_p && g.child_p == h.child_m && g.degq == h.degq) return q;
  }
  return -1;
}

/* Take the elements of `cand` the templated hp kernels can run (instance exists, neighbour N <= HP_MAXNP,
 * mortar points <= HP_MAXNM, one warp's scratch fits) into (N, Nq) buckets with resolved per-record
 * metadata; the others are appended to `rest` (completeness kernel).  Elements whose neighbours are all local
 * (and have a trace instance) take the two-pass path (measured faster down to 300-element levels; the option
 * two_pass_min_elements can demand a minimum count); the rest (ghost neighbours) the single-pass kernel. */
int d4b_hp_build(d4b_mesh *m, const d4b_mesh_desc_t *D, const std::vector<int32_t> &cand, std::vector<int32_t> &rest)
{
  m->n_hp = 0; m->n_tp1 = m->n_tp2 = 0; m->tpW = nullptr; m->tpW_doubles = 0;
  if (!D->hm) { rest.insert(rest.end(), cand.begin(), cand.end()); return D4B_OK; }
  std::vector<hp_bucket_host> hb(D4B_NDEG * D4B_NDEGQ);
  const int nl = m->n_local;
  struct elig { int32_t e; int nmx, mx; };
  std::vector<elig> el;
  for (int32_t e : cand) {
    const int N = D->deg[e] + 1, Nq = D->degq[e] + 1;
    bool ok = d4b_hp_supported(N, Nq);
    int nmx = Nq, mx = N;
    for (int r = D->hm_begin[e]; ok && r < D->hm_begin[e + 1]; r++) {
      const d4b_hm_t &h = D->hm[r];
      const int Nm = h.degq + 1;
      if (Nm > HP_MAXNM || Nm < N) ok = false;
      if (h.nbr >= 0) { const int Np = D->deg[h.nbr] + 1; if (Np > HP_MAXNP || Np > Nm) ok = false; if (Np > mx) mx = Np; }
      if (Nm > nmx) nmx = Nm;
    }
    if (ok && hp_warp_doubles_rt(N, Nq, nmx, mx) * 8 > 200 * 1024) ok = false;
    if (!ok) { rest.push_back(e); continue; }
    el.push_back({e, nmx, mx});
  }
  /* ---- which of them take the two-pass path */
  /* option two_pass_min_elements (D4B_TP_MIN; D4B_NO_TP = never).  Default (-1): always on a level without neighbour
   * ranks (measured faster down to 300 elements: 61 against 79 us per apply); with neighbour ranks only from 4 096
   * two-pass elements up -- there the elements along the rank boundary stay single-pass and small levels pay for
   * the extra launches (MG-FCG on the 50 k-element hp mesh: 6.15 / 5.01 s at 2 / 4 GPUs, 6.48 / 5.36 s with 0) */
  int tp_min = d4b_opt_tp_min();
  if (tp_min < 0) tp_min = D->n_peers > 0 ? 4096 : 0;
  const bool no_tp = tp_min == 0x7fffffff;
  std::vector<char> is_tp(nl, 0);
  std::vector<int32_t> twin(D->hm_begin[nl], -1);
  int n_tp = 0;
  if (!no_tp) {
    for (const elig &x : el) {
      const int e = x.e;
      bool ok = true;
      for (int r = D->hm_begin[e]; ok && r < D->hm_begin[e + 1]; r++) {
        const d4b_hm_t &h = D->hm[r];
        if (h.nbr < 0) continue;
        if (h.nbr >= nl || !d4b_hp_supported(D->deg[h.nbr] + 1, D->degq[h.nbr] + 1)) { ok = false; break; }
        if ((twin[r] = tp_find_twin(D, e, r)) < 0) ok = false;
      }
      if (ok) { is_tp[e] = 1; n_tp++; }
    }
    if (n_tp < tp_min) { std::fill(is_tp.begin(), is_tp.end(), 0); n_tp = 0; }
  }
  /* ---- single-pass buckets */
  for (const elig &x : el) {
    const int e = x.e;
    if (is_tp[e]) continue;
    const int N = D->deg[e] + 1, Nq = D->degq[e] + 1, nmx = x.nmx, mx = x.mx;
    hp_bucket_host &b = hb[(N - 1) * D4B_NDEGQ + (Nq - 1)];
    hp_slot s; s.node_off = D->node_off[e]; s.quad_off = D->quad_off[e];
    s.rec_begin = (int32_t)b.recs.size(); s.nrec = D->hm_begin[e + 1] - D->hm_begin[e]; s.pad = 0;
    for (int r = D->hm_begin[e]; r < D->hm_begin[e + 1]; r++) {
      const d4b_hm_t &h = D->hm[r];
      hp_rec R; R.fac_off = h.fac_off; R.g_off = h.g_off; R.nbr_u = 0;
      const int bnd = h.nbr < 0;
      int Np = N;
      if (!bnd) {
        Np = D->deg[h.nbr] + 1;
        R.nbr_u = h.nbr < nl ? D->node_off[h.nbr] : -(D->node_off[h.nbr] - D->node_off[nl]) - 1;
      }
      R.code = (h.f_m & 7) | (((bnd ? 0 : h.f_p) & 7) << 3) | ((h.tcode & 7) << 6) | ((h.child_m + 1) << 9) |
               (((bnd ? -1 : h.child_p) + 1) << 12) | (bnd << 15) | ((bnd && h.pad == 1) << 16);
      R.dims = (h.degq + 1) | (Np << 8);
      b.recs.push_back(R);
    }
    b.slots.push_back(s);
    if (nmx > b.NMX) b.NMX = nmx;
    if (mx > b.MX) b.MX = mx;
  }
  for (int k = 0; k < D4B_NDEG * D4B_NDEGQ; k++) {
    hp_bucket_host &b = hb[k];
    if (b.slots.empty()) continue;
    if (m->n_hp == D4B_MAX_BUCKETS) { d4b_set_error("more than %d (deg, deg_quad) pairs on one level", D4B_MAX_BUCKETS); return D4B_EINVAL; }
    d4b_hp_bucket &o = m->hp[m->n_hp++];
    o.N = k / D4B_NDEGQ + 1; o.Nq = k % D4B_NDEGQ + 1; o.count = (int)b.slots.size(); o.NMX = b.NMX; o.MX = b.MX;
    o.n_rec = (int)b.recs.size();
    D4B_CUDA(cudaMalloc(&o.slots, b.slots.size() * sizeof(hp_slot)));
    D4B_CUDA(cudaMemcpy(o.slots, b.slots.data(), b.slots.size() * sizeof(hp_slot), cudaMemcpyHostToDevice));
    D4B_CUDA(cudaMalloc(&o.recs, b.recs.size() * sizeof(hp_rec)));
    D4B_CUDA(cudaMemcpy(o.recs, b.recs.data(), b.recs.size() * sizeof(hp_rec), cudaMemcpyHostToDevice));
  }
  if (!n_tp) return D4B_OK;
  /* ---- two-pass tables.  W slots: every record of a two-pass element + the twins of its interior records */
  std::vector<int64_t> w_of(D->hm_begin[nl], -1);
  int64_t wn = 0;
  auto want = [&](int r) { if (w_of[r] < 0) { const int nm = D->hm[r].degq + 1; w_of[r] = wn; wn += 2 * nm * nm; } };
  for (int e = 0; e < nl; e++) {
    if (!is_tp[e]) continue;
    for (int r = D->hm_begin[e]; r < D->hm_begin[e + 1]; r++) { want(r); if (twin[r] >= 0) want(twin[r]); }
  }
  std::vector<tp_bucket_host> b1(D4B_NDEG * D4B_NDEGQ), b2(D4B_NDEG * D4B_NDEGQ);
  for (int e = 0; e < nl; e++) {
    int nw = 0;
    for (int r = D->hm_begin[e]; r < D->hm_begin[e + 1]; r++) nw += w_of[r] >= 0;
    if (!nw) continue;
    const int N = D->deg[e] + 1, Nq = D->degq[e] + 1;
    tp_bucket_host &t1 = b1[(N - 1) * D4B_NDEGQ + (Nq - 1)];
    tp_slot s1; s1.node_off = D->node_off[e]; s1.quad_off = D->quad_off[e]; s1.rec_begin = (int32_t)t1.recs.size(); s1.nrec = nw;
    s1.flags = is_tp[e]; s1.pad = 0;
    tp_bucket_host *t2 = is_tp[e] ? &b2[(N - 1) * D4B_NDEGQ + (Nq - 1)] : nullptr;
    tp_slot s2 = s1;
    if (t2) { s2.rec_begin = (int32_t)t2->recs.size(); s2.nrec = D->hm_begin[e + 1] - D->hm_begin[e]; }
    if (Nq > t1.NMX) t1.NMX = Nq;
    for (int r = D->hm_begin[e]; r < D->hm_begin[e + 1]; r++) {
      if (w_of[r] < 0) continue;
      const d4b_hm_t &h = D->hm[r];
      const int bnd = h.nbr < 0, Nm = h.degq + 1;
      tp_rec R; R.fac_off = h.fac_off; R.g_off = h.g_off; R.w_own = w_of[r]; R.w_twin = bnd ? -1 : (twin[r] >= 0 ? w_of[twin[r]] : -1);
      R.code = (h.f_m & 7) | ((h.tcode & 7) << 6) | ((h.child_m + 1) << 9) | (bnd << 15) | ((bnd && h.pad == 1) << 16);
      R.dims = Nm; R.pad = 0;
      if (Nm < N || Nm > HP_MAXNM) { d4b_set_error("two-pass hp path: mortar of %d points on an element with N = %d", Nm, N); return D4B_EINVAL; }
      t1.recs.push_back(R);
      if (Nm > t1.NMX) t1.NMX = Nm;
      if (t2) { t2->recs.push_back(R); if (Nm > t2->NMX) t2->NMX = Nm; }
    }
    t1.slots.push_back(s1);
    if (t2) t2->slots.push_back(s2);
  }
  int rc;
  if ((rc = tp_upload(b1.data(), m->tp1, &m->n_tp1)) != D4B_OK) return rc;
  if ((rc = tp_upload(b2.data(), m->tp2, &m->n_tp2)) != D4B_OK) return rc;
  D4B_CUDA(cudaMalloc((void **)&m->tpW, (size_t)(wn ? wn : 1) * sizeof(double)));
  m->tpW_doubles = wn;
  return D4B_OK;
}

static int hp_bind_ops(int qt)
{
  static bool bound[2] = {false, false};
  if (bound[qt]) return D4B_OK;
  const d4b_ops_dev *ops = d4b_ops_device(qt);
  if (!ops) return D4B_ECUDA;
  D4B_CUDA(cudaMemcpyToSymbol(c_hp_ops, ops, sizeof(d4b_ops_dev), qt * sizeof(d4b_ops_dev), cudaMemcpyHostToDevice));
  D4B_CUDA(cudaDeviceSynchronize());
  bound[qt] = true;
  return D4B_OK;
}

/* the element's own tables live in a __constant__ block of the (N, Nq) instance: written once; a level with the
 * other quadrature type waits for everything in flight before it is replaced */
template <int N, int Nq> static int hp_ensure_tab(int qt)
{
  static int have_qt = -1;
  if (have_qt == qt) return D4B_OK;
  if (have_qt != -1) D4B_CUDA(cudaDeviceSynchronize());
  hp_tab<N, Nq> t;
  double I[D4B_OPS_MAXN * D4B_OPS_MAXN], Dm[D4B_OPS_MAXN * D4B_OPS_MAXN];
  d4b_ops_interp(N - 1, Nq - 1, qt, -1, I); d4b_ops_diff(N - 1, Dm);
  for (int q = 0; q < Nq; q++)
    for (int j = 0; j < N; j++) {
      t.B[q * N + j] = I[q * N + j];
      double s = 0.;
      for (int k = 0; k < N; k++) s += I[q * N + k] * Dm[k * N + j];
      t.C[q * N + j] = s;
    }
  for (int i = 0; i < N * N; i++) t.D[i] = Dm[i];
  D4B_CUDA((hp_const<N, Nq>::put(&t)));
  D4B_CUDA(cudaDeviceSynchronize());   /* the staged copy has landed before any stream launches */
  have_qt = qt;
  return D4B_OK;
}

template <int N, int Nq>
static int hp_launch_bucket(const d4b_mesh *m, const d4b_hp_bucket &b, const double *u, double *Au, cudaStream_t st)
{
  static size_t attr_smem = 0;
  { int rct = hp_ensure_tab<N, Nq>(m->quad_type); if (rct != D4B_OK) return rct; }
  const size_t per_warp = hp_warp_doubles<N, Nq>(b.NMX, b.MX);
  int wpb = 4;
  while (wpb > 1 && wpb * per_warp * 8 > 110 * 1024) wpb >>= 1;   /* at least two CTAs per SM */
  const size_t smem = wpb * per_warp * 8;
  if (smem > attr_smem) {
    D4B_CUDA(cudaFuncSetAttribute(d4b_apply_hp_kernel<N, Nq>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_smem = smem;
  }
  hp_args A;
  A.slots = (const hp_slot *)b.slots; A.recs = (const hp_rec *)b.recs; A.G = m->G; A.fac = m->fac; A.bc_g = m->bc_g;
  A.u = u; A.u_ghost = m->u_ghost; A.Au = Au; A.n_list = b.count; A.qt = m->quad_type; A.NMX = b.NMX; A.MX = b.MX;
  A.per_warp = (int)per_warp;
  d4b_apply_hp_kernel<N, Nq><<<(b.count + wpb - 1) / wpb, 32 * wpb, smem, st>>>(A);
  D4B_CUDA(cudaGetLastError());
  d4b_count_launch(1);
  return D4B_OK;
}

int d4b_launch_hp_bucket(const d4b_mesh *m, int i, const double *u, double *Au, cudaStream_t st)
{
  int rc = hp_bind_ops(m->quad_type);
  if (rc != D4B_OK) return rc;
  const d4b_hp_bucket &b = m->hp[i];
  rc = D4B_EINVAL;
#define X(a, c, g) if (b.N == a && b.Nq == c) rc = hp_launch_bucket<a, c>(m, b, u, Au, st);
  HP_PAIRS(X)
#undef X
  return rc;
}

template <int N, int Nq>
static int tp_launch_bucket(const d4b_mesh *m, int pass, const d4b_tp_bucket &b, const double *u, double *Au, cudaStream_t st)
{
  using L = HL<N, Nq>;
  static size_t attr_smem[2] = {0, 0};
  { int rct = hp_ensure_tab<N, Nq>(m->quad_type); if (rct != D4B_OK) return rct; }
  size_t per_warp;
  if (pass == 1) { int scr = tp1_scratch(N, b.NMX); if (scr < L::VOL) scr = L::VOL; per_warp = (size_t)L::UA + scr; }
  else per_warp = (size_t)L::UA + 12 * L::N2 + tp2_scratch(N, b.NMX);
  int wpb = 4;
  while (wpb > 1 && wpb * per_warp * 8 > 110 * 1024) wpb >>= 1;
  const size_t smem = wpb * per_warp * 8;
  if (smem > attr_smem[pass - 1]) {
    if (pass == 1) D4B_CUDA(cudaFuncSetAttribute(d4b_tp_trace_kernel<N, Nq>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else D4B_CUDA(cudaFuncSetAttribute(d4b_tp_flux_kernel<N, Nq>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_smem[pass - 1] = smem;
  }
  tp_args A;
  A.slots = (const tp_slot *)b.slots; A.recs = (const tp_rec *)b.recs; A.G = m->G; A.fac = m->fac; A.bc_g = m->bc_g;
  A.u = u; A.Au = Au; A.W = m->tpW; A.n_list = b.count; A.qt = m->quad_type; A.NMX = b.NMX; A.per_warp = (int)per_warp;
  const int grid = (b.count + wpb - 1) / wpb;
  if (pass == 1) d4b_tp_trace_kernel<N, Nq><<<grid, 32 * wpb, smem, st>>>(A);
  else d4b_tp_flux_kernel<N, Nq><<<grid, 32 * wpb, smem, st>>>(A);
  D4B_CUDA(cudaGetLastError());
  d4b_count_launch(1);
  return D4B_OK;
}

int d4b_launch_tp_bucket(const d4b_mesh *m, int pass, int i, const double *u, double *Au, cudaStream_t st)
{
  int rc = hp_bind_ops(m->quad_type);
  if (rc != D4B_OK) return rc;
  const d4b_tp_bucket &b = pass == 1 ? m->tp1[i] : m->tp2[i];
  rc = D4B_EINVAL;
#define X(a, c, g) if (b.N == a && b.Nq == c) rc = tp_launch_bucket<a, c>(m, pass, b, u, Au, st);
  HP_PAIRS(X)
#undef X
  return rc;
}
