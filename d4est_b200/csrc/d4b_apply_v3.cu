/* d4b_apply_v3.cu -- register-blocked SIPG apply for "regular" elements (same element class as
 * d4b_apply_fast.cu: every face a physical boundary or a conforming interface with an
 * equal-(deg,deg_quad) neighbour, any inter-tree orientation).
 *
 * Formulation (same operator as dGMath/d4est_laplacian_with_opt.c:289, restated in the Lagrange
 * basis ON the quadrature points, exact because deg_quad >= deg):
 *     ut = (B x B x B) u                       interpolation GLL -> quad points
 *     g_l = Dq_l ut                            collocated reference gradient
 *     h   = G g                                metric (Quadrature/d4est_quadrature.c:339-375)
 *     face trace of field x on face (d,+-) = sum_m e+-(m) x(..m..)   (rank-1 along the normal),
 *     normal derivative = sum_m ed+-(m) ut(..m..), ed = Dq^T e
 *     flux at mortar points as in ..._flux_sipg.c:645-677 (interior) / :139-191 (Dirichlet)
 *     rt  = sum_l Dq_l^T (h_l + lift H_l) + lift F ;   Au = (B x B x B)^T rt
 * Thread mapping: Nq threads per element, each holding one PLANE of the element in registers, in
 * two layouts -- A: thread = r-index, plane (s,t);  C: thread = t-index, plane (r,s).  Two of the
 * three contractions are register-local in each layout; the third moves through a per-warp
 * shared-memory transposition (__syncwarp only, no block barriers).
 * Neighbour side of a face: if the neighbour is in the same warp batch (opposite face, identity
 * orientation -- 24 of the 48 faces of a Morton family of 8) its own trace and flux come by warp
 * shuffle from the threads that hold them and a+ = -a- of the neighbour; otherwise one lane stages
 * the neighbour's nodal values (cp.async.bulk + mbarrier), turns them into the trace and reference
 * gradient at the neighbour's face points, and the flux threads fold those with the a+ rows of MY
 * record.  Batches with few enough external faces are staged in one round, the others in three.
 */
#include <vector>
#include "d4b_internal.h"
#include "d4b_ops.h"
#include <cstdlib>
#include <cstring>

#define V3_MAXN 8
#define V3_MAXQ 8
struct v3_tab {
  double B[V3_MAXQ * V3_MAXN];   /* Nq x N  GLL -> quad interpolation                 */
  double Dq[V3_MAXQ * V3_MAXQ];  /* Nq x Nq derivative of the quad-point Lagrange basis */
  double D[V3_MAXN * V3_MAXN];   /* N x N   GLL derivative (neighbour normal lines)    */
  double e[2][V3_MAXQ];          /* quad-point basis at -1 / +1                        */
  double ed[2][V3_MAXQ];         /* its derivative at -1 / +1  (= Dq^T e)              */
  double w[V3_MAXQ];
};
/* one immutable table per template instance: written once (per quadrature type), so launches of different
 * instances on different streams never race on it */
#define V3_PAIRS(X) X(2, 2) X(2, 3) X(2, 4) X(3, 3) X(3, 4) X(4, 4)
enum {
#define X(a, b) V3I_##a##_##b,
  V3_PAIRS(X)
#undef X
  V3_NINST
};
template <int N, int Nq> struct v3_index;
#define X(a, b) template <> struct v3_index<a, b> { static constexpr int value = V3I_##a##_##b; };
V3_PAIRS(X)
#undef X
__constant__ v3_tab c_v3s[V3_NINST];

#define FULL 0xffffffffu


/* ---- small dense contractions on register planes p[x + X*y] ------------------------------ */
/* along x:  out[o + O*y] = sum_i M(o,i) in[i + I*y],  M(o,i) = TR ? M[i*O+o] : M[o*I+i]      */
template <int I, int O, int Y, bool TR>
__device__ __forceinline__ void cx(const double *M, const double (&in)[I * Y], double (&out)[O * Y])
{
#pragma unroll
  for (int y = 0; y < Y; y++)
#pragma unroll
    for (int o = 0; o < O; o++) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < I; i++) s = fma(TR ? M[i * O + o] : M[o * I + i], in[i + I * y], s);
      out[o + O * y] = s;
    }
}
/* along y:  out[x + X*o] = sum_i M(o,i) in[x + X*i] */
template <int I, int O, int X, bool TR>
__device__ __forceinline__ void cy(const double *M, const double (&in)[X * I], double (&out)[X * O])
{
#pragma unroll
  for (int o = 0; o < O; o++)
#pragma unroll
    for (int x = 0; x < X; x++) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < I; i++) s = fma(TR ? M[i * O + o] : M[o * I + i], in[x + X * i], s);
      out[x + X * o] = s;
    }
}
template <int I, int O, int X, bool TR>
__device__ __forceinline__ void cy_acc(const double *M, const double (&in)[X * I], double (&out)[X * O])
{
#pragma unroll
  for (int o = 0; o < O; o++)
#pragma unroll
    for (int x = 0; x < X; x++) {
      double s = out[x + X * o];
#pragma unroll
      for (int i = 0; i < I; i++) s = fma(TR ? M[i * O + o] : M[o * I + i], in[x + X * i], s);
      out[x + X * o] = s;
    }
}
template <int I, int O, int Y, bool TR>
__device__ __forceinline__ void cx_acc(const double *M, const double (&in)[I * Y], double (&out)[O * Y])
{
#pragma unroll
  for (int y = 0; y < Y; y++)
#pragma unroll
    for (int o = 0; o < O; o++) {
      double s = out[o + O * y];
#pragma unroll
      for (int i = 0; i < I; i++) s = fma(TR ? M[i * O + o] : M[o * I + i], in[i + I * y], s);
      out[o + O * y] = s;
    }
}

/* ---- mbarrier / bulk-copy PTX ------------------------------------------------------------ */
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void *bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void *bar, uint32_t parity)
{
  asm volatile(
      "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(
          smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, void *bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}


__global__ void v3_fill_slot_of(int n, const int32_t *__restrict__ elem_list, int32_t *__restrict__ slot_of)
{
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) slot_of[elem_list[s]] = s;
}

__global__ void v3_build_slots(int n, const int32_t *__restrict__ elem_list, const int32_t *__restrict__ frec,
                               const int32_t *__restrict__ nrec, const d4b_hm_t *__restrict__ hm,
                               const int64_t *__restrict__ node_off, const int64_t *__restrict__ quad_off,
                               int n_local, int64_t local_nodes, int Q2, const int32_t *__restrict__ slot_of, int epw,
                               int share, v3_slot *__restrict__ out)
{
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const int e = elem_list[s];
  v3_slot S;
  S.node_off = node_off[e]; S.quad_off = quad_off[e];
  for (int f = 0; f < 6; f++) {
    const d4b_hm_t h = hm[frec[6 * s + f]];
    const int nr = nrec ? nrec[6 * s + f] : -2;
    v3_face F;
    F.fac_off = h.fac_off; F.g_off = h.g_off; F.nbr_u = 0; F.spare = 0; F.pad = 0;
    int mode = 0;
    if (h.nbr >= 0) {
      mode = nr >= 0 ? 1 : 2;
      F.nbr_u = h.nbr < n_local ? node_off[h.nbr] : -(node_off[h.nbr] - local_nodes) - 1;
    }
    F.code = (h.tcode & 7) | ((h.f_p & 7) << 3) | (mode << 6);
    if (mode == 1 && h.tcode == 1 && h.f_p == (f ^ 1) && share) {
      const int ns = slot_of[h.nbr];
      if (ns >= 0 && ns / epw == s / epw) F.code |= (1 << 8) | ((ns % epw) << 9);
    }
    S.f[f] = F;
  }
  out[s] = S;
}

/* Staging plan of one batch of `epw` slots: tasks[round][lane] = -1 (none) or the source of the copy in one
 * word, (offset of the neighbour's nodal values << 8) | ghost << 3 | f_p, so that issuing a round costs a
 * single coalesced load.
 * External interior faces (neighbour outside the batch) need their neighbour's nodal values staged and
 * turned into a trace by one lane each.  If the whole batch has at most `nslot` of them they all go
 * into round 0 (a Morton-ordered family of 8 has 24); otherwise round d carries faces 2d, 2d+1 of
 * every element in slot 2 * element + (face & 1), consumed before round d + 1 overwrites them. */
#define V3_TASK_WORDS 96
__global__ void v3_build_tasks(int n, int epw, int nslot, v3_slot *__restrict__ slots, int64_t *__restrict__ tasks)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int nb = (n + epw - 1) / epw;
  if (b >= nb) return;
  int64_t *T = tasks + (size_t)b * V3_TASK_WORDS;
  for (int i = 0; i < V3_TASK_WORDS; i++) T[i] = -1;
  int next = 0;
  for (int le = 0; le < epw && b * epw + le < n; le++)
    for (int f = 0; f < 6; f++) {
      const int c = slots[b * epw + le].f[f].code;
      if (((c >> 6) & 3) && !((c >> 8) & 1)) next++;
    }
  const bool one_round = next <= nslot;
  next = 0;
  for (int le = 0; le < epw && b * epw + le < n; le++)
    for (int f = 0; f < 6; f++) {
      v3_face &F = slots[b * epw + le].f[f];
      int &c = F.code;
      if (!((c >> 6) & 3) || ((c >> 8) & 1)) continue;
      const int round = one_round ? 0 : (f >> 1), sl = one_round ? next++ : 2 * le + (f & 1);
      const int64_t nu = F.nbr_u;
      T[32 * round + sl] = ((nu >= 0 ? nu : -nu - 1) << 8) | (nu < 0 ? 8 : 0) | ((c >> 3) & 7);
      c = (c & ~(31 << 14)) | (sl << 14);
    }
}

struct v3_args {
  const v3_slot *slots; const int64_t *tasks; const double *G; const double *fac; const double *bc_g;
  const double *u, *u_ghost; double *Au; int n_list; /* one past the last slot */ int slot_begin;
};

/* read-only 8-byte load that the compiler may not sink towards its use: a group of these, issued
 * back to back, keeps a whole row of independent loads in flight */
__device__ __forceinline__ double ldg_pin(const double *p)
{
  double v;
  asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}

template <int N, int Nq> struct V3 {
  static constexpr int TPE = Nq, EPW = 32 / Nq, Q2 = Nq * Nq, Q3 = Q2 * Nq, N2 = N * N, N3 = N2 * N;
  static constexpr bool SWZ = (N == 4 && Nq == 4);   /* XOR-swizzled 64-double element blocks      */
  static constexpr int PS = Q2 + 1;                  /* padded plane stride (generic layout)       */
  static constexpr int EST = SWZ ? 64 : ((Nq * PS + 3) / 4) * 4; /* element stride of X/Y/Z buffers */
  static constexpr int BUF = EPW * EST;
  static constexpr int HTS = EPW * 2 * Q2;
  static constexpr int TASKD = N3;                    /* doubles staged per external face: u of the neighbour */
  static constexpr int TROUT = 4 * Q2;               /* trace written over them: u+, du+/dr, du+/ds, du+/dt  */
  static constexpr int SD = TASKD > TROUT ? TASKD : TROUT;
  static constexpr int SS = SD + 2;                  /* even (16-byte bulk copies), SS/2 odd for SD = 16, 64 */
  static constexpr int SM_BYTES = 233472, CTA_RESERVED = 1024;
  /* staging slots: as many as two CTAs x 4 warps per SM leave room for, at least one round of the
   * three-round plan (2 per element), at most one per lane */
  static constexpr int AVAIL = (SM_BYTES / 2 - CTA_RESERVED) / 32 - (3 * BUF + HTS);
  static constexpr int NS_FIT = AVAIL > 0 ? AVAIL / SS : 0;
  static constexpr int NSLOT = NS_FIT > 32 ? 32 : (NS_FIT < 2 * EPW ? 2 * EPW : NS_FIT);
  static constexpr int NSS = NSLOT * SS;
  static constexpr int BARPAD = 0;                   /* the mbarrier sits in the pad of slot 0     */
  static constexpr int PERWARP = NSS + 3 * BUF + HTS + BARPAD; /* doubles                          */
  static constexpr bool TWO = 2 * (4 * PERWARP * 8 + CTA_RESERVED) <= SM_BYTES;   /* 2 CTAs x 4 warps */
  static constexpr int WPB_MAX = (SM_BYTES - CTA_RESERVED) / (PERWARP * 8);
  static constexpr int WPB = TWO ? 4 : (WPB_MAX >= 6 ? 6 : (WPB_MAX >= 1 ? WPB_MAX : 1));
  static constexpr int CTAS = TWO ? 2 : 1;
  static constexpr int THREADS = 32 * WPB;
  static constexpr size_t SMEM = (size_t)WPB * PERWARP * 8;
  static constexpr bool BULK = (N3 % 2 == 0);
  /* index of (a,b,c) inside the element block of a transposition buffer.  Swizzled form: both the
   * layout-A access (lanes vary a and element) and the layout-C access (lanes vary c and element)
   * touch 16 distinct 8-byte bank pairs per half-warp. */
  static __device__ __forceinline__ int ti(int a, int b, int c, int le)
  {
    return SWZ ? ((a ^ c) + 4 * (b ^ (le & 3)) + 16 * c) : (a + Nq * b + PS * c);
  }
};

/* my face point (ma,mb) -> index of the same physical point in the neighbour's face order
 * (inverse of the map used by d4b_apply_fast.cu; tcode per include/d4est_b200.h) */
template <int Nq> __device__ __forceinline__ int to_nbr_point(int tc, int ma, int mb)
{
  int xx = (tc & 1) ? ma : mb, yy = (tc & 1) ? mb : ma;
  int qa = (tc & 2) ? Nq - 1 - xx : xx, qb = (tc & 4) ? Nq - 1 - yy : yy;
  return qa + Nq * qb;
}
/* neighbour face point (qa,qb) -> index in MY face order */
template <int Nq> __device__ __forceinline__ int to_my_point(int tc, int qa, int qb)
{
  int xx = (tc & 2) ? Nq - 1 - qa : qa, yy = (tc & 4) ? Nq - 1 - qb : qb;
  return (tc & 1) ? xx + Nq * yy : yy + Nq * xx;
}

/* One lane turns the staged nodal values S = u_nbr[N^3] of a neighbour into its trace on the shared
 * face f_p, at the neighbour's face quadrature points q' (neighbour order), written over S:
 * out[q'] = u+,  out[(1 + l) Q2 + q'] = du+/dr_l in the neighbour's reference directions l = 0..2. */
template <int N, int Nq>
__device__ __forceinline__ void v3_nbr_trace(double *S, int f_p)
{
  constexpr int N2 = N * N, Q2 = Nq * Nq;
  const v3_tab &c_v3 = c_v3s[v3_index<N, Nq>::value];
  const int dp = f_p >> 1;
  const int sa = dp == 0 ? N : 1, sb = dp == 2 ? N : N2, sn = dp == 0 ? 1 : (dp == 1 ? N : N2);
  const int lay = (f_p & 1) ? (N - 1) * sn : 0;
  const double *Dl = c_v3.D + ((f_p & 1) ? (N - 1) * N : 0);
  const int ia = dp == 0 ? 1 : 0, ib = dp == 2 ? 1 : 2;
  double U[Q2], DN[Q2];
  {
    double dn[N2], T1[Nq * N];
#pragma unroll
    for (int xb = 0; xb < N; xb++)
#pragma unroll
      for (int xa = 0; xa < N; xa++) {
        const double *p = S + xa * sa + xb * sb;
        double s = 0.;
#pragma unroll
        for (int m = 0; m < N; m++) s = fma(Dl[m], p[m * sn], s);
        dn[xa + N * xb] = s;
      }
    cx<N, Nq, N, false>(c_v3.B, dn, T1); cy<N, Nq, Nq, false>(c_v3.B, T1, DN);
  }
  {
    double X[N2], T1[Nq * N];
#pragma unroll
    for (int xb = 0; xb < N; xb++)
#pragma unroll
      for (int xa = 0; xa < N; xa++) X[xa + N * xb] = S[xa * sa + xb * sb + lay];
    cx<N, Nq, N, false>(c_v3.B, X, T1); cy<N, Nq, Nq, false>(c_v3.B, T1, U);
  }
  /* every read of S is done: the slot is overwritten from here on */
#pragma unroll
  for (int q = 0; q < Q2; q++) { S[q] = U[q]; S[(1 + dp) * Q2 + q] = DN[q]; }
  {
    double DA[Q2];
    cx<Nq, Nq, Nq, false>(c_v3.Dq, U, DA);
#pragma unroll
    for (int q = 0; q < Q2; q++) S[(1 + ia) * Q2 + q] = DA[q];
    cy<Nq, Nq, Nq, false>(c_v3.Dq, U, DA);
#pragma unroll
    for (int q = 0; q < Q2; q++) S[(1 + ib) * Q2 + q] = DA[q];
  }
}

/* face classes of a code word */
__device__ __forceinline__ bool v3_is_bnd(int code) { return ((code >> 6) & 3) == 0; }
__device__ __forceinline__ bool v3_is_int(int code) { return (code >> 8) & 1; }
__device__ __forceinline__ bool v3_is_ext(int code) { return !v3_is_bnd(code) && !v3_is_int(code); }

/* SIPG flux at the Nq face points a thread owns (…_flux_sipg.c:645-677, Dirichlet :139-191).
 * Points i = 0..Nq-1 have face-local coordinates (ma,mb) = FIRST_FIXED ? (t,i) : (i,t).
 * a[4][Nq]: a-_r, a-_s, a-_t, sj*sigma at the points (loaded by the caller ahead of use).
 * in: own trace ut and own flux qm = sum_l a-_l du-/dr_l;  neighbour side up = u+, qp = sum_l a+_l du+/dr_l
 * at the same points (ignored on the boundary).
 * out: F (tested with v), Hr/Hs/Ht (tested with dv/dr_l). */
template <int Nq, bool FIRST_FIXED>
__device__ __forceinline__ void v3_flux(const v3_tab &c_v3, const v3_args &A, int code, const v3_face *fc, int t,
                                        const double (&a)[4][Nq], const double (&ut)[Nq], const double (&qm)[Nq],
                                        const double (&up)[Nq], const double (&qp)[Nq], double (&F)[Nq],
                                        double (&Hr)[Nq], double (&Hs)[Nq], double (&Ht)[Nq])
{
  const bool bnd = v3_is_bnd(code);
  const int64_t g_off = bnd ? __ldg(&fc->g_off) : -1;
  const double wt = c_v3.w[t];
#pragma unroll
  for (int i = 0; i < Nq; i++) {
    const int ma = FIRST_FIXED ? t : i, mb = FIRST_FIXED ? i : t;
    const double w = wt * c_v3.w[i];
    double jump, t1, hs;
    if (!bnd) {
      jump = ut[i] - up[i]; t1 = -0.5 * (qm[i] + qp[i]); hs = -0.5;
    } else {
      const double g = (g_off >= 0) ? __ldg(A.bc_g + g_off + ma + Nq * mb) : 0.;
      jump = ut[i] - g; t1 = -qm[i]; hs = -1.0;
    }
    F[i] = w * (t1 + a[3][i] * jump);
    const double c = w * hs * jump;
    Hr[i] = c * a[0][i]; Hs[i] = c * a[1][i]; Ht[i] = c * a[2][i];
  }
}
/* neighbour side of an EXTERNAL face from its staged trace NB (v3_nbr_trace) and my record's a+ at my points */
template <int Nq, bool FIRST_FIXED>
__device__ __forceinline__ void v3_ext_side(const double *NB, int code, int t, const double (&ap)[3][Nq],
                                            double (&up)[Nq], double (&qp)[Nq])
{
  constexpr int Q2 = Nq * Nq;
  const int tc = code & 7;
  /* neighbour-order index of my point (ma,mb) is affine: q0 + ma*sa + mb*sb (see to_nbr_point) */
  const int f2 = (tc & 2) ? -1 : 1, f4 = (tc & 4) ? -Nq : Nq;
  const int sa = (tc & 1) ? f2 : f4, sb = (tc & 1) ? f4 : f2;
  const int q0 = ((tc & 2) ? Nq - 1 : 0) + ((tc & 4) ? Nq * (Nq - 1) : 0);
  const double *NBt = NB + q0 + t * (FIRST_FIXED ? sa : sb);
  const int st = FIRST_FIXED ? sb : sa;
#pragma unroll
  for (int i = 0; i < Nq; i++) {
    const double *nb = NBt + i * st;
    up[i] = nb[0];
    qp[i] = ap[0][i] * nb[Q2] + ap[1][i] * nb[2 * Q2] + ap[2][i] * nb[3 * Q2];
  }
}
/* the own-side factors of a face at my points: issued early so that the loads overlap the arithmetic
 * before their use */
template <int Nq, bool FIRST_FIXED>
__device__ __forceinline__ void v3_load_fac(const v3_args &A, int64_t fac_off, int code, int t, double (&a)[4][Nq])
{
  constexpr int Q2 = Nq * Nq;
  const double *Fp = A.fac + fac_off;
  const int sj = v3_is_bnd(code) ? 3 : 6;
#pragma unroll
  for (int i = 0; i < Nq; i++) {
    const int q = FIRST_FIXED ? t + Nq * i : i + Nq * t;
    a[0][i] = ldg_pin(Fp + q); a[1][i] = ldg_pin(Fp + Q2 + q); a[2][i] = ldg_pin(Fp + 2 * Q2 + q);
    a[3][i] = ldg_pin(Fp + sj * Q2 + q);
  }
}
/* neighbour side of the two faces (s = 0, 1) of one direction: INTERNAL faces take the partner
 * element's own trace and flux of its opposite face from the threads that hold them (same t, same
 * points: identity orientation), a+ = -a- of the neighbour;  EXTERNAL faces read their staged trace */
template <int Nq, int TPE, bool FIRST_FIXED>
__device__ __forceinline__ void v3_nbr_side(const v3_args &A, unsigned mask, int lane, int t, const double *NS, int SS,
                                            const int (&code)[2], const int64_t (&foff)[2], const double (&ut)[2][Nq],
                                            const double (&qm)[2][Nq], double (&up)[2][Nq], double (&qp)[2][Nq])
{
  constexpr int Q2 = Nq * Nq;
#pragma unroll
  for (int s = 0; s < 2; s++) {
    const int src = v3_is_int(code[s]) ? ((code[s] >> 9) & 31) * TPE + t : lane;
#pragma unroll
    for (int i = 0; i < Nq; i++) {
      up[s][i] = __shfl_sync(mask, ut[s ^ 1][i], src);
      qp[s][i] = -__shfl_sync(mask, qm[s ^ 1][i], src);
    }
    if (v3_is_ext(code[s])) {
      /* a+ of my record at my points (L2-resident: prefetched at the start of the batch) */
      double ap[3][Nq];
      const double *Fp = A.fac + foff[s] + 3 * Q2;
#pragma unroll
      for (int i = 0; i < Nq; i++) {
        const int q = FIRST_FIXED ? t + Nq * i : i + Nq * t;
        ap[0][i] = ldg_pin(Fp + q); ap[1][i] = ldg_pin(Fp + Q2 + q); ap[2][i] = ldg_pin(Fp + 2 * Q2 + q);
      }
      v3_ext_side<Nq, FIRST_FIXED>(NS + ((code[s] >> 14) & 31) * SS, code[s], t, ap, up[s], qp[s]);
    }
  }
}

template <int N, int Nq>
__global__ void __launch_bounds__(V3<N, Nq>::THREADS, V3<N, Nq>::CTAS) d4b_apply_v3_kernel(const v3_args A)
{
  using L = V3<N, Nq>;
  constexpr int TPE = L::TPE, EPW = L::EPW, Q2 = L::Q2, Q3 = L::Q3, N2 = L::N2, N3 = L::N3;
  constexpr int EST = L::EST, SS = L::SS;
  extern __shared__ __align__(128) unsigned char smraw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *W = reinterpret_cast<double *>(smraw) + (size_t)warp * L::PERWARP;
  double *NS = W, *BX = NS + L::NSS, *BY = BX + L::BUF, *BZ = BY + L::BUF, *HT = BZ + L::BUF;
  uint64_t *bar = reinterpret_cast<uint64_t *>(NS + L::SD);
  const int le = lane / TPE, t = lane % TPE;
  const v3_tab &c_v3 = c_v3s[v3_index<N, Nq>::value];
  const double *B = c_v3.B, *Dq = c_v3.Dq;
#define TI(a, b, c) L::ti(a, b, c, le)
  if (lane == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncwarp();
  uint32_t npar = 0;
  /* persistent: every warp walks its own batches of EPW elements; warps never synchronise with
   * each other, so their load phases drift apart and overlap with the other warps' arithmetic */
  const int nbatch = (A.n_list - A.slot_begin + EPW - 1) / EPW;
#pragma unroll 1
  for (int batch = blockIdx.x * L::WPB + warp; batch < nbatch; batch += gridDim.x * L::WPB) {
  const int slot0 = A.slot_begin + batch * EPW;
  const int slot = slot0 + le;
  const bool act = le < EPW && slot < A.n_list;
  const unsigned amask = __ballot_sync(FULL, act);
  const v3_slot *sl = A.slots + (act ? slot : 0);
  const int64_t noff = act ? __ldg(&sl->node_off) : -1;
  /* record offsets and codes of all six faces and the metric offset, loaded once up front: the factor
   * loads of a phase then do not wait for a pointer fetch first */
  int64_t foffs[6]; int fcodes[6];
#pragma unroll
  for (int f = 0; f < 6; f++) { foffs[f] = __ldg(&sl->f[f].fac_off); fcodes[f] = __ldg(&sl->f[f].code); }
  const int64_t qoff = __ldg(&sl->quad_off);

  /* ---- neighbour staging: round r covers faces 2r, 2r+1 of every element of the batch ---- */
  const int64_t *tkb = A.tasks + (size_t)(slot0 / EPW) * V3_TASK_WORDS;
  int64_t tk = -1;
  auto issue_round = [&](int r) -> uint32_t {
    tk = __ldg(tkb + 32 * r + lane);
    uint32_t bytes = 0;
    const double *su = nullptr;
    if (tk >= 0) {
      su = ((tk & 8) ? A.u_ghost : A.u) + (tk >> 8);
      if (L::BULK && (((uintptr_t)su) & 15) == 0) bytes = N3 * 8;
    }
    const uint32_t total = __reduce_add_sync(FULL, bytes);
    if (lane == 0 && total) mbar_expect_tx(bar, total);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (tk >= 0) {
      double *dst = NS + lane * SS;
      if (bytes) bulk_g2s(dst, su, N3 * 8, bar);
      else /* unaligned source (odd sizes): plain copy by the owning lane */
        for (int i = 0; i < N3; i++) dst[i] = __ldg(su + i);
    }
    return total;
  };
  /* sign words of the round-1 and round-2 tasks: a batch staged in one round (the common case) skips
   * the later rounds without touching memory again */
  const int tk1_hi = __ldg(reinterpret_cast<const int *>(tkb + 32 + lane) + 1);
  const int tk2_hi = __ldg(reinterpret_cast<const int *>(tkb + 64 + lane) + 1);
  uint32_t tot = issue_round(0);

  /* L2 prefetch of the own-side factors of faces 0..3.  Requests are issued close to their use: with 64
   * elements in flight per SM, everything prefetched at once would not stay in L2 until it is read */
  if (act) {
    for (int f = t; f < 4; f += TPE) {
      const char *p = reinterpret_cast<const char *>(A.fac + __ldg(&sl->f[f].fac_off));
      const bool ext = v3_is_ext(__ldg(&sl->f[f].code));
      for (int o = 0; o < 7 * Q2 * 8; o += 128)
        if (ext || o < 3 * Q2 * 8 || o + 128 > 6 * Q2 * 8) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
    }
  }

  /* ---- own u -> BX, coalesced; all loads of the batch in flight before the first store ---- */
  {
    constexpr int PL = (N3 + 31) / 32;
    double uv[EPW][PL];
#pragma unroll
    for (int ee = 0; ee < EPW; ee++) {
      const int64_t base = __shfl_sync(FULL, noff, ee * TPE);
#pragma unroll
      for (int m = 0; m < PL; m++) {
        const int n = lane + 32 * m;
        uv[ee][m] = (base >= 0 && n < N3) ? __ldg(A.u + base + n) : 0.;
      }
    }
#pragma unroll
    for (int ee = 0; ee < EPW; ee++)
#pragma unroll
      for (int m = 0; m < PL; m++) {
        const int n = lane + 32 * m;
        if (n < N3) BX[ee * EST + L::ti(n % N, (n / N) % N, n / N2, ee)] = uv[ee][m];
      }
  }
  __syncwarp();
  double *Xe = BX + le * EST, *Ye = BY + le * EST, *Ze = BZ + le * EST;

  /* P1 (layout C, nodal): thread k = t < N holds u(:,:,k); B_r, B_s in registers -> v(a,b,k) -> BY */
  if (act && t < N) {
    double X[N2], V1[Nq * N], V[Q2];
#pragma unroll
    for (int j = 0; j < N; j++)
#pragma unroll
      for (int i = 0; i < N; i++) X[i + N * j] = Xe[TI(i, j, t)];
    cx<N, Nq, N, false>(B, X, V1); cy<N, Nq, Nq, false>(B, V1, V);
#pragma unroll
    for (int b = 0; b < Nq; b++)
#pragma unroll
      for (int a = 0; a < Nq; a++) Ye[TI(a, b, t)] = V[a + Nq * b];
  }
  __syncwarp();
  /* A0 (layout A): thread a = t holds (b,c); B_t -> ut; write ut -> BX (T2) and g_t -> BZ (T3) */
  if (act) {
    double V[Nq * N], ut[Q2], gt[Q2];
#pragma unroll
    for (int k = 0; k < N; k++)
#pragma unroll
      for (int b = 0; b < Nq; b++) V[b + Nq * k] = Ye[TI(t, b, k)];
    cy<N, Nq, Nq, false>(B, V, ut);
    cy<Nq, Nq, Nq, false>(Dq, ut, gt);
#pragma unroll
    for (int c = 0; c < Nq; c++)
#pragma unroll
      for (int b = 0; b < Nq; b++) { Xe[TI(t, b, c)] = ut[b + Nq * c]; Ze[TI(t, b, c)] = gt[b + Nq * c]; }
  }
  __syncwarp();

  /* ---- three rounds of neighbour traces, each consumed from its slots right after it is built:
   *      faces 0,1 in layout C (C0), faces 2,3 in layout A (A1a), faces 4,5 after the loop (A1b) ---- */
  double Hr23[2][Nq];
#pragma unroll 1
  for (int r = 0; r < 3; r++) {
    if (tot) { mbar_wait(bar, npar); npar ^= 1; }
    __syncwarp();
    if (tk >= 0) v3_nbr_trace<N, Nq>(NS + lane * SS, (int)(tk & 7));
    __syncwarp();
    if (r == 0) {
      /* C0 (layout C): thread c = t holds (a,b): faces 0,1 (normal r), then g_r -> BY (T4) */
      if (act) {
        double uC[Q2], rC[Q2];
        int64_t foff[2]; int fcode[2];
#pragma unroll
        for (int s = 0; s < 2; s++) { foff[s] = foffs[s]; fcode[s] = fcodes[s]; }
#pragma unroll
        for (int b = 0; b < Nq; b++)
#pragma unroll
          for (int a = 0; a < Nq; a++) { uC[a + Nq * b] = Xe[TI(a, b, t)]; rC[a + Nq * b] = 0.; }
        double fa[2][4][Nq], ut2[2][Nq], qm2[2][Nq], up[2][Nq], qp[2][Nq];
#pragma unroll
        for (int s = 0; s < 2; s++) v3_load_fac<Nq, false>(A, foff[s], fcode[s], t, fa[s]);
        { /* g_r = Dq_r ut -> BY (T4), while the factor loads are in flight */
          double gr[Q2];
          cx<Nq, Nq, Nq, false>(Dq, uC, gr);
#pragma unroll
          for (int b = 0; b < Nq; b++)
#pragma unroll
            for (int a = 0; a < Nq; a++) Ye[TI(a, b, t)] = gr[a + Nq * b];
        }
#pragma unroll
        for (int s = 0; s < 2; s++) {
          double dr[Nq], ds[Nq], dt[Nq];
#pragma unroll
          for (int b = 0; b < Nq; b++) {
            double s0 = 0., s1 = 0., s2 = 0.;
#pragma unroll
            for (int a = 0; a < Nq; a++) {
              s0 = fma(c_v3.e[s][a], uC[a + Nq * b], s0); s1 = fma(c_v3.ed[s][a], uC[a + Nq * b], s1);
              s2 = fma(c_v3.e[s][a], Ze[TI(a, b, t)], s2);
            }
            ut2[s][b] = s0; dr[b] = s1; dt[b] = s2;
          }
#pragma unroll
          for (int b = 0; b < Nq; b++) {
            double s0 = 0.;
#pragma unroll
            for (int m = 0; m < Nq; m++) s0 = fma(Dq[b * Nq + m], ut2[s][m], s0);
            ds[b] = s0;
          }
#pragma unroll
          for (int b = 0; b < Nq; b++) qm2[s][b] = fa[s][0][b] * dr[b] + fa[s][1][b] * ds[b] + fa[s][2][b] * dt[b];
        }
        v3_nbr_side<Nq, TPE, false>(A, amask, lane, t, NS, SS, fcode, foff, ut2, qm2, up, qp);
#pragma unroll
        for (int s = 0; s < 2; s++) {
          double F[Nq], Hr[Nq], Hs[Nq], Ht[Nq];
          v3_flux<Nq, false>(c_v3, A, fcode[s], &sl->f[s], t, fa[s], ut2[s], qm2[s], up[s], qp[s], F, Hr, Hs, Ht);
#pragma unroll
          for (int b = 0; b < Nq; b++) {
            double s0 = F[b];
#pragma unroll
            for (int m = 0; m < Nq; m++) s0 = fma(Dq[m * Nq + b], Hs[m], s0);
#pragma unroll
            for (int a = 0; a < Nq; a++)
              rC[a + Nq * b] = fma(c_v3.e[s][a], s0, fma(c_v3.ed[s][a], Hr[b], rC[a + Nq * b]));
            HT[(le * 2 + s) * Q2 + b + Nq * t] = Ht[b];
          }
        }
        /* rC parked in BZ over g_t (same thread, same addresses) */
#pragma unroll
        for (int b = 0; b < Nq; b++)
#pragma unroll
          for (int a = 0; a < Nq; a++) Ze[TI(a, b, t)] = rC[a + Nq * b];
      }
      __syncwarp();
      /* L2 prefetch of the factors of faces 4,5 (the others were requested at the start of the batch) */
      if (act) {
    for (int f = 4 + t; f < 6; f += TPE) {
      const char *p = reinterpret_cast<const char *>(A.fac + __ldg(&sl->f[f].fac_off));
      const bool ext = v3_is_ext(__ldg(&sl->f[f].code));
      for (int o = 0; o < 7 * Q2 * 8; o += 128)
        if (ext || o < 3 * Q2 * 8 || o + 128 > 6 * Q2 * 8) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
    }
      }
    }
    if (r == 1) {
      /* A1a (layout A): thread a = t holds (b,c): faces 2,3; their lift goes straight into the
       * residual accumulator BZ (which already holds the layout-C part), H_r stays in registers */
      if (act) {
        double ut[Q2], gr[Q2], rA[Q2];
        int64_t foff[2]; int fcode[2];
#pragma unroll
        for (int f = 0; f < 2; f++) { foff[f] = foffs[2 + f]; fcode[f] = fcodes[2 + f]; }
        double fa[2][4][Nq];
#pragma unroll
        for (int s = 0; s < 2; s++) v3_load_fac<Nq, true>(A, foff[s], fcode[s], t, fa[s]);
#pragma unroll
        for (int c = 0; c < Nq; c++)
#pragma unroll
          for (int b = 0; b < Nq; b++) {
            ut[b + Nq * c] = Xe[TI(t, b, c)]; gr[b + Nq * c] = Ye[TI(t, b, c)]; rA[b + Nq * c] = 0.;
          }
        /* faces 2,3: normal s (index b); my points (a=t, c) */
        double ut2[2][Nq], qm2[2][Nq], up[2][Nq], qp[2][Nq];
#pragma unroll
        for (int s = 0; s < 2; s++) {
          double dr[Nq], ds[Nq], dt[Nq];
#pragma unroll
          for (int c = 0; c < Nq; c++) {
            double s0 = 0., s1 = 0., s2 = 0.;
#pragma unroll
            for (int b = 0; b < Nq; b++) {
              s0 = fma(c_v3.e[s][b], ut[b + Nq * c], s0); s1 = fma(c_v3.ed[s][b], ut[b + Nq * c], s1);
              s2 = fma(c_v3.e[s][b], gr[b + Nq * c], s2);
            }
            ut2[s][c] = s0; ds[c] = s1; dr[c] = s2;
          }
#pragma unroll
          for (int c = 0; c < Nq; c++) {
            double s0 = 0.;
#pragma unroll
            for (int m = 0; m < Nq; m++) s0 = fma(Dq[c * Nq + m], ut2[s][m], s0);
            dt[c] = s0;
          }
#pragma unroll
          for (int c = 0; c < Nq; c++) qm2[s][c] = fa[s][0][c] * dr[c] + fa[s][1][c] * ds[c] + fa[s][2][c] * dt[c];
        }
        v3_nbr_side<Nq, TPE, true>(A, amask, lane, t, NS, SS, fcode, foff, ut2, qm2, up, qp);
#pragma unroll
      for (int s = 0; s < 2; s++) {
        double F[Nq], Hr[Nq], Hs[Nq], Ht[Nq];
        v3_flux<Nq, true>(c_v3, A, fcode[s], &sl->f[2 + s], t, fa[s], ut2[s], qm2[s], up[s], qp[s], F, Hr, Hs, Ht);
#pragma unroll
        for (int c = 0; c < Nq; c++) {
          double s0 = F[c];
#pragma unroll
          for (int m = 0; m < Nq; m++) s0 = fma(Dq[m * Nq + c], Ht[m], s0);
          Hr23[s][c] = Hr[c];
#pragma unroll
          for (int b = 0; b < Nq; b++) {
            rA[b + Nq * c] = fma(c_v3.e[s][b], s0, fma(c_v3.ed[s][b], Hs[c], rA[b + Nq * c]));
          }
        }
      }
#pragma unroll
        for (int c = 0; c < Nq; c++)
#pragma unroll
          for (int b = 0; b < Nq; b++) Ze[TI(t, b, c)] += rA[b + Nq * c];
      }
      __syncwarp();
      /* L2 prefetch of the metric block, consumed in the next phase */
      if (act) {
        const char *g = reinterpret_cast<const char *>(A.G + 6 * qoff);
        for (int o = t * 128; o < 6 * Q3 * 8; o += TPE * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + o));
      }
    }
    if (r < 2) {
      if (__any_sync(FULL, (tk1_hi & tk2_hi) >= 0)) tot = issue_round(r + 1);
      else { tot = 0; tk = -1; }
    }
  }

  /* A1b (layout A): faces 4,5, metric, s/t transposed derivatives */
  {
    double rA[Q2];
    if (act) {
      double ut[Q2], gr[Q2], hrF[Q2];
      int64_t foff[2]; int fcode[2];
#pragma unroll
      for (int f = 0; f < 2; f++) { foff[f] = foffs[4 + f]; fcode[f] = fcodes[4 + f]; }
      double fa[2][4][Nq];
#pragma unroll
      for (int s = 0; s < 2; s++) v3_load_fac<Nq, true>(A, foff[s], fcode[s], t, fa[s]);
#pragma unroll
      for (int c = 0; c < Nq; c++)
#pragma unroll
        for (int b = 0; b < Nq; b++) {
          ut[b + Nq * c] = Xe[TI(t, b, c)]; gr[b + Nq * c] = Ye[TI(t, b, c)];
          rA[b + Nq * c] = 0.;
          hrF[b + Nq * c] = c_v3.e[0][b] * Hr23[0][c] + c_v3.e[1][b] * Hr23[1][c];
        }
      /* faces 4,5: normal t (index c); my points (a=t, b) */
      double ut2[2][Nq], qm2[2][Nq], up[2][Nq], qp[2][Nq];
#pragma unroll
      for (int s = 0; s < 2; s++) {
        double dr[Nq], ds[Nq], dt[Nq];
#pragma unroll
        for (int b = 0; b < Nq; b++) {
          double s0 = 0., s1 = 0., s2 = 0.;
#pragma unroll
          for (int c = 0; c < Nq; c++) {
            s0 = fma(c_v3.e[s][c], ut[b + Nq * c], s0); s1 = fma(c_v3.ed[s][c], ut[b + Nq * c], s1);
            s2 = fma(c_v3.e[s][c], gr[b + Nq * c], s2);
          }
          ut2[s][b] = s0; dt[b] = s1; dr[b] = s2;
        }
#pragma unroll
        for (int b = 0; b < Nq; b++) {
          double s0 = 0.;
#pragma unroll
          for (int m = 0; m < Nq; m++) s0 = fma(Dq[b * Nq + m], ut2[s][m], s0);
          ds[b] = s0;
        }
#pragma unroll
        for (int b = 0; b < Nq; b++) qm2[s][b] = fa[s][0][b] * dr[b] + fa[s][1][b] * ds[b] + fa[s][2][b] * dt[b];
      }
      v3_nbr_side<Nq, TPE, true>(A, amask, lane, t, NS, SS, fcode, foff, ut2, qm2, up, qp);
#pragma unroll
      for (int s = 0; s < 2; s++) {
        double F[Nq], Hr[Nq], Hs[Nq], Ht[Nq];
        v3_flux<Nq, true>(c_v3, A, fcode[s], &sl->f[4 + s], t, fa[s], ut2[s], qm2[s], up[s], qp[s], F, Hr, Hs, Ht);
#pragma unroll
        for (int b = 0; b < Nq; b++) {
          double s0 = F[b];
#pragma unroll
          for (int m = 0; m < Nq; m++) s0 = fma(Dq[m * Nq + b], Hs[m], s0);
#pragma unroll
          for (int c = 0; c < Nq; c++) {
            rA[b + Nq * c] = fma(c_v3.e[s][c], s0, fma(c_v3.ed[s][c], Ht[b], rA[b + Nq * c]));
            hrF[b + Nq * c] = fma(c_v3.e[s][c], Hr[b], hrF[b + Nq * c]);
          }
        }
      }
      /* metric: h = G g at my points (a=t,b,c), one c-row of G loads in flight at a time;
       * h_t also receives the lift of faces 0,1 */
#pragma unroll
      for (int c = 0; c < Nq; c++)
#pragma unroll
        for (int b = 0; b < Nq; b++) { Ze[TI(t, b, c)] += rA[b + Nq * c]; rA[b + Nq * c] = 0.; }
      /* rows of G are double-buffered: row 0 is requested before g_s, g_t are computed, row c + 1 before
       * row c is used */
      const double *Ge = A.G + 6 * qoff + t;
      double Gv[2][6][Nq];
      auto load_row = [&](double (&R)[6][Nq], int c) {
#pragma unroll
        for (int k = 0; k < 6; k++)
#pragma unroll
          for (int b = 0; b < Nq; b++) R[k][b] = ldg_pin(Ge + k * Q3 + Nq * (b + Nq * c));
      };
      load_row(Gv[0], 0);
      asm volatile("" ::: "memory");
      double gs[Q2], gt[Q2];
      cx<Nq, Nq, Nq, false>(Dq, ut, gs); cy<Nq, Nq, Nq, false>(Dq, ut, gt);
      const double e0 = c_v3.e[0][t], e1 = c_v3.e[1][t];
      const double *HTe = HT + le * 2 * Q2;
#pragma unroll
      for (int c = 0; c < Nq; c++) {
        if (c + 1 < Nq) load_row(Gv[(c + 1) & 1], c + 1);
        asm volatile("" ::: "memory");
        const double (&R)[6][Nq] = Gv[c & 1];
#pragma unroll
        for (int b = 0; b < Nq; b++) {
          const int i = b + Nq * c;
          const double a = gr[i], bb = gs[i], cc = gt[i];
          gr[i] = R[0][b] * a + R[1][b] * bb + R[2][b] * cc + hrF[i];
          gs[i] = R[1][b] * a + R[3][b] * bb + R[4][b] * cc;
          gt[i] = R[2][b] * a + R[4][b] * bb + R[5][b] * cc + e0 * HTe[i] + e1 * HTe[Q2 + i];
        }
      }
      cx_acc<Nq, Nq, Nq, true>(Dq, gs, rA); cy_acc<Nq, Nq, Nq, true>(Dq, gt, rA);
      /* h_r -> BX (T5; same addresses this thread read ut from); residual part -> accumulator BZ */
#pragma unroll
      for (int c = 0; c < Nq; c++)
#pragma unroll
        for (int b = 0; b < Nq; b++) { Xe[TI(t, b, c)] = gr[b + Nq * c]; Ze[TI(t, b, c)] += rA[b + Nq * c]; }
    }
    __syncwarp();
    /* C1 (layout C): residual accumulator (BZ) + Dq_r^T h_r -> BY (T7) */
    if (act) {
      double hr[Q2], rC[Q2];
#pragma unroll
      for (int b = 0; b < Nq; b++)
#pragma unroll
        for (int a = 0; a < Nq; a++) { hr[a + Nq * b] = Xe[TI(a, b, t)]; rC[a + Nq * b] = Ze[TI(a, b, t)]; }
      cx_acc<Nq, Nq, Nq, true>(Dq, hr, rC);
#pragma unroll
      for (int b = 0; b < Nq; b++)
#pragma unroll
        for (int a = 0; a < Nq; a++) Ye[TI(a, b, t)] = rC[a + Nq * b];
    }
    __syncwarp();
    /* A2 (layout A): rt = rA + rC ; B_t^T, B_s^T in registers -> y(a; j,k) -> BX (T8) */
    if (act) {
      double Y1[Nq * N], Y[N2];
#pragma unroll
      for (int c = 0; c < Nq; c++)
#pragma unroll
        for (int b = 0; b < Nq; b++) rA[b + Nq * c] = Ye[TI(t, b, c)];
      cy<Nq, N, Nq, true>(B, rA, Y1);   /* contract c -> k : Y1[b + Nq k] */
      cx<Nq, N, N, true>(B, Y1, Y);     /* contract b -> j : Y[j + N k]   */
#pragma unroll
      for (int k = 0; k < N; k++)
#pragma unroll
        for (int j = 0; j < N; j++) Xe[TI(t, j, k)] = Y[j + N * k];
    }
  }
  __syncwarp();
  /* C2 (layout C, nodal): thread k = t < N: B_r^T -> Au(i,j,k) -> BZ */
  if (act && t < N) {
    double Yc[Nq * N], O[N2];
#pragma unroll
    for (int j = 0; j < N; j++)
#pragma unroll
      for (int a = 0; a < Nq; a++) Yc[a + Nq * j] = Xe[TI(a, j, t)];
    cx<Nq, N, N, true>(B, Yc, O);
#pragma unroll
    for (int j = 0; j < N; j++)
#pragma unroll
      for (int i = 0; i < N; i++) Ze[TI(i, j, t)] = O[i + N * j];
  }
  __syncwarp();
  {
    constexpr int PL = (N3 + 31) / 32;
    double ov[EPW][PL];
#pragma unroll
    for (int ee = 0; ee < EPW; ee++)
#pragma unroll
      for (int m = 0; m < PL; m++) {
        const int n = lane + 32 * m;
        ov[ee][m] = n < N3 ? BZ[ee * EST + L::ti(n % N, (n / N) % N, n / N2, ee)] : 0.;
      }
#pragma unroll
    for (int ee = 0; ee < EPW; ee++) {
      const int64_t base = __shfl_sync(FULL, noff, ee * TPE);
#pragma unroll
      for (int m = 0; m < PL; m++) {
        const int n = lane + 32 * m;
        if (base >= 0 && n < N3) A.Au[base + n] = ov[ee][m];
      }
    }
  }
  __syncwarp();
  } /* batch loop */
#undef TI
}

/* ------------------------------------------------------------------ host side */

bool d4b_v3_supported(int N, int Nq)
{
  if (getenv("D4B_NO_V3") != nullptr) return false;   /* test switch: route the bucket to the line-task kernel */
#define X(a, b) if (N == a && Nq == b) return true;
  V3_PAIRS(X)
#undef X
  return false;
}

static int v3_upload_tab(int qt, int N, int Nq, cudaStream_t st)
{
  /* each instance has its own __constant__ slot, written once; a level with the other quadrature type waits
   * for everything in flight on the device before the slot is rewritten */
  static int have_qt[V3_NINST];
  static bool init = false;
  if (!init) { for (int i = 0; i < V3_NINST; i++) have_qt[i] = -1; init = true; }
  int idx = -1;
#define X(a, c) if (N == a && Nq == c) idx = V3I_##a##_##c;
  V3_PAIRS(X)
#undef X
  if (idx < 0) { d4b_set_error("no register-blocked instance for N=%d Nq=%d", N, Nq); return D4B_EINVAL; }
  if (have_qt[idx] == qt) return D4B_OK;
  if (have_qt[idx] != -1) D4B_CUDA(cudaDeviceSynchronize());
  v3_tab tt, *t = &tt;
  memset(t, 0, sizeof *t);
  double I[D4B_OPS_MAXN * D4B_OPS_MAXN], Dm[D4B_OPS_MAXN * D4B_OPS_MAXN], Dqm[D4B_OPS_MAXN * D4B_OPS_MAXN];
  double x[D4B_OPS_MAXN], w[D4B_OPS_MAXN], e0[D4B_OPS_MAXN], e1[D4B_OPS_MAXN];
  d4b_ops_interp(N - 1, Nq - 1, qt, -1, I); d4b_ops_diff(N - 1, Dm); d4b_ops_quad(Nq - 1, qt, x, w);
  d4b_ops_quadbasis(Nq - 1, qt, Dqm, e0, e1);
  for (int i = 0; i < Nq * N; i++) t->B[i] = I[i];
  for (int i = 0; i < Nq * Nq; i++) t->Dq[i] = Dqm[i];
  for (int i = 0; i < N * N; i++) t->D[i] = Dm[i];
  for (int m = 0; m < Nq; m++) {
    t->e[0][m] = e0[m]; t->e[1][m] = e1[m]; t->w[m] = w[m];
    double s0 = 0., s1 = 0.;
    for (int a = 0; a < Nq; a++) { s0 += e0[a] * Dqm[a * Nq + m]; s1 += e1[a] * Dqm[a * Nq + m]; }
    t->ed[0][m] = s0; t->ed[1][m] = s1;
  }
  D4B_CUDA(cudaMemcpyToSymbol(c_v3s, t, sizeof(v3_tab), (size_t)idx * sizeof(v3_tab), cudaMemcpyHostToDevice));
  D4B_CUDA(cudaDeviceSynchronize());   /* the staged copy has landed before any stream launches */
  (void)st;
  have_qt[idx] = qt;
  return D4B_OK;
}

static int v3_prepare_slots(const d4b_mesh *m, const d4b_fast_bucket &b, cudaStream_t st);
int d4b_v3_prepare_slots(const d4b_mesh *m, const d4b_fast_bucket &b, cudaStream_t st) { return v3_prepare_slots(m, b, st); }

static int v3_prepare_slots(const d4b_mesh *m, const d4b_fast_bucket &b, cudaStream_t st)
{
  if (b.v3_slots) return D4B_OK;   /* resolved once: the tables are immutable after create */
  v3_slot *s = nullptr;
  int32_t *slot_of = nullptr;
  int epw = 0, nslot = 0;
#define X(a, c) if (b.N == a && b.Nq == c) { epw = V3<a, c>::EPW; nslot = V3<a, c>::NSLOT; }
  V3_PAIRS(X)
#undef X
  const bool plan = epw != 0;   /* staging plan only for the pairs the register-blocked kernel has; the slot table
                                   alone also serves the line-task kernel (d4b_apply_fast.cu) */
  if (!plan) epw = 1;
  const int nbatch = plan ? (b.count + epw - 1) / epw : 0;
  /* test switch: no shuffle exchange inside a batch, every interior face is staged (three rounds) */
  const int share = plan && getenv("D4B_V3_NO_SHARE") == nullptr;
  D4B_CUDA(cudaMalloc((void **)&s, (size_t)b.count * sizeof(v3_slot) + (size_t)nbatch * V3_TASK_WORDS * sizeof(int64_t)));
  D4B_CUDA(cudaMalloc((void **)&slot_of, (size_t)(m->n_local + 1) * sizeof(int32_t)));
  D4B_CUDA(cudaMemsetAsync(slot_of, 0xff, (size_t)(m->n_local + 1) * sizeof(int32_t), st));
  v3_fill_slot_of<<<(b.count + 255) / 256, 256, 0, st>>>(b.count, b.elem_list, slot_of);
  v3_build_slots<<<(b.count + 127) / 128, 128, 0, st>>>(b.count, b.elem_list, b.frec, b.nrec, m->hm, m->node_off,
                                                      m->quad_off, m->n_local, m->local_nodes, b.Nq * b.Nq, slot_of,
                                                      epw, share, s);
  if (plan) v3_build_tasks<<<(nbatch + 127) / 128, 128, 0, st>>>(b.count, epw, nslot, s, reinterpret_cast<int64_t *>(s + b.count));
  D4B_CUDA(cudaGetLastError());
  D4B_CUDA(cudaStreamSynchronize(st));
  cudaFree(slot_of);
  const_cast<d4b_fast_bucket &>(b).v3_slots = s;
  return D4B_OK;
}

/* Slot ranges of the host-buffer pipeline: range k = slots [b[k], b[k+1]) */
struct v3_bounds { int n; int b[D4B_PIPE_MAX + 1]; };
__device__ __forceinline__ int v3_chunk_of(const v3_bounds &B, int slot)
{
  int k = 0;
  while (k + 1 < B.n && slot >= B.b[k + 1]) k++;
  return k;
}
static v3_bounds v3_make_bounds(int n_ranges, const int *bounds)
{
  v3_bounds B; B.n = n_ranges;
  for (int k = 0; k <= n_ranges; k++) B.b[k] = bounds[k];
  return B;
}

/* For the host-buffer pipeline: need[k] = index of the last u chunk that range k of the bucket reads (its own
 * elements and their local face neighbours), chunks and ranges being the same element ranges. */
__global__ void v3_range_need(const v3_slot *__restrict__ slots, v3_bounds B, int64_t N3, int *__restrict__ need)
{
  const int k = blockIdx.x, s0 = B.b[k], s1 = B.b[k + 1];
  long long mx = 0;
  for (int s = s0 + threadIdx.x; s < s1; s += blockDim.x) {
    long long v = slots[s].node_off;
    for (int f = 0; f < 6; f++) if (((slots[s].f[f].code >> 6) & 3) && slots[s].f[f].nbr_u > v) v = slots[s].f[f].nbr_u;
    mx = max(mx, v);
  }
  __shared__ long long red[256];
  red[threadIdx.x] = mx; __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] = max(red[threadIdx.x], red[threadIdx.x + o]); __syncthreads(); }
  if (threadIdx.x == 0) need[k] = v3_chunk_of(B, (int)(red[0] / N3));
}

int d4b_v3_range_need(const d4b_mesh *m, const d4b_fast_bucket &b, int n_ranges, const int *bounds, int *need_host, cudaStream_t st)
{
  int rc = v3_prepare_slots(m, b, st);
  if (rc != D4B_OK) return rc;
  int *d = nullptr;
  D4B_CUDA(cudaMalloc((void **)&d, n_ranges * sizeof(int)));
  v3_range_need<<<n_ranges, 256, 0, st>>>((const v3_slot *)b.v3_slots, v3_make_bounds(n_ranges, bounds), (int64_t)b.N * b.N * b.N, d);
  D4B_CUDA(cudaGetLastError());
  D4B_CUDA(cudaMemcpyAsync(need_host, d, n_ranges * sizeof(int), cudaMemcpyDeviceToHost, st));
  D4B_CUDA(cudaStreamSynchronize(st));
  cudaFree(d);
  return D4B_OK;
}

/* For the host-buffer pipeline: the elements of the bucket that read a face neighbour more than `lookahead`
 * chunks ahead of their own chunk (thin layers at the seams of the space-filling curve) are collected into a
 * bucket of their own, so that every other element of range k can run as soon as chunk k + lookahead is there.
 * Elements with a ghost neighbour join them (N > 1: the halo exchange follows the last upload). */
__global__ void v3_mark_late(const v3_slot *__restrict__ slots, int n, v3_bounds B, int64_t N3, int lookahead,
                             int32_t *__restrict__ flag)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  int late = 0;
  const int mine = v3_chunk_of(B, s);
  for (int f = 0; f < 6; f++) {
    const v3_face &F = slots[s].f[f];
    if (!((F.code >> 6) & 3)) continue;
    if (F.nbr_u < 0) late = 1;   /* ghost neighbour: its values arrive with the halo exchange, after the uploads */
    else if (v3_chunk_of(B, (int)(F.nbr_u / N3)) > mine + lookahead) late = 1;
  }
  flag[s] = late;
}
__global__ void v3_gather_late(int n, const int32_t *__restrict__ list, const int32_t *__restrict__ elem,
                               const int32_t *__restrict__ frec, const int32_t *__restrict__ nrec,
                               int32_t *__restrict__ elem_o, int32_t *__restrict__ frec_o, int32_t *__restrict__ nrec_o)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = list[i];
  elem_o[i] = elem[s];
  for (int f = 0; f < 6; f++) { frec_o[6 * i + f] = frec[6 * s + f]; nrec_o[6 * i + f] = nrec[6 * s + f]; }
}

int d4b_v3_build_late(const d4b_mesh *m, const d4b_fast_bucket &b, int n_ranges, const int *bounds, int lookahead,
                      d4b_fast_bucket *late, cudaStream_t st)
{
  int rc = v3_prepare_slots(m, b, st);
  if (rc != D4B_OK) return rc;
  late->N = b.N; late->Nq = b.Nq; late->count = 0;
  late->elem_list = late->frec = late->nrec = nullptr; late->v3_slots = nullptr;
  int32_t *flag = nullptr;
  D4B_CUDA(cudaMalloc((void **)&flag, (size_t)b.count * sizeof(int32_t)));
  v3_mark_late<<<(b.count + 255) / 256, 256, 0, st>>>((const v3_slot *)b.v3_slots, b.count, v3_make_bounds(n_ranges, bounds),
                                                      (int64_t)b.N * b.N * b.N, lookahead, flag);
  D4B_CUDA(cudaGetLastError());
  std::vector<int32_t> h(b.count), list;
  D4B_CUDA(cudaMemcpyAsync(h.data(), flag, (size_t)b.count * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  D4B_CUDA(cudaStreamSynchronize(st));
  for (int s = 0; s < b.count; s++) if (h[s]) list.push_back(s);
  const int n = (int)list.size();
  if (n) {
    D4B_CUDA(cudaMemcpyAsync(flag, list.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    D4B_CUDA(cudaMalloc((void **)&late->elem_list, (size_t)n * sizeof(int32_t)));
    D4B_CUDA(cudaMalloc((void **)&late->frec, (size_t)n * 6 * sizeof(int32_t)));
    D4B_CUDA(cudaMalloc((void **)&late->nrec, (size_t)n * 6 * sizeof(int32_t)));
    v3_gather_late<<<(n + 255) / 256, 256, 0, st>>>(n, flag, b.elem_list, b.frec, b.nrec, late->elem_list, late->frec,
                                                    late->nrec);
    D4B_CUDA(cudaGetLastError());
    D4B_CUDA(cudaStreamSynchronize(st));
    late->count = n;
  }
  cudaFree(flag);
  return D4B_OK;
}

/* bucket made of the slots `list` of b (host list, kept in its order) */
static int v3_subset(const d4b_fast_bucket &b, const std::vector<int32_t> &list, d4b_fast_bucket *o, cudaStream_t st)
{
  o->N = b.N; o->Nq = b.Nq; o->count = 0;
  o->elem_list = o->frec = o->nrec = nullptr; o->v3_slots = nullptr;
  const int n = (int)list.size();
  if (!n) return D4B_OK;
  int32_t *dl = nullptr;
  D4B_CUDA(cudaMalloc((void **)&dl, (size_t)n * sizeof(int32_t)));
  D4B_CUDA(cudaMemcpyAsync(dl, list.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  D4B_CUDA(cudaMalloc((void **)&o->elem_list, (size_t)n * sizeof(int32_t)));
  D4B_CUDA(cudaMalloc((void **)&o->frec, (size_t)n * 6 * sizeof(int32_t)));
  D4B_CUDA(cudaMalloc((void **)&o->nrec, (size_t)n * 6 * sizeof(int32_t)));
  v3_gather_late<<<(n + 255) / 256, 256, 0, st>>>(n, dl, b.elem_list, b.frec, b.nrec, o->elem_list, o->frec, o->nrec);
  D4B_CUDA(cudaGetLastError());
  D4B_CUDA(cudaStreamSynchronize(st));
  cudaFree(dl);
  o->count = n;
  return D4B_OK;
}

/* For the halo / compute overlap of a distributed apply: split b into the slots whose aligned group of `group`
 * slots (the warp batch of the kernel that will run them, so the batches -- Morton families -- stay intact and
 * keep their internal faces) touches no ghost element, and the rest. */
int d4b_v3_split_ghost(const d4b_mesh *m, const d4b_fast_bucket &b, int group, d4b_fast_bucket *early,
                       d4b_fast_bucket *late, cudaStream_t st)
{
  int rc = v3_prepare_slots(m, b, st);
  if (rc != D4B_OK) return rc;
  int32_t *flag = nullptr;
  D4B_CUDA(cudaMalloc((void **)&flag, (size_t)b.count * sizeof(int32_t)));
  const int one[2] = {0, b.count};
  v3_mark_late<<<(b.count + 255) / 256, 256, 0, st>>>((const v3_slot *)b.v3_slots, b.count, v3_make_bounds(1, one),
                                                      (int64_t)b.N * b.N * b.N, 1 << 30, flag);
  D4B_CUDA(cudaGetLastError());
  std::vector<int32_t> h(b.count), le, ll;
  D4B_CUDA(cudaMemcpyAsync(h.data(), flag, (size_t)b.count * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  D4B_CUDA(cudaStreamSynchronize(st));
  cudaFree(flag);
  if (group < 1) group = 1;
  for (int g0 = 0; g0 < b.count; g0 += group) {
    const int g1 = g0 + group < b.count ? g0 + group : b.count;
    bool any = false;
    for (int s = g0; s < g1; s++) any = any || h[s];
    for (int s = g0; s < g1; s++) (any ? ll : le).push_back(s);
  }
  if ((rc = v3_subset(b, le, early, st)) != D4B_OK) return rc;
  return v3_subset(b, ll, late, st);
}

template <int N, int Nq>
static int v3_launch_t(const d4b_mesh *m, const d4b_fast_bucket &b, const double *u, double *Au, cudaStream_t st,
                       int slot_begin, int slot_end)
{
  using L = V3<N, Nq>;
  static bool attr = false;
  if (!attr) {
    D4B_CUDA(cudaFuncSetAttribute(d4b_apply_v3_kernel<N, Nq>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::SMEM));
    attr = true;
  }
  { int rcs = v3_prepare_slots(m, b, st); if (rcs != D4B_OK) return rcs; }
  v3_args A;
  A.slots = (const v3_slot *)b.v3_slots; A.tasks = reinterpret_cast<const int64_t *>(A.slots + b.count); A.G = m->G; A.fac = m->fac; A.bc_g = m->bc_g;
  A.u = u; A.u_ghost = m->u_ghost; A.Au = Au; A.n_list = slot_end; A.slot_begin = slot_begin;
  static int n_sm = 0;
  if (!n_sm) { int dev = 0; D4B_CUDA(cudaGetDevice(&dev)); D4B_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)); }
  const int per = L::EPW * L::WPB;
  int grid = (slot_end - slot_begin + per - 1) / per;
  if (grid <= 0) return D4B_OK;
  /* one batch per warp by default: the warps of a CTA then run in phase and share the instruction
   * cache (the kernel is ~100 KB of SASS); D4B_V3_PERSISTENT=1 caps the grid at one CTA per SM */
  static int persist = -1;
  if (persist < 0) persist = getenv("D4B_V3_PERSISTENT") != nullptr;
  if (persist && grid > n_sm * L::CTAS) grid = n_sm * L::CTAS;
  d4b_apply_v3_kernel<N, Nq><<<grid, L::THREADS, L::SMEM, st>>>(A);
  D4B_CUDA(cudaGetLastError());
  d4b_count_launch(1);
  return D4B_OK;
}

int d4b_launch_v3_range(const d4b_mesh *m, const d4b_fast_bucket &b, const double *u, double *Au, cudaStream_t st,
                        int slot_begin, int slot_end)
{
  int rc = v3_upload_tab(m->quad_type, b.N, b.Nq, st);
  if (rc != D4B_OK) return rc;
#define X(a, c) if (b.N == a && b.Nq == c) return v3_launch_t<a, c>(m, b, u, Au, st, slot_begin, slot_end);
  V3_PAIRS(X)
#undef X
  d4b_set_error("d4b_launch_v3: unsupported (N,Nq)"); return D4B_EINVAL;
}

int d4b_launch_v3(const d4b_mesh *m, const d4b_fast_bucket &b, const double *u, double *Au, cudaStream_t st)
{ return d4b_launch_v3_range(m, b, u, Au, st, 0, b.count); }

int d4b_v3_elems_per_cta(int N, int Nq)
{
#define X(a, c) if (N == a && Nq == c) return V3<a, c>::EPW * V3<a, c>::WPB;
  V3_PAIRS(X)
#undef X
  return 32;
}
