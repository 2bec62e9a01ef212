"""CPU restatement (numpy) of d4est's matrix-free SIPG Laplacian apply on the flat tables.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg -- never by the product path (d4est_b200/ fails loudly without CUDA).

Parity is PINNED: tests/test_oracle.py checks this file against Au produced by the
unmodified reference (oracle/_ref/d4b_harness running d4est_laplacian_apply_aij and
d4est_laplacian_with_opt_apply_aij) on the committed fixtures under tests/golden/, and
against the reference unit test's golden sum (Au sum = 9.04698104235131).

Each function cites the reference file:line (under the d4est tree) it restates.  1-D tables
are rebuilt here with numpy.polynomial (independent of d4est_b200/csrc/d4b_ops.c).
"""
import functools
import numpy as np
from numpy.polynomial import legendre as L


# ---------------------------------------------------------------- 1-D tables
@functools.lru_cache(None)
def gll(deg):
    """GLL nodes/weights (dGMath/d4est_operators.c:727-829 tabulated values)."""
    n = deg
    c = np.zeros(n + 1); c[n] = 1.0
    xi = np.sort(np.real(L.legroots(L.legder(c)))) if n > 1 else np.array([])
    x = np.concatenate([[-1.0], xi, [1.0]])
    for _ in range(3):  # Newton polish on P_n'
        d1 = L.legval(x[1:-1], L.legder(c)); d2 = L.legval(x[1:-1], L.legder(c, 2))
        x[1:-1] -= d1 / d2
    x = 0.5 * (x - x[::-1])
    w = 2.0 / (n * (n + 1) * L.legval(x, c) ** 2)
    return x, w


@functools.lru_cache(None)
def quad_rule(degq, quad_type):
    """Quadrature/d4est_quadrature_legendre.c:23-92 (Gauss) / _lobatto.c (GLL)."""
    if quad_type == 1:
        return gll(degq)
    x, w = L.leggauss(degq + 1)
    return x, w


def _lagrange(xn, y):
    """M[q, j] = l_j(y_q) on nodes xn (explicit product form)."""
    M = np.ones((len(y), len(xn)))
    for j in range(len(xn)):
        for k in range(len(xn)):
            if k != j:
                M[:, j] *= (y - xn[k]) / (xn[j] - xn[k])
    return M


@functools.lru_cache(None)
def diff_matrix(deg):
    """D = V_r V^-1 (dGMath/d4est_operators.c:855-872): D[i, j] = l_j'(x_i)."""
    x, _ = gll(deg)
    N = deg + 1
    D = np.zeros((N, N))
    for i in range(N):
        for j in range(N):
            s = 0.0
            for m in range(N):
                if m == j:
                    continue
                p = 1.0 / (x[j] - x[m])
                for k in range(N):
                    if k != j and k != m:
                        p *= (x[i] - x[k]) / (x[j] - x[k])
                s += p
            D[i, j] = s
    return D


@functools.lru_cache(None)
def interp(deg, degq, quad_type, child=-1):
    """GLL(deg) -> quadrature points (dGMath/d4est_operators.c:411-438); child 0/1 =
    composition with the hp-prolongation to the lower/upper half (:944-993, reference.c:36-47)."""
    x, _ = gll(deg)
    y, _ = quad_rule(degq, quad_type)
    if child == 0:
        y = 0.5 * y - 0.5
    elif child == 1:
        y = 0.5 * y + 0.5
    return _lagrange(x, y)


# ---------------------------------------------------------------- helpers
def face_index(N, f):
    """Volume node ids of face f in face-local order (Kron/d4est_kron.h:583-636 slicer)."""
    idx = np.arange(N ** 3).reshape(N, N, N)  # [t, s, r]
    layer = N - 1 if (f & 1) else 0
    d = f >> 1
    if d == 0:
        return idx[:, :, layer].reshape(-1)  # (t, s): s fastest
    if d == 1:
        return idx[:, layer, :].reshape(-1)  # (t, r): r fastest
    return idx[layer, :, :].reshape(-1)      # (s, r): r fastest


def tcode_perm(tcode, n):
    """m-order point -> p-order point (dGMath/d4est_operators.c:1993-2087)."""
    b, a = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    x, y = (a, b) if (tcode & 1) else (b, a)
    if tcode & 2:
        x = n - 1 - x
    if tcode & 4:
        y = n - 1 - y
    return (x + n * y).reshape(-1)


def grad_ref(u3, D):
    """dudr_l = D_l u (dGMath/d4est_operators.c:1385-1406); u3 indexed [t, s, r]."""
    return [np.einsum("ij,tsj->tsi", D, u3), np.einsum("ij,tjr->tir", D, u3), np.einsum("ij,jsr->isr", D, u3)]


def _face_fields(u_el, N, f, D):
    u3 = u_el.reshape(N, N, N)
    g = grad_ref(u3, D)
    fi = face_index(N, f)
    return np.stack([u_el[fi]] + [gl.reshape(-1)[fi] for gl in g])  # [4, N^2]


def _E(deg, degq, qt, child):
    ca = -1 if child < 0 else (child & 1)
    cb = -1 if child < 0 else ((child >> 1) & 1)
    return interp(deg, degq, qt, ca), interp(deg, degq, qt, cb)


# ---------------------------------------------------------------- the apply
def twin_records(fx):
    """twin[r] = the record of the neighbour that shares the mortar of record r with the roles swapped (same faces,
    child indices exchanged, same mortar degree), or -1 (boundary, ghost neighbour)."""
    nl = int(fx["hdr"][0]); hm, hb = fx["hm"], fx["hm_begin"]
    key = {}
    for e in range(nl):
        for r in range(hb[e], hb[e + 1]):
            h = hm[r]
            key[(e, int(h["nbr"]), int(h["f_m"]), int(h["f_p"]), int(h["child_m"]), int(h["child_p"]), int(h["degq"]))] = r
    twin = np.full(int(hb[nl]), -1, np.int64)
    for e in range(nl):
        for r in range(hb[e], hb[e + 1]):
            h = hm[r]; p = int(h["nbr"])
            if 0 <= p < nl:
                twin[r] = key.get((p, e, int(h["f_p"]), int(h["f_m"]), int(h["child_p"]), int(h["child_m"]), int(h["degq"])), -1)
    return twin


def own_traces(fx, u):
    """Pass 1 of the two-pass form (d4b_tp_trace_kernel): per record the own-side trace at the mortar points,
    (u-, sum_l a-_l du-/dr_l), in the element's own point order."""
    nl, _, qt, _ = [int(v) for v in fx["hdr"]]
    deg, node_off, hm, hb, fac = fx["deg"], fx["node_off"], fx["hm"], fx["hm_begin"], fx["fac"]
    W = [None] * int(hb[nl])
    for e in range(nl):
        dg = int(deg[e]); N = dg + 1; D = diff_matrix(dg)
        ue = u[node_off[e]:node_off[e + 1]]
        for r in range(hb[e], hb[e + 1]):
            h = hm[r]; mdq = int(h["degq"]); Nm = mdq + 1; n2 = Nm * Nm
            Ea, Eb = _E(dg, mdq, qt, int(h["child_m"]))
            fm = _face_fields(ue, N, int(h["f_m"]), D)
            mq = np.einsum("qa,pb,cba->cpq", Ea, Eb, fm.reshape(4, N, N)).reshape(4, n2)
            am = fac[h["fac_off"]:][:3 * n2].reshape(3, n2)
            W[r] = (mq[0], (am * mq[1:]).sum(0))
    return W


def apply(fx, u, u_ghost=None, twin_traces=False):
    """Au = A u.  Restates d4est_laplacian_with_opt_apply_aij
    (dGMath/d4est_laplacian_with_opt.c:289-368) in the one-side-at-a-time form of its twin
    d4est_laplacian_apply_aij (dGMath/d4est_laplacian.c:317; the two agree to 1e-13,
    Tests/Unit/d4est_test_laplacian_speedup.c:485).
    twin_traces=True: the neighbour side of an interior record with a local neighbour is taken from the own-side
    trace of the neighbour's twin record, u+ = W[twin].u, sum a+ du+ = -W[twin].q (reordered by the orientation
    code) -- the formulation of the two-pass device kernels; a+ = -a-(twin) holds at the same physical points."""
    nl, ng, qt, _ = [int(v) for v in fx["hdr"]]
    twin = twin_records(fx) if twin_traces else None
    Wtr = own_traces(fx, u) if twin_traces else None
    deg, degq, node_off, quad_off = fx["deg"], fx["degq"], fx["node_off"], fx["quad_off"]
    hm, hb, fac, bcg, G = fx["hm"], fx["hm_begin"], fx["fac"], fx["bc_g"], fx["G"]
    local_nodes = int(node_off[nl])
    uext = u if u_ghost is None or ng == 0 else np.concatenate([u, u_ghost])
    Au = np.zeros(local_nodes)
    for e in range(nl):
        dg, dq = int(deg[e]), int(degq[e])
        N, Nq = dg + 1, dq + 1
        D, I = diff_matrix(dg), interp(dg, dq, qt)
        ue = u[node_off[e]:node_off[e + 1]]
        # volume: Quadrature/d4est_quadrature.c:263-375
        g = grad_ref(ue.reshape(N, N, N), D)
        gq = [np.einsum("at,bs,cr,tsr->abc", I, I, I, gl).reshape(-1) for gl in g]
        Ge = G[6 * quad_off[e]:6 * quad_off[e + 1]].reshape(6, Nq ** 3)
        sym = [[0, 1, 2], [1, 3, 4], [2, 4, 5]]
        W = []
        for l in range(3):
            fl = sum(Ge[sym[l][k]] * gq[k] for k in range(3)).reshape(Nq, Nq, Nq)
            W.append(np.einsum("at,bs,cr,abc->tsr", I, I, I, fl).reshape(-1))
        acc = np.zeros(N ** 3)
        # faces: dGMath/d4est_laplacian_with_opt_flux.c:212-900 + …_flux_sipg.c:529-1251
        for r in range(hb[e], hb[e + 1]):
            h = hm[r]
            mdq = int(h["degq"]); Nm = mdq + 1; n2 = Nm * Nm
            _, wq = quad_rule(mdq, qt)
            w2 = np.outer(wq, wq).reshape(-1)
            Ea, Eb = _E(dg, mdq, qt, int(h["child_m"]))
            fm = _face_fields(ue, N, int(h["f_m"]), D)
            mq = np.einsum("qa,pb,cba->cpq", Ea, Eb, fm.reshape(4, N, N)).reshape(4, n2)
            F = fac[h["fac_off"]:]
            am = F[:3 * n2].reshape(3, n2)
            qm = (am * mq[1:]).sum(0)
            if h["nbr"] >= 0 and twin_traces and twin[r] >= 0:
                perm = tcode_perm(int(h["tcode"]), Nm)
                sjs = F[6 * n2:7 * n2]
                jump = mq[0] - Wtr[twin[r]][0][perm]
                t1 = -0.5 * (qm - Wtr[twin[r]][1][perm]); hs = -0.5
            elif h["nbr"] >= 0:
                nb = int(h["nbr"]); dp = int(deg[nb]); Np = dp + 1
                un = uext[node_off[nb]:node_off[nb + 1]]
                Pa, Pb = _E(dp, mdq, qt, int(h["child_p"]))
                fp = _face_fields(un, Np, int(h["f_p"]), diff_matrix(dp))
                pq = np.einsum("qa,pb,cba->cpq", Pa, Pb, fp.reshape(4, Np, Np)).reshape(4, n2)
                pq = pq[:, tcode_perm(int(h["tcode"]), Nm)]
                ap = F[3 * n2:6 * n2].reshape(3, n2); sjs = F[6 * n2:7 * n2]
                qp = (ap * pq[1:]).sum(0)
                jump = mq[0] - pq[0]
                t1 = -0.5 * (qm + qp); hs = -0.5         # sipg.c:651-675
            else:
                sjs = F[3 * n2:4 * n2]
                gb = bcg[h["g_off"]:h["g_off"] + n2] if h["g_off"] >= 0 else 0.0
                jump = mq[0] - gb
                t1 = -qm; hs = -1.0                       # sipg.c:139-191 (Dirichlet)
            if h["nbr"] < 0 and h["pad"] == 1:            # Robin, sipg.c:377-476: sj (alpha u - beta), no gradient terms
                FH = np.stack([w2 * (sjs * mq[0] - gb)] + [np.zeros(n2)] * 3)
            else:
                FH = np.stack([w2 * (t1 + sjs * jump)] + [w2 * hs * am[l] * jump for l in range(3)])
            back = np.einsum("qa,pb,cpq->cba", Ea, Eb, FH.reshape(4, Nm, Nm)).reshape(4, N * N)
            fi = face_index(N, int(h["f_m"]))
            acc[fi] += back[0]
            for l in range(3):
                W[l][fi] += back[1 + l]
        W3 = [w.reshape(N, N, N) for w in W]
        out = (np.einsum("ij,tsi->tsj", D, W3[0]) + np.einsum("ij,tir->tjr", D, W3[1]) +
               np.einsum("ij,isr->jsr", D, W3[2])).reshape(-1)
        Au[node_off[e]:node_off[e + 1]] = acc + out
    return Au


def mass_apply(fx, coef, v):
    """out = I^T diag(w (x) w (x) w * c) I v per element, c = J f(x, u0) at the volume quadrature points: what
    d4est_quadrature_apply_fofufofvlilj computes element by element (Quadrature/d4est_quadrature.c:593-760, ending in
    d4est_quadrature_apply_mass_matrix).  Pinned against the reference's own output (fixture key mass_ref)."""
    nl, _, qt, _ = [int(x) for x in fx["hdr"]]
    no, qo = fx["node_off"], fx["quad_off"]
    out = np.zeros(int(no[nl]))
    for e in range(nl):
        dg, dq = int(fx["deg"][e]), int(fx["degq"][e]); N, Nq = dg + 1, dq + 1
        I = interp(dg, dq, qt); _, w = quad_rule(dq, qt)
        vq = np.einsum("at,bs,cr,tsr->abc", I, I, I, v[no[e]:no[e + 1]].reshape(N, N, N))
        c = coef[qo[e]:qo[e + 1]].reshape(Nq, Nq, Nq) * np.einsum("a,b,c->abc", w, w, w)
        out[no[e]:no[e + 1]] = np.einsum("at,bs,cr,abc->tsr", I, I, I, c * vq).reshape(-1)
    return out
