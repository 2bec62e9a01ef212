"""CPU restatement (numpy) of the solver pieces around the apply.  TEST INFRASTRUCTURE ONLY.

PINNED for cg_eigs and the Chebyshev iteration against the unmodified reference (the fixtures hold
sol_eig / sol_u_after_eigs / sol_u_after_cheby / sol_r_after_cheby written by oracle/_ref/d4b_harness
calling the reference's cg_eigs and d4est_solver_multigrid_smoother_cheby_iterate_aux).  The flexible
CG, plain CG and the V-cycle are restated from source ("parity unpinned" for those three: the
reference's own drivers need PETSc headers / a live p4est hierarchy that the serial oracle build does
not have; they are pinned only through their building blocks above).
`A` is any callable u -> A u (normally lambda u: dg_oracle.apply(fx, u)).
"""
import numpy as np
import dg_oracle


def cheby_iterate(A, u, rhs, it, lmin, lmax, residual_at_end=True):
    """Solver/d4est_solver_multigrid_smoother_cheby.c:81-176."""
    d, c = (lmax + lmin) * .5, (lmax - lmin) * .5
    p = np.zeros_like(u)
    alpha = 0.0
    for i in range(it):
        r = rhs - A(u)
        if i == 0:
            alpha = 1. / d
        elif i == 1:
            alpha = 2. * d / (2 * d * d - c * c)
        else:
            alpha = 1. / (d - (alpha * c * c / 4.))
        beta = alpha * d - 1.
        p = alpha * r + beta * p
        u = u + p
    r = rhs - A(u) if residual_at_end else None
    return u, r


def _gershgorin(i, n, a0, b0, a1, b1, use_new):
    """Solver/d4est_solver_cg_eigs.c:7-59."""
    if not use_new:
        if i != 0 and i < n - 1:
            return 1. / a1 + b0 / a0 + abs(np.sqrt(b1) / a1) + abs(np.sqrt(b0) / a0)
        if i == 0:
            return 1. / a1 + np.sqrt(b1) / a1
        return 1. / a1 + b0 / a0 + abs(np.sqrt(b0) / a0)
    if i != 0:
        return 1. / a1 + b0 / a0 + abs(np.sqrt(b0) / a0)
    return 1. / a1 + np.sqrt(b1) / a1


def cg_eigs(A, u, rhs, imax, use_new=False):
    """Solver/d4est_solver_cg_eigs.c:115-276; returns (bound, u advanced by the CG steps)."""
    r = rhs - A(u)
    d = r.copy()
    delta_new = r @ r
    alpha = beta = -1.
    bound = None
    for i in range(imax):
        Ad = A(d)
        alpha_old, alpha = alpha, delta_new / (d @ Ad)
        u = u + alpha * d
        r = r - alpha * Ad
        delta_old, delta_new = delta_new, r @ r
        beta_old, beta = beta, delta_new / delta_old
        d = r + beta * d
        t = _gershgorin(i, u.size, alpha_old, beta_old, alpha, beta, use_new)
        bound = t if i == 0 else max(bound, t)
    return bound, u


def cg_solve(A, u, rhs, imax, atol, rtol):
    """Solver/d4est_solver_cg.c:76-190."""
    r = rhs - A(u); d = r.copy(); dn = r @ r; d0 = dn; i = 0
    while i < imax and dn > atol * atol + d0 * rtol * rtol:
        Ad = A(d); alpha = dn / (d @ Ad)
        u = u + alpha * d; r = r - alpha * Ad
        do, dn = dn, r @ r
        d = r + (dn / do) * d
        i += 1
    return u, i, dn


def fcg_solve(A, u, rhs, imax, atol, rtol, pc=None):
    """Solver/d4est_solver_fcg_improved.c:97-346."""
    r = rhs - A(u)
    tol = atol + rtol * np.sqrt(r @ r)
    rho = None
    k = 0
    for k in range(imax):
        v = pc(r) if pc is not None else r.copy()
        w = A(v)
        alpha, beta = v @ r, v @ w
        if k > 0:
            gamma, rr = v @ q, r @ r
            d = -gamma / rho * d + v
            q = -gamma / rho * q + w
            rho = beta - gamma * gamma / rho
        else:
            rho, d, q = beta, v.copy(), w.copy()
        u = u + alpha / rho * d
        r = r - alpha / rho * q
        if k > 0 and np.sqrt(rr) <= tol:
            break
    return u, k


def prolong_matrix(degH, degh, child=-1):
    """dGMath/d4est_operators.c:995 (p) and :944-993 (hp, child 0/1)."""
    xH, _ = dg_oracle.gll(degH); xh, _ = dg_oracle.gll(degh)
    y = xh if child < 0 else (0.5 * xh - 0.5 if child == 0 else 0.5 * xh + 0.5)
    return dg_oracle._lagrange(xH, y)


def transfer(tab, src, n_dst, restrict):
    """Solver/d4est_solver_multigrid_callbacks.h:109 (restriction = P^T) / :236 (prolongation = P)."""
    dst = np.zeros(n_dst)
    for e in range(len(tab["kind"])):
        dH = int(tab["degH"][e]); NH = dH + 1; co = int(tab["coarse_off"][e])
        nchild = 8 if tab["kind"][e] == 1 else 1
        for c in range(nchild):
            dh = int(tab["child_deg"][8 * e + c]); Nh = dh + 1; fo = int(tab["child_off"][8 * e + c])
            v = (lambda b: -1) if nchild == 1 else (lambda b: (c >> b) & 1)
            Px, Py, Pz = (prolong_matrix(dH, dh, v(b)) for b in range(3))
            if restrict:
                xh = src[fo:fo + Nh ** 3].reshape(Nh, Nh, Nh)
                dst[co:co + NH ** 3] += np.einsum("kK,jJ,iI,kji->KJI", Pz, Py, Px, xh).reshape(-1)
            else:
                xH = src[co:co + NH ** 3].reshape(NH, NH, NH)
                dst[fo:fo + Nh ** 3] = np.einsum("kK,jJ,iI,KJI->kji", Pz, Py, Px, xH).reshape(-1)
    return dst


def brick_transfer_table(n_coarse_per_dir, deg):
    """Transfer table between Morton-ordered uniform bricks 2n -> n with equal degree: the 8 children
    of coarse element e are fine elements 8e .. 8e+7 (child c = bits r,s,t = Morton order)."""
    nH = n_coarse_per_dir ** 3; N3 = (deg + 1) ** 3
    return {"kind": np.ones(nH, np.int32), "degH": np.full(nH, deg, np.int32),
            "coarse_off": (np.arange(nH, dtype=np.int64) * N3),
            "child_deg": np.full(8 * nH, deg, np.int32),
            "child_off": (np.arange(8 * nH, dtype=np.int64) * N3)}
