"""CPU restatement (numpy) of the solver pieces around the apply.  TEST INFRASTRUCTURE ONLY.

PARITY PINNED against the unmodified reference for every function (tests/test_oracle.py):
  cg_eigs, cheby_iterate   fixtures hold sol_eig / sol_u_after_eigs / sol_u_after_cheby / sol_r_after_cheby written by
                           oracle/_ref/d4b_harness calling the reference's cg_eigs and ..._smoother_cheby_iterate_aux
  fcg_solve, cg_solve      the reference's d4est_solver_fcg_solve / d4est_solver_cg_solve run for a fixed 25
                           iterations on the curved hp mesh (krylov_fcg25_u / krylov_cg25_u in the multigrid
                           fixture written by oracle/_ref/d4b_mg_harness; the drivers compile against typedef-only
                           PETSc/HDF5 headers, oracle/ref_build/include/)
  transfer                 the reference's d4est_operators_apply_{p,hp}_prolong[_transpose] on probe vectors over
                           the hierarchy captured from its own coarsening (T*.ref_restricted / T*.ref_prolonged)
  mg_vcycle / mg_pc        the reference's d4est_krylov_pc_multigrid_apply = d4est_solver_multigrid_solve with one
                           V-cycle: first preconditioner application of its MG-FCG solve (mg_r0 -> mg_z0)
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


def fcg_solve(A, u, rhs, imax, atol, rtol, pc=None, textbook=False):
    """Solver/d4est_solver_fcg_improved.c:97-346, INCLUDING what its direction update really computes:
    d4est_linalg_vec_axpyeqz(a, x, y, z) copies y into z and then adds a*x (LinearAlgebra/d4est_linalg.c:217-221),
    and the driver calls it with z aliasing x (:262 `axpyeqz(-gamma/rho, d, v, d)`, :266 same for q), so the old
    direction is overwritten before it is used: d_k = (1 - gamma/rho) v_k, q_k = (1 - gamma/rho) w_k.  That is the
    arithmetic the reference's iteration counts come from (pinned by krylov_fcg25_u); textbook=True gives the
    update the comments in the source describe (d_k = v_k - gamma/rho d_{k-1})."""
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
            if textbook:
                d = -gamma / rho * d + v
                q = -gamma / rho * q + w
            else:
                d = (1. - gamma / rho) * v
                q = (1. - gamma / rho) * w
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


def amr_sweep_table(deg_in, code):
    """Transfer table of one sweep of d4est_amr_interpolate_field (hpAMR/d4est_amr.c:406-485): element i of degree
    |deg_in[i]| with code[i] < 0 -> 8 children of the same degree (hp-prolongation, :366), code[i] >= deg ->
    p-prolongation (:379); coarsening aborts in the reference (:395)."""
    n = len(deg_in)
    tab = {"kind": np.zeros(n, np.int32), "degH": np.zeros(n, np.int32), "coarse_off": np.zeros(n, np.int64),
           "child_deg": np.zeros(8 * n, np.int32), "child_off": np.zeros(8 * n, np.int64)}
    ni = no = 0
    for e in range(n):
        d, c = abs(int(deg_in[e])), int(code[e])
        tab["degH"][e] = d; tab["coarse_off"][e] = ni; ni += (d + 1) ** 3
        if c < 0:
            assert -c == d, "h-refinement keeps the degree"
            tab["kind"][e] = 1
            for k in range(8):
                tab["child_deg"][8 * e + k] = d; tab["child_off"][8 * e + k] = no; no += (d + 1) ** 3
        else:
            assert c >= d, "coarsening is not supported"
            tab["child_deg"][8 * e] = c; tab["child_off"][8 * e] = no; no += (c + 1) ** 3
    return tab, ni, no


def amr_interpolate_field(initial_log, refinement_log, balance_log, field):
    """d4est_amr_interpolate_field: initial grid -> refined grid -> 2:1-balanced grid."""
    t1, n_in, n_aux = amr_sweep_table(initial_log, refinement_log)
    assert n_in == field.size
    aux = transfer(t1, field, n_aux, restrict=False)
    t2, n_aux2, n_out = amr_sweep_table(balance_log, balance_log)
    assert n_aux2 == n_aux
    return transfer(t2, aux, n_out, restrict=False)


def brick_transfer_table(n_coarse_per_dir, deg):
    """Transfer table between Morton-ordered uniform bricks 2n -> n with equal degree: the 8 children
    of coarse element e are fine elements 8e .. 8e+7 (child c = bits r,s,t = Morton order)."""
    nH = n_coarse_per_dir ** 3; N3 = (deg + 1) ** 3
    return {"kind": np.ones(nH, np.int32), "degH": np.full(nH, deg, np.int32),
            "coarse_off": (np.arange(nH, dtype=np.int64) * N3),
            "child_deg": np.full(8 * nH, deg, np.int32),
            "child_off": (np.arange(8 * nH, dtype=np.int64) * N3)}


def mg_params(pd):
    """d4b_mg_params_t fields from the "mg_params" array of a multigrid fixture (order: d4b_mg_harness.c)."""
    keys = ["vcycle_imax", "vcycle_rtol", "vcycle_atol", "cheby_imax", "cheby_eigs_cg_imax", "cheby_eigs_lmax_lmin_ratio",
            "cheby_eigs_max_multiplier", "eigs_reuse_fromdownvcycle", "eigs_reuse_fromlastvcycle", "use_new_cg_eigs",
            "bottom_kind", "bottom_imax", "bottom_eigs_cg_imax", "bottom_eigs_lmax_lmin_ratio",
            "bottom_eigs_max_multiplier", "bottom_atol", "bottom_rtol", "fcg_atol", "fcg_rtol", "ref_seconds"]
    ints = {"vcycle_imax", "cheby_imax", "cheby_eigs_cg_imax", "eigs_reuse_fromdownvcycle", "eigs_reuse_fromlastvcycle",
            "use_new_cg_eigs", "bottom_kind", "bottom_imax", "bottom_eigs_cg_imax"}
    return {k: (int(v) if k in ints else float(v)) for k, v in zip(keys, pd)}


def mg_vcycle(A, tabs, sizes, u, rhs, par, eigs, vcycles_done=0):
    """One V-cycle: Solver/d4est_solver_multigrid.c:751-1330.  A[l]: operator of level l (0 = coarsest),
    tabs[l]: transfer table between l and l+1, sizes[l]: nodes of level l.  Smoother = cg_eigs (when the reuse
    flags ask for it, ..._smoother_cheby.c:221-262, 264-360) + Chebyshev; bottom = ..._bottom_solver_cheby.c:59-113.
    Returns (u, r_top)."""
    top = len(A) - 1
    us, fs, rs = [None] * (top + 1), [None] * (top + 1), [None] * (top + 1)
    us[top], fs[top] = u, rhs

    def smooth(l, down):
        if down:
            compute = not (par["eigs_reuse_fromlastvcycle"] and vcycles_done != 0)
        else:
            compute = not (par["eigs_reuse_fromdownvcycle"] or (par["eigs_reuse_fromlastvcycle"] and vcycles_done != 0))
        if compute:
            eigs[l], us[l] = cg_eigs(A[l], us[l], fs[l], par["cheby_eigs_cg_imax"], bool(par["use_new_cg_eigs"]))
            eigs[l] *= par["cheby_eigs_max_multiplier"]
        us[l], rs[l] = cheby_iterate(A[l], us[l], fs[l], par["cheby_imax"], eigs[l] / par["cheby_eigs_lmax_lmin_ratio"], eigs[l])

    for l in range(top, 0, -1):
        if l != top:
            us[l] = np.zeros(sizes[l])
        smooth(l, True)
        fs[l - 1] = transfer(tabs[l - 1], rs[l], sizes[l - 1], True)
    us[0] = np.zeros(sizes[0])
    if par["bottom_kind"] == 1:
        us[0], _, _ = cg_solve(A[0], us[0], fs[0], par["bottom_imax"], par["bottom_atol"], par["bottom_rtol"])
    else:
        e, us[0] = cg_eigs(A[0], us[0], fs[0], par["bottom_eigs_cg_imax"], bool(par["use_new_cg_eigs"]))
        e *= par["bottom_eigs_max_multiplier"]; eigs[0] = e
        us[0], rs[0] = cheby_iterate(A[0], us[0], fs[0], par["bottom_imax"], e / par["bottom_eigs_lmax_lmin_ratio"], e)
    for l in range(top):
        us[l + 1] = us[l + 1] + transfer(tabs[l], us[l], sizes[l + 1], False)
        smooth(l + 1, False)
    return us[top], rs[top]


def mg_pc(A, tabs, sizes, r, par):
    """z = MG(r) from a zero guess with vcycle_imax cycles: Solver/d4est_krylov_pc_multigrid.c:41-75 +
    d4est_solver_multigrid_solve (multigrid.c:1420-1500)."""
    z = np.zeros(sizes[-1]); eigs = [0.] * len(A)
    r2 = r2_last = float(r @ r)
    stop = par["vcycle_rtol"] ** 2 * r2 + par["vcycle_atol"] ** 2
    done = 0
    while done < par["vcycle_imax"] and r2 > stop:
        z, rt = mg_vcycle(A, tabs, sizes, z, r, par, eigs, done)
        r2 = float(rt @ rt); done += 1
        if np.sqrt(r2 / r2_last) >= .99:
            break
        r2_last = r2
    return z
