"""Shared-memory bank-conflict model of d4b_apply_fast_kernel<N,Nq> (d4b_apply_fast.cu).

Every shared access of the kernel is listed in accesses() as the address each thread touches; the model groups
the CTA's threads into half-warps (a 64-bit access is served one half-warp per wavefront when its 16 addresses
fall in distinct 8-byte banks) and counts wavefronts against the ideal.  Used to choose the paddings of
struct fast_pad (d4b_fast_pads.inc).
usage: fast_bank_sim.py N Nq            print the conflicting stages of the current layout
       fast_bank_sim.py N Nq search [r] coordinate descent over the paddings, r random restarts
"""
import ast, random, re, sys, os


class Layout:
    def __init__(s, N, Nq, SJ=None, AS=None, BS=None, TB=None, PERPAD=0, QP=None):
        s.N, s.Nq = N, Nq
        s.N2, s.Q2, s.NQ = N * N, Nq * Nq, N * Nq
        s.SJ = SJ if SJ is not None else N + 1
        s.SK = N * s.SJ
        s.UA = N * s.SK
        s.AS = AS if AS is not None else s.N2 + (4 if N == 4 else 0)
        s.AB = Nq * s.AS
        s.BS = BS if BS is not None else s.NQ + 1
        s.BB = Nq * s.BS
        s.TB = TB if TB is not None else s.N2 + 1
        s.QP = QP if QP is not None else s.Q2      # stride of a Q2 block in the face scratch
        VOL = 2 * s.AB + 3 * s.BB
        s.RA = max(8 * s.TB, 8 * s.QP, 6 * s.NQ)
        RB = max(12 * s.NQ, 8 * s.QP, 4 * s.N2)
        s.PER = 2 * s.UA + max(VOL, s.RA + RB) + PERPAD
        s.TPE = s.N2
        s.EPB = max(128 // s.TPE, 1)


def accesses(L):
    """list of (stage, f(lt) -> per loop trip, the sequence of element-relative addresses thread lt touches)"""
    N, Nq, N2, Q2, NQ = L.N, L.Nq, L.N2, L.Q2, L.NQ
    SJ, SK, AS, BS, TB, TPE, QP = L.SJ, L.SK, L.AS, L.BS, L.TB, L.TPE, L.QP
    U, ACC, S = 0, L.UA, 2 * L.UA
    V0, V1, V2 = S, S + L.AB, S + 2 * L.AB
    V3, V4 = V2 + L.BB, V2 + 2 * L.BB
    RA, RB = S, S + L.RA
    out = []
    add = lambda name, f: out.append((name, f))
    add("load U", lambda lt: [[U + (lt % N) + SJ * (lt // N) + SK * m for m in range(N)]])
    add("vol t ld", lambda lt: [[U + lt % N + SJ * (lt // N) + SK * m for m in range(N)]])
    add("vol t st", lambda lt: [[V + lt + AS * q for V in (V0, V1) for q in range(Nq)]])
    add("vol s ld", lambda lt: [[V + t % N + AS * (t // N) + N * m for V in (V0, V1) for m in range(N)] for t in range(lt, NQ, TPE)])
    add("vol s st", lambda lt: [[V + t % N + N * (t // N) + BS * q for V in (V2, V3, V4) for q in range(Nq)] for t in range(lt, NQ, TPE)])
    add("vol r ld/st", lambda lt: [[V + BS * (t % Nq) + N * (t // Nq) + m for V in (V2, V3, V4) for m in range(N)] * 2 for t in range(lt, Q2, TPE)])
    add("vol sT ld", lambda lt: [[V + t % N + N * (t // N) + BS * q for V in (V2, V3, V4) for q in range(Nq)] for t in range(lt, NQ, TPE)])
    add("vol sT st", lambda lt: [[V + t % N + AS * (t // N) + N * m for V in (V0, V1) for m in range(N)] for t in range(lt, NQ, TPE)])
    add("vol tT ld", lambda lt: [[V + lt + AS * q for V in (V0, V1) for q in range(Nq)]])
    add("vol tT st", lambda lt: [[ACC + lt % N + SJ * (lt // N) + SK * m for m in range(N)]])
    for d in range(3):
        sa = SJ if d == 0 else 1
        sb = SJ if d == 2 else SK
        sn = 1 if d == 0 else (SJ if d == 1 else SK)
        add(f"d{d} T ld", lambda lt, sa=sa, sb=sb, sn=sn: [[U + (lt % N) * sa + (lt // N) * sb + sn * m for m in range(N)]])
        add(f"d{d} T st", lambda lt: [[RA + k * TB + lt for k in range(8)]])
        add(f"d{d} S1 ld", lambda lt: [[RA + ((t // (2 * N)) * 2 + ((t // N) & 1)) * TB + N * (t % N) + m for m in range(N)] for t in range(lt, 8 * N, TPE)])
        add(f"d{d} S1 st", lambda lt: [[RB + (t // (2 * N)) * 3 * NQ + Nq * (t % N) + (2 * NQ if (t // N) & 1 else 0) + q for q in range(Nq)]
                                       + ([] if (t // N) & 1 else [RB + (t // (2 * N)) * 3 * NQ + Nq * (t % N) + NQ + q for q in range(Nq)])
                                       for t in range(lt, 8 * N, TPE)])
        add(f"d{d} QV ld", lambda lt: [[RB + (t // Nq) * 3 * NQ + t % Nq + k * NQ + Nq * m for k in range(3) for m in range(N)] for t in range(lt, 4 * Nq, TPE)])
        add(f"d{d} QV st", lambda lt: [[RA + (t // Nq) * 2 * QP + f * QP + t % Nq + Nq * qb for qb in range(Nq) for f in range(2)] for t in range(lt, 4 * Nq, TPE)])
        add(f"d{d} FLUX ld", lambda lt: [[RA + ((t // Q2) * 2 + f) * QP + t % Q2 for f in (0, 1)] + [RA + (4 + 2 * (t // Q2) + f) * QP + t % Q2 for f in (0, 1)] for t in range(lt, 2 * Q2, TPE)])
        add(f"d{d} FLUX st", lambda lt: [[RB + (t // Q2) * 4 * QP + t % Q2 + k * QP for k in range(4)] for t in range(lt, 2 * Q2, TPE)])

        def back1_ld(lt):
            trips = []
            for t in range(lt, 6 * Nq, TPE):
                qa, k, s = t % Nq, (t // Nq) % 3, t // (3 * Nq)
                fh = RB + s * 4 * QP + qa
                blocks = (0, 2) if k == 0 else ((1,) if k == 1 else (3,))
                trips.append([fh + b * QP + Nq * m for b in blocks for m in range(Nq)])
            return trips
        add(f"d{d} BACK1 ld", back1_ld)
        add(f"d{d} BACK1 st", lambda lt: [[RA + ((t // (3 * Nq)) * 3 + (t // Nq) % 3) * NQ + t % Nq + Nq * m for m in range(N)] for t in range(lt, 6 * Nq, TPE)])

        def back2_ld(lt):
            trips = []
            for t in range(lt, 4 * N, TPE):
                b, k, s = t % N, (t // N) & 1, t // (2 * N)
                r1 = RA + s * 3 * NQ + Nq * b
                trips.append([r1 + bb * NQ + q for bb in ((0, 1) if k == 0 else (2,)) for q in range(Nq)])
            return trips
        add(f"d{d} BACK2 ld", back2_ld)
        add(f"d{d} BACK2 st", lambda lt: [[RB + ((t // (2 * N)) * 2 + ((t // N) & 1)) * N2 + N * (t % N) + m for m in range(N)] for t in range(lt, 4 * N, TPE)])
        add(f"d{d} LIFT", lambda lt, sa=sa, sb=sb, sn=sn: [[RB + lt + k * N2 for k in range(4)] + [ACC + (lt % N) * sa + (lt // N) * sb + m * sn for m in range(N)] * 2])
    add("store ACC", lambda lt: [[ACC + (lt % N) + SJ * (lt // N) + SK * m for m in range(N)]])
    return out


def simulate(L, verbose=False):
    nthr = L.TPE * L.EPB
    tot = ideal = 0
    rows = []
    for name, f in accesses(L):
        per_thread = [f(t % L.TPE) for t in range(nthr)]
        ntrip = max(len(p) for p in per_thread)
        w = i = 0
        for trip in range(ntrip):
            ninst = max((len(p[trip]) for p in per_thread if len(p) > trip), default=0)
            for k in range(ninst):
                for h0 in range(0, nthr, 16):
                    banks = {}
                    for t in range(h0, min(h0 + 16, nthr)):
                        p = per_thread[t]
                        if len(p) > trip and len(p[trip]) > k:
                            a = (t // L.TPE) * L.PER + p[trip][k]
                            banks.setdefault(a % 16, set()).add(a)
                    if banks:
                        w += max(len(v) for v in banks.values())
                        i += 1
        rows.append((name, w, i))
        tot += w
        ideal += i
    if verbose:
        for name, w, i in rows:
            if w > i:
                print(f"  {name:14s} wavefronts {w:5d} ideal {i:5d}  x{w / i:.2f}")
    return tot, ideal


def current_pads(N, Nq):
    """the FAST_PAD line of (N, Nq) in d4b_fast_pads.inc, if any"""
    inc = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "d4est_b200", "csrc", "d4b_fast_pads.inc")
    if os.path.exists(inc):
        for m in re.finditer(r"^FAST_PAD\(([^)]*)\)", open(inc).read(), re.M):
            v = [int(x) for x in m.group(1).split(",")]
            if v[0] == N and v[1] == Nq:
                return dict(SJ=v[2], AS=v[3], BS=v[4], TB=v[5], QP=v[6], PERPAD=v[7])
    return {}


def search(N, Nq, restarts):
    rng = random.Random(1)
    space = dict(SJ=range(N, N + 4), AS=range(N * N, N * N + 9), BS=range(N * Nq, N * Nq + 9), TB=range(N * N, N * N + 6),
                 PERPAD=range(16), QP=range(Nq * Nq, Nq * Nq + 6))

    def ev(p):
        L = Layout(N, Nq, **p)
        t, i = simulate(L)
        return t, L.PER, i
    base = dict(SJ=N + 1, AS=N * N + (4 if N == 4 else 0), BS=N * Nq + 1, TB=N * N + 1, PERPAD=0, QP=Nq * Nq)
    best = (ev(base), base)
    print("default", best)
    for rnd in range(restarts):
        cur = dict(best[1]) if rnd == 0 else {k: rng.choice(list(v)) for k, v in space.items()}
        curv = ev(cur)
        improved = True
        while improved:
            improved = False
            for k in space:
                for v in space[k]:
                    if v == cur[k]:
                        continue
                    p = dict(cur); p[k] = v
                    e = ev(p)
                    if e[:2] < curv[:2]:
                        cur, curv, improved = p, e, True
        print("round", rnd, curv, cur, round(curv[0] / curv[2], 3), flush=True)
        if curv[:2] < best[0][:2]:
            best = (curv, cur)
    p = best[1]
    print(f"FAST_PAD({N}, {Nq}, {p['SJ']}, {p['AS']}, {p['BS']}, {p['TB']}, {p['QP']}, {p['PERPAD']}) "
          f"/* modelled wavefronts -> {best[0][0]} (ideal {best[0][2]}) */")


if __name__ == "__main__":
    N, Nq = int(sys.argv[1]), int(sys.argv[2])
    if len(sys.argv) > 3:
        search(N, Nq, int(sys.argv[4]) if len(sys.argv) > 4 else 6)
    else:
        for tag, pads in (("default", {}), ("d4b_fast_pads.inc", current_pads(N, Nq))):
            L = Layout(N, Nq, **pads)
            t, i = simulate(L, tag == "default")
            print(f"{tag}: N={N} Nq={Nq} PER={L.PER} wavefronts {t} ideal {i} ratio {t / i:.3f}")
