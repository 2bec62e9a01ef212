"""Rank-local flat tables from serial flat tables (host-side logic of the N>1 path).

d4est shards the mesh by p4est's space-filling-curve partition: rank r owns a contiguous range of
the global element order, sees the face neighbours owned by other ranks as *ghost* elements
(d4est_ghost_t, Mesh/d4est_ghost.h:7; ordered by owner rank, then by position on the curve, as in
p4est's ghost layer) and ships the nodal values of its *mirror* elements to the ranks that hold them
as ghosts (d4est_ghost_data_exchange, Mesh/d4est_ghost_data.c:143-255).  Under MPI the glue
(d4est_b200/glue/d4est_b200_glue.c) produces exactly these rank-local tables from p4est's ghost
layer; this module produces them from a SERIAL table set (an oracle fixture), so that the N-rank
path can be tested against the 1-rank path on any mesh, with or without GPUs.

Table names and conventions are those of d4b_mesh_desc_t (include/d4est_b200.h): ghosts follow the
locals in deg/node_off, hm.nbr = local id | n_local + ghost id | -1, and per peer p the mirror
elements send_elem[send_begin[p]:send_begin[p+1]] fill the peer's ghost range
recv_begin[p]:recv_begin[p+1] in the same (global) order.
"""
import numpy as np


def split_ranges(n_elem, nranks):
    """Equal-count contiguous ranges of the curve (p4est_partition with uniform weights)."""
    cuts = [(n_elem * r) // nranks for r in range(nranks + 1)]
    return np.asarray(cuts, dtype=np.int64)


def partition(fx, nranks, cuts=None, only=None):
    """Return a list of per-rank table dicts (same keys as a fixture, plus 'owned', 'ghost_gid'); with only = r the
    list holds None for the other ranks (each process of an N-rank run builds just its own tables)."""
    nl, ng, qt, _ = [int(v) for v in fx["hdr"]]
    assert ng == 0, "partition() expects serial tables"
    cuts = split_ranges(nl, nranks) if cuts is None else np.asarray(cuts, dtype=np.int64)
    deg, degq, node_off, quad_off = fx["deg"], fx["degq"], fx["node_off"], fx["quad_off"]
    hm, hb, fac, bcg, G = fx["hm"], fx["hm_begin"], fx["fac"], fx["bc_g"], fx["G"]
    owner = np.searchsorted(cuts, np.arange(nl), side="right") - 1
    parts = []
    for r in range(nranks):
        if only is not None and r != only:
            parts.append(None)
            continue
        e0, e1 = int(cuts[r]), int(cuts[r + 1])
        recs = hm[hb[e0]:hb[e1]].copy()
        nbr = recs["nbr"].astype(np.int64)
        off = (nbr >= 0) & ((nbr < e0) | (nbr >= e1))
        ghost_gid = np.unique(nbr[off])                      # sorted: owner-major, curve order inside
        gidx = {int(g): i for i, g in enumerate(ghost_gid)}
        n_loc, n_gh = e1 - e0, len(ghost_gid)
        new_nbr = nbr.copy()
        inside = (nbr >= e0) & (nbr < e1)
        new_nbr[inside] -= e0
        new_nbr[off] = [n_loc + gidx[int(g)] for g in nbr[off]]
        # factor and boundary-value pools rebuilt record by record (offsets re-based)
        fac_parts, g_parts, fo, go = [], [], 0, 0
        for i in range(len(recs)):
            n2 = (int(recs["degq"][i]) + 1) ** 2
            ln = (7 if nbr[i] >= 0 else 4) * n2
            o = int(recs["fac_off"][i])
            fac_parts.append(fac[o:o + ln]); recs["fac_off"][i] = fo; fo += ln
            if nbr[i] < 0 and recs["g_off"][i] >= 0:
                g0 = int(recs["g_off"][i])
                g_parts.append(bcg[g0:g0 + n2]); recs["g_off"][i] = go; go += n2
        recs["nbr"] = new_nbr.astype(np.int32)
        d_loc = deg[e0:e1]; d_gh = deg[ghost_gid] if n_gh else np.zeros(0, deg.dtype)
        d_all = np.concatenate([d_loc, d_gh]).astype(np.int32)
        noff = np.concatenate([[0], np.cumsum((d_all.astype(np.int64) + 1) ** 3)]).astype(np.int64)
        gh_owner = owner[ghost_gid] if n_gh else np.zeros(0, np.int64)
        # mirrors: my elements that have a record whose neighbour lives on another rank
        mirrors = {}
        el_of_rec = np.repeat(np.arange(e0, e1), np.diff(hb[e0:e1 + 1]))
        for e, g in zip(el_of_rec[off], nbr[off]):
            mirrors.setdefault(int(owner[g]), set()).add(int(e))
        peers = sorted(set(int(p) for p in gh_owner) | set(mirrors))
        send_begin, send_elem, recv_begin = [0], [], [0]
        for p in peers:
            send_elem += [e - e0 for e in sorted(mirrors.get(p, ()))]
            send_begin.append(len(send_elem))
            recv_begin.append(int(np.searchsorted(gh_owner, p, side="right")))
        t = {
            "hdr": np.array([n_loc, n_gh, qt, len(peers)], np.int32),
            "deg": d_all, "node_off": noff,
            "degq": degq[e0:e1].astype(np.int32), "quad_off": (quad_off[e0:e1 + 1] - quad_off[e0]).astype(np.int64),
            "G": G[6 * quad_off[e0]:6 * quad_off[e1]].copy(),
            "hm_begin": (hb[e0:e1 + 1] - hb[e0]).astype(np.int32), "hm": recs,
            "fac": np.concatenate(fac_parts) if fac_parts else np.zeros(0),
            "bc_g": np.concatenate(g_parts) if g_parts else np.zeros(0),
            "peer_rank": np.array(peers, np.int32), "send_begin": np.array(send_begin, np.int32),
            "send_elem": np.array(send_elem, np.int32), "recv_begin": np.array(recv_begin, np.int32),
            "owned": (e0, e1), "ghost_gid": ghost_gid.astype(np.int64),
            "node_range": (int(node_off[e0]), int(node_off[e1])),
        }
        parts.append(t)
    return parts


def local_vector(fx, part, v):
    """Slice of a global nodal vector owned by the rank."""
    a, b = part["node_range"]
    return np.ascontiguousarray(v[a:b])


def ghost_vector(fx, part, v):
    """Ghost values of a global nodal vector in the rank's ghost order (what the halo delivers)."""
    no = fx["node_off"]
    if len(part["ghost_gid"]) == 0:
        return np.zeros(0)
    return np.concatenate([v[no[g]:no[g + 1]] for g in part["ghost_gid"]])


def pack_mirrors(part, u_local, p):
    """Send buffer for peer index p: nodal values of the mirror elements, concatenated
    (Mesh/d4est_ghost_data.c:9-29 lays the receive side out the same way)."""
    no, sb, se = part["node_off"], part["send_begin"], part["send_elem"]
    el = se[sb[p]:sb[p + 1]]
    return np.concatenate([u_local[no[e]:no[e + 1]] for e in el]) if len(el) else np.zeros(0)


def recv_slice(part, p):
    """Range of the ghost vector filled by peer index p."""
    nl = int(part["hdr"][0]); no, rb = part["node_off"], part["recv_begin"]
    base = int(no[nl])
    return int(no[nl + rb[p]]) - base, int(no[nl + rb[p + 1]]) - base


def exchange_halo_torch(part, u_local, dist):
    """Forward halo over a torch.distributed process group (gloo on CPU): the host-side mirror of
    d4b_halo_exchange (csrc/d4b_halo.cu), used by the world_size-2 CPU tests."""
    import torch
    nl = int(part["hdr"][0])
    ug = torch.zeros(int(part["node_off"][-1] - part["node_off"][nl]), dtype=torch.float64)
    ops, keep = [], []
    for p, peer in enumerate(part["peer_rank"]):
        sbuf = torch.from_numpy(np.ascontiguousarray(pack_mirrors(part, u_local, p)))
        a, b = recv_slice(part, p)
        keep.append(sbuf)
        if sbuf.numel():
            ops.append(dist.P2POp(dist.isend, sbuf, int(peer)))
        if b > a:
            ops.append(dist.P2POp(dist.irecv, ug[a:b], int(peer)))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return ug.numpy()


# ---------------------------------------------------------------------------------------------------------------
# multigrid hierarchy: d4est keeps families together when it partitions (p4est_partition(p4est, 1, NULL),
# driver.c:135-152) so that restriction / prolongation stay rank-local, and drops the coarsest levels as the rank
# count grows (Solver/d4est_solver_multigrid.c:253-277).  The same here: cuts are chosen on a base level and every
# finer level is cut at the first descendant of the base cut elements.
def hierarchy_cuts(levels, tabs, nranks, base=0):
    """levels[l]: serial tables of level l (0 = coarsest), tabs[l]: transfer table between l and l+1.
    Returns cuts[l] (element indices, length nranks+1) for l >= base; the base level is cut so that the FINEST-level
    nodes below each range are balanced."""
    top = len(levels) - 1
    # first fine element of every coarse element, level by level
    first_child = []
    for l in range(top):
        no_f = levels[l + 1]["node_off"][:int(levels[l + 1]["hdr"][0]) + 1]
        fc = np.searchsorted(no_f, tabs[l]["child_off"][0::8]).astype(np.int64)
        first_child.append(np.concatenate([fc, [int(levels[l + 1]["hdr"][0])]]))
    nb = int(levels[base]["hdr"][0])
    # finest-level nodes below each base element
    lo = np.arange(nb + 1, dtype=np.int64)
    for l in range(base, top):
        lo = first_child[l][lo]
    no_top = levels[top]["node_off"]
    weight = np.diff(no_top[lo])
    cum = np.concatenate([[0], np.cumsum(weight)])
    cuts_b = [0]
    for r in range(1, nranks):
        c = int(np.searchsorted(cum, cum[-1] * r / nranks))
        cuts_b.append(min(max(c, cuts_b[-1] + 1), nb - (nranks - r)))
    cuts_b.append(nb)
    cuts = {base: np.asarray(cuts_b, dtype=np.int64)}
    for l in range(base, top):
        cuts[l + 1] = first_child[l][cuts[l]]
    return cuts


def partition_transfer(tab, coarse, fine, r):
    """Rows of a transfer table that belong to rank r, offsets re-based to the rank's node ranges.
    coarse / fine: the rank-local tables of the two levels (from partition())."""
    e0, e1 = coarse["owned"]
    c0, f0 = coarse["node_range"][0], fine["node_range"][0]
    ch = tab["child_off"][8 * e0:8 * e1].astype(np.int64).copy()
    kind = tab["kind"][e0:e1]
    used = np.zeros(ch.size, bool)
    used[0::8] = True
    for c in range(1, 8):
        used[c::8] = kind == 1
    ch[used] -= f0
    ch[~used] = 0
    return {"kind": kind.astype(np.int32).copy(), "degH": tab["degH"][e0:e1].astype(np.int32).copy(),
            "coarse_off": (tab["coarse_off"][e0:e1] - c0).astype(np.int64),
            "child_deg": tab["child_deg"][8 * e0:8 * e1].astype(np.int32).copy(), "child_off": ch}


def partition_hierarchy(levels, tabs, nranks, base=0, only=None):
    """Per rank: (list of rank-local level tables for levels base..top, list of rank-local transfer tables);
    with only = r the other entries are None."""
    cuts = hierarchy_cuts(levels, tabs, nranks, base)
    top = len(levels) - 1
    parts = {l: partition(levels[l], nranks, cuts[l], only) for l in range(base, top + 1)}
    out = []
    for r in range(nranks):
        if only is not None and r != only:
            out.append(None)
            continue
        lv = [parts[l][r] for l in range(base, top + 1)]
        tr = [partition_transfer(tabs[l], parts[l][r], parts[l + 1][r], r) for l in range(base, top)]
        out.append((lv, tr))
    return out
