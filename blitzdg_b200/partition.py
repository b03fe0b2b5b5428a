"""Element partitioning for multi-GPU runs."""
import numpy as np


def rcb_partition(cx, cy, nparts):
    """Recursive coordinate bisection of element centroids into `nparts` balanced, contiguous parts."""
    K = len(cx)
    part = np.zeros(K, dtype=np.int32)

    def rec(idx, p0, n):
        if n == 1:
            part[idx] = p0
            return
        nl = n // 2
        xs, ys = cx[idx], cy[idx]
        coord = xs if (xs.max() - xs.min()) >= (ys.max() - ys.min()) else ys
        order = np.argsort(coord, kind="stable")
        cut = (len(idx) * nl) // n
        rec(idx[order[:cut]], p0, nl)
        rec(idx[order[cut:]], p0 + nl, n - nl)

    rec(np.arange(K), 0, int(nparts))
    return part


def strip_partition(cy, nparts):
    """Balanced horizontal strips by centroid y (stable): the natural split of the structured meshes."""
    order = np.argsort(cy, kind="stable")
    part = np.empty(len(cy), dtype=np.int32)
    K = len(cy)
    for r in range(nparts):
        part[order[(K * r) // nparts:(K * (r + 1)) // nparts]] = r
    return part


def make_partition(mesh, nparts, how="metis", seed=20260101):
    """ElementPartitionMap of `mesh` for `nparts` ranks.  how: "metis" = MeshManager.partitionMesh (METIS_PartMeshNodal
    with the reference's options, src/MeshManager.cpp:491-544; fixed seed so that every rank of a job computes the same
    map), "rcb" = recursive coordinate bisection, "strips" = balanced horizontal strips.  Returns (epart, description)."""
    K, Nf = mesh.get_NumElements(), mesh.get_NumFaces()
    if nparts <= 1:
        return np.zeros(K, dtype=np.int32), "single part"
    if how == "metis":
        ep = mesh.partitionMesh(nparts, method="auto", seed=seed)
        return np.asarray(ep, dtype=np.int32), "%s (k-way, reference options src/MeshManager.cpp:498-514, seed %d)" % (mesh.PartitionInfo["method"], seed)
    cen = mesh.get_Vertices().reshape(-1, 3)[mesh.get_Elements().reshape(K, Nf)].mean(axis=1)
    if how == "rcb":
        return rcb_partition(cen[:, 0], cen[:, 1], nparts), "recursive coordinate bisection"
    if how == "strips":
        return strip_partition(cen[:, 1], nparts), "%d horizontal strips" % nparts
    raise ValueError(how)


def partition_stats(mesh, epart):
    """Per-rank element counts, neighbour counts and cut faces of an element partition (for reports)."""
    K, Nf = mesh.get_NumElements(), mesh.get_NumFaces()
    epart = np.asarray(epart)
    P = int(epart.max()) + 1
    e2e = mesh.get_EToE().reshape(K, Nf)
    nb = epart[e2e]
    off = nb != epart[:, None]
    pair = np.unique(np.stack([np.repeat(epart, Nf)[off.reshape(-1)], nb[off]], axis=1), axis=0)
    return {"elements": np.bincount(epart, minlength=P).tolist(),
            "neighbours": np.bincount(pair[:, 0], minlength=P).tolist() if len(pair) else [0] * P,
            "cut_faces": np.bincount(np.repeat(epart, Nf)[off.reshape(-1)], minlength=P).tolist(),
            "cut_faces_total": int(off.sum()) // 2}


def build_local(mesh, epart, rank, N):
    """Provision rank `rank` of an element-partitioned mesh.

    Local elements are renumbered interior-first, then elements with an off-rank neighbour; off-rank
    '+' traces address ghost slots Np*K_loc + g, grouped by owner rank and ordered by (global element id,
    local node) -- the same key the owner sorts its send list by, so no negotiation is needed.

    Returns (ctx_local: DGContext2D, halo: dict for capi.pack_halo, info: dict with 'elements' =
    global ids of the local elements in local order).
    """
    from .mesh_manager import MeshManager
    from .synthetic import provision

    epart = np.asarray(epart)
    K, Nf = mesh.get_NumElements(), mesh.get_NumFaces()
    e2e = mesh.get_EToE().reshape(K, Nf)
    etov = mesh.get_Elements().reshape(K, Nf)
    vert = mesh.get_Vertices().reshape(-1, 3)
    bct = mesh.get_BCType().reshape(K, Nf)
    mine = np.nonzero(epart == rank)[0]
    nb_part = epart[e2e[mine]]  # [Kloc][Nf]
    is_cut = (nb_part != rank).any(axis=1)
    n_touch = int(is_cut.sum())
    # Elements wedged between cut elements (>= 2 of their neighbours are cut, e.g. the upper triangles of a
    # structured boundary row) join the cut group: left among the interior elements they would form tiles whose
    # members are not neighbours of each other.  "Interior touches no ghost slot" still holds.
    cut_glob = np.zeros(K, dtype=bool)
    cut_glob[mine[is_cut]] = True
    nb_cut = (cut_glob[e2e[mine]] & (e2e[mine] != mine[:, None])).sum(axis=1)
    is_cut = is_cut | (nb_cut >= 2)
    local = np.concatenate([mine[~is_cut], mine[is_cut]])
    n_interior = int((~is_cut).sum())
    Kloc = len(local)
    ring = np.unique(e2e[mine][nb_part != rank])
    sub = np.concatenate([local, ring])
    uv, inv = np.unique(etov[sub].reshape(-1), return_inverse=True)
    m = MeshManager()
    m.buildMesh(inv.reshape(len(sub), Nf), vert[uv])
    bc_sub = m.get_BCType().reshape(len(sub), Nf).copy()
    bc_sub[:Kloc] = bct[local]  # true boundary tags of the local elements; ring/local faces stay 0
    m.set_BCType(bc_sub)
    prov = provision(m, N)
    ctx = prov.get_DGContext()
    Np, Nfp = ctx.Np, ctx.Nfp
    Nfn = Nf * Nfp
    # ---- cut to the local elements
    cut = lambda a: np.ascontiguousarray(a[:, :Kloc])
    for name in ("rx", "ry", "sx", "sy", "nx", "ny", "Fscale", "J", "x", "y"):
        setattr(ctx, name, cut(getattr(ctx, name)))
    vM = ctx.vmapM[:Kloc * Nfn].copy()
    vP = ctx.vmapP[:Kloc * Nfn].astype(np.int64)
    ctx.K = Kloc
    off = vP >= Np * Kloc
    ring_elem_sub = vP[off] // Np          # index into sub
    ring_node = vP[off] % Np
    gelem = sub[ring_elem_sub]             # global element id
    owner = epart[gelem]
    key = (owner.astype(np.int64) * K + gelem) * Np + ring_node
    ukey, ginv = np.unique(key, return_inverse=True)
    vP[off] = Np * Kloc + ginv
    ctx.vmapM, ctx.vmapP = vM, vP.astype(np.int32)
    ctx.BCmap = {t: ids[ids < Kloc * Nfn] for t, ids in ctx.BCmap.items()}
    ctx.BCmap = {t: ids for t, ids in ctx.BCmap.items() if len(ids)}
    g_owner = (ukey // Np) // K
    nbrs = np.unique(g_owner).astype(np.int32)
    recv_offset = np.concatenate([[0], np.cumsum([(g_owner == r).sum() for r in nbrs])]).astype(np.int32)
    # ---- send lists: my face nodes on faces whose neighbour lives on rank r, sorted by (global elem, node)
    loc_of = {}
    glob_local = local
    fm = prov.Fmask.astype(np.int64)  # [Nfp][Nf]
    nbp = epart[e2e[local]]  # [Kloc][Nf]
    send_nodes, send_offset, nbr_list = [], [0], []
    for r in np.unique(nbp[nbp != rank]):
        kk, ff = np.nonzero(nbp == r)
        nodes = fm[:, ff].T  # [nfaces][Nfp]
        gk = np.repeat(glob_local[kk], Nfp)
        lk = np.repeat(kk, Nfp)
        nn = nodes.reshape(-1)
        skey = gk.astype(np.int64) * Np + nn
        _, first = np.unique(skey, return_index=True)
        send_nodes.append((lk[first] * Np + nn[first]).astype(np.int32))
        send_offset.append(send_offset[-1] + len(first))
        nbr_list.append(int(r))
    assert nbr_list == nbrs.tolist(), (nbr_list, nbrs.tolist())
    halo = dict(rank=int(rank), nranks=int(epart.max()) + 1, n_ghost=int(len(ukey)), nbr_rank=nbrs,
                send_offset=np.array(send_offset, dtype=np.int32),
                send_nodes=np.concatenate(send_nodes) if send_nodes else np.zeros(0, dtype=np.int32),
                recv_offset=recv_offset, n_interior=n_interior)
    info = dict(elements=local, ghost_keys=ukey, Np=Np, K_global=K)
    return ctx, halo, info
