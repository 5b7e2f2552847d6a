"""CPU laboratory for the multilevel preconditioner of the random-walker solve (design experiments, not product code).

Builds the random-walker system of a synthetic scene with the oracle (graph + FAST features + class weights), then
runs float64 PCG over all label columns with candidate preconditioners and prints iteration counts.
  python scripts/amg_lab.py c5 0            # 640x480 scene, 8 labels
  python scripts/amg_lab.py crop            # 1024x768 centre crop of C4, 16 labels
"""
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def build_system(kind, idx=0, cache=True):
    path = "/tmp/amg_lab_%s_%d.npz" % (kind, idx)
    if cache and os.path.exists(path):
        d = np.load(path)
        return sp.csr_matrix((d["data"], d["indices"], d["indptr"]), shape=tuple(d["shape"])), d["B"]
    import orc
    from tcs_b200 import synth
    from tcs_b200.seeds import seeds_from_labels
    if kind == "c5":
        xyz, rgba, label = synth.make_c5(idx)
        res, per = 0.006, 5
    elif kind == "crop":
        w, h = 4480, 2880
        xyz, rgba, label = synth.make_scene(w, h, 15, 20261104 + idx)
        cw, ch = 1024, 768
        r0, c0 = (h - ch) // 2, (w - cw) // 2
        xyz = xyz.reshape(h, w, 3)[r0:r0 + ch, c0:c0 + cw].reshape(-1, 3).copy()
        rgba = rgba.reshape(h, w)[r0:r0 + ch, c0:c0 + cw].reshape(-1).copy()
        label = label.reshape(h, w)[r0:r0 + ch, c0:c0 + cw].reshape(-1).copy()
        res, per = 0.005, 10
    elif kind == "mid":
        w, h = 2240, 1440
        xyz, rgba, label = synth.make_scene(w, h, 15, 20261104 + idx)
        res, per = 0.005, 10
    else:
        raise SystemExit("unknown kind")
    sx, sl, _ = seeds_from_labels(xyz, label, idx, per)
    g = orc.Graph.from_cloud(xyz, rgba, res)
    g.normals(False, orc.MODE_FAST)
    g.signed_curvatures()
    g.edge_weights(orc.PRESET_CLASS, 1e-5, orc.SMALL_COERCE, orc.MODE_FAST)
    s, t, w = g.edges()
    V = g.V
    near = g.nearest_vertices(sx)
    labels = np.unique(sl)
    colour = np.zeros(V, np.int64)
    for i, v in enumerate(near):
        colour[v] = int(np.searchsorted(labels, sl[i])) + 1
    K = len(labels)
    W = sp.coo_matrix((w.astype(np.float64), (s.astype(np.int64), t.astype(np.int64))), shape=(V, V)).tocsr()
    W = W + W.T
    deg = np.asarray(W.sum(axis=1)).ravel()
    free = (colour == 0) & (deg >= np.finfo(np.float32).eps)
    # drop seedless components
    nc, comp = sp.csgraph.connected_components(W, directed=False)
    has_seed = np.zeros(nc, bool)
    has_seed[comp[colour > 0]] = True
    free &= has_seed[comp]
    rows = np.nonzero(free)[0]
    Lfull = sp.diags(deg) - W
    L = Lfull[rows][:, rows].tocsr()
    B = np.zeros((len(rows), K))
    Wf = W[rows].tocsr()
    for k in range(K):
        B[:, k] = Wf @ (colour == k + 1).astype(np.float64)
    L.sort_indices()
    if cache:
        np.savez(path, data=L.data, indices=L.indices, indptr=L.indptr, shape=np.array(L.shape), B=B)
    return L, B


def pcg(A, B, M, tol=1e-6, maxit=20000, flexible=False, verbose=False):
    n, K = B.shape
    X = np.zeros_like(B)
    R = B.copy()
    bn = np.linalg.norm(B, axis=0)
    bn[bn == 0] = 1
    Z = M(R)
    P = Z.copy()
    rz = np.einsum("ij,ij->j", R, Z)
    Rold = None
    for it in range(1, maxit + 1):
        AP = A @ P
        pAp = np.einsum("ij,ij->j", P, AP)
        al = np.where(pAp > 0, rz / np.where(pAp > 0, pAp, 1), 0)
        X += al * P
        if flexible:
            Rold = R.copy()
        R -= al * AP
        rel = np.linalg.norm(R, axis=0) / bn
        if verbose and it % 10 == 0:
            print(it, rel.max())
        if rel.max() <= tol:
            return X, it, rel
        Z = M(R)
        rz2 = np.einsum("ij,ij->j", R, Z)
        if flexible:
            num = rz2 - np.einsum("ij,ij->j", Rold, Z)
        else:
            num = rz2
        be = np.where(rz > 0, num / np.where(rz > 0, rz, 1), 0)
        P = Z + be * P
        rz = rz2
    return X, maxit, rel


# ---------------------------------------------------------------- aggregation
def strength_pairs(A, agg_w=None):
    """one pass of heavy-edge handshaking matching on the graph of A (off-diagonal negatives).  Returns agg id per row, count."""
    n = A.shape[0]
    Ao = A.tocoo()
    mask = Ao.row != Ao.col
    r, c, v = Ao.row[mask], Ao.col[mask], -Ao.data[mask]
    d = A.diagonal()
    # strength measure: w_ij / sqrt(d_i d_j)  (symmetric, scale-free)
    s = v / np.sqrt(d[r] * d[c])
    return r, c, s


def pairwise_match(A, rounds=4, kappa_thresh=0.0):
    n = A.shape[0]
    r, c, s = strength_pairs(A)
    matched = np.full(n, -1, np.int64)
    for rd in range(rounds):
        ok = (matched[r] < 0) & (matched[c] < 0) & (s > kappa_thresh)
        if not ok.any():
            break
        rr, cc, ss = r[ok], c[ok], s[ok]
        # strongest available neighbour per row (ties: lowest column)
        order = np.lexsort((cc, -ss, rr))
        rr, cc, ss = rr[order], cc[order], ss[order]
        first = np.ones(len(rr), bool)
        first[1:] = rr[1:] != rr[:-1]
        best = np.full(n, -1, np.int64)
        best[rr[first]] = cc[first]
        i = np.nonzero(best >= 0)[0]
        mutual = i[best[best[i]] == i]
        matched[mutual] = best[mutual]
    agg = np.full(n, -1, np.int64)
    # pair id = min of the two
    rep = np.where(matched >= 0, np.minimum(np.arange(n), matched), np.arange(n))
    uniq, agg = np.unique(rep, return_inverse=True)
    return agg, len(uniq)


def galerkin(A, agg, m):
    n = A.shape[0]
    Z = sp.csr_matrix((np.ones(n), (np.arange(n), agg)), shape=(n, m))
    Ac = (Z.T @ A @ Z).tocsr()
    Ac.sort_indices()
    return Ac, Z


class Hierarchy:
    def __init__(self, A, passes=2, min_rows=200, max_levels=20, verbose=True, thresh=0.0, rounds=4):
        self.A = [A.tocsr()]
        self.Z = []
        t = time.time()
        while self.A[-1].shape[0] > min_rows and len(self.A) < max_levels:
            Af = self.A[-1]
            Zt = None
            Ac = Af
            for p in range(passes):
                agg, m = pairwise_match(Ac, rounds=rounds, kappa_thresh=thresh)
                Ac, Zp = galerkin(Ac, agg, m)
                Zt = Zp if Zt is None else (Zt @ Zp).tocsr()
            if Ac.shape[0] > 0.9 * Af.shape[0]:
                break
            self.A.append(Ac)
            self.Z.append(Zt.tocsr())
        if verbose:
            print("  hierarchy: rows", [a.shape[0] for a in self.A], "nnz", [a.nnz for a in self.A],
                  "op complexity %.2f" % (sum(a.nnz for a in self.A) / self.A[0].nnz), "(%.1fs)" % (time.time() - t))
        self.D = [a.diagonal() for a in self.A]
        Al = self.A[-1].toarray()
        Al += 1e-10 * np.diag(np.diag(Al))
        self.coarse_inv = np.linalg.pinv(Al)
        self.bottom_sweeps = 0

    # ---- multiplicative V-cycle, nu Jacobi sweeps (weight w), correction scaled by `over`
    def vcycle(self, R, l=0, nu=1, w=0.7, over=1.0, gamma=1):
        A, D = self.A[l], self.D[l]
        if l == len(self.A) - 1:
            if self.bottom_sweeps <= 0:
                return self.coarse_inv @ R
            Dg = np.where(D > 1e-30, D, np.inf)
            X = (w / Dg)[:, None] * R
            for _ in range(self.bottom_sweeps - 1):
                X += (w / Dg)[:, None] * (R - A @ X)
            return X
        X = (w / D)[:, None] * R
        for _ in range(nu - 1):
            X += (w / D)[:, None] * (R - A @ X)
        Rc = self.Z[l].T @ (R - A @ X)
        g = gamma[l + 1] if isinstance(gamma, (list, tuple)) and l + 1 < len(gamma) else (1 if isinstance(gamma, (list, tuple)) else gamma)
        Ec = self.vcycle(Rc, l + 1, nu, w, over, gamma)
        for _ in range(g - 1):
            Ec += self.vcycle(Rc - self.A[l + 1] @ Ec, l + 1, nu, w, over, gamma)
        X += over * (self.Z[l] @ Ec)
        for _ in range(nu):
            X += (w / D)[:, None] * (R - A @ X)
        return X

    # ---- additive multilevel: z = sum_l Z_0..l D_l^-1 Z^T r  (BPX-like), exact on the last level
    def additive(self, R, scale=1.0):
        out = R / self.D[0][:, None]
        Rl = R
        stack = []
        for l in range(len(self.Z)):
            Rl = self.Z[l].T @ Rl
            stack.append(Rl)
        E = self.coarse_inv @ stack[-1]
        for l in range(len(self.Z) - 1, -1, -1):
            if l < len(self.Z) - 1:
                E = E + scale * stack[l] / self.D[l + 1][:, None]
            E = self.Z[l] @ E if False else E
            # prolongate one level down
            E = self.Z[l] @ E
            if l > 0:
                pass
        return out + E

    # ---- additive at the fine level (Jacobi + coarse correction), multiplicative V-cycle below
    def hybrid(self, R, nu=1, w=0.7, over=1.0, gamma=1, top=1.0):
        out = R / self.D[0][:, None]
        Rc = self.Z[0].T @ R
        g = gamma[1] if isinstance(gamma, (list, tuple)) else gamma
        Ec = self.vcycle(Rc, 1, nu, w, over, gamma)
        for _ in range(g - 1):
            Ec += self.vcycle(Rc - self.A[1] @ Ec, 1, nu, w, over, gamma)
        return out + top * (self.Z[0] @ Ec)


if __name__ == "__main__":
    kind = sys.argv[1] if len(sys.argv) > 1 else "c5"
    idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    t = time.time()
    L, B = build_system(kind, idx)
    n, K = B.shape
    print("system %s %d: n=%d nnz=%d K=%d (%.1fs)" % (kind, idx, n, L.nnz, K, time.time() - t))
    d = L.diagonal()
    which = sys.argv[3].split(",") if len(sys.argv) > 3 else ["jacobi", "v11", "hybrid"]
    if "jacobi" in which:
        t = time.time()
        X, it, rel = pcg(L, B, lambda R: R / d[:, None])
        print("jacobi: %d iterations (%.1fs)" % (it, time.time() - t))
    H = None
    for passes in (2, 3):
        if not any(w in which for w in ("v11", "v22", "hybrid", "additive", "k11")):
            break
        print("pairwise passes per level:", passes)
        H = Hierarchy(L, passes=passes)
        for name, M in [
            ("v11 w0.7", lambda R: H.vcycle(R, 0, 1, 0.7)),
            ("v11 w0.7 over1.5", lambda R: H.vcycle(R, 0, 1, 0.7, 1.5)),
            ("v22 w0.7", lambda R: H.vcycle(R, 0, 2, 0.7)),
            ("w11 w0.7", lambda R: H.vcycle(R, 0, 1, 0.7, 1.0, 2)),
            ("hybrid v11", lambda R: H.hybrid(R, 1, 0.7)),
            ("hybrid w11", lambda R: H.hybrid(R, 1, 0.7, 1.0, 2)),
        ]:
            key = name.split()[0]
            if key not in which and not (key == "w11" and "v11" in which):
                continue
            t = time.time()
            X, it, rel = pcg(L, B, M, maxit=3000)
            print("  %-22s %5d iterations (%.1fs)" % (name, it, time.time() - t))
