"""Synthetic input producer: the reference's *input format* without FreeFem++.

The reference obtains its 26 component matrices, the right-hand-side vector and the
read-out (interpolation) matrices from a FreeFem++ subprocess
(`source/jax_plate/pyFFInterface.py:169-509`, `load_matrices_unsymm`).  FreeFem++ is not
available here, so this module assembles the same forms natively on a structured
triangulation of a rectangular cantilever plate:

* `Lh`  = P1 on the vertices (in-plane displacements u, v)              (`pyFFInterface.py:178-179`)
* `Mh`  = Morley element (vertex values of w + edge-midpoint normal derivatives)
                                                                        (`pyFFInterface.py:181-182`)
* forms `Sxx..Syy`, `SxxL..SyyL`, `M11`, `M11Correction` (`:217-230`), `Rxxx..Ryxy`
  (`:232-241`), `Txxxx..Tyyyy`, `M33`, `M33Correction`, `M33I2`, `M33I2Correction`
  (`:243-259`), Dirichlet rows with `tgv=-1` (row := unit row, columns untouched, `:176`),
* block placement / symmetrisation / row removal of `:337-498`,
* return tuple in the order of `:503-509`.

The output is what `Problem.__init__` (`source/jax_plate/Problem.py:310-374`) consumes:
26 scipy COO matrices (N x N, N = 2*Lh + Mh), `rhs_vec`, dense interpolation matrices and
the block sizes.  Nothing here is on the timed hot path.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import scipy.sparse as sps

# names in the order of pyFFInterface.py:503-507
COMPONENT_NAMES = (
    "KA11", "KA12", "KA16", "KA22", "KA26", "KA66",
    "KB11", "KB12", "KB16", "KB22", "KB26", "KB66",
    "KD11", "KD12", "KD16", "KD22", "KD26", "KD66",
    "KM11", "KM11Corr", "KM22", "KM22Corr", "KM33",
    "KM33Corr", "KM33I2", "KM33I2Corr",
)

# Dunavant degree-4 rule (6 points) on the reference triangle (barycentric coordinates)
_A1, _B1, _W1 = 0.445948490915965, 0.108103018168070, 0.223381589678011
_A2, _B2, _W2 = 0.091576213509771, 0.816847572980459, 0.109951743655322
_QUAD_BARY = np.array([
    [_B1, _A1, _A1], [_A1, _B1, _A1], [_A1, _A1, _B1],
    [_B2, _A2, _A2], [_A2, _B2, _A2], [_A2, _A2, _B2],
])
_QUAD_W = np.array([_W1, _W1, _W1, _W2, _W2, _W2])


@dataclass
class PlateMesh:
    points: np.ndarray      # [nv, 2]
    tris: np.ndarray        # [nt, 3] vertex ids (counter-clockwise)
    edges: np.ndarray       # [ne, 2] vertex ids, edges[:,0] < edges[:,1]
    tri_edges: np.ndarray   # [nt, 3] edge ids; local edge e is opposite local vertex e
    clamped_vertices: np.ndarray  # bool [nv]
    clamped_edges: np.ndarray     # bool [ne]
    Lx: float
    Ly: float


def rectangle_mesh(nx: int, ny: int, Lx: float, Ly: float) -> PlateMesh:
    """Structured triangulation; the short edge x = Lx carries label 1 (clamped),
    as border C3 in `source/jax_plate/geometry/sh_i.edp:25`."""
    xs = np.linspace(0.0, Lx, nx + 1)
    ys = np.linspace(-Ly / 2, Ly / 2, ny + 1)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    points = np.stack([X.ravel(), Y.ravel()], axis=1)

    def vid(i, j):
        return i * (ny + 1) + j

    I, J = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    I, J = I.ravel(), J.ravel()
    v00, v10, v11, v01 = vid(I, J), vid(I + 1, J), vid(I + 1, J + 1), vid(I, J + 1)
    tris = np.concatenate([np.stack([v00, v10, v11], 1), np.stack([v00, v11, v01], 1)], 0)
    tris = tris.astype(np.int64)

    # local edge e is opposite local vertex e
    loc = np.stack([tris[:, [1, 2]], tris[:, [2, 0]], tris[:, [0, 1]]], axis=1)  # [nt,3,2]
    loc_sorted = np.sort(loc, axis=2)
    nv = points.shape[0]
    key = loc_sorted[..., 0] * nv + loc_sorted[..., 1]
    ukey, inv = np.unique(key.ravel(), return_inverse=True)
    edges = np.stack([ukey // nv, ukey % nv], axis=1)
    tri_edges = inv.reshape(-1, 3)

    tol = 1e-12 * max(Lx, Ly)
    clamped_v = np.abs(points[:, 0] - Lx) < tol
    clamped_e = clamped_v[edges[:, 0]] & clamped_v[edges[:, 1]]
    return PlateMesh(points, tris, edges, tri_edges, clamped_v, clamped_e, Lx, Ly)


def _p1_gradients(mesh: PlateMesh):
    p = mesh.points[mesh.tris]                     # [nt,3,2]
    x, y = p[..., 0], p[..., 1]
    det = (x[:, 1] - x[:, 0]) * (y[:, 2] - y[:, 0]) - (x[:, 2] - x[:, 0]) * (y[:, 1] - y[:, 0])
    area = 0.5 * det
    gx = np.stack([y[:, 1] - y[:, 2], y[:, 2] - y[:, 0], y[:, 0] - y[:, 1]], 1) / det[:, None]
    gy = np.stack([x[:, 2] - x[:, 1], x[:, 0] - x[:, 2], x[:, 1] - x[:, 0]], 1) / det[:, None]
    return area, gx, gy


def _morley_coefficients(mesh: PlateMesh):
    """Per element: coefficients C[t, m, j] of basis j on monomials m = (1,x,y,x2,xy,y2) of
    the scaled local coordinate ((X - Xc)/hs), together with Xc and hs.  Basis j<3: vertex
    value at local vertex j; j>=3: normal derivative at the midpoint of local edge j-3 with
    the *global* edge normal (rotate edge tangent low-id -> high-id by -90 degrees)."""
    pts = mesh.points[mesh.tris]                   # [nt,3,2]
    xc = pts.mean(axis=1)                          # [nt,2]
    area, _, _ = _p1_gradients(mesh)
    hs = np.sqrt(2.0 * np.abs(area))               # element length scale
    q = (pts - xc[:, None, :]) / hs[:, None, None]  # scaled vertex coords

    def mono(x, y):
        return np.stack([np.ones_like(x), x, y, x * x, x * y, y * y], axis=-1)

    def dmono(x, y):
        z, o = np.zeros_like(x), np.ones_like(x)
        dx = np.stack([z, o, z, 2 * x, y, z], axis=-1)
        dy = np.stack([z, z, o, z, x, 2 * y], axis=-1)
        return dx, dy

    nt = pts.shape[0]
    V = np.zeros((nt, 6, 6))
    V[:, 0:3, :] = mono(q[..., 0], q[..., 1])      # vertex-value functionals
    e = mesh.edges[mesh.tri_edges]                 # [nt,3,2] global (low, high) vertex ids
    pa, pb = mesh.points[e[..., 0]], mesh.points[e[..., 1]]
    tang = pb - pa
    tang /= np.linalg.norm(tang, axis=-1, keepdims=True)
    normal = np.stack([tang[..., 1], -tang[..., 0]], axis=-1)      # [nt,3,2]
    mid = (0.5 * (pa + pb) - xc[:, None, :]) / hs[:, None, None]
    dx, dy = dmono(mid[..., 0], mid[..., 1])       # derivative wrt scaled coords
    # d/dX = (1/hs) d/dx  -> functional = n . grad_X
    V[:, 3:6, :] = (normal[..., 0:1] * dx + normal[..., 1:2] * dy) / hs[:, None, None]
    C = np.linalg.inv(V)                           # columns = basis coefficient vectors
    return C, xc, hs


def _scatter(nrow, ncol, rows, cols, vals):
    m = sps.coo_matrix((vals.ravel(), (rows.ravel(), cols.ravel())), shape=(nrow, ncol))
    m = m.tocsr()
    m.sum_duplicates()
    return m


def _dirichlet_rows(mat: sps.csr_matrix, idx: np.ndarray) -> sps.csr_matrix:
    """`on(1, ...)` with tgv = -1: constrained rows become unit rows; entries stay in the
    pattern as explicit zeros."""
    mat = mat.tocsr(copy=True)
    flag = np.zeros(mat.shape[0], dtype=bool)
    flag[idx] = True
    row_of = np.repeat(np.arange(mat.shape[0]), np.diff(mat.indptr))
    kill = flag[row_of]
    mat.data[kill] = 0.0
    diag = kill & (mat.indices == row_of)
    mat.data[diag] = 1.0
    return mat


@dataclass
class SyntheticPlate:
    """What `load_matrices_unsymm` returns (`pyFFInterface.py:503-509`), plus the mesh."""
    mats: list                 # 26 scipy COO (N, N) float64
    rhs_vec: np.ndarray        # [N]
    interp_mat: np.ndarray     # [npts, Mh]   w at the read-out points
    interp_mat_Lh: np.ndarray  # [npts, Lh]
    Lh_size: int
    Mh_size: int
    mesh: PlateMesh
    interp_mat_Wx: np.ndarray  # [npts, Mh]
    interp_mat_Wy: np.ndarray  # [npts, Mh]
    test_point: tuple

    @property
    def N(self) -> int:
        return 2 * self.Lh_size + self.Mh_size

    def as_tuple(self):
        return (self.mats, self.rhs_vec, self.interp_mat, self.interp_mat_Lh, self.Lh_size,
                self.Mh_size, self.mesh, self.interp_mat_Wx, self.interp_mat_Wy)


def _locate(mesh: PlateMesh, pts: np.ndarray, nx: int, ny: int):
    """Triangle id and barycentric coordinates of points in the structured mesh."""
    Lx, Ly = mesh.Lx, mesh.Ly
    fx = np.clip(pts[:, 0] / Lx * nx, 0, nx - 1e-9)
    fy = np.clip((pts[:, 1] + Ly / 2) / Ly * ny, 0, ny - 1e-9)
    i, j = fx.astype(int), fy.astype(int)
    lx, ly = fx - i, fy - j
    cell = i * ny + j
    lower = ly <= lx                      # triangle (v00, v10, v11)
    tri = np.where(lower, cell, cell + nx * ny)
    p = mesh.points[mesh.tris[tri]]
    T = np.stack([p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]], axis=2)   # [n,2,2]
    rhs = pts - p[:, 0]
    lam12 = np.linalg.solve(T, rhs[..., None])[..., 0]
    lam = np.concatenate([1 - lam12.sum(1, keepdims=True), lam12], axis=1)
    return tri, lam


def make_plate(nx: int, ny: int, Lx: float = 99.9e-3, Ly: float = 20e-3,
               r_accel: float = 3.8e-3, n_readout: int = 16) -> SyntheticPlate:
    mesh = rectangle_mesh(nx, ny, Lx, Ly)
    nv, ne, nt = mesh.points.shape[0], mesh.edges.shape[0], mesh.tris.shape[0]
    Lh, Mh = nv, nv + ne
    N = 2 * Lh + Mh

    area, gx, gy = _p1_gradients(mesh)
    xtest, ytest = r_accel, Ly / 2 - r_accel          # sh_i.edp:11-15
    cent = mesh.points[mesh.tris].mean(axis=1)
    ind = ((cent[:, 0] - xtest) ** 2 + (cent[:, 1] - ytest) ** 2 <= r_accel ** 2 + 1e-8)
    ind = ind.astype(np.float64)                      # indAccel, sh_i.edp:34

    # ---------------- P1 x P1 forms (rows = test r, cols = trial u) ----------------
    ldof = mesh.tris                                  # [nt,3]
    Rl = np.repeat(ldof[:, :, None], 3, axis=2)
    Cl = np.repeat(ldof[:, None, :], 3, axis=1)
    a = area[:, None, None]

    def p1(gr, gu):                                   # area * d(r)_i * d(u)_j
        return _scatter(Lh, Lh, Rl, Cl, a * gr[:, :, None] * gu[:, None, :])

    SxxL, SxyL = p1(gx, gx), p1(gx, gy)               # Sxy = dx(r)*dy(u)
    SyxL, SyyL = p1(gy, gx), p1(gy, gy)               # Syx = dx(u)*dy(r)
    mloc = (np.ones((3, 3)) + np.eye(3)) / 12.0
    M11 = _scatter(Lh, Lh, Rl, Cl, a * mloc[None])
    M11C = _scatter(Lh, Lh, Rl, Cl, (a * ind[:, None, None]) * mloc[None])

    dL = np.nonzero(mesh.clamped_vertices)[0]
    Sxx, Sxy = _dirichlet_rows(SxxL, dL), _dirichlet_rows(SxyL, dL)
    Syx, Syy = _dirichlet_rows(SyxL, dL), _dirichlet_rows(SyyL, dL)
    M11, M11C = _dirichlet_rows(M11, dL), _dirichlet_rows(M11C, dL)

    # ---------------- Morley forms ----------------
    C, xc, hs = _morley_coefficients(mesh)            # [nt,6(mono),6(basis)]
    mdof = np.concatenate([mesh.tris, nv + mesh.tri_edges], axis=1)   # [nt,6]
    wxx = 2.0 * C[:, 3, :] / hs[:, None] ** 2
    wxy = C[:, 4, :] / hs[:, None] ** 2
    wyy = 2.0 * C[:, 5, :] / hs[:, None] ** 2

    Rm = np.repeat(mdof[:, :, None], 6, axis=2)
    Cm = np.repeat(mdof[:, None, :], 6, axis=1)

    def tform(dt, dw):                                # area * d2(t)_i * d2(w)_j
        return _scatter(Mh, Mh, Rm, Cm, a * dt[:, :, None] * dw[:, None, :])

    Txxxx, Txxyy, Tyyxx = tform(wxx, wxx), tform(wyy, wxx), tform(wxx, wyy)
    Txxxy, Txyxx = tform(wxy, wxx), tform(wxx, wxy)
    Txyyy, Tyyxy = tform(wyy, wxy), tform(wxy, wyy)
    Txyxy, Tyyyy = tform(wxy, wxy), tform(wyy, wyy)

    # basis values / gradients at quadrature points
    pts = mesh.points[mesh.tris]
    qp = np.einsum("qk,tkd->tqd", _QUAD_BARY, pts)    # [nt,6q,2]
    qs = (qp - xc[:, None, :]) / hs[:, None, None]
    x, y = qs[..., 0], qs[..., 1]
    one, zero = np.ones_like(x), np.zeros_like(x)
    mono = np.stack([one, x, y, x * x, x * y, y * y], axis=-1)             # [nt,q,6]
    mdx = np.stack([zero, one, zero, 2 * x, y, zero], axis=-1) / hs[:, None, None]
    mdy = np.stack([zero, zero, one, zero, x, 2 * y], axis=-1) / hs[:, None, None]
    val = mono @ C                                     # [nt,q,6 basis]
    vdx, vdy = mdx @ C, mdy @ C
    wq = _QUAD_W[None, :] * area[:, None]              # [nt,q]
    m33 = np.einsum("tq,tqi,tqj->tij", wq, val, val)
    m33i2 = np.einsum("tq,tqi,tqj->tij", wq, vdx, vdx) + np.einsum("tq,tqi,tqj->tij", wq, vdy, vdy)
    M33 = _scatter(Mh, Mh, Rm, Cm, m33)
    M33C = _scatter(Mh, Mh, Rm, Cm, m33 * ind[:, None, None])
    M33I2 = _scatter(Mh, Mh, Rm, Cm, m33i2)
    M33I2C = _scatter(Mh, Mh, Rm, Cm, m33i2 * ind[:, None, None])

    dM = np.concatenate([np.nonzero(mesh.clamped_vertices)[0],
                         nv + np.nonzero(mesh.clamped_edges)[0]])
    (Txxxx, Txxyy, Tyyxx, Txxxy, Txyxx, Txyyy, Tyyxy, Txyxy, Tyyyy, M33, M33C, M33I2,
     M33I2C) = [_dirichlet_rows(m, dM) for m in
                (Txxxx, Txxyy, Tyyxx, Txxxy, Txyxx, Txyyy, Tyyxy, Txyxy, Tyyyy, M33, M33C,
                 M33I2, M33I2C)]

    # ---------------- P1 (test r, rows) x Morley (trial w, cols) ----------------
    Rr = np.repeat(ldof[:, :, None], 6, axis=2)
    Cr = np.repeat(mdof[:, None, :], 3, axis=1)

    def rform(gr, dw):                                # area * d(r)_i * d2(w)_j
        return _scatter(Lh, Mh, Rr, Cr, a * gr[:, :, None] * dw[:, None, :])

    Rxxx, Rxyy, Rxxy = rform(gx, wxx), rform(gx, wyy), rform(gx, wxy)
    Ryxx, Ryyy, Ryxy = rform(gy, wxx), rform(gy, wyy), rform(gy, wxy)

    # ---------------- block placement (pyFFInterface.py:337-498) ----------------
    def place(mat, nrow_blk, ncol_blk):
        m = mat.tocoo()
        return sps.coo_matrix((m.data, (m.row + nrow_blk * Lh, m.col + ncol_blk * Lh)),
                              shape=(N, N))

    def transp(m):                                    # (mat + mat^T).tocoo()
        return (m + m.transpose(copy=True)).tocoo()

    def rmrows(m, rows):
        m = m.tocoo()
        flag = np.zeros(N, dtype=bool)
        flag[rows] = True
        m.data[flag[m.row]] = 0.0
        return m

    rows_u, rows_v, rows_w = dL, dL + Lh, dM + 2 * Lh

    def coupling_uv(mat):                             # move(.,1,0) -> transp -> rmrows 1,2
        return rmrows(rmrows(transp(place(mat, 0, 1)), rows_u), rows_v)

    def coupling_w(mat, blk):                         # move(.,2,blk) -> transp -> rmrows
        return rmrows(rmrows(transp(place(mat, blk, 2)), rows_u if blk == 0 else rows_v),
                      rows_w)

    def add(*ms):
        # scipy's '+' prunes nothing but may drop explicit zeros; rebuild on the union of
        # the stored patterns so explicit zeros survive like in the reference's COO data.
        rows = np.concatenate([m.tocoo().row for m in ms])
        cols = np.concatenate([m.tocoo().col for m in ms])
        data = np.concatenate([m.tocoo().data for m in ms])
        key = rows.astype(np.int64) + N * cols.astype(np.int64)
        uk, inv = np.unique(key, return_inverse=True)
        acc = np.zeros(uk.size)
        np.add.at(acc, inv, data)
        return sps.coo_matrix((acc, (uk % N, uk // N)), shape=(N, N))

    KA11 = place(Sxx, 0, 0)
    KA22 = place(Syy, 1, 1)
    KA12 = coupling_uv(SxyL)
    KA16 = add(place(Sxy + Syx, 0, 0), coupling_uv(SxxL))
    KA26 = add(coupling_uv(SyyL), place(Sxy + Syx, 1, 1))
    KA66 = add(place(Syy, 0, 0), coupling_uv(SyxL), place(Sxx, 1, 1))

    KB11 = coupling_w(-Rxxx, 0)
    KB22 = coupling_w(-Ryyy, 1)
    KB12 = add(coupling_w(-Rxyy, 0), coupling_w(-Ryxx, 1))
    KB16 = add(coupling_w(-Ryxx, 0), coupling_w(-Rxxx, 1))
    KB26 = add(coupling_w(-2 * Rxxy - Ryyy, 0), coupling_w(-Rxyy - 2 * Ryxy, 1))
    KB66 = add(coupling_w(-2 * Ryxy, 0), coupling_w(-2 * Rxxy, 1))

    KD11 = place(Txxxx, 2, 2)
    KD12 = place(Txxyy + Tyyxx, 2, 2)
    KD16 = place(2 * (Txxxy + Txyxx), 2, 2)
    KD22 = place(Tyyyy, 2, 2)
    KD26 = place(2 * (Txyyy + Tyyxy), 2, 2)
    KD66 = place(4 * Txyxy, 2, 2)

    KM11, KM11Corr = place(M11, 0, 0), place(M11C, 0, 0)
    KM22, KM22Corr = place(M11, 1, 1), place(M11C, 1, 1)
    KM33, KM33Corr = place(M33, 2, 2), place(M33C, 2, 2)
    KM33I2, KM33I2Corr = place(M33I2, 2, 2), place(M33I2C, 2, 2)

    mats = [KA11, KA12, KA16, KA22, KA26, KA66, KB11, KB12, KB16, KB22, KB26, KB66,
            KD11, KD12, KD16, KD22, KD26, KD66, KM11, KM11Corr, KM22, KM22Corr, KM33,
            KM33Corr, KM33I2, KM33I2Corr]
    mats = [m.tocoo() for m in mats]

    # rhs: vBCMh = BCM(0, Mh, tgv=-1) -> funcBC = 1 on clamped vertex dofs, 0 on the
    # clamped normal-derivative dofs (pyFFInterface.py:185-197, 500-501)
    rhs_vec = np.zeros(N)
    rhs_vec[2 * Lh + np.nonzero(mesh.clamped_vertices)[0]] = 1.0

    # read-out points: a ring of radius 0.3 r_accel around the test point plus the centre
    # (stand-in for the P1 disc mesh `accTh`, pyFFInterface.py:199-212)
    th = np.linspace(0.0, 2 * np.pi, n_readout, endpoint=False)
    rp = np.stack([xtest + 0.3 * r_accel * np.cos(th), ytest + 0.3 * r_accel * np.sin(th)], 1)
    rp = np.concatenate([rp, [[xtest, ytest]]], axis=0)
    tri, lam = _locate(mesh, rp, nx, ny)
    npts = rp.shape[0]
    interp_L = np.zeros((npts, Lh))
    np.add.at(interp_L, (np.arange(npts)[:, None], mesh.tris[tri]), lam)
    s = (rp - xc[tri]) / hs[tri][:, None]
    x, y = s[:, 0], s[:, 1]
    one, zero = np.ones_like(x), np.zeros_like(x)
    mono = np.stack([one, x, y, x * x, x * y, y * y], axis=-1)
    mdx = np.stack([zero, one, zero, 2 * x, y, zero], axis=-1) / hs[tri][:, None]
    mdy = np.stack([zero, zero, one, zero, x, 2 * y], axis=-1) / hs[tri][:, None]
    Ct = C[tri]
    interp_W = np.zeros((npts, Mh))
    interp_Wx = np.zeros((npts, Mh))
    interp_Wy = np.zeros((npts, Mh))
    rr = np.arange(npts)[:, None]
    np.add.at(interp_W, (rr, mdof[tri]), np.einsum("pm,pmj->pj", mono, Ct))
    np.add.at(interp_Wx, (rr, mdof[tri]), np.einsum("pm,pmj->pj", mdx, Ct))
    np.add.at(interp_Wy, (rr, mdof[tri]), np.einsum("pm,pmj->pj", mdy, Ct))

    return SyntheticPlate(mats, rhs_vec, interp_W, interp_L, Lh, Mh, mesh, interp_Wx,
                          interp_Wy, (xtest, ytest))


# named sizes of SURVEY.md section 8(d)
NAMED_SIZES = {
    "tiny": (6, 2),          # N = 102
    "small": (30, 4),        # N ~ 1.0e3   (config 0 scale: sh_i.edp has ny=2, nx=30)
    "cfg0": (60, 6),         # N ~ 2.7e3
    "cfg1": (200, 16),       # N = 20 067  (BASELINE.json configs[1])
    "40k": (130, 50),        # N = 39 723
    "cfg4": (210, 80),       # N = 101 963 (configs[3])
    "cfg5": (160, 50),       # N ~ 4.9e4   (configs[4])
}
