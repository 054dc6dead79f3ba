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


def mesh_from_triangulation(points, tris, clamped_vertex_mask) -> PlateMesh:
    """A `PlateMesh` from ANY conforming triangulation (what FreeFem++'s `buildmesh` hands the
    reference for an arbitrary `.edp` geometry, `pyFFInterface.py:279-336`): orients the
    triangles counter-clockwise, numbers the edges, drops unused vertices, and flags as clamped
    the edges whose two vertices are clamped AND that lie on the boundary."""
    points = np.asarray(points, dtype=np.float64)
    tris = np.asarray(tris, dtype=np.int64)
    used = np.zeros(points.shape[0], dtype=bool)
    used[tris.ravel()] = True
    remap = np.cumsum(used) - 1
    points, tris = points[used], remap[tris]
    clamped_v = np.asarray(clamped_vertex_mask, dtype=bool)[used]
    p = points[tris]
    det = ((p[:, 1, 0] - p[:, 0, 0]) * (p[:, 2, 1] - p[:, 0, 1])
           - (p[:, 2, 0] - p[:, 0, 0]) * (p[:, 1, 1] - p[:, 0, 1]))
    if np.any(det == 0):
        raise ValueError("degenerate triangle in the mesh")
    flip = det < 0
    tris[flip] = tris[flip][:, [0, 2, 1]]
    loc = np.stack([tris[:, [1, 2]], tris[:, [2, 0]], tris[:, [0, 1]]], axis=1)
    loc_sorted = np.sort(loc, axis=2)
    nv = points.shape[0]
    key = loc_sorted[..., 0] * nv + loc_sorted[..., 1]
    ukey, inv, cnt = np.unique(key.ravel(), return_inverse=True, return_counts=True)
    if np.any(cnt > 2):
        raise ValueError("non-manifold mesh: an edge belongs to more than two triangles")
    edges = np.stack([ukey // nv, ukey % nv], axis=1)
    boundary = cnt == 1
    clamped_e = clamped_v[edges[:, 0]] & clamped_v[edges[:, 1]] & boundary
    ext = points.max(axis=0) - points.min(axis=0)
    return PlateMesh(points, tris, edges, inv.reshape(-1, 3), clamped_v, clamped_e, float(ext[0]), float(ext[1]))


def shaped_plate_mesh(nx: int, ny: int, Lx: float, Ly: float, keep) -> PlateMesh:
    """Structured triangulation of the rectangle with the cells whose centre fails
    `keep(xc, yc)` removed (L-shapes, notches, holes with a staircase boundary): a "complex
    geometry" without a mesh generator.  The edge x = Lx stays the clamped one."""
    base = rectangle_mesh(nx, ny, Lx, Ly)
    cent = base.points[base.tris].mean(axis=1)
    # both triangles of a cell share the cell centre up to the diagonal split: decide per cell
    I, J = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    xc = (I.ravel() + 0.5) * Lx / nx
    yc = -Ly / 2 + (J.ravel() + 0.5) * Ly / ny
    keep_cell = np.asarray(keep(xc, yc), dtype=bool)
    keep_tri = np.concatenate([keep_cell, keep_cell])
    del cent
    tol = 1e-12 * max(Lx, Ly)
    return mesh_from_triangulation(base.points, base.tris[keep_tri], np.abs(base.points[:, 0] - Lx) < tol)


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


def _locate(mesh: PlateMesh, pts: np.ndarray):
    """Triangle id and barycentric coordinates of points in ANY triangulation: the triangle
    whose smallest barycentric coordinate is largest (inside: >= 0; a point in a hole or
    outside snaps to the nearest triangle in that sense)."""
    p = mesh.points[mesh.tris]                                     # [nt,3,2]
    T = np.stack([p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]], axis=2)   # [nt,2,2]
    Tinv = np.linalg.inv(T)
    tri = np.zeros(pts.shape[0], dtype=np.int64)
    lam = np.zeros((pts.shape[0], 3))
    for k, x in enumerate(pts):
        l12 = np.einsum("tij,tj->ti", Tinv, x[None, :] - p[:, 0])
        lall = np.concatenate([1 - l12.sum(1, keepdims=True), l12], axis=1)
        t = int(np.argmax(lall.min(axis=1)))
        tri[k], lam[k] = t, lall[t]
    return tri, lam


def make_plate(nx: int, ny: int, Lx: float = 99.9e-3, Ly: float = 20e-3,
               r_accel: float = 3.8e-3, n_readout: int = 16, keep=None) -> SyntheticPlate:
    """Rectangular cantilever plate (`keep=None`) or a shaped one: `keep(xc, yc) -> bool` per
    cell centre (see `shaped_plate_mesh`).  The accelerometer sits in the corner of the free
    end, `sh_i.edp:11-15`."""
    mesh = rectangle_mesh(nx, ny, Lx, Ly) if keep is None else shaped_plate_mesh(nx, ny, Lx, Ly, keep)
    return plate_from_mesh(mesh, (r_accel, Ly / 2 - r_accel), r_accel, n_readout)


def plate_from_mesh(mesh: PlateMesh, test_point, r_accel: float = 3.8e-3, n_readout: int = 16) -> SyntheticPlate:
    """The 26 component matrices, right-hand side and read-out operators of
    `load_matrices_unsymm` (`pyFFInterface.py:169-509`) on an ARBITRARY conforming
    triangulation: P1 for (u, v), Morley for w, Dirichlet rows with tgv = -1 on the clamped
    vertices / edges of `mesh`, accelerometer disc of radius `r_accel` around `test_point`."""
    Ly = mesh.Ly
    nv, ne, nt = mesh.points.shape[0], mesh.edges.shape[0], mesh.tris.shape[0]
    Lh, Mh = nv, nv + ne
    N = 2 * Lh + Mh

    area, gx, gy = _p1_gradients(mesh)
    xtest, ytest = test_point
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
    tri, lam = _locate(mesh, rp)
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


def _lshape_with_hole(Lx, Ly):
    """L-shaped plate (the quadrant x < 0.45 Lx, y < -0.1 Ly removed) with a circular hole."""
    def keep(xc, yc):
        notch = (xc < 0.45 * Lx) & (yc < -0.1 * Ly)
        hole = (xc - 0.7 * Lx) ** 2 + (yc + 0.1 * Ly) ** 2 < (0.18 * Ly) ** 2
        return ~(notch | hole)
    return keep


# "complex-geometry" plates (BASELINE configs[3]): same outer box, cells removed
SHAPED_SIZES = {
    "tinyL": (12, 6),        # N ~ 5e2
    "smallL": (40, 12),      # N ~ 2.6e3
    "cfg4L": (240, 96),      # N ~ 1.0e5   (configs[3], complex geometry)
}


def make_named(name: str) -> SyntheticPlate:
    """`make_plate` for a name of NAMED_SIZES (rectangles) or SHAPED_SIZES (L-shape + hole)."""
    if name in SHAPED_SIZES:
        nx, ny = SHAPED_SIZES[name]
        return make_plate(nx, ny, keep=_lshape_with_hole(99.9e-3, 20e-3))
    return make_plate(*NAMED_SIZES[name])


def save_plate(path, plate: SyntheticPlate) -> None:
    """On-disk format of the hot path's inputs (the tuple of `pyFFInterface.py:503-509` plus
    the mesh): one .npz with the 26 component matrices as concatenated COO triplets
    (`coo_row`, `coo_col`, `coo_data`, `coo_ptr[27]`; explicit zeros kept, they are part of the
    pattern), `rhs_vec`, the dense read-out operators and the mesh arrays."""
    rows = [m.row.astype(np.int64) for m in plate.mats]
    ptr = np.concatenate([[0], np.cumsum([r.size for r in rows])])
    np.savez_compressed(
        path, format=np.array("jax_plate_b200.plate.v1"), names=np.array(COMPONENT_NAMES),
        N=np.array(plate.N), Lh=np.array(plate.Lh_size), Mh=np.array(plate.Mh_size), coo_ptr=ptr,
        coo_row=np.concatenate(rows), coo_col=np.concatenate([m.col.astype(np.int64) for m in plate.mats]),
        coo_data=np.concatenate([m.data for m in plate.mats]), rhs_vec=plate.rhs_vec,
        interp_mat=plate.interp_mat, interp_mat_Lh=plate.interp_mat_Lh, interp_mat_Wx=plate.interp_mat_Wx,
        interp_mat_Wy=plate.interp_mat_Wy, test_point=np.array(plate.test_point),
        mesh_points=plate.mesh.points, mesh_tris=plate.mesh.tris, mesh_edges=plate.mesh.edges,
        mesh_tri_edges=plate.mesh.tri_edges, mesh_clamped_vertices=plate.mesh.clamped_vertices,
        mesh_clamped_edges=plate.mesh.clamped_edges, mesh_extent=np.array([plate.mesh.Lx, plate.mesh.Ly]))


def load_plate(path) -> SyntheticPlate:
    z = np.load(path, allow_pickle=False)
    if str(z["format"]) != "jax_plate_b200.plate.v1":
        raise ValueError(f"{path}: not a plate file of this package")
    N, ptr = int(z["N"]), z["coo_ptr"]
    mats = [sps.coo_matrix((z["coo_data"][a:b], (z["coo_row"][a:b], z["coo_col"][a:b])), shape=(N, N))
            for a, b in zip(ptr[:-1], ptr[1:])]
    mesh = PlateMesh(z["mesh_points"], z["mesh_tris"], z["mesh_edges"], z["mesh_tri_edges"],
                     z["mesh_clamped_vertices"], z["mesh_clamped_edges"], float(z["mesh_extent"][0]),
                     float(z["mesh_extent"][1]))
    return SyntheticPlate(mats, z["rhs_vec"], z["interp_mat"], z["interp_mat_Lh"], int(z["Lh"]), int(z["Mh"]),
                          mesh, z["interp_mat_Wx"], z["interp_mat_Wy"], tuple(z["test_point"]))


# ------------------------------------------------------------------------------------------
# Mid-plane symmetric (bending-only) plates: the output of `load_matrices_symm`
# (`pyFFInterface.py:10-167`) without FreeFem++.  One Morley field w; the clamped DOFs are
# ELIMINATED (the reference keeps the free rows / columns and moves the boundary values to the
# right-hand side, `evaluateMatrixAndRHS` `:81-117`).  The reference densifies every matrix
# (`.todense()`, `:99`); here they stay sparse (CSR) -- `dense=True` gives the reference's arrays.

SYMM_MODULI = ("11", "12", "16", "22", "26", "66")      # MODULI_INDICES of the reference


def bending_plate_from_mesh(mesh: PlateMesh, test_point, r_accel: float = 3.8e-3, dense: bool = False) -> dict:
    """Keys as returned by `load_matrices_symm` (`pyFFInterface.py:149-166`): `Ks` [6] and `M`,
    `L`, `MCorrection`, `LCorrection` on the free DOFs, their boundary vectors `fKs` [6, n],
    `fM`, `fL`, ..., `interpolation_vector` [n], `interpolation_value_from_bc`,
    `test_point_coord`, `constrained_idx` (bool over all DOFs), `mesh`, `boundary_value`."""
    nv, ne = mesh.points.shape[0], mesh.edges.shape[0]
    Mh = nv + ne
    area, _, _ = _p1_gradients(mesh)
    xtest, ytest = test_point
    cent = mesh.points[mesh.tris].mean(axis=1)
    ind = ((cent[:, 0] - xtest) ** 2 + (cent[:, 1] - ytest) ** 2 <= r_accel ** 2 + 1e-8).astype(np.float64)
    a = area[:, None, None]
    C, xc, hs = _morley_coefficients(mesh)
    mdof = np.concatenate([mesh.tris, nv + mesh.tri_edges], axis=1)
    wxx = 2.0 * C[:, 3, :] / hs[:, None] ** 2
    wxy = C[:, 4, :] / hs[:, None] ** 2
    wyy = 2.0 * C[:, 5, :] / hs[:, None] ** 2
    Rm = np.repeat(mdof[:, :, None], 6, axis=2)
    Cm = np.repeat(mdof[:, None, :], 6, axis=1)

    def form(dv, du):                                  # area * d2(v)_i * d2(u)_j, rows = test function v
        return _scatter(Mh, Mh, Rm, Cm, a * dv[:, :, None] * du[:, None, :])

    # varf K11 .. K66 of pyFFInterface.py:50-55
    total = {
        "K11": form(wxx, wxx),
        "K12": form(wxx, wyy) + form(wyy, wxx),
        "K16": 2.0 * (form(wxx, wxy) + form(wxy, wxx)),
        "K22": form(wyy, wyy),
        "K26": 2.0 * (form(wyy, wxy) + form(wxy, wyy)),
        "K66": 4.0 * form(wxy, wxy),
    }
    pts = mesh.points[mesh.tris]
    qp = np.einsum("qk,tkd->tqd", _QUAD_BARY, pts)
    qs = (qp - xc[:, None, :]) / hs[:, None, None]
    x, y = qs[..., 0], qs[..., 1]
    one, zero = np.ones_like(x), np.zeros_like(x)
    mono = np.stack([one, x, y, x * x, x * y, y * y], axis=-1)
    mdx = np.stack([zero, one, zero, 2 * x, y, zero], axis=-1) / hs[:, None, None]
    mdy = np.stack([zero, zero, one, zero, x, 2 * y], axis=-1) / hs[:, None, None]
    val, vdx, vdy = mono @ C, mdx @ C, mdy @ C
    wq = _QUAD_W[None, :] * area[:, None]
    m = np.einsum("tq,tqi,tqj->tij", wq, val, val)
    l = np.einsum("tq,tqi,tqj->tij", wq, vdx, vdx) + np.einsum("tq,tqi,tqj->tij", wq, vdy, vdy)
    total["M"] = _scatter(Mh, Mh, Rm, Cm, m)
    total["L"] = _scatter(Mh, Mh, Rm, Cm, l)
    total["MCorrection"] = _scatter(Mh, Mh, Rm, Cm, m * ind[:, None, None])
    total["LCorrection"] = _scatter(Mh, Mh, Rm, Cm, l * ind[:, None, None])

    # boundary values: funcBC = 1 on the clamped vertex values, 0 on the clamped normal derivatives
    constrained = np.concatenate([mesh.clamped_vertices, mesh.clamped_edges]).astype(bool)
    free = ~constrained
    f_bc = np.zeros(Mh)
    f_bc[:nv][mesh.clamped_vertices] = 1.0

    def reduce(Kt):                                    # evaluateMatrixAndRHS, :81-117
        Kt = Kt.tocsr()
        K = Kt[free][:, free].tocsr()
        rhs = -np.asarray(Kt[constrained][:, free].T @ f_bc[constrained]).ravel()
        return (np.asarray(K.todense()) if dense else K), rhs

    out = {}
    for name in ("M", "L", "MCorrection", "LCorrection"):
        out[name], out["f" + name] = reduce(total[name])
    Ks, fKs = zip(*[reduce(total["K" + ss]) for ss in SYMM_MODULI])
    out["Ks"] = np.array(Ks) if dense else list(Ks)
    out["fKs"] = np.array(fKs)
    # value of w at the test point (row 0 of `interpC`, :35-36)
    tri, _ = _locate(mesh, np.array([[xtest, ytest]]))
    s = (np.array([xtest, ytest]) - xc[tri[0]]) / hs[tri[0]]
    mono_t = np.array([1.0, s[0], s[1], s[0] * s[0], s[0] * s[1], s[1] * s[1]])
    iv = np.zeros(Mh)
    np.add.at(iv, mdof[tri[0]], mono_t @ C[tri[0]])
    out["interpolation_value_from_bc"] = float(iv[constrained] @ f_bc[constrained])
    out["interpolation_vector"] = iv[free]
    out["test_point_coord"] = np.array([xtest, ytest])
    out["constrained_idx"] = constrained
    out["mesh"] = mesh
    out["boundary_value"] = f_bc
    return out


def make_bending_plate(nx: int, ny: int, Lx: float = 99.9e-3, Ly: float = 20e-3, r_accel: float = 3.8e-3,
                       keep=None, dense: bool = False) -> dict:
    """`load_matrices_symm` output for the cantilever strip of `make_plate` (same meshes)."""
    mesh = rectangle_mesh(nx, ny, Lx, Ly) if keep is None else shaped_plate_mesh(nx, ny, Lx, Ly, keep)
    return bending_plate_from_mesh(mesh, (r_accel, Ly / 2 - r_accel), r_accel, dense)
