"""Structured synthetic meshes of BASELINE.json's configs, in the reference's conventions.

Domain [-1,1]^2, nx*ny rectangles, vertices numbered row-major; triangles: each cell split into two
CCW triangles along the same diagonal, element order cell-major; quads: CCW cells.  All boundary
faces become BCTag.Wall through MeshManager.buildBCTable, as after readMesh.  Optional interior-vertex
jitter (+-0.2 h_cell, rng seed 20260101) turns quads genuinely bilinear.
"""
import numpy as np

from .mesh_manager import MeshManager

CONFIG_GRIDS = {
    "tri_1M": ("tri", 1000, 500),
    "tri_2M": ("tri", 1000, 1000),
    "tri_4M": ("tri", 2000, 1000),
    "quad_2M": ("quad", 2000, 1000),
    "tri_16M": ("tri", 4000, 2000),
}


def structured_arrays(shape, nx, ny, jitter=0.0, x0=-1.0, x1=1.0, y0=-1.0, y1=1.0, block=None):
    """Returns (EToV [K][Nf] int64, Vert [Nv][3]).  block=(bx, by): elements ordered by blocks of bx x by cells
    (cell-major inside a block, blocks row-major) instead of plain cell-major -- the locality study of SURVEY.md §8d;
    the mesh is the same up to an element permutation."""
    xs = np.linspace(x0, x1, nx + 1)
    ys = np.linspace(y0, y1, ny + 1)
    X, Y = np.meshgrid(xs, ys)  # [ny+1][nx+1], vertex id = j*(nx+1)+i
    if jitter > 0:
        rng = np.random.default_rng(20260101)
        hx, hy = (x1 - x0) / nx, (y1 - y0) / ny
        X = X.copy()
        Y = Y.copy()
        X[1:-1, 1:-1] += jitter * hx * rng.uniform(-1, 1, size=(ny - 1, nx - 1))
        Y[1:-1, 1:-1] += jitter * hy * rng.uniform(-1, 1, size=(ny - 1, nx - 1))
    vert = np.stack([X.reshape(-1), Y.reshape(-1), np.zeros(X.size)], axis=1)
    i, j = np.meshgrid(np.arange(nx, dtype=np.int64), np.arange(ny, dtype=np.int64))
    if block:
        bx, by = block
        if nx % bx or ny % by:
            raise ValueError("block %s does not divide the %dx%d grid" % (block, nx, ny))
        # [ny/by][by][nx/bx][bx] -> [ny/by][nx/bx][by][bx]
        i = i.reshape(ny // by, by, nx // bx, bx).transpose(0, 2, 1, 3)
        j = j.reshape(ny // by, by, nx // bx, bx).transpose(0, 2, 1, 3)
    v00 = (j * (nx + 1) + i).reshape(-1)
    v10, v01, v11 = v00 + 1, v00 + nx + 1, v00 + nx + 2
    if shape == "tri":
        lower = np.stack([v00, v10, v11], axis=1)
        upper = np.stack([v00, v11, v01], axis=1)
        etov = np.stack([lower, upper], axis=1).reshape(-1, 3)
    elif shape == "quad":
        etov = np.stack([v00, v10, v11, v01], axis=1)
    else:
        raise ValueError(shape)
    return etov, vert


def shuffle_numbering(etov, vert, seed):
    """Random element order, vertex numbering and starting vertex per element (orientation kept): what an unstructured
    mesh generator's output looks like to the setup code and to the device element order."""
    rng = np.random.default_rng(seed)
    K, Nf = etov.shape
    vperm = rng.permutation(len(vert))  # new id of old vertex
    inv = np.empty_like(vperm)
    inv[vperm] = np.arange(len(vert))
    etov = vperm[etov]
    rot = rng.integers(0, Nf, K)
    etov = np.take_along_axis(etov, (np.arange(Nf)[None, :] + rot[:, None]) % Nf, axis=1)
    return np.ascontiguousarray(etov[rng.permutation(K)]), np.ascontiguousarray(vert[inv])


def structured_mesh(shape, nx, ny, jitter=0.0, block=None, shuffle=None):
    etov, vert = structured_arrays(shape, nx, ny, jitter, block=block)
    if shuffle is not None:
        etov, vert = shuffle_numbering(etov, vert, shuffle)
    m = MeshManager()
    m.buildMesh(etov, vert)
    return m


def provision(mesh, N):
    from .quad_nodes import QuadNodesProvisioner
    from .triangle_nodes import TriangleNodesProvisioner

    P = QuadNodesProvisioner if mesh.get_NumFaces() == 4 else TriangleNodesProvisioner
    return P(N, mesh)
