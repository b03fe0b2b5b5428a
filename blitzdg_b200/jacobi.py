"""Jacobi polynomials, Gauss / Gauss-Lobatto nodes and 1-D Vandermonde matrices.

Restates the conventions of /root/reference/src/JacobiBuilders.cpp:18-127 and
/root/reference/include/VandermondeBuilders.hpp:45-74.  Setup-time host code.
"""
from math import gamma, sqrt

import numpy as np

from . import lapack


def jacobi_polynomial(x, alpha, beta, N):
    """Orthonormal Jacobi P_N^{(alpha,beta)}(x) (JacobiBuilders.cpp:18-52)."""
    x = np.asarray(x, dtype=np.float64)
    gamma0 = 2.0 ** (alpha + beta + 1) / (alpha + beta + 1) * gamma(alpha + 1) * gamma(beta + 1) / gamma(alpha + beta + 1)
    p0 = np.full_like(x, 1.0 / sqrt(gamma0))
    if N == 0:
        return p0
    gamma1 = (alpha + 1) * (beta + 1) / (alpha + beta + 3) * gamma0
    p1 = ((alpha + beta + 2) * x / 2 + (alpha - beta) / 2) / sqrt(gamma1)
    if N == 1:
        return p1
    aold = 2 / (2 + alpha + beta) * sqrt((alpha + 1) * (beta + 1) / (alpha + beta + 3))
    pm, p = p0, p1
    for i in range(1, N):
        h1 = 2 * i + alpha + beta
        anew = 2 / (h1 + 2) * sqrt((i + 1) * (i + 1 + alpha + beta) * (i + 1 + alpha) * (i + 1 + beta) / (h1 + 1) / (h1 + 3))
        bnew = -(alpha * alpha - beta * beta) / h1 / (h1 + 2)
        pn = 1 / anew * (-aold * pm + (x - bnew) * p)
        pm, p = p, pn
        aold = anew
    return p


def grad_jacobi(x, alpha, beta, N):
    """d/dx of the orthonormal Jacobi polynomial (JacobiBuilders.cpp:117-127)."""
    x = np.asarray(x, dtype=np.float64)
    if N == 0:
        return np.zeros_like(x)
    return sqrt(N * (N + alpha + beta + 1)) * jacobi_polynomial(x, alpha + 1, beta + 1, N - 1)


def jacobi_gauss(alpha, beta, N):
    """Gauss-Jacobi nodes and weights via the symmetric tridiagonal eigenproblem (JacobiBuilders.cpp:54-95)."""
    if N == 0:
        return np.array([-(alpha - beta) / (alpha + beta + 2)]), np.array([2.0])
    eps = np.finfo(np.float64).eps
    J = np.zeros((N + 1, N + 1))
    for i in range(N + 1):
        h1 = 2.0 * i + alpha + beta
        # the reference evaluates this 0/0 for alpha+beta=0 at i=0 and then overwrites J(0,0)=0
        with np.errstate(all="ignore"):
            J[i, i] = np.float64(-0.5 * (alpha * alpha - beta * beta)) / np.float64(h1 + 2.0) / np.float64(h1)
        if i < N:
            J[i, i + 1] = 2.0 / (h1 + 2) * sqrt((i + 1) * ((i + 1) + alpha + beta) * ((i + 1) + alpha) * ((i + 1) + beta) / (h1 + 1) / (h1 + 3))
    if (alpha + beta) < 10 * eps:
        J[0, 0] = 0.0
    J = J + J.T
    x, vecs = lapack.eigh(J)
    gamma0 = 2.0 ** (alpha + beta + 1) / (alpha + beta + 1) * gamma(alpha + 1) * gamma(beta + 1) / gamma(alpha + beta + 1)
    w = vecs[0, :] ** 2 * gamma0
    return x, w


def jacobi_gauss_lobatto(alpha, beta, N):
    """Gauss-Lobatto nodes (JacobiBuilders.cpp:97-115)."""
    x = np.zeros(N + 1)
    x[0], x[N] = -1.0, 1.0
    if N == 1:
        return x
    xint, _ = jacobi_gauss(alpha + 1.0, beta + 1.0, N - 2)
    x[1:N] = xint
    return x


def vandermonde_1d(r, ncols=None):
    """V[i][j] = P_j^{(0,0)}(r_i) (VandermondeBuilders.hpp:45-58)."""
    r = np.asarray(r, dtype=np.float64)
    n = len(r) if ncols is None else ncols
    V = np.zeros((len(r), n))
    for j in range(n):
        V[:, j] = jacobi_polynomial(r, 0.0, 0.0, j)
    return V


def grad_vandermonde_1d(r):
    """Column i = d/dr P_i^{(0,0)}(r) (VandermondeBuilders.hpp:65-74)."""
    r = np.asarray(r, dtype=np.float64)
    n = len(r)
    DV = np.zeros((n, n))
    for i in range(n):
        DV[:, i] = grad_jacobi(r, 0.0, 0.0, i)
    return DV
