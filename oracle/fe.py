"""ORACLE (test infrastructure, not product code): 1-D finite element data of FE_Q(p) + QGauss(p+1).

CPU restatement of what deal.II's ShapeInfo provides to the reference's FEEvaluation
(include/operator.h:22, 450-472).  deal.II itself is not vendored in /root/reference and
cannot be built here, so parity is UNPINNED against deal.II binaries; the pins are the analytic
known-answer tests in tests/ (SURVEY.md 8(c)).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import this package.
"""
import numpy as np
from numpy.polynomial import legendre as L


def gauss_rule(n):
    """Gauss-Legendre points/weights on [0,1] (QGauss<1>(n))."""
    x, w = L.leggauss(n)
    return 0.5 * (x + 1.0), 0.5 * w


def gauss_lobatto_nodes(n):
    """Gauss-Lobatto points on [0,1]: support points of FE_Q(n-1) (QGaussLobatto<1>(n))."""
    if n == 1:
        return np.array([0.5])
    if n == 2:
        return np.array([0.0, 1.0])
    N = n - 1
    c = np.zeros(N + 1)
    c[N] = 1.0
    interior = L.legroots(L.legder(c))  # roots of P_N'
    x = np.concatenate(([-1.0], np.sort(interior), [1.0]))
    x = 0.5 * (x - x[::-1])  # symmetrise
    return 0.5 * (x + 1.0)


def lagrange_basis(nodes, x):
    """V[q, i] = l_i(x_q) and D[q, i] = l_i'(x_q) for the Lagrange basis on `nodes`."""
    nodes = np.asarray(nodes, dtype=np.longdouble)
    x = np.atleast_1d(np.asarray(x, dtype=np.longdouble))
    n = nodes.size
    V = np.ones((x.size, n), dtype=np.longdouble)
    D = np.zeros((x.size, n), dtype=np.longdouble)
    for i in range(n):
        for m in range(n):
            if m != i:
                V[:, i] *= (x - nodes[m]) / (nodes[i] - nodes[m])
        for k in range(n):
            if k == i:
                continue
            t = np.ones(x.size, dtype=np.longdouble) / (nodes[i] - nodes[k])
            for m in range(n):
                if m != i and m != k:
                    t *= (x - nodes[m]) / (nodes[i] - nodes[m])
            D[:, i] += t
    return V.astype(np.float64), D.astype(np.float64)


class FE1D:
    """Everything 1-D the cell kernels need, in deal.II's terms (SURVEY.md A.3 'shape_info')."""

    def __init__(self, degree):
        self.degree = p = degree
        self.n = n = p + 1
        self.nodes = gauss_lobatto_nodes(n)
        self.xq, self.wq = gauss_rule(n)
        # shape_values[q][i], shape_gradients[q][i]
        self.B, self.G = lagrange_basis(self.nodes, self.xq)
        # collocation derivative: Lagrange basis on the Gauss points, differentiated at the Gauss points
        _, self.Dc = lagrange_basis(self.xq, self.xq)
        # 1-D mass / stiffness (for the closed-form checks)
        self.M = self.B.T @ (self.wq[:, None] * self.B)
        self.K = self.G.T @ (self.wq[:, None] * self.G)
        # subface interpolation matrices: parent basis at the child's nodes
        self.S = [lagrange_basis(self.nodes, 0.5 * c + 0.5 * self.nodes)[0] for c in (0, 1)]
        # h-embedding: parent basis at the union of both children's nodes, (2p+1) x (p+1)
        self.Ph = np.vstack([self.S[0], self.S[1][1:]])


def p_embedding(degree_fine, degree_coarse):
    """coarse 1-D basis at the fine support points, (pf+1) x (pc+1)"""
    return lagrange_basis(gauss_lobatto_nodes(degree_coarse + 1), gauss_lobatto_nodes(degree_fine + 1))[0]
