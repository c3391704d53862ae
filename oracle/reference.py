"""ORACLE (test infrastructure, not product code): CPU restatement of the reference's hot path.

Restates, in numpy/scipy, what peterrum/dealii-multigrid computes through deal.II on the path
  Operator::vmult & co.        include/operator.h:152-242, 450-493
  PreconditionChebyshev        multigrid_throughput.cc:847-883       (deal.II lac/precondition.h)
  MGTwoLevelTransfer           multigrid_throughput.cc:1600-1621     (deal.II mg_transfer_global_coarsening)
  Multigrid / PreconditionMG   multigrid_throughput.cc:1091-1133     (deal.II multigrid.h)
  SolverCG + ReductionControl  multigrid_throughput.cc:1143-1144, 1625-1635
following SURVEY.md Appendix A.  deal.II is a third-party dependency that is NOT vendored in
/root/reference (README.md:26 clones deal.II master, ~9.4); it cannot be compiled or imported here, and
the reference ships no tests or golden vectors for this path => PARITY UNPINNED against deal.II binaries.
Substitute pins live in tests/: assembled-matrix cross-checks, polynomial reproduction,
symmetry, R = P^T, explicit Chebyshev polynomials, scipy eigenvalues.

The index maps (cell -> DoF entity tables, hanging-node masks, constrained lists, transfer cell maps)
are INPUTS shared with the product, exactly as they would be taken from deal.II's MatrixFree in a real
integration; this module verifies them through geometry (polynomial reproduction), never by trust.

Formulations here are deliberately different from the CUDA kernels: dense element matrices from Gauss
quadrature, per-node Kronecker constraint rows, explicit scipy sparse matrices, deal.II's 12-sweep
collocation evaluation for the matrix-free variant.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import this package.
"""
import numpy as np
import scipy.sparse as sp

from .fe import FE1D, p_embedding

FLAG_BIT = np.uint32(0x80000000)
INDEX_MASK = np.uint32(0x7FFFFFFF)


# ----------------------------------------------------------------------------------------- index maps
def node_entity_tables(p):
    """entity index (ternary) and offset inside the entity for every lexicographic node of FE_Q(p)"""
    n = p + 1
    ent = np.zeros(n ** 3, dtype=np.int64)
    off = np.zeros(n ** 3, dtype=np.int64)
    for k in range(n):
        for j in range(n):
            for i in range(n):
                t = [0 if a == 0 else (2 if a == p else 1) for a in (i, j, k)]
                stride, o = 1, 0
                for a, ta in zip((i, j, k), t):
                    if ta == 1:
                        o += (a - 1) * stride
                        stride *= (p - 1)
                ent[(k * n + j) * n + i] = t[0] + 3 * t[1] + 9 * t[2]
                off[(k * n + j) * n + i] = o
    return ent, off


def expand_table(table, p):
    """(n_cells, 27) entity table -> (index (n_cells, n^3) int64, flag (n_cells, n^3) bool)"""
    ent, off = node_entity_tables(p)
    t = np.asarray(table, dtype=np.uint32)[:, ent]
    return (t & INDEX_MASK).astype(np.int64) + off[None, :], (t & FLAG_BIT) != 0


def cell_constraint_matrix(mask, fe):
    """dense C_e (n^3 x n^3): local nodal values = C_e @ slot values, for one hanging-node mask.

    Per-node Kronecker rows: along direction d the 1-D subface matrix S[subcell_d]
    applies to a node iff the node lies in a constrained face whose normal is not d, or on the
    constrained edge along d."""
    p, n = fe.degree, fe.n
    s = [(mask >> d) & 1 for d in range(3)]
    face = [(mask >> (3 + d)) & 1 for d in range(3)]
    edge = [(mask >> (6 + d)) & 1 for d in range(3)]
    eye = np.eye(n)
    C = np.zeros((n ** 3, n ** 3))
    for k in range(n):
        for j in range(n):
            for i in range(n):
                idx = (i, j, k)
                rows = []
                for d in range(3):
                    a, b = [x for x in range(3) if x != d]
                    on_a, on_b = idx[a] == s[a] * p, idx[b] == s[b] * p
                    applies = (face[a] and on_a) or (face[b] and on_b) or (edge[d] and on_a and on_b)
                    rows.append(fe.S[s[d]][idx[d]] if applies else eye[idx[d]])
                C[(k * n + j) * n + i] = np.einsum("k,j,i->kji", rows[2], rows[1], rows[0]).ravel()
    return C


def slot_replaced(mask, p):
    """bool (27,): entity slots that hold the coarse neighbour's (parent's) entity for this mask"""
    s = [(mask >> d) & 1 for d in range(3)]
    out = np.zeros(27, dtype=bool)
    for e in range(27):
        t = (e % 3, (e // 3) % 3, e // 9)
        for d in range(3):
            a, b = (d + 1) % 3, (d + 2) % 3
            if ((mask >> (3 + d)) & 1) and t[d] == 2 * s[d]:
                out[e] = True
            if ((mask >> (6 + d)) & 1) and t[a] == 2 * s[a] and t[b] == 2 * s[b]:
                out[e] = True
    return out


class LevelMaps:
    """Index maps of one level (inputs): what MatrixFree::reinit would provide."""

    def __init__(self, degree, eval_table, own_table, mask, cell_level, cell_xyz, constrained, n_dofs, root_length=2.0):
        self.degree = degree
        self.fe = FE1D(degree)
        self.n_cells = len(cell_level)
        self.n_dofs = int(n_dofs)
        self.eval_table = np.asarray(eval_table, dtype=np.uint32)
        self.own_table = np.asarray(own_table, dtype=np.uint32)
        self.eval_idx, self.eval_flag = expand_table(eval_table, degree)
        self.own_idx, self.own_flag = expand_table(own_table, degree)
        self.mask = np.asarray(mask, dtype=np.int64)
        self.cell_level = np.asarray(cell_level, dtype=np.int64)
        self.cell_xyz = np.asarray(cell_xyz, dtype=np.int64)
        self.h = root_length / (2.0 ** self.cell_level)
        self.constrained = np.asarray(constrained, dtype=np.int64)
        self._C = {}

    @classmethod
    def from_dofmap(cls, d):
        return cls(d.degree, d.eval_table(), d.own_table(), d.mask(), d.cell_level(), d.cell_xyz(), d.constrained(),
                   d.n_owned + d.n_ghost)

    def C(self, mask):
        mask = int(mask)
        if mask not in self._C:
            self._C[mask] = cell_constraint_matrix(mask, self.fe)
        return self._C[mask]

    def node_coordinates(self):
        """(n_cells, n^3, 3): coordinates of each cell's own nodes"""
        n = self.fe.n
        x1 = self.fe.nodes
        ref = np.stack(np.meshgrid(x1, x1, x1, indexing="ij"), axis=-1)  # [i,j,k,:]
        ref = ref.transpose(2, 1, 0, 3).reshape(n ** 3, 3)               # lexicographic, x fastest
        origin = -1.0 + self.h[:, None] * self.cell_xyz
        return origin[:, None, :] + self.h[:, None, None] * ref[None, :, :]

    def dof_coordinates(self):
        """(n_dofs, 3) support point of every DoF, from the cells' own tables"""
        xyz = np.full((self.n_dofs, 3), np.nan)
        xyz[self.own_idx.ravel()] = self.node_coordinates().reshape(-1, 3)
        return xyz


# ----------------------------------------------------------------------------------------- operator
def element_matrix_reference(fe):
    """dense Laplace element matrix on the unit cube from Gauss quadrature of grad phi_i . grad phi_j"""
    B, G, w = fe.B, fe.G, fe.wq
    n = fe.n
    W = np.einsum("a,b,c->abc", w, w, w).ravel()  # weights, z slowest
    # values of d/dx, d/dy, d/dz of basis (k,j,i) at quadrature point (c,b,a)
    dx = np.einsum("ck,bj,ai->cbakji", B, B, G).reshape(n ** 3, n ** 3)
    dy = np.einsum("ck,bj,ai->cbakji", B, G, B).reshape(n ** 3, n ** 3)
    dz = np.einsum("ck,bj,ai->cbakji", G, B, B).reshape(n ** 3, n ** 3)
    return dx.T @ (W[:, None] * dx) + dy.T @ (W[:, None] * dy) + dz.T @ (W[:, None] * dz)


def cell_vertices(maps, deformation=None):
    """(n_cells, 8, 3) vertex coordinates of the cells (vertices lexicographic, x fastest), optionally moved by
    deformation(xyz) -> xyz (applied to the vertices; the geometry inside a cell stays trilinear, MappingQ1)"""
    corner = np.array([[(v >> d) & 1 for d in range(3)] for v in range(8)], dtype=float)
    origin = -1.0 + maps.h[:, None] * maps.cell_xyz
    xyz = origin[:, None, :] + maps.h[:, None, None] * corner[None, :, :]
    return xyz if deformation is None else deformation(xyz.reshape(-1, 3)).reshape(-1, 8, 3)


def element_matrices_mapped(fe, vertices):
    """dense Laplace element matrices (n_cells, n^3, n^3) of trilinearly mapped hexahedra by Gauss quadrature:
    A_ij = sum_q w_q det J_q (J_q^-T grad phi_i) . (J_q^-T grad phi_j) -- what FEEvaluation computes with a general
    MatrixFree mapping (include/operator.h:450-493), written with dense reference-gradient matrices instead of sweeps"""
    B, G, w = fe.B, fe.G, fe.wq
    n = fe.n
    xq = fe.xq
    W = np.einsum("a,b,c->abc", w, w, w).ravel()                                   # quadrature point (c,b,a), a fastest
    d = [np.einsum("ck,bj,ai->cbakji", *m).reshape(n ** 3, n ** 3) for m in ((B, B, G), (B, G, B), (G, B, B))]
    Q = np.stack(np.meshgrid(xq, xq, xq, indexing="ij"), axis=-1).transpose(2, 1, 0, 3).reshape(n ** 3, 3)
    # derivatives of the 8 trilinear vertex functions at the quadrature points: (nq, 8, 3)
    dN = np.zeros((n ** 3, 8, 3))
    for v in range(8):
        c = [(v >> a) & 1 for a in range(3)]
        val = [Q[:, a] if c[a] else 1.0 - Q[:, a] for a in range(3)]
        der = [1.0 if c[a] else -1.0 for a in range(3)]
        dN[:, v, 0] = der[0] * val[1] * val[2]
        dN[:, v, 1] = val[0] * der[1] * val[2]
        dN[:, v, 2] = val[0] * val[1] * der[2]
    vertices = np.asarray(vertices, dtype=float)
    out = np.zeros((len(vertices), n ** 3, n ** 3))
    for c, X in enumerate(vertices):
        J = np.einsum("vc,qvd->qcd", X, dN)                                         # J[q, c, d] = d x_c / d xi_d
        Jinv = np.linalg.inv(J)
        det = np.linalg.det(J)
        if np.any(det <= 0):
            raise ValueError("inverted cell")
        M = np.einsum("q,qdc,qec->qde", W * det, Jinv, Jinv)                        # w det J^-1 J^-T
        for a in range(3):
            for b in range(3):
                out[c] += d[a].T @ (M[:, a, b][:, None] * d[b])
    return out


def assemble_matrix_mapped(maps, vertices, identity_rows=True):
    """scipy CSR of the constrained operator on mapped cells (cf. assemble_matrix)"""
    Ae = element_matrices_mapped(maps.fe, vertices)
    n3 = maps.fe.n ** 3
    for m in np.unique(maps.mask):
        if (m >> 3) != 0:
            C = maps.C(m)
            sel = maps.mask == m
            Ae[sel] = C.T[None] @ Ae[sel] @ C[None]
    keep = ~maps.eval_flag
    blk = Ae * (keep[:, :, None] & keep[:, None, :])
    rows = np.repeat(maps.eval_idx, n3, axis=1).ravel()
    cols = np.tile(maps.eval_idx, (1, n3)).ravel()
    A = sp.coo_matrix((blk.ravel(), (rows, cols)), shape=(maps.n_dofs, maps.n_dofs)).tocsr()
    if identity_rows:
        dd = np.zeros(maps.n_dofs)
        dd[maps.constrained] = 1.0
        A = A + sp.diags(dd)
    A.eliminate_zeros()
    return A


def assemble_matrix(maps, identity_rows=True):
    """scipy CSR of the constrained operator: sum_e C_e^T (h A_ref) C_e on unconstrained DoFs (+ identity
    rows on constrained DoFs as Operator::vmult adds them, include/operator.h:170-172)"""
    A_ref = element_matrix_reference(maps.fe)
    n3 = maps.fe.n ** 3
    rows, cols, vals = [], [], []
    for m in np.unique(maps.mask):
        cells = np.nonzero(maps.mask == m)[0]
        Ae = A_ref if (m >> 3) == 0 else maps.C(m).T @ A_ref @ maps.C(m)
        idx = maps.eval_idx[cells]
        keep = ~maps.eval_flag[cells]
        blk = maps.h[cells][:, None, None] * Ae[None, :, :] * (keep[:, :, None] & keep[:, None, :])
        rows.append(np.repeat(idx, n3, axis=1).ravel())
        cols.append(np.tile(idx, (1, n3)).ravel())
        vals.append(blk.ravel())
    A = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(maps.n_dofs, maps.n_dofs)).tocsr()
    if identity_rows:
        d = np.zeros(maps.n_dofs)
        d[maps.constrained] = 1.0
        A = A + sp.diags(d)
    A.eliminate_zeros()
    return A


def cell_apply_collocation(fe, u):
    """deal.II's evaluate(gradients) -> quadrature loop -> integrate(gradients) on the unit cube, for a batch
    u[c, k, j, i]: 12 one-dimensional sweeps, returns A_ref u without the h factor."""
    B, Dc, w = fe.B, fe.Dc, fe.wq
    v = np.einsum("ai,ckji->ckja", B, u)
    v = np.einsum("bj,ckja->ckba", B, v)
    v = np.einsum("dk,ckba->cdba", B, v)
    gx = np.einsum("qa,cdba->cdbq", Dc, v)
    gy = np.einsum("qb,cdba->cdqa", Dc, v)
    gz = np.einsum("qd,cdba->cqba", Dc, v)
    W = np.einsum("d,b,a->dba", w, w, w)
    gx, gy, gz = gx * W, gy * W, gz * W
    r = np.einsum("qa,cdbq->cdba", Dc, gx) + np.einsum("qb,cdqa->cdba", Dc, gy) + np.einsum("qd,cqba->cdba", Dc, gz)
    r = np.einsum("dk,cdba->ckba", B, r)
    r = np.einsum("bj,ckba->ckja", B, r)
    r = np.einsum("ai,ckja->ckji", B, r)
    return r


class Operator:
    """Operator<3,1,double> restated (include/operator.h).  edge_* only for local-smoothing levels."""

    def __init__(self, maps, edge_constrained=None, edge_cells=None, has_edge_global=False, rnd=None):
        self.maps = maps
        # rnd: rounding applied to every vector a kernel writes -- identity for FP64 levels, round_f32 models
        # MGNumberType=float (multigrid_throughput.cc:2432-2433: level vectors and operators in float)
        self.rnd = rnd if rnd is not None else (lambda v: v)
        self.edge = np.zeros(0, dtype=np.int64) if edge_constrained is None else np.asarray(edge_constrained, dtype=np.int64)
        self.edge_cells = None if edge_cells is None else np.asarray(edge_cells, dtype=np.int64)
        self.has_edge_global = has_edge_global

    def m(self):
        return self.maps.n_dofs

    def cell_loop(self, src, cells=None):
        """dst = sum over cells of scatter(C^T A_e C gather(src)) -- MatrixFree::cell_loop + do_cell_integral_global"""
        mp = self.maps
        n = mp.fe.n
        dst = np.zeros(mp.n_dofs)
        sel = np.arange(mp.n_cells) if cells is None else cells
        for m in np.unique(mp.mask[sel]):
            c = sel[mp.mask[sel] == m]
            idx, flag = mp.eval_idx[c], mp.eval_flag[c]
            u = np.where(flag, 0.0, src[idx])                      # gather, Dirichlet -> 0
            if m >> 3:
                u = u @ mp.C(m).T                                  # hanging-node interpolation
            r = cell_apply_collocation(mp.fe, u.reshape(-1, n, n, n)).reshape(-1, n ** 3) * mp.h[c][:, None]
            if m >> 3:
                r = r @ mp.C(m)                                    # transposed interpolation
            np.add.at(dst, idx[~flag], r[~flag])
        return dst

    def vmult_interface_down(self, src):
        dst = self.cell_loop(src)
        dst[self.maps.constrained] = src[self.maps.constrained]    # include/operator.h:198-200
        return self.rnd(dst)

    def vmult(self, src):
        src = src.copy()
        saved = src[self.edge].copy()
        src[self.edge] = 0.0                                       # :156-164
        dst = self.cell_loop(src)
        dst[self.maps.constrained] = src[self.maps.constrained]    # :170-172
        dst[self.edge] = saved                                     # :175-182
        return self.rnd(dst)

    Tvmult = vmult

    def vmult_interface_up(self, src):
        if not self.has_edge_global:
            return np.zeros(self.maps.n_dofs)                      # :206-210
        cpy = np.zeros_like(src)
        cpy[self.edge] = src[self.edge]                            # :216-221
        return self.rnd(self.cell_loop(cpy, self.edge_cells))      # :224-225

    def compute_inverse_diagonal(self):
        """MatrixFreeTools::compute_diagonal + threshold inversion (include/operator.h:228-242)"""
        mp = self.maps
        A_ref = element_matrix_reference(mp.fe)
        diag = np.zeros(mp.n_dofs)
        for m in np.unique(mp.mask):
            c = np.nonzero(mp.mask == m)[0]
            Ae = A_ref if (m >> 3) == 0 else mp.C(m).T @ A_ref @ mp.C(m)
            idx, flag = mp.eval_idx[c], mp.eval_flag[c]
            contrib = mp.h[c][:, None] * np.diag(Ae)[None, :]
            np.add.at(diag, idx[~flag], contrib[~flag])
        diag[self.edge] = 0.0
        return np.where(np.abs(diag) > 1.0e-10, 1.0 / np.where(diag == 0, 1.0, diag), 1.0)


class MappedOperator:
    """the same operator on trilinearly mapped cells (general MatrixFree mapping), from the assembled matrix"""

    def __init__(self, maps, vertices):
        self.maps = maps
        self.A0 = assemble_matrix_mapped(maps, vertices, identity_rows=False)
        d = np.zeros(maps.n_dofs)
        d[maps.constrained] = 1.0
        self.A = (self.A0 + sp.diags(d)).tocsr()

    def m(self):
        return self.maps.n_dofs

    def vmult(self, x):
        return self.A @ x

    Tvmult = vmult_interface_down = vmult

    def compute_inverse_diagonal(self):
        diag = self.A0.diagonal()
        return np.where(np.abs(diag) > 1.0e-10, 1.0 / np.where(diag == 0, 1.0, diag), 1.0)


# ----------------------------------------------------------------------------------------- solver control / CG
class ReductionControl:
    def __init__(self, maxiter=10000, abstol=1e-20, reltol=1e-4):
        self.maxiter, self.abstol, self.reltol = maxiter, abstol, reltol
        self.last_step, self.last_value, self.initial_value = 0, 0.0, 0.0

    def check(self, step, value):
        self.last_step, self.last_value = step, value
        if step == 0:
            self.initial_value, self.reduced_tol = value, value * self.reltol
        if value <= self.reduced_tol or value <= self.abstol:
            return "success"
        if step >= self.maxiter or np.isnan(value):
            return "failure"
        return "iterate"


class IterationNumberControl:
    def __init__(self, maxiter, tol=1e-10):
        self.maxiter, self.tol = maxiter, tol
        self.last_step, self.last_value = 0, 0.0

    def check(self, step, value):
        self.last_step, self.last_value = step, value
        if step >= self.maxiter or value <= self.tol:
            return "success"
        return "iterate"


def solver_cg(A, x, b, P, control, lanczos=None, rnd=None):
    """SolverCG::solve. A, P: callables (P None = PreconditionIdentity). x modified in place.
    lanczos: optional (diag, offdiag) lists receiving the projected tridiagonal matrix entries.
    rnd: rounding of the vectors the solver updates (round_f32 for a solver running on float vectors)"""
    if rnd is None:
        rnd = lambda v: v
    if np.any(x != 0):
        g = rnd(A(x) - b)
    else:
        g = -b.copy()
    res = np.sqrt(g @ g)
    if control.check(0, res) != "iterate":
        return 0
    it, gh, eba, prev_alpha = 0, 0.0, 0.0, 0.0
    d = None
    while True:
        it += 1
        if it > 1:
            gh_old = gh
            if P is not None:
                h = P(g)
                gh = g @ h
            else:
                h = g
                gh = res * res
            beta = gh / gh_old
            d = rnd(-h + beta * d)
            if lanczos is not None:
                lanczos[0].append(1.0 / prev_alpha + eba)
                lanczos[1].append(np.sqrt(beta) / prev_alpha)
                eba = beta / prev_alpha
        else:
            if P is not None:
                h = P(g)
                d = -h
                gh = g @ h
            else:
                d = -g
                gh = res * res
        h = A(d)
        alpha = gh / (d @ h)
        prev_alpha = alpha
        x[:] = rnd(x + alpha * d)
        g = rnd(g + alpha * h)
        res = np.sqrt(abs(g @ g))
        if control.check(it, res) != "iterate":
            return it


# ----------------------------------------------------------------------------------------- Chebyshev
class PreconditionChebyshev:
    """PreconditionChebyshev<Operator, Vector, DiagonalMatrix>"""

    def __init__(self, op, degree=3, smoothing_range=20.0, eig_cg_n_iterations=20, owned_begin=0):
        self.op, self.degree, self.smoothing_range, self.eig_its = op, degree, smoothing_range, eig_cg_n_iterations
        self.rnd = getattr(op, "rnd", lambda v: v)
        self.dinv = self.rnd(op.compute_inverse_diagonal())
        self.owned_begin = owned_begin
        self.done = False

    def estimate_eigenvalues(self):
        if self.done:
            return
        n = self.op.m()
        z = ((np.arange(n) + self.owned_begin) % 11).astype(float)
        z -= z.sum() / n
        diag, off = [], []
        x = np.zeros(n)
        solver_cg(self.op.vmult, x, self.rnd(z), lambda g: self.rnd(self.dinv * g), IterationNumberControl(self.eig_its, 1e-10), (diag, off),
                  rnd=self.rnd)
        if not diag:
            self.lmin = self.lmax = 1.0
        else:
            T = np.diag(diag) + np.diag(off[:len(diag) - 1], 1) + np.diag(off[:len(diag) - 1], -1)
            ev = np.linalg.eigvalsh(T)
            self.lmin, self.lmax = ev[0], 1.2 * ev[-1]
        alpha = self.lmax / self.smoothing_range if self.smoothing_range > 1 else min(0.9 * self.lmax, self.lmin)
        self.delta, self.theta = 0.5 * (self.lmax - alpha), 0.5 * (self.lmax + alpha)
        self.done = True

    def vmult(self, rhs):
        self.estimate_eigenvalues()
        x = self.rnd((1.0 / self.theta) * (self.dinv * rhs))
        if self.degree < 2 or abs(self.delta) < 1e-40:
            return x
        x_old = np.zeros_like(x)
        rhok, sigma = self.delta / self.theta, self.theta / self.delta
        for k in range(self.degree - 1):
            t = self.op.vmult(x)
            rhokp = 1.0 / (2.0 * sigma - rhok)
            f1, f2 = rhokp * rhok, 2.0 * rhokp / self.delta
            rhok = rhokp
            if k == 0:
                new = (1.0 + f1) * x + f2 * (self.dinv * (rhs - t))
            else:
                new = (1.0 + f1) * x - f1 * x_old + f2 * (self.dinv * (rhs - t))
            x_old, x = x, self.rnd(new)
        return x

    def step(self, x, rhs):
        self.estimate_eigenvalues()
        t = self.op.vmult(x)
        new = x + (1.0 / self.theta) * (self.dinv * (rhs - t))
        x_old, x = x, self.rnd(new)
        if self.degree < 2 or abs(self.delta) < 1e-40:
            return x
        rhok, sigma = self.delta / self.theta, self.theta / self.delta
        for k in range(self.degree - 1):
            t = self.op.vmult(x)
            rhokp = 1.0 / (2.0 * sigma - rhok)
            f1, f2 = rhokp * rhok, 2.0 * rhokp / self.delta
            rhok = rhokp
            new = (1.0 + f1) * x - f1 * x_old + f2 * (self.dinv * (rhs - t))
            x_old, x = x, self.rnd(new)
        return x


def round_f32(v):
    """round to the nearest float32, keep float64 storage (models vectors stored in MGNumberType=float)"""
    return np.asarray(v, dtype=np.float32).astype(np.float64)


# ----------------------------------------------------------------------------------------- transfers
class MGTwoLevelTransfer:
    """MGTwoLevelTransfer as an explicit sparse prolongation matrix P; restriction = P^T."""

    def __init__(self, fine, coarse, fine_first, n_children, weights=None):
        pf, pc = fine.degree, coarse.degree
        nc, nf = pc + 1, pf + 1
        fec = coarse.fe
        if weights is None:
            weights = transfer_weights(fine, coarse, fine_first, n_children)
        self.weights = weights  # per fine DoF
        rows, cols, vals = [], [], []
        P1 = {8: fec.Ph, 1: (np.eye(nc) if pf == pc else p_embedding(pf, pc))}
        for c in range(coarse.n_cells):
            nch = int(n_children[c])
            if nch == 0:
                continue  # local smoothing: parent without children on the fine level
            P = P1[nch]
            P3 = np.einsum("Kk,Jj,Ii->KJIkji", P, P, P).reshape(P.shape[0] ** 3, nc ** 3)
            if coarse.mask[c] >> 3:
                P3 = P3 @ coarse.C(coarse.mask[c])
            cidx, cflag = coarse.eval_idx[c], coarse.eval_flag[c]
            NF = P.shape[0]
            if nch == 1:
                fidx = fine.own_idx[fine_first[c]]
            else:
                fidx = np.zeros(NF ** 3, dtype=np.int64)
                for K in range(NF):
                    for J in range(NF):
                        for I in range(NF):
                            cx, cy, cz = int(I > pf), int(J > pf), int(K > pf)
                            i, j, k = I - cx * pf, J - cy * pf, K - cz * pf
                            fidx[(K * NF + J) * NF + I] = fine.own_idx[fine_first[c] + cx + 2 * cy + 4 * cz][(k * nf + j) * nf + i]
            blk = weights[fidx][:, None] * P3 * (~cflag)[None, :]
            rows.append(np.repeat(fidx, nc ** 3))
            cols.append(np.tile(cidx, fidx.size))
            vals.append(blk.ravel())
        self.P = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))),
                               shape=(fine.n_dofs, coarse.n_dofs)).tocsr()

    rnd = staticmethod(lambda v: v)   # set to round_f32 for transfers between float level vectors

    def prolongate_and_add(self, dst_fine, src_coarse):
        dst_fine[:] = self.rnd(dst_fine + self.P @ src_coarse)

    def restrict_and_add(self, dst_coarse, src_fine):
        dst_coarse[:] = self.rnd(dst_coarse + self.P.T @ src_fine)


def transfer_weights(fine, coarse, fine_first, n_children):
    """1/valence per fine DoF (number of coarse cells whose fine patch contains it), 0 on constrained fine DoFs"""
    val = np.zeros(fine.n_dofs)
    for c in range(coarse.n_cells):
        cells = np.arange(fine_first[c], fine_first[c] + n_children[c])
        val[np.unique(fine.own_idx[cells].ravel())] += 1
    w = np.where(val > 0, 1.0 / np.maximum(val, 1), 0.0)
    w[fine.constrained] = 0.0
    return w


def transfer_weights_uniform(fine, coarse, fine_first, n_children):
    """transfer_weights for fine meshes WITHOUT hanging nodes, vectorised (the L=7 host baseline of bench.py needs it): a
    coarse cell's patch is the cell itself or its 8 children in Morton order (x fastest); node (i,j,k) of child (a,b,c)
    also belongs to an earlier child of the same patch iff it lies on a low face towards one (a=1,i=0 / b=1,j=0 / c=1,k=0),
    so counting every (patch, DoF) pair once is a static mask per child.  Checked against transfer_weights in
    tests/test_oracle.py."""
    if np.any(fine.mask >> 3):
        raise ValueError("transfer_weights_uniform: fine mesh has hanging nodes")
    n = fine.fe.n
    i, j, k = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    i, j, k = (a.transpose(2, 1, 0).ravel() for a in (i, j, k))   # lexicographic, x fastest
    first = np.ones((8, n ** 3), dtype=bool)
    for ch in range(8):
        a, b, c = ch & 1, (ch >> 1) & 1, (ch >> 2) & 1
        first[ch] = ~(((a == 1) & (i == 0)) | ((b == 1) & (j == 0)) | ((c == 1) & (k == 0)))
    nch = np.asarray(n_children).astype(np.int64)
    if not np.all((nch == 1) | (nch == 8)):
        raise ValueError("transfer_weights_uniform: patches of 1 or 8 cells only")
    parent = np.repeat(np.arange(coarse.n_cells), nch)
    child = np.arange(fine.n_cells) - np.asarray(fine_first).astype(np.int64)[parent]
    val = np.zeros(fine.n_dofs)
    for ch in range(8):
        cells = np.nonzero((child == ch) & (nch[parent] == 8))[0]
        if len(cells):
            val += np.bincount(fine.own_idx[cells][:, first[ch]].ravel(), minlength=fine.n_dofs)
    single = np.nonzero(nch[parent] == 1)[0]
    if len(single):
        val += np.bincount(fine.own_idx[single].ravel(), minlength=fine.n_dofs)
    w = np.where(val > 0, 1.0 / np.maximum(val, 1), 0.0)
    w[fine.constrained] = 0.0
    return w


# ----------------------------------------------------------------------------------------- multigrid
class Multigrid:
    """Multigrid<Vector>::cycle / level_v_step + PreconditionMG::vmult for a global-coarsening hierarchy
    (multigrid_throughput.cc:1091-1133)."""

    def __init__(self, operators, smoothers, transfers, coarse, edge_in=False, copy_indices=None):
        """copy_indices: None (global coarsening: the top level IS the outer vector) or, for local smoothing, one
        (active_idx, level_idx) pair of index arrays per level (MGLevelGlobalTransfer::copy_indices)"""
        self.ops, self.smoothers, self.transfers, self.coarse, self.edge_in = operators, smoothers, transfers, coarse, edge_in
        self.copy_indices = copy_indices
        self.max_level = len(operators) - 1
        self.rnd = getattr(operators[-1], "rnd", lambda v: v)   # round_f32: level vectors in float (the casts of
                                                                 # multigrid_throughput.cc:528-550 at copy_to_mg / copy_from_mg)

    def vmult(self, src):
        """PreconditionMG::vmult: copy_to_mg, V-cycle, copy_from_mg"""
        self.defect = [np.zeros(op.m()) for op in self.ops]
        if self.copy_indices is None:
            self.defect[-1][:] = self.rnd(src)
        else:
            for l, (ai, li) in enumerate(self.copy_indices):
                self.defect[l][li] = self.rnd(src[ai])
        self.solution = [np.zeros(op.m()) for op in self.ops]
        self.t = [np.zeros(op.m()) for op in self.ops]
        self.level_v_step(self.max_level)
        if self.copy_indices is None:
            return self.solution[-1].copy()
        dst = np.zeros_like(src)
        for l, (ai, li) in enumerate(self.copy_indices):
            dst[ai] = self.solution[l][li]
        return dst

    def level_v_step(self, l):
        if l == 0:
            self.solution[0] = self.coarse(self.defect[0])
            return
        self.solution[l] = self.smoothers[l].vmult(self.defect[l])
        self.t[l] = self.rnd(self.defect[l] - self.ops[l].vmult_interface_down(self.solution[l]))
        self.transfers[l].restrict_and_add(self.defect[l - 1], self.t[l])
        self.level_v_step(l - 1)
        self.transfers[l].prolongate_and_add(self.solution[l], self.solution[l - 1])
        if self.edge_in:
            self.t[l] = self.ops[l].vmult_interface_up(self.solution[l])
            self.defect[l] = self.rnd(self.defect[l] - self.t[l])
        self.solution[l] = self.smoothers[l].step(self.solution[l], self.defect[l])


class CoarseGridCG:
    """MGCoarseGridIterativeSolver<SolverCG, Operator, Identity|Chebyshev> (multigrid_throughput.cc:911-944)"""

    def __init__(self, op, precond=None, maxiter=10000, abstol=1e-20, reltol=1e-4):
        self.op, self.precond = op, precond
        self.ctl_args = (maxiter, abstol, reltol)
        self.total_iterations = 0

    def __call__(self, rhs):
        x = np.zeros_like(rhs)
        ctl = ReductionControl(*self.ctl_args)
        P = None if self.precond is None else self.precond.vmult
        self.total_iterations += solver_cg(self.op.vmult, x, rhs, P, ctl, rnd=getattr(self.op, "rnd", None))
        return x


def assemble_rhs_constant(maps):
    """Operator::rhs for f = 1, u_D = 0 (include/operator.h:362-413): b_i = int phi_i, constrained rows dropped"""
    m1 = maps.fe.M.sum(axis=1)
    loc = np.einsum("k,j,i->kji", m1, m1, m1).ravel()
    rhs = np.zeros(maps.n_dofs)
    for m in np.unique(maps.mask):
        c = np.nonzero(maps.mask == m)[0]
        r = np.tile(loc, (len(c), 1)) * (maps.h[c] ** 3)[:, None]
        if m >> 3:
            r = r @ maps.C(m)
        np.add.at(rhs, maps.eval_idx[c][~maps.eval_flag[c]], r[~maps.eval_flag[c]])
    return rhs


GAUSSIAN_CENTER, GAUSSIAN_WIDTH = np.array([-0.5, -0.5, -0.5]), 0.1   # multigrid_throughput.cc:2294-2295


def gaussian_solution(x):
    """GaussianSolution::value (multigrid_throughput.cc:60-89) at points x[..., 3]"""
    r2 = np.sum((np.asarray(x) - GAUSSIAN_CENTER) ** 2, axis=-1)
    return np.exp(-r2 / GAUSSIAN_WIDTH ** 2) / (np.sqrt(2.0 * np.pi) * GAUSSIAN_WIDTH) ** 3


def gaussian_rhs(x):
    """GaussianRightHandSide::value (multigrid_throughput.cc:92-127) = -laplace(GaussianSolution)"""
    r2 = np.sum((np.asarray(x) - GAUSSIAN_CENTER) ** 2, axis=-1)
    w2 = GAUSSIAN_WIDTH ** 2
    return (2 * 3 - 4 * r2 / w2) / w2 * np.exp(-r2 / w2) / (np.sqrt(2.0 * np.pi) * GAUSSIAN_WIDTH) ** 3


def assemble_rhs_gaussian(maps, return_lift=False):
    """Operator::rhs for SimulationType=Gaussian (include/operator.h:362-447, multigrid_throughput.cc:2292-2299):
      b_i = int f phi_i  with QGauss(p+1), constraints resolved in the scatter (:371-413),
      b  -= A_hn x_D     (:415-446): x_D holds GaussianSolution at the support points of the Dirichlet DoFs
                         (interpolate_boundary_values, :2309-2310; constraints.distribute), A_hn is the cell loop of a
                         MatrixFree that knows the hanging-node constraints only -- Dirichlet entries are READ here, and a
                         hanging node on the boundary is read through its parents' Dirichlet values -- and rows of
                         constrained DoFs are dropped (constraints.set_zero)."""
    fe, n = maps.fe, maps.fe.n
    # quadrature points of every cell, lexicographic with x fastest
    q1 = fe.xq
    refq = np.stack(np.meshgrid(q1, q1, q1, indexing="ij"), axis=-1).transpose(2, 1, 0, 3).reshape(n ** 3, 3)
    origin = -1.0 + maps.h[:, None] * maps.cell_xyz
    xq = origin[:, None, :] + maps.h[:, None, None] * refq[None, :, :]
    w3 = np.einsum("k,j,i->kji", fe.wq, fe.wq, fe.wq).ravel()
    fq = gaussian_rhs(xq) * w3[None, :] * (maps.h ** 3)[:, None]                        # f(x_q) JxW
    B = fe.B                                                                            # B[q, i] = phi_i(x_q)
    loc = fq @ np.einsum("ak,bj,ci->abckji", B, B, B).reshape(n ** 3, n ** 3)          # sum_q fq[q] phi_i(x_q), (qz,qy,qx) x (k,j,i)
    # Dirichlet values at the support points of the flagged entries
    xd = np.zeros(maps.n_dofs)
    coords = maps.dof_coordinates()
    bdofs = boundary_dofs(maps)
    xd[bdofs] = gaussian_solution(coords[bdofs])
    A_ref = element_matrix_reference(fe)
    rhs = np.zeros(maps.n_dofs)
    lift_total = np.zeros(maps.n_dofs)
    for m in np.unique(maps.mask):
        c = np.nonzero(maps.mask == m)[0]
        idx, flag = maps.eval_idx[c], maps.eval_flag[c]
        r = loc[c]
        u = xd[idx]                                                                     # Dirichlet entries are read
        if m >> 3:
            u = u @ maps.C(m).T
        lift = (u @ A_ref.T) * maps.h[c][:, None]
        if m >> 3:
            r, lift = r @ maps.C(m), lift @ maps.C(m)
        np.add.at(rhs, idx[~flag], r[~flag])
        np.add.at(lift_total, idx[~flag], lift[~flag])
    rhs -= lift_total
    return (rhs, xd) if return_lift else rhs


def boundary_dofs(maps):
    """DoFs on the boundary that carry a Dirichlet value: the entries flagged in the evaluation table (bit 31)"""
    return np.unique(maps.eval_idx[maps.eval_flag])


def polynomial_coarsening_sequence_bisect(degree):
    d = [degree]
    while d[-1] > 1:
        d.append(max(d[-1] // 2, 1))
    return d[::-1]
