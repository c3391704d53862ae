/* mgcpu.c -- TEST INFRASTRUCTURE, not product code: a threaded C restatement of the reference's CPU algorithm for
 * the hot path on meshes without hanging nodes (global-coarsening hierarchies of the hypercube), used only as the timed
 * host baseline of bench.py (cpu_baseline, --impl reference) and checked against oracle/reference.py in tests/.
 * PARITY UNPINNED against deal.II binaries (deal.II cannot be built here); pinned against the numpy
 * restatement, which carries the substitute pins.
 *
 * What it follows:
 *   cell kernel      FEEvaluation::gather_evaluate / submit_gradient / integrate_scatter, include/operator.h:461-472:
 *                    basis change to the Gauss collocation space (3 sweeps), collocation derivative (3), quadrature
 *                    weights x Cartesian Jacobian, and the transposes -- deal.II's 12 sweeps
 *   Operator::vmult  identity rows on constrained DoFs, include/operator.h:152-183
 *   Chebyshev        PreconditionChebyshev as configured in multigrid_throughput.cc:847-883
 *   transfers        MGTwoLevelTransfer::prolongate_and_add / restrict_and_add, h-transfer
 *   V-cycle, CG      Multigrid::level_v_step, PreconditionMG, SolverCG + ReductionControl,
 *                    coarse solver = CG with reltol 1e-4 (multigrid_throughput.cc:888-921)
 * SIMD: the cell kernel runs on 8 cells at once, one cell per lane of a 512-bit vector (GCC vector types; two 256-bit
 * operations on AVX2 hosts) -- deal.II's VectorizedArray<double> over the cells of a MatrixFree batch; lane by lane it is the
 * scalar kernel, bit for bit (tests/test_oracle.py), and mgcpu_set_batched(0) switches back to it.  The 1-D sweeps run in
 * even-odd form (struct EO; deal.II: EvaluatorVariant::evaluate_evenodd), mgcpu_set_even_odd(0) gives the full products.
 * Threads: the operator loop runs over coloured BLOCKS of cells (Morton-aligned cubes of 8^k cells, coloured by the parity of
 * the block coordinates: blocks of one colour share no DoF; a thread walks through whole blocks in order) -- one cache-friendly
 * pass over the vectors per product, the analogue of deal.II's one MPI rank per core working on its contiguous share of the
 * cells; small levels and the transfers use the 8 colours of the cells themselves.  Reductions in a fixed static schedule.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define FLAG 0x80000000u
#define MAXN 8

/* Even-odd decomposition of a 1-D matrix with M[n-1-q][n-1-i] = sign * M[q][i] (shape values at symmetric points: +1,
 * collocation derivative: -1): with e_i = in_i + in_{n-1-i}, o_i = in_i - in_{n-1-i},
 *   out_q = A + B, out_{n-1-q} = sign (A - B),  A = sum_i Me[q][i] e_i (+ Mc[q] in_mid),  B = sum_i Mo[q][i] o_i,
 * which is how deal.II's FEEvaluation applies its 1-D kernels (EvaluatorVariant::evaluate_evenodd): n = 5 needs 13 multiply-adds
 * and 8 additions per line instead of 25 multiply-adds. */
struct EO
{
  double Me[(MAXN / 2) * (MAXN / 2)], Mo[(MAXN / 2) * (MAXN / 2)], Mc[MAXN / 2], Mmid[MAXN / 2], Mmm;
  int    sign;
};

typedef struct
{
  int       p, n, n3;
  long      n_cells, n_dofs, n_con;
  uint32_t *idx; /* n_cells*n3, bit 31 = Dirichlet */
  double   *h;
  uint32_t *con;
  int       n_colors;
  long     *color_begin;
  uint32_t *order;
  /* operator loop: colouring of Morton-aligned blocks of vb_block cells (mgcpu_set_vmult_blocks); a thread takes whole blocks */
  int       vb_colors;
  long     *vb_begin, vb_block;
  uint32_t *vb_order;
  double    B[MAXN * MAXN], Dc[MAXN * MAXN], w[MAXN];
  struct EO eo[4]; /* even-odd forms of B, B^T, Dc, Dc^T */
  int       eo_ok; /* the 1-D matrices have the point symmetry the decomposition needs */
  /* smoother */
  double *dinv, theta, delta, lmax;
  /* transfer from the next coarser level */
  long      n_coarse;
  uint32_t *fine_first;
  uint8_t  *nch;
  double   *wfine;
  uint32_t *own; /* n_cells*n3 own indices (no flags) */
  double    P1[(2 * MAXN) * MAXN];
  /* work */
  double *defect, *sol, *t, *tmp, *xold;
} Level;

typedef struct
{
  int    nl, cheb_degree, eig_its;
  double range;
  Level *L;
  long   coarse_iterations;
} MG;

/* ------------------------------------------------------------------------------------------------ BLAS-1 */
static double dot(const double *a, const double *b, long n)
{
  double s = 0;
#pragma omp parallel for reduction(+ : s) schedule(static)
  for (long i = 0; i < n; ++i)
    s += a[i] * b[i];
  return s;
}
static void axpy(double *y, double a, const double *x, long n)
{
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; ++i)
    y[i] += a * x[i];
}
static void sadd(double *y, double s, double a, const double *x, long n)
{
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; ++i)
    y[i] = s * y[i] + a * x[i];
}
static void equ(double *y, double a, const double *x, long n)
{
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; ++i)
    y[i] = a * x[i];
}
static void zero(double *y, long n)
{
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n; ++i)
    y[i] = 0.0;
}

/* ------------------------------------------------------------------------------------------------ cell kernel */
static int g_blocks  = 1; /* mgcpu_set_blocks: operator loop over coloured blocks of cells (default) or coloured cells */
static int g_batched = 1; /* mgcpu_set_batched: cell kernel on VL cells at once (default) or one cell at a time */

/* one 1-D sweep along direction d of a n x n x n array: out[.., q, ..] = sum_i M[q*n+i] in[.., i, ..] (or M^T) */
static inline __attribute__((always_inline)) void sweep(const double *M, const int transpose, const int n, const int d, const double *in,
                                                        double *out)
{
  const int stride = d == 0 ? 1 : (d == 1 ? n : n * n);
  for (int a = 0; a < n; ++a)
    for (int b = 0; b < n; ++b)
      {
        const int base = d == 0 ? (a * n + b) * n : (d == 1 ? a * n * n + b : a * n + b);
        for (int q = 0; q < n; ++q)
          {
            double s = 0;
            for (int i = 0; i < n; ++i)
              s += (transpose ? M[i * n + q] : M[q * n + i]) * in[base + i * stride];
            out[base + q * stride] = s;
          }
      }
}

static int g_even_odd = 1; /* mgcpu_set_even_odd: 1-D sweeps in even-odd form (default) or as full n x n products */

/* the even-odd sweep for scalars and for vectors of cells: one body, so that both kernels round identically */
#define DEFINE_SWEEP_EO(NAME, T, ZERO)                                                                                       \
  static inline __attribute__((always_inline)) void NAME(const struct EO *E, const int n, const int d, const T *in, T *out)   \
  {                                                                                                                          \
    const int stride = d == 0 ? 1 : (d == 1 ? n : n * n), h = n / 2, odd = n & 1, hs = MAXN / 2;                              \
    for (int a = 0; a < n; ++a)                                                                                              \
      for (int b = 0; b < n; ++b)                                                                                            \
        {                                                                                                                    \
          const int base = d == 0 ? (a * n + b) * n : (d == 1 ? a * n * n + b : a * n + b);                                   \
          T         e[MAXN / 2], o[MAXN / 2], m = ZERO;                                                                      \
          for (int i = 0; i < h; ++i)                                                                                        \
            {                                                                                                                \
              const T x0 = in[base + i * stride], x1 = in[base + (n - 1 - i) * stride];                                       \
              e[i] = x0 + x1;                                                                                                \
              o[i] = x0 - x1;                                                                                                \
            }                                                                                                                \
          if (odd)                                                                                                           \
            m = in[base + h * stride];                                                                                       \
          for (int q = 0; q < h; ++q)                                                                                        \
            {                                                                                                                \
              T A = ZERO, Bv = ZERO;                                                                                         \
              if (odd)                                                                                                       \
                A = E->Mc[q] * m;                                                                                            \
              for (int i = 0; i < h; ++i)                                                                                    \
                {                                                                                                            \
                  A += E->Me[q * hs + i] * e[i];                                                                             \
                  Bv += E->Mo[q * hs + i] * o[i];                                                                            \
                }                                                                                                            \
              out[base + q * stride]           = A + Bv;                                                                     \
              out[base + (n - 1 - q) * stride] = E->sign > 0 ? A - Bv : Bv - A;                                              \
            }                                                                                                                \
          if (odd)                                                                                                           \
            {                                                                                                                \
              T A = E->Mmm * m;                                                                                              \
              for (int i = 0; i < h; ++i)                                                                                    \
                A += E->Mmid[i] * (E->sign > 0 ? e[i] : o[i]);                                                               \
              out[base + h * stride] = A;                                                                                    \
            }                                                                                                                \
        }                                                                                                                    \
  }
DEFINE_SWEEP_EO(sweep_eo, double, 0.0)

/* fills E for M (or M^T); returns 0 if M lacks the symmetry */
static int make_even_odd(struct EO *E, const double *M, const int transpose, const int n, const int sign)
{
  const int h = n / 2, hs = MAXN / 2;
  double    A[MAXN * MAXN], scale = 0;
  for (int q = 0; q < n; ++q)
    for (int i = 0; i < n; ++i)
      {
        A[q * n + i] = transpose ? M[i * n + q] : M[q * n + i];
        scale        = fmax(scale, fabs(A[q * n + i]));
      }
  for (int q = 0; q < n; ++q)
    for (int i = 0; i < n; ++i)
      if (fabs(A[(n - 1 - q) * n + (n - 1 - i)] - sign * A[q * n + i]) > 1e-12 * scale)
        return 0;
  memset(E, 0, sizeof(*E));
  E->sign = sign;
  for (int q = 0; q < h; ++q)
    {
      for (int i = 0; i < h; ++i)
        {
          E->Me[q * hs + i] = 0.5 * (A[q * n + i] + A[q * n + (n - 1 - i)]);
          E->Mo[q * hs + i] = 0.5 * (A[q * n + i] - A[q * n + (n - 1 - i)]);
        }
      if (n & 1)
        E->Mc[q] = A[q * n + h];
    }
  if (n & 1)
    {
      for (int i = 0; i < h; ++i)
        E->Mmid[i] = A[h * n + i];
      E->Mmm = A[h * n + h];
    }
  return 1;
}

/* r = A_ref u on the unit cube (12 sweeps), scaled by h afterwards by the caller */
static inline __attribute__((always_inline)) void cell_apply_n(const Level *L, const double *u, double *r, const int n)
{
  const int n3 = n * n * n;
  double    v[MAXN * MAXN * MAXN], t1[MAXN * MAXN * MAXN], g[MAXN * MAXN * MAXN], acc[MAXN * MAXN * MAXN];
  const int eo = g_even_odd && L->eo_ok;
#define SW(WHICH, M, TR, D, IN, OUT)            \
  do                                            \
    {                                           \
      if (eo)                                   \
        sweep_eo(&L->eo[WHICH], n, D, IN, OUT); \
      else                                      \
        sweep(M, TR, n, D, IN, OUT);            \
    }                                           \
  while (0)
  SW(0, L->B, 0, 0, u, t1);
  SW(0, L->B, 0, 1, t1, v);
  SW(0, L->B, 0, 2, v, t1); /* t1 = values at the quadrature points */
  for (int q = 0; q < n3; ++q)
    acc[q] = 0.0;
  for (int d = 0; d < 3; ++d)
    {
      SW(2, L->Dc, 0, d, t1, g); /* d/dx_d at the quadrature points */
      for (int c = 0; c < n; ++c)
        for (int b = 0; b < n; ++b)
          for (int a = 0; a < n; ++a)
            g[(c * n + b) * n + a] *= L->w[a] * L->w[b] * L->w[c];
      SW(3, L->Dc, 1, d, g, v);
      for (int q = 0; q < n3; ++q)
        acc[q] += v[q];
    }
  SW(1, L->B, 1, 2, acc, v);
  SW(1, L->B, 1, 1, v, t1);
  SW(1, L->B, 1, 0, t1, r);
#undef SW
}

/* the sweeps are compiled once per 1-D size so that the loops unroll and vectorise */
static void cell_apply(const Level *L, const double *u, double *r)
{
  switch (L->n)
    {
      case 2: cell_apply_n(L, u, r, 2); break;
      case 3: cell_apply_n(L, u, r, 3); break;
      case 4: cell_apply_n(L, u, r, 4); break;
      case 5: cell_apply_n(L, u, r, 5); break;
      case 6: cell_apply_n(L, u, r, 6); break;
      case 7: cell_apply_n(L, u, r, 7); break;
      default: cell_apply_n(L, u, r, L->n);
    }
}

/* ---- the same kernel on VL cells at once: lane l of every vector belongs to cell l of the batch (what deal.II's
 * VectorizedArray<double> does over the cells of a MatrixFree batch).  GCC vector types: AVX-512 or 2 x AVX2 per operation,
 * chosen by -march; the arithmetic per lane is the scalar kernel's, operation for operation. */
#define VL 8
typedef double vd __attribute__((vector_size(VL * sizeof(double)), aligned(64)));

static inline __attribute__((always_inline)) void sweep_v(const double *M, const int transpose, const int n, const int d, const vd *in, vd *out)
{
  const int stride = d == 0 ? 1 : (d == 1 ? n : n * n);
  for (int a = 0; a < n; ++a)
    for (int b = 0; b < n; ++b)
      {
        const int base = d == 0 ? (a * n + b) * n : (d == 1 ? a * n * n + b : a * n + b);
        for (int q = 0; q < n; ++q)
          {
            vd s = {0};
            for (int i = 0; i < n; ++i)
              s += (transpose ? M[i * n + q] : M[q * n + i]) * in[base + i * stride];
            out[base + q * stride] = s;
          }
      }
}

DEFINE_SWEEP_EO(sweep_eo_v, vd, ((vd){0}))

static inline __attribute__((always_inline)) void cell_apply_v_n(const Level *L, const vd *u, vd *r, const int n)
{
  const int n3 = n * n * n;
  vd        v[MAXN * MAXN * MAXN], t1[MAXN * MAXN * MAXN], g[MAXN * MAXN * MAXN], acc[MAXN * MAXN * MAXN];
  const int eo = g_even_odd && L->eo_ok;
#define SW(WHICH, M, TR, D, IN, OUT)              \
  do                                              \
    {                                             \
      if (eo)                                     \
        sweep_eo_v(&L->eo[WHICH], n, D, IN, OUT); \
      else                                        \
        sweep_v(M, TR, n, D, IN, OUT);            \
    }                                             \
  while (0)
  SW(0, L->B, 0, 0, u, t1);
  SW(0, L->B, 0, 1, t1, v);
  SW(0, L->B, 0, 2, v, t1);
  for (int q = 0; q < n3; ++q)
    acc[q] = (vd){0};
  for (int d = 0; d < 3; ++d)
    {
      SW(2, L->Dc, 0, d, t1, g);
      for (int c = 0; c < n; ++c)
        for (int b = 0; b < n; ++b)
          for (int a = 0; a < n; ++a)
            g[(c * n + b) * n + a] *= L->w[a] * L->w[b] * L->w[c];
      SW(3, L->Dc, 1, d, g, v);
      for (int q = 0; q < n3; ++q)
        acc[q] += v[q];
    }
  SW(1, L->B, 1, 2, acc, v);
  SW(1, L->B, 1, 1, v, t1);
  SW(1, L->B, 1, 0, t1, r);
#undef SW
}

static void cell_apply_v(const Level *L, const vd *u, vd *r)
{
  switch (L->n)
    {
      case 2: cell_apply_v_n(L, u, r, 2); break;
      case 3: cell_apply_v_n(L, u, r, 3); break;
      case 4: cell_apply_v_n(L, u, r, 4); break;
      case 5: cell_apply_v_n(L, u, r, 5); break;
      case 6: cell_apply_v_n(L, u, r, 6); break;
      case 7: cell_apply_v_n(L, u, r, 7); break;
      default: cell_apply_v_n(L, u, r, L->n);
    }
}

/* VL cells at once: gather, apply, scatter lane by lane (the lanes may share DoFs: the scatter is sequential) */
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wmaybe-uninitialized" /* only the first n^3 entries of u are used, and they are set */
static inline void cell_batch(const Level *L, const uint32_t *cells, const long first, double *dst, const double *src)
{
  const int       n3 = L->n3;
  vd              u[MAXN * MAXN * MAXN], r[MAXN * MAXN * MAXN];
  const uint32_t *id[VL];
  double          h[VL];
  for (int l = 0; l < VL; ++l)
    {
      const long c = cells ? (long)cells[first + l] : first + l;
      id[l]        = L->idx + c * n3;
      h[l]         = L->h[c];
    }
  for (int q = 0; q < n3; ++q)
    for (int l = 0; l < VL; ++l)
      u[q][l] = (id[l][q] & FLAG) ? 0.0 : src[id[l][q]];
  cell_apply_v(L, u, r);
  for (int l = 0; l < VL; ++l)
    for (int q = 0; q < n3; ++q)
      if (!(id[l][q] & FLAG))
        dst[id[l][q]] += h[l] * r[q][l];
}

#pragma GCC diagnostic pop

/* cells of ONE colour (no two of them share a DoF), VL at a time; the tail goes through the scalar kernel */
static void cell_range_batched(const Level *L, const uint32_t *cells, long nc, double *dst, const double *src)
{
  const long n_batch = nc / VL;
#pragma omp parallel for schedule(static)
  for (long bt = 0; bt < n_batch; ++bt)
    cell_batch(L, cells, bt * VL, dst, src);
}

static void cell_range(const Level *L, const uint32_t *cells, long nc, double *dst, const double *src, int atomic)
{
  if (!atomic && g_batched && nc >= VL)
    {
      const long done = (nc / VL) * VL;
      cell_range_batched(L, cells, done, dst, src);
      if (done == nc)
        return;
      cells = cells ? cells + done : NULL;
      if (!cells)
        {
          /* contiguous numbering: run the tail through the scalar loop below with explicit cell numbers */
          uint32_t tail[VL];
          for (long k = done; k < nc; ++k)
            tail[k - done] = (uint32_t)k;
          cell_range(L, tail, nc - done, dst, src, 0);
          return;
        }
      nc -= done;
    }
  const int n3 = L->n3;
#pragma omp parallel for schedule(static)
  for (long k = 0; k < nc; ++k)
    {
      const long      c  = cells ? (long)cells[k] : k;
      const uint32_t *id = L->idx + c * n3;
      double          u[MAXN * MAXN * MAXN], r[MAXN * MAXN * MAXN];
      for (int q = 0; q < n3; ++q)
        u[q] = (id[q] & FLAG) ? 0.0 : src[id[q]];
      cell_apply(L, u, r);
      const double h = L->h[c];
      for (int q = 0; q < n3; ++q)
        if (!(id[q] & FLAG))
          {
            if (atomic)
              {
#pragma omp atomic
                dst[id[q]] += h * r[q];
              }
            else
              dst[id[q]] += h * r[q];
          }
    }
}

/* Blocks of ONE colour: `block` consecutive entries of `cells` are a Morton-aligned cube of cells; two blocks of a colour
 * share no DoF, the cells inside a block do, so a block is the unit of work of a thread and is walked through in order.  This
 * is the cache-friendly schedule (one pass over the vectors per product instead of one per cell colour), what an MPI rank of
 * the reference does on its contiguous share of the cells. */
static void cell_range_blocks(const Level *L, const uint32_t *cells, long nc, long block, double *dst, const double *src)
{
  const long n_blocks = nc / block;
#pragma omp parallel for schedule(static)
  for (long blk = 0; blk < n_blocks; ++blk)
    {
      if (g_batched && block % VL == 0)
        for (long c = blk * block; c < (blk + 1) * block; c += VL)
          cell_batch(L, cells, c, dst, src);
      else
        {
          const int n3 = L->n3;
          double    u[MAXN * MAXN * MAXN], r[MAXN * MAXN * MAXN];
          for (long k = blk * block; k < (blk + 1) * block; ++k)
            {
              const long      c  = (long)cells[k];
              const uint32_t *id = L->idx + c * n3;
              for (int q = 0; q < n3; ++q)
                u[q] = (id[q] & FLAG) ? 0.0 : src[id[q]];
              cell_apply(L, u, r);
              for (int q = 0; q < n3; ++q)
                if (!(id[q] & FLAG))
                  dst[id[q]] += L->h[c] * r[q];
            }
        }
    }
}

/* Operator::vmult (no refinement edges on global-coarsening levels) */
static void vmult(const Level *L, double *dst, const double *src)
{
  zero(dst, L->n_dofs);
  if (g_blocks && L->vb_colors > 0)
    for (int col = 0; col < L->vb_colors; ++col)
      cell_range_blocks(L, L->vb_order + L->vb_begin[col], L->vb_begin[col + 1] - L->vb_begin[col], L->vb_block, dst, src);
  else if (L->n_colors > 0)
    for (int col = 0; col < L->n_colors; ++col)
      cell_range(L, L->order + L->color_begin[col], L->color_begin[col + 1] - L->color_begin[col], dst, src, 0);
  else
    cell_range(L, NULL, L->n_cells, dst, src, 1);
  for (long i = 0; i < L->n_con; ++i)
    dst[L->con[i]] = src[L->con[i]];
}

/* MatrixFreeTools::compute_diagonal + threshold inversion (include/operator.h:228-242) */
static void compute_inverse_diagonal(Level *L)
{
  const int n3 = L->n3;
  double    dloc[MAXN * MAXN * MAXN], u[MAXN * MAXN * MAXN], r[MAXN * MAXN * MAXN];
  for (int q = 0; q < n3; ++q)
    {
      memset(u, 0, sizeof(double) * n3);
      u[q] = 1.0;
      cell_apply(L, u, r);
      dloc[q] = r[q];
    }
  zero(L->dinv, L->n_dofs);
  for (long c = 0; c < L->n_cells; ++c)
    for (int q = 0; q < n3; ++q)
      {
        const uint32_t id = L->idx[c * n3 + q];
        if (!(id & FLAG))
          L->dinv[id] += L->h[c] * dloc[q];
      }
  for (long i = 0; i < L->n_dofs; ++i)
    L->dinv[i] = fabs(L->dinv[i]) > 1e-10 ? 1.0 / L->dinv[i] : 1.0;
}

/* ------------------------------------------------------------------------------------------------ solver control */
typedef struct
{
  int    kind; /* 0 ReductionControl, 1 IterationNumberControl */
  int    maxiter;
  double abstol, reltol, initial, last;
  int    last_step;
} Control;
/* 0 iterate, 1 success, 2 failure */
static int check(Control *c, int step, double v)
{
  c->last_step = step;
  c->last      = v;
  if (c->kind == 0)
    {
      if (step == 0)
        c->initial = v;
      if (v <= c->initial * c->reltol || v <= c->abstol)
        return 1;
      if (step >= c->maxiter || isnan(v))
        return 2;
      return 0;
    }
  return (step >= c->maxiter || v <= c->abstol) ? 1 : 0;
}

typedef void (*apply_fn)(void *ctx, double *dst, const double *src);

/* SolverCG::solve; x = 0 on entry; P == NULL: identity; lanczos arrays may be NULL */
static int cg(apply_fn A, void *actx, apply_fn P, void *pctx, double *x, const double *b, long n, Control *ctl, double *g, double *d,
              double *h, double *ldiag, double *loff, int *nl)
{
  equ(g, -1.0, b, n);
  double res = sqrt(dot(g, g, n));
  if (check(ctl, 0, res) != 0)
    return 0;
  int    it = 0;
  double gh = 0, eba = 0, prev_alpha = 0;
  while (1)
    {
      ++it;
      if (it > 1)
        {
          const double gh_old = gh;
          if (P)
            {
              P(pctx, h, g);
              gh = dot(g, h, n);
            }
          else
            gh = res * res;
          const double beta = gh / gh_old;
          sadd(d, beta, -1.0, P ? h : g, n);
          if (ldiag)
            {
              ldiag[*nl] = 1.0 / prev_alpha + eba;
              loff[*nl]  = sqrt(beta) / prev_alpha;
              ++*nl;
              eba = beta / prev_alpha;
            }
        }
      else
        {
          if (P)
            {
              P(pctx, h, g);
              equ(d, -1.0, h, n);
              gh = dot(g, h, n);
            }
          else
            {
              equ(d, -1.0, g, n);
              gh = res * res;
            }
        }
      A(actx, h, d);
      const double alpha = gh / dot(d, h, n);
      prev_alpha         = alpha;
      axpy(x, alpha, d, n);
      axpy(g, alpha, h, n);
      res = sqrt(fabs(dot(g, g, n)));
      if (check(ctl, it, res) != 0)
        return it;
    }
}

/* largest eigenvalue of a symmetric tridiagonal matrix by bisection on the Sturm count */
static double tridiag_max_eig(const double *dg, const double *off, int m)
{
  double lo = dg[0], hi = dg[0];
  for (int i = 0; i < m; ++i)
    {
      const double r = (i > 0 ? fabs(off[i - 1]) : 0) + (i + 1 < m ? fabs(off[i]) : 0);
      if (dg[i] - r < lo)
        lo = dg[i] - r;
      if (dg[i] + r > hi)
        hi = dg[i] + r;
    }
  for (int it = 0; it < 200; ++it)
    {
      const double mid = 0.5 * (lo + hi);
      int          cnt = 0; /* eigenvalues < mid */
      double       q   = dg[0] - mid;
      if (q < 0)
        ++cnt;
      for (int i = 1; i < m; ++i)
        {
          if (q == 0)
            q = 1e-300;
          q = dg[i] - mid - off[i - 1] * off[i - 1] / q;
          if (q < 0)
            ++cnt;
        }
      if (cnt >= m)
        hi = mid;
      else
        lo = mid;
    }
  return 0.5 * (lo + hi);
}

/* ------------------------------------------------------------------------------------------------ Chebyshev */
static void A_level(void *ctx, double *dst, const double *src) { vmult((const Level *)ctx, dst, src); }
static void P_jacobi(void *ctx, double *dst, const double *src)
{
  const Level *L = (const Level *)ctx;
#pragma omp parallel for schedule(static)
  for (long i = 0; i < L->n_dofs; ++i)
    dst[i] = L->dinv[i] * src[i];
}

static void estimate_eigenvalues(Level *L, int eig_its, double range)
{
  const long n = L->n_dofs;
  double    *z = malloc(sizeof(double) * n), *x = calloc(n, sizeof(double)), *g = malloc(sizeof(double) * n),
         *d = malloc(sizeof(double) * n), *h = malloc(sizeof(double) * n);
  double sum = 0;
  for (long i = 0; i < n; ++i)
    {
      z[i] = (double)(i % 11);
      sum += z[i];
    }
  for (long i = 0; i < n; ++i)
    z[i] -= sum / n;
  double  ldiag[64], loff[64];
  int     nl  = 0;
  Control ctl = {1, eig_its, 1e-10, 0, 0, 0, 0};
  cg(A_level, L, P_jacobi, L, x, z, n, &ctl, g, d, h, ldiag, loff, &nl);
  double lmax = 1.0;
  if (nl > 0)
    lmax = 1.2 * tridiag_max_eig(ldiag, loff, nl);
  const double alpha = lmax / range;
  L->lmax            = lmax;
  L->delta           = 0.5 * (lmax - alpha);
  L->theta           = 0.5 * (lmax + alpha);
  free(z), free(x), free(g), free(d), free(h);
}

/* one Chebyshev update: out = (1+f1) x - f1 xold + f2 dinv (rhs - t) */
static void cheb_update(const Level *L, double f1, double f2, const double *rhs, const double *t, const double *x, const double *xold,
                        double *out, int have_old)
{
#pragma omp parallel for schedule(static)
  for (long i = 0; i < L->n_dofs; ++i)
    out[i] = (1.0 + f1) * x[i] - (have_old ? f1 * xold[i] : 0.0) + f2 * L->dinv[i] * (rhs[i] - t[i]);
}

/* PreconditionChebyshev::vmult (zero_guess) / step: x is overwritten / updated */
static void chebyshev(Level *L, int degree, double *x, const double *rhs, int zero_guess)
{
  const long n = L->n_dofs;
  double    *xo = L->xold, *t = L->tmp;
  if (zero_guess)
    {
#pragma omp parallel for schedule(static)
      for (long i = 0; i < n; ++i)
        {
          x[i]  = L->dinv[i] * rhs[i] / L->theta;
          xo[i] = 0.0;
        }
    }
  else
    {
      vmult(L, t, x);
#pragma omp parallel for schedule(static)
      for (long i = 0; i < n; ++i)
        {
          xo[i] = x[i];
          x[i]  = x[i] + L->dinv[i] * (rhs[i] - t[i]) / L->theta;
        }
    }
  if (degree < 2 || fabs(L->delta) < 1e-40)
    return;
  double rhok = L->delta / L->theta;
  const double sigma = L->theta / L->delta;
  for (int k = 0; k < degree - 1; ++k)
    {
      vmult(L, t, x);
      const double rhokp = 1.0 / (2.0 * sigma - rhok);
      const double f1 = rhokp * rhok, f2 = 2.0 * rhokp / L->delta;
      rhok = rhokp;
      /* new value into xo, then swap roles by copying (keeps the caller's pointer) */
      cheb_update(L, f1, f2, rhs, t, x, xo, xo, !(zero_guess && k == 0));
#pragma omp parallel for schedule(static)
      for (long i = 0; i < n; ++i)
        {
          const double nv = xo[i];
          xo[i]           = x[i];
          x[i]            = nv;
        }
    }
}

/* ------------------------------------------------------------------------------------------------ transfers */
/* fine index / weight of node (I,J,K) of the (2p+1)^3 patch of coarse cell c (children in z-order) */
static inline uint32_t patch_index(const Level *F, long first, int I, int J, int K)
{
  const int p = F->p, n = F->n;
  const int cx = I > p, cy = J > p, cz = K > p;
  const int i = I - cx * p, j = J - cy * p, k = K - cz * p;
  return F->own[(first + cx + 2 * cy + 4 * cz) * F->n3 + (k * n + j) * n + i];
}

/* one sweep with the 1-D embedding P1 (NF x n): in has extent n along d, out NF; extents of the other directions given */
static void embed_sweep(const double *P1, int transpose, int n, int NF, int d, const int ext[3], const double *in, double *out)
{
  /* ext = extents of `in` (x,y,z) for the forward sweep, of `out` for the transposed sweep */
  int ein[3] = {ext[0], ext[1], ext[2]}, eout[3] = {ext[0], ext[1], ext[2]};
  if (!transpose)
    eout[d] = NF;
  else
    {
      ein[d]  = NF;
      eout[d] = n;
    }
  const int sin[3]  = {1, ein[0], ein[0] * ein[1]};
  const int sout[3] = {1, eout[0], eout[0] * eout[1]};
  const int a = (d + 1) % 3, b = (d + 2) % 3;
  for (int ib = 0; ib < ein[b]; ++ib)
    for (int ia = 0; ia < ein[a]; ++ia)
      {
        const int bi = ia * sin[a] + ib * sin[b], bo = ia * sout[a] + ib * sout[b];
        if (!transpose)
          for (int I = 0; I < NF; ++I)
            {
              double s = 0;
              for (int l = 0; l < n; ++l)
                s += P1[I * n + l] * in[bi + l * sin[d]];
              out[bo + I * sout[d]] = s;
            }
        else
          for (int l = 0; l < n; ++l)
            {
              double s = 0;
              for (int I = 0; I < NF; ++I)
                s += P1[I * n + l] * in[bi + I * sin[d]];
              out[bo + l * sout[d]] = s;
            }
      }
}

#define MAXNF (2 * MAXN - 1)
static void transfer_cells(const Level *C, const Level *F, const uint32_t *cells, long nc, double *dst, const double *src, int prolongate,
                           int atomic)
{
  const int n = C->n, n3 = C->n3, NF = 2 * C->p + 1;
#pragma omp parallel for schedule(static)
  for (long kk = 0; kk < nc; ++kk)
    {
      const long      c   = cells ? (long)cells[kk] : kk;
      const int       nch = F->nch[c];
      const uint32_t *cid = C->idx + c * n3;
      const long      ff  = F->fine_first[c];
      double          a0[MAXNF * MAXNF * MAXNF], a1[MAXNF * MAXNF * MAXNF];
      if (nch == 0)
        continue;
      if (prolongate)
        {
          for (int q = 0; q < n3; ++q)
            a0[q] = (cid[q] & FLAG) ? 0.0 : src[cid[q]];
          if (nch == 1)
            {
              for (int q = 0; q < n3; ++q)
                {
                  const uint32_t f = F->own[ff * n3 + q];
                  const double   v = F->wfine[f] * a0[q];
                  if (atomic)
                    {
#pragma omp atomic
                      dst[f] += v;
                    }
                  else
                    dst[f] += v;
                }
              continue;
            }
          int e[3] = {n, n, n};
          embed_sweep(F->P1, 0, n, NF, 0, e, a0, a1);
          e[0] = NF;
          embed_sweep(F->P1, 0, n, NF, 1, e, a1, a0);
          e[1] = NF;
          embed_sweep(F->P1, 0, n, NF, 2, e, a0, a1);
          for (int K = 0; K < NF; ++K)
            for (int J = 0; J < NF; ++J)
              for (int I = 0; I < NF; ++I)
                {
                  const uint32_t f = patch_index(F, ff, I, J, K);
                  const double   v = F->wfine[f] * a1[(K * NF + J) * NF + I];
                  if (atomic)
                    {
#pragma omp atomic
                      dst[f] += v;
                    }
                  else
                    dst[f] += v;
                }
        }
      else
        {
          if (nch == 1)
            for (int q = 0; q < n3; ++q)
              {
                const uint32_t f = F->own[ff * n3 + q];
                a0[q]            = F->wfine[f] * src[f];
              }
          else
            {
              for (int K = 0; K < NF; ++K)
                for (int J = 0; J < NF; ++J)
                  for (int I = 0; I < NF; ++I)
                    {
                      const uint32_t f           = patch_index(F, ff, I, J, K);
                      a1[(K * NF + J) * NF + I] = F->wfine[f] * src[f];
                    }
              int e[3] = {NF, NF, n};
              embed_sweep(F->P1, 1, n, NF, 2, e, a1, a0); /* e = extents of the output */
              e[1] = n;
              embed_sweep(F->P1, 1, n, NF, 1, e, a0, a1);
              e[0] = n;
              embed_sweep(F->P1, 1, n, NF, 0, e, a1, a0);
            }
          for (int q = 0; q < n3; ++q)
            if (!(cid[q] & FLAG))
              {
                if (atomic)
                  {
#pragma omp atomic
                    dst[cid[q]] += a0[q];
                  }
                else
                  dst[cid[q]] += a0[q];
              }
        }
    }
}
static void transfer(const Level *C, const Level *F, double *dst, const double *src, int prolongate)
{
  if (C->n_colors > 0)
    for (int col = 0; col < C->n_colors; ++col)
      transfer_cells(C, F, C->order + C->color_begin[col], C->color_begin[col + 1] - C->color_begin[col], dst, src, prolongate, 0);
  else
    transfer_cells(C, F, NULL, C->n_cells, dst, src, prolongate, 1);
}

/* ------------------------------------------------------------------------------------------------ multigrid */
typedef struct
{
  MG     *mg;
  double *g, *d, *h; /* coarse CG work */
} VCtx;

static void level_v_step(MG *mg, int l, double *cg_g, double *cg_d, double *cg_h)
{
  Level *L = &mg->L[l];
  if (l == 0)
    {
      /* MGCoarseGridIterativeSolver<SolverCG, Operator, PreconditionIdentity>, ReductionControl(10000, 1e-20, 1e-4) */
      Control ctl = {0, 10000, 1e-20, 1e-4, 0, 0, 0};
      zero(L->sol, L->n_dofs);
      mg->coarse_iterations += cg(A_level, L, NULL, NULL, L->sol, L->defect, L->n_dofs, &ctl, cg_g, cg_d, cg_h, NULL, NULL, NULL);
      return;
    }
  Level *C = &mg->L[l - 1];
  chebyshev(L, mg->cheb_degree, L->sol, L->defect, 1); /* pre-smoothing, zero initial guess */
  vmult(L, L->t, L->sol);                              /* vmult_interface_down == vmult without refinement edges */
  sadd(L->t, -1.0, 1.0, L->defect, L->n_dofs);         /* t = defect - A sol */
  transfer(C, L, C->defect, L->t, 0);
  level_v_step(mg, l - 1, cg_g, cg_d, cg_h);
  transfer(C, L, L->sol, C->sol, 1);
  chebyshev(L, mg->cheb_degree, L->sol, L->defect, 0); /* post-smoothing */
}

static void vcycle(void *ctx, double *dst, const double *src)
{
  VCtx *v  = (VCtx *)ctx;
  MG   *mg = v->mg;
  for (int l = 0; l < mg->nl; ++l)
    zero(mg->L[l].defect, mg->L[l].n_dofs);
  Level *T = &mg->L[mg->nl - 1];
  memcpy(T->defect, src, sizeof(double) * T->n_dofs);
  level_v_step(mg, mg->nl - 1, v->g, v->d, v->h);
  memcpy(dst, T->sol, sizeof(double) * T->n_dofs);
}

/* ------------------------------------------------------------------------------------------------ C API (ctypes) */
MG *mgcpu_create(int n_levels, int cheb_degree, double smoothing_range, int eig_its)
{
  MG *mg          = calloc(1, sizeof(MG));
  mg->nl          = n_levels;
  mg->L           = calloc(n_levels, sizeof(Level));
  mg->cheb_degree = cheb_degree;
  mg->range       = smoothing_range;
  mg->eig_its     = eig_its;
  return mg;
}

static void *dup_mem(const void *p, size_t bytes)
{
  void *q = malloc(bytes ? bytes : 1);
  if (bytes)
    memcpy(q, p, bytes);
  return q;
}

/* idx: n_cells*(p+1)^3 lexicographic DoF indices (bit 31 = Dirichlet), h per cell, cells grouped by colour in `order`
 * (n_colors = 0: no colouring, atomics) */
void mgcpu_set_level(MG *mg, int l, int p, long n_cells, long n_dofs, const uint32_t *idx, const double *h, long n_con, const uint32_t *con,
                     const double *B, const double *Dc, const double *w, int n_colors, const long *color_begin, const uint32_t *order)
{
  Level *L   = &mg->L[l];
  L->p       = p;
  L->n       = p + 1;
  L->n3      = L->n * L->n * L->n;
  L->n_cells = n_cells;
  L->n_dofs  = n_dofs;
  L->n_con   = n_con;
  L->idx     = dup_mem(idx, sizeof(uint32_t) * n_cells * L->n3);
  L->h       = dup_mem(h, sizeof(double) * n_cells);
  L->con     = dup_mem(con, sizeof(uint32_t) * n_con);
  memcpy(L->B, B, sizeof(double) * L->n * L->n);
  memcpy(L->Dc, Dc, sizeof(double) * L->n * L->n);
  L->eo_ok = make_even_odd(&L->eo[0], L->B, 0, L->n, +1) && make_even_odd(&L->eo[1], L->B, 1, L->n, +1) &&
             make_even_odd(&L->eo[2], L->Dc, 0, L->n, -1) && make_even_odd(&L->eo[3], L->Dc, 1, L->n, -1);
  memcpy(L->w, w, sizeof(double) * L->n);
  L->n_colors = n_colors;
  if (n_colors > 0)
    {
      L->color_begin = dup_mem(color_begin, sizeof(long) * (n_colors + 1));
      L->order       = dup_mem(order, sizeof(uint32_t) * n_cells);
    }
  L->dinv   = malloc(sizeof(double) * n_dofs);
  L->defect = calloc(n_dofs, sizeof(double));
  L->sol    = calloc(n_dofs, sizeof(double));
  L->t      = calloc(n_dofs, sizeof(double));
  L->tmp    = calloc(n_dofs, sizeof(double));
  L->xold   = calloc(n_dofs, sizeof(double));
}

/* transfer between level l-1 and l (same degree): per coarse cell first fine cell and number of children (1 or 8), weights per
 * fine DoF, own indices of the fine cells, 1-D embedding (2p+1) x (p+1) */
void mgcpu_set_transfer(MG *mg, int l, const uint32_t *fine_first, const uint8_t *nch, const double *wfine, const uint32_t *own, const double *P1)
{
  Level *F = &mg->L[l], *C = &mg->L[l - 1];
  F->n_coarse   = C->n_cells;
  F->fine_first = dup_mem(fine_first, sizeof(uint32_t) * C->n_cells);
  F->nch        = dup_mem(nch, C->n_cells);
  F->wfine      = dup_mem(wfine, sizeof(double) * F->n_dofs);
  F->own        = dup_mem(own, sizeof(uint32_t) * F->n_cells * F->n3);
  memcpy(F->P1, P1, sizeof(double) * (2 * F->p + 1) * F->n);
}

/* inverse diagonals and Chebyshev eigenvalue estimates of all levels */
void mgcpu_setup(MG *mg)
{
  for (int l = 0; l < mg->nl; ++l)
    {
      compute_inverse_diagonal(&mg->L[l]);
      if (l > 0)
        estimate_eigenvalues(&mg->L[l], mg->eig_its, mg->range);
    }
}

void mgcpu_vmult(MG *mg, int l, double *dst, const double *src) { vmult(&mg->L[l], dst, src); }
double mgcpu_lambda_max(MG *mg, int l) { return mg->L[l].lmax; }
void mgcpu_transfer(MG *mg, int l, double *dst, const double *src, int prolongate) { transfer(&mg->L[l - 1], &mg->L[l], dst, src, prolongate); }

void mgcpu_vcycle(MG *mg, double *dst, const double *src)
{
  const long n = mg->L[0].n_dofs;
  VCtx       v = {mg, malloc(sizeof(double) * n), malloc(sizeof(double) * n), malloc(sizeof(double) * n)};
  vcycle(&v, dst, src);
  free(v.g), free(v.d), free(v.h);
}

/* SolverCG + PreconditionMG from a zero initial guess; returns the number of iterations */
int mgcpu_solve(MG *mg, double *x, const double *b, double reltol, double *final_residual, long *coarse_iterations)
{
  Level     *T = &mg->L[mg->nl - 1];
  const long n = T->n_dofs, n0 = mg->L[0].n_dofs;
  VCtx       v = {mg, malloc(sizeof(double) * n0), malloc(sizeof(double) * n0), malloc(sizeof(double) * n0)};
  double    *g = malloc(sizeof(double) * n), *d = malloc(sizeof(double) * n), *h = malloc(sizeof(double) * n);
  Control    ctl = {0, 10000, 1e-20, reltol, 0, 0, 0};
  zero(x, n);
  mg->coarse_iterations = 0;
  const int its         = cg(A_level, T, vcycle, &v, x, b, n, &ctl, g, d, h, NULL, NULL, NULL);
  if (final_residual)
    *final_residual = ctl.last;
  if (coarse_iterations)
    *coarse_iterations = mg->coarse_iterations;
  free(g), free(d), free(h), free(v.g), free(v.d), free(v.h);
  return its;
}

void mgcpu_set_batched(int on) { g_batched = on; }
void mgcpu_set_blocks(int on) { g_blocks = on; }
/* order: the cells grouped by block colour, every block (block_cells consecutive entries) a Morton-aligned cube of cells */
int mgcpu_set_vmult_blocks(void *h, int level, long block_cells, int n_colors, const long *begin, const uint32_t *order)
{
  Level *L = &((MG *)h)->L[level];
  for (int c = 0; c < n_colors; ++c)
    if ((begin[c + 1] - begin[c]) % block_cells != 0)
      return 1;
  free(L->vb_begin), free(L->vb_order);
  L->vb_colors = n_colors;
  L->vb_block  = block_cells;
  L->vb_begin  = dup_mem(begin, sizeof(long) * (n_colors + 1));
  L->vb_order  = dup_mem(order, sizeof(uint32_t) * L->n_cells);
  return 0;
}
void mgcpu_set_even_odd(int on) { g_even_odd = on; }
int  mgcpu_even_odd_available(void *h, int level) { return ((MG *)h)->L[level].eo_ok; }
int  mgcpu_vector_lanes(void) { return VL; }
void mgcpu_set_threads(int n)
{
#ifdef _OPENMP
  omp_set_num_threads(n > 0 ? n : 1);
#else
  (void)n;
#endif
}

int mgcpu_num_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void mgcpu_destroy(MG *mg)
{
  for (int l = 0; l < mg->nl; ++l)
    {
      Level *L = &mg->L[l];
      free(L->idx), free(L->h), free(L->con), free(L->color_begin), free(L->order), free(L->vb_begin), free(L->vb_order), free(L->dinv), free(L->fine_first), free(L->nch);
      free(L->wfine), free(L->own), free(L->defect), free(L->sol), free(L->t), free(L->tmp), free(L->xold);
    }
  free(mg->L);
  free(mg);
}
