#include "fe1d.h"

#include <algorithm>
#include <cmath>

namespace mgb
{
  namespace
  {
    // Legendre P_n(x) and P_n'(x) on [-1,1]
    void legendre(int n, long double x, long double &p, long double &dp)
    {
      long double p0 = 1.0L, p1 = x;
      if (n == 0)
        {
          p  = 1.0L;
          dp = 0.0L;
          return;
        }
      for (int k = 2; k <= n; ++k)
        {
          const long double pk = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k;
          p0                   = p1;
          p1                   = pk;
        }
      p  = p1;
      dp = n * (x * p1 - p0) / (x * x - 1.0L);
    }
  } // namespace

  void gauss_rule(int n, std::vector<double> &x, std::vector<double> &w)
  {
    x.resize(n);
    w.resize(n);
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int i = 0; i < n; ++i)
      {
        long double z = std::cos(pi * (i + 0.75L) / (n + 0.5L));
        long double p, dp;
        for (int it = 0; it < 100; ++it)
          {
            legendre(n, z, p, dp);
            const long double dz = p / dp;
            z -= dz;
            if (std::fabs((double)dz) < 1e-19)
              break;
          }
        legendre(n, z, p, dp);
        // map [-1,1] -> [0,1], ascending order
        x[n - 1 - i] = (double)(0.5L * (z + 1.0L));
        w[n - 1 - i] = (double)(1.0L / ((1.0L - z * z) * dp * dp));
      }
  }

  std::vector<double> gauss_lobatto_nodes(int n)
  {
    std::vector<double> x(n);
    if (n == 1)
      {
        x[0] = 0.5;
        return x;
      }
    x[0]     = 0.0;
    x[n - 1] = 1.0;
    const int         N  = n - 1; // interior nodes are the roots of P_N'
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int i = 1; i < N; ++i)
      {
        long double z = -std::cos(pi * i / N); // Chebyshev-Lobatto start
        for (int it = 0; it < 100; ++it)
          {
            long double p, dp;
            legendre(N, z, p, dp);
            // P_N'' from the Legendre ODE: (1-z^2)P'' = 2 z P' - N(N+1) P
            const long double ddp = (2 * z * dp - N * (N + 1) * p) / (1.0L - z * z);
            const long double dz  = dp / ddp;
            z -= dz;
            if (std::fabs((double)dz) < 1e-19)
              break;
          }
        x[i] = (double)(0.5L * (z + 1.0L));
      }
    // enforce exact symmetry about 1/2
    for (int i = 0; i < n / 2; ++i)
      {
        const double a = 0.5 * (x[i] + (1.0 - x[n - 1 - i]));
        x[i]           = a;
        x[n - 1 - i]   = 1.0 - a;
      }
    if (n % 2 == 1)
      x[n / 2] = 0.5;
    return x;
  }

  double lagrange_value(const std::vector<double> &nodes, int j, double x)
  {
    long double v = 1.0L;
    const int   n = (int)nodes.size();
    for (int m = 0; m < n; ++m)
      if (m != j)
        v *= ((long double)x - nodes[m]) / ((long double)nodes[j] - nodes[m]);
    return (double)v;
  }

  double lagrange_derivative(const std::vector<double> &nodes, int j, double x)
  {
    const int   n = (int)nodes.size();
    long double s = 0.0L;
    for (int k = 0; k < n; ++k)
      if (k != j)
        {
          long double t = 1.0L / ((long double)nodes[j] - nodes[k]);
          for (int m = 0; m < n; ++m)
            if (m != j && m != k)
              t *= ((long double)x - nodes[m]) / ((long double)nodes[j] - nodes[m]);
          s += t;
        }
    return (double)s;
  }

  FE1D::FE1D(int degree_)
    : degree(degree_)
    , n(degree_ + 1)
  {
    nodes = gauss_lobatto_nodes(n);
    gauss_rule(n, xq, wq);
    Bq.assign(n * n, 0.0);
    Dq.assign(n * n, 0.0);
    for (int q = 0; q < n; ++q)
      for (int i = 0; i < n; ++i)
        {
          Bq[q * n + i] = lagrange_value(nodes, i, xq[q]);
          Dq[q * n + i] = lagrange_derivative(xq, i, xq[q]);
        }
    M.assign(n * n, 0.0);
    K.assign(n * n, 0.0);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j)
        {
          long double m = 0, k = 0;
          for (int q = 0; q < n; ++q)
            {
              m += (long double)wq[q] * lagrange_value(nodes, i, xq[q]) *
                   lagrange_value(nodes, j, xq[q]);
              k += (long double)wq[q] * lagrange_derivative(nodes, i, xq[q]) *
                   lagrange_derivative(nodes, j, xq[q]);
            }
          M[i * n + j] = (double)m;
          K[i * n + j] = (double)k;
        }
    for (int c = 0; c < 2; ++c)
      {
        S[c].assign(n * n, 0.0);
        for (int i = 0; i < n; ++i)
          for (int j = 0; j < n; ++j)
            {
              double v = lagrange_value(nodes, j, 0.5 * c + 0.5 * nodes[i]);
              if (std::fabs(v) < 1e-15)
                v = 0.0;
              if (std::fabs(v - 1.0) < 1e-15)
                v = 1.0;
              S[c][i * n + j] = v;
            }
      }
    const int nf = 2 * degree + 1;
    Ph.assign(nf * n, 0.0);
    for (int I = 0; I < nf; ++I)
      for (int j = 0; j < n; ++j)
        Ph[I * n + j] = (I <= degree) ? S[0][I * n + j] : S[1][(I - degree) * n + j];
    compute_fdm();
  }

  namespace
  {
    // generalised symmetric eigenproblem Kb y = lambda Mb y (m x m, Mb SPD): Cholesky + cyclic Jacobi.
    // Returns eigenvectors as columns of Y (Y^T Mb Y = I) and eigenvalues in ascending order.
    void generalized_eig(const std::vector<long double> &Kb, const std::vector<long double> &Mb, int m, std::vector<long double> &Y,
                         std::vector<long double> &lam)
    {
      std::vector<long double> L(m * m, 0.0L);
      for (int i = 0; i < m; ++i)
        for (int j = 0; j <= i; ++j)
          {
            long double s = Mb[i * m + j];
            for (int k = 0; k < j; ++k)
              s -= L[i * m + k] * L[j * m + k];
            L[i * m + j] = (i == j) ? std::sqrt(s) : s / L[j * m + j];
          }
      // C = L^-1 Kb L^-T
      std::vector<long double> X(m * m), C(m * m);
      for (int c = 0; c < m; ++c) // X = L^-1 Kb (column by column)
        for (int i = 0; i < m; ++i)
          {
            long double s = Kb[i * m + c];
            for (int k = 0; k < i; ++k)
              s -= L[i * m + k] * X[k * m + c];
            X[i * m + c] = s / L[i * m + i];
          }
      for (int r = 0; r < m; ++r) // C^T = L^-1 X^T
        for (int i = 0; i < m; ++i)
          {
            long double s = X[r * m + i];
            for (int k = 0; k < i; ++k)
              s -= L[i * m + k] * C[r * m + k];
            C[r * m + i] = s / L[i * m + i];
          }
      std::vector<long double> Q(m * m, 0.0L);
      for (int i = 0; i < m; ++i)
        Q[i * m + i] = 1.0L;
      for (int sweep = 0; sweep < 100; ++sweep)
        {
          long double off = 0;
          for (int i = 0; i < m; ++i)
            for (int j = 0; j < i; ++j)
              off += C[i * m + j] * C[i * m + j];
          if (off < 1e-60L)
            break;
          for (int p = 0; p < m; ++p)
            for (int q = p + 1; q < m; ++q)
              {
                if (std::fabs((double)C[p * m + q]) < 1e-300)
                  continue;
                const long double theta = (C[q * m + q] - C[p * m + p]) / (2.0L * C[p * m + q]);
                const long double t     = (theta >= 0 ? 1.0L : -1.0L) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0L));
                const long double c = 1.0L / std::sqrt(t * t + 1.0L), s = t * c;
                for (int k = 0; k < m; ++k)
                  {
                    const long double a = C[k * m + p], b = C[k * m + q];
                    C[k * m + p] = c * a - s * b;
                    C[k * m + q] = s * a + c * b;
                  }
                for (int k = 0; k < m; ++k)
                  {
                    const long double a = C[p * m + k], b = C[q * m + k];
                    C[p * m + k] = c * a - s * b;
                    C[q * m + k] = s * a + c * b;
                  }
                for (int k = 0; k < m; ++k)
                  {
                    const long double a = Q[k * m + p], b = Q[k * m + q];
                    Q[k * m + p] = c * a - s * b;
                    Q[k * m + q] = s * a + c * b;
                  }
              }
        }
      // Y = L^-T Q, sorted by eigenvalue
      std::vector<int> order(m);
      for (int i = 0; i < m; ++i)
        order[i] = i;
      std::sort(order.begin(), order.end(), [&](int a, int b) { return C[a * m + a] < C[b * m + b]; });
      Y.assign(m * m, 0.0L);
      lam.assign(m, 0.0L);
      for (int cc = 0; cc < m; ++cc)
        {
          const int c = order[cc];
          lam[cc]     = C[c * m + c];
          for (int i = m - 1; i >= 0; --i)
            {
              long double s = Q[i * m + c];
              for (int k = i + 1; k < m; ++k)
                s -= L[k * m + i] * Y[k * m + cc];
              Y[i * m + cc] = s / L[i * m + i];
            }
        }
    }
  } // namespace

  void FE1D::compute_fdm()
  {
    // work in the symmetric / antisymmetric subspaces so that the modes have exact parity
    const int ne = (n + 1) / 2, no = n / 2;
    Tm.assign(n * n, 0.0);
    lambda.assign(n, 0.0);
    const long double r2 = std::sqrt(0.5L);
    for (int parity = 0; parity < 2; ++parity)
      {
        const int m = parity == 0 ? ne : no;
        if (m == 0)
          continue;
        // basis vectors b_c (c < m) of the subspace, as full-length vectors
        std::vector<long double> Bv(m * n, 0.0L);
        for (int c = 0; c < m; ++c)
          {
            if (n % 2 == 1 && c == n / 2)
              Bv[c * n + c] = 1.0L; // middle node (even subspace only)
            else
              {
                Bv[c * n + c]         = r2;
                Bv[c * n + n - 1 - c] = parity == 0 ? r2 : -r2;
              }
          }
        std::vector<long double> Kb(m * m), Mb(m * m);
        for (int a = 0; a < m; ++a)
          for (int b = 0; b < m; ++b)
            {
              long double k = 0, mm = 0;
              for (int i = 0; i < n; ++i)
                for (int j = 0; j < n; ++j)
                  {
                    k += Bv[a * n + i] * (long double)K[i * n + j] * Bv[b * n + j];
                    mm += Bv[a * n + i] * (long double)M[i * n + j] * Bv[b * n + j];
                  }
              Kb[a * m + b] = k;
              Mb[a * m + b] = mm;
            }
        std::vector<long double> Y, lam;
        generalized_eig(Kb, Mb, m, Y, lam);
        for (int c = 0; c < m; ++c)
          {
            const int row = parity == 0 ? c : ne + c;
            lambda[row]   = (double)std::max(lam[c], 0.0L);
            if (parity == 0 && c == 0)
              lambda[row] = 0.0; // constant mode: K 1 = 0 exactly
            // s = sum_b Y[b][c] b_b ;  Tm[row,:] = s^T M
            std::vector<long double> s(n, 0.0L);
            for (int b = 0; b < m; ++b)
              for (int i = 0; i < n; ++i)
                s[i] += Y[b * m + c] * Bv[b * n + i];
            for (int l = 0; l < n; ++l)
              {
                long double v = 0;
                for (int i = 0; i < n; ++i)
                  v += s[i] * (long double)M[i * n + l];
                Tm[row * n + l] = (double)v;
              }
            // exact parity
            for (int l = 0; l < n / 2; ++l)
              {
                const double a = 0.5 * (Tm[row * n + l] + (parity == 0 ? 1.0 : -1.0) * Tm[row * n + n - 1 - l]);
                Tm[row * n + l]         = a;
                Tm[row * n + n - 1 - l] = parity == 0 ? a : -a;
              }
            if (parity == 1 && n % 2 == 1)
              Tm[row * n + n / 2] = 0.0;
          }
      }
  }

  std::vector<double> p_embedding(int degree_fine, int degree_coarse)
  {
    const auto          xf = gauss_lobatto_nodes(degree_fine + 1);
    const auto          xc = gauss_lobatto_nodes(degree_coarse + 1);
    const int           nf = degree_fine + 1, nc = degree_coarse + 1;
    std::vector<double> P(nf * nc);
    for (int i = 0; i < nf; ++i)
      for (int j = 0; j < nc; ++j)
        {
          double v = lagrange_value(xc, j, xf[i]);
          if (std::fabs(v) < 1e-15)
            v = 0.0;
          if (std::fabs(v - 1.0) < 1e-15)
            v = 1.0;
          P[i * nc + j] = v;
        }
    return P;
  }
} // namespace mgb
