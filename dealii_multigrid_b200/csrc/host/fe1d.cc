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
    std::vector<double> xq, wq;
    gauss_rule(n, xq, wq);
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
