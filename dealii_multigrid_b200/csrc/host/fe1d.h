// 1-D finite element data for FE_Q(p) with QGauss(p+1), the pair the reference
// instantiates (multigrid_throughput.cc:1560-1562, 2269-2270).
//
// Everything is defined on the unit interval [0,1]:
//   nodes   : Gauss-Lobatto support points of FE_Q(p)            (n = p+1)
//   M, K    : 1-D mass / stiffness matrices integrated with Gauss(n) -- the
//             tensor factors of the Cartesian cell matrix
//             A_e = h^(d-2) (K x M x M + M x K x M + M x M x K)
//             (same operator as operator.h:461-472 on a cube)
//   S[0|1]  : subface interpolation (parent 1-D basis at the child's nodes)
//             used by the hanging-node gather/scatter
//   Ph      : h-embedding (2p+1) x (p+1), parent basis at the union of the two
//             children's nodes (MGTwoLevelTransfer)
#pragma once
#include <vector>

namespace mgb
{
  std::vector<double> gauss_lobatto_nodes(int n); // on [0,1]
  void                gauss_rule(int n, std::vector<double> &x, std::vector<double> &w); // on [0,1]

  // value of the j-th Lagrange polynomial on `nodes` at x
  double lagrange_value(const std::vector<double> &nodes, int j, double x);
  double lagrange_derivative(const std::vector<double> &nodes, int j, double x);

  struct FE1D
  {
    int                 degree = 0;
    int                 n      = 0;  // degree+1
    std::vector<double> nodes;       // n
    std::vector<double> M, K;        // n*n row-major
    std::vector<double> S[2];        // n*n row-major, S[c][i*n+j] = l_j(c/2 + x_i/2)
    std::vector<double> Ph;          // (2p+1)*n row-major
    // general (non-Cartesian) cells: evaluation at the Gauss points in collocation form
    std::vector<double> xq, wq;      // Gauss(n) points / weights on [0,1]
    std::vector<double> Bq;          // n*n, Bq[q*n+i] = l_i(x_q): nodal values -> values at the Gauss points
    std::vector<double> Dq;          // n*n, Dq[q*n+r] = derivative at x_q of the r-th Lagrange polynomial ON the Gauss points
    // fast diagonalisation of the pencil (K, M):  K = Tm^T diag(lambda) Tm,  M = Tm^T Tm, so that
    //   A_e = h (Tm x Tm x Tm)^T diag(lambda_i + lambda_j + lambda_k) (Tm x Tm x Tm)
    // Rows of Tm are the modes; the first (n+1)/2 rows are symmetric in the node index, the remaining
    // n/2 antisymmetric (the nodes are symmetric about 1/2), which the device kernels use as an
    // even-odd decomposition of every 1-D sweep.
    std::vector<double> Tm;          // n*n row-major, Tm[a*n+l]
    std::vector<double> lambda;      // n
    explicit FE1D(int degree);

  private:
    void compute_fdm();
  };

  // p-embedding: (pf+1) x (pc+1), coarse basis at the fine nodes
  std::vector<double> p_embedding(int degree_fine, int degree_coarse);
} // namespace mgb
