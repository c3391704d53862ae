// instantiates every adapter class of include/mgb200_dealii.h against the deal.II stubs (syntax and mgb200.h usage only)
#include <mgb200_dealii.h>
using namespace mgb::dealii_adapter;
template class mgb::dealii_adapter::VectorB200<double>;
template class mgb::dealii_adapter::VectorB200<float>;
template class mgb::dealii_adapter::OperatorB200<3, double>;
template class mgb::dealii_adapter::OperatorB200<3, float>;
template class mgb::dealii_adapter::ChebyshevB200<3, float>;
template class mgb::dealii_adapter::MGTwoLevelTransferB200<3, double>;
template class mgb::dealii_adapter::MGTransferB200<3, float>;
void use(const dealii::DoFHandler<3> &dof, const dealii::Mapping<3> &mapping, const dealii::Quadrature<3> &quad,
         const dealii::AffineConstraints<double> &constraints)
{
  OperatorB200<3, double> fine, coarse;
  fine.reinit(mapping, dof, quad, constraints);
  coarse.reinit(mapping, dof, quad, constraints, 2);
  VectorB200<double> x, y;
  fine.initialize_dof_vector(x);
  y.reinit(x);
  fine.vmult(y, x);
  ChebyshevB200<3, double> smoother;
  smoother.initialize(fine);
  smoother.step(y, x);
  MGTwoLevelTransferB200<3, double> t;
  std::vector<std::vector<dealii::DoFHandler<3>::cell_iterator>> patches;
  t.reinit(patches, 4, 4, constraints, fine, coarse);
  t.prolongate_and_add(y, x);
  MGTransferB200<3, double> ls;
  ls.set_copy_indices(0, {{0u, 0u}}, coarse);
  std::vector<VectorB200<double>> levels(1);
  ls.copy_to_mg(levels, x);
  ls.copy_from_mg(y, levels);
  const double r = y * x + y.l2_norm() + y.add_and_dot(1.0, x, y);
  (void)r;
}
