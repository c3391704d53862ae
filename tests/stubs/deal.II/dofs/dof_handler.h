#pragma once
#include <deal.II/base/subscriptor.h>
namespace dealii
{
  template <int dim> class FiniteElement
  {
  public:
    unsigned int degree;
  };
  template <int dim> class DoFCellAccessor
  {
  public:
    int        level() const;
    Point<dim> vertex(unsigned int) const;
    void       get_dof_indices(std::vector<types::global_dof_index> &) const;
    void       get_mg_dof_indices(std::vector<types::global_dof_index> &) const;
  };
  template <int dim> class DoFHandler
  {
  public:
    using cell_iterator = const DoFCellAccessor<dim> *;
    const FiniteElement<dim> &get_fe() const;
    types::global_dof_index   n_dofs() const;
    types::global_dof_index   n_dofs(unsigned int level) const;
    const IndexSet           &locally_owned_mg_dofs(unsigned int level) const;
    MPI_Comm                  get_communicator() const;
  };
  template <typename Number> class AffineConstraints;
  namespace DoFTools
  {
    template <int dim> IndexSet extract_locally_relevant_dofs(const DoFHandler<dim> &);
    template <int dim, typename Number> void make_hanging_node_constraints(const DoFHandler<dim> &, AffineConstraints<Number> &);
  } // namespace DoFTools
} // namespace dealii
