#pragma once
#include <deal.II/base/partitioner.h>
#include <deal.II/dofs/dof_handler.h>
#include <deal.II/lac/affine_constraints.h>
#include <cstdint>
namespace dealii
{
  namespace internal
  {
    namespace MatrixFreeFunctions
    {
      using compressed_constraint_kind = std::uint16_t;
      struct DoFInfo
      {
        std::vector<unsigned int>                          dof_indices;
        std::vector<std::pair<unsigned int, unsigned int>> row_starts;
        std::vector<compressed_constraint_kind>            hanging_node_constraint_masks;
      };
    } // namespace MatrixFreeFunctions
  }   // namespace internal
  template <int dim, typename Number> class MatrixFree
  {
  public:
    struct AdditionalData
    {
      UpdateFlags  mapping_update_flags;
      unsigned int mg_level = numbers::invalid_unsigned_int;
    };
    void reinit(const Mapping<dim> &, const DoFHandler<dim> &, const AffineConstraints<Number> &, const Quadrature<dim> &, const AdditionalData &);
    unsigned int                                        n_cell_batches() const;
    unsigned int                                        n_active_entries_per_cell_batch(unsigned int) const;
    typename DoFHandler<dim>::cell_iterator             get_cell_iterator(unsigned int batch, unsigned int lane) const;
    const internal::MatrixFreeFunctions::DoFInfo       &get_dof_info(unsigned int = 0) const;
    std::shared_ptr<const Utilities::MPI::Partitioner>  get_vector_partitioner(unsigned int = 0) const;
  };
} // namespace dealii
/* include/operator.h:539-556 of the reference (the adapter is used next to that header) */
template <int dim, typename Number> dealii::IndexSet get_refinement_edges(const dealii::MatrixFree<dim, Number> &);
