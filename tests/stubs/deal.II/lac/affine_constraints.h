#pragma once
#include <deal.II/base/subscriptor.h>
namespace dealii
{
  template <typename Number> class AffineConstraints
  {
  public:
    void reinit(const IndexSet &);
    void close();
    bool is_constrained(types::global_dof_index) const;
  };
} // namespace dealii
