#pragma once
#include <deal.II/base/subscriptor.h>
namespace dealii
{
  namespace FETools
  {
    template <int dim> std::vector<unsigned int> lexicographic_to_hierarchic_numbering(unsigned int degree);
  }
} // namespace dealii
