/* STUB of the deal.II declarations include/mgb200_dealii.h touches (deal.II is not installed in the build container).
 * Signatures follow deal.II 9.4; only tests/test_adapter_header.py uses these files, to syntax-check the adapter. */
#pragma once
#include <cstddef>
#include <memory>
#include <utility>
#include <vector>
namespace dealii
{
  class Subscriptor
  {
  public:
    virtual ~Subscriptor() = default;
  };
  namespace types
  {
    using global_dof_index = unsigned int;
  }
  namespace numbers
  {
    static const unsigned int invalid_unsigned_int = static_cast<unsigned int>(-1);
  }
  using MPI_Comm = int;
  class IndexSet
  {
  public:
    bool                    is_element(types::global_dof_index) const;
    types::global_dof_index index_within_set(types::global_dof_index) const;
    void                    fill_index_vector(std::vector<types::global_dof_index> &) const;
  };
  template <int dim> class Point
  {
  public:
    double operator[](unsigned int) const;
  };
  template <int dim> class Mapping
  {};
  template <int dim> class Quadrature
  {};
  template <typename Number> struct VectorizedArray
  {
    static constexpr std::size_t size() { return 8; }
  };
  enum UpdateFlags
  {
    update_values            = 1,
    update_gradients         = 2,
    update_quadrature_points = 4
  };
  inline UpdateFlags operator|(UpdateFlags a, UpdateFlags b) { return UpdateFlags(int(a) | int(b)); }
} // namespace dealii
