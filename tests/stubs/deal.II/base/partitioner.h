#pragma once
#include <deal.II/base/subscriptor.h>
namespace dealii
{
  namespace Utilities
  {
    namespace MPI
    {
      template <typename T> T max(const T &, const MPI_Comm &);
      class Partitioner
      {
      public:
        std::pair<types::global_dof_index, types::global_dof_index>    local_range() const;
        unsigned int                                                    locally_owned_size() const;
        unsigned int                                                    n_ghost_indices() const;
        types::global_dof_index                                         local_to_global(unsigned int) const;
        unsigned int                                                    global_to_local(types::global_dof_index) const;
        bool                                                            in_local_range(types::global_dof_index) const;
        bool                                                            is_ghost_entry(types::global_dof_index) const;
        const std::vector<std::pair<unsigned int, unsigned int>>       &ghost_targets() const;
        const std::vector<std::pair<unsigned int, unsigned int>>       &import_targets() const;
        const std::vector<std::pair<unsigned int, unsigned int>>       &import_indices() const;
      };
    } // namespace MPI
  }   // namespace Utilities
} // namespace dealii
