/* mgb200_dealii.h -- header-only adapters that put libmgb200.so (include/mgb200.h) behind the deal.II concepts the
 * reference composes, so that multigrid_throughput.cc runs its hot path on the GPU with these substitutions:
 *
 *   Operator<dim, 1, Number>            (include/operator.h:11-557)                  -> mgb::dealii_adapter::OperatorB200
 *   LinearAlgebra::distributed::Vector  (multigrid_throughput.cc:2031)               -> VectorB200
 *   PreconditionChebyshev<...>          (multigrid_throughput.cc:847-883)            -> ChebyshevB200
 *   MGTwoLevelTransfer<dim, VectorType> (multigrid_throughput.cc:1531-1533, 1600-1604) -> MGTwoLevelTransferB200
 *   MGTransferMatrixFree<dim, Number>   (multigrid_throughput.cc:1800-1802)          -> MGTransferB200 (copy_indices part)
 *
 * Everything deal.II computes on the host stays deal.II's: triangulation, DoF numbering, constraints, MatrixFree::reinit.
 * The adapters only EXTRACT the index maps those objects already hold and hand them to the C ABI as plain arrays.
 *
 * This header cannot be compiled in the build container (deal.II is absent); tests/test_adapter_header.py compiles it
 * against tests/stubs/deal.II (declarations of exactly the deal.II members used here), which keeps its syntax and its use
 * of mgb200.h honest.  Requirements on the DoF numbering: the DoFs of one mesh entity are consecutive (true for
 * DoFHandler::distribute_dofs and every DoFRenumbering that moves whole entities); the cells are axis-aligned cubes of
 * GridGenerator::hyper_cube(-1, 1) (all meshes of include/grid_generator.h), or the cell vertices are passed
 * (OperatorB200::set_general_geometry). */
#ifndef MGB200_DEALII_H
#define MGB200_DEALII_H

#include <deal.II/base/partitioner.h>
#include <deal.II/base/subscriptor.h>
#include <deal.II/dofs/dof_handler.h>
#include <deal.II/fe/fe_tools.h>
#include <deal.II/lac/affine_constraints.h>
#include <deal.II/matrix_free/matrix_free.h>

#include <mgb200.h>

#include <cstdint>
#include <memory>
#include <stdexcept>
#include <type_traits>
#include <vector>

namespace mgb
{
  namespace dealii_adapter
  {
    inline void check(const int rc)
    {
      if (rc != 0)
        throw std::runtime_error(std::string("mgb200: ") + mgb_last_error()); /* deal.II: AssertThrow(..., ExcMessage) */
    }
    template <typename Number> constexpr int number_type() { return std::is_same<Number, double>::value ? MGB_DOUBLE : MGB_FLOAT; }

    /* Utilities::MPI::Partitioner -> the partitioner fields of mgb_operator_desc (local range, ghosts grouped by owner,
     * import indices grouped by the rank that needs them) */
    struct PartitionerArrays
    {
      int64_t               owned_begin = 0;
      uint32_t              n_owned = 0, n_ghost = 0;
      std::vector<int64_t>  ghost_global;
      std::vector<int32_t>  ghost_ranks, import_ranks;
      std::vector<uint32_t> ghost_offsets, import_offsets, import_indices;

      explicit PartitionerArrays(const dealii::Utilities::MPI::Partitioner &p)
      {
        owned_begin = int64_t(p.local_range().first);
        n_owned     = uint32_t(p.locally_owned_size());
        n_ghost     = uint32_t(p.n_ghost_indices());
        for (unsigned int i = 0; i < n_ghost; ++i)
          ghost_global.push_back(int64_t(p.local_to_global(n_owned + i)));
        ghost_offsets.push_back(0);
        for (const auto &t : p.ghost_targets()) /* (rank, count), ascending ranks */
          {
            ghost_ranks.push_back(int32_t(t.first));
            ghost_offsets.push_back(ghost_offsets.back() + uint32_t(t.second));
          }
        import_offsets.push_back(0);
        for (const auto &t : p.import_targets())
          {
            import_ranks.push_back(int32_t(t.first));
            import_offsets.push_back(import_offsets.back() + uint32_t(t.second));
          }
        for (const auto &range : p.import_indices()) /* compressed [begin, end) ranges of owned local indices */
          for (unsigned int i = range.first; i < range.second; ++i)
            import_indices.push_back(uint32_t(i));
      }
    };

    /* LinearAlgebra::distributed::Vector<Number> as far as SolverCG, PreconditionChebyshev, Multigrid and the reference's
     * driver use it */
    template <typename Number> class VectorB200
    {
    public:
      using value_type = Number;
      VectorB200() = default;
      VectorB200(const VectorB200 &o) { *this = o; }
      ~VectorB200()
      {
        if (h)
          mgb_vector_destroy(h);
      }
      VectorB200 &operator=(const VectorB200 &o)
      {
        reinit(o, true);
        check(mgb_vector_equ(h, 1.0, o.h));
        return *this;
      }
      void reinit(const VectorB200 &o, const bool /*omit_zeroing*/ = false)
      {
        if (h)
          mgb_vector_destroy(h);
        h       = nullptr;
        n_owned = o.n_owned;
        n_ghost = o.n_ghost;
        check(mgb_vector_create(n_owned, n_ghost, number_type<Number>(), &h));
      }
      /* Operator::initialize_dof_vector hands the handle over */
      void adopt(mgb_vector *v, const uint32_t owned, const uint32_t ghost)
      {
        if (h)
          mgb_vector_destroy(h);
        h       = v;
        n_owned = owned;
        n_ghost = ghost;
      }
      VectorB200 &operator=(const Number a)
      {
        check(mgb_vector_set(h, double(a)));
        return *this;
      }
      void   add(const Number a, const VectorB200 &x) { check(mgb_vector_add(h, double(a), x.h)); }
      void   sadd(const Number s, const Number a, const VectorB200 &x) { check(mgb_vector_sadd(h, double(s), double(a), x.h)); }
      void   equ(const Number a, const VectorB200 &x) { check(mgb_vector_equ(h, double(a), x.h)); }
      Number operator*(const VectorB200 &y) const
      {
        double r = 0;
        check(mgb_vector_dot(h, y.h, &r));
        return Number(r);
      }
      Number l2_norm() const
      {
        double r = 0;
        check(mgb_vector_l2_norm(h, &r));
        return Number(r);
      }
      Number add_and_dot(const Number a, const VectorB200 &v, const VectorB200 &w) const
      {
        /* SolverCG calls g.add_and_dot(alpha, h, g): this += a v; return this . w  with w == *this */
        if (&w != this)
          throw std::runtime_error("VectorB200::add_and_dot: only the form SolverCG uses (w == *this) is provided");
        double r = 0;
        check(mgb_vector_add_and_dot(h, double(a), v.h, &r));
        return Number(r);
      }
      void update_ghost_values() const { check(mgb_vector_update_ghost_values(h)); }
      void compress(int /*VectorOperation::add*/) { check(mgb_vector_compress_add(h)); }
      void zero_out_ghost_values() const { check(mgb_vector_zero_out_ghost_values(h)); }
      unsigned int locally_owned_size() const { return n_owned; }
      /* host <-> device copies for set-up and output (e.g. the right-hand side assembled by Operator::rhs on the host) */
      void from_host(const Number *values) { check(mgb_vector_copy_from_host(h, values, n_owned)); }
      void to_host(Number *values) const { check(mgb_vector_copy_to_host(h, values, n_owned)); }

      mgb_vector *h       = nullptr;
      uint32_t    n_owned = 0, n_ghost = 0;
    };
  } // namespace dealii_adapter
} // namespace mgb

#include "mgb200_dealii_operator.h"
#include "mgb200_dealii_multigrid.h"

#endif
