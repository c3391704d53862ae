// Internal to the C-ABI translation units (capi.cc, capi_mg.cc): the opaque handle types of include/mgb200.h and the
// exception -> status code guard.
#pragma once
#include "mgb200.h"

#include "problem.h"

#include <memory>
#include <string>

namespace mgb
{
  std::string &capi_last_error(); // thread-local
  template <class F> int guarded(F &&f)
  {
    try
      {
        f();
        return 0;
      }
    catch (const std::exception &e)
      {
        capi_last_error() = e.what();
        return 1;
      }
    catch (...)
      {
        capi_last_error() = "unknown exception";
        return 2;
      }
  }
  inline void require_same_type(int a, int b)
  {
    if (a != b)
      throw std::runtime_error("number types of the arguments differ");
  }
} // namespace mgb

struct mgb_mesh
{
  mgb::Mesh m;
};
struct mgb_dofmap
{
  const mgb::DoFMap           *d = nullptr;
  std::unique_ptr<mgb::DoFMap> owned;
};
struct mgb_transfer_map
{
  mgb::TransferMap t;
  long long   n_fine_cells = 0;
};
struct mgb_level_edges
{
  mgb::LevelEdges e;
};
struct mgb_copy_indices
{
  mgb::CopyIndices c;
};
struct mgb_vector
{
  int                           type = MGB_DOUBLE;
  mgb::DVec<double>                 *pd   = nullptr;
  mgb::DVec<float>                  *pf   = nullptr;
  std::unique_ptr<mgb::DVec<double>> od;
  std::unique_ptr<mgb::DVec<float>>  of;
};
struct mgb_operator
{
  int                                type = MGB_DOUBLE;
  mgb::OperatorT<double>                 *pd   = nullptr;
  mgb::OperatorT<float>                  *pf   = nullptr;
  std::unique_ptr<mgb::OperatorT<double>> od;
  std::unique_ptr<mgb::OperatorT<float>>  of;
};
struct mgb_chebyshev
{
  int                                 type = MGB_DOUBLE;
  mgb::ChebyshevT<double>                 *pd   = nullptr;
  mgb::ChebyshevT<float>                  *pf   = nullptr;
  std::unique_ptr<mgb::ChebyshevT<double>> od;
  std::unique_ptr<mgb::ChebyshevT<float>>  of;
};
struct mgb_transfer
{
  int                                        type = MGB_DOUBLE;
  mgb::TwoLevelTransferT<double>                 *pd   = nullptr;
  mgb::TwoLevelTransferT<float>                  *pf   = nullptr;
  std::unique_ptr<mgb::TwoLevelTransferT<double>> od;
  std::unique_ptr<mgb::TwoLevelTransferT<float>>  of;
};
