// JSON-driven problem setup and solve: the equivalent of run<3,1,double,LevelNumber>() +
// solve_with_global_coarsening / solve_with_local_smoothing + mg_solve of multigrid_throughput.cc.
#pragma once
#include "partition.h"
#include "solver.h"

#include <map>
#include <string>

namespace mgb
{
  // multigrid_throughput.cc:297-334
  struct MultigridParameters
  {
    struct
    {
      std::string type    = "amg";
      unsigned    maxiter = 10000;
      double      abstol  = 1e-20;
      double      reltol  = 1e-4;
      unsigned    n_cycles = 1;
    } coarse_solver;
    struct
    {
      std::string type                = "chebyshev";
      double      smoothing_range     = 20;
      unsigned    degree              = 5;
      unsigned    eig_cg_n_iterations = 20;
    } smoother;
    struct
    {
      unsigned maxiter = 10000;
      double   abstol  = 1e-20;
      double   reltol  = 1e-4;
    } cg_normal;
    unsigned n_repetitions = 5;
  };

  // multigrid_throughput.cc:1970-2015
  struct RunParameters
  {
    std::string  type            = "PMG";
    std::string  geometry_type   = "quadrant_flexible";
    unsigned     n_ref_global    = 6;
    unsigned     n_ref_local     = 0;
    unsigned     fe_degree_fine  = 4;
    bool         paraview        = false;
    bool         verbose         = true;
    unsigned     p               = 0;
    std::string  policy_name     = "";
    std::string  mg_number_type  = "float";
    std::string  simulation_type = "Constant";
    int          min_level       = -1;
    int          min_n_cells     = -1;
    MultigridParameters mg_data;
    void parse_json(const std::string &text); // unknown keys are skipped, like parse_input_from_json(file, true)
  };

  struct Result
  {
    int       dim = 3, degree = 0, n_ref_global = 0, n_ref_local = 0, n_levels = 0, n_iterations = 0, sub_comm_size = 1;
    long long n_cells = 0, n_cells_hn = 0, n_cells_n = 0, n_dofs = 0;
    double    time = 0, time_cg = 0, throughput = 0;
    double    stage[N_STAGES + 2] = {};
    double    initial_residual = 0, final_residual = 0;
    long long coarse_iterations = 0;
  };

  class ProblemBase
  {
  public:
    virtual ~ProblemBase()                                  = default;
    virtual int            n_levels() const                 = 0;
    virtual const DoFMap  &level_dofmap(int l) const        = 0;
    virtual const Mesh    &level_mesh(int l) const          = 0;
    virtual void          *level_operator(int l)            = 0; // OperatorT<LevelNumber>*
    virtual void          *level_smoother(int l)            = 0; // ChebyshevT<LevelNumber>*
    virtual void          *level_transfer(int l)            = 0; // TwoLevelTransferT<LevelNumber>* (l-1 <-> l)
    virtual OperatorT<double> &fine_operator()              = 0;
    virtual const DoFMap  &fine_dofmap() const              = 0; // active-mesh DoFs of the outer operator
    virtual DVec<double>  &rhs()                            = 0;
    virtual DVec<double>  &solution()                       = 0;
    virtual int            level_number_type() const        = 0;
    virtual void           vcycle(DVec<double> &dst, const DVec<double> &src) = 0;
    virtual Result         solve_once(bool with_timers)     = 0;
    virtual Result         run(unsigned n_repetitions)      = 0;
    // MGTools::print_multigrid_statistics of the partitioned hierarchy (ref :1659, :1866)
    virtual HierarchyStatistics statistics() const          = 0;
    virtual size_t         statistics_levels() const        = 0; // meshes the statistics cover
    RunParameters          prm;
  };

  ProblemBase *create_problem(const std::string &json_text, int n_ranks, int rank);

  // right-hand side of the reference (include/operator.h:362-447) assembled on the host (input generation)
  void assemble_rhs(const Mesh &mesh, const DoFMap &dofs, const std::string &simulation_type, std::vector<double> &rhs_local);
} // namespace mgb
