// multigrid_throughput -- same CLI contract as the reference's main() (multigrid_throughput.cc:2398-2470):
// any number of JSON parameter files; one table row per file, columns named as in the reference's
// ConvergenceTable.  Single process = one GPU; multi-GPU runs go through bench.py /
// the Python launcher, which bootstraps NCCL.
#include "mgb200.h"

#include <cstdint>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

int main(int argc, char **argv)
{
  if (argc == 1)
    {
      printf("ERROR: No .json parameter files has been provided!\n");
      return 1;
    }
  if (mgb_init(0) != 0)
    {
      fprintf(stderr, "Exception on processing:\n%s\nAborting!\n", mgb_last_error());
      return 1;
    }
  std::vector<mgb_result>     rows;
  std::vector<mgb_statistics> stats; // appended as columns for inputs with Verbosity=true (ref :1657-1665, 1863-1872)
  const auto                  print_table = [&rows, &stats]() {
    bool any_stats = false;
    for (const mgb_statistics &s : stats)
      any_stats = any_stats || s.verbose;
    printf("%4s %10s %10s %10s %6s %12s %11s %11s %13s %8s %12s %11s %11s %11s %11s %11s %11s %11s %11s %13s %11s %11s %14s", "dim",
           "n_cells", "n_cells_hn", "n_cells_n", "degree", "n_ref_global", "n_ref_local", "n_dofs", "sub_comm_size", "n_levels",
           "n_iterations", "time", "time_cg", "throughput", "time_pre", "time_residuum", "time_res", "time_cs", "time_pro",
           "time_edge_pro", "time_post", "time_to_mg", "time_to_global");
    if (any_stats)
      printf(" %12s %17s %12s %14s %11s", "workload_eff", "workload_path_max", "vertical_eff", "horizontal_eff", "mem_total");
    printf("\n");
    for (size_t k = 0; k < rows.size(); ++k)
      {
        const mgb_result &r = rows[k];
        printf("%4d %10lld %10lld %10lld %6d %12d %11d %11lld %13d %8d %12d %11.4e %11.4e %11.4e %11.4e %13.4e %11.4e %11.4e %11.4e "
               "%13.4e %11.4e %11.4e %14.4e",
               r.dim, (long long)r.n_cells, (long long)r.n_cells_hn, (long long)r.n_cells_n, r.degree, r.n_ref_global, r.n_ref_local,
               (long long)r.n_dofs, r.sub_comm_size, r.n_levels, r.n_iterations, r.time, r.time_cg, r.throughput, r.time_pre,
               r.time_residuum, r.time_res, r.time_cs, r.time_pro, r.time_edge_pro, r.time_post, r.time_to_mg, r.time_to_global);
        if (any_stats)
          {
            const mgb_statistics &s = stats[k];
            printf(" %12.4e %17.4e %12.4e %14.4e %11.4e", s.workload_eff, s.workload_path_max, s.vertical_eff, s.horizontal_eff, s.mem_total);
          }
        printf("\n");
      }
  };
  for (int i = 1; i < argc; ++i)
    {
      printf("%s\n", argv[i]);
      std::ifstream f(argv[i]);
      if (!f)
        {
          fprintf(stderr, "cannot open %s\n", argv[i]);
          return 1;
        }
      std::stringstream ss;
      ss << f.rdbuf();
      mgb_problem *p = nullptr;
      mgb_result   r;
      if (mgb_problem_create(ss.str().c_str(), 1, 0, &p) != 0 || mgb_problem_run(p, 5, &r) != 0)
        {
          fprintf(stderr, "\n\n----------------------------------------------------\nException on processing: \n%s\nAborting!\n"
                          "----------------------------------------------------\n",
                  mgb_last_error());
          return 1;
        }
      mgb_statistics       st = {};
      const int            nm = mgb_problem_n_meshes(p);
      std::vector<int64_t> lv(size_t(nm) * MGB_STATISTICS_LEVEL_VALUES);
      if (mgb_problem_statistics(p, &st, lv.data()) != 0)
        st = mgb_statistics{};
      if (st.verbose) // the per-level tables of MGTools::print_multigrid_statistics (include/mg_tools.h:310-470), sums over ranks
        {
          printf("Workload:\n  efficiency: %g max path: %g\n", st.workload_eff, st.workload_path_max);
          printf("Vertical communication:\n  efficiency: %g\nHorizontal communication:\n  efficiency: %g\n", st.vertical_eff,
                 st.horizontal_eff);
          printf("| %5s | %11s | %11s | %11s | %13s | %14s | %13s | %14s |\n", "level", "n_cells_min", "n_cells_max", "n_cells_sum",
                 "n_local_child", "n_remote_child", "n_local_cells", "n_remote_cells");
          for (int l = 0; l < nm; ++l)
            {
              const int64_t *v = &lv[size_t(l) * MGB_STATISTICS_LEVEL_VALUES];
              printf("| %5d | %11lld | %11lld | %11lld | %13lld | %14lld | %13lld | %14lld |\n", l, (long long)v[0], (long long)v[1],
                     (long long)v[2], (long long)v[3], (long long)v[4], (long long)v[5], (long long)v[6]);
            }
          printf("Memory consumption:\n  total: %g\n\n", st.mem_total);
        }
      mgb_problem_destroy(p);
      rows.push_back(r);
      stats.push_back(st);
      print_table();
    }
  print_table();
  return 0;
}
