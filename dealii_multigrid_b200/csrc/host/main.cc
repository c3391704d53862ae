// multigrid_throughput -- same CLI contract as the reference's main() (multigrid_throughput.cc:2398-2470):
// any number of JSON parameter files; one table row per file, columns named as in the reference's
// ConvergenceTable.  Single process = one GPU; multi-GPU runs go through bench.py /
// the Python launcher, which bootstraps NCCL.
#include "mgb200.h"

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
  std::vector<mgb_result> rows;
  const auto              print_table = [&rows]() {
    printf("%4s %10s %10s %10s %6s %12s %11s %11s %13s %8s %12s %11s %11s %11s %11s %11s %11s %11s %11s %13s %11s %11s %14s\n", "dim",
           "n_cells", "n_cells_hn", "n_cells_n", "degree", "n_ref_global", "n_ref_local", "n_dofs", "sub_comm_size", "n_levels",
           "n_iterations", "time", "time_cg", "throughput", "time_pre", "time_residuum", "time_res", "time_cs", "time_pro",
           "time_edge_pro", "time_post", "time_to_mg", "time_to_global");
    for (const mgb_result &r : rows)
      printf("%4d %10lld %10lld %10lld %6d %12d %11d %11lld %13d %8d %12d %11.4e %11.4e %11.4e %11.4e %13.4e %11.4e %11.4e %11.4e %13.4e "
             "%11.4e %11.4e %14.4e\n",
             r.dim, (long long)r.n_cells, (long long)r.n_cells_hn, (long long)r.n_cells_n, r.degree, r.n_ref_global, r.n_ref_local,
             (long long)r.n_dofs, r.sub_comm_size, r.n_levels, r.n_iterations, r.time, r.time_cg, r.throughput, r.time_pre,
             r.time_residuum, r.time_res, r.time_cs, r.time_pro, r.time_edge_pro, r.time_post, r.time_to_mg, r.time_to_global);
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
      mgb_problem_destroy(p);
      rows.push_back(r);
      print_table();
    }
  print_table();
  return 0;
}
