/* Part of mgb200_dealii.h: Operator and Chebyshev smoother adapters. */
#ifndef MGB200_DEALII_OPERATOR_H
#define MGB200_DEALII_OPERATOR_H

namespace mgb
{
  namespace dealii_adapter
  {
    /* Operator<dim, 1, Number> of include/operator.h with the same reinit signature (:24-29).  The index maps are read
     * from a MatrixFree that knows the hanging-node constraints only -- the object the reference itself builds in
     * Operator::rhs (:415-433): its DoFInfo::dof_indices hold, per cell and in lexicographic order, exactly the
     * (p+1)^dim entries FEEvaluation reads, with the parent's DoFs in the slots of a hanging face or edge and the mask
     * in DoFInfo::hanging_node_constraint_masks.  Dirichlet DoFs are flagged in bit 31 (read as 0, never written). */
    template <int dim, typename Number> class OperatorB200 : public dealii::Subscriptor
    {
    public:
      using value_type = Number;
      using VectorType = VectorB200<Number>;
      using size_type  = dealii::types::global_dof_index;
      static_assert(dim == 3, "libmgb200 implements the 3-D path of the reference");

      ~OperatorB200()
      {
        if (op)
          mgb_operator_destroy(op);
      }

      void reinit(const dealii::Mapping<dim> &mapping, const dealii::DoFHandler<dim> &dof_handler, const dealii::Quadrature<dim> &quad,
                  const dealii::AffineConstraints<Number> &constraints, const unsigned int mg_level = dealii::numbers::invalid_unsigned_int)
      {
        /* (1) hanging-node-only constraints and the MatrixFree on them, as include/operator.h:415-433 */
        dealii::AffineConstraints<Number> constraints_without_dbc;
        if (mg_level == dealii::numbers::invalid_unsigned_int)
          {
            constraints_without_dbc.reinit(dealii::DoFTools::extract_locally_relevant_dofs(dof_handler));
            dealii::DoFTools::make_hanging_node_constraints(dof_handler, constraints_without_dbc);
          }
        constraints_without_dbc.close();
        typename dealii::MatrixFree<dim, Number>::AdditionalData data;
        data.mapping_update_flags = dealii::update_quadrature_points | dealii::update_gradients | dealii::update_values;
        data.mg_level             = mg_level;
        matrix_free.reinit(mapping, dof_handler, constraints_without_dbc, quad, data);

        const unsigned int degree = dof_handler.get_fe().degree, n1 = degree + 1, n_dofs_cell = n1 * n1 * n1;
        const auto        &di     = matrix_free.get_dof_info(0);
        const auto         part   = matrix_free.get_vector_partitioner(0);
        constexpr unsigned int n_lanes = dealii::VectorizedArray<Number>::size();

        /* (2) per physical cell, in MatrixFree's order: lexicographic index map, mask, refinement level, vertices */
        std::vector<uint32_t> dof_indices;
        std::vector<uint16_t> masks;
        std::vector<uint8_t>  levels;
        cell_vertices.clear();
        for (unsigned int b = 0; b < matrix_free.n_cell_batches(); ++b)
          for (unsigned int v = 0; v < matrix_free.n_active_entries_per_cell_batch(b); ++v)
            {
              const unsigned int  row = b * n_lanes + v;
              const unsigned int *idx = di.dof_indices.data() + di.row_starts[row].first;
              if (di.row_starts[row + 1].first - di.row_starts[row].first != n_dofs_cell)
                throw std::runtime_error("OperatorB200: MatrixFree stores a compressed index list for a cell; only hanging-node "
                                         "constraints may be given to this MatrixFree");
              for (unsigned int i = 0; i < n_dofs_cell; ++i)
                {
                  const size_type g       = part->local_to_global(idx[i]);
                  const bool      dirichlet = constraints.is_constrained(g) && !constraints_without_dbc.is_constrained(g);
                  dof_indices.push_back(uint32_t(idx[i]) | (dirichlet ? 0x80000000u : 0u));
                }
              masks.push_back(di.hanging_node_constraint_masks.empty() ? uint16_t(0) : uint16_t(di.hanging_node_constraint_masks[row]));
              const auto cell = matrix_free.get_cell_iterator(b, v);
              levels.push_back(uint8_t(cell->level()));
              for (unsigned int vertex = 0; vertex < 8; ++vertex)
                for (unsigned int d = 0; d < 3; ++d)
                  cell_vertices.push_back(cell->vertex(vertex)[d]); /* deal.II vertex order = x + 2y + 4z */
            }

        /* (3) constrained DoFs: identity rows of vmult (include/operator.h:44-47, 170-172): Dirichlet + hanging */
        constrained_indices.clear();
        for (unsigned int i = 0; i < part->locally_owned_size(); ++i)
          if (constraints.is_constrained(part->local_to_global(i)))
            constrained_indices.push_back(i);

        /* (4) refinement-edge data of a level operator (include/operator.h:49-99) */
        edge_constrained_indices.clear();
        edge_cells.clear();
        bool has_edge_global = false;
        if (mg_level != dealii::numbers::invalid_unsigned_int)
          {
            const dealii::IndexSet  edges = get_refinement_edges(matrix_free); /* include/operator.h:539-556 */
            const dealii::IndexSet &owned = dof_handler.locally_owned_mg_dofs(mg_level);
            std::vector<size_type>  interface_indices;
            edges.fill_index_vector(interface_indices);
            std::vector<char> is_edge(part->locally_owned_size() + part->n_ghost_indices(), 0);
            for (const size_type g : interface_indices)
              {
                if (owned.is_element(g))
                  edge_constrained_indices.push_back(uint32_t(owned.index_within_set(g)));
                if (part->in_local_range(g) || part->is_ghost_entry(g))
                  is_edge[part->global_to_local(g)] = 1;
              }
            has_edge_global = dealii::Utilities::MPI::max(edge_constrained_indices.size(), dof_handler.get_communicator()) > 0;
            /* cells touching an edge DoF (the reference finds them by reading a marker vector, :71-99) */
            const unsigned int n_cells = unsigned(levels.size());
            for (unsigned int c = 0; c < n_cells; ++c)
              for (unsigned int i = 0; i < n_dofs_cell; ++i)
                if (is_edge[dof_indices[size_t(c) * n_dofs_cell + i] & 0x7fffffffu])
                  {
                    edge_cells.push_back(c);
                    break;
                  }
          }

        /* (5) hand the arrays over */
        const PartitionerArrays pa(*part);
        mgb_operator_desc       d{};
        d.degree                      = int32_t(degree);
        d.number_type                 = number_type<Number>();
        d.n_cells                     = int64_t(levels.size());
        d.dof_indices                 = dof_indices.data();
        d.mask                        = masks.data();
        d.cell_level                  = levels.data();
        d.root_length                 = 2.0; /* GridGenerator::hyper_cube(-1, 1), include/grid_generator.h */
        d.constrained                 = constrained_indices.data();
        d.n_constrained               = uint32_t(constrained_indices.size());
        d.edge_constrained            = edge_constrained_indices.data();
        d.n_edge_constrained          = uint32_t(edge_constrained_indices.size());
        d.has_edge_constrained_global = has_edge_global ? 1 : 0;
        d.edge_cells                  = edge_cells.data();
        d.n_edge_cells                = uint32_t(edge_cells.size());
        d.n_dofs_global               = int64_t(mg_level == dealii::numbers::invalid_unsigned_int ? dof_handler.n_dofs() : dof_handler.n_dofs(mg_level));
        d.owned_begin                 = pa.owned_begin;
        d.n_owned                     = pa.n_owned;
        d.n_ghost                     = pa.n_ghost;
        d.ghost_global                = pa.ghost_global.data();
        d.ghost_ranks                 = pa.ghost_ranks.data();
        d.ghost_offsets               = pa.ghost_offsets.data();
        d.n_ghost_ranks               = uint32_t(pa.ghost_ranks.size());
        d.import_ranks                = pa.import_ranks.data();
        d.import_offsets              = pa.import_offsets.data();
        d.n_import_ranks              = uint32_t(pa.import_ranks.size());
        d.import_indices              = pa.import_indices.data();
        if (op)
          mgb_operator_destroy(op);
        op = nullptr;
        check(mgb_operator_create(&d, &op));
        n_owned = pa.n_owned;
        n_ghost = pa.n_ghost;
      }

      /* meshes whose cells are not axis-aligned cubes: the metric is evaluated from the cell vertices (MappingQ1) */
      void set_general_geometry() { check(mgb_operator_set_cell_vertices(op, cell_vertices.data(), int64_t(cell_vertices.size() / 24))); }

      size_type m() const { return size_type(mgb_operator_m(op)); }
      Number    el(unsigned int, unsigned int) const { throw std::runtime_error("OperatorB200::el is not implemented (the reference's neither)"); }
      void      initialize_dof_vector(VectorType &vec) const
      {
        mgb_vector *v = nullptr;
        check(mgb_operator_initialize_dof_vector(op, &v));
        vec.adopt(v, n_owned, n_ghost);
      }
      void vmult(VectorType &dst, const VectorType &src) const { check(mgb_operator_vmult(op, dst.h, src.h)); }
      void Tvmult(VectorType &dst, const VectorType &src) const { check(mgb_operator_tvmult(op, dst.h, src.h)); }
      void vmult_interface_down(VectorType &dst, const VectorType &src) const { check(mgb_operator_vmult_interface_down(op, dst.h, src.h)); }
      void vmult_interface_up(VectorType &dst, const VectorType &src) const { check(mgb_operator_vmult_interface_up(op, dst.h, src.h)); }
      void compute_inverse_diagonal(VectorType &diagonal) const
      {
        initialize_dof_vector(diagonal);
        check(mgb_operator_compute_inverse_diagonal(op, diagonal.h));
      }
      const dealii::MatrixFree<dim, Number> &get_matrix_free() const { return matrix_free; }

      mgb_operator *op = nullptr;

    private:
      dealii::MatrixFree<dim, Number> matrix_free;
      std::vector<uint32_t>           constrained_indices, edge_constrained_indices, edge_cells;
      std::vector<double>             cell_vertices;
      uint32_t                        n_owned = 0, n_ghost = 0;
    };

    /* PreconditionChebyshev<Operator, VectorType, DiagonalMatrix<VectorType>> as mg::SmootherRelaxation uses it
     * (multigrid_throughput.cc:847-883): the inverse diagonal is computed by the library */
    template <int dim, typename Number> class ChebyshevB200
    {
    public:
      struct AdditionalData
      {
        unsigned int degree = 3, eig_cg_n_iterations = 20;
        double       smoothing_range = 20.0;
      };
      using VectorType = VectorB200<Number>;
      ~ChebyshevB200()
      {
        if (c)
          mgb_chebyshev_destroy(c);
      }
      void initialize(const OperatorB200<dim, Number> &A, const AdditionalData &ad = AdditionalData())
      {
        if (c)
          mgb_chebyshev_destroy(c);
        c = nullptr;
        check(mgb_chebyshev_create(A.op, int(ad.degree), ad.smoothing_range, ad.eig_cg_n_iterations, &c));
      }
      void vmult(VectorType &dst, const VectorType &src) const { check(mgb_chebyshev_vmult(c, dst.h, src.h)); } /* pre-smoothing  */
      void step(VectorType &dst, const VectorType &src) const { check(mgb_chebyshev_step(c, dst.h, src.h)); }   /* post-smoothing */
      void Tvmult(VectorType &dst, const VectorType &src) const { vmult(dst, src); }
      void Tstep(VectorType &dst, const VectorType &src) const { step(dst, src); }
      void clear() {}

      mgb_chebyshev *c = nullptr;
    };
  } // namespace dealii_adapter
} // namespace mgb
#endif
