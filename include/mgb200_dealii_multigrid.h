/* Part of mgb200_dealii.h: two-level transfers and the local-smoothing copy_indices. */
#ifndef MGB200_DEALII_MULTIGRID_H
#define MGB200_DEALII_MULTIGRID_H

namespace mgb
{
  namespace dealii_adapter
  {
    /* ternary entity index e = tx + 3 ty + 9 tz (t = 0 low, 1 interior, 2 high) and its first lexicographic node */
    inline unsigned int entity_first_node(const unsigned int e, const unsigned int p)
    {
      const unsigned int t[3] = {e % 3, (e / 3) % 3, e / 9}, n = p + 1;
      unsigned int       ijk[3];
      for (unsigned int d = 0; d < 3; ++d)
        ijk[d] = t[d] == 0 ? 0 : (t[d] == 2 ? p : 1);
      return (ijk[2] * n + ijk[1]) * n + ijk[0];
    }
    inline unsigned int entity_size(const unsigned int e, const unsigned int p)
    {
      unsigned int s = 1;
      for (unsigned int d = 0, q = e; d < 3; ++d, q /= 3)
        if (q % 3 == 1)
          s *= p - 1;
      return s;
    }

    /* own DoFs of one cell as the 27-entry entity table of mgb_transfer_map_view::fine_own: first DoF of every mesh entity
     * (the DoFs of an entity are consecutive), bit 31 on constrained entities */
    template <int dim, typename Number, typename CellIterator>
    void own_entity_table(const CellIterator &cell, const unsigned int degree, const std::vector<unsigned int> &lex_to_hier,
                          const dealii::Utilities::MPI::Partitioner &part, const dealii::AffineConstraints<Number> &constraints,
                          const bool level_dofs, uint32_t *table /* 27 */)
    {
      std::vector<dealii::types::global_dof_index> dofs(lex_to_hier.size());
      if (level_dofs)
        cell->get_mg_dof_indices(dofs);
      else
        cell->get_dof_indices(dofs);
      for (unsigned int e = 0; e < 27; ++e)
        {
          if (entity_size(e, degree) == 0)
            {
              table[e] = 0x80000000u;
              continue;
            }
          const auto g = dofs[lex_to_hier[entity_first_node(e, degree)]];
          table[e]     = uint32_t(part.global_to_local(g)) | (constraints.is_constrained(g) ? 0x80000000u : 0u);
        }
    }

    /* MGTwoLevelTransfer<dim, VectorType>: reinit_geometric_transfer / reinit_polynomial_transfer
     * (multigrid_throughput.cc:1600-1604).  Covers the case where the children of every locally owned coarse cell are
     * locally owned as well (one rank, or the reference's FirstChildPolicy); otherwise deal.II's own transfer has to
     * provide the ghosted fine vector the library expects (mgb_transfer_map_view::distributed). */
    template <int dim, typename Number> class MGTwoLevelTransferB200
    {
    public:
      using VectorType = VectorB200<Number>;
      ~MGTwoLevelTransferB200()
      {
        if (t)
          mgb_transfer_destroy(t);
      }

      /* fine_cells[c] = the fine-mesh cells covering coarse cell c: the cell itself (p-transfer, or a cell that was not
       * refined) or its 2^dim children in deal.II's child order (= Morton order, x fastest) */
      template <typename CellIterator>
      void reinit(const std::vector<std::vector<CellIterator>> &fine_cells, const unsigned int degree_fine, const unsigned int degree_coarse,
                  const dealii::AffineConstraints<Number> &constraints_fine, const OperatorB200<dim, Number> &op_fine,
                  const OperatorB200<dim, Number> &op_coarse, const bool level_dofs = false)
      {
        const auto lex_to_hier = dealii::FETools::lexicographic_to_hierarchic_numbering<dim>(degree_fine);
        const auto part        = op_fine.get_matrix_free().get_vector_partitioner(0);
        std::vector<uint32_t> fine_first, fine_own;
        std::vector<uint8_t>  n_children;
        for (const auto &patch : fine_cells)
          {
            if (patch.size() != 0 && patch.size() != 1 && patch.size() != 8)
              throw std::runtime_error("MGTwoLevelTransferB200: a coarse cell is covered by 0 (local smoothing), 1 or 8 fine cells");
            fine_first.push_back(uint32_t(fine_own.size() / 27));
            n_children.push_back(uint8_t(patch.size()));
            for (const auto &cell : patch)
              {
                fine_own.resize(fine_own.size() + 27);
                own_entity_table<dim, Number>(cell, degree_fine, lex_to_hier, *part, constraints_fine, level_dofs, &fine_own[fine_own.size() - 27]);
              }
          }
        /* weights: 1 / (number of patches touching the entity), 0 on constrained entities (MGTwoLevelTransfer's weights) */
        const size_t          n_slots = fine_own.size() / 27;
        std::vector<uint8_t>  weight_code(fine_own.size(), 0);
        std::vector<uint16_t> valence(size_t(part->locally_owned_size()) + part->n_ghost_indices(), 0);
        std::vector<uint32_t> seen_by(valence.size(), 0xffffffffu);
        for (size_t c = 0; c < fine_first.size(); ++c)
          for (unsigned int k = 0; k < n_children[c]; ++k)
            for (unsigned int e = 0; e < 27; ++e)
              {
                const uint32_t w = fine_own[(size_t(fine_first[c]) + k) * 27 + e];
                if ((w & 0x80000000u) || seen_by[w] == uint32_t(c))
                  continue;
                seen_by[w] = uint32_t(c);
                ++valence[w];
              }
        for (size_t s = 0; s < n_slots * 27; ++s)
          if (!(fine_own[s] & 0x80000000u))
            weight_code[s] = uint8_t(valence[fine_own[s]]);

        mgb_transfer_map_view v{};
        v.degree_fine    = int32_t(degree_fine);
        v.degree_coarse  = int32_t(degree_coarse);
        v.n_coarse_cells = int64_t(fine_first.size());
        v.n_fine_cells   = int64_t(n_slots);
        v.fine_first     = fine_first.data();
        v.n_children     = n_children.data();
        v.weight_code    = weight_code.data();
        v.fine_own       = fine_own.data();
        v.n_fine_slots   = int64_t(n_slots);
        v.distributed    = 0;
        if (t)
          mgb_transfer_destroy(t);
        t = nullptr;
        check(mgb_transfer_create_from_arrays(&v, op_fine.op, op_coarse.op, &t));
      }
      void prolongate_and_add(VectorType &dst, const VectorType &src) const { check(mgb_transfer_prolongate_and_add(t, dst.h, src.h)); }
      void prolongate(VectorType &dst, const VectorType &src) const { check(mgb_transfer_prolongate(t, dst.h, src.h)); }
      void restrict_and_add(VectorType &dst, const VectorType &src) const { check(mgb_transfer_restrict_and_add(t, dst.h, src.h)); }

      mgb_transfer *t = nullptr;
    };

    /* MGTransferMatrixFree<dim, Number> (multigrid_throughput.cc:1800-1817): level transfers as two-level transfers over
     * the refined parents + MGLevelGlobalTransfer::copy_indices between the active-mesh vector and the level vectors.
     * copy_indices[level] comes from deal.II (MGLevelGlobalTransfer fills it in fill_copy_indices); the adapter needs
     * the pairs in LOCAL indices of the two partitioners. */
    template <int dim, typename Number> class MGTransferB200
    {
    public:
      using VectorType = VectorB200<Number>;
      ~MGTransferB200()
      {
        for (mgb_level_copy *c : copies)
          if (c)
            mgb_level_copy_destroy(c);
      }
      std::vector<MGTwoLevelTransferB200<dim, Number>> transfers; /* [l] between l-1 and l */
      std::vector<mgb_level_copy *>                    copies;    /* [l] */

      /* copy_indices[level] of MGLevelGlobalTransfer as (active local index, level local index) pairs */
      void set_copy_indices(const unsigned int level, const std::vector<std::pair<unsigned int, unsigned int>> &pairs,
                            const OperatorB200<dim, Number> &level_operator)
      {
        std::vector<uint32_t> a, l;
        for (const auto &p : pairs)
          {
            a.push_back(uint32_t(p.first));
            l.push_back(uint32_t(p.second));
          }
        if (copies.size() <= level)
          copies.resize(level + 1, nullptr);
        if (copies[level])
          mgb_level_copy_destroy(copies[level]);
        copies[level] = nullptr;
        check(mgb_level_copy_create_from_arrays(uint32_t(a.size()), a.data(), l.data(), level_operator.op, &copies[level]));
      }

      void prolongate(const unsigned int to_level, VectorType &dst, const VectorType &src) const { transfers[to_level].prolongate(dst, src); }
      void prolongate_and_add(const unsigned int to_level, VectorType &dst, const VectorType &src) const
      {
        transfers[to_level].prolongate_and_add(dst, src);
      }
      void restrict_and_add(const unsigned int from_level, VectorType &dst, const VectorType &src) const
      {
        transfers[from_level].restrict_and_add(dst, src);
      }
      /* dst: one vector per level (MGLevelObject); src / the result of copy_from_mg: double vectors of the active mesh */
      template <typename LevelVectors> void copy_to_mg(LevelVectors &dst, const VectorB200<double> &src) const
      {
        for (unsigned int l = 0; l < copies.size(); ++l)
          check(mgb_level_copy_to_mg(copies[l], dst[l].h, src.h));
      }
      template <typename LevelVectors> void copy_from_mg(VectorB200<double> &dst, const LevelVectors &src) const
      {
        dst = 0.0;
        for (unsigned int l = 0; l < copies.size(); ++l)
          check(mgb_level_copy_from_mg(copies[l], dst.h, src[l].h));
      }
    };
  } // namespace dealii_adapter
} // namespace mgb
#endif
