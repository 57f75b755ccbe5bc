// Graph -> block-sparse SoA buffers (the `bsnls_problem` of include/bsnls.h), iterating the very containers
// fuse_graphs::HashGraph::createProblem iterates (hash_graph.cpp:475-522): variables_ first, then
// constraints_, both in unordered_map order. That is what makes the parameter / residual-block ordering handed
// to the GPU identical to what Ceres would be handed for the same in-memory graph.
#pragma once
#include <map>
#include <unordered_map>
#include <vector>

#include "../../include/bsnls.h"
#include "fuse_mirror.h"

namespace bsfuse {

struct FlatProblem {
  std::vector<int32_t> block_size, block_manifold, block_group;
  std::vector<uint8_t> block_constant;
  std::vector<double> block_values;
  std::vector<int64_t> block_offset;
  std::vector<Variable*> block_variable;  // scatter-back targets (Ceres writes Variable::data() in place)
  std::vector<UUID> block_uuid;
  std::vector<double> cameras;            // camera classes; the table only grows (indices cached in the constraints stay valid)
  std::map<std::vector<double>, int32_t> camera_index;
  struct Group {
    int type = 0;
    std::vector<int32_t> blocks, aux, loss_type, res_index;
    std::vector<double> consts, loss_param;
  };
  std::vector<Group> groups;              // indexed by bsnls_factor_type, empty ones skipped in `c`
  std::vector<bsnls_factor_group> cgroups;
  bsnls_problem c{};
  void Finalize();                        // (re)build `c` after the vectors changed
  void ScatterValues(const double* values) const;  // -> Variable::data()
  void GatherValues();                    // Variable::data() -> block_values (structure unchanged since the last Flatten)
};

// Throws std::invalid_argument for constraint types the GPU path does not know (type() string dispatch,
// as the reference itself does at bs_models/src/lib/global_mapping/submap_refinement.cpp:72-73).
// `out` is meant to be reused across calls for one graph: per-constraint lookups (type dispatch, variable objects, camera
// class) are resolved once and cached in the constraint (Constraint::flat_hint), so a re-flatten after a small change of
// a sliding window costs a pass over the containers and no hashing.
void Flatten(const HashGraph& graph, FlatProblem* out);

}  // namespace bsfuse
