// fuse_constraints::marginalizeVariables for the GPU graph.
//
// Mirrors dependencies/fuse/fuse_constraints/src/marginalize_variables.cpp:
//   :166-258  marginalizeVariables   the constraints of the variables to remove are linearised at the current values,
//                                    the variables are eliminated one at a time, what is left becomes MarginalConstraints
//   :281-399  detail::linearize      -> bsnls_linearize: the residuals and tangent Jacobians (loss correction included)
//                                    come from the device kernels that the solve itself uses (K1-K3), one launch for all
//   :401-535  detail::marginalizeNext  dense [A | b] of one variable's terms, Householder QR, the rows below the variable's
//                                    own columns are the marginal on the others (host: the matrices are a few dozen wide)
//   marginal_constraint.h:85-96      MarginalConstraint: A blocks, b and the linearisation points x_bar
// The elimination order is the order of `marginalized_variables` (the reference's second overload, :176-181; its first one
// asks CCOLAMD for an order, which only changes fill, not the marginal).
#pragma once
#include <string>
#include <vector>

#include "gpu_hash_graph.h"

namespace bsfuse {

// fuse_core::Transaction, the part marginalizeVariables fills (transaction.h)
struct Transaction {
  std::vector<UUID> removed_variables, removed_constraints;
  std::vector<Constraint::SharedPtr> added_constraints;
};

// detail::LinearTerm (marginalize_variables.h:62-68): A[i] is rows x cols[i], row-major
struct LinearTerm {
  std::vector<unsigned int> variables;
  std::vector<std::vector<double>> A; std::vector<int> cols;
  std::vector<double> b;
};
LinearTerm marginalizeNext(const std::vector<LinearTerm>& linear_terms);

// throws std::invalid_argument when a marginal would need more than BSNLS_MARGINAL_MAX_BLOCKS blocks (a marginal with more
// than BSNLS_MARGINAL_MAX_ROWS rows is emitted as several constraints, one per block of rows), std::runtime_error when the
// device linearisation fails
Transaction marginalizeVariables(const std::string& source, const std::vector<UUID>& marginalized_variables, GpuHashGraph& graph);

// fuse_core::Graph::update (fuse_core/src/graph.cpp:69-95): added constraints, then removed constraints, then removed variables
void UpdateGraph(HashGraph& graph, const Transaction& transaction);

namespace constraints {
// fuse_constraints::MarginalConstraint (marginal_constraint.h:85-96): x_bar is read from the variables as they are now.
// A[i]: rows x 3 row-major (tangent columns of variable i). payload = the BSNLS_MARGINAL_STRIDE layout of include/bsnls.h
Constraint::SharedPtr Marginal(const std::string& source, const std::vector<const Variable*>& variables,
                               const std::vector<std::vector<double>>& A, const std::vector<double>& b);
}  // namespace constraints

}  // namespace bsfuse
