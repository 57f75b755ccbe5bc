#include "marginalize.h"

#include <algorithm>
#include <cmath>
#include <stdexcept>
#include <unordered_map>
#include <unordered_set>

namespace bsfuse {

namespace constraints {
Constraint::SharedPtr Marginal(const std::string& source, const std::vector<const Variable*>& variables,
                               const std::vector<std::vector<double>>& A, const std::vector<double>& b) {
  const size_t n = variables.size(), rows = b.size();
  if (n < 1 || n > BSNLS_MARGINAL_MAX_BLOCKS) throw std::invalid_argument("MarginalConstraint: between 1 and 10 variables are supported by the GPU path");
  if (rows > BSNLS_MARGINAL_MAX_ROWS) throw std::invalid_argument("MarginalConstraint: at most 15 residual rows are supported by the GPU path");
  if (A.size() != n) throw std::invalid_argument("MarginalConstraint: one A block per variable");   // marginal_constraint.h:146-152
  std::vector<double> pl(BSNLS_MARGINAL_STRIDE, 0.0);
  std::vector<UUID> uuids;
  pl[0] = static_cast<double>(n);
  size_t off = 0;
  for (size_t i = 0; i < n; ++i) {
    const Variable& v = *variables[i];
    if (v.size() != 3 && v.size() != 4) throw std::invalid_argument("MarginalConstraint: variable type not supported by the GPU path: " + v.type());
    if (A[i].size() != rows * 3) throw std::invalid_argument("MarginalConstraint: A block must be rows x local size");
    if (off + v.size() > 32) throw std::invalid_argument("MarginalConstraint: more than 32 parameters");
    pl[1 + i] = static_cast<double>(v.size());
    for (size_t k = 0; k < v.size(); ++k) pl[26 + off + k] = v.data()[k];
    off += v.size();
    for (size_t r = 0; r < rows; ++r) for (int k = 0; k < 3; ++k) pl[58 + r * 30 + 3 * i + k] = A[i][r * 3 + k];
    uuids.push_back(v.uuid());
  }
  for (size_t r = 0; r < rows; ++r) pl[11 + r] = b[r];
  return std::make_shared<Constraint>(NextConstraintUuid(), "fuse_constraints::MarginalConstraint", source, uuids, pl);
}
}  // namespace constraints

void UpdateGraph(HashGraph& graph, const Transaction& t) {
  for (const auto& c : t.added_constraints) graph.addConstraint(c);
  for (const auto& u : t.removed_constraints) graph.removeConstraint(u);
  for (const auto& u : t.removed_variables) graph.removeVariable(u);
}

// Householder QR in place, R in the upper triangle, the reflectors' signs as Eigen's makeHouseholder picks them
// (beta = -sign(c0) |x|; Householder.h:66-95), which is what the reference's golden values contain.
static void HouseholderR(std::vector<double>& M, int rows, int cols) {
  const int size = std::min(rows, cols);
  std::vector<double> v(rows);
  for (int k = 0; k < size; ++k) {
    double tail = 0.0;
    for (int i = k + 1; i < rows; ++i) tail += M[i * cols + k] * M[i * cols + k];
    const double c0 = M[k * cols + k];
    if (tail <= std::numeric_limits<double>::min()) continue;   // already triangular in this column: tau = 0
    double beta = std::sqrt(c0 * c0 + tail);
    if (c0 >= 0.0) beta = -beta;
    v[k] = 1.0;
    for (int i = k + 1; i < rows; ++i) v[i] = M[i * cols + k] / (c0 - beta);
    const double tau = (beta - c0) / beta;
    for (int j = k; j < cols; ++j) {       // apply H = I - tau v v^T to the remaining columns
      double s = 0.0;
      for (int i = k; i < rows; ++i) s += v[i] * M[i * cols + j];
      s *= tau;
      for (int i = k; i < rows; ++i) M[i * cols + j] -= s * v[i];
    }
    M[k * cols + k] = beta;
    for (int i = k + 1; i < rows; ++i) M[i * cols + k] = 0.0;
  }
  for (int i = 1; i < rows; ++i) for (int j = 0; j < std::min(i, cols); ++j) M[i * cols + j] = 0.0;
}

LinearTerm marginalizeNext(const std::vector<LinearTerm>& linear_terms) {   // marginalize_variables.cpp:401-535
  if (linear_terms.empty()) return {};
  std::vector<unsigned int> dense_to_index;
  for (const auto& t : linear_terms) dense_to_index.insert(dense_to_index.end(), t.variables.begin(), t.variables.end());
  std::sort(dense_to_index.begin(), dense_to_index.end());
  dense_to_index.erase(std::unique(dense_to_index.begin(), dense_to_index.end()), dense_to_index.end());
  std::vector<unsigned int> index_to_dense(dense_to_index.back() + 1, 0), index_to_cols(dense_to_index.back() + 1, 0);
  for (size_t d = 0; d < dense_to_index.size(); ++d) index_to_dense[dense_to_index[d]] = static_cast<unsigned int>(d);
  std::vector<int> row_offsets{0};
  for (const auto& t : linear_terms) {
    row_offsets.push_back(row_offsets.back() + static_cast<int>(t.b.size()));
    for (size_t i = 0; i < t.variables.size(); ++i) index_to_cols[t.variables[i]] = static_cast<unsigned int>(t.cols[i]);
  }
  std::vector<int> column_offsets{0};
  for (size_t d = 0; d < dense_to_index.size(); ++d) column_offsets.push_back(column_offsets.back() + static_cast<int>(index_to_cols[dense_to_index[d]]));
  const int rows = row_offsets.back(), cols = column_offsets.back() + 1;
  std::vector<double> Ab(static_cast<size_t>(rows) * cols, 0.0);
  for (size_t ti = 0; ti < linear_terms.size(); ++ti) {
    const auto& t = linear_terms[ti];
    const int r0 = row_offsets[ti], nr = static_cast<int>(t.b.size());
    for (size_t i = 0; i < t.variables.size(); ++i) {
      const int c0 = column_offsets[index_to_dense[t.variables[i]]], nc = t.cols[i];
      for (int r = 0; r < nr; ++r) for (int c = 0; c < nc; ++c) Ab[static_cast<size_t>(r0 + r) * cols + c0 + c] = t.A[i][static_cast<size_t>(r) * nc + c];
    }
    for (int r = 0; r < nr; ++r) Ab[static_cast<size_t>(r0 + r) * cols + cols - 1] = t.b[r];
  }
  HouseholderR(Ab, rows, cols);
  // the first row block is the conditional on the marginalised variable, the rows below it the marginal on the others
  const int min_row = column_offsets[1], max_row = std::min(rows, cols - 1), marginal_rows = max_row - min_row;
  LinearTerm out;
  if (marginal_rows > 0) {
    for (size_t d = 1; d < dense_to_index.size(); ++d) {
      const unsigned int index = dense_to_index[d]; const int nc = static_cast<int>(index_to_cols[index]), c0 = column_offsets[d];
      out.variables.push_back(index); out.cols.push_back(nc);
      std::vector<double> A(static_cast<size_t>(marginal_rows) * nc);
      for (int r = 0; r < marginal_rows; ++r) for (int c = 0; c < nc; ++c) A[static_cast<size_t>(r) * nc + c] = Ab[static_cast<size_t>(min_row + r) * cols + c0 + c];
      out.A.push_back(std::move(A));
    }
    out.b.resize(marginal_rows);
    for (int r = 0; r < marginal_rows; ++r) out.b[r] = Ab[static_cast<size_t>(min_row + r) * cols + cols - 1];
  }
  return out;
}

Transaction marginalizeVariables(const std::string& source, const std::vector<UUID>& marginalized_variables, GpuHashGraph& graph) {
  Transaction transaction;
  for (const auto& u : marginalized_variables) { graph.getVariable(u); transaction.removed_variables.push_back(u); }   // std::out_of_range for an unknown one
  // UuidOrdering (uuid_ordering.h): the variables to remove first, in the given order; connected ones as they appear
  std::vector<UUID> variable_order(marginalized_variables);
  std::unordered_map<UUID, unsigned int, UuidHash> order_of;
  for (size_t i = 0; i < variable_order.size(); ++i) if (!order_of.emplace(variable_order[i], static_cast<unsigned int>(i)).second) throw std::invalid_argument("marginalizeVariables: a variable is listed twice");
  auto push_back = [&](const UUID& u) { if (order_of.emplace(u, static_cast<unsigned int>(variable_order.size())).second) variable_order.push_back(u); };
  // the constraints to linearise, each with the lowest-ordered marginalised variable that uses it (:203-224)
  std::unordered_set<UUID, UuidHash> used;
  std::vector<std::pair<size_t, const Constraint*>> work;
  for (size_t i = 0; i < marginalized_variables.size(); ++i)
    for (const UUID& cu : graph.getConnectedConstraints(variable_order[i])) {
      if (!used.insert(cu).second) continue;
      const Constraint& c = graph.getConstraint(cu);
      for (const UUID& vu : c.variables()) push_back(vu);
      work.push_back({i, &c});
      transaction.removed_constraints.push_back(cu);
    }
  std::vector<std::vector<LinearTerm>> linear_terms(variable_order.size());
  if (!work.empty()) {
    // detail::linearize of all of them in one device evaluation (bsnls_linearize)
    std::vector<const Constraint*> cs; for (const auto& w : work) cs.push_back(w.second);
    std::vector<double> res, jac;
    graph.Linearize(cs, &res, &jac);
    for (size_t q = 0; q < work.size(); ++q) {
      const Constraint& c = *work[q].second;
      const int nr = c.flat_hint.nres;
      LinearTerm t;
      t.b.assign(res.begin() + q * BSNLS_MARGINAL_MAX_ROWS, res.begin() + q * BSNLS_MARGINAL_MAX_ROWS + nr);
      for (size_t bidx = 0; bidx < c.variables().size(); ++bidx) {
        t.variables.push_back(order_of.at(c.variables()[bidx])); t.cols.push_back(3);
        std::vector<double> A(static_cast<size_t>(nr) * 3);
        for (int r = 0; r < nr; ++r) for (int k = 0; k < 3; ++k) A[r * 3 + k] = jac[(q * BSNLS_MARGINAL_MAX_ROWS + r) * 30 + 3 * bidx + k];
        t.A.push_back(std::move(A));
      }
      linear_terms[work[q].first].push_back(std::move(t));
    }
  }
  // eliminate in order; a marginal goes to the bucket of its lowest-ordered variable (:231-241)
  for (size_t i = 0; i < marginalized_variables.size(); ++i) {
    LinearTerm m = marginalizeNext(linear_terms[i]);
    if (!m.variables.empty()) { const unsigned int lowest = m.variables.front(); linear_terms[lowest].push_back(std::move(m)); }
  }
  // what is left on the remaining variables becomes marginal constraints (:243-252, createMarginalConstraint :537-555)
  for (size_t i = marginalized_variables.size(); i < linear_terms.size(); ++i)
    for (const LinearTerm& t : linear_terms[i]) {
      std::vector<const Variable*> vars;
      for (unsigned int idx : t.variables) vars.push_back(&graph.getVariable(variable_order[idx]));
      // the GPU factor holds BSNLS_MARGINAL_MAX_ROWS rows: a taller marginal leaves as several constraints over the same
      // variables, one per block of rows -- the same cost, since a marginal constraint carries no loss function
      const int rows = static_cast<int>(t.b.size());
      for (int r0 = 0; r0 < rows; r0 += BSNLS_MARGINAL_MAX_ROWS) {
        const int nr = std::min<int>(BSNLS_MARGINAL_MAX_ROWS, rows - r0);
        std::vector<std::vector<double>> A;
        for (size_t i = 0; i < t.A.size(); ++i) A.emplace_back(t.A[i].begin() + static_cast<size_t>(r0) * t.cols[i], t.A[i].begin() + static_cast<size_t>(r0 + nr) * t.cols[i]);
        transaction.added_constraints.push_back(constraints::Marginal(source, vars, A, std::vector<double>(t.b.begin() + r0, t.b.begin() + r0 + nr)));
      }
    }
  return transaction;
}

}  // namespace bsfuse
