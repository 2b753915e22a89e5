// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into or called from the
// product path (towr2025_b200/).  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may use anything in oracle/.
//
// sparse_shim.h: a tiny dependency-free stand-in for the slice of Eigen 3.3 the
// reference's hot path uses (Eigen is NOT installed here):
//   * V3 / M3                      ~ Eigen::Vector3d / Matrix3d
//   * SpVec                        ~ Eigen::SparseVector<double,RowMajor> (a Jacobian row)
//   * SpMat                        ~ Eigen::SparseMatrix<double,RowMajor>
// What matters for parity is Eigen's STRUCTURAL semantics, because the stored
// entries at x0 are the sparsity pattern Ipopt sees:
//   - coeffRef(r,c) inserts an explicit entry even if the value stays 0.0
//   - scalar*sparse and unary minus keep the pattern
//   - sparse +/- sparse stores the union of both patterns (only-lhs: a-0, only-rhs: 0-b)
//   - sparse*sparse is the conservative product: every structurally reachable
//     (i,j) is stored; value = sum over k ascending of lhs(i,k)*rhs(k,j),
//     first term assigned, later terms +=   (no value-based pruning)
//   - dense.sparseView() (default args) drops exact zeros;
//     dense.sparseView(1.0,-1.0) keeps everything
//   - middleRows / row assignment replaces the rows' stored entries
// Dense 3-vector reductions (dot, squaredNorm, small products) are evaluated
// left to right; the reference's true order depends on Eigen's vectorisation
// flags, which nothing in the reference pins (SURVEY.md §8c "parity unpinned").
#pragma once
#include <cmath>
#include <cstddef>
#include <utility>
#include <vector>

namespace orc {

struct V3 {
  double d[3];
  V3() : d{0.0, 0.0, 0.0} {}
  V3(double x, double y, double z) : d{x, y, z} {}
  double& operator()(int i) { return d[i]; }
  double operator()(int i) const { return d[i]; }
  double x() const { return d[0]; }
  double y() const { return d[1]; }
  double z() const { return d[2]; }
  static V3 Unit(int i) { V3 u; u.d[i] = 1.0; return u; }
};

inline V3 operator+(const V3& a, const V3& b) { return V3(a.d[0]+b.d[0], a.d[1]+b.d[1], a.d[2]+b.d[2]); }
inline V3 operator-(const V3& a, const V3& b) { return V3(a.d[0]-b.d[0], a.d[1]-b.d[1], a.d[2]-b.d[2]); }
inline V3 operator-(const V3& a) { return V3(-a.d[0], -a.d[1], -a.d[2]); }
inline V3 operator*(double s, const V3& a) { return V3(s*a.d[0], s*a.d[1], s*a.d[2]); }
inline V3 operator*(const V3& a, double s) { return V3(a.d[0]*s, a.d[1]*s, a.d[2]*s); }
inline V3 operator/(const V3& a, double s) { return V3(a.d[0]/s, a.d[1]/s, a.d[2]/s); }
inline V3& operator+=(V3& a, const V3& b) { a = a + b; return a; }
inline double dot(const V3& a, const V3& b) { return a.d[0]*b.d[0] + a.d[1]*b.d[1] + a.d[2]*b.d[2]; }
inline double squaredNorm(const V3& a) { return dot(a, a); }
inline double norm(const V3& a) { return std::sqrt(squaredNorm(a)); }
inline V3 normalized(const V3& a) {
  double z = squaredNorm(a);
  return z > 0.0 ? a / std::sqrt(z) : a;
}
inline V3 cross(const V3& a, const V3& b) {
  return V3(a.d[1]*b.d[2] - a.d[2]*b.d[1],
            a.d[2]*b.d[0] - a.d[0]*b.d[2],
            a.d[0]*b.d[1] - a.d[1]*b.d[0]);
}
inline V3 cwiseProduct(const V3& a, const V3& b) { return V3(a.d[0]*b.d[0], a.d[1]*b.d[1], a.d[2]*b.d[2]); }

struct M3 {
  double m[3][3];
  M3() { for (auto& r : m) for (double& v : r) v = 0.0; }
  double& operator()(int i, int j) { return m[i][j]; }
  double operator()(int i, int j) const { return m[i][j]; }
  M3 transpose() const { M3 t; for (int i=0;i<3;++i) for (int j=0;j<3;++j) t.m[i][j] = m[j][i]; return t; }
};
inline V3 operator*(const M3& A, const V3& v) {
  V3 r;
  for (int i=0;i<3;++i) r.d[i] = A.m[i][0]*v.d[0] + A.m[i][1]*v.d[1] + A.m[i][2]*v.d[2];
  return r;
}

using Entry = std::pair<int, double>;   // (inner index, value), rows kept sorted by index

struct SpVec {
  int size = 0;
  std::vector<Entry> e;
  SpVec() = default;
  explicit SpVec(int n) : size(n) {}
};

// merge two sorted rows with a binary functor f(a,b); missing side = 0.0
template <class F>
inline std::vector<Entry> merge_rows(const std::vector<Entry>& a, const std::vector<Entry>& b, F f) {
  std::vector<Entry> out;
  out.reserve(a.size() + b.size());
  size_t i = 0, j = 0;
  while (i < a.size() || j < b.size()) {
    if (j >= b.size() || (i < a.size() && a[i].first < b[j].first)) { out.emplace_back(a[i].first, f(a[i].second, 0.0)); ++i; }
    else if (i >= a.size() || b[j].first < a[i].first)              { out.emplace_back(b[j].first, f(0.0, b[j].second)); ++j; }
    else { out.emplace_back(a[i].first, f(a[i].second, b[j].second)); ++i; ++j; }
  }
  return out;
}

inline SpVec operator+(const SpVec& a, const SpVec& b) { SpVec r(a.size); r.e = merge_rows(a.e, b.e, [](double x, double y){ return x + y; }); return r; }
inline SpVec operator-(const SpVec& a, const SpVec& b) { SpVec r(a.size); r.e = merge_rows(a.e, b.e, [](double x, double y){ return x - y; }); return r; }
inline SpVec operator*(double s, const SpVec& a) { SpVec r = a; for (auto& en : r.e) en.second = s * en.second; return r; }
inline SpVec operator*(const SpVec& a, double s) { SpVec r = a; for (auto& en : r.e) en.second = en.second * s; return r; }
inline SpVec operator-(const SpVec& a) { SpVec r = a; for (auto& en : r.e) en.second = -en.second; return r; }

// dense row vector -> sparse row keeping every coefficient: v.transpose().sparseView(1.0,-1.0)
inline SpVec sparseViewAll(const V3& v) { SpVec r(3); for (int i=0;i<3;++i) r.e.emplace_back(i, v.d[i]); return r; }

struct SpMat {
  int rows = 0, cols = 0;
  std::vector<std::vector<Entry>> r;
  SpMat() = default;
  SpMat(int nr, int nc) : rows(nr), cols(nc), r(nr) {}
  void resize(int nr, int nc) { rows = nr; cols = nc; r.assign(nr, {}); }

  double& coeffRef(int i, int j) {
    auto& row = r[i];
    size_t k = 0;
    while (k < row.size() && row[k].first < j) ++k;
    if (k == row.size() || row[k].first != j) row.insert(row.begin() + k, Entry(j, 0.0));
    return row[k].second;
  }
  SpVec row(int i) const { SpVec v(cols); v.e = r[i]; return v; }
  void setRow(int i, const SpVec& v) { r[i] = v.e; }
  // jac.middleRows(start, B.rows) = B
  void setMiddleRows(int start, const SpMat& B) { for (int i = 0; i < B.rows; ++i) r[start + i] = B.r[i]; }
  size_t nonZeros() const { size_t n = 0; for (auto& row : r) n += row.size(); return n; }
  SpMat transpose() const {
    SpMat t(cols, rows);
    for (int i = 0; i < rows; ++i) for (auto& en : r[i]) t.r[en.first].emplace_back(i, en.second);
    return t;
  }
};

inline SpMat operator+(const SpMat& A, const SpMat& B) {
  SpMat R(A.rows, A.cols);
  for (int i = 0; i < A.rows; ++i) R.r[i] = merge_rows(A.r[i], B.r[i], [](double x, double y){ return x + y; });
  return R;
}
inline SpMat operator-(const SpMat& A, const SpMat& B) {
  SpMat R(A.rows, A.cols);
  for (int i = 0; i < A.rows; ++i) R.r[i] = merge_rows(A.r[i], B.r[i], [](double x, double y){ return x - y; });
  return R;
}
inline SpMat operator-(const SpMat& A) { SpMat R = A; for (auto& row : R.r) for (auto& en : row) en.second = -en.second; return R; }
inline SpMat operator*(double s, const SpMat& A) { SpMat R = A; for (auto& row : R.r) for (auto& en : row) en.second = s * en.second; return R; }
inline SpMat& operator+=(SpMat& A, const SpMat& B) { A = A + B; return A; }

// one result row of the conservative product: sum_k ascending lhs(k)*B.row(k)
inline std::vector<Entry> row_times(const std::vector<Entry>& lhs, const SpMat& B) {
  std::vector<Entry> acc;   // kept sorted by column
  for (const auto& lk : lhs) {
    const auto& brow = B.r[lk.first];
    for (const auto& bj : brow) {
      size_t k = 0;
      while (k < acc.size() && acc[k].first < bj.first) ++k;
      if (k == acc.size() || acc[k].first != bj.first) acc.insert(acc.begin() + k, Entry(bj.first, lk.second * bj.second));
      else acc[k].second += lk.second * bj.second;
    }
  }
  return acc;
}
inline SpMat operator*(const SpMat& A, const SpMat& B) {
  SpMat R(A.rows, B.cols);
  for (int i = 0; i < A.rows; ++i) R.r[i] = row_times(A.r[i], B);
  return R;
}
inline SpVec operator*(const SpVec& v, const SpMat& B) { SpVec r(B.cols); r.e = row_times(v.e, B); return r; }

// sparse (row-major) * dense vector: tmp = 0; tmp += a_ik * v_k; res = 0 + 1*tmp
inline V3 operator*(const SpMat& A, const V3& v) {
  V3 out;
  for (int i = 0; i < 3; ++i) {
    double tmp = 0.0;
    for (const auto& en : A.r[i]) tmp += en.second * v.d[en.first];
    out.d[i] += 1.0 * tmp;
  }
  return out;
}
// sparse (row-major) * dense 3x3
inline M3 operator*(const SpMat& A, const M3& B) {
  M3 out;
  for (int i = 0; i < 3; ++i)
    for (int c = 0; c < 3; ++c) {
      double tmp = 0.0;
      for (const auto& en : A.r[i]) tmp += en.second * B.m[en.first][c];
      out.m[i][c] += 1.0 * tmp;
    }
  return out;
}

// dense.sparseView()  : prune exact zeros     (|x| <= 0*eps)
// dense.sparseView(1.0,-1.0): keep everything (|x| <= -1 never)
inline SpMat sparseView(const M3& A, bool keep_zeros) {
  SpMat S(3, 3);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      if (keep_zeros || !(std::fabs(A.m[i][j]) <= 0.0)) S.r[i].emplace_back(j, A.m[i][j]);
  return S;
}
inline M3 toDense(const SpMat& S) { M3 A; for (int i = 0; i < 3; ++i) for (auto& en : S.r[i]) A.m[i][en.first] = en.second; return A; }

}  // namespace orc
