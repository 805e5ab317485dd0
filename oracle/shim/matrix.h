// matrix.h -- shim for the header /root/reference/spmm.cc includes but the
// reference repository does not ship.  TEST INFRASTRUCTURE (oracle/_ref build
// only).  The interface is inferred from its uses: spmm.cc:9-20,47-51,56,64-66
// and spmm.cu:133-146 (CSRMatrix(rows, cols, indptr, indices), COOMatrix(rows,
// cols, nnz, row, col), DenseMatrix(rows, cols, data) + dump(), vec2ptr
// overloads, random_csr / random_coo / random_dense).
#ifndef SPMMDB_ORACLE_MATRIX_SHIM_H
#define SPMMDB_ORACLE_MATRIX_SHIM_H
#include <cstdint>
#include <cstdlib>
#include <initializer_list>
#include <iostream>
#include <random>
#include <vector>

struct CSRMatrix {
  int64_t num_rows, num_cols;
  int64_t* indptr;
  int64_t* indices;
  CSRMatrix(int64_t r, int64_t c, int64_t* ip, int64_t* ix)
      : num_rows(r), num_cols(c), indptr(ip), indices(ix) {}
};

struct COOMatrix {
  int64_t num_rows, num_cols, nnz;
  int64_t* row;
  int64_t* col;
  COOMatrix(int64_t r, int64_t c, int64_t z, int64_t* rw, int64_t* cl)
      : num_rows(r), num_cols(c), nnz(z), row(rw), col(cl) {}
};

struct DenseMatrix {
  int64_t num_rows, num_cols;
  double* data;
  DenseMatrix(int64_t r, int64_t c, double* d)
      : num_rows(r), num_cols(c), data(d) {}
  void dump() const {
    for (int64_t i = 0; i < num_rows; ++i) {
      for (int64_t j = 0; j < num_cols; ++j) std::cout << data[i * num_cols + j] << " ";
      std::cout << std::endl;
    }
  }
};

inline int64_t* vec2ptr(std::initializer_list<int64_t> v) {
  int64_t* p = (int64_t*)malloc(v.size() * sizeof(int64_t));
  size_t i = 0;
  for (int64_t x : v) p[i++] = x;
  return p;
}

inline double* vec2ptr(std::initializer_list<std::initializer_list<double>> v) {
  size_t n = 0;
  for (auto& r : v) n += r.size();
  double* p = (double*)malloc(n * sizeof(double));
  size_t i = 0;
  for (auto& r : v) for (double x : r) p[i++] = x;
  return p;
}

inline std::mt19937_64& shim_gen() { static std::mt19937_64 g(1234); return g; }

inline CSRMatrix random_csr(int64_t l, int64_t m, double p = 0.01) {
  std::uniform_real_distribution<double> flip(0, 1);
  std::vector<int64_t> ip(1, 0), ix;
  for (int64_t i = 0; i < l; ++i) {
    for (int64_t j = 0; j < m; ++j) if (flip(shim_gen()) < p) ix.push_back(j);
    ip.push_back((int64_t)ix.size());
  }
  int64_t* a = (int64_t*)malloc(ip.size() * sizeof(int64_t));
  int64_t* b = (int64_t*)malloc((ix.size() + 1) * sizeof(int64_t));
  for (size_t i = 0; i < ip.size(); ++i) a[i] = ip[i];
  for (size_t i = 0; i < ix.size(); ++i) b[i] = ix[i];
  return CSRMatrix(l, m, a, b);
}

inline COOMatrix random_coo(int64_t l, int64_t m, double p = 0.01) {
  CSRMatrix c = random_csr(l, m, p);
  int64_t nnz = c.indptr[l];
  int64_t* row = (int64_t*)malloc((nnz + 1) * sizeof(int64_t));
  for (int64_t i = 0; i < l; ++i)
    for (int64_t j = c.indptr[i]; j < c.indptr[i + 1]; ++j) row[j] = i;
  return COOMatrix(l, m, nnz, row, c.indices);
}

inline DenseMatrix random_dense(int64_t m, int64_t n) {
  std::uniform_real_distribution<double> dist(-1, 1);
  double* d = (double*)malloc((size_t)(m * n) * sizeof(double));
  for (int64_t i = 0; i < m * n; ++i) d[i] = dist(shim_gen());
  return DenseMatrix(m, n, d);
}
#endif
