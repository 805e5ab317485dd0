// ref_host_wrap.cc -- C entry points around the reference's OWN host helpers
// (utility.cc calculateNnzb / convertGraphToCSR, reorder_strategy.cc permutate,
// load_data.cc text formats), which are compiled from /root/reference where
// they lie and linked next to this file.  TEST INFRASTRUCTURE (oracle/_ref).
#include <cstring>
#include <string>
#include <vector>
#include "load_data.h"          // /root/reference/load_data.h
#include "reorder_strategy.h"   // /root/reference/reorder_strategy.h
#include "utility.h"            // /root/reference/utility.h

static std::vector<std::vector<int>> to_adj(int n, const int* rowptr, const int* colind) {
  std::vector<std::vector<int>> e(n);
  for (int i = 0; i < n; ++i) e[i].assign(colind + rowptr[i], colind + rowptr[i + 1]);
  return e;
}

static void from_adj(const std::vector<std::vector<int>>& e, int* rowptr, int* colind) {
  int cnt = 0;
  rowptr[0] = 0;
  for (size_t i = 0; i < e.size(); ++i) {
    for (int c : e[i]) colind[cnt++] = c;
    rowptr[i + 1] = cnt;
  }
}

extern "C" {

int ref_calculate_nnzb(int n, const int* rowptr, const int* colind, int bs) {
  return calculateNnzb(to_adj(n, rowptr, colind), bs);       // utility.cc:47-69
}

void ref_permutate(int n, const int* rowptr, const int* colind, const int* old2new,
                   int* new_rowptr, int* new_colind) {
  std::vector<int> p(old2new, old2new + n);
  from_adj(permutate(p, to_adj(n, rowptr, colind)), new_rowptr, new_colind);  // reorder_strategy.cc:26-39
}

// reverseCuthillMcKee (reorder_strategy.cc:57-66 -> BFSTraversal :68-98 -> permutate):
// returns the REORDERED graph (the reference never exposes the permutation).
void ref_rcmk_graph(int n, const int* rowptr, const int* colind, int* new_rowptr,
                    int* new_colind) {
  from_adj(reverseCuthillMcKee(to_adj(n, rowptr, colind)), new_rowptr, new_colind);
}

// BFSTraversal (reorder_strategy.cc:68-98) and maxDegreeSort (:41-55): the reordered graphs
void ref_bfs_graph(int n, const int* rowptr, const int* colind, int* new_rowptr, int* new_colind) {
  from_adj(BFSTraversal(to_adj(n, rowptr, colind)), new_rowptr, new_colind);
}

void ref_max_degree_sort_graph(int n, const int* rowptr, const int* colind, int* new_rowptr, int* new_colind) {
  from_adj(maxDegreeSort(to_adj(n, rowptr, colind)), new_rowptr, new_colind);
}

void ref_dump_csr(const char* prefix, int n, int nnz, int* rowptr, int* colind) {
  dumpCSRToFile(prefix, n, nnz, rowptr, colind);              // load_data.cc:125-141
}

// returns n; *nnz_out = nnz; arrays malloc'ed by the reference, copied out
int ref_load_csr(const char* prefix, int* rowptr_out, int cap_rows, int* colind_out,
                 int cap_nnz, int* nnz_out) {
  int *rp = nullptr, *ci = nullptr;
  std::pair<int, int> r = loadCSRFromFile(prefix, &rp, &ci);  // load_data.cc:143-165
  if (r.first + 1 <= cap_rows && r.second <= cap_nnz) {
    std::memcpy(rowptr_out, rp, sizeof(int) * (size_t)(r.first + 1));
    std::memcpy(colind_out, ci, sizeof(int) * (size_t)r.second);
  }
  free(rp); free(ci);
  *nnz_out = r.second;
  return r.first;
}

// edge-list file -> CSR (loadGraphFromFile load_data.cc:167-184 + convertGraphToCSR utility.cc:31-45)
int ref_load_graph_csr(const char* filename, int* rowptr_out, int cap_rows,
                       int* colind_out, int cap_nnz, int* nnz_out) {
  std::vector<std::vector<int>> e;
  int nnz = loadGraphFromFile(filename, e);
  int n = (int)e.size();
  if (n + 1 <= cap_rows && nnz <= cap_nnz) from_adj(e, rowptr_out, colind_out);
  *nnz_out = nnz;
  return n;
}

}  // extern "C"
