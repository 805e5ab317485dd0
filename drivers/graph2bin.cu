// graph2bin -- one-off conversion of a graph of the reference's pipeline into the binary cache format the
// drivers prefer.  No reference counterpart: the reference re-parses the text files
// (load_data.cc:143-184) on every run.
//   graph2bin <prefix>                 tmp-style text pair <prefix>_indptr.txt/_indices.txt -> <prefix>.csrbin
//   graph2bin --edges <file> <prefix>  edge list "n nnz / x y ..." -> <prefix>.csrbin and the text pair
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>

#include "spmmdb.h"

int main(int argc, char** argv) {
  const bool edges = argc == 4 && !strcmp(argv[1], "--edges");
  if (argc != 2 && !edges) {
    fprintf(stderr, "usage: %s <prefix> | --edges <edge list file> <prefix>\n", argv[0]);
    return 2;
  }
  const std::string src = edges ? argv[2] : argv[1], dst = edges ? argv[3] : argv[1];
  int n = 0;
  int64_t nnz = 0;
  int *rp = nullptr, *ci = nullptr;
  auto t0 = std::chrono::steady_clock::now();
  int rc = spmmdb_graph_read(src.c_str(), edges ? SPMMDB_GRAPH_EDGELIST : SPMMDB_GRAPH_TEXT, &n, &nnz, &rp, &ci);
  if (rc != SPMMDB_OK) {
    fprintf(stderr, "cannot read %s (status %d)\n", src.c_str(), rc);
    return 1;
  }
  auto t1 = std::chrono::steady_clock::now();
  rc = spmmdb_graph_write(dst.c_str(), SPMMDB_GRAPH_BINARY, n, nnz, rp, ci);
  if (rc == SPMMDB_OK && edges) rc = spmmdb_graph_write(dst.c_str(), SPMMDB_GRAPH_TEXT, n, nnz, rp, ci);
  auto t2 = std::chrono::steady_clock::now();
  spmmdb_graph_free(rp);
  spmmdb_graph_free(ci);
  if (rc != SPMMDB_OK) {
    fprintf(stderr, "cannot write %s (status %d)\n", dst.c_str(), rc);
    return 1;
  }
  printf("n = %d nnz = %lld read %.3f s write %.3f s\n", n, (long long)nnz,
         std::chrono::duration<double>(t1 - t0).count(), std::chrono::duration<double>(t2 - t1).count());
  return 0;
}
