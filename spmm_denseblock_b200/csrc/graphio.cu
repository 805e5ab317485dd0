// graphio.cu -- graph files on the host side of the path.  HOST code only:
// no CUDA call, so it works (and is tested) without a GPU.
//
//   text      <prefix>_indptr.txt  = "<n+1>\n" + n+1 ints each followed by ' ' + "\n"
//             <prefix>_indices.txt = "<nnz>\n" + nnz ints each followed by ' ' + "\n"
//             (dumpCSRToFile / loadCSRFromFile, load_data.cc:125-165) -- read and written byte for byte
//   edge list "<n> <nnz>" then nnz pairs "x y" (loadGraphFromFile, load_data.cc:167-184: neighbours sorted,
//             duplicates kept, nnz taken from the header) -> CSR (convertGraphToCSR, utility.cc:31-45)
//   binary    <prefix>.csrbin = 32-byte header {"SPMMDBG1", int64 rows, int64 cols, int64 nnz},
//             int32 rowptr[rows+1], int32 colind[nnz], little endian: one read() per array
//
// The text parsers mmap the file, cut it into one byte range per thread at token boundaries, count the
// tokens of every range, and parse each range straight into its final position.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "spmmdb.h"

namespace {

struct Mapped {
  const char* p = nullptr;
  size_t len = 0;
  int fd = -1;
  bool open(const std::string& path) {
    fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0) return false;
    len = (size_t)st.st_size;
    if (len == 0) return true;
    void* m = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
    if (m == MAP_FAILED) return false;
    madvise(m, len, MADV_SEQUENTIAL);
    p = (const char*)m;
    return true;
  }
  ~Mapped() {
    if (p) munmap((void*)p, len);
    if (fd >= 0) close(fd);
  }
};

inline bool is_tok(char c) { return (c >= '0' && c <= '9') || c == '-' || c == '+'; }

int io_threads(size_t bytes) {
  const char* e = getenv("SPMMDB_IO_THREADS");
  int t = e ? atoi(e) : (int)std::thread::hardware_concurrency();
  t = std::max(1, std::min(t, 32));
  return (int)std::max<size_t>(1, std::min<size_t>((size_t)t, bytes / (1u << 20) + 1));
}

template <class F>
void parallel_for(int parts, F f) {
  if (parts == 1) { f(0); return; }
  std::vector<std::thread> th;
  for (int i = 1; i < parts; ++i) th.emplace_back(f, i);
  f(0);
  for (auto& t : th) t.join();
}

size_t count_tokens(const char* b, const char* e) {
  size_t n = 0;
  bool in = false;
  for (const char* q = b; q < e; ++q) {
    bool t = is_tok(*q);
    n += (t && !in);
    in = t;
  }
  return n;
}

// parses every integer token in [b, e) into out (at most cap), returns false on overflow of int range
bool parse_tokens(const char* b, const char* e, long long* out, size_t cap) {
  size_t n = 0;
  const char* q = b;
  while (q < e) {
    while (q < e && !is_tok(*q)) ++q;
    if (q >= e) break;
    bool neg = false;
    if (*q == '-' || *q == '+') { neg = (*q == '-'); ++q; }
    long long v = 0;
    while (q < e && *q >= '0' && *q <= '9') { v = v * 10 + (*q - '0'); ++q; }
    if (n < cap) out[n] = neg ? -v : v;
    ++n;
  }
  return n <= cap;
}

// all integer tokens of a whole text file
bool read_all_ints(const std::string& path, std::vector<long long>& out) {
  Mapped f;
  if (!f.open(path)) return false;
  const int parts = io_threads(f.len);
  std::vector<size_t> cut(parts + 1, f.len);
  cut[0] = 0;
  for (int i = 1; i < parts; ++i) {
    size_t c = f.len / parts * i;
    while (c < f.len && is_tok(f.p[c])) ++c;  // never cut inside a token
    cut[i] = c;
  }
  for (int i = 1; i <= parts; ++i) cut[i] = std::max(cut[i], cut[i - 1]);
  std::vector<size_t> cnt(parts + 1, 0);
  parallel_for(parts, [&](int i) { cnt[i + 1] = count_tokens(f.p + cut[i], f.p + cut[i + 1]); });
  for (int i = 0; i < parts; ++i) cnt[i + 1] += cnt[i];
  out.resize(cnt[parts]);
  parallel_for(parts, [&](int i) {
    parse_tokens(f.p + cut[i], f.p + cut[i + 1], out.data() + cnt[i], cnt[i + 1] - cnt[i]);
  });
  return true;
}

// "<count>\n v v v ..." -> the first <count> values as int32 (malloc)
int read_counted_file(const std::string& path, int** out, long long* count) {
  std::vector<long long> v;
  if (!read_all_ints(path, v)) return SPMMDB_ERR_IO;
  if (v.empty() || v[0] < 0 || (size_t)v[0] > v.size() - 1) return SPMMDB_ERR_IO;  // short file
  const long long n = v[0];
  int* a = (int*)malloc(sizeof(int) * (size_t)std::max<long long>(n, 1));
  if (!a) return SPMMDB_ERR_ALLOC;
  const int parts = io_threads((size_t)n * 8);
  parallel_for(parts, [&](int i) {
    for (long long j = n / parts * i, e = (i == parts - 1) ? n : n / parts * (i + 1); j < e; ++j) a[j] = (int)v[j + 1];
  });
  *out = a;
  *count = n;
  return SPMMDB_OK;
}

// Graph files are square adjacency matrices (n rows, n columns): a column id outside [0, n) would become an
// out-of-bounds gather of a B row on the device, so it is refused here.
int check_csr(int n, long long nnz, const int* rowptr, const int* colind) {
  if (n < 0 || rowptr[0] != 0 || rowptr[n] != nnz) return SPMMDB_ERR_INVALID_VALUE;
  for (int i = 0; i < n; ++i)
    if (rowptr[i + 1] < rowptr[i]) return SPMMDB_ERR_INVALID_VALUE;
  const int parts = io_threads((size_t)nnz * 4);
  std::vector<int> bad(parts, 0);
  parallel_for(parts, [&](int i) {
    for (long long j = nnz / parts * i, e = (i == parts - 1) ? nnz : nnz / parts * (i + 1); j < e; ++j)
      if ((unsigned)colind[j] >= (unsigned)n) { bad[i] = 1; break; }
  });
  for (int b : bad)
    if (b) return SPMMDB_ERR_INVALID_VALUE;
  return SPMMDB_OK;
}

int read_text(const std::string& prefix, int* n, int64_t* nnz, int** rowptr, int** colind) {
  int *rp = nullptr, *ci = nullptr;
  long long np1 = 0, nz = 0;
  int rc = read_counted_file(prefix + "_indptr.txt", &rp, &np1);
  if (rc == SPMMDB_OK) rc = read_counted_file(prefix + "_indices.txt", &ci, &nz);
  if (rc == SPMMDB_OK && np1 < 1) rc = SPMMDB_ERR_IO;
  if (rc == SPMMDB_OK) rc = check_csr((int)np1 - 1, nz, rp, ci);
  if (rc != SPMMDB_OK) { free(rp); free(ci); return rc; }
  *n = (int)np1 - 1; *nnz = nz; *rowptr = rp; *colind = ci;
  return SPMMDB_OK;
}

int read_edgelist(const std::string& path, int* n_out, int64_t* nnz_out, int** rowptr, int** colind) {
  std::vector<long long> v;
  if (!read_all_ints(path, v)) return SPMMDB_ERR_IO;
  if (v.size() < 2 || v[0] < 0 || v[1] < 0) return SPMMDB_ERR_IO;
  const long long n = v[0], nnz = v[1];
  if (v.size() < 2 + 2 * (size_t)nnz || n > 0x7fffffffLL || nnz > 0x7fffffffLL) return SPMMDB_ERR_IO;
  int* rp = (int*)calloc((size_t)n + 1, sizeof(int));
  int* ci = (int*)malloc(sizeof(int) * (size_t)std::max<long long>(nnz, 1));
  if (!rp || !ci) { free(rp); free(ci); return SPMMDB_ERR_ALLOC; }
  for (long long e = 0; e < nnz; ++e) {
    long long x = v[2 + 2 * e], y = v[3 + 2 * e];
    if (x < 0 || x >= n || y < 0 || y >= n) { free(rp); free(ci); return SPMMDB_ERR_INVALID_VALUE; }
    ++rp[x + 1];
  }
  for (long long i = 0; i < n; ++i) rp[i + 1] += rp[i];
  std::vector<int> fill(rp, rp + n);
  for (long long e = 0; e < nnz; ++e) ci[fill[v[2 + 2 * e]]++] = (int)v[3 + 2 * e];
  const int parts = io_threads((size_t)nnz * 8);
  parallel_for(parts, [&](int i) {  // std::sort per neighbour list, as the reference does (duplicates stay)
    for (long long r = n / parts * i, e = (i == parts - 1) ? n : n / parts * (i + 1); r < e; ++r)
      std::sort(ci + rp[r], ci + rp[r + 1]);
  });
  *n_out = (int)n; *nnz_out = nnz; *rowptr = rp; *colind = ci;
  return SPMMDB_OK;
}

const char kMagic[8] = {'S', 'P', 'M', 'M', 'D', 'B', 'G', '1'};

bool read_fully(int fd, void* dst, size_t bytes) {
  char* d = (char*)dst;
  while (bytes) {
    ssize_t r = ::read(fd, d, std::min<size_t>(bytes, (size_t)1 << 30));
    if (r <= 0) return false;
    d += r; bytes -= (size_t)r;
  }
  return true;
}

int read_binary(const std::string& prefix, int* n, int64_t* nnz, int** rowptr, int** colind) {
  int fd = ::open((prefix + ".csrbin").c_str(), O_RDONLY);
  if (fd < 0) return SPMMDB_ERR_IO;
  struct { char magic[8]; int64_t rows, cols, nnz; } h;
  int rc = SPMMDB_OK;
  int *rp = nullptr, *ci = nullptr;
  if (!read_fully(fd, &h, sizeof(h)) || memcmp(h.magic, kMagic, 8) != 0 || h.rows < 0 || h.rows > 0x7ffffffeLL ||
      h.nnz < 0 || h.nnz > 0x7fffffffLL)
    rc = SPMMDB_ERR_IO;
  if (rc == SPMMDB_OK) {
    rp = (int*)malloc(sizeof(int) * ((size_t)h.rows + 1));
    ci = (int*)malloc(sizeof(int) * (size_t)std::max<int64_t>(h.nnz, 1));
    if (!rp || !ci) rc = SPMMDB_ERR_ALLOC;
  }
  if (rc == SPMMDB_OK && (!read_fully(fd, rp, sizeof(int) * ((size_t)h.rows + 1)) ||
                          !read_fully(fd, ci, sizeof(int) * (size_t)h.nnz)))
    rc = SPMMDB_ERR_IO;
  close(fd);
  if (rc == SPMMDB_OK) rc = check_csr((int)h.rows, h.nnz, rp, ci);
  if (rc != SPMMDB_OK) { free(rp); free(ci); return rc; }
  *n = (int)h.rows; *nnz = h.nnz; *rowptr = rp; *colind = ci;
  return SPMMDB_OK;
}

// "<count>\n" + "v " per value + "\n": what operator<< and std::endl produce in dumpCSRToFile
int write_counted_file(const std::string& path, const int* a, long long count) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) return SPMMDB_ERR_IO;
  const int parts = io_threads((size_t)count * 8);
  std::vector<std::string> chunk(parts);
  parallel_for(parts, [&](int i) {
    long long b = count / parts * i, e = (i == parts - 1) ? count : count / parts * (i + 1);
    std::string& s = chunk[i];
    s.reserve((size_t)(e - b) * 8);
    char tmp[16];
    for (long long j = b; j < e; ++j) {
      long long v = a[j];
      bool neg = v < 0;
      if (neg) v = -v;
      int k = 0;
      do { tmp[k++] = (char)('0' + v % 10); v /= 10; } while (v);
      if (neg) s.push_back('-');
      while (k) s.push_back(tmp[--k]);
      s.push_back(' ');
    }
  });
  bool ok = fprintf(f, "%lld\n", count) > 0;
  for (auto& s : chunk) ok = ok && (s.empty() || fwrite(s.data(), 1, s.size(), f) == s.size());
  ok = ok && fputc('\n', f) != EOF;
  ok = (fclose(f) == 0) && ok;
  return ok ? SPMMDB_OK : SPMMDB_ERR_IO;
}

bool exists(const std::string& p) {
  struct stat st;
  return stat(p.c_str(), &st) == 0;
}

}  // namespace

extern "C" {

int spmmdb_graph_read(const char* path, int format, int* n, int64_t* nnz, int** rowptr, int** colind) {
  if (!path || !n || !nnz || !rowptr || !colind) return SPMMDB_ERR_INVALID_VALUE;
  const std::string p(path);
  switch (format) {
    case SPMMDB_GRAPH_TEXT: return read_text(p, n, nnz, rowptr, colind);
    case SPMMDB_GRAPH_BINARY: return read_binary(p, n, nnz, rowptr, colind);
    case SPMMDB_GRAPH_EDGELIST: return read_edgelist(p, n, nnz, rowptr, colind);
    case SPMMDB_GRAPH_AUTO:
      return exists(p + ".csrbin") ? read_binary(p, n, nnz, rowptr, colind) : read_text(p, n, nnz, rowptr, colind);
    default: return SPMMDB_ERR_INVALID_VALUE;
  }
}

int spmmdb_graph_write(const char* prefix, int format, int n, int64_t nnz, const int* rowptr, const int* colind) {
  if (!prefix || n < 0 || nnz < 0 || !rowptr || (nnz > 0 && !colind)) return SPMMDB_ERR_INVALID_VALUE;
  const std::string p(prefix);
  if (format == SPMMDB_GRAPH_TEXT) {
    int rc = write_counted_file(p + "_indptr.txt", rowptr, (long long)n + 1);
    return rc != SPMMDB_OK ? rc : write_counted_file(p + "_indices.txt", colind, nnz);
  }
  if (format != SPMMDB_GRAPH_BINARY) return SPMMDB_ERR_INVALID_VALUE;
  FILE* f = fopen((p + ".csrbin").c_str(), "wb");
  if (!f) return SPMMDB_ERR_IO;
  struct { char magic[8]; int64_t rows, cols, nnz; } h;
  memcpy(h.magic, kMagic, 8);
  h.rows = h.cols = n;
  h.nnz = nnz;
  bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && fwrite(rowptr, sizeof(int), (size_t)n + 1, f) == (size_t)n + 1 &&
            (nnz == 0 || fwrite(colind, sizeof(int), (size_t)nnz, f) == (size_t)nnz);
  ok = (fclose(f) == 0) && ok;
  return ok ? SPMMDB_OK : SPMMDB_ERR_IO;
}

void spmmdb_graph_free(void* p) { free(p); }

}  // extern "C"
