// reorder.cu -- the simple orderings of the reference produced on the device: the permutation
// old2new only; applying it is spmmdb_permute_csr_f32 / spmmdb_permute_rows_f32 (permute.cu).
//
//   spmmdb_order_by_degree   maxDegreeSort (reorder_strategy.cc:41-55): vertices by descending degree.  The reference
//                            orders equal degrees however std::sort leaves them (unstable); here ties keep ascending
//                            id, so the permuted graph has the reference's row lengths always and the reference's
//                            rows whenever the degrees are distinct.
//   spmmdb_order_bfs         BFSTraversal (reorder_strategy.cc:68-98): the numbering a FIFO breadth-first search gives,
//                            started at the lowest unvisited id, neighbours taken in adjacency order.  Exactly the
//                            reference's permutation, for any pattern (symmetric or not).
//   spmmdb_order_rcm         reverseCuthillMcKee (:57-66): that search on adjacency lists sorted by descending degree.
//
// Integer work, deterministic.  The search runs level by level for ALL roots at once.  A sequential FIFO search
// restarts at the lowest unvisited id whenever its queue runs empty; vertex v is therefore numbered by the search of
// root label(v) = the lowest id among the vertices that can reach v (if a smaller vertex reached v, that vertex or its
// own root would have claimed v first), the roots are the vertices with label(v) = v, and the search of root r visits
// exactly T_r = { v : label(v) = r }: an edge out of T_r leads into T_r or into the territory of a smaller root, which
// is already numbered, and every vertex of T_r is reached from r through T_r alone.  (For a symmetric pattern T_r is a
// connected component and r its lowest id.)  Inside T_r the search numbers level by level, a level in the order
// (position of the parent in its level, position in the parent's adjacency list), the parent being the first frontier
// vertex that lists the child.  So
//   old2new[v] = (vertices with a smaller label) + (vertices of T_label(v) in earlier levels)
//                + (rank of v in its level among the vertices with its label),
// and the frontier array of a level, built by stable compaction of the winning (parent, position) slots, is ordered
// by label first because the roots are.
#include <limits.h>

#include "common.cuh"

namespace spmmdb {
int exclusive_scan_i32(const int* counts, int n, int* sums, int* out, cudaStream_t st);  // convert.cu
int permute_csr_impl(int n, const int* rowptr, const int* colind, const float* val, const int* old2new, int rows_stay,
                     int* new_rowptr, int* new_colind, float* new_val, cudaStream_t st);  // permute.cu
namespace {

constexpr unsigned FULL = 0xffffffffu;

// ---- degree order: stable binary radix passes on key = maxdeg - degree -------------------------------------------
__global__ void __launch_bounds__(256) deg_key_kernel(int n, const int* __restrict__ rowptr, int* __restrict__ deg,
                                                      int* __restrict__ maxdeg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  int d = 0;
  if (i < n) { d = __ldg(rowptr + i + 1) - __ldg(rowptr + i); deg[i] = d; }
  d = __reduce_max_sync(FULL, d);
  if ((threadIdx.x & 31) == 0 && d > 0) atomicMax(maxdeg, d);
}

__global__ void __launch_bounds__(256) radix_init_kernel(int n, const int* __restrict__ deg, int maxdeg, int* __restrict__ key,
                                                         int* __restrict__ id) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { key[i] = maxdeg - deg[i]; id[i] = i; }
}

__global__ void __launch_bounds__(256) radix_flag_kernel(int n, const int* __restrict__ key, int bit, int* __restrict__ zero) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) zero[i] = ((key[i] >> bit) & 1) ^ 1;
}

// stable split: entries whose bit is 0 first (in order), then the others (in order); zpos = exclusive scan of `zero`
__global__ void __launch_bounds__(256) radix_scatter_kernel(int n, const int* __restrict__ key, const int* __restrict__ id,
                                                            int bit, const int* __restrict__ zpos, int* __restrict__ key_out,
                                                            int* __restrict__ id_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int k = key[i], z = zpos[i];
  const int dst = ((k >> bit) & 1) ? zpos[n] + (i - z) : z;
  key_out[dst] = k;
  id_out[dst] = id[i];
}

__global__ void __launch_bounds__(256) invert_kernel(int n, const int* __restrict__ id, int* __restrict__ old2new) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) old2new[id[i]] = i;
}

// ---- breadth-first order ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) iota_fill_kernel(int n, int* __restrict__ label, int* __restrict__ key, int key_init) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { label[i] = i; key[i] = key_init; }
}

// label[v] = lowest id among the vertices that reach v (v included): the root whose search numbers v.  An edge u -> v
// hands u's label to v; labels only decrease and label[label[v]] reaches v too, so jumping to it is valid.  The fixed
// point is reached when no edge and no jump lowers anything.  (On a symmetric pattern these are the connected
// components, labelled by their lowest id.)  A warp per row.
__global__ void __launch_bounds__(256) cc_hook_kernel(int n, const int* __restrict__ rowptr, const int* __restrict__ colind,
                                                      int* label, int* __restrict__ changed) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  bool any = false;
  for (int64_t u = warp; u < n; u += nwarps) {
    const int s = __ldg(rowptr + u), e = __ldg(rowptr + u + 1);
    const int lu = __ldcg(label + u);
    for (int p = s + lane; p < e; p += 32) {
      const int v = __ldg(colind + p);
      if (lu < __ldcg(label + v)) {
        atomicMin(label + v, lu);
        any = true;
      }
    }
  }
  if (__any_sync(FULL, any) && lane == 0) *changed = 1;
}

__global__ void __launch_bounds__(256) cc_jump_kernel(int n, int* label, int* __restrict__ changed) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n) return;
  const int l0 = __ldcg(label + v);
  int l = l0;
  for (;;) {
    const int ll = __ldcg(label + l);
    if (ll == l) break;
    l = ll;
  }
  if (l < l0) { atomicMin(label + v, l); *changed = 1; }
}

// out[i] = 1 where vertex i is the lowest id of its component; size[label] counts the component
__global__ void __launch_bounds__(256) cc_roots_kernel(int n, const int* __restrict__ label, int* __restrict__ is_root,
                                                       int* __restrict__ size) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n) return;
  const int l = label[v];
  is_root[v] = (l == v);
  atomicAdd(size + l, 1);
}

// level 0: the roots in id order; a root is the first vertex of its component
__global__ void __launch_bounds__(256) bfs_roots_kernel(int n, const int* __restrict__ label, const int* __restrict__ rpos,
                                                        int* __restrict__ frontier, int* __restrict__ base,
                                                        int* __restrict__ old2new) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n || label[v] != v) return;
  frontier[rpos[v]] = v;
  old2new[v] = base[v];
  base[v] += 1;
}

__global__ void __launch_bounds__(256) bfs_degree_kernel(int nf, const int* __restrict__ frontier, const int* __restrict__ rowptr,
                                                         int* __restrict__ deg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nf) { const int u = frontier[i]; deg[i] = __ldg(rowptr + u + 1) - __ldg(rowptr + u); }
}

// MODE 0: every unvisited neighbour is offered its slot (offset of the parent in the level's edge list + position in
//         the parent's list); the lowest slot is the one a FIFO search would have taken.
// MODE 1: count the slots that won, per frontier vertex.
// MODE 2: write the winners to the next frontier, in slot order (woff = exclusive scan of the counts).
template <int MODE>
__global__ void __launch_bounds__(256) bfs_expand_kernel(int nf, const int* __restrict__ frontier, const int* __restrict__ rowptr,
                                                         const int* __restrict__ colind, const int* __restrict__ eoff,
                                                         const int* __restrict__ label, const int* old2new, int* key,
                                                         int* __restrict__ cnt, const int* __restrict__ woff,
                                                         int* __restrict__ next) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp; i < nf; i += nwarps) {
    const int u = frontier[i];
    const int s = __ldg(rowptr + u), d = __ldg(rowptr + u + 1) - s;
    const int slot0 = eoff[i], lu = label[u];
    int run = 0;
    for (int j0 = 0; j0 < d; j0 += 32) {
      const int j = j0 + lane;
      int v = -1;
      bool open = false;
      // a neighbour with another label belongs to an earlier root's search (numbered before this one started)
      if (j < d) { v = __ldg(colind + s + j); open = __ldcg(old2new + v) < 0 && label[v] == lu; }
      if (MODE == 0) {
        if (open) atomicMin(key + v, slot0 + j);
      } else {
        const bool win = open && __ldcg(key + v) == slot0 + j;
        const unsigned mask = __ballot_sync(FULL, win);
        if (MODE == 2 && win) next[woff[i] + run + __popc(mask & ((1u << lane) - 1))] = v;
        run += __popc(mask);
      }
    }
    if (MODE == 1 && lane == 0) cnt[i] = run;
  }
}

// positions of a level: STEP 0 notes where each component's run starts in the frontier, STEP 1 numbers the
// vertices, STEP 2 moves the component's next free position past them
template <int STEP>
__global__ void __launch_bounds__(256) bfs_number_kernel(int nf, const int* __restrict__ frontier, const int* __restrict__ label,
                                                         int* __restrict__ firsts, int* __restrict__ base, int* __restrict__ old2new) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nf) return;
  const int v = frontier[i], c = label[v];
  if (STEP == 0) {
    if (i == 0 || label[frontier[i - 1]] != c) firsts[c] = i;
  } else if (STEP == 1) {
    old2new[v] = base[c] + i - firsts[c];
  } else {
    if (i == nf - 1 || label[frontier[i + 1]] != c) base[c] += i - firsts[c] + 1;
  }
}

__global__ void __launch_bounds__(256) map_kernel(int64_t cnt, const int* __restrict__ table, int* __restrict__ x) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cnt; i += stride) x[i] = __ldg(table + x[i]);
}

inline int blocks_for(int64_t n, int per = 256) { return (int)((n + per - 1) / per); }

}  // namespace
}  // namespace spmmdb

using namespace spmmdb;

extern "C" int spmmdb_order_by_degree(int n, const int* rowptr, int* old2new, void* stream) {
  if (n < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (n == 0) return SPMMDB_OK;
  if (rowptr == nullptr || old2new == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int ntiles = (n + 2047) / 2048 + 1;
  // scratch: maxdeg[8] | deg n | key a, b n | id a, b n | zero n | zpos n + 1 | scan sums
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * (8 + 7 * (size_t)n + 1 + ntiles + 8), st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* maxdeg_dev = ws;
  int* deg = ws + 8;
  int* key[2] = {deg + n, deg + 2 * (size_t)n};
  int* id[2] = {deg + 3 * (size_t)n, deg + 4 * (size_t)n};
  int* zero = deg + 5 * (size_t)n;
  int* zpos = zero + n;
  int* sums = zpos + n + 1;
  const int grid = blocks_for(n);
  SPMMDB_CUDA_OK(cudaMemsetAsync(maxdeg_dev, 0, sizeof(int), st));
  deg_key_kernel<<<grid, 256, 0, st>>>(n, rowptr, deg, maxdeg_dev);
  int maxdeg = 0;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(&maxdeg, maxdeg_dev, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  radix_init_kernel<<<grid, 256, 0, st>>>(n, deg, maxdeg, key[0], id[0]);
  int cur = 0;
  for (int bit = 0; (maxdeg >> bit) != 0; ++bit) {
    radix_flag_kernel<<<grid, 256, 0, st>>>(n, key[cur], bit, zero);
    int rc = exclusive_scan_i32(zero, n, sums, zpos, st);
    if (rc != SPMMDB_OK) return rc;
    radix_scatter_kernel<<<grid, 256, 0, st>>>(n, key[cur], id[cur], bit, zpos, key[cur ^ 1], id[cur ^ 1]);
    cur ^= 1;
  }
  invert_kernel<<<grid, 256, 0, st>>>(n, id[cur], old2new);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

extern "C" int spmmdb_order_bfs(int n, const int* rowptr, const int* colind, int* old2new, void* stream) {
  if (n < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (n == 0) return SPMMDB_OK;
  if (rowptr == nullptr || old2new == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const int ntiles = (n + 2047) / 2048 + 1;
  // scratch: flags[8] | label | key | frontier a, b | cnt | off (n + 1) | firsts (component sizes first) | base (n + 1)
  //          | scan sums
  const size_t N = (size_t)n;
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * (8 + 8 * N + 3 + ntiles + 8), st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* flags = ws;
  int* label = ws + 8;
  int* key = label + N;
  int* fr[2] = {key + N, key + 2 * N};
  int* cnt = key + 3 * N;
  int* off = cnt + N;             // n + 1
  int* firsts = off + N + 1;      // n + 1 (also holds the scan of a level's winner counts)
  int* base = firsts + N + 1;     // n + 1
  int* sums = base + N + 1;
  const int grid = blocks_for(n);
  const int wgrid = (int)(((int64_t)n + 7) / 8 < (int64_t)nsm * 16 ? ((int64_t)n + 7) / 8 : (int64_t)nsm * 16);

  // 1. components, labelled by their lowest id
  iota_fill_kernel<<<grid, 256, 0, st>>>(n, label, key, INT_MAX);
  for (int it = 0;; ++it) {
    SPMMDB_CUDA_OK(cudaMemsetAsync(flags, 0, sizeof(int), st));
    cc_hook_kernel<<<wgrid, 256, 0, st>>>(n, rowptr, colind, label, flags);
    cc_jump_kernel<<<grid, 256, 0, st>>>(n, label, flags);
    int changed = 0;
    SPMMDB_CUDA_OK(cudaMemcpyAsync(&changed, flags, sizeof(int), cudaMemcpyDeviceToHost, st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    if (!changed) break;
  }
  // 2. roots in id order = level 0; base[c] = vertices in components with a lower lowest id
  SPMMDB_CUDA_OK(cudaMemsetAsync(firsts, 0, sizeof(int) * N, st));
  SPMMDB_CUDA_OK(cudaMemsetAsync(old2new, 0xff, sizeof(int) * N, st));
  cc_roots_kernel<<<grid, 256, 0, st>>>(n, label, cnt, firsts);
  int rc = exclusive_scan_i32(firsts, n, sums, base, st);
  if (rc != SPMMDB_OK) return rc;
  rc = exclusive_scan_i32(cnt, n, sums, off, st);
  if (rc != SPMMDB_OK) return rc;
  int nf = 0;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(&nf, off + n, sizeof(int), cudaMemcpyDeviceToHost, st));
  bfs_roots_kernel<<<grid, 256, 0, st>>>(n, label, off, fr[0], base, old2new);
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  // 3. level by level
  int64_t numbered = nf;
  int cur = 0;
  while (nf > 0) {
    const int fg = blocks_for(nf);
    const int fw = (int)(((int64_t)nf + 7) / 8 < (int64_t)nsm * 16 ? ((int64_t)nf + 7) / 8 : (int64_t)nsm * 16);
    bfs_degree_kernel<<<fg, 256, 0, st>>>(nf, fr[cur], rowptr, cnt);
    rc = exclusive_scan_i32(cnt, nf, sums, off, st);
    if (rc != SPMMDB_OK) return rc;
    bfs_expand_kernel<0><<<fw, 256, 0, st>>>(nf, fr[cur], rowptr, colind, off, label, old2new, key, nullptr, nullptr, nullptr);
    bfs_expand_kernel<1><<<fw, 256, 0, st>>>(nf, fr[cur], rowptr, colind, off, label, old2new, key, cnt, nullptr, nullptr);
    // the counts' scan goes to `firsts` (free between levels); the slot offsets in `off` are still needed
    rc = exclusive_scan_i32(cnt, nf, sums, firsts, st);
    if (rc != SPMMDB_OK) return rc;
    int nn = 0;
    SPMMDB_CUDA_OK(cudaMemcpyAsync(&nn, firsts + nf, sizeof(int), cudaMemcpyDeviceToHost, st));
    bfs_expand_kernel<2><<<fw, 256, 0, st>>>(nf, fr[cur], rowptr, colind, off, label, old2new, key, nullptr, firsts, fr[cur ^ 1]);
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    if (nn > 0) {
      const int ng = blocks_for(nn);
      bfs_number_kernel<0><<<ng, 256, 0, st>>>(nn, fr[cur ^ 1], label, firsts, base, old2new);
      bfs_number_kernel<1><<<ng, 256, 0, st>>>(nn, fr[cur ^ 1], label, firsts, base, old2new);
      bfs_number_kernel<2><<<ng, 256, 0, st>>>(nn, fr[cur ^ 1], label, firsts, base, old2new);
    }
    SPMMDB_CUDA_OK(cudaGetLastError());
    numbered += nn;
    nf = nn;
    cur ^= 1;
  }
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  if (numbered != n) {  // cannot happen with column indices in [0, n): every vertex is reached from its label
    set_last_error("spmmdb_order_bfs: vertices left unnumbered (column index out of range?)", cudaErrorInvalidValue, __FILE__,
                   __LINE__);
    return SPMMDB_ERR_INVALID_VALUE;
  }
  return SPMMDB_OK;
}

// reverseCuthillMcKee (reorder_strategy.cc:57-66): every adjacency list sorted by descending degree of the neighbour,
// then BFSTraversal.  (Despite the name the reference never reverses the order.)  Sorting a list by descending degree
// with ties in ascending id is sorting it by the neighbours' rank in spmmdb_order_by_degree, which the relabel-and-
// sort pass of permute.cu does with the rows left in place.  The reference sorts with std::sort, which leaves equal
// degrees in an unspecified order: lists of at most 16 entries come out stable there (insertion sort), so on graphs of
// maximum degree <= 16 this is the reference's permutation exactly; beyond that it is the same search with the ties
// among equal-degree neighbours taken in ascending id.
extern "C" int spmmdb_order_rcm(int n, const int* rowptr, const int* colind, int* old2new, void* stream) {
  if (n < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (n == 0) return SPMMDB_OK;
  if (rowptr == nullptr || old2new == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  int nnz = 0;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(&nnz, rowptr + n, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  if (nnz > 0 && colind == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  int* tmp = (int*)scratch(kScratchLayoutC, sizeof(int) * (2 * (size_t)n + (size_t)nnz + 8), st);
  if (tmp == nullptr) return SPMMDB_ERR_ALLOC;
  int* rank = tmp;           // rank[v]: position of v by (descending degree, ascending id)
  int* by_rank = tmp + n;    // its inverse
  int* lists = by_rank + n;  // the adjacency lists, each sorted by rank
  int rc = spmmdb_order_by_degree(n, rowptr, rank, stream);
  if (rc != SPMMDB_OK) return rc;
  invert_kernel<<<blocks_for(n), 256, 0, st>>>(n, rank, by_rank);
  if (nnz > 0) {
    rc = permute_csr_impl(n, rowptr, colind, nullptr, rank, /*rows_stay=*/1, const_cast<int*>(rowptr), lists, nullptr, st);
    if (rc != SPMMDB_OK) return rc;
    const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
    map_kernel<<<nsm * 8, 256, 0, st>>>(nnz, by_rank, lists);
  }
  SPMMDB_CUDA_OK(cudaGetLastError());
  return spmmdb_order_bfs(n, rowptr, lists, old2new, stream);
}
