/*
 * spmmdb.h -- C ABI of the B200-native SpMM engine (libspmmdb.so).
 *
 * This is the drop-in boundary for the reference's one data-parallel hot path
 * (sparse adjacency A in CSR / BSR times a dense feature matrix B, plus the
 * csr2bsr / bsr2csr conversions).  Every entry point names the reference
 * interface it replaces (file:line in xuyifangreeneyes/spmm-denseblock).
 *
 * Conventions
 *   - plain C: pointers + sizes, no C++/torch types.  Unless a function says
 *     "host", every pointer is a DEVICE pointer owned by the caller; the
 *     callee allocates nothing the caller sees and frees nothing.
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream, as in
 *     the reference: gespmm_csrmm.h:385-417 `<<<..., 0, 0>>>`).  Calls are
 *     asynchronous unless stated otherwise.
 *   - indices are 0-based int32, values fp32 (the only instantiation the
 *     reference uses: run_csrmm.cu:144, run_bsrmm.cu:156).
 *   - return value: SPMMDB_OK (0) or a negative spmmdb_status.  The C++
 *     drop-in headers (include/gespmm_csrmm.h, include/rocsparse_bsrmm.h) map
 *     these to the reference's void / cusparseStatus_t conventions.
 *   - there is NO CPU fallback: without a CUDA device every compute entry
 *     point returns SPMMDB_ERR_CUDA.
 */
#ifndef SPMMDB_H
#define SPMMDB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  SPMMDB_OK = 0,
  SPMMDB_ERR_INVALID_VALUE = -1,  /* cusparse: CUSPARSE_STATUS_INVALID_VALUE            */
  SPMMDB_ERR_NOT_SUPPORTED = -2,  /* cusparse: CUSPARSE_STATUS_MATRIX_TYPE_NOT_SUPPORTED */
  SPMMDB_ERR_CUDA = -3,           /* a CUDA runtime call failed (see spmmdb_last_error)  */
  SPMMDB_ERR_ALLOC = -4,          /* internal workspace allocation failed                */
  SPMMDB_ERR_ARCH = -5,           /* device is not sm_100 (tcgen05 path requested)       */
  SPMMDB_ERR_IO = -6              /* graph file missing, short or malformed              */
} spmmdb_status;

#define SPMMDB_ROW_MAJOR 0
#define SPMMDB_COL_MAJOR 1

#define SPMMDB_DIR_ROW 0     /* CUSPARSE_DIRECTION_ROW: row-major bs x bs blocks    */
#define SPMMDB_DIR_COLUMN 1  /* CUSPARSE_DIRECTION_COLUMN: column-major blocks      */

#define SPMMDB_MATH_FP32 0   /* CUDA-core fp32 FMA, fp32 accumulate (exact path)    */
#define SPMMDB_MATH_TF32 1   /* tcgen05 kind::tf32, fp32 accumulate in TMEM         */
#define SPMMDB_MATH_BF16 2   /* tcgen05 kind::f16 (bf16 inputs), fp32 accumulate    */
#define SPMMDB_MATH_BF16_PACKED 3 /* as _BF16, but bsr_val already points at bf16 blocks
                                     made once by spmmdb_bsr_pack_bf16 (same block order and
                                     direction): half the A stream per call, no per-call
                                     conversion.  Unit adjacency weights are exact in bf16. */

/* Library / build information. */
const char* spmmdb_version(void);
const char* spmmdb_last_error(void);          /* text of the last CUDA error seen */
int spmmdb_device_sm(void);                   /* e.g. 100 on B200; <0 = no device */

/* Concurrency.  Every compute entry point is asynchronous on the caller's stream.  Library scratch (long-row
 * queue, partial sums, layout copies, conversion counters) is kept per (device, stream): calls on one stream
 * are ordered by the stream; calls on different streams or devices share nothing and may be in flight together
 * (at most 8 streams per device hold scratch at a time, the least recently used one is synchronised and
 * released).  Host threads may call concurrently as long as they do not issue on the SAME stream at once.
 *
 * Tuning knobs (experiments, forced kernel paths in the tests): value of spmmdb_set_tuning(name, v) if set,
 * else the environment variable SPMMDB_<name>, else the built-in default.  INT_MIN as value clears an
 * override.  Names are listed in DESIGN.md ("Environment knobs"); every variant produces results inside the
 * documented tolerance, most of them the same bits. */
int spmmdb_set_tuning(const char* name, int value);
int spmmdb_get_tuning(const char* name, int dflt);

/* --------------------------------------------------------------------------
 * CSR x dense:  C = alpha * A * B + beta * C
 * Replaces gespmm_csrmm<float> / spmmWrapper<float> (gespmm_csrmm.h:379-426:
 * row-major B and C, alpha = 1, beta = 0) and the cuSPARSE csrmm / csrmm2 call
 * sites (run_csrmm.cu:133-142, check_result.cu:211-216: column-major C, B
 * column-major for op(B)=N or row-major for op(B)=T).
 *   m        rows of A (and C)           k   columns of A = rows of B
 *   K        dense width (B_ncols / dim) nnz rowptr[m], or -1 if unknown
 *   b_layout / c_layout  SPMMDB_ROW_MAJOR (ld >= K) or SPMMDB_COL_MAJOR
 *   With row-major B/C, alpha == 1, beta == 0 each output element is the fp32
 *   FMA chain in CSR order, i.e. bit-identical to the reference kernel, for
 *   every row of at most spmmdb_csrmm_segment_for(m, nnz) non-zeros; longer rows are
 *   summed per fixed-size segment and the segment sums added in order
 *   (deterministic, independent of how rows are sharded over GPUs).
 * ------------------------------------------------------------------------ */
int spmmdb_csrmm_f32(int m, int k, int K, int64_t nnz, const int* rowptr, const int* colind,
                     const float* val, const float* B, int64_t ldb, int b_layout, float* C,
                     int64_t ldc, int c_layout, float alpha, float beta, void* stream);

/* --------------------------------------------------------------------------
 * Extended CSR entry (row-major B and C only): the operand and cache-residency levers BASELINE.json's
 * north_star names for this bandwidth-bound op, next to the drop-in above.
 *   b_dtype     SPMMDB_DTYPE_F32, or SPMMDB_DTYPE_BF16: B holds bf16 rows (made once per feature matrix by
 *               spmmdb_dense_pack_bf16; ldb in elements, a multiple of 8, K % 8 == 0, K <= 1024).  The B-row
 *               gathers -- nnz * K * 4 bytes through L2, what bounds the op on graphs -- are halved; A's values,
 *               the accumulation and C stay fp32.  Error against the fp32 result <= 2^-9 relative per
 *               product (stated tolerance 1e-2, measured ~2e-3 relative to sum |a||b|).
 *   hot_bitmap  NULL, or k bits from spmmdb_csr_hot_columns: B rows of marked columns are loaded with an L2
 *               evict_last policy, all others evict_first, so the most re-read rows stay resident instead of
 *               being flushed by rows that are read once.  Same bits as without it.
 * spmmdb_csr_hot_columns marks the most frequent columns of `colind`, as many as fit `max_hot_columns`
 * (e.g. 60 % of the L2 size / bytes per B row); whole frequency classes only, so fewer may be marked.
 * One pass over colind per graph (0.3 ms for 62 M non-zeros); hot_bitmap holds (k + 31) / 32 words;
 * n_hot_host (HOST, may be NULL; non-NULL synchronises) receives the number marked.
 * ------------------------------------------------------------------------ */
#define SPMMDB_DTYPE_F32 0
#define SPMMDB_DTYPE_BF16 1
int spmmdb_csrmm_ex_f32(int m, int k, int K, int64_t nnz, const int* rowptr, const int* colind, const float* val,
                        const void* B, int b_dtype, int64_t ldb, float* C, int64_t ldc, float alpha, float beta,
                        const unsigned* hot_bitmap, void* stream);
int spmmdb_csr_hot_columns(int k, int64_t nnz, const int* colind, int64_t max_hot_columns, unsigned* hot_bitmap,
                           int* n_hot_host, void* stream);
int spmmdb_dense_pack_bf16(int64_t rows, int K, const float* in, int64_t ldi, void* out_bf16, int64_t ldo,
                           void* stream);

/* Row-segment length: rows of at most this many non-zeros are one sequential FMA chain.
 *   spmmdb_csrmm_segment(nnz)         lower bound that depends on nnz alone (64 / 256 / 512)
 *   spmmdb_csrmm_segment_for(m, nnz)  the length a call with these sizes uses: the above, raised to
 *                                     the power of two >= 4 * average row length
 *   spmmdb_set_csrmm_segment(seg)     process-wide override (power of two >= 32; 0 = automatic).  A row-
 *                                     sharded run sets spmmdb_csrmm_segment_for(m, nnz) of the WHOLE matrix
 *                                     on every rank, so that a row is cut the same way on any shard and
 *                                     the concatenated shards equal the single-GPU result bit for bit. */
int spmmdb_csrmm_segment(int64_t nnz);
int spmmdb_csrmm_segment_for(int m, int64_t nnz);
int spmmdb_set_csrmm_segment(int seg);  /* SPMMDB_OK or SPMMDB_ERR_INVALID_VALUE */
int spmmdb_get_csrmm_segment(void);     /* the override, 0 = automatic */

/* --------------------------------------------------------------------------
 * BSR x dense:  C = alpha * A_bsr * op(B) + beta * C
 * Replaces rocsparse_bsrmm_template<float> (rocsparse_bsrmm.h:102-256) and the
 * cusparseSbsrmm call sites (run_bsrmm.cu:161, check_result.cu:223-228).
 *   dir      SPMMDB_DIR_ROW / SPMMDB_DIR_COLUMN (block storage)
 *   trans_b  0: B column-major, ldb >= kb*bs;  1: B row-major [k][ldb], ldb >= n
 *   C        column-major (ldc >= mb*bs) like the reference, or row-major
 *            [mb*bs][ldc] with c_layout = SPMMDB_ROW_MAJOR (extension)
 *   math     SPMMDB_MATH_FP32 (CUDA cores), _TF32 or _BF16 (tcgen05 tensor
 *            cores: block_dim 16 or 32, and 64 with bf16; every other shape runs the fp32 kernel)
 * Returns the mathematical product -- the reference kernel's output-column
 * bug (rocsparse_bsrmm_impl.h:27,326,413) is not reproduced.
 * ------------------------------------------------------------------------ */
int spmmdb_bsrmm_f32(int dir, int trans_b, int mb, int n, int kb, int64_t nnzb, int block_dim,
                     float alpha, const float* bsr_val, const int* bsr_rowptr,
                     const int* bsr_colind, const float* B, int64_t ldb, float beta, float* C,
                     int64_t ldc, int c_layout, int math, void* stream);

/* fp32 blocks -> bf16 blocks for SPMMDB_MATH_BF16_PACKED (nnzb * block_dim^2 values, round to nearest even);
 * out_bf16 holds 2 bytes per value, both pointers 16-byte aligned.  With _PACKED the tensor-core shapes
 * (block_dim 16 / 32 / 64) are the only ones served: others return SPMMDB_ERR_NOT_SUPPORTED. */
int spmmdb_bsr_pack_bf16(int64_t nnzb, int block_dim, const float* bsr_val, void* out_bf16, void* stream);

/* Every block transposed (column-major blocks -> row-major blocks or back), out != bsr_val.  The tf32 tensor-core
 * kernel reads row-major blocks in place; column-major blocks (CUSPARSE_DIRECTION_COLUMN, the block_cublas layout) cost
 * it a transposed copy of all blocks per call.  A caller that multiplies the same matrix many times converts once and
 * calls with SPMMDB_DIR_ROW (the grouped block GEMM on C4: 2.1 -> 0.85 ms per call). */
int spmmdb_bsr_transpose_blocks_f32(int64_t nnzb, int block_dim, const float* bsr_val, float* out, void* stream);

/* Process-wide default arithmetic of the drop-in rocsparse_bsrmm_template<>
 * (its signature has no `math` argument).  Initial value: the SPMMDB_BSRMM_MATH
 * environment variable ("fp32" | "tf32" | "bf16"), else SPMMDB_MATH_TF32: the tensor-core path is THE BSR
 * path for block_dim 16 / 32 (relative error <= 1e-2 stated by BASELINE.json, measured 5e-4; the reference's
 * own check is an absolute 1e-4 on O(1) values, check_result.cu:233-246, which tf32 does not meet: callers that
 * need it select fp32).  Shapes the tensor-core kernel does not take run the fp32 kernel. */
int spmmdb_set_bsrmm_math(int math);  /* SPMMDB_OK or SPMMDB_ERR_INVALID_VALUE */
int spmmdb_get_bsrmm_math(void);

/* --------------------------------------------------------------------------
 * Per-block GEMM formulation of BSR SpMM (block_cublas.cu:123-136,
 * test_block_cublas.cu:273-284): z(col-major, ld m) += A_blk(column-major
 * bs x bs) * Y[col*bs.., :](row-major, ld dim) for every non-zero block, as ONE
 * grouped launch instead of nnzb cublasSgemm calls.  alpha = beta = 1.
 * ------------------------------------------------------------------------ */
int spmmdb_block_gemm_f32(int mb, int dim, int64_t nnzb, int block_dim, const int* bsr_rowptr,
                          const int* bsr_colind, const float* bsr_val, const float* Y, float* Z,
                          int math, void* stream);

/* --------------------------------------------------------------------------
 * csr2bsr, phase 1: replaces cusparseXcsr2bsrNnz (run_bsrmm.cu:121,
 * check_result.cu:173, csr2bsr.cu:176).  Writes bsr_rowptr[mb+1] (device) and
 * returns nnzb through a HOST pointer; synchronises `stream` to do so.
 * Column indices must be ascending inside each row (load_data.cc:179-182).
 * ------------------------------------------------------------------------ */
int spmmdb_csr2bsr_nnz(int m, int n, const int* rowptr, const int* colind, int block_dim,
                       int* bsr_rowptr, int* nnzb_host, void* stream);

/* csr2bsr, phase 2: replaces cusparseScsr2bsr (run_bsrmm.cu:139).  bsr_colind
 * ascending per block row, blocks zero filled, ragged edges zero padded. */
int spmmdb_csr2bsr_f32(int dir, int m, int n, const int* rowptr, const int* colind,
                       const float* val, int block_dim, const int* bsr_rowptr, int* bsr_colind,
                       float* bsr_val, void* stream);

/* bsr2csr: replaces cusparseSbsr2csr (bsr2csr.cu:186-188).  Every block
 * expands to bs*bs explicit entries (zeros kept): nnz = nnzb*bs*bs. */
int spmmdb_bsr2csr_f32(int dir, int mb, int nb, const int* bsr_rowptr, const int* bsr_colind,
                       const float* bsr_val, int block_dim, int* csr_rowptr, int* csr_colind,
                       float* csr_val, void* stream);

/* --------------------------------------------------------------------------
 * Hybrid dense-block split: replaces divide_matrix (divide.cu:52-127), the CPU
 * pre-pass of the reference's `divide` driver.  Per block row, every column
 * block whose fill (float)(count * 1.0 / bs^2) >= density becomes a BSR block
 * (ascending block columns, zero filled); every other non-zero stays, in its
 * original order, in a remainder CSR matrix of the same shape.  A = A_bsr +
 * A_csr, so  C = A_csr * B  then  C += A_bsr * B  (divide.cu:348-373: both calls
 * with beta = 1 into a zeroed C) is the whole product, with the dense part on
 * the tensor-core BSR path and the rest on the CSR path.
 *   val      NULL = unit weights (the reference stores 1.0f, divide.cu:113,230)
 *   density  empty blocks never qualify (the reference emits them for density <= 0)
 * Two phases like csr2bsr: _nnz writes both row-pointer arrays (device) and
 * returns nnzb / the remainder nnz through HOST pointers (synchronises), the
 * caller allocates, _f32 fills.  spmmdb_divide_threshold = the smallest count
 * per block that qualifies (host function, no CUDA).
 * ------------------------------------------------------------------------ */
int spmmdb_divide_threshold(int block_dim, float density);
int spmmdb_divide_nnz(int m, int n, const int* rowptr, const int* colind, int block_dim, float density,
                      int* bsr_rowptr, int* csr_rowptr, int* nnzb_host, int* csr_nnz_host, void* stream);
int spmmdb_divide_f32(int dir, int m, int n, const int* rowptr, const int* colind, const float* val,
                      int block_dim, float density, const int* bsr_rowptr, int* bsr_colind, float* bsr_val,
                      const int* csr_rowptr, int* csr_colind, float* csr_val, void* stream);

/* The hybrid product as one object -- the BSR entry to use on graph adjacency matrices (at the 1-3 % block fill a
 * reordering leaves, plain BSR multiplies mostly zeros; the split sends the dense blocks to the tensor cores and the
 * rest to the CSR kernel).  What the reference's `divide` program does around divide_matrix (divide.cu:205-391):
 *   spmmdb_hybrid_create     splits A (device CSR, val NULL = unit weights) once and keeps both parts in device memory
 *                            owned by the handle; synchronises `stream`
 *   spmmdb_hybrid_spmm_f32   C = alpha * A * B + beta * C as  C = alpha * A_csr * B + beta * C;  C += alpha * A_bsr * B
 *                            (divide.cu:348-373).  Row-major B [n][ldb] and C [m][ldc]; m and n need NOT be multiples of
 *                            block_dim and neither B nor C is padded (the reference drivers pad both).  math: SPMMDB_MATH_*
 *                            of the block part.
 *   spmmdb_hybrid_info       number of dense blocks and of remainder entries */
typedef struct spmmdb_hybrid spmmdb_hybrid;
int spmmdb_hybrid_create(int m, int n, const int* rowptr, const int* colind, const float* val, int block_dim,
                         float density, spmmdb_hybrid** out, void* stream);
int spmmdb_hybrid_info(const spmmdb_hybrid* h, int64_t* nnzb, int64_t* csr_nnz);
int spmmdb_hybrid_spmm_f32(const spmmdb_hybrid* h, int K, const float* B, int64_t ldb, float* C, int64_t ldc,
                           float alpha, float beta, int math, void* stream);
int spmmdb_hybrid_destroy(spmmdb_hybrid* h);

/* --------------------------------------------------------------------------
 * Apply a node permutation on the device (the reordering algorithms themselves stay on
 * the CPU and only produce old2new; one new id per old id, a bijection of 0..n-1).
 *   spmmdb_permute_csr_f32   A' = P A P^T, replaces permutate + sortNeighbors
 *                            (reorder_strategy.cc:15-39): column j -> old2new[j], row i -> position
 *                            old2new[i], every row re-sorted ascending.  val / new_val may both be NULL
 *                            (the reference's adjacency lists carry no values); values follow their
 *                            entries.  Writes new_rowptr[n+1], new_colind[nnz]; synchronises `stream`.
 *   spmmdb_permute_rows_f32  out[old2new[r]][0..K) = in[r][0..K): the feature rows of B (or C back
 *                            with the inverse permutation) in the new numbering; out != in.
 *   spmmdb_heatmap           heat[(i / bs) * nb + j / bs] += 1 per non-zero, nb = ceil(n / bs),
 *                            heat cleared first: getHeatmap (utility.cc:71-88; the reference uses 256).
 * Block counts / density / utilisation (calculateNnzb utility.cc:47-69, reorder_graph.cc:12-24)
 * are spmmdb_csr2bsr_nnz's nnzb:  density = nnzb / nb^2, utilisation = nnz / (nnzb * bs^2).
 * ------------------------------------------------------------------------ */
int spmmdb_permute_csr_f32(int n, const int* rowptr, const int* colind, const float* val,
                           const int* old2new, int* new_rowptr, int* new_colind, float* new_val,
                           void* stream);
int spmmdb_permute_rows_f32(int rows, int K, const float* in, int64_t ldi, const int* old2new,
                            float* out, int64_t ldo, void* stream);
int spmmdb_heatmap(int n, const int* rowptr, const int* colind, int block_dim, int* heat, void* stream);
/* Fill-decile histogram of the non-empty block_dim x block_dim blocks: calculate_block_density_dist
 * (block_density_dist.cpp:47-86).  dist10_host (HOST, 10 entries): dist[d] = blocks whose fill
 * (float)(count * 1.0 / bs^2) lands in decile (int)(fill * 10), 10 folded into 9 -- the reference's expression,
 * bit for bit.  Synchronises `stream`. */
int spmmdb_block_fill_histogram(int m, int n, const int* rowptr, const int* colind, int block_dim,
                                int64_t* dist10_host, void* stream);

/* --------------------------------------------------------------------------
 * The reference's simple orderings produced on the device.  Each writes the permutation
 * old2new[n] (DEVICE) that spmmdb_permute_csr_f32 / spmmdb_permute_rows_f32 apply, and synchronises `stream`.
 *   spmmdb_order_by_degree  maxDegreeSort (reorder_strategy.cc:41-55): descending degree.  The reference leaves
 *                           equal degrees in whatever order std::sort gives; here ties keep ascending id, so the
 *                           permuted graph has the reference's row lengths always, its rows when degrees differ.
 *   spmmdb_order_bfs        BFSTraversal (reorder_strategy.cc:68-98): FIFO breadth-first numbering, started at the
 *                           lowest unvisited id, neighbours in adjacency order (colind need not be sorted).  The
 *                           reference's permutation exactly, for any pattern, symmetric or not.
 * The other front-ends (Gorder, Rabbit, METIS, greedy) stay on the CPU (BASELINE.json north_star).
 * ------------------------------------------------------------------------ */
int spmmdb_order_by_degree(int n, const int* rowptr, int* old2new, void* stream);
int spmmdb_order_bfs(int n, const int* rowptr, const int* colind, int* old2new, void* stream);
/* reverseCuthillMcKee (reorder_strategy.cc:57-66): adjacency lists sorted by descending degree of the neighbour, then
 * BFSTraversal (the reference never reverses the order).  Equal-degree neighbours are taken in ascending id; the
 * reference's std::sort leaves them unspecified (stable for lists of <= 16 entries: on graphs of maximum degree <= 16
 * this is the reference's permutation exactly). */
int spmmdb_order_rcm(int n, const int* rowptr, const int* colind, int* old2new, void* stream);

/* --------------------------------------------------------------------------
 * Host-buffer variant of spmmdb_csrmm_f32 (row-major B and C, ld = K): all
 * pointers are HOST pointers (pinned memory gives full PCIe rate).  Copies A and
 * B to the device, runs the device path, copies C back and synchronises -- the
 * state of the reference drivers before run_csrmm.cu:94-116 / after :161.
 * ------------------------------------------------------------------------ */
int spmmdb_csrmm_host_f32(int m, int k, int K, int64_t nnz, const int* h_rowptr, const int* h_colind,
                          const float* h_val, const float* h_B, float* h_C, float alpha, float beta,
                          void* stream);

/* --------------------------------------------------------------------------
 * Allocation-free CSR x dense:  C = alpha * A * B + beta * C, row-major B (ld >= K) and C (ld >= K).
 * No counterpart in the reference: added for callers that cannot
 * have the library call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The entry point calls no cudaMalloc / cudaFree and no allocating library routine: the scratch
 * its kernels need (long-row list, per-segment partial sums) is one fixed __device__ array inside the
 * library, SPMMDB_NOALLOC_SCRATCH_BYTES per GPU that loads libspmmdb.so, held until the process exits.
 *   limit     K <= SPMMDB_NOALLOC_MAX_K; a larger K returns SPMMDB_ERR_NOT_SUPPORTED and spmmdb_last_error()
 *             names the limit.  m, k and nnz are not limited: long rows beyond the array's capacity
 *             (16384 rows of more than 1024 non-zeros, min(16384, 2^21 / K) segments of 512 non-zeros) are
 *             summed in place, with the same result.
 *   sharing   every call on a device uses the same array and finds the last call's contents in it (the
 *             entry point clears what must start at zero).  Calls on one stream are ordered by the stream;
 *             calls that OVERLAP on the GPU (two streams, two host threads) race on the array and give
 *             wrong sums: serialise them.
 *   result    within 1e-5 of spmmdb_csrmm_f32; rows of <= 1024 non-zeros are the same FMA chain in CSR
 *             order, longer rows are summed per 512-entry segment, segment sums added in order.
 *   val must not be NULL; asynchronous on `stream`.
 * ------------------------------------------------------------------------ */
#define SPMMDB_NOALLOC_MAX_K 512
#define SPMMDB_NOALLOC_SCRATCH_BYTES 8782080
int spmmdb_csrmm_noalloc_f32(int m, int k, int K, const int* rowptr, const int* colind, const float* val,
                             const float* B, int64_t ldb, float* C, int64_t ldc, float alpha, float beta,
                             void* stream);

/* --------------------------------------------------------------------------
 * Multi-GPU row sharding (new behaviour defined by BASELINE.json; the reference
 * is single-GPU).  HOST function, no CUDA call: splits the rows of a CSR (or
 * block rows of a BSR) matrix into `parts` contiguous ranges of ~equal nnz.
 *   h_rowptr  HOST pointer, m + 1 entries
 *   align     boundaries are multiples of `align` rows (1 for CSR; the block size
 *             when a CSR matrix is cut before csr2bsr)
 *   bounds    HOST out, parts + 1 entries: range g = [bounds[g], bounds[g+1]);
 *             boundary g = first row r with rowptr[r] >= g*nnz/parts, rounded down to `align`
 * Output rows are disjoint, so the sharded SpMM needs no reduction; the long-row
 * split of spmmdb_csrmm_f32 depends only on the row and the segment length, so with
 * the same spmmdb_set_csrmm_segment() on every rank concatenated shards are
 * bit-identical to the single-GPU result.
 * ------------------------------------------------------------------------ */
int spmmdb_shard_plan(const int* h_rowptr, int m, int parts, int align, int* bounds);

/* --------------------------------------------------------------------------
 * Multi-GPU execution, one process per GPU (BASELINE.json north_star: nnz-balanced row ranges of A, B replicated
 * by an NCCL all-gather over NVLink, no reduction).  For a square adjacency matrix (m == k): rank g owns rows
 * [bounds[g], bounds[g+1]) of A, of C and of B (bounds from spmmdb_shard_plan, HOST array of world + 1 entries).
 * NCCL is loaded at run time (libnccl.so.2; the copy already in the process if there is one); without it these
 * return SPMMDB_ERR_NOT_SUPPORTED.
 *   spmmdb_mg_unique_id      128-byte ncclUniqueId made on one rank; the caller hands the bytes to every rank
 *                            (MPI, torch.distributed, a file ...)
 *   spmmdb_mg_create         collective: ncclCommInitRank on the current device + the library's streams / events
 *   spmmdb_mg_create_from_comm  adopt the caller's ncclComm_t instead
 *   spmmdb_mg_allgather_rows_f32  all-gather-v, in place, of the row ranges of a full-size row-major B: one NCCL
 *                            group of send / recv pairs (every rank sends to every peer at once)
 *   spmmdb_sharded_csrmm_f32  C_local = alpha * A_local * B + beta * C_local where every rank passes ITS rows of
 *                            B (B_local, [rows][ldb]) and its shard of A (rowptr rebased to 0, global column
 *                            ids).  K is cut into `slabs` feature slabs (0 = default 2); slab s + 1 is gathered
 *                            on a communication stream while the SpMM of slab s runs on `stream`.  The replica of
 *                            B lives in library memory (k * K floats per rank, slab-major).  Same bits as the
 *                            single-GPU call on the whole matrix when every rank has set the whole matrix's segment length
 *                            (spmmdb_set_csrmm_segment(spmmdb_csrmm_segment_for(m, nnz))).
 * ------------------------------------------------------------------------ */
typedef struct spmmdb_mg spmmdb_mg;
int spmmdb_mg_unique_id(void* id128);
int spmmdb_mg_create(int rank, int world, const void* id128, spmmdb_mg** out);
int spmmdb_mg_create_from_comm(void* nccl_comm, int rank, int world, spmmdb_mg** out);
int spmmdb_mg_destroy(spmmdb_mg* g);
int spmmdb_mg_allgather_rows_f32(spmmdb_mg* g, float* B, int64_t ldb, int k, const int* bounds, void* stream);
int spmmdb_sharded_csrmm_f32(spmmdb_mg* g, int k, int K, int64_t nnz_local, const int* rowptr, const int* colind,
                             const float* val, const float* B_local, int64_t ldb, const int* bounds, float* C,
                             int64_t ldc, float alpha, float beta, int slabs, void* stream);

/* --------------------------------------------------------------------------
 * Graph files.  HOST functions, no CUDA call.
 *   SPMMDB_GRAPH_TEXT      <path>_indptr.txt + <path>_indices.txt, the reference's cache format
 *                          (dumpCSRToFile / loadCSRFromFile, load_data.cc:125-165): read, and written
 *                          byte for byte as the reference writes it
 *   SPMMDB_GRAPH_EDGELIST  <path> = "<n> <nnz>" then nnz pairs "x y" (loadGraphFromFile
 *                          load_data.cc:167-184 + convertGraphToCSR utility.cc:31-45: neighbour lists
 *                          sorted ascending, duplicate edges kept, nnz taken from the header); read only
 *   SPMMDB_GRAPH_BINARY    <path>.csrbin: 32-byte header {"SPMMDBG1", int64 rows, int64 cols, int64 nnz},
 *                          int32 rowptr[rows+1], int32 colind[nnz] (little endian)
 *   SPMMDB_GRAPH_AUTO      read only: the binary file if <path>.csrbin exists, else the text pair
 * spmmdb_graph_read returns malloc'ed arrays (release each with spmmdb_graph_free); the text
 * parsers run on all host cores (SPMMDB_IO_THREADS overrides).  Where the reference reads garbage
 * from a missing or short file, these return SPMMDB_ERR_IO.
 * ------------------------------------------------------------------------ */
#define SPMMDB_GRAPH_TEXT 0
#define SPMMDB_GRAPH_BINARY 1
#define SPMMDB_GRAPH_EDGELIST 2
#define SPMMDB_GRAPH_AUTO 3
int spmmdb_graph_read(const char* path, int format, int* n, int64_t* nnz, int** rowptr, int** colind);
int spmmdb_graph_write(const char* prefix, int format, int n, int64_t nnz, const int* rowptr,
                       const int* colind);
void spmmdb_graph_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* SPMMDB_H */
