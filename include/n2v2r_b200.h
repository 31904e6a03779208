/*
 * n2v2r_b200.h -- C ABI of the B200-native UASE embed-and-rank hot path.
 *
 * Drop-in boundary for pmandros/node2vec2rank's hot path.  The
 * reference has no FFI of its own (pure Python); each entry point below replaces the
 * Python function cited next to it.  A maintainer binds these with ctypes (see
 * INTEGRATION.md); node2vec2rank_b200/_native.py is that binding.
 *
 * Conventions
 *   - every pointer named d_* is DEVICE memory owned by the caller; h_* is HOST memory;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream);
 *   - all functions return n2v2r_status_t; n2v2r_last_error() gives the message of the
 *     last failure on the calling thread;
 *   - no function allocates device memory behind the caller's back: scratch space is
 *     passed as (d_ws, ws_bytes) and sized with the matching *_workspace_bytes call;
 *   - functions are asynchronous on `stream` unless a host scalar must be returned
 *     (select_nonzero, uase, planes_needed, is_symmetric), which synchronise the stream.
 *
 * Storage of a graph layer ("planes"): the (rows x cols) matrix is held as P in {1,2,3}
 * bf16 planes, value = plane0 + plane1 + plane2, each plane row-major with leading
 * dimension ld (elements, multiple of 64, padding zero-filled).  P=1 is exact for
 * binarised layers (and any bf16-representable weights), P=2 keeps 16 mantissa bits,
 * P=3 is exact fp32.
 */
#ifndef N2V2R_B200_H
#define N2V2R_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define N2V2R_ABI_VERSION 3

typedef enum {
  N2V2R_OK = 0,
  N2V2R_EINVAL = 1,        /* bad argument                                            */
  N2V2R_ENOMEM = 2,        /* workspace too small                                     */
  N2V2R_ECUDA = 3,         /* CUDA runtime / driver error                             */
  N2V2R_ECOMM = 4,         /* all-reduce callback failed                              */
  N2V2R_ENOTCONV = 5,      /* solver hit max_iters before tol (results still written) */
  N2V2R_EUNSUPPORTED = 6   /* device is not sm_100 / shape outside kernel limits      */
} n2v2r_status_t;

typedef enum { N2V2R_F32 = 0, N2V2R_F64 = 1 } n2v2r_dtype_t;
typedef enum { N2V2R_EUCLIDEAN = 0, N2V2R_COSINE = 1, N2V2R_CORRELATION = 2 } n2v2r_metric_t;
typedef enum { N2V2R_SEQUENTIAL = 0, N2V2R_ONE_VS_BEFORE = 1, N2V2R_ONE_VS_REST = 2 } n2v2r_strategy_t;
typedef enum { N2V2R_ALGO_TCGEN05 = 0, N2V2R_ALGO_SIMT = 1 } n2v2r_algo_t;
/* how a layer is held in HBM: dense bf16 planes; compressed tiles (sparse layers, expanded in shared memory for the tensor
 * cores); ONE e4m3 plane of 1 byte per entry (binarised layers and any layer whose entries e4m3 represents exactly); or row
 * lists (sparse layers, exact fp32 values, multiplied entry by entry on the FMA pipes: "K1d row lists" below) */
typedef enum { N2V2R_STORE_PLANES = 0, N2V2R_STORE_TILES = 1, N2V2R_STORE_E4M3 = 2, N2V2R_STORE_ROWS = 3 } n2v2r_storage_t;

int n2v2r_abi_version(void);
const char* n2v2r_last_error(void);
/* sm count, compute capability (major*10+minor) and total memory of `device`. */
int n2v2r_device_info(int device, int* sm_count, int* cc, size_t* total_mem);

/* ------------------------------------------------------------------------------------
 * K0  synthetic stochastic block model layer (benchmark input).
 * Counter-based: entry (i,j), i<j, of layer `layer` is a pure function of
 * (seed, layer, i, j); A[j,i] = A[i,j]; zero diagonal.  community(i) = i % K.
 * h_B is the K*K row-major (symmetric) block probability matrix on the HOST.
 * weighted=0: A in {0,1}; weighted=1: A = edge * w, w in (0,1] a second counter draw with 16 bits ((k+1)/65536: two bf16
 * planes are exact); weighted=2: 24-bit weights fp32(u') + 2^-24 (three planes).
 * NumPy twin: node2vec2rank_b200/synthetic.py (bit-identical).
 * ---------------------------------------------------------------------------------- */
int n2v2r_sbm_generate(void* stream, float* d_out, int64_t n, int64_t ld, int K,
                       const float* h_B, uint64_t seed, int layer, int weighted);

/* ------------------------------------------------------------------------------------
 * K1a  order statistics for the percentile cut.
 * Replaces np.percentile(x[x.nonzero()], q) at preprocessing_utils.py:134.
 * Considers y = transform(x) (abs if `absolute`, then y < thr -> 0 if has_thr) and only
 * y != 0.  Returns the number of nonzeros and, for position pos = frac*(nnz-1) (evaluated in
 * float32 when index_f32 != 0, as NumPy does for float32 data; returned in *h_pos), the two
 * adjacent order statistics y_(floor(pos)) and y_(floor(pos)+1) (ascending, 0-based),
 * exactly (multi-pass radix select on the order-preserving bit pattern).  The caller does
 * the linear interpolation in the input dtype like NumPy does.  Synchronises `stream`.
 * d_in must be contiguous (ld == cols) and 16-byte aligned (the passes read 16-byte vectors).
 * ---------------------------------------------------------------------------------- */
size_t n2v2r_select_workspace_bytes(void);
int n2v2r_select_nonzero(void* stream, const void* d_in, int dtype, int64_t rows, int64_t cols,
                         int64_t ld, int absolute, int has_thr, double thr, double frac, int index_f32,
                         void* d_ws, size_t ws_bytes,
                         int64_t* h_nnz, double* h_lo, double* h_hi, double* h_pos);

/* ------------------------------------------------------------------------------------
 * K1b  fused load pass.  Replaces preprocessing_utils.py:124-141 (abs -> threshold ->
 * percentile cut -> binarize -> fp32) and, as a by-product, the column sums that
 * model.py:293-294 (degree_difference_ranking) needs.
 * Any of d_out_f32 / d_planes / d_colsum may be NULL.  d_planes holds n_planes planes of
 * rows*ld_p bf16 each (plane stride rows*ld_p); padding columns [cols, ld_p) are zeroed.
 * d_colsum (cols doubles) must be zeroed by the caller; sums are of the fp32 output values.
 * Fast path (4 columns per thread, cp.async ring): cols % 4 == 0, ld_in % 4 == 0 (fp32) and 16-byte
 * aligned d_in / d_out_f32; any other layout runs the general 2-column kernel.
 * ---------------------------------------------------------------------------------- */
int n2v2r_transform(void* stream, const void* d_in, int dtype, int64_t rows, int64_t cols, int64_t ld_in,
                    int absolute, int has_thr, double thr, int has_cut, double cut, int binarize,
                    float* d_out_f32, int64_t ld_out,
                    void* d_planes, int n_planes, int64_t ld_p,
                    double* d_colsum);

/* Minimum number of bf16 planes (1..3) that represents every entry of the fp32 matrix
 * exactly.  Synchronises `stream`. */
int n2v2r_planes_needed(void* stream, const float* d_in, int64_t rows, int64_t cols, int64_t ld,
                        int* h_planes);
/* One bf16 plane -> e4m3 bytes (rows x ld8, ld8 a multiple of 128, padding zero): the N2V2R_STORE_E4M3 form of a layer.
 * *h_exact = 1 when every value is representable (then the layer is exact); synchronises `stream`. */
int n2v2r_plane_to_e4m3(void* stream, const void* d_plane, int64_t rows, int64_t cols, int64_t ld_p,
                        void* d_out, int64_t ld8, int* h_exact);
/* 1 if the square plane set is symmetric (columns >= n of the padding are ignored).  ld_p must be a
 * multiple of 64 and d_planes 16-byte aligned.  Synchronises `stream`. */
int n2v2r_is_symmetric(void* stream, const void* d_planes, int n_planes, int64_t n, int64_t ld_p,
                       int* h_symmetric);
/* out = in^T for every plane (used only for non-symmetric layers); padding columns of out are written as
 * zeros.  ld_p must be a multiple of 64, both pointers 16-byte aligned. */
int n2v2r_transpose_planes(void* stream, const void* d_planes, int n_planes, int64_t n, int64_t ld_p,
                           void* d_out);

/* ------------------------------------------------------------------------------------
 * K1c  compressed tiles: the storage of a SPARSE layer (after the threshold / top-percent cut of
 * preprocessing_utils.py:127-135 most entries are zero).  Plays the role of the reference's dense ->
 * csc_matrix conversion (N2V2R.__fit, model.py:51-53), in a form the tensor-core product can expand in
 * shared memory.  The (n x n) matrix is cut into tiles of tile_rows x 64 (tile_rows = 128 or 256; tile
 * (rt, kb) has index rt * ceil(n/64) + kb); a tile is a list of 8-byte entries
 *     { u16 pos, bf16 p0, bf16 p1, bf16 p2 },   value = p0 + p1 + p2 (exact fp32),
 * pos = element offset of (r, c) inside a K-major SWIZZLE_128B bf16 tile:
 *     pos = (r>>3)*512 + (r&7)*64 + (((c>>3) ^ (r&7))<<3) + (c&7);
 * lists are padded to a multiple of 4 entries with pos = 0xFFFF and hold their entries in the order the product kernel's
 * converter warps want (shared-memory-bank interleaved); d_tile_ptr[t] is the first entry of tile t,
 * d_tile_ptr[n_tiles] the total.  Build: count (non-zeros per 128-row tile of A and of A^T, planes the values
 * need, symmetry; synchronises) -> scan (tile_rows chosen by the caller; synchronises, returns the number of
 * entries to allocate) -> fill (transpose=1 writes the tiles of A^T from the same dense input).
 * ---------------------------------------------------------------------------------- */
int n2v2r_tiles_count(void* stream, const float* d_in, int64_t n, int64_t ld,
                      uint32_t* d_counts /* ceil(n/128) * ceil(n/64) */, uint32_t* d_counts_t /* same, for A^T */,
                      int* h_planes, int* h_symmetric);
int n2v2r_tiles_scan(void* stream, const uint32_t* d_counts, int64_t n, int tile_rows,
                     uint32_t* d_tile_ptr /* ceil(n/tile_rows) * ceil(n/64) + 1 */, int64_t* h_entries);
int n2v2r_tiles_fill(void* stream, const float* d_in, int64_t n, int64_t ld, int transpose, int tile_rows,
                     const uint32_t* d_tile_ptr, void* d_entries /* 16-byte aligned, h_entries * 8 bytes */);
/* d_out (n x ld_out fp32, zero-filled by the caller) = the matrix the tiles hold */
int n2v2r_tiles_to_dense(void* stream, const uint32_t* d_tile_ptr, const void* d_entries, int64_t n, int tile_rows,
                         float* d_out, int64_t ld_out);

/* ------------------------------------------------------------------------------------
 * K2  bipartite -> unipartite projection on the tensor cores.
 * Replaces bipartite_to_unipartite_projection (preprocessing_utils.py:16-32):
 * on_rows=0: out = W^T W (cols x cols); on_rows=1: out = W W^T (rows x rows).
 * W is (rows x cols) row-major, dtype F32 or F64; operands are split into three bf16
 * terms (24 mantissa bits) and accumulated in fp32 (tcgen05.mma, TMA-fed) in K chunks of
 * 512 that are summed in fixed order.  Only tiles touching the upper triangle are computed
 * and mirrored: the output is exactly symmetric.  Error <= ~4e-6 of the largest entry.
 * ---------------------------------------------------------------------------------- */
size_t n2v2r_project_workspace_bytes(int64_t rows, int64_t cols, int on_rows);
int n2v2r_project(void* stream, const void* d_W, int dtype, int64_t rows, int64_t cols, int64_t ld,
                  int on_rows, float* d_out, int64_t ld_out, void* d_ws, size_t ws_bytes);

/* ------------------------------------------------------------------------------------
 * K1d  row lists: the storage of a SPARSE layer for the FMA-pipe product (k3_rows.cu).  The rows are cut into blocks of
 * rows_per_block (one CTA each; <= 528), the columns into panels of panel_cols (multiple of 32).  The non-zeros of
 * (block b, panel p) form one list: a u16 header of rows_per_block + 1 offsets counted in PAIRS of entries (padded to 16
 * bytes; a row with an odd number of non-zeros ends with the padding entry {0xFFFFFFFF, 0}) followed by the
 * 8-byte entries {u32 column inside the panel, fp32 value} sorted by row; inside a row the columns alternate in parity
 * (even, odd, even, ... when floor(row_in_block / ceil(rows_per_block / 176)) is even; odd, even, odd, ... otherwise;
 * ascending inside each parity class, the surplus of the larger class last: the order keeps the gathers of paired lane
 * groups in opposite shared-memory bank halves); padded to 16 bytes.
 * d_list_ptr[b n_panels + p] is the offset of the list in 16-byte units, d_list_ptr[n_blocks n_panels] the total.
 * Build: count (per-row counts, list offsets; returns the total size in 16-byte units, the longest list in bytes and the
 * number of non-zeros; synchronises) -> allocate -> fill.  The lists of A^T are built from a transposed copy of the input.
 * ---------------------------------------------------------------------------------- */
size_t n2v2r_rows_counts_bytes(int64_t n, int rows_per_block, int panel_cols);
int n2v2r_rows_count(void* stream, const float* d_in, int64_t n, int64_t ld, int rows_per_block, int panel_cols,
                     void* d_counts, uint32_t* d_list_ptr, int64_t* h_units, int32_t* h_max_list_bytes, int64_t* h_nnz);
int n2v2r_rows_fill(void* stream, const float* d_in, int64_t n, int64_t ld, int rows_per_block, int panel_cols,
                    const void* d_counts, const uint32_t* d_list_ptr, void* d_lists);
/* d_out (n x ld_out fp32, zero-filled by the caller) = the matrix the lists hold */
int n2v2r_rows_to_dense(void* stream, const uint32_t* d_list_ptr, const void* d_lists, int64_t n, int rows_per_block,
                        int panel_cols, float* d_out, int64_t ld_out);

/* ------------------------------------------------------------------------------------
 * K3-K8  UASE: truncated SVD of the unfolded matrix [A_1 | ... | A_T] and the per-layer
 * right embeddings V_t * S^(1/2).  Replaces spectral_embedding.UASE as called at
 * model.py:54-55 (scipy svds / ARPACK underneath).
 * Chebyshev-filtered block subspace iteration with Rayleigh-Ritz, locking and deflation;
 * CholeskyQR2; single-CTA Jacobi for the l x l eigenproblem; the only O(n^2) work is the
 * streaming skinny products Y += A_t X and Z_t = A_t^T Q.
 * Layer sharding: this process owns T_local layers [layer_offset, layer_offset+T_local)
 * of T_global; `allreduce` (may be NULL when T_local == T_global) sums a device buffer
 * over all processes, in place, ordered on `stream`.
 * ---------------------------------------------------------------------------------- */
typedef struct {
  const void* d_planes;    /* PLANES: n_planes planes, each n x ld bf16; E4M3: one plane n x ld bytes (ld % 128 == 0) */
  const void* d_planes_t;  /* planes of A^T, or NULL when the layer is symmetric         */
  int32_t n_planes;        /* planes stored (PLANES) / plane terms the values need (TILES) / unused, 0 (ROWS: fp32 values) */
  int32_t storage;         /* n2v2r_storage_t                                            */
  int64_t ld;
  /* N2V2R_STORE_TILES (see "K1c compressed tiles" below) */
  const uint32_t* d_tile_ptr;    /* n_tiles + 1 entry offsets                            */
  const void* d_entries;         /* 8-byte entries                                       */
  const uint32_t* d_tile_ptr_t;  /* tiles of A^T, or NULL when the layer is symmetric    */
  const void* d_entries_t;
  int64_t n_entries, n_entries_t;
  int32_t tile_rows;             /* TILES: 128 or 256; ROWS: rows per block                */
  int32_t panel_cols;            /* ROWS: columns per panel (ABI 2: reserved)              */
  /* ROW SLAB of a layer (bf16 planes only; sharding of ONE layer over several processes when there are fewer layers
   * than GPUs): rows > 0 means this process holds rows [row_offset, row_offset + rows) of the n x n layer:
   * d_planes = the slab (rows x ld), d_planes_t = its transpose (n x ld_t, ld_t a multiple of 64 >= rows; required). */
  int64_t row_offset, rows, ld_t;
  /* N2V2R_STORE_ROWS (ABI 3): d_tile_ptr / d_tile_ptr_t = the list offsets (16-byte units), d_entries / d_entries_t = the
   * lists, n_entries / n_entries_t = non-zeros; the longest list of A and of A^T in bytes (n2v2r_rows_count) */
  int32_t list_cap, list_cap_t;
} n2v2r_layer_t;

typedef struct {
  int64_t n;               /* nodes                                                      */
  int32_t T_local, T_global, layer_offset;
  int32_t k;               /* embedding dimension (max of the sweep), k <= l             */
  int32_t l;               /* block width: 16, 32, 48, 64 or 80                          */
  int32_t max_iters;       /* cap on operator applications (A A^T)                       */
  int32_t min_iters;
  int32_t cheb_degree;     /* cap on the Chebyshev filter degree per outer iteration (0 = default 32,
                              1 = plain subspace iteration)                                */
  int32_t algo;            /* n2v2r_algo_t                                               */
  int32_t k_chunk;         /* columns of A per accumulation unit (0 = default)           */
  int32_t n_check;         /* number of entries of h_check_dims                          */
  double tol;              /* stop when the subspace-angle bound at every checked d <= tol */
  uint64_t seed;
  const int32_t* h_check_dims; /* swept dimensions d (gap-aware stopping rule); may be NULL */
  /* solver tuning; 0 selects the default in every field (ABI 2: these replace the former environment variables) */
  double floor_mult;       /* residual floor as a multiple of eps32 * lambda_1 (default 0.02) */
  double rel_floor;        /* residual target relative to lambda_j (default 8 eps32)     */
  int32_t first_degree;    /* cap on the degree of the first filter (default 2)          */
  int32_t no_early_lock;   /* 1 = do not lock outliers early                             */
  int32_t trace;           /* 1 = one line per outer iteration on stderr                 */
  int32_t comm_mode;       /* how Y is summed over the processes: 0 = `allreduce` callback, 1 = `user` is an
                              ncclComm_t and the library calls ncclAllReduce itself on `stream` */
  void* group_comm;        /* row slabs: communicator (n2v2r_comm_init) of the processes that hold slabs of the SAME layers
                              as this one; Z_t = sum over the slabs of slab^T Q[slab rows] is all-reduced over it        */
  void* batch;             /* seed batching: handle of n2v2r_batch_create shared by concurrent solves, or NULL */
  int32_t batch_slot;      /* this solve's slot in the batch, 0 .. participants-1                              */
  int32_t reserved2;
} n2v2r_uase_params_t;

#define N2V2R_MAX_CHECK 32
typedef struct {
  int32_t iterations;      /* applications of the operator A A^T to the l-column block      */
  int32_t layer_passes;    /* number of (layer x skinny product) launches                */
  int32_t converged;       /* 0 = max_iters reached; 1 = the angle bound met tol at every checked d, or every wanted
                              residual is at its target (ARPACK's tol = 0 criterion); 2 = stopped because the residuals
                              stagnated above the target (the fp32 floor of this matrix)                              */
  int32_t kernel_launches;
  int32_t outer_iterations;/* Rayleigh-Ritz steps                                           */
  int32_t locked;          /* Ritz pairs at the residual floor when the solver stopped      */
  double stream_ms;        /* CUDA-event time of the streaming products (split + MMA + reduce) */
  double mma_ms;           /* CUDA-event time of the tcgen05 MMA kernel launches alone      */
  int32_t mma_launches;
  int32_t bound_met;       /* 1 when the angle bound is <= tol at every checked d           */
  double total_ms;
  double stream_bytes;     /* algorithmic bytes of those launches            */
  double max_angle_bound;  /* final value of the stopping quantity                       */
  double resid[96];        /* final residual norm per Ritz pair, relative to lambda_1       */
  double ritz[96];         /* Ritz values (sigma^2), descending                          */
  double angle_bounds[N2V2R_MAX_CHECK];  /* final Davis-Kahan bound per checked d (order of h_check_dims) */
} n2v2r_uase_stats_t;

typedef int (*n2v2r_allreduce_fn)(void* user, void* d_buf, int64_t count, int dtype, void* stream);

/* NCCL communicator owned by the library (comm_mode = 1 of n2v2r_uase): the library binds the NCCL the process has
 * already loaded (dlopen of libnccl.so.2, no link-time dependency).  One process creates the 128-byte unique id, the
 * host side distributes it (torch.distributed broadcast, MPI, a file ...), every process calls n2v2r_comm_init with
 * its rank; the handle is passed as `user` to n2v2r_uase.  n2v2r_comm_allreduce_f32 sums a device buffer in place. */
int n2v2r_comm_unique_id(void* h_id128);
int n2v2r_comm_init(const void* h_id128, int n_ranks, int rank, void** comm);
int n2v2r_comm_destroy(void* comm);
int n2v2r_comm_allreduce_f32(void* stream, void* comm, float* d_buf, int64_t count);

size_t n2v2r_uase_workspace_bytes(const n2v2r_uase_params_t* p);
int n2v2r_uase(void* stream, const n2v2r_uase_params_t* p, const n2v2r_layer_t* h_layers,
               void* d_ws, size_t ws_bytes, n2v2r_allreduce_fn allreduce, void* user,
               float* d_embed /* T_local x n x k */, float* d_left /* n x k or NULL */,
               double* h_sigma /* k */, n2v2r_uase_stats_t* h_stats);

/* ------------------------------------------------------------------------------------
 * Seed batching (stability runs: README.md:54, :173; BASELINE configs[4]).  `participants` solves on the same resident,
 * unsharded layers run concurrently -- one host thread and one stream each, each calling n2v2r_uase with params.batch =
 * the handle and its own params.batch_slot, seed, workspace and outputs -- and every product M X they need is served for
 * all waiting solves with ONE pass over each layer per up-to-three operands (bf16-plane layers; other storages are served
 * one by one).  Every solve returns exactly the bits of the unbatched solve.  The handle owns a stream and events; d_ws is
 * the batch's own operand / partial-sum workspace.  n2v2r_batch_destroy synchronises and reports
 * {products served, layer passes, MMA milliseconds, MMA launches}.
 * ---------------------------------------------------------------------------------- */
size_t n2v2r_batch_workspace_bytes(const n2v2r_uase_params_t* p, int participants);
int n2v2r_batch_create(const n2v2r_uase_params_t* p, const n2v2r_layer_t* h_layers, int participants,
                       void* d_ws, size_t ws_bytes, void** handle);
int n2v2r_batch_destroy(void* handle, double* h_stats4);

/* One streaming skinny product, exposed for tests / roofline measurement:
 * d_out (n x l fp32) = A X  (transpose=0) or A^T X (transpose=1), X = d_x (n x l fp32). */
size_t n2v2r_apply_workspace_bytes(int64_t n, int l, int k_chunk);
int n2v2r_layer_apply(void* stream, const n2v2r_layer_t* h_layer, int64_t n, int l, int transpose,
                      const float* d_x, float* d_out, int algo, int k_chunk,
                      void* d_ws, size_t ws_bytes);

/* ------------------------------------------------------------------------------------
 * K9  per-node distances for every swept d.  Replaces N2V2R.__rank (model.py:57-96) +
 * compute_pairwise_distances (model_utils.py:39-67).
 * d_embed: (T x n x k), dtype F32 or F64.  Column c uses the leading h_col_dim[c] components and metric
 * h_col_metric[c].  Comparisons follow `strategy`: sequential / one_vs_before give
 * T-1 comparisons (layers 1..T-1), one_vs_rest gives T.  float64 arithmetic.
 * d_scores: (n_comparisons x n_cols x n) doubles, each column contiguous.
 * ---------------------------------------------------------------------------------- */
int n2v2r_distances(void* stream, const void* d_embed, int dtype, int T, int64_t n, int k,
                    const int32_t* h_col_dim, const int32_t* h_col_metric, int n_cols,
                    int strategy, double* d_scores);

/* ------------------------------------------------------------------------------------
 * K10+K11  descending rank of every score column (segmented radix sort on the
 * order-preserving 64-bit pattern; ties stable by node position; NaN last) and Borda
 * points sum_c (n - position_c).  Replaces aggregate_transform (model.py:167-185) +
 * borda_aggregate_parallel/_get_ranking (model_utils.py:22-36).
 * d_scores: (n_cols x n) doubles, n_cols = n_groups * columns per comparison; d_positions:
 * (n_cols x n) int32 or NULL; d_borda: (n_groups x n) int64, one Borda table per group of
 * consecutive columns (all comparisons of a run are ranked by one launch).
 * ---------------------------------------------------------------------------------- */
size_t n2v2r_rank_workspace_bytes(int64_t n, int n_cols);
int n2v2r_rank_borda(void* stream, const double* d_scores, int64_t n, int n_cols, int n_groups,
                     int32_t* d_positions, int64_t* d_borda, void* d_ws, size_t ws_bytes);
/* The same rank + Borda for callers that cannot allocate device memory (code under CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc): no workspace argument, the sort buffers are a static device array inside the library
 * (24 MiB held by every GPU that loads it, until exit), and the call allocates nothing.  No counterpart in the reference.
 * Limit: n * n_cols <= N2V2R_RANK_NOALLOC_MAX_ELEMS, larger shapes are
 * rejected with N2V2R_EINVAL (use n2v2r_rank_borda).  All calls share the one array: calls that overlap on the GPU race --
 * order them on one stream.  Same arguments and results (bit-identical) as n2v2r_rank_borda otherwise. */
#define N2V2R_RANK_NOALLOC_MAX_ELEMS (1 << 20)
int n2v2r_rank_borda_noalloc(void* stream, const double* d_scores, int64_t n, int n_cols, int n_groups,
                             int32_t* d_positions, int64_t* d_borda);

/* ------------------------------------------------------------------------------------
 * K11b  sign transfer for the signed ranks.  Replaces signed_transform_single (model_utils.py:7-19) as
 * N2V2R.signed_ranks_transform applies it to every column of a comparison (model.py:236-256): the score of
 * node i keeps its sign where d_prior[i] > 0 and is negated otherwise (zero / negative / NaN priors).
 * d_scores, d_out: (n_cols x n), float64 (is_int64 = 0) or int64 (is_int64 = 1: the Borda table); may alias.
 * ---------------------------------------------------------------------------------- */
int n2v2r_signed_scores(void* stream, const void* d_scores, int is_int64, int64_t n, int n_cols,
                        const double* d_prior, void* d_out);

/* ------------------------------------------------------------------------------------
 * K0b  weighted edge list -> dense symmetric matrix on the device.  Replaces
 * nx.read_weighted_edgelist + nx.to_numpy_array in DataLoader.__load_graph
 * (dataloader.py:100-104) after the host has parsed the file and resolved duplicate
 * edges (undirected graph: the last line of an edge wins).  d_u / d_v: endpoints as
 * positions in the FINAL node order (negative = node not in the common node set, edge
 * skipped); d_w: weights; d_out: zero-filled (n x ld) matrix of `dtype`.
 * ---------------------------------------------------------------------------------- */
int n2v2r_scatter_edges(void* stream, const int32_t* d_u, const int32_t* d_v, const double* d_w, int64_t n_edges,
                        void* d_out, int dtype, int64_t n, int64_t ld);

/* ------------------------------------------------------------------------------------
 * K0d  device side of match_networks for adjacency files.  Replaces the `.loc[common_rows, common_cols]` reindex
 * (preprocessing_utils.py:291-306) and the optional `.T` (dataloader.py:103) of DataLoader.__load_graph's CSV branch
 * (dataloader.py:86-91): the file is parsed in row chunks on the host, each chunk goes through pinned memory into a
 * float64 device matrix in FILE order, and this kernel writes the matrix of the common nodes in their final order:
 * out[i][j] = in[row_idx[i]][col_idx[j]], or with transpose != 0 out[i][j] = in[col_idx[j]][row_idx[i]]
 * (row_idx then indexes the file's columns and col_idx its rows).  Indices must lie inside `in`.
 * ---------------------------------------------------------------------------------- */
int n2v2r_gather_submatrix(void* stream, const double* d_in, int64_t in_rows, int64_t in_cols, int64_t ld_in,
                           const int32_t* d_row_idx, const int32_t* d_col_idx, int64_t out_rows, int64_t out_cols,
                           int transpose, double* d_out, int64_t ld_out);

/* ------------------------------------------------------------------------------------
 * K0c  sparse host->device transfer of a dense fp32 layer.  The reference converts every
 * dense float32 graph to CSC on the host before the SVD (N2V2R.__fit, model.py:51-55); here
 * the host side of that conversion only serves the PCIe copy: n2v2r_host_pack_nonzeros
 * (HOST code, `threads` std::threads, no CUDA call) writes a bitmap (one bit per entry,
 * ceil(cols/32) 32-bit words per row, bit j of word w = entry 32 w + j) and the non-zero
 * entries of each row in column order; h_row_start[r] is the position of row r's first
 * value in h_values (rows are claimed by the threads in blocks, so the positions are not
 * monotone in r).  "Zero" means all 32 bits zero, so the round trip is bit-exact.
 * *nnz = number of values written, or -1 when they do not fit `capacity` (the caller then
 * copies the layer dense).  n2v2r_unpack_nonzeros rebuilds the dense (rows x ld) fp32
 * matrix on the device from device copies of the three arrays (columns cols..ld-1 are not
 * written).
 * ---------------------------------------------------------------------------------- */
int n2v2r_host_pack_nonzeros(const void* h_a, int64_t rows, int64_t cols, int64_t ld, uint32_t* h_bitmap,
                             void* h_values, int64_t capacity, int64_t* h_row_start, int threads, int64_t* nnz);
int n2v2r_unpack_nonzeros(void* stream, const uint32_t* d_bitmap, const void* d_values, const int64_t* d_row_start,
                          int64_t rows, int64_t cols, void* d_out, int64_t ld);

/* ------------------------------------------------------------------------------------
 * K12  ranking-quality read-out: binary-relevance DCG and nDCG of n_rankings rankings of
 * `len` items each (one CTA per ranking).  Replaces relevance_and_ndcg /
 * normalized_discounted_cumulative_gain / discounted_cumulative_gain (post_utils.py:45-122,
 * defaults binary_relevance=True, denominator=log2(pos+2), ideal = the relevant items that
 * occur in the ranking first) and the loop of parallel_empirical_p_ndcg (post_utils.py:125-128).
 * d_rankings: (n_rankings x len) int32 item ids in rank order (ids outside [0, n_items) count
 * as not relevant); d_relevant: n_items bytes (non-zero = relevant); d_ndcg: n_rankings
 * doubles (0 when no relevant item occurs); d_dcg: n_rankings doubles or NULL.
 * ---------------------------------------------------------------------------------- */
int n2v2r_ndcg(void* stream, const int32_t* d_rankings, int64_t n_rankings, int64_t len,
               const uint8_t* d_relevant, int64_t n_items, double* d_ndcg, double* d_dcg);

#ifdef __cplusplus
}
#endif
#endif /* N2V2R_B200_H */
