/*
 * bspgemm.h -- C ABI of the B200-native block-sparse chain-product library (libbspgemm.so).
 *
 * Drop-in boundary for the one hot path of
 * UmeshK2005/Sparse-Matrix-Multiplication-using-OpenMP-MPI-and-CUDA:
 *     helper2 -> helper -> matrix_multiplyKernel          (sparse_matrix_mult.cu, "cu:N" below)
 * i.e. C = A*B on block-sparse matrices of k x k uint64 tiles, and the pairwise-tree product
 * of a chain of them, with the reference's exact arithmetic:
 *     acc = ((acc + ((a*b mod 2^64) % (2^64-1))) mod 2^64) % (2^64-1)        (cu:59, cu:61)
 * folded per output element over the matching tile pairs in ascending inner block index and,
 * inside a pair, ascending inner element index (cu:54-62, cu:149-155).
 *
 * Conventions: plain C types, opaque handles; the caller owns handles, the library owns the
 * device memory behind them; every call returns a status (0 = BSP_OK), the library never
 * exits the process (the reference's CUDA_CHECK, cu:33-41, calls exit()).  One context = one
 * device + one stream; a context is not thread-safe; different contexts may be used from
 * different host threads.  There is no CPU fallback: without a CUDA device every compute
 * call fails with BSP_ERR_CUDA.
 */
#ifndef BSPGEMM_H
#define BSPGEMM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bsp_ctx bsp_ctx;
typedef struct bsp_mat bsp_mat;

enum {
    BSP_OK = 0,
    BSP_ERR_ARG = 1,      /* bad argument (null, negative coordinate, k mismatch, ...) */
    BSP_ERR_CUDA = 2,     /* a CUDA runtime call failed; see bsp_last_error */
    BSP_ERR_NOMEM = 3,    /* device or host allocation failed */
    BSP_ERR_LIMIT = 4,    /* a size exceeds what this build supports (stated in the message) */
    BSP_ERR_IO = 5        /* file could not be read / written / parsed */
};

/* Per-product counters and device timings (CUDA events on the context's stream). */
typedef struct bsp_stats {
    int64_t nnzb_a, nnzb_b, nnzb_c; /* tiles in A, B and C */
    int64_t nprod;                  /* tile products = sum over A tiles (i,j) of |B row j| */
    double ms_symbolic;             /* pattern of C + pair lists */
    double ms_numeric;              /* tile multiply-accumulate */
    double ms_total;                /* whole call, device time */
    int64_t launches;               /* kernels launched by this call */
    int64_t exact_refolds;          /* output elements recomputed with the per-operation path */
    int32_t numeric_path;           /* numeric kernel that ran: 0 generic exact, 4 / 8 / 16 / 32 = specialised u64; 100 = fp32 SIMT, 132 / 164 = tensor cores k=32 / 64 */
    int32_t reserved;
} bsp_stats;

/* ---- context ------------------------------------------------------------------------ */

/* Replaces the process-wide set-up of main (cu:421-436: the four cudaMalloc'd staging
 * buffers lent to every helper call).  `stream` is a cudaStream_t to run on, or NULL for a
 * stream owned by the context. */
int bsp_ctx_create(int device, void *stream, bsp_ctx **out);
void bsp_ctx_destroy(bsp_ctx *ctx);                       /* cu:667-675 */
const char *bsp_last_error(const bsp_ctx *ctx);           /* message of the last failing call */
int bsp_ctx_sync(bsp_ctx *ctx);                           /* cudaStreamSynchronize */
int64_t bsp_ctx_launch_count(const bsp_ctx *ctx);         /* kernels launched so far */
const char *bsp_version(void);

/* ---- matrices (struct one_matrix, cu:26-32) ----------------------------------------- */

/* Upload a coordinate list from HOST memory: brow/bcol [nnzb], vals [nnzb*k*k] tile-major,
 * each tile row-major.  Keys may be unsorted; on duplicate keys the LAST one wins, as the
 * reader's m[{row,col}] = v does (cu:383).  Coordinates must be >= 0. */
int bsp_mat_from_coo(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int64_t nnzb,
                     const int32_t *brow, const int32_t *bcol, const uint64_t *vals, bsp_mat **out);

/* float32 tiles (BASELINE.json configs[4]; an extension -- the reference reads and multiplies
 * uint64 only, cu:375-376, cu:44-66, so there is no reference path to match): same pattern
 * handling, products accumulated in fp32, by tensor cores for k = 32 / 64 (TF32 inputs) or by
 * an fp32 SIMT kernel.  bsp_spgemm / bsp_chain / bsp_prune_zero / panels work on either type;
 * both operands must have the same type. */
int bsp_mat_from_coo_f32(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int64_t nnzb,
                         const int32_t *brow, const int32_t *bcol, const float *vals, bsp_mat **out);
int bsp_mat_to_coo_f32(bsp_ctx *ctx, const bsp_mat *m, int32_t *brow, int32_t *bcol, float *vals);
int bsp_mat_is_f32(const bsp_mat *m);
/* float32 products: 0 = auto (tensor cores -- TF32 inputs, fp32 accumulate -- for k = 32 and 64, fp32 SIMT otherwise),
 * 1 = fp32 SIMT, 2 = tensor cores (k = 32 or 64 only).  The tensor-core path repacks both operands into two scratch
 * buffers of their size that stay with the context (they only grow) until bsp_ctx_destroy. */
int bsp_ctx_set_f32_path(bsp_ctx *ctx, int path);

/* Same, from DEVICE memory (arrays are copied, not adopted: the copies are complete on return). */
int bsp_mat_from_device_coo(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int64_t nnzb,
                            const int32_t *d_brow, const int32_t *d_bcol, const uint64_t *d_vals,
                            bsp_mat **out);

/* Header of a matrix: rows/cols as carried by the reference (cu:280-282), tile edge, tiles. */
int bsp_mat_info(const bsp_mat *m, int32_t *rows, int32_t *cols, int32_t *k, int64_t *nnzb);

/* Download as a coordinate list sorted by (brow, bcol) -- the iteration order of the
 * reference's std::map that the writer relies on (cu:598).  Host buffers sized from
 * bsp_mat_info. */
int bsp_mat_to_coo(bsp_ctx *ctx, const bsp_mat *m, int32_t *brow, int32_t *bcol, uint64_t *vals);

/* Device view (CSR of blocks) for callers that move partial products between GPUs:
 * rowptr [nbrows+1] int32, colidx [nnzb] int32, vals [nnzb*k*k] uint64.  Pointers stay
 * valid until the matrix is freed. */
int bsp_mat_device_view(const bsp_mat *m, int32_t *nbrows, int32_t *nbcols,
                        const int32_t **d_rowptr, const int32_t **d_colidx, const uint64_t **d_vals);

/* Copy the CSR-of-blocks arrays into caller-owned DEVICE buffers (sizes as in
 * bsp_mat_device_view): the sending side of the partial-product exchange that replaces
 * MPI_Send of header / keys / values (cu:471-506); the buffers can be handed to NCCL. */
int bsp_mat_export_device(bsp_ctx *ctx, const bsp_mat *m, int32_t *d_rowptr, int32_t *d_colidx,
                          uint64_t *d_vals);

/* Copy a matrix into another context, possibly on another GPU of the same process (peer copy
 * over NVLink when available): how a4 gathers the partial products on device 0 (replaces
 * MPI_Send / MPI_Recv + map rebuild, cu:471-551). */
int bsp_mat_copy_to_ctx(bsp_ctx *src_ctx, const bsp_mat *m, bsp_ctx *dst_ctx, bsp_mat **out);

/* Build a matrix from a device CSR-of-blocks view (arrays are copied).  The receiving side
 * of the partial-product exchange that replaces MPI_Recv + map rebuild (cu:513-551).  The copies
 * are complete on return. */
int bsp_mat_from_device_bcsr(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t nbrows,
                             int32_t nbcols, int64_t nnzb, const int32_t *d_rowptr,
                             const int32_t *d_colidx, const uint64_t *d_vals, bsp_mat **out);
/* Same for float32 tiles (extension, see bsp_mat_from_coo_f32; no reference counterpart). */
int bsp_mat_from_device_bcsr_f32(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t nbrows,
                                 int32_t nbcols, int64_t nnzb, const int32_t *d_rowptr,
                                 const int32_t *d_colidx, const float *d_vals, bsp_mat **out);

void bsp_mat_free(bsp_ctx *ctx, bsp_mat *m);

/* ---- the hot path ------------------------------------------------------------------- */

/* C = A*B  (= helper, cu:97-286).  A and B are borrowed and not modified; *C is new.
 * All-zero result tiles are kept, as the reference keeps them in intermediates
 * (cu:259-269).  stats may be NULL. */
int bsp_spgemm(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, bsp_mat **C, bsp_stats *stats);

/* Block-row panel [row_lo, row_hi) of C = A*B: the unit of row-panel sharding across GPUs
 * (no counterpart in the reference, which cannot split one product).  The result has the
 * full block-row space of A with tiles only in the panel. */
int bsp_spgemm_panel(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t row_lo, int32_t row_hi,
                     bsp_mat **C, bsp_stats *stats);

/* Concatenate n block-row panels (results of bsp_spgemm_panel over consecutive, disjoint
 * row ranges of the same product, in ascending row order) into one matrix: the gather step
 * of row-panel sharding.  No counterpart in the reference. */
int bsp_mat_concat_panels(bsp_ctx *ctx, bsp_mat *const *panels, int32_t n, bsp_mat **out);

/* Product of mats[0..n-1] by the reference's pairwise tree: adjacent pairs at every level,
 * an odd tail is carried to the next level unchanged (= helper2, cu:287-327).  Inputs are
 * borrowed; *out is new (a copy of mats[0] when n == 1).  If log is non-NULL it receives the
 * "multiplying a b\n" lines the reference prints (cu:301), with `log_base` added to the
 * indices (the sub-chain start index of cu:442/450).  stats (may be NULL) is summed over the
 * products. */
int bsp_chain(bsp_ctx *ctx, bsp_mat *const *mats, int32_t n, bsp_mat **out, bsp_stats *stats,
              char *log, size_t log_cap, int32_t log_base);

/* Remove tiles whose k*k entries are all zero, in place (the intended semantics of
 * cu:577-592; the shipped loop erases while iterating and crashes). */
int bsp_prune_zero(bsp_ctx *ctx, bsp_mat *m);

/* One call from host buffers to host buffers: upload A and B, multiply, download C sorted.
 * The shape of the reference's helper seen from its caller (one_matrix in, one_matrix out).
 * Pass NULL output arrays to only obtain *nnzb_c (then call again with buffers).
 * cap_c = capacity of the output arrays in tiles. */
int bsp_multiply_host(bsp_ctx *ctx, int32_t k,
                      int32_t rows_a, int32_t cols_a, int64_t nnzb_a, const int32_t *brow_a,
                      const int32_t *bcol_a, const uint64_t *vals_a,
                      int32_t rows_b, int32_t cols_b, int64_t nnzb_b, const int32_t *brow_b,
                      const int32_t *bcol_b, const uint64_t *vals_b,
                      int64_t cap_c, int64_t *nnzb_c, int32_t *brow_c, int32_t *bcol_c,
                      uint64_t *vals_c, bsp_stats *stats);

/* The same compaction as bsp_prune_zero for callers that must not allocate device memory:
 * processes that gave all of the GPU's memory to their own arena at start-up, or that cannot
 * afford the device-wide stall of cudaMalloc/cudaFree inside a latency-bound loop.  Has no
 * counterpart in the reference.  The caller passes the input CSR-of-blocks arrays and output
 * arrays it already owns (d_colidx_out / d_vals_out sized for nnzb tiles, d_rowptr_out for
 * nbrows+1 entries, none aliasing an input); *nnzb_out receives the number of tiles kept.
 * The flags and positions the compaction needs live in an array built into the library
 * (about 8 MiB, resident on every GPU that loads the library until the process exits), so the
 * call performs no cudaMalloc, cudaFree or pool allocation.  At most BSP_NOALLOC_MAX_TILES
 * tiles; more returns BSP_ERR_LIMIT.  All calls on one GPU share that array: do not let two of
 * them overlap on the device (they race).  The first call on each GPU looks up the array's
 * address; make it before starting any stream capture. */
#define BSP_NOALLOC_MAX_TILES (1 << 20)
int bsp_prune_zero_noalloc(bsp_ctx *ctx, int32_t k, int32_t nbrows, int64_t nnzb,
                           const int32_t *d_rowptr, const int32_t *d_colidx, const uint64_t *d_vals,
                           int32_t *d_rowptr_out, int32_t *d_colidx_out, uint64_t *d_vals_out,
                           int64_t *nnzb_out);

/* ---- several GPUs: NCCL transport and the distributed chain (replaces cu:438-571) ------ */

/* One communicator per GPU, held by the process (or host thread) that drives it.  NCCL is bound
 * at run time (libnccl.so.2); without it these calls return BSP_ERR_CUDA.  A communicator is used
 * with one context; its calls are collective where stated and run on the context's stream. */
typedef struct bsp_comm bsp_comm;
#define BSP_COMM_ID_BYTES 128
/* Rank 0 makes the rendezvous token (ncclGetUniqueId) and hands it to the other ranks by any
 * means (the reference's MPI_Init, cu:404, has no counterpart here); cap >= BSP_COMM_ID_BYTES. */
int bsp_comm_unique_id(void *id, size_t cap);
/* Collective over all ranks (ncclCommInitRank). */
int bsp_comm_create(bsp_ctx *ctx, const void *id, int nranks, int rank, bsp_comm **out);
/* Wrap an ncclComm_t the caller already owns (it is not destroyed with the handle). */
int bsp_comm_adopt(bsp_ctx *ctx, void *nccl_comm, int nranks, int rank, bsp_comm **out);
void bsp_comm_destroy(bsp_ctx *ctx, bsp_comm *comm);
int bsp_comm_info(const bsp_comm *comm, int *nranks, int *rank, int64_t *bytes_sent);

/* Move one matrix (uint64 or float32 tiles) between two ranks: a fixed-size header, then row
 * pointers, block columns and tile values in one NCCL group.  Replaces MPI_Send of header / keys /
 * values (cu:471-506) and MPI_Recv + map rebuild (cu:513-551).  bsp_mat_send returns when the
 * matrix may be changed; bsp_mat_recv waits once on the host to size the body, as cu:513-520 does. */
int bsp_mat_send(bsp_ctx *ctx, bsp_comm *comm, const bsp_mat *m, int peer);
int bsp_mat_recv(bsp_ctx *ctx, bsp_comm *comm, int peer, bsp_mat **out);
/* Collective: the root passes `in`, every other rank receives a copy in *out (no reference
 * counterpart: the reference never replicates an operand). */
int bsp_mat_bcast(bsp_ctx *ctx, bsp_comm *comm, int root, const bsp_mat *in, bsp_mat **out);

/* Collective: the ranks' block-row panels of one product (bsp_spgemm_panel over disjoint,
 * rank-ascending row ranges) are concatenated on `root` (*out there, NULL elsewhere); received
 * straight into the result's arrays.  The "gather for file output" step of row-panel sharding. */
int bsp_gather_panels(bsp_ctx *ctx, bsp_comm *comm, const bsp_mat *panel, int root, bsp_mat **out);

/* Collective: the whole multi-rank chain product of main (cu:438-571).  Every rank passes ITS
 * sub-chain (mats[0..n_local-1], the split of cu:438-456 is the caller's: rank r loads matrices
 * [r*(N/P), (r+1)*(N/P)), the last rank the remainder; when N < P rank 0 passes all and the others
 * none, cu:612-616); log / log_base as in bsp_chain.  The sub-chain is reduced locally, then the P
 * partial products are reduced with the association tree of rank 0's helper2 (cu:569-571) as a
 * step-doubling tree in which EVERY product is sharded by block rows over the ranks of its subtree
 * and stays distributed between levels (the reference multiplies the partials on rank 0 alone).
 * Same tree, same per-element fold order: the result on rank 0 (*out; NULL elsewhere) is
 * bit-identical to the P-rank reference run.  Zero tiles are kept (prune on rank 0 afterwards). */
int bsp_chain_distributed(bsp_ctx *ctx, bsp_comm *comm, bsp_mat *const *mats, int32_t n_local, int32_t log_base,
                          bsp_mat **out, bsp_stats *stats, char *log, size_t log_cap);

/* Contiguous block-row panels of A*B for `nparts` GPUs, balanced by the work of every block row
 * (tile products, from the operands' row pointers, computed on the device; block rows too long for the
 * shared-memory classes of the symbolic pass count 40 % more): bounds[0..nparts], host array.  Panel p is rows [bounds[p], bounds[p+1]); feed it to bsp_spgemm_panel.  No reference
 * counterpart (the reference cannot split one product). */
int bsp_panel_bounds(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t nparts, int32_t *bounds);

/* ---- text format of the reference (either side of the path) -------------------------- */

/* Parse one "matrixI" file (rows cols / blocks / {brow bcol / k*k values}), cu:345-383. */
int bsp_mat_read_text(bsp_ctx *ctx, const char *path, int32_t k, bsp_mat **out);

/* Write the output layout of cu:595-608 ("R C", count, then per tile "i j" and k rows of k
 * values, single spaces, no trailing space, '\n' line ends), tiles in (brow, bcol) order. */
int bsp_mat_write_text(bsp_ctx *ctx, const bsp_mat *m, const char *path);

/* Parse n files at once: host threads parse into pinned staging buffers while the calling
 * thread uploads (asynchronous copies) and builds the finished ones on the device, so upload
 * and ingestion hide behind parsing.  The reference loads its files with one OpenMP task per
 * file into pageable memory (cu:334-339); the pinned/asynchronous ingestion its report
 * describes is not in its code.  out[i] receives the matrix of paths[i]; on error every out[i]
 * is NULL and nothing is left allocated. */
int bsp_mat_read_text_many(bsp_ctx *ctx, const char *const *paths, int32_t n, int32_t k, bsp_mat **out);

/* ---- binary container + content checksum (no counterpart in the reference) ------------ */

/* 64-bit checksum of the matrix content (header, tile coordinates, tile values), computed on
 * the device in one pass over the tiles.  It depends on the content only, not on how the
 * matrix was produced (one GPU or row panels on several), so full-size results -- 19.5 GB for
 * BASELINE.json configs[1], where the reference's only comparison is a diff of 35-50 GB text
 * files (cu:595-608) -- can be compared by value.  Defined in csrc/bsp_checksum.cu. */
int bsp_mat_checksum(bsp_ctx *ctx, const bsp_mat *m, uint64_t *out);
/* The two halves of the definition, for a matrix that is spread over several GPUs as row panels:
 * the order-independent sum over this panel's tiles (add the ranks' sums modulo 2^64), and the
 * final mix of that sum with the header of the whole matrix. */
int bsp_mat_checksum_body(bsp_ctx *ctx, const bsp_mat *m, uint64_t *body);
uint64_t bsp_checksum_finish(int32_t rows, int32_t cols, int32_t k, int32_t esize, int64_t nnzb, uint64_t body);

/* Binary file: 64-byte header ("BSPMAT01", rows, cols, k, element size, tile count, checksum),
 * then brow[], bcol[] (int32, tiles in (brow, bcol) order) and the tile values.  The reader
 * verifies the checksum on the device and returns BSP_ERR_IO on a mismatch.  Works for uint64
 * and float32 matrices. */
int bsp_mat_write_bin(bsp_ctx *ctx, const bsp_mat *m, const char *path);
int bsp_mat_read_bin(bsp_ctx *ctx, const char *path, bsp_mat **out);

#ifdef __cplusplus
}
#endif
#endif /* BSPGEMM_H */
