/*
 * bsp_oracle.h -- CPU restatement of the reference's block-sparse chain product.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the shipped product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and there only as the checker or the timed CPU arm.
 *
 * Parity status: the reference ships no tests or golden vectors, so this
 * restatement is pinned against outputs of the reference's own translation unit built
 * host-only (oracle/_ref, recipe in oracle/build_ref.sh) on the known-answer cases of
 * tests/golden/ and on seeded random inputs (tests/test_oracle_vs_ref.py).
 *
 * "cu:N" = /root/reference/sparse_matrix_mult.cu line N.
 */
#ifndef BSP_ORACLE_H
#define BSP_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Block-sparse matrix as sorted coordinate list (the reference keeps an ordered
 * std::map keyed by (block row, block col), cu:26-32; we keep the same order). */
typedef struct orc_mat {
    int32_t rows, cols; /* header values, carried but never used in arithmetic (cu:280-282) */
    int32_t k;          /* tile edge */
    int64_t nnzb;       /* number of tiles */
    int32_t *brow;      /* [nnzb] */
    int32_t *bcol;      /* [nnzb] */
    uint64_t *vals;     /* [nnzb * k * k], tile-major, each tile row-major */
} orc_mat;

/* Build from an unsorted coordinate list. Duplicate keys: the last one wins (cu:383). */
orc_mat *orc_from_coo(int32_t rows, int32_t cols, int32_t k, int64_t nnzb,
                      const int32_t *brow, const int32_t *bcol, const uint64_t *vals);
orc_mat *orc_clone(const orc_mat *m);
void orc_free(orc_mat *m);

/* One product A*B (= helper, cu:97-286). Zero tiles are kept (cu:259-269).
 * nprod_out (optional) receives the number of tile products. */
orc_mat *orc_multiply(const orc_mat *a, const orc_mat *b, int64_t *nprod_out);

/* Same, with the emulated kernel loop spread over OpenMP threads (each output tile is
 * still accumulated by one thread in reference order, so the result is identical). */
orc_mat *orc_multiply_mt(const orc_mat *a, const orc_mat *b, int64_t *nprod_out, int threads);

/* Pairwise tree over arr[0..n-1] (= helper2, cu:287-327). Consumes nothing; returns a new
 * matrix. If log is non-NULL, "multiplying a b\n" lines are appended as the reference
 * prints them (cu:301). nprod_total (optional) accumulates tile products. */
orc_mat *orc_chain(orc_mat *const *arr, int n, int64_t *nprod_total, char *log, int64_t log_cap);

/* Whole-program semantics of main (cu:402-682) for P ranks: split the chain as
 * cu:438-456, reduce each sub-chain, then reduce the P partials on rank 0 (cu:569-571).
 * If n < P only rank 0 works (cu:612-616). No prune. */
orc_mat *orc_chain_ranks(orc_mat *const *arr, int n, int P, int64_t *nprod_total);

/* Drop tiles whose k*k values are all zero (intended semantics of cu:577-592). */
void orc_prune_zero(orc_mat *m);

/* Text format of the reference: reader cu:345-383, writer cu:595-608. */
orc_mat *orc_read(const char *path, int32_t k);
int orc_write(const orc_mat *m, const char *path);

/* The arithmetic of cu:59 and cu:61 on single values (exposed for unit tests). */
uint64_t orc_mulmod(uint64_t a, uint64_t b);
uint64_t orc_addmod(uint64_t s, uint64_t p);

#ifdef __cplusplus
}
#endif
#endif
