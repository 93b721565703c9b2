// bsp_symbolic.cu -- GPU symbolic pass of C = A*B on CSR-of-blocks.
//
// Replaces the reference's single-threaded host symbolic phase: the B row index
// (cu:141-145), the map d: (i,c) -> [j...] (cu:147-156), the key array (cu:167-176) and the
// per-round pack loop that copies every A/B tile pair into a staging buffer (cu:189-218).
// Here nothing is copied: the pass produces, per output tile, a list of (A tile index,
// B tile index) pairs in ascending inner block index j -- the order in which the reference's
// kernel folds them (cu:54-62) -- and the numeric pass gathers tiles by index.
//
// One warp per block row i of A:
//   count: set bit c in a per-warp shared-memory bitmap for every (j in A row i, c in B row j)
//          (warp lanes run over B row j, atomicOr into shared memory), popcount -> |C row i|;
//          the bitmap makes the merged column set sorted and duplicate-free for free.
//   fill : rebuild the bitmap, prefix-popcount its words with warp shuffles (rank structure),
//          emit C's sorted colidx, count pairs per output tile, scan them in place into
//          pair_ptr, then append the pairs walking A row i in ascending j (lanes over B row j;
//          inside one j all target tiles differ, so the cursor updates never collide and the
//          append order per tile is ascending j).
// HBM-bound integer work; algorithmic bytes:
//   4*nnzb(A) + 4*nprod + 4*nnzb(C) + 12*(nbr+1), plus the pair lists this design adds:
//   8*nprod + 8*(nnzb(C)+1).
#include <cstdlib>
#include <vector>

#include "bsp_internal.h"

namespace {

constexpr int SYM_WARPS = 8;
constexpr int SYM_THREADS = SYM_WARPS * 32;

__device__ __forceinline__ int warp_sum(int v)
{
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

__device__ __forceinline__ int warp_excl_scan(int v, int *total)
{
    const int lane = threadIdx.x & 31;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    *total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}

// Set, in the warp's bitmap, the block columns reachable from block row `row` of A.
// Returns (warp-uniform) the number of tile products of the row.
__device__ __forceinline__ int build_row_bitmap(uint32_t *bm, int words, const int32_t *__restrict__ a_rowptr,
                                                const int32_t *__restrict__ a_colidx, int b_nbr,
                                                const int32_t *__restrict__ b_rowptr,
                                                const int32_t *__restrict__ b_colidx, int row)
{
    const int lane = threadIdx.x & 31;
    for (int w = lane; w < words; w += 32) bm[w] = 0u;
    __syncwarp();
    const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
    int nprod = 0;
    for (int p = a0; p < a1; p++) {
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
        nprod += b1 - b0;
        for (int q = b0 + lane; q < b1; q += 32) {
            const int c = b_colidx[q];
            atomicOr(&bm[c >> 5], 1u << (c & 31));
        }
    }
    __syncwarp();
    return nprod;
}

// Walk the tiles of one B row with the whole warp, four independent index loads per lane and round
// (a row of a few hundred tiles otherwise costs one exposed L2 round trip per 32 tiles).
template <typename F>
__device__ __forceinline__ void for_each_tile_of_row(const int32_t *__restrict__ b_colidx, int b0, int b1, int lane, F f)
{
    for (int q0 = b0 + lane; q0 < b1; q0 += 128) {
        int c[4];
#pragma unroll
        for (int u = 0; u < 4; u++) c[u] = q0 + 32 * u < b1 ? b_colidx[q0 + 32 * u] : -1;
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (c[u] >= 0) f(q0 + 32 * u, c[u]);
    }
}

// Light rows (few tile products) are counted by one warp each; a row whose product count
// exceeds SYM_HEAVY (R-MAT hubs, dense chain levels) is appended to a list and counted by a
// whole CTA in sym_count_heavy_kernel, so that one warp never walks 10^5 products alone.
constexpr int SYM_HEAVY = 4096;

__global__ void __launch_bounds__(SYM_THREADS)
sym_count_kernel(int a_nbr, const int32_t *__restrict__ a_rowptr, const int32_t *__restrict__ a_colidx, int b_nbr,
                 const int32_t *__restrict__ b_rowptr, const int32_t *__restrict__ b_colidx, int words, int row_lo,
                 int row_hi, int32_t *__restrict__ row_cnt, int32_t *__restrict__ row_prod,
                 int32_t *__restrict__ heavy_list, int32_t *__restrict__ heavy_count,
                 int32_t *__restrict__ overflow_flag)
{
    extern __shared__ uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *bm = smem + (size_t)warp * words;
    for (int row = blockIdx.x * SYM_WARPS + warp; row < a_nbr; row += gridDim.x * SYM_WARPS) {
        int cnt = 0, nprod = 0;
        if (row >= row_lo && row < row_hi && a_rowptr[row + 1] > a_rowptr[row]) {
            // tile products of the row: lanes over the A entries.  Summed in 64 bits: the per-row counts
            // are kept as int32, a row beyond 2^31-1 products raises the flag (BSP_ERR_LIMIT on the host)
            const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
            long long np64 = 0;
            for (int p = a0 + lane; p < a1; p += 32) {
                const int j = a_colidx[p];
                if (j < b_nbr) np64 += b_rowptr[j + 1] - b_rowptr[j];
            }
#pragma unroll
            for (int d = 16; d; d >>= 1) np64 += __shfl_xor_sync(0xffffffffu, np64, d);
            if (np64 > 0x7fffffffLL) {
                if (lane == 0) *overflow_flag = 1;
                np64 = 0; // the product is refused; keep the remaining passes in range
            }
            nprod = (int)np64;
            if (nprod > SYM_HEAVY) {
                if (lane == 0) heavy_list[atomicAdd(heavy_count, 1)] = row;
                cnt = -1; // filled in by sym_count_heavy_kernel
            } else if (nprod > 0) {
                build_row_bitmap(bm, words, a_rowptr, a_colidx, b_nbr, b_rowptr, b_colidx, row);
                for (int w = lane; w < words; w += 32) cnt += __popc(bm[w]);
                cnt = warp_sum(cnt);
                __syncwarp();
            }
        }
        if (lane == 0) {
            if (cnt >= 0) row_cnt[row] = cnt;
            row_prod[row] = nprod;
        }
    }
}

constexpr int HEAVY_THREADS = 512;

__global__ void __launch_bounds__(HEAVY_THREADS)
sym_count_heavy_kernel(const int32_t *__restrict__ heavy_list, const int32_t *__restrict__ heavy_count,
                       const int32_t *__restrict__ a_rowptr, const int32_t *__restrict__ a_colidx, int b_nbr,
                       const int32_t *__restrict__ b_rowptr, const int32_t *__restrict__ b_colidx, int words,
                       int32_t *__restrict__ row_cnt)
{
    extern __shared__ uint32_t bm[];
    __shared__ int warp_cnt[HEAVY_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = HEAVY_THREADS / 32;
    const int nheavy = *heavy_count;
    for (int idx = blockIdx.x; idx < nheavy; idx += gridDim.x) {
        const int row = heavy_list[idx];
        for (int w = tid; w < words; w += HEAVY_THREADS) bm[w] = 0u;
        __syncthreads();
        const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
        for (int p = a0 + warp; p < a1; p += nwarp) { // warps over A entries, lanes over B row j
            const int j = a_colidx[p];
            if (j >= b_nbr) continue;
            for_each_tile_of_row(b_colidx, b_rowptr[j], b_rowptr[j + 1], lane,
                                 [&](int, int c) { atomicOr(&bm[c >> 5], 1u << (c & 31)); });
        }
        __syncthreads();
        int cnt = 0;
        for (int w = tid; w < words; w += HEAVY_THREADS) cnt += __popc(bm[w]);
        cnt = warp_sum(cnt);
        if (lane == 0) warp_cnt[warp] = cnt;
        __syncthreads();
        if (tid == 0) {
            int t = 0;
            for (int i = 0; i < nwarp; i++) t += warp_cnt[i];
            row_cnt[row] = t;
        }
        __syncthreads();
    }
}

// ---- fill, shared-memory variant: one CTA per block row of C ----------------------------------
// Everything the row needs lives in shared memory: the column bitmap and its rank table, one
// counter per output tile, and the row's whole pair list (4 B per tile product, packed).  Pairs are
// placed in ascending j by walking the A row entry by entry with the whole CTA spread over
// B row j (inside one j every target tile is distinct, so the cursor bump needs no atomic),
// then streamed to global memory with coalesced stores -- DRAM sees each pair exactly once.
constexpr int FILL_THREADS = 512;
constexpr int FILL_SORT_MAX = 16; // rows whose tiles all have <= this many pairs: parallel placement + tiny sorts

__device__ __forceinline__ int block_excl_scan_i32(int v, int *sm33, int *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) sm33[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < nwarp ? sm33[lane] : 0;
        int ws = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int o = __shfl_up_sync(0xffffffffu, ws, d);
            if (lane >= d) ws += o;
        }
        sm33[lane] = ws - w;
        if (lane == 31) sm33[32] = ws;
    }
    __syncthreads();
    const int res = inc - v + sm33[warp];
    *total = sm33[32];
    __syncthreads();
    return res;
}

// Shared-memory budget of one row: 8*words (bitmap + rank table) + 8*(acap+2) (A entries)
// + 4*(tcap+4) (tile counters) + 4*pcap (the row's pairs, packed) + 2*(pcap+8) (tile of every product).
__host__ __device__ inline size_t fill_smem_bytes(int words, int pcap, int tcap, int acap)
{
    return 8 * (size_t)words + 8 * ((size_t)acap + 2) + 4 * ((size_t)tcap + 4) + 4 * (size_t)pcap + 2 * ((size_t)pcap + 8);
}

// A pair is kept in shared memory as one word: bit 31 = last pair of its tile, bits 22..30 = ordinal of the A entry
// in the row (below FILL_ACAP), bits 0..21 = offset of the B tile in its block row.  Half the bytes of the (A tile,
// B tile) index pair that goes to global memory -- a third more rows per SM -- and the word orders a tile's pairs
// by A entry as it is.  Block rows of B longer than 2^22 tiles keep the row out of this kernel (bsp_symbolic).
constexpr int FILL_ACAP = 512;
constexpr int FILL_IBITS = 22;

// BSP_SYM_PROF=1: cycles thread 0 of every CTA spends in each phase of sym_fill_smem_kernel (summed over rows)
__device__ unsigned long long sym_fill_prof[8];

template <bool PROF>
__global__ void __launch_bounds__(FILL_THREADS)
sym_fill_smem_kernel(const int32_t *__restrict__ rowlist, const int32_t *__restrict__ a_rowptr,
                     const int32_t *__restrict__ a_colidx, int b_nbr, const int32_t *__restrict__ b_rowptr,
                     const int32_t *__restrict__ b_colidx, int words, int pcap, int tcap, int acap,
                     const int32_t *__restrict__ c_rowptr, const int64_t *__restrict__ row_prod_ptr,
                     int32_t *__restrict__ c_colidx, int64_t *__restrict__ pair_ptr, int2 *__restrict__ pairs)
{
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ int sm33[33];
    uint2 *bmr = (uint2 *)smem;                                  // [words] .x = bitmap word of the row's block columns, .y = its rank
    int2 *aent = (int2 *)(bmr + words);                          // [acap+2] (first B tile, product offset) per A entry
    int *cur = (int *)(aent + acap + 2);                         // [tcap+4] per-tile counter, then cursor
    uint32_t *sp = (uint32_t *)(cur + tcap + 4);                 // [pcap]   the row's pairs, packed
    unsigned short *prodt = (unsigned short *)(sp + pcap);       // [pcap+8] output tile (rank in the row) of every product
    const int tid = threadIdx.x, lane = threadIdx.x & 31;
    const int row = rowlist[blockIdx.x];
    const int cbase = c_rowptr[row], T = c_rowptr[row + 1] - cbase;
    const int64_t pbase = row_prod_ptr[row];
    const int P = (int)(row_prod_ptr[row + 1] - pbase);
    const int a0 = a_rowptr[row], nA = a_rowptr[row + 1] - a0;

    long long tk = 0;
    auto tick = [&](int slot) {
        if (PROF && tid == 0) { const long long now = clock64(); atomicAdd(&sym_fill_prof[slot], (unsigned long long)(now - tk)); tk = now; }
    };
    if (PROF && tid == 0) tk = clock64();
    for (int w = tid; w < words; w += FILL_THREADS) bmr[w] = make_uint2(0u, 0u);
    for (int t = tid; t < T + 1; t += FILL_THREADS) cur[t] = 0;
    // 0. A entries: first tile of B row j and (after the scan) the offset of its products
    int carry = 0;
    for (int e0 = 0; e0 < nA; e0 += FILL_THREADS) {
        const int e = e0 + tid;
        int b0 = 0, len = 0;
        if (e < nA) {
            const int j = a_colidx[a0 + e];
            if (j < b_nbr) { b0 = b_rowptr[j]; len = b_rowptr[j + 1] - b0; }
        }
        int tot;
        const int ex = block_excl_scan_i32(len, sm33, &tot);
        if (e < nA) aent[e] = make_int2(b0, carry + ex);
        carry += tot;
    }
    if (tid == 0) aent[nA] = make_int2(0, P);
    __syncthreads();
    tick(0);
    // 1. one pass over the row's products, flattened over the CTA (x = product number): find
    //    its A entry by binary search in the offset table, read its block column from B once,
    //    keep it in shared memory (in the area the pairs will occupy later) and set its bit
    int *tmpc = (int *)sp;
    // every thread owns a contiguous run of products [x0, x1): one binary search for the A entry
    // of x0, then it walks forward (runs are ~P/512 long, so consecutive lanes read consecutive
    // memory and the run length being odd keeps the shared-memory stores conflict-free)
    int per = (P + FILL_THREADS - 1) / FILL_THREADS;
    per |= 1;
    const int x0 = min(tid * per, P), x1 = min(x0 + per, P);
    int e_run = 0;
    if (x0 < x1) {
        int lo = 0, hi = nA; // last entry with offset <= x0
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (aent[mid].y <= x0) lo = mid; else hi = mid;
        }
        e_run = lo;
    }
    {
        // the loads of one batch are independent, so their L2 latencies overlap
        constexpr int UB = 8;
        int e = e_run;
        int2 ae = aent[e];
        int next_off = aent[e + 1].y;
        for (int xb = x0; xb < x1; xb += UB) {
            int addr[UB], cv[UB];
#pragma unroll
            for (int u = 0; u < UB; u++) {
                const int x = xb + u;
                addr[u] = -1;
                if (x < x1) {
                    while (x >= next_off) { e++; ae = aent[e]; next_off = aent[e + 1].y; }
                    addr[u] = ae.x + (x - ae.y);
                }
            }
#pragma unroll
            for (int u = 0; u < UB; u++) cv[u] = addr[u] >= 0 ? b_colidx[addr[u]] : 0;
#pragma unroll
            for (int u = 0; u < UB; u++)
                if (addr[u] >= 0) {
                    tmpc[xb + u] = cv[u];
                    atomicOr(&bmr[cv[u] >> 5].x, 1u << (cv[u] & 31));
                }
        }
    }
    __syncthreads();
    tick(1);
    // 2. rank table; the sorted block columns of the C row go straight to global memory (a thread writes the
    //    columns of its bitmap word to consecutive slots, neighbouring threads to neighbouring runs)
    carry = 0;
    for (int w0 = 0; w0 < words; w0 += FILL_THREADS) {
        const int w = w0 + tid;
        const uint32_t bits = w < words ? bmr[w].x : 0u;
        int tot;
        const int ex = block_excl_scan_i32(__popc(bits), sm33, &tot);
        if (w < words) {
            bmr[w].y = (uint32_t)(carry + ex);
            uint32_t b = bits;
            int32_t *o = c_colidx + cbase + carry + ex;
            while (b) {
                const int bit = __ffs(b) - 1;
                b &= b - 1;
                *o++ = w * 32 + bit;
            }
        }
        carry += tot;
    }
    __syncthreads();
    tick(2);
    // 3. output tile of every product (16-bit rank in the row) + pairs per tile
    for (int x = x0; x < x1; x++) {
        const int c = tmpc[x];
        const uint2 br = bmr[c >> 5]; // bitmap word and its rank in one 8-byte read
        const int t = (int)br.y + __popc(br.x & ((1u << (c & 31)) - 1u));
        prodt[x] = (unsigned short)t;
        atomicAdd(&cur[t], 1);
    }
    __syncthreads();
    tick(3);
    // 4. exclusive scan -> first slot of every tile; publish pair_ptr; cur[] becomes the cursor
    //    (every thread scans a contiguous chunk of tiles serially; one block scan joins the chunks)
    int maxcnt = 0;
    {
        const int ch = (T + FILL_THREADS - 1) / FILL_THREADS;
        const int t0 = min(tid * ch, T), t1 = min(t0 + ch, T);
        int sum = 0;
        for (int t = t0; t < t1; t++) { const int v = cur[t]; sum += v; maxcnt = max(maxcnt, v); }
        int tot;
        int ex = block_excl_scan_i32(sum, sm33, &tot);
        for (int t = t0; t < t1; t++) { const int v = cur[t]; cur[t] = ex; ex += v; }
        __syncthreads();
        for (int t = tid; t < T; t += FILL_THREADS) pair_ptr[(size_t)cbase + t] = pbase + cur[t];
    }
    maxcnt = __syncthreads_or(maxcnt > FILL_SORT_MAX);
    tick(4);
    constexpr uint32_t LAST = 0x80000000u;
    if (!maxcnt) {
        // 5a. every tile has few pairs: place all products at once (cursor bumped atomically, any
        //     order), then put each tile's pairs in ascending A entry -- which is ascending j,
        //     because the A row is sorted.
        if (tid == 0) sm33[0] = 0; // length of the sort work list (published by the barrier after the placement)
        {
            int e = e_run;
            int2 ae = aent[e];
            int next_off = aent[e + 1].y;
            for (int x = x0; x < x1; x++) {
                while (x >= next_off) { e++; ae = aent[e]; next_off = aent[e + 1].y; }
                const int pos = atomicAdd(&cur[prodt[x]], 1);
                sp[pos] = ((uint32_t)e << FILL_IBITS) | (uint32_t)(x - ae.y);
            }
        }
        __syncthreads();
        tick(5);
        // One pair: nothing to order.  Two pairs: one compare.  Both without a branch: first and last pair of the
        // tile are read, put in order and written back with the flag on the last.  Three or more (about one tile in ten
        // on a uniform 1 % pattern): the tile goes on a work list -- the product-tile table is free now -- so that the
        // insertion sorts run on full warps afterwards instead of on a few lanes of every warp.
        unsigned short *work = prodt;
        for (int t0 = 0; t0 < T; t0 += FILL_THREADS) {
            const int t = t0 + tid;
            int cnt = 0, s1 = 0;
            if (t < T) {
                s1 = t ? cur[t - 1] : 0; // cursors now hold tile ends
                cnt = cur[t] - s1;
            }
            if (cnt == 1 || cnt == 2) {
                const uint32_t p0 = sp[s1], p1 = sp[s1 + cnt - 1];
                sp[s1] = min(p0, p1);                // same word twice when cnt == 1 ...
                sp[s1 + cnt - 1] = max(p0, p1) | LAST; // ... and this store comes second
            }
            const unsigned many = __ballot_sync(0xffffffffu, cnt > 2);
            if (many) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&sm33[0], __popc(many));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (cnt > 2) work[base + __popc(many & ((1u << lane) - 1u))] = (unsigned short)t;
            }
        }
        __syncthreads();
        const int nwork = sm33[0];
        for (int w = tid; w < nwork; w += FILL_THREADS) {
            const int t = work[w];
            const int e1 = cur[t], s1 = t ? cur[t - 1] : 0;
            for (int i = s1 + 1; i < e1; i++) {
                const uint32_t v = sp[i];
                int j = i - 1;
                while (j >= s1 && sp[j] > v) { sp[j + 1] = sp[j]; j--; }
                sp[j + 1] = v;
            }
            sp[e1 - 1] |= LAST; // 6. last pair of the tile
        }
        __syncthreads();
    } else {
        // 5b. ordered placement: A row in ascending j, the CTA over B row j.  Everything the loop
        //     touches is in shared memory; inside one j all target tiles differ, so the cursor bump
        //     needs no atomic, and the barrier between two j keeps the append order.
        for (int e = 0; e < nA; e++) {
            const int2 ae = aent[e];
            const int len = aent[e + 1].y - ae.y;
            if (len == 0) continue; // uniform
            for (int i = tid; i < len; i += FILL_THREADS) {
                const int t = prodt[ae.y + i];
                const int pos = cur[t];
                cur[t] = pos + 1;
                sp[pos] = ((uint32_t)e << FILL_IBITS) | (uint32_t)i;
            }
            __syncthreads();
        }
        // 6. mark the last pair of every tile: the numeric pass finds tile ends in the pair stream itself
        for (int t = tid; t < T; t += FILL_THREADS) sp[cur[t] - 1] |= LAST;
        __syncthreads();
    }
    tick(6);
    // 7. stream the row's pairs out as (A tile, B tile | last): 8 bytes per pair, coalesced
    for (int x = tid; x < P; x += FILL_THREADS) {
        const uint32_t w = sp[x];
        const int e = (int)((w >> FILL_IBITS) & (FILL_ACAP - 1));
        pairs[pbase + x] = make_int2(a0 + e, (aent[e].x + (int)(w & ((1u << FILL_IBITS) - 1u))) | (int)(w & LAST));
    }
    tick(7);
}

// ---- fill, long rows: one CTA per row, every warp owns a range of output tiles -----------------------
// Rows beyond the shared-memory classes (15 000 .. 2^18 tile products: two thirds of the products of an
// R-MAT pattern).  sym_fill_global_kernel below keeps their per-tile counters in global memory and, when
// some tile has more than a few pairs (almost every such row has one), places the pairs entry by entry
// with a CTA barrier after each: ncu showed 10 % of the issue slots used and the barrier as the top stall.
// Here the counters live in shared memory (4 bytes per OUTPUT TILE, not per product: a row with 250 000
// products has 16 000 tiles) and the placement needs no CTA barrier: after counting, the row's tiles are cut
// into 16 ranges of equal pair counts; the B rows are sorted by block column, so a tile range is a column
// range and a contiguous piece of every B row (two binary searches per entry, 32 entries at a time, one per
// lane).  Every warp walks the A row alone, in ascending inner index, over its own piece of each B row: it
// is the only writer of its tiles' cursors, the pieces of one B row fall into different tiles, and a
// __syncwarp() between two entries keeps the append order -- the order the arithmetic needs, by construction.
constexpr int BIG_WARPS_LONG = 16; // long rows: 512 threads
// (a 4-warp instance of the same kernel for rows of a few thousand products -- configs[1]: 6.7 k per row -- was measured:
// 1.40 ms against 0.94 ms for the CTA-per-row shared-memory kernel; the pieces of a B row per warp are too short there)

__host__ __device__ inline size_t big_smem_bytes(int words, int tcap) { return 8 * (size_t)words + 4 * ((size_t)tcap + 4); }

template <int BIG_WARPS>
__global__ void __launch_bounds__(BIG_WARPS * 32, BIG_WARPS >= 16 ? 2 : 8)
sym_fill_big_kernel(const int32_t *__restrict__ rowlist, const int32_t *__restrict__ a_rowptr,
                    const int32_t *__restrict__ a_colidx, int b_nbr, const int32_t *__restrict__ b_rowptr,
                    const int32_t *__restrict__ b_colidx, int nbc, int words, const int32_t *__restrict__ c_rowptr,
                    const int64_t *__restrict__ row_prod_ptr, int32_t *c_colidx, int64_t *__restrict__ pair_ptr,
                    int2 *pairs)
{
    constexpr int BIG_THREADS = BIG_WARPS * 32;
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ int sm33[33];
    __shared__ int wt0[BIG_WARPS + 1], wc0[BIG_WARPS + 1]; // first tile / first block column of every warp's range
    uint32_t *bm = smem;
    uint32_t *wrank = bm + words;
    int *cur = (int *)(wrank + words); // [T + 1] pairs per tile, then first slot, then cursor (row-local)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int row = rowlist[blockIdx.x];
    const int cbase = c_rowptr[row], T = c_rowptr[row + 1] - cbase;
    const int64_t pbase = row_prod_ptr[row];
    const int P = (int)(row_prod_ptr[row + 1] - pbase);
    const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
    for (int w = tid; w < words; w += BIG_THREADS) bm[w] = 0u;
    for (int t = tid; t < T + 1; t += BIG_THREADS) cur[t] = 0;
    __syncthreads();
    // A. column bitmap of the row (warps over A entries, lanes over B row j)
    for (int p = a0 + warp; p < a1; p += BIG_WARPS) {
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        for_each_tile_of_row(b_colidx, b_rowptr[j], b_rowptr[j + 1], lane,
                             [&](int, int c) { atomicOr(&bm[c >> 5], 1u << (c & 31)); });
    }
    __syncthreads();
    // B. rank table + sorted colidx of the C row
    int carry = 0;
    for (int w0 = 0; w0 < words; w0 += BIG_THREADS) {
        const int w = w0 + tid;
        const uint32_t bits = w < words ? bm[w] : 0u;
        int tot;
        const int ex = block_excl_scan_i32(__popc(bits), sm33, &tot);
        if (w < words) {
            wrank[w] = (uint32_t)(carry + ex);
            uint32_t b = bits;
            int o = cbase + carry + ex;
            while (b) {
                const int bit = __ffs(b) - 1;
                b &= b - 1;
                c_colidx[o++] = w * 32 + bit;
            }
        }
        carry += tot;
    }
    __syncthreads();
    // C. pairs per tile
    for (int p = a0 + warp; p < a1; p += BIG_WARPS) {
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        for_each_tile_of_row(b_colidx, b_rowptr[j], b_rowptr[j + 1], lane, [&](int, int c) {
            const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
            atomicAdd(&cur[t], 1);
        });
    }
    __syncthreads();
    // D. counts -> first slot of every tile (every thread scans a contiguous piece serially, one block scan joins
    //    the pieces); pair_ptr of the row
    {
        const int ch = (T + BIG_THREADS - 1) / BIG_THREADS;
        const int t0 = min(tid * ch, T), t1 = min(t0 + ch, T);
        int sum = 0;
        for (int t = t0; t < t1; t++) sum += cur[t];
        int tot;
        int ex = block_excl_scan_i32(sum, sm33, &tot);
        for (int t = t0; t < t1; t++) { const int v = cur[t]; cur[t] = ex; ex += v; }
        if (tid == 0) cur[T] = P;
        __syncthreads();
        for (int t = tid; t < T; t += BIG_THREADS) pair_ptr[(size_t)cbase + t] = pbase + cur[t];
    }
    // E. the warps' tile ranges: equal shares of the row's pairs; the range of warp w starts at the first tile whose
    //    first slot is at or beyond w * P / 16
    if (tid <= BIG_WARPS) {
        int t = T;
        if (tid < BIG_WARPS) {
            const int want = (int)((long long)P * tid / BIG_WARPS);
            int lo = 0, hi = T;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (cur[mid] >= want) hi = mid; else lo = mid + 1;
            }
            t = tid == 0 ? 0 : lo;
        }
        wt0[tid] = t;
    }
    __threadfence_block(); // c_colidx of this row was written by this CTA in step B
    __syncthreads();
    if (tid <= BIG_WARPS) wc0[tid] = wt0[tid] < T ? c_colidx[cbase + wt0[tid]] : nbc;
    __syncthreads();
    // F. ordered placement, warp by warp
    {
        const int c_lo = wc0[warp], c_hi = wc0[warp + 1];
        if (c_lo < c_hi) {
            for (int e0 = a0; e0 < a1; e0 += 32) {
                // this lane's entry: the piece [lo, hi) of its B row that falls into the warp's column range
                int lo = 0, hi = 0;
                if (e0 + lane < a1) {
                    const int j = a_colidx[e0 + lane];
                    if (j < b_nbr) {
                        const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
                        int l = b0, h = b1;
                        while (l < h) { const int m = (l + h) >> 1; if (b_colidx[m] < c_lo) l = m + 1; else h = m; }
                        lo = l;
                        h = b1;
                        if (c_hi < nbc) { while (l < h) { const int m = (l + h) >> 1; if (b_colidx[m] < c_hi) l = m + 1; else h = m; } hi = l; }
                        else hi = b1;
                    }
                }
                unsigned todo = __ballot_sync(0xffffffffu, lo < hi); // entries with a non-empty piece, ascending
                while (todo) {
                    const int ee = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const int plo = __shfl_sync(0xffffffffu, lo, ee), phi = __shfl_sync(0xffffffffu, hi, ee);
                    const int p = e0 + ee;
                    for (int q = plo + lane; q < phi; q += 32) {
                        const int c = b_colidx[q];
                        const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
                        const int pos = cur[t];
                        cur[t] = pos + 1;
                        pairs[pbase + pos] = make_int2(p, q);
                    }
                    __syncwarp(); // the next entry may append to the same tiles
                }
            }
        }
    }
    __threadfence_block();
    __syncthreads();
    // G. the cursors are the tile ends now: flag the last pair of every tile (sign bit of .y)
    for (int t = tid; t < T; t += BIG_THREADS) pairs[pbase + cur[t] - 1].y |= (int)0x80000000;
}

// ---- fill, global-memory variant: rows whose pair list does not fit in shared memory ----------
// Same algorithm with the per-tile counters/cursors kept in the row's own pair_ptr slots: slot t
// holds the count, then the first free slot of tile t; after the placement it holds the END of
// tile t, and a final in-row shift turns ends into starts.  A row touches only its own slots.
constexpr int FILLG_THREADS = 512;

__global__ void __launch_bounds__(FILLG_THREADS)
sym_fill_global_kernel(const int32_t *__restrict__ rowlist, const int32_t *__restrict__ a_rowptr,
                       const int32_t *__restrict__ a_colidx, int b_nbr, const int32_t *__restrict__ b_rowptr,
                       const int32_t *__restrict__ b_colidx, int words, const int32_t *__restrict__ c_rowptr,
                       const int64_t *__restrict__ row_prod_ptr, int32_t *__restrict__ c_colidx,
                       unsigned long long *pair_ptr, int2 *__restrict__ pairs)
{
    extern __shared__ uint32_t smem[];
    __shared__ int sm33[33];
    uint32_t *bm = smem;
    uint32_t *wrank = bm + words;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = FILLG_THREADS / 32;
    const int row = rowlist[blockIdx.x];
    const int cbase = c_rowptr[row], cend = c_rowptr[row + 1], T = cend - cbase;
    const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
    for (int w = tid; w < words; w += FILLG_THREADS) bm[w] = 0u;
    for (int t = tid; t < T; t += FILLG_THREADS) pair_ptr[(size_t)cbase + t] = 0ull;
    __syncthreads();
    for (int p = a0 + warp; p < a1; p += nwarp) {
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
        for (int q = b0 + lane; q < b1; q += 32) {
            const int c = b_colidx[q];
            atomicOr(&bm[c >> 5], 1u << (c & 31));
        }
    }
    __syncthreads();
    int carry = 0;
    for (int w0 = 0; w0 < words; w0 += FILLG_THREADS) {
        const int w = w0 + tid;
        const uint32_t bits = w < words ? bm[w] : 0u;
        int tot;
        const int ex = block_excl_scan_i32(__popc(bits), sm33, &tot);
        if (w < words) {
            wrank[w] = (uint32_t)(carry + ex);
            uint32_t b = bits;
            int o = cbase + carry + ex;
            while (b) {
                const int bit = __ffs(b) - 1;
                b &= b - 1;
                c_colidx[o++] = w * 32 + bit;
            }
        }
        carry += tot;
    }
    __syncthreads();
    for (int p = a0 + warp; p < a1; p += nwarp) {
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
        for (int q = b0 + lane; q < b1; q += 32) {
            const int c = b_colidx[q];
            const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
            atomicAdd(&pair_ptr[(size_t)cbase + t], 1ull);
        }
    }
    __threadfence_block();
    __syncthreads();
    // largest pair count of a tile (decides how the pairs are put in order below)
    __shared__ int maxcnt_sh;
    if (tid == 0) maxcnt_sh = 0;
    __syncthreads();
    {
        int m = 0;
        for (int t = tid; t < T; t += FILLG_THREADS) m = max(m, (int)pair_ptr[(size_t)cbase + t]);
        m = max(m, __shfl_xor_sync(0xffffffffu, m, 16)); m = max(m, __shfl_xor_sync(0xffffffffu, m, 8));
        m = max(m, __shfl_xor_sync(0xffffffffu, m, 4)); m = max(m, __shfl_xor_sync(0xffffffffu, m, 2));
        m = max(m, __shfl_xor_sync(0xffffffffu, m, 1));
        if (lane == 0) atomicMax(&maxcnt_sh, m);
    }
    __syncthreads();
    const bool few = maxcnt_sh <= FILL_SORT_MAX;
    // in-place exclusive scan of the row's counts by warp 0 (rows this large are rare)
    if (warp == 0) {
        unsigned long long run = (unsigned long long)row_prod_ptr[row];
        for (int t0 = 0; t0 < T; t0 += 32) {
            const int t = t0 + lane;
            const unsigned long long v = t < T ? pair_ptr[(size_t)cbase + t] : 0ull;
            unsigned long long inc = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            if (t < T) pair_ptr[(size_t)cbase + t] = run + inc - v;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    __threadfence_block();
    __syncthreads();
    if (few) {
        // every tile has few pairs: place everything at once (atomic cursor bump, any order),
        // then order each tile's pairs by A index (= ascending j) with a tiny insertion sort
        for (int p = a0 + warp; p < a1; p += nwarp) {
            const int j = a_colidx[p];
            if (j >= b_nbr) continue;
            const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
            for (int q = b0 + lane; q < b1; q += 32) {
                const int c = b_colidx[q];
                const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
                const unsigned long long pos = atomicAdd(&pair_ptr[(size_t)cbase + t], 1ull);
                pairs[pos] = make_int2(p, q);
            }
        }
        __threadfence_block();
        __syncthreads();
        for (int t = tid; t < T; t += FILLG_THREADS) {
            const long long e1 = (long long)((volatile unsigned long long *)pair_ptr)[(size_t)cbase + t];
            const long long s1 = t ? (long long)((volatile unsigned long long *)pair_ptr)[(size_t)cbase + t - 1]
                                   : (long long)row_prod_ptr[row];
            for (long long i = s1 + 1; i < e1; i++) {
                const int2 v = pairs[i];
                long long jx = i - 1;
                while (jx >= s1 && pairs[jx].x > v.x) { pairs[jx + 1] = pairs[jx]; jx--; }
                pairs[jx + 1] = v;
            }
        }
        __threadfence_block();
        __syncthreads();
    } else {
        for (int p = a0; p < a1; p++) {
            const int j = a_colidx[p];
            if (j >= b_nbr) continue;
            const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
            for (int q = b0 + tid; q < b1; q += FILLG_THREADS) {
                const int c = b_colidx[q];
                const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
                volatile unsigned long long *cu = &pair_ptr[(size_t)cbase + t];
                const unsigned long long pos = *cu;
                *cu = pos + 1ull;
                pairs[pos] = make_int2(p, q);
            }
            __threadfence_block();
            __syncthreads();
        }
    }
    // slots now hold tile ends: mark the last pair of every tile (sign bit of .y) ...
    for (int t = tid; t < T; t += FILLG_THREADS) {
        const unsigned long long e = ((volatile unsigned long long *)pair_ptr)[(size_t)cbase + t];
        pairs[e - 1].y |= (int)0x80000000;
    }
    __syncthreads();
    // ... then shift by one inside the row (last chunk first) to turn ends into starts
    const int nchunk = (T + FILLG_THREADS - 1) / FILLG_THREADS;
    for (int ch = nchunk - 1; ch >= 0; ch--) {
        const int t = ch * FILLG_THREADS + tid;
        unsigned long long prev = 0ull;
        if (t < T) prev = t == 0 ? (unsigned long long)row_prod_ptr[row]
                                 : ((volatile unsigned long long *)pair_ptr)[(size_t)cbase + t - 1];
        __syncthreads();
        if (t < T) pair_ptr[(size_t)cbase + t] = prev;
        __threadfence_block();
        __syncthreads();
    }
}

// ---- fill, rows with many pairs per output tile: bit matrix instead of ordered placement ----------
// Chain levels (banded factors: every tile of a row of C folds 17..105 pairs) spend their time in the
// ordered placement above: one barrier per A entry, half the CTA idle, 6..50 products per ns.  For a
// row with T output tiles and nA entries in its A row this kernel keeps the T x nA bit matrix
// M[t][e] = "A entry e contributes to tile t" in shared memory.  Every product sets its bit (any
// order); the pairs of tile t are then its set bits in ascending e, which IS ascending inner index j,
// so a product's slot is a popcount prefix and the last pair of a tile is its highest bit.  No per-
// product shared memory (pairs go straight to global memory), so the row may be of any length: it
// serves rows of every class whose matrix fits (two sizes, DENSE_M0 / DENSE_M1 words).
constexpr int DENSE_THREADS = 512;
constexpr int DENSE_M0 = 6 * 1024;   // 24 KB: several CTAs per SM
constexpr int DENSE_M1 = 40 * 1024;  // 160 KB: one CTA per SM
constexpr int DENSE_MIN_RATIO = 8;   // rows with at least this many tile products per output tile

__host__ __device__ inline int dense_row_words(int nA) { return ((nA + 31) / 32) | 1; } // odd stride: no bank conflicts

// shared-memory words a row needs next to the column bitmap: the T x W bit matrix and T+1 tile starts
__host__ __device__ inline int dense_need_words(int T, int nA) { return T * (dense_row_words(nA) + 1) + 1; }

__host__ __device__ inline size_t dense_smem_bytes(int words, int mcap)
{
    return 8 * (size_t)words + 4 * (size_t)mcap + 16; // bitmap + ranks, then M and the tile starts
}

__global__ void __launch_bounds__(DENSE_THREADS)
sym_fill_dense_kernel(const int32_t *__restrict__ rowlist, const int32_t *__restrict__ a_rowptr,
                      const int32_t *__restrict__ a_colidx, int b_nbr, const int32_t *__restrict__ b_rowptr,
                      const int32_t *__restrict__ b_colidx, int words, int mcap,
                      const int32_t *__restrict__ c_rowptr, const int64_t *__restrict__ row_prod_ptr,
                      int32_t *__restrict__ c_colidx, int64_t *__restrict__ pair_ptr, int2 *__restrict__ pairs)
{
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ int sm33[33];
    uint32_t *bm = smem;
    uint32_t *wrank = bm + words;
    uint32_t *M = wrank + words;          // [T * W]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = DENSE_THREADS / 32;
    const int row = rowlist[blockIdx.x];
    const int cbase = c_rowptr[row], T = c_rowptr[row + 1] - cbase;
    const int64_t pbase = row_prod_ptr[row];
    const int a0 = a_rowptr[row], nA = a_rowptr[row + 1] - a0;
    const int W = dense_row_words(nA);
    int *tstart = (int *)(M + T * W);     // [T + 1]; T * (W + 1) + 1 <= mcap by classification
    (void)mcap;
    for (int w = tid; w < words; w += DENSE_THREADS) bm[w] = 0u;
    for (int x = tid; x < T * W; x += DENSE_THREADS) M[x] = 0u;
    __syncthreads();
    // 1. column bitmap of the row
    for (int e = warp; e < nA; e += nwarp) {
        const int j = a_colidx[a0 + e];
        if (j >= b_nbr) continue;
        for_each_tile_of_row(b_colidx, b_rowptr[j], b_rowptr[j + 1], lane,
                             [&](int, int c) { atomicOr(&bm[c >> 5], 1u << (c & 31)); });
    }
    __syncthreads();
    // 2. rank table + sorted colidx of the C row
    int carry = 0;
    for (int w0 = 0; w0 < words; w0 += DENSE_THREADS) {
        const int w = w0 + tid;
        const uint32_t bits = w < words ? bm[w] : 0u;
        int tot;
        const int ex = block_excl_scan_i32(__popc(bits), sm33, &tot);
        if (w < words) {
            wrank[w] = (uint32_t)(carry + ex);
            uint32_t b = bits;
            int o = cbase + carry + ex;
            while (b) {
                const int bit = __ffs(b) - 1;
                b &= b - 1;
                c_colidx[o++] = w * 32 + bit;
            }
        }
        carry += tot;
    }
    __syncthreads();
    // 3. every product sets its bit
    for (int e = warp; e < nA; e += nwarp) {
        const int j = a_colidx[a0 + e];
        if (j >= b_nbr) continue;
        const uint32_t ebit = 1u << (e & 31);
        const int ew = e >> 5;
        for_each_tile_of_row(b_colidx, b_rowptr[j], b_rowptr[j + 1], lane, [&](int, int c) {
            const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
            atomicOr(&M[t * W + ew], ebit);
        });
    }
    __syncthreads();
    // 4. pairs per tile = popcount of its row of M; scan -> first slot of every tile
    carry = 0;
    for (int t0 = 0; t0 < T; t0 += DENSE_THREADS) {
        const int t = t0 + tid;
        int cnt = 0;
        if (t < T)
            for (int w = 0; w < W; w++) cnt += __popc(M[t * W + w]);
        int tot;
        const int ex = block_excl_scan_i32(cnt, sm33, &tot);
        if (t < T) {
            tstart[t] = carry + ex;
            pair_ptr[(size_t)cbase + t] = pbase + carry + ex;
        }
        carry += tot;
    }
    if (tid == 0) tstart[T] = carry;
    __syncthreads();
    // 5. slot of a product = first slot of its tile + number of set bits below its own
    for (int e = warp; e < nA; e += nwarp) {
        const int j = a_colidx[a0 + e];
        if (j >= b_nbr) continue;
        const uint32_t below = (1u << (e & 31)) - 1u;
        const int ew = e >> 5;
        for_each_tile_of_row(b_colidx, b_rowptr[j], b_rowptr[j + 1], lane, [&](int q, int c) {
            const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
            const uint32_t *Mt = M + t * W;
            int r = __popc(Mt[ew] & below);
            for (int w = 0; w < ew; w++) r += __popc(Mt[w]);
            const int pos = tstart[t] + r;
            const int last = (pos + 1 == tstart[t + 1]) ? (int)0x80000000 : 0;
            pairs[pbase + pos] = make_int2(a0 + e, q | last);
        });
    }
}

// ---- fill, very long rows: several CTAs per row, one phase per launch --------------------------------
// sym_fill_global_kernel gives a row that does not fit in shared memory ONE CTA, which handles 0.3-0.4
// tile products per ns: fine for 10^4..10^5 products per row (all rows run side by side), but an R-MAT
// hub row or a dense chain level (10^6..10^7 products in one row) becomes the tail of the whole pass.
// Such rows (more than hv_min products) take this path instead: the same steps as separate launches,
// those that touch every product spread over the whole GPU, one warp per (row, A entry):
//   hv_pattern   one CTA per row: column bitmap -> rank table (both kept in global memory), colidx of C
//   hv_products<false>  counts per output tile (global atomics)
//   hv_scan      one CTA per row: counts -> first slot of every tile (pair_ptr and the cursor array)
//   hv_products<true>   every product takes the next free slot of its tile (atomic cursor): any order
//   hv_sort_small / hv_sort_big   each tile's pairs sorted by A index, which is ascending inner block
//                index j because the A row is sorted -- the order the arithmetic needs; the last pair
//                of a tile gets its end flag here
// It moves every pair twice (scattered 8-byte stores, then the sort), about 2x the one-CTA kernel's
// total cost, so it is used only where that kernel's latency, not its throughput, is the problem.
constexpr int HV_THREADS = 256;
constexpr int HV_SORT_CAP = 8192; // pairs of one tile sorted in shared memory (64 KB); longer lists are sorted in place

__global__ void __launch_bounds__(HV_THREADS)
hv_pattern_kernel(const int32_t *__restrict__ rowlist, const int32_t *__restrict__ a_rowptr,
                  const int32_t *__restrict__ a_colidx, int b_nbr, const int32_t *__restrict__ b_rowptr,
                  const int32_t *__restrict__ b_colidx, int words, const int32_t *__restrict__ c_rowptr,
                  int32_t *__restrict__ c_colidx, uint32_t *__restrict__ gbm, uint32_t *__restrict__ grank,
                  int32_t *__restrict__ ent_cnt)
{
    extern __shared__ uint32_t bm[];
    __shared__ int sm33[33];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = HV_THREADS / 32;
    const int h = blockIdx.x, row = rowlist[h];
    for (int w = tid; w < words; w += HV_THREADS) bm[w] = 0u;
    __syncthreads();
    const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
    for (int p = a0 + warp; p < a1; p += nwarp) {
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
        for (int q = b0 + lane; q < b1; q += 32) {
            const int c = b_colidx[q];
            atomicOr(&bm[c >> 5], 1u << (c & 31));
        }
    }
    __syncthreads();
    const int cbase = c_rowptr[row];
    int carry = 0;
    for (int w0 = 0; w0 < words; w0 += HV_THREADS) {
        const int w = w0 + tid;
        const uint32_t bits = w < words ? bm[w] : 0u;
        int tot;
        const int ex = block_excl_scan_i32(__popc(bits), sm33, &tot);
        if (w < words) {
            gbm[(size_t)h * words + w] = bits;
            grank[(size_t)h * words + w] = (uint32_t)(carry + ex);
            uint32_t b = bits;
            int o = cbase + carry + ex;
            while (b) {
                const int bit = __ffs(b) - 1;
                b &= b - 1;
                c_colidx[o++] = w * 32 + bit;
            }
        }
        carry += tot;
    }
    if (tid == 0) ent_cnt[h] = a1 - a0;
}

// one warp per (heavy row, A entry), lanes over the tiles of B row j
template <bool PLACE>
__global__ void __launch_bounds__(HV_THREADS)
hv_products_kernel(const int32_t *__restrict__ rowlist, int nheavy, const int32_t *__restrict__ ent_ptr,
                   const int32_t *__restrict__ a_rowptr, const int32_t *__restrict__ a_colidx, int b_nbr,
                   const int32_t *__restrict__ b_rowptr, const int32_t *__restrict__ b_colidx, int words,
                   const uint32_t *__restrict__ gbm, const uint32_t *__restrict__ grank,
                   const int32_t *__restrict__ c_rowptr, unsigned long long *__restrict__ cur, int2 *__restrict__ pairs)
{
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * HV_THREADS + threadIdx.x) >> 5, nw = (gridDim.x * HV_THREADS) >> 5;
    const int total = ent_ptr[nheavy];
    for (int item = gw; item < total; item += nw) {
        int lo = 0, hi = nheavy; // heavy row of the item: last h with ent_ptr[h] <= item
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (ent_ptr[mid] <= item) lo = mid; else hi = mid;
        }
        const int row = rowlist[lo];
        const int p = a_rowptr[row] + (item - ent_ptr[lo]);
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
        const size_t cbase = (size_t)c_rowptr[row];
        const uint32_t *bmh = gbm + (size_t)lo * words, *rkh = grank + (size_t)lo * words;
        for (int q = b0 + lane; q < b1; q += 32) {
            const int c = b_colidx[q];
            const int t = (int)rkh[c >> 5] + __popc(bmh[c >> 5] & ((1u << (c & 31)) - 1u));
            const unsigned long long pos = atomicAdd(&cur[cbase + t], 1ull);
            if (PLACE) pairs[pos] = make_int2(p, q);
        }
    }
}

__global__ void __launch_bounds__(HV_THREADS)
hv_scan_kernel(const int32_t *__restrict__ rowlist, const int32_t *__restrict__ c_rowptr,
               const int64_t *__restrict__ row_prod_ptr, unsigned long long *__restrict__ cur,
               int64_t *__restrict__ pair_ptr)
{
    __shared__ int sm33[33];
    const int tid = threadIdx.x;
    const int row = rowlist[blockIdx.x];
    const int cbase = c_rowptr[row], T = c_rowptr[row + 1] - cbase;
    const int64_t pbase = row_prod_ptr[row];
    int carry = 0;
    for (int t0 = 0; t0 < T; t0 += HV_THREADS) {
        const int t = t0 + tid;
        const int v = t < T ? (int)cur[(size_t)cbase + t] : 0;
        int tot;
        const int ex = block_excl_scan_i32(v, sm33, &tot);
        if (t < T) {
            const int64_t start = pbase + carry + ex;
            pair_ptr[(size_t)cbase + t] = start;
            cur[(size_t)cbase + t] = (unsigned long long)start;
        }
        carry += tot;
    }
}

// tiles with at most 32 pairs: one warp per tile, bitonic network in registers; longer ones are listed
__global__ void __launch_bounds__(HV_THREADS)
hv_sort_small_kernel(const int32_t *__restrict__ rowlist, const int32_t *__restrict__ c_rowptr,
                     const int64_t *__restrict__ pair_ptr, const unsigned long long *__restrict__ cur,
                     int2 *__restrict__ pairs, int32_t *__restrict__ big_list, int32_t *__restrict__ big_count)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = HV_THREADS / 32;
    const int row = rowlist[blockIdx.x];
    const int cbase = c_rowptr[row], T = c_rowptr[row + 1] - cbase;
    for (int t = blockIdx.y * nwarp + warp; t < T; t += gridDim.y * nwarp) {
        const int64_t s0 = pair_ptr[(size_t)cbase + t];
        const int L = (int)((int64_t)cur[(size_t)cbase + t] - s0);
        if (L > 32) {
            if (lane == 0) big_list[atomicAdd(big_count, 1)] = cbase + t;
            continue;
        }
        int2 v = make_int2(0x7fffffff, 0);
        if (lane < L) v = pairs[s0 + lane];
        if (L > 1) {
#pragma unroll
            for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
                for (int stride = size >> 1; stride > 0; stride >>= 1) {
                    const int ox = __shfl_xor_sync(0xffffffffu, v.x, stride);
                    const int oy = __shfl_xor_sync(0xffffffffu, v.y, stride);
                    const bool up = (lane & size) == 0;          // ascending block
                    const bool lower = (lane & stride) == 0;     // this lane keeps the smaller of the two
                    const bool take = (lower == up) ? (ox < v.x) : (ox > v.x);
                    if (take) { v.x = ox; v.y = oy; }
                }
            }
        }
        if (lane == L - 1) v.y |= (int)0x80000000;
        if (lane < L) pairs[s0 + lane] = v;
    }
}

// longer lists: one CTA per tile, bitonic sort in shared memory, or in a per-CTA global scratch area of
// `scratch_stride` pairs when a list exceeds HV_SORT_CAP (a tile with more than 8192 pairs needs an A
// row with more than 8192 entries; the host sizes the scratch from the column count of A)
__global__ void __launch_bounds__(HV_THREADS)
hv_sort_big_kernel(const int32_t *__restrict__ big_list, const int32_t *__restrict__ big_count,
                   const int64_t *__restrict__ pair_ptr, const unsigned long long *__restrict__ cur,
                   int2 *pairs, int2 *scratch, size_t scratch_stride)
{
    extern __shared__ __align__(16) uint32_t hv_smem[];
    const int tid = threadIdx.x;
    const int n = *big_count;
    for (int i = blockIdx.x; i < n; i += gridDim.x) {
        const int tile = big_list[i];
        const int64_t s0 = pair_ptr[tile];
        const int L = (int)((int64_t)cur[tile] - s0);
        int N = 64;
        while (N < L) N <<= 1;
        int2 *a = N <= HV_SORT_CAP ? (int2 *)hv_smem : scratch + (size_t)blockIdx.x * scratch_stride;
        for (int x = tid; x < N; x += HV_THREADS) a[x] = x < L ? pairs[s0 + x] : make_int2(0x7fffffff, 0);
        __syncthreads();
        for (int size = 2; size <= N; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int x = tid; x < N / 2; x += HV_THREADS) {
                    const int lo = ((x / stride) * stride * 2) + (x % stride), hi = lo + stride;
                    const bool up = (lo & size) == 0;
                    const int2 u = a[lo], w = a[hi];
                    if ((u.x > w.x) == up) { a[lo] = w; a[hi] = u; }
                }
                __syncthreads();
            }
        }
        for (int x = tid; x < L; x += HV_THREADS) {
            int2 v = a[x];
            if (x == L - 1) v.y |= (int)0x80000000;
            pairs[s0 + x] = v;
        }
        __syncthreads();
    }
}

// ---- fill, tiny rows: one WARP per block row, no bitmap ------------------------------------------
// A row with at most TINY_P tile products and TINY_A entries in its A row is sorted instead of
// bit-mapped: every product gets the key (block column << 5 | ordinal of its A entry); a bitonic
// sort of the <= 256 keys in shared memory groups the products by output tile with ascending j
// inside a tile, which IS the pair list; tile heads are found with a ballot, their ranks with
// a popcount prefix.  This is the common case of chain levels with few factors, of most R-MAT
// rows and of small inputs, where a 512-thread CTA per row mostly waits at barriers.
constexpr int TINY_P = 256;
constexpr int TINY_A = 32;
constexpr int TINY_WARPS = 8;

__global__ void __launch_bounds__(TINY_WARPS * 32)
sym_fill_tiny_kernel(const int32_t *__restrict__ rowlist, int nrows, const int32_t *__restrict__ a_rowptr,
                     const int32_t *__restrict__ a_colidx, int b_nbr, const int32_t *__restrict__ b_rowptr,
                     const int32_t *__restrict__ b_colidx, const int32_t *__restrict__ c_rowptr,
                     const int64_t *__restrict__ row_prod_ptr, int32_t *__restrict__ c_colidx,
                     int64_t *__restrict__ pair_ptr, int2 *__restrict__ pairs)
{
    __shared__ uint32_t keys_sh[TINY_WARPS][TINY_P];
    __shared__ int32_t vals_sh[TINY_WARPS][TINY_P];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *keys = keys_sh[warp];
    int32_t *vals = vals_sh[warp];
    for (int li = blockIdx.x * TINY_WARPS + warp; li < nrows; li += gridDim.x * TINY_WARPS) {
        const int row = rowlist[li];
        const int cbase = c_rowptr[row];
        const int64_t pbase = row_prod_ptr[row];
        const int P = (int)(row_prod_ptr[row + 1] - pbase);
        const int a0 = a_rowptr[row], nA = a_rowptr[row + 1] - a0;
        // lane e holds A entry e: first B tile and number of products
        int b0 = 0, len = 0;
        if (lane < nA) {
            const int j = a_colidx[a0 + lane];
            if (j < b_nbr) { b0 = b_rowptr[j]; len = b_rowptr[j + 1] - b0; }
        }
        int tot;
        const int off = warp_excl_scan(len, &tot);
        // keys: (block column, A entry ordinal); values: index of the B tile
        for (int e = 0; e < nA; e++) {
            const int eb0 = __shfl_sync(0xffffffffu, b0, e), elen = __shfl_sync(0xffffffffu, len, e);
            const int eoff = __shfl_sync(0xffffffffu, off, e);
            for (int i = lane; i < elen; i += 32) {
                keys[eoff + i] = ((uint32_t)b_colidx[eb0 + i] << 5) | (uint32_t)e;
                vals[eoff + i] = eb0 + i;
            }
        }
        int N = 32;
        while (N < P) N <<= 1;
        for (int x = P + lane; x < N; x += 32) keys[x] = 0xffffffffu;
        __syncwarp();
        // bitonic sort of N (<= 256) keys with their values
        for (int size = 2; size <= N; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int x = lane; x < N / 2; x += 32) {
                    const int lo = ((x / stride) * stride * 2) + (x % stride), hi = lo + stride;
                    const bool up = ((lo & size) == 0);
                    const uint32_t k0 = keys[lo], k1 = keys[hi];
                    if ((k0 > k1) == up) {
                        keys[lo] = k1; keys[hi] = k0;
                        const int32_t v0 = vals[lo];
                        vals[lo] = vals[hi]; vals[hi] = v0;
                    }
                }
                __syncwarp();
            }
        }
        // tile heads, ranks, outputs
        int tiles_before = 0;
        for (int x0 = 0; x0 < P; x0 += 32) {
            const int x = x0 + lane;
            const bool in = x < P;
            const uint32_t key = in ? keys[x] : 0xffffffffu;
            const bool head = in && (x == 0 || (keys[x - 1] >> 5) != (key >> 5));
            const bool last = in && (x == P - 1 || (keys[x + 1] >> 5) != (key >> 5));
            const unsigned hm = __ballot_sync(0xffffffffu, head);
            if (head) {
                const int t = tiles_before + __popc(hm & ((1u << lane) - 1u));
                c_colidx[cbase + t] = (int32_t)(key >> 5);
                pair_ptr[(size_t)cbase + t] = pbase + x;
            }
            if (in) pairs[pbase + x] = make_int2(a0 + (int)(key & 31u), vals[x] | (last ? (int)0x80000000 : 0));
            tiles_before += __popc(hm);
        }
        __syncwarp();
    }
}

// ---- rows -> capacity classes -----------------------------------------------------------------
// class c rows go to lists[c * nbr ...]; counts[c] = rows in class c.  A row with P tile
// products fits class c when P <= caps[c]; larger than every cap -> class NCLASS-1 (global).
constexpr int NCLASS = 9; // 8: long rows, one CTA per row with per-warp tile ranges (counters in shared memory); 6, 7: rows with many pairs per tile whose bit matrix fits DENSE_M0 / DENSE_M1 words;
                          // 0: tiny rows (warp + sort), 1..3: shared-memory CTA classes, 4: global-memory CTA,
                          // 5: rows so long that one CTA would be the tail of the pass -> multi-launch path

__global__ void classify_rows_kernel(int nbr, const int32_t *__restrict__ a_rowptr,
                                     const int32_t *__restrict__ row_cnt, const int32_t *__restrict__ row_prod,
                                     int cap0, int cap1, int cap2, int tcap0, int tcap1, int tcap2, int acap,
                                     int tiny_ok, int hv_min, int dense_ok, int big_tcap, int32_t *__restrict__ lists,
                                     int32_t *__restrict__ counts)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nbr || row_cnt[row] == 0) return;
    const int P = row_prod[row], T = row_cnt[row];
    const int nA = a_rowptr[row + 1] - a_rowptr[row];
    int cls = (P <= cap0 && T <= tcap0) ? 1 : ((P <= cap1 && T <= tcap1) ? 2 : ((P <= cap2 && T <= tcap2) ? 3 : 4));
    if (nA > acap) cls = 4;                                  // A row too long for the shared-memory entry table
    if (tiny_ok && P <= TINY_P && nA <= TINY_A) cls = 0;     // warp-per-row sort
    // spread over the whole GPU -- unless the row's counters fit the one-CTA kernel with per-warp tile ranges, which
    // is the cheaper of the two up to a few million products (R-MAT hub row, 3.3e5 products: 0.5 ms there, 1.0 ms here)
    if (cls == 4 && P > hv_min && (T > big_tcap || P > 8 * (long long)hv_min)) cls = 5;
    if (dense_ok && cls != 0 && P >= dense_ok * T) {   // dense_ok = minimum tile products per output tile  // many pairs per tile: bit-matrix placement
        const int mw = T <= DENSE_M1 ? dense_need_words(T, nA) : DENSE_M1 + 1;
        if (mw <= DENSE_M0) cls = 6;
        else if (mw <= DENSE_M1) cls = 7;
    }
    if (cls == 4 && T <= big_tcap) cls = 8;                  // counters fit in shared memory: per-warp tile ranges
    if (cls >= 1 && cls <= 3) { // what the class really needs: its launch is sized from these
        atomicMax(&counts[NCLASS + 2 + (cls - 1)], P);
        atomicMax(&counts[NCLASS + 5 + (cls - 1)], T);
    }
    const int slot = atomicAdd(&counts[cls], 1);
    lists[(size_t)cls * nbr + slot] = row;
}

// multi-launch fill of the rows that do not fit the shared-memory classes (see "heavy rows" above)
int fill_heavy_rows(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const int32_t *rowlist, int nheavy, int words,
                    const bsp_symbolic_out *out, const int64_t *row_prod_ptr)
{
    bsp_scratch sc(ctx);
    unsigned long long *cur = nullptr;
    uint32_t *gbm = nullptr, *grank = nullptr;
    int32_t *ent_cnt = nullptr, *ent_ptr = nullptr, *big_list = nullptr, *big_count = nullptr;
    int2 *scratch = nullptr;
    // heavy rows are processed in batches so that their bitmaps and rank tables stay below 256 MB
    int hb = (int)(((size_t)256 << 20) / ((size_t)words * 8));
    if (hb < 1) hb = 1;
    if (hb > nheavy) hb = nheavy;
    int64_t big_cap = out->nprod / 33 + 1;
    if (big_cap > out->nnzb_c) big_cap = out->nnzb_c;
    BSP_TRY(sc.alloc(&cur, (size_t)out->nnzb_c + 1));
    BSP_TRY(sc.alloc(&gbm, (size_t)hb * words));
    BSP_TRY(sc.alloc(&grank, (size_t)hb * words));
    BSP_TRY(sc.alloc(&ent_cnt, (size_t)hb + 1));
    BSP_TRY(sc.alloc(&ent_ptr, (size_t)hb + 1));
    BSP_TRY(sc.alloc(&big_list, (size_t)big_cap));
    BSP_TRY(sc.alloc(&big_count, 1));
    const int grid_big = ctx->sm_count * 2;
    size_t scratch_stride = 0;
    {
        size_t n = 64;
        while (n < (size_t)A->nbc) n <<= 1; // a tile cannot have more pairs than the A row has entries
        if (n > (size_t)HV_SORT_CAP) {
            scratch_stride = n;
            BSP_TRY(sc.alloc(&scratch, scratch_stride * grid_big));
        }
    }
    BSP_CUDA(ctx, cudaMemsetAsync(cur, 0, ((size_t)out->nnzb_c + 1) * sizeof(unsigned long long), ctx->stream));
    BSP_CUDA(ctx, cudaFuncSetAttribute(hv_sort_big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HV_SORT_CAP * 8));
    const int grid = ctx->sm_count * 8;
    for (int h0 = 0; h0 < nheavy; h0 += hb) {
        const int n = nheavy - h0 < hb ? nheavy - h0 : hb;
        const int32_t *rl = rowlist + h0;
        BSP_CUDA(ctx, cudaMemsetAsync(ent_cnt + n, 0, sizeof(int32_t), ctx->stream));
        BSP_CUDA(ctx, cudaMemsetAsync(big_count, 0, sizeof(int32_t), ctx->stream));
        hv_pattern_kernel<<<n, HV_THREADS, (size_t)words * sizeof(uint32_t), ctx->stream>>>(
            rl, A->rowptr, A->colidx, B->nbr, B->rowptr, B->colidx, words, out->c_rowptr, out->c_colidx, gbm, grank, ent_cnt);
        BSP_LAUNCH_CHECK(ctx);
        BSP_TRY(bsp_scan_i32(ctx, ent_cnt, ent_ptr, (int64_t)n + 1, nullptr));
        hv_products_kernel<false><<<grid, HV_THREADS, 0, ctx->stream>>>(rl, n, ent_ptr, A->rowptr, A->colidx, B->nbr, B->rowptr,
                                                                       B->colidx, words, gbm, grank, out->c_rowptr, cur, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
        hv_scan_kernel<<<n, HV_THREADS, 0, ctx->stream>>>(rl, out->c_rowptr, row_prod_ptr, cur, out->pair_ptr);
        BSP_LAUNCH_CHECK(ctx);
        hv_products_kernel<true><<<grid, HV_THREADS, 0, ctx->stream>>>(rl, n, ent_ptr, A->rowptr, A->colidx, B->nbr, B->rowptr,
                                                                      B->colidx, words, gbm, grank, out->c_rowptr, cur, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
        int y = ctx->sm_count * 16 / n;
        if (y < 1) y = 1;
        if (y > 64) y = 64;
        hv_sort_small_kernel<<<dim3((unsigned)n, (unsigned)y), HV_THREADS, 0, ctx->stream>>>(rl, out->c_rowptr, out->pair_ptr, cur,
                                                                                            out->pairs, big_list, big_count);
        BSP_LAUNCH_CHECK(ctx);
        hv_sort_big_kernel<<<grid_big, HV_THREADS, HV_SORT_CAP * 8, ctx->stream>>>(big_list, big_count, out->pair_ptr, cur,
                                                                                  out->pairs, scratch, scratch_stride);
        BSP_LAUNCH_CHECK(ctx);
    }
    return BSP_OK;
}

} // namespace

int bsp_symbolic(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t row_lo, int32_t row_hi,
                 bsp_symbolic_out *out)
{
    const int64_t pair_budget = out->pair_budget; // the one input field
    memset(out, 0, sizeof(*out));
    out->pair_budget = pair_budget;
    const int nbr = A->nbr;
    const int words = (B->nbc + 31) / 32 > 0 ? (B->nbc + 31) / 32 : 1;
    const size_t smem_count = (size_t)SYM_WARPS * words * sizeof(uint32_t);
    const size_t bitmap_bytes = 2 * (size_t)words * sizeof(uint32_t);
    if (smem_count > 200 * 1024)
        BSP_FAIL(ctx, BSP_ERR_LIMIT,
                 "symbolic pass: block-column space %d needs %zu B of shared memory per CTA (limit 204800 B, "
                 "i.e. at most 204800 block columns)", B->nbc, smem_count);

    int32_t *row_cnt = nullptr, *row_prod = nullptr, *lists = nullptr, *counts = nullptr, *heavy = nullptr;
    int64_t *row_prod_ptr = nullptr;
    bsp_scratch sc(ctx); // releases everything below on any return; the outputs are handed over at the end
    BSP_TRY(sc.alloc(&row_cnt, (size_t)nbr + 1));
    BSP_TRY(sc.alloc(&row_prod, (size_t)nbr + 1));
    BSP_TRY(sc.alloc(&row_prod_ptr, (size_t)nbr + 1));
    BSP_TRY(sc.alloc(&lists, (size_t)NCLASS * (nbr > 0 ? nbr : 1)));
    BSP_TRY(sc.alloc(&counts, (size_t)NCLASS + 8)); // [NCLASS] = heavy rows of the count pass, [NCLASS+1] = per-row overflow flag,
                                                    // [NCLASS+2 .. +4] / [NCLASS+5 .. +7] = most products / tiles of a row in classes 1..3
    BSP_TRY(sc.alloc(&heavy, (size_t)(nbr > 0 ? nbr : 1)));
    BSP_TRY(sc.alloc(&out->c_rowptr, (size_t)nbr + 1));
    BSP_CUDA(ctx, cudaMemsetAsync(row_cnt + nbr, 0, sizeof(int32_t), ctx->stream));
    BSP_CUDA(ctx, cudaMemsetAsync(row_prod + nbr, 0, sizeof(int32_t), ctx->stream));
    BSP_CUDA(ctx, cudaMemsetAsync(counts, 0, (NCLASS + 8) * sizeof(int32_t), ctx->stream));

    int grid = (nbr + SYM_WARPS - 1) / SYM_WARPS;
    const int max_grid = ctx->sm_count * 8;
    if (grid > max_grid) grid = max_grid;
    if (grid < 1) grid = 1;
    // capacity classes of the shared-memory fill (tile products / output tiles of a row):
    // about 8 / 2 / 1 CTAs per SM; larger rows use the global-memory variant
    // (with 4 B per pair in shared memory: three CTAs per SM up to 7680 products / 5632 tiles, two up to 12288 / 8192;
    // every class is then launched with the capacity its own rows need -- their maxima come back with the totals)
    const int acap = FILL_ACAP;
    int caps[3] = {1536, 7680, 12288}, tcaps[3] = {1536, 5632, 8192};
    if (const char *e = getenv("BSP_SYM_CAPS")) { // test hook: force rows into the larger classes
        int a, b, c;
        if (sscanf(e, "%d,%d,%d", &a, &b, &c) == 3 && a > 0 && a <= b && b <= c) {
            caps[0] = tcaps[0] = a; caps[1] = tcaps[1] = b; caps[2] = tcaps[2] = c;
        }
    }
    // A row above hv_min tile products is not given to one CTA (0.3-0.4 products/ns: 1 ms per 3*10^5) but to the
    // multi-launch path, whose cost is spread over the GPU; below it the one-CTA kernel is cheaper in total
    // (measured: chain top product, 66 k products per row: 3.3 ms against 38 ms).  BSP_HV_MIN_P is a test hook.
    int hv_min = 1 << 18;
    if (const char *e = getenv("BSP_HV_MIN_P")) { const int v = atoi(e); if (v > 0) hv_min = v; }
    // 0 = rows keep their capacity class (BSP_SYM_NO_DENSE, a comparison hook), else the minimum number of tile
    // products per output tile for the bit-matrix fill (BSP_SYM_DENSE_RATIO overrides the default)
    int dense_ok = getenv("BSP_SYM_NO_DENSE") ? 0 : DENSE_MIN_RATIO;
    if (dense_ok)
        if (const char *e = getenv("BSP_SYM_DENSE_RATIO")) { const int v = atoi(e); if (v > 0) dense_ok = v; }
    // the tiny-row kernel packs the block column into 27 bits; BSP_SYM_CAPS (a test hook) also switches it off
    const int tiny_ok = (B->nbc < (1 << 27) && !getenv("BSP_SYM_CAPS")) ? 1 : 0;
    // long rows with at most big_tcap output tiles keep their per-tile counters in shared memory (two CTAs per SM);
    // BSP_SYM_BIG=0 sends them to the global-memory kernel instead (comparison / test hook)
    int big_tcap = 0;
    if (8 * (size_t)words + 64 < 110 * 1024) big_tcap = (int)((110 * 1024 - 8 * (size_t)words) / 4) - 8;
    if (const char *e = getenv("BSP_SYM_BIG")) { if (atoi(e) == 0) big_tcap = 0; }
    if (B->nbc > (1 << FILL_IBITS)) caps[0] = caps[1] = caps[2] = tcaps[0] = tcaps[1] = tcaps[2] = 0; // packed pair: 22-bit offset in a B row
    for (int c = 0; c < 3; c++) {
        if (caps[c] > 60000) caps[c] = 60000; // tile ranks are kept as 16-bit values in shared memory
        if (tcaps[c] > caps[c]) tcaps[c] = caps[c];
        while (caps[c] > 2 && fill_smem_bytes(words, caps[c], tcaps[c], acap) > 220 * 1024) {
            caps[c] /= 2;
            if (tcaps[c] > caps[c]) tcaps[c] = caps[c];
        }
        caps[c] &= ~1; tcaps[c] &= ~1; // keeps the int2 arrays 8-byte aligned
    }
    if (nbr > 0) {
        BSP_CUDA(ctx, cudaFuncSetAttribute(sym_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem_count));
        sym_count_kernel<<<grid, SYM_THREADS, smem_count, ctx->stream>>>(
            nbr, A->rowptr, A->colidx, B->nbr, B->rowptr, B->colidx, words, row_lo, row_hi, row_cnt, row_prod, heavy,
            counts + NCLASS, counts + NCLASS + 1);
        BSP_LAUNCH_CHECK(ctx);
        {
            const size_t hb = (size_t)words * sizeof(uint32_t);
            BSP_CUDA(ctx, cudaFuncSetAttribute(sym_count_heavy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hb));
            int hgrid = nbr < ctx->sm_count * 4 ? nbr : ctx->sm_count * 4; // CTAs loop over the (device-side) list
            sym_count_heavy_kernel<<<hgrid, HEAVY_THREADS, hb, ctx->stream>>>(heavy, counts + NCLASS, A->rowptr, A->colidx,
                                                                              B->nbr, B->rowptr, B->colidx, words, row_cnt);
            BSP_LAUNCH_CHECK(ctx);
        }
        classify_rows_kernel<<<(nbr + 255) / 256, 256, 0, ctx->stream>>>(nbr, A->rowptr, row_cnt, row_prod, caps[0],
                                                                        caps[1], caps[2], tcaps[0], tcaps[1], tcaps[2],
                                                                        acap, tiny_ok, hv_min, dense_ok, big_tcap, lists, counts);
        BSP_LAUNCH_CHECK(ctx);
    }
    bsp_trace_mark(ctx, "sym_count");
    BSP_TRY(bsp_scan_i32(ctx, row_cnt, out->c_rowptr, (int64_t)nbr + 1, nullptr));
    BSP_TRY(bsp_scan_i32_to_i64(ctx, row_prod, row_prod_ptr, (int64_t)nbr + 1));

    // totals -> host (the one unavoidable sync of the pass: C and the pair lists are sized from them); a kernel
    // store into mapped pinned memory, not a copy -- see bsp_publish
    int32_t *h_nnzb = (int32_t *)ctx->pinned;
    int64_t *h_nprod = (int64_t *)((char *)ctx->pinned + 8);
    int32_t *h_counts = (int32_t *)((char *)ctx->pinned + 128); // clear of the numeric pass's counter at +64
    {
        bsp_publish_list pl = {};
        pl.n = 3;
        pl.src[0] = out->c_rowptr + nbr; pl.offset[0] = 0;  pl.words[0] = 1;
        pl.src[1] = row_prod_ptr + nbr;  pl.offset[1] = 8;  pl.words[1] = 2;
        pl.src[2] = counts;              pl.offset[2] = 128; pl.words[2] = NCLASS + 8;
        BSP_TRY(bsp_publish(ctx, pl));
    }
    bsp_trace_mark(ctx, "sym_scans");
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    out->nnzb_c = *h_nnzb;
    out->nprod = *h_nprod;
    int ncls[NCLASS];
    for (int c = 0; c < NCLASS; c++) ncls[c] = h_counts[c];
    int cmax_p[3], cmax_t[3]; // most tile products / output tiles of a row in each shared-memory class
    for (int c = 0; c < 3; c++) { cmax_p[c] = h_counts[NCLASS + 2 + c]; cmax_t[c] = h_counts[NCLASS + 5 + c]; }
    if (h_counts[NCLASS + 1])
        BSP_FAIL(ctx, BSP_ERR_LIMIT, "a block row of the product has more than 2^31-1 tile products");
    if (out->nnzb_c < 0) BSP_FAIL(ctx, BSP_ERR_LIMIT, "product has more than 2^31-1 tiles");
    if (out->nprod > 0x7fffffffLL) {
        // nnzb(C) <= nprod, so the int32 row-pointer scan can only have wrapped past zero in this case:
        // total the per-row tile counts again in 64 bits
        int64_t *cnt64 = nullptr;
        BSP_TRY(sc.alloc(&cnt64, (size_t)nbr + 1));
        BSP_TRY(bsp_scan_i32_to_i64(ctx, row_cnt, cnt64, (int64_t)nbr + 1));
        {
            bsp_publish_list pl = {};
            pl.n = 1;
            pl.src[0] = cnt64 + nbr; pl.offset[0] = 8; pl.words[0] = 2;
            BSP_TRY(bsp_publish(ctx, pl));
        }
        BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (*h_nprod > 0x7fffffffLL) BSP_FAIL(ctx, BSP_ERR_LIMIT, "product has more than 2^31-1 tiles");
    }
    out->n_split = 0;
    out->split = nullptr;
    if (out->pair_budget > 0 && out->nprod > out->pair_budget) {
        // too many tile products for one pair list: hand back row panels that fit (greedy over the
        // per-row product counts, which are already on the device as a prefix sum)
        std::vector<int64_t> pp((size_t)nbr + 1);
        BSP_CUDA(ctx, cudaMemcpyAsync(pp.data(), row_prod_ptr, ((size_t)nbr + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
        BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        const int lo = row_lo < 0 ? 0 : row_lo, hi = row_hi > nbr ? nbr : row_hi;
        std::vector<int32_t> cuts;
        cuts.push_back(lo);
        int start = lo;
        for (int r = lo; r < hi; r++) {
            if (pp[r + 1] - pp[start] > out->pair_budget && r > start) { cuts.push_back(r); start = r; }
        }
        cuts.push_back(hi);
        if (cuts.size() > 2) {
            out->n_split = (int32_t)cuts.size() - 1;
            out->split = (int32_t *)malloc(cuts.size() * sizeof(int32_t));
            if (!out->split) BSP_FAIL(ctx, BSP_ERR_NOMEM, "symbolic pass: panel list");
            memcpy(out->split, cuts.data(), cuts.size() * sizeof(int32_t));
            out->c_rowptr = nullptr; // released by the scratch guard
            out->c_colidx = nullptr; out->pair_ptr = nullptr; out->pairs = nullptr;
            return BSP_OK;
        }
    }

    BSP_TRY(sc.alloc(&out->c_colidx, (size_t)out->nnzb_c));
    BSP_TRY(sc.alloc(&out->pair_ptr, (size_t)out->nnzb_c + 1));
    BSP_TRY(sc.alloc(&out->pairs, (size_t)out->nprod));
    // every row writes its own slots; the closing entry pair_ptr[nnzb_c] = nprod comes from the row scan
    BSP_CUDA(ctx, cudaMemcpyAsync(out->pair_ptr + out->nnzb_c, row_prod_ptr + nbr, sizeof(int64_t),
                                  cudaMemcpyDeviceToDevice, ctx->stream));
    bsp_trace_mark(ctx, "sym_sync_alloc");
    if (ncls[0] > 0) {
        int g = (ncls[0] + TINY_WARPS - 1) / TINY_WARPS;
        if (g > ctx->sm_count * 16) g = ctx->sm_count * 16;
        sym_fill_tiny_kernel<<<g, TINY_WARPS * 32, 0, ctx->stream>>>(lists, ncls[0], A->rowptr, A->colidx, B->nbr, B->rowptr,
                                                                    B->colidx, out->c_rowptr, row_prod_ptr, out->c_colidx,
                                                                    out->pair_ptr, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
    }
    for (int c = 0; c < 3; c++) {
        if (ncls[c + 1] == 0) continue;
        // capacity the rows of this class need (even: keeps the arrays behind aligned), not the class bound
        int pcap = (cmax_p[c] + 65) & ~63, tcap = (cmax_t[c] + 65) & ~63;
        if (pcap > caps[c] || getenv("BSP_SYM_CAPS")) pcap = caps[c];
        if (tcap > tcaps[c] || getenv("BSP_SYM_CAPS")) tcap = tcaps[c];
        if (pcap < 2) pcap = 2;
        if (tcap < 2) tcap = 2;
        const size_t bytes = fill_smem_bytes(words, pcap, tcap, acap);
        const char *spe = getenv("BSP_SYM_PROF");
        const bool sprof = spe && spe[0] == '1';
        auto fill_kernel = sprof ? sym_fill_smem_kernel<true> : sym_fill_smem_kernel<false>;
        BSP_CUDA(ctx, cudaFuncSetAttribute(fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        if (sprof) {
            unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            BSP_CUDA(ctx, cudaMemcpyToSymbolAsync(sym_fill_prof, zero, sizeof(zero), 0, cudaMemcpyHostToDevice, ctx->stream));
        }
        fill_kernel<<<ncls[c + 1], FILL_THREADS, bytes, ctx->stream>>>(
            lists + (size_t)(c + 1) * nbr, A->rowptr, A->colidx, B->nbr, B->rowptr, B->colidx, words, pcap, tcap, acap,
            out->c_rowptr, row_prod_ptr, out->c_colidx, out->pair_ptr, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
        if (sprof) {
            unsigned long long h[8];
            BSP_CUDA(ctx, cudaMemcpyFromSymbolAsync(h, sym_fill_prof, sizeof(h), 0, cudaMemcpyDeviceToHost, ctx->stream));
            BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            static const char *names[8] = {"init+A_entries", "columns+bitmap", "ranks+sorted_cols", "tile_of_product+counts", "scan+pair_ptr",
                                           "placement", "order+flags", "stream_out"};
            fprintf(stderr, "[bsp sym prof] class %d: %d rows, smem %zu B, cycles per row:", c + 1, ncls[c + 1], bytes);
            for (int i = 0; i < 8; i++) fprintf(stderr, " %s=%.0f", names[i], (double)h[i] / ncls[c + 1]);
            fprintf(stderr, "\n");
        }
    }
    if (ncls[4] > 0) {
        BSP_CUDA(ctx, cudaFuncSetAttribute(sym_fill_global_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)bitmap_bytes));
        sym_fill_global_kernel<<<ncls[4], FILLG_THREADS, bitmap_bytes, ctx->stream>>>(
            lists + (size_t)4 * nbr, A->rowptr, A->colidx, B->nbr, B->rowptr, B->colidx, words, out->c_rowptr,
            row_prod_ptr, out->c_colidx, (unsigned long long *)out->pair_ptr, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
    }
    for (int c = 6; c <= 7; c++) {
        if (ncls[c] == 0) continue;
        const int mcap = c == 6 ? DENSE_M0 : DENSE_M1;
        const size_t bytes = dense_smem_bytes(words, mcap);
        BSP_CUDA(ctx, cudaFuncSetAttribute(sym_fill_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        sym_fill_dense_kernel<<<ncls[c], DENSE_THREADS, bytes, ctx->stream>>>(
            lists + (size_t)c * nbr, A->rowptr, A->colidx, B->nbr, B->rowptr, B->colidx, words, mcap, out->c_rowptr,
            row_prod_ptr, out->c_colidx, out->pair_ptr, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
    }
    if (ncls[8] > 0) {
        const size_t bytes = big_smem_bytes(words, big_tcap);
        BSP_CUDA(ctx, cudaFuncSetAttribute(sym_fill_big_kernel<BIG_WARPS_LONG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        sym_fill_big_kernel<BIG_WARPS_LONG><<<ncls[8], BIG_WARPS_LONG * 32, bytes, ctx->stream>>>(lists + (size_t)8 * nbr, A->rowptr, A->colidx, B->nbr,
                                                                         B->rowptr, B->colidx, B->nbc, words, out->c_rowptr,
                                                                         row_prod_ptr, out->c_colidx, out->pair_ptr, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
    }
    if (ncls[5] > 0) BSP_TRY(fill_heavy_rows(ctx, A, B, lists + (size_t)5 * nbr, ncls[5], words, out, row_prod_ptr));
    bsp_trace_mark(ctx, "sym_fill");
    sc.keep(out->c_rowptr); sc.keep(out->c_colidx); sc.keep(out->pair_ptr); sc.keep(out->pairs); // now the caller's
    return BSP_OK;
}
