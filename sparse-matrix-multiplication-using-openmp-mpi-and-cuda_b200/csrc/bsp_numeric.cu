// bsp_numeric.cu -- numeric pass of C = A*B: multiply-accumulate of k x k uint64 tiles.
//
// Replaces matrix_multiplyKernel (cu:44-66) and the host loop around it that stages tile
// pairs in rounds of 500 output keys (cu:181-270).  Output-stationary: the threads that own
// an output tile walk its pair list (built by the symbolic pass in ascending inner block
// index) and gather the A and B tiles by index straight from the CSR-of-blocks arrays, so no
// tile is ever copied into a staging buffer and no atomics are needed.
//
// Arithmetic (cu:59, cu:61), per output element, pairs in list order, inner index ascending:
//     p   = phi(a * b mod 2^64)        phi(x) = x % (2^64-1) = (x == 2^64-1 ? 0 : x)
//     acc = phi(acc + p mod 2^64)
// phi is not a ring homomorphism, so the fold order is part of the result (tests/golden t7/t8).
//
// Fast path (k = 4, 8, 16).  phi changes a value only when an intermediate equals 2^64-1.
// As long as that has not happened, the running value equals the plain wrapping sum
// s' = sum a*b mod 2^64 (3 integer multiply-adds per step instead of ~12 instructions).  The
// kernel therefore accumulates s' and, for every step, folds a conservative detector into one
// register:   flag = max(flag, lo32(s'_new), lo32(s'_new) - lo32(s'_old))
// lo32(s'_new) - lo32(s'_old) is the low word of the product, so flag reaches 0xffffffff
// whenever the product or the new sum has an all-ones low word -- a superset of "equals
// 2^64-1".  By induction s' is exact up to the first such step; if the flag fired anywhere in
// a thread's row of an output tile (probability ~2^-31 per step on random data, certain on the
// edge-case tests), that row is recomputed with the per-operation arithmetic in reference
// order.  The result is bit-exact in every case.
//
// Work distribution.  A group of K threads owns one output tile at a time, thread r holding
// row r of the tile (K accumulators).  Work is cut in the PAIR index space, not the tile
// space: chunk g owns the tiles whose first pair falls in [g*Q, (g+1)*Q), so every group
// performs the same number of tile products whatever the distribution of pairs per tile
// (1.5 on average for 1 % density, thousands in a dense chain level, power-law for R-MAT).
// All groups of a warp step through their pair lists in lockstep: one tile product per step.
// Per step a thread loads its A row (K u64) into registers and the group stages the B tile in
// shared memory with coalesced 128-bit loads (double buffered, one __syncwarp per step); the
// next step's A row and B chunk are prefetched into registers before the multiply-adds of the
// current one.
#include <cstdlib>

#include "bsp_internal.h"

namespace {

__device__ __forceinline__ uint64_t phi(uint64_t x) { return x == ~0ull ? 0ull : x; }

// exact per-operation fold of one output element
__device__ __forceinline__ uint64_t fold_exact(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                                               const int2 *__restrict__ pairs, int64_t p0, int64_t p1, int k, int r,
                                               int c)
{
    const size_t kk = (size_t)k * k;
    uint64_t acc = 0;
    for (int64_t p = p0; p < p1; p++) {
        const int2 pr = pairs[p];
        const uint64_t *A = a_vals + (size_t)pr.x * kk + (size_t)r * k;
        const uint64_t *B = b_vals + (size_t)(pr.y & 0x7fffffff) * kk + c; // sign bit = last pair of the tile
        for (int j = 0; j < k; j++) {
            const uint64_t t = phi(A[j] * B[(size_t)j * k]);
            acc = phi(acc + t);
        }
    }
    return acc;
}

// ---- any tile edge k: exact arithmetic throughout -------------------------------------------
// 256-thread CTAs; when k*k <= 256 a CTA works on 256/(k*k) output tiles at a time with one
// thread per element, otherwise on one tile with threads striding over its elements.
constexpr int GEN_THREADS = 256;

__global__ void __launch_bounds__(GEN_THREADS)
numeric_generic_kernel(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                       const int64_t *__restrict__ pair_ptr, const int2 *__restrict__ pairs, int64_t nnzb_c, int k,
                       uint64_t *__restrict__ c_vals)
{
    const int E = k * k;
    const int slots = E <= GEN_THREADS ? GEN_THREADS / E : 1;
    const int slot = E <= GEN_THREADS ? threadIdx.x / E : 0;
    const int e0 = E <= GEN_THREADS ? threadIdx.x % E : threadIdx.x;
    if (slot >= slots) return;
    for (int64_t t = (int64_t)blockIdx.x * slots + slot; t < nnzb_c; t += (int64_t)gridDim.x * slots) {
        const int64_t p0 = pair_ptr[t], p1 = pair_ptr[t + 1];
        for (int e = e0; e < E; e += GEN_THREADS) {
            c_vals[(size_t)t * E + e] = fold_exact(a_vals, b_vals, pairs, p0, p1, k, e / k, e % k);
            if (E <= GEN_THREADS) break;
        }
    }
}

// ---- fast path ------------------------------------------------------------------------------

// s += a*b (mod 2^64) and the conservative 2^64-1 detector on the low words.
// Kept as a plain 64-bit multiply-add: nvcc lowers it to IMAD.WIDE.U32 onto the accumulator
// + two IMAD for the cross terms + one add into the high word, and the detector to
// VIADDMNMX (max(lo - lo_old, lo)) + half a VIMNMX3 -- 5.5 integer instructions per step.
// (Hand-split 32-bit halves, separate cross-term accumulators and PTX mad.wide chains were all
// tried: ptxas re-associates them into longer sequences with extra register moves.)
__device__ __forceinline__ void mac_flag(uint64_t &s, uint32_t &flag, uint64_t a, uint64_t b)
{
    const uint32_t lo_old = (uint32_t)s;
    s += a * b;
    const uint32_t lo = (uint32_t)s;
    flag = max(flag, max(lo, lo - lo_old));
}

template <int K>
__device__ __noinline__ void refold_row_exact(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                                              const int2 *__restrict__ pairs, int64_t p0, int64_t p1, int r,
                                              uint64_t *out_row)
{
    for (int c = 0; c < K; c++) out_row[c] = fold_exact(a_vals, b_vals, pairs, p0, p1, K, r, c);
}

// -- chunk schedule ---------------------------------------------------------------------------
// The block columns of C are cut into `npanels` panels so that the part of B a panel touches
// (|B| / npanels) stays resident in the 126 MB L2 while all block rows are swept; a segment is
// (panel, block row) = a contiguous range of output tiles; a chunk is at most Q consecutive tile
// products of one segment (whole tiles).  Chunks are numbered panel-major, so the CTAs that
// run together gather from the same panel of B.

__device__ __forceinline__ int lower_bound_i32(const int32_t *__restrict__ a, int lo, int hi, int key)
{
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ int lower_bound_i64(const int64_t *__restrict__ a, int lo, int hi, int64_t key)
{
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void seg_bounds_kernel(int nbr, int npanels, int panel_w, const int32_t *__restrict__ c_rowptr,
                                  const int32_t *__restrict__ c_colidx, const int64_t *__restrict__ pair_ptr, int Q,
                                  int32_t *__restrict__ seg_ts, int32_t *__restrict__ seg_nchunks)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int nseg = nbr * npanels;
    if (x > nseg) return;
    if (x == nseg) { seg_nchunks[x] = 0; seg_ts[x] = c_rowptr[nbr]; return; }
    const int pnl = x / nbr, row = x - pnl * nbr;
    const int r0 = c_rowptr[row], r1 = c_rowptr[row + 1];
    const int ts = pnl == 0 ? r0 : lower_bound_i32(c_colidx, r0, r1, pnl * panel_w);
    const int te = pnl == npanels - 1 ? r1 : lower_bound_i32(c_colidx, r0, r1, (pnl + 1) * panel_w);
    seg_ts[x] = ts;
    const int64_t np = pair_ptr[te] - pair_ptr[ts];
    seg_nchunks[x] = (int)((np + Q - 1) / Q);
}

// logarithmic size class of a chunk of n tile products: 8 classes per octave, 0 for an empty chunk
__device__ __forceinline__ unsigned char chunk_bucket(int64_t n)
{
    if (n <= 0) return 0;
    if (n >= ((int64_t)1 << 31)) return 255;
    const unsigned v = (unsigned)n;
    const int l = 31 - __clz(v);
    const int sub = l >= 3 ? (int)((v >> (l - 3)) & 7u) : 0;
    return (unsigned char)min(255, l * 8 + sub + 1);
}

__global__ void chunk_fill_kernel(int nbr, int npanels, int panel_w, const int32_t *__restrict__ c_rowptr,
                                  const int32_t *__restrict__ c_colidx, const int64_t *__restrict__ pair_ptr, int Q,
                                  const int32_t *__restrict__ seg_ts, const int32_t *__restrict__ seg_off,
                                  int2 *__restrict__ chunk_tiles, unsigned char *__restrict__ chunk_key)
{
    const int nseg = nbr * npanels;
    const int total = seg_off[nseg];
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= total) return;
    // segment of chunk x: last seg with seg_off[seg] <= x
    int lo = 0, hi = nseg;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (seg_off[mid] <= x) lo = mid; else hi = mid;
    }
    const int seg = lo, l = x - seg_off[seg], n = seg_off[seg + 1] - seg_off[seg];
    const int pnl = seg / nbr, row = seg - pnl * nbr;
    const int ts = seg_ts[seg];
    const int r1 = c_rowptr[row + 1];
    const int te = pnl == npanels - 1 ? r1 : lower_bound_i32(c_colidx, ts, r1, (pnl + 1) * panel_w);
    const int64_t base = pair_ptr[ts];
    const int t0 = l == 0 ? ts : lower_bound_i64(pair_ptr, ts, te, base + (int64_t)l * Q);
    const int t1 = l == n - 1 ? te : lower_bound_i64(pair_ptr, ts, te, base + (int64_t)(l + 1) * Q);
    chunk_tiles[x] = make_int2(t0, t1);
    if (chunk_key) chunk_key[x] = chunk_bucket(pair_ptr[t1] - pair_ptr[t0]);
}

// -- chunks ordered by size ---------------------------------------------------------------------
// The groups of a warp step through their chunks in lockstep, so a warp is as slow as its longest chunk.
// A chunk is "whole tiles until Q products are reached": with a skewed pattern (R-MAT: 3 % of the products
// sit in tiles of more than 96 pairs, one tile has 1887) the eight chunks of a warp differ widely and the
// lockstep wastes 17 % of the lanes' steps over the whole product, 30 % in the hub rows.  The chunks are
// therefore put in descending order of a logarithmic size class (8 classes per octave) by a counting
// sort: warps get chunks of nearly equal length (98 % of the steps useful in the same simulation) and the
// long ones start first.  Inside a class the order of the rows is kept block by block, so chunks of one
// block row of C still run side by side and share the A tiles in L1/L2.
constexpr int CHUNK_CLASSES = 256;
constexpr int CSORT_THREADS = 256;

__global__ void __launch_bounds__(CSORT_THREADS)
chunk_hist_kernel(const unsigned char *__restrict__ key, const int32_t *__restrict__ total_ptr, int32_t *__restrict__ hist)
{
    __shared__ int h[CHUNK_CLASSES];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int x = blockIdx.x * CSORT_THREADS + threadIdx.x;
    if (x < *total_ptr) atomicAdd(&h[key[x]], 1);
    __syncthreads();
    if (h[threadIdx.x]) atomicAdd(&hist[threadIdx.x], h[threadIdx.x]);
}

__global__ void __launch_bounds__(CSORT_THREADS)
chunk_scatter_kernel(const unsigned char *__restrict__ key, const int2 *__restrict__ chunk_tiles,
                     const int32_t *__restrict__ total_ptr, const int32_t *__restrict__ hist, int32_t *__restrict__ cursor,
                     int2 *__restrict__ sorted)
{
    __shared__ int base[CHUNK_CLASSES], cnt[CHUNK_CLASSES], resv[CHUNK_CLASSES];
    const int tid = threadIdx.x;
    // first slot of class b in descending class order = chunks in the classes above b
    {
        int v = hist[CHUNK_CLASSES - 1 - tid]; // tid-th largest class
        base[tid] = v;
        __syncthreads();
        for (int d = 1; d < CHUNK_CLASSES; d <<= 1) { // inclusive scan over 256 entries
            const int o = tid >= d ? base[tid - d] : 0;
            __syncthreads();
            base[tid] += o;
            __syncthreads();
        }
        const int excl = base[tid] - v;
        __syncthreads();
        base[tid] = excl; // indexed by position in descending order: class b sits at CHUNK_CLASSES-1-b
    }
    cnt[tid] = 0;
    __syncthreads();
    const int x = blockIdx.x * CSORT_THREADS + tid;
    const bool in = x < *total_ptr;
    int b = 0, local = 0;
    if (in) { b = key[x]; local = atomicAdd(&cnt[b], 1); }
    __syncthreads();
    if (cnt[tid]) resv[tid] = atomicAdd(&cursor[tid], cnt[tid]);
    __syncthreads();
    if (in) sorted[base[CHUNK_CLASSES - 1 - b] + resv[b] + local] = chunk_tiles[x];
}


// -- the kernel -------------------------------------------------------------------------------

__device__ __forceinline__ void cp_async_ca16(void *smem_dst, const void *gsrc)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_cg16(void *smem_dst, const void *gsrc)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// The multiply-add with the 64-bit accumulator SPLIT: s collects the low x low partial products (one
// IMAD.WIDE.U32 with the accumulator as addend) and a separate 32-bit x collects the two cross terms (two
// IMAD with x as addend); the true value is s + (x << 32), formed once per output tile.  nvcc's own
// lowering of s += a*b (MACV = 0, mac_flag) needs a fourth instruction per step, the add of the cross
// terms into the high word, which ptxas often issues as IMAD.IADD on the multiplier pipe.  The detector
// only looks at low words, which x never touches, so it is unchanged.  The low product is spelled
// mul.wide.u32 in PTX: written as C++ 64-bit arithmetic nvcc emits a 64 x 64 multiply of zero-extended
// values and ptxas leaves additions of zero behind; written as mad.wide.u32 ptxas splits the accumulate
// into IMAD.WIDE + IADD3 + IADD3.X.  configs[1]: 7.68 -> 6.82 warp instructions per warp-MAC, 11.47 ->
// 11.01 ms under ncu; the chain's k=4 top product 5.83 -> 5.36 ms.
template <int MACV>
__device__ __forceinline__ void mac_split(uint64_t &s, uint32_t &x, uint32_t &flag, uint64_t a, uint64_t b)
{
    if (MACV == 0) { mac_flag(s, flag, a, b); return; }
    const uint32_t al = (uint32_t)a, ah = (uint32_t)(a >> 32), bl = (uint32_t)b, bh = (uint32_t)(b >> 32);
    const uint32_t lo_old = (uint32_t)s;
    uint64_t t;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(t) : "r"(al), "r"(bl));
    s += t;
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(x) : "r"(al), "r"(bh));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(x) : "r"(ah), "r"(bl));
    const uint32_t lo = (uint32_t)s;
    flag = max(flag, max(lo, lo - lo_old));
}

// A group of TPT threads owns one output tile at a time; thread h of the group holds rows
// h, h+TPT, ... (RPT = K/TPT rows, K accumulators each).  Fewer threads per tile means fewer
// redundant shared-memory reads of the B tile (every thread of a group needs all of it): the
// first version of this kernel used K threads per tile and spent as long in the shared-memory
// pipe (56 wavefronts per tile product) as in the integer pipe; K/2 threads per tile halve that.
template <int K>
struct RowsCfg {
    static constexpr int TPT = K / 2;                          // threads per output tile
    static constexpr int RPT = K / TPT;                        // rows per thread (= 2)
    static constexpr int GROUPS = 32 / TPT;                    // tiles in flight per warp
    static constexpr int TILE_BYTES = K * K * 8;
    static constexpr int CHUNKS = TILE_BYTES / 16;             // 16-byte chunks per tile (32 for k=8, 8 for k=4)
    static constexpr int TPI = CHUNKS >= 32 ? 1 : 32 / CHUNKS; // tiles moved by one warp-wide cp.async
    static constexpr int IPT = CHUNKS >= 32 ? CHUNKS / 32 : 1; // warp-wide cp.async instructions per tile
    static constexpr int GROUP_STRIDE = TILE_BYTES + 32;       // one tile per group; +32 B shifts the next group by 8 banks
    static constexpr int STAGE_BYTES = GROUPS * GROUP_STRIDE;  // per warp: one set of tiles (A: single buffer, B: STAGES buffers)
    __host__ __device__ static constexpr int WARP_BYTES(int stages) { return (1 + stages) * STAGE_BYTES; } // A + B stages
};

// position of 16-byte chunk i of row r inside the A tile: XOR swizzle so that the TPT threads
// of a group (rows h, h+TPT of 64 B each for k=8) read different bank quads with LDS.128
template <int K>
__device__ __forceinline__ int a_chunk_offset(int r, int i)
{
    if (K == 8) return r * 64 + (((i ^ (r >> 1)) & 3) << 4);
    return r * (K * 8) + i * 16;
}

// 16-byte asynchronous copies; nbytes = 16 copies, nbytes = 0 reads nothing and zero-fills
// (used for groups that have run out of tile products: no branch, no global traffic)
__device__ __forceinline__ void cp_async16_ca(unsigned smem_addr, const void *gsrc, unsigned nbytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_addr), "l"(gsrc), "r"(nbytes) : "memory");
}
__device__ __forceinline__ void cp_async16_cg(unsigned smem_addr, const void *gsrc, unsigned nbytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_addr), "l"(gsrc), "r"(nbytes) : "memory");
}
__device__ __forceinline__ ulonglong2 lds128(unsigned smem_addr)
{
    ulonglong2 v;
    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];\n" : "=l"(v.x), "=l"(v.y) : "r"(smem_addr));
    return v;
}

template <int K, int STAGES, int WARPS, int MINB, int MACV>
__global__ void __launch_bounds__(WARPS * 32, MINB)
numeric_rows_kernel(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                    const int64_t *__restrict__ pair_ptr, const int2 *__restrict__ pairs,
                    const int2 *__restrict__ chunk_tiles, const int32_t *__restrict__ nchunks_ptr,
                    uint64_t *__restrict__ c_vals, unsigned long long *__restrict__ refold_count)
{
    using Cfg = RowsCfg<K>;
    // the A tiles have ONE buffer (rows go to registers at step start): a deeper prologue would overwrite them
    static_assert(STAGES == 2, "numeric_rows_kernel: A tiles are single-buffered, only two B stages are valid");
    constexpr int TPT = Cfg::TPT, RPT = Cfg::RPT;
    constexpr int NCOPY = 2 * Cfg::GROUPS * Cfg::IPT / Cfg::TPI; // warp-wide cp.async instructions per step
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane / TPT, h = lane % TPT;
    // 32-bit shared-memory addresses, computed once
    const unsigned wbase = (unsigned)__cvta_generic_to_shared(smem) + (unsigned)warp * Cfg::WARP_BYTES(STAGES);

    const int nchunks = *nchunks_ptr;
    const int g = (blockIdx.x * WARPS + warp) * Cfg::GROUPS + grp;
    int t = 0, n = 0;        // current output tile, tile products (steps) of this group's chunk
    const int2 *prs = pairs; // this chunk's slice of the pair stream
    if (g < nchunks) {
        const int2 ct = chunk_tiles[g];
        if (ct.x < ct.y) {
            const int64_t p0 = pair_ptr[ct.x];
            t = ct.x;
            n = (int)(pair_ptr[ct.y] - p0);
            prs = pairs + p0;
        }
    }
    int steps = n;
#pragma unroll
    for (int d = 16; d; d >>= 1) steps = max(steps, __shfl_xor_sync(0xffffffffu, steps, d));
    if (steps == 0) return; // warp-uniform

    // What this lane moves in each warp-wide cp.async: chunk `cn` of a tile of group
    // (q0 + cq), to a fixed offset inside the stage (A tiles XOR-swizzled, B tiles as they are).
    const int cn = Cfg::TPI > 1 ? lane % Cfg::CHUNKS : lane;
    const int cq = Cfg::TPI > 1 ? lane / Cfg::CHUNKS : 0;
    const unsigned char *a_lane = reinterpret_cast<const unsigned char *>(a_vals) + cn * 16;
    const unsigned char *b_lane = reinterpret_cast<const unsigned char *>(b_vals) + cn * 16;
    unsigned dstA[Cfg::IPT], dstB[Cfg::IPT];
#pragma unroll
    for (int part = 0; part < Cfg::IPT; part++) {
        const int nn = cn + 32 * part;
        dstA[part] = cq * Cfg::GROUP_STRIDE + a_chunk_offset<K>(nn / (K / 2), nn % (K / 2));
        dstB[part] = cq * Cfg::GROUP_STRIDE + nn * 16;
    }

    // The whole warp copies the tiles of all its groups for one step into stage ST: every
    // cp.async instruction moves 512 contiguous bytes (one k=8 tile, or four k=4 tiles).
    // (ai, bi) = this group's pair for that step; a negative ai marks a group that has run out of
    // steps: its copies become zero-size (nothing is read; thousands of idle groups re-reading
    // one dummy tile would serialise on a single L2 line).
    auto issue = [&](int ST, int ai, int bi) {
        const unsigned abase = wbase, bbase = wbase + (1 + ST) * Cfg::STAGE_BYTES; // A: single buffer, B: stage ST
#pragma unroll
        for (int q0 = 0; q0 < Cfg::GROUPS; q0 += Cfg::TPI) {
            const int src_lane = (q0 + cq) * TPT;
            const int qa = __shfl_sync(0xffffffffu, ai, src_lane);
            const int qb = __shfl_sync(0xffffffffu, bi, src_lane);
            const unsigned nbytes = qa < 0 ? 0u : 16u;
            const size_t ia = (size_t)(unsigned)(qa & 0x7fffffff), ib = (size_t)(unsigned)qb;
#pragma unroll
            for (int part = 0; part < Cfg::IPT; part++) {
                cp_async16_ca(abase + q0 * Cfg::GROUP_STRIDE + dstA[part], a_lane + ia * Cfg::TILE_BYTES + part * 512, nbytes);
                cp_async16_cg(bbase + q0 * Cfg::GROUP_STRIDE + dstB[part], b_lane + ib * Cfg::TILE_BYTES + part * 512, nbytes);
            }
        }
    };
    (void)NCOPY;

    // prologue: fill STAGES-1 stages.  (nx, ny) = pair of the next step to be issued;
    // lastq bit i = "the pair of step k+i closes its tile" (sign bit set by the symbolic pass)
    // Pair loads are unconditional (index clamped to the chunk): a predicated load would make
    // the loop-carried registers wait for it on the spot instead of one step later.
    const int n_last = max(n - 1, 0);
    int nx, ny;
    unsigned lastq = 0;
    { const int2 pr = __ldcs(prs); nx = pr.x; ny = pr.y; }
#pragma unroll
    for (int st = 0; st < STAGES - 1; st++) {
        issue(st, st < n ? nx : (nx | (int)0x80000000), ny & 0x7fffffff);
        lastq |= (st < n && ny < 0 ? 1u : 0u) << st;
        { const int2 pr = __ldcs(prs + min(st + 1, n_last)); nx = pr.x; ny = pr.y; }
        cp_async_commit();
    }

    uint64_t acc[RPT][K];
    uint32_t xs[RPT][K]; // cross-term sums of the split form (MACV != 0), else unused
    uint32_t flag = 0;
#pragma unroll
    for (int rr = 0; rr < RPT; rr++)
#pragma unroll
        for (int c = 0; c < K; c++) { acc[rr][c] = 0; xs[rr][c] = 0; }

    // per-thread read addresses inside a stage: two A rows (swizzled chunks) and the B tile
    const unsigned gbase = wbase + grp * Cfg::GROUP_STRIDE;
    const unsigned hb16 = K == 8 ? ((h >> 1) & 1) * 16 : 0;

    auto step = [&](int ST, int k) {
        cp_async_wait<STAGES - 2>(); // this thread's copies for step k have landed
        __syncwarp();                // ... and everybody else's; everybody is done reading step k-1's B stage
        // A rows of this step go to registers right away: the A tiles have a single buffer
        // (12.5 instead of 16.5 KB of shared memory per warp -> one more CTA per SM)
        uint64_t a_cur[RPT][K];
        {
            const unsigned st = gbase;
#pragma unroll
            for (int rr = 0; rr < RPT; rr++) {
                const int r = h + TPT * rr;
                const unsigned row = st + r * (K * 8);
#pragma unroll
                for (int i = 0; i < K / 2; i++) {
                    // chunk i of row r sits at ((i ^ (r >> 1)) & 3) * 16 for k=8 (see a_chunk_offset);
                    // r >> 1 = (h >> 1) + 2*rr, so the offset is ((i ^ 2rr) * 16) ^ hb16
                    const unsigned off = K == 8 ? ((unsigned)(((i ^ (2 * rr)) & 3) * 16) ^ hb16) : (unsigned)(i * 16);
                    const ulonglong2 v = lds128(row + off);
                    a_cur[rr][2 * i] = v.x;
                    a_cur[rr][2 * i + 1] = v.y;
                }
            }
        }
        __syncwarp(); // everybody holds its A rows: the A buffer may be overwritten
        {
            const int kk = k + (STAGES - 1);
            issue((ST + STAGES - 1) % STAGES, kk < n ? nx : (nx | (int)0x80000000), ny & 0x7fffffff);
            lastq |= (kk < n && ny < 0 ? 1u : 0u) << (STAGES - 1);
            { const int2 pr = __ldcs(prs + min(kk + 1, n_last)); nx = pr.x; ny = pr.y; }
            cp_async_commit();
        }
        if (k < n) {
            const unsigned bsm = gbase + (1 + ST) * Cfg::STAGE_BYTES;
            // B row j+1 is fetched from shared memory while row j is being multiplied
            ulonglong2 bnext[K / 2];
#pragma unroll
            for (int c2 = 0; c2 < K / 2; c2++) bnext[c2] = lds128(bsm + c2 * 16);
#pragma unroll
            for (int j = 0; j < K; j++) {
                ulonglong2 bcur[K / 2];
#pragma unroll
                for (int c2 = 0; c2 < K / 2; c2++) bcur[c2] = bnext[c2];
                if (j + 1 < K) {
#pragma unroll
                    for (int c2 = 0; c2 < K / 2; c2++) bnext[c2] = lds128(bsm + ((j + 1) * (K / 2) + c2) * 16);
                }
#pragma unroll
                for (int c2 = 0; c2 < K / 2; c2++) {
#pragma unroll
                    for (int rr = 0; rr < RPT; rr++) {
                        mac_split<MACV>(acc[rr][2 * c2], xs[rr][2 * c2], flag, a_cur[rr][j], bcur[c2].x);
                        mac_split<MACV>(acc[rr][2 * c2 + 1], xs[rr][2 * c2 + 1], flag, a_cur[rr][j], bcur[c2].y);
                    }
                }
            }
            if (lastq & 1u) { // last pair of the tile: write this thread's rows, move on
                if (MACV != 0) {
#pragma unroll
                    for (int rr = 0; rr < RPT; rr++)
#pragma unroll
                        for (int c = 0; c < K; c++) acc[rr][c] += (uint64_t)xs[rr][c] << 32;
                }
                if (flag == 0xffffffffu) {
                    const int64_t p0 = pair_ptr[t], p1 = (prs - pairs) + k + 1;
#pragma unroll
                    for (int rr = 0; rr < RPT; rr++) {
                        uint64_t exact[K]; // only this copy lives in local memory, not the accumulators
                        refold_row_exact<K>(a_vals, b_vals, pairs, p0, p1, h + TPT * rr, exact);
#pragma unroll
                        for (int c = 0; c < K; c++) acc[rr][c] = exact[c];
                    }
                    atomicAdd(refold_count, (unsigned long long)RPT);
                }
#pragma unroll
                for (int rr = 0; rr < RPT; rr++) {
                    ulonglong2 *out = reinterpret_cast<ulonglong2 *>(c_vals + (size_t)t * (K * K) +
                                                                     (size_t)(h + TPT * rr) * K);
#pragma unroll
                    for (int i = 0; i < K / 2; i++)
                        __stcs(&out[i], make_ulonglong2(acc[rr][2 * i], acc[rr][2 * i + 1]));
#pragma unroll
                    for (int c = 0; c < K; c++) { acc[rr][c] = 0; xs[rr][c] = 0; }
                }
                flag = 0;
                t++;
            }
        }
        lastq >>= 1;
    };

    for (int k0 = 0; k0 < steps; k0 += STAGES) {
#pragma unroll
        for (int st = 0; st < STAGES; st++)
            if (k0 + st < steps) step(st, k0 + st); // warp-uniform
    }
    cp_async_wait<0>();
}

// ---- fast path for large tiles (k = 16, 32) --------------------------------------------------
// One thread per output-tile ROW (K accumulators in registers), 32/K tiles in flight per warp.
// Per step the warp stages the B tile(s) in shared memory with cp.async (double buffered) and
// every thread streams its own row of the A tile straight from global memory, two entries per
// 128-bit load, two loads ahead of use.  Same arithmetic, detector and exact refold as above.
template <int K>
struct BigCfg {
    static constexpr int GROUPS = 32 / K;                 // tiles in flight per warp (1 for k=32, 2 for k=16)
    static constexpr int TILE_BYTES = K * K * 8;
    static constexpr int CHUNKS = TILE_BYTES / 16;        // 16-byte chunks per tile
    static constexpr int STAGE_BYTES = GROUPS * TILE_BYTES;
};

template <int K, int WARPS, int MACV>
__global__ void __launch_bounds__(WARPS * 32)
numeric_bigtile_kernel(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                       const int64_t *__restrict__ pair_ptr, const int2 *__restrict__ pairs,
                       const int2 *__restrict__ chunk_tiles, const int32_t *__restrict__ nchunks_ptr,
                       uint64_t *__restrict__ c_vals, unsigned long long *__restrict__ refold_count)
{
    using Cfg = BigCfg<K>;
    constexpr int STAGES = 2;
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane / K, r = lane % K;
    const unsigned wbase = (unsigned)__cvta_generic_to_shared(smem) + (unsigned)warp * (STAGES * Cfg::STAGE_BYTES);

    const int nchunks = *nchunks_ptr;
    const int g = (blockIdx.x * WARPS + warp) * Cfg::GROUPS + grp;
    int t = 0, n = 0;
    const int2 *prs = pairs;
    if (g < nchunks) {
        const int2 ct = chunk_tiles[g];
        if (ct.x < ct.y) {
            const int64_t p0 = pair_ptr[ct.x];
            t = ct.x;
            n = (int)(pair_ptr[ct.y] - p0);
            prs = pairs + p0;
        }
    }
    int steps = n;
#pragma unroll
    for (int d = 16; d; d >>= 1) steps = max(steps, __shfl_xor_sync(0xffffffffu, steps, d));
    if (steps == 0) return;

    // the warp copies the B tiles of its groups for one step: every cp.async moves 512 contiguous bytes
    auto issue = [&](int ST, int bi, bool valid) {
        const unsigned sbase = wbase + ST * Cfg::STAGE_BYTES;
#pragma unroll
        for (int q = 0; q < Cfg::GROUPS; q++) {
            const int qb = __shfl_sync(0xffffffffu, bi, q * K);
            const unsigned nbytes = __shfl_sync(0xffffffffu, (int)valid, q * K) ? 16u : 0u;
            const unsigned char *src = reinterpret_cast<const unsigned char *>(b_vals) + (size_t)(unsigned)qb * Cfg::TILE_BYTES;
#pragma unroll
            for (int x = 0; x < Cfg::CHUNKS / 32; x++)
                cp_async16_cg(sbase + q * Cfg::TILE_BYTES + (x * 32 + lane) * 16, src + (x * 32 + lane) * 16, nbytes);
        }
    };

    const int n_last = max(n - 1, 0);
    int2 cur = __ldcs(prs); // pair of the current step, nxt = pair of the next one
    int2 nxt = __ldcs(prs + min(1, n_last));
    issue(0, cur.y & 0x7fffffff, n > 0);
    cp_async_commit();

    uint64_t acc[K];
    uint32_t xs[K]; // cross-term sums of the split multiply-add (mac_split)
    uint32_t flag = 0;
#pragma unroll
    for (int c = 0; c < K; c++) { acc[c] = 0; xs[c] = 0; }

    for (int k = 0; k < steps; k++) {
        cp_async_wait<0>();
        __syncwarp();
        issue((k + 1) & 1, nxt.y & 0x7fffffff, k + 1 < n);
        cp_async_commit();
        const int2 nn = __ldcs(prs + min(k + 2, n_last));
        if (k < n) {
            const unsigned bsm = wbase + (k & 1) * Cfg::STAGE_BYTES + grp * Cfg::TILE_BYTES;
            const ulonglong2 *arow = reinterpret_cast<const ulonglong2 *>(a_vals + (size_t)cur.x * (K * K) + (size_t)r * K);
            ulonglong2 a0 = arow[0], a1 = arow[1]; // two loads in flight ahead of use
            // k = 32: 1024 multiply-adds per step; unrolling all of them costs the compiler 200 s (one
            // 1024-long dependency chain through `flag`) and buys nothing over 128 per loop body
            constexpr int J2_UNROLL = K >= 32 ? 2 : K / 2;
#pragma unroll J2_UNROLL
            for (int j2 = 0; j2 < K / 2; j2++) {
                const ulonglong2 av = a0;
                a0 = a1;
                if (j2 + 2 < K / 2) a1 = arow[j2 + 2];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const uint64_t a = h ? av.y : av.x;
                    const int j = 2 * j2 + h;
#pragma unroll
                    for (int c2 = 0; c2 < K / 2; c2++) {
                        const ulonglong2 bv = lds128(bsm + (j * (K / 2) + c2) * 16); // broadcast inside the group
                        mac_split<MACV>(acc[2 * c2], xs[2 * c2], flag, a, bv.x);
                        mac_split<MACV>(acc[2 * c2 + 1], xs[2 * c2 + 1], flag, a, bv.y);
                    }
                }
            }
            if (cur.y < 0) { // last pair of the tile
                if (MACV != 0) {
#pragma unroll
                    for (int c = 0; c < K; c++) acc[c] += (uint64_t)xs[c] << 32;
                }
                if (flag == 0xffffffffu) {
                    uint64_t exact[K];
                    refold_row_exact<K>(a_vals, b_vals, pairs, pair_ptr[t], (prs - pairs) + k + 1, r, exact);
#pragma unroll
                    for (int c = 0; c < K; c++) acc[c] = exact[c];
                    atomicAdd(refold_count, 1ull);
                }
                ulonglong2 *out = reinterpret_cast<ulonglong2 *>(c_vals + (size_t)t * (K * K) + (size_t)r * K);
#pragma unroll
                for (int i = 0; i < K / 2; i++) __stcs(&out[i], make_ulonglong2(acc[2 * i], acc[2 * i + 1]));
#pragma unroll
                for (int c = 0; c < K; c++) { acc[c] = 0; xs[c] = 0; }
                flag = 0;
                t++;
            }
        }
        cur = nxt;
        nxt = nn;
    }
    cp_async_wait<0>();
}

struct NumericTuning { int Q; int panel_mb; int stages; int warps; bool q_forced; int macv; bool sort_chunks; };

NumericTuning numeric_tuning()
{
    NumericTuning tn = {96, 1 << 20, 2, 4, false, 3, true}; // panels off by default: measured no gain on cfg2 (see profiles/)
    if (const char *e = getenv("BSP_NUM_Q")) { int v = atoi(e); if (v >= 1) { tn.Q = v; tn.q_forced = true; } }
    if (const char *e = getenv("BSP_NUM_PANEL_MB")) { int v = atoi(e); if (v >= 1) tn.panel_mb = v; }
    if (const char *e = getenv("BSP_NUM_SORT")) tn.sort_chunks = atoi(e) != 0; // 0: chunks stay in row order (comparison hook)
    if (const char *e = getenv("BSP_NUM_MACV")) tn.macv = atoi(e) == 0 ? 0 : 3; // 0: the first kernel's plain 64-bit form
    if (const char *e = getenv("BSP_NUM_WARPS")) { int v = atoi(e); if (v == 4 || v == 8) tn.warps = v; }
    return tn;
}

// chunk schedule shared by the fast kernels; returns the device arrays (caller releases them)
struct ChunkSchedule { int32_t *seg_ts, *seg_nch, *seg_off; int2 *chunk_tiles; int max_chunks; int64_t nseg; };

int build_schedule(bsp_ctx *ctx, const bsp_mat *B, const bsp_symbolic_out *sym, int nbr, const int32_t *c_rowptr,
                   int tile_bytes, const NumericTuning &tn, ChunkSchedule *cs)
{
    // panels: keep the touched part of B well inside L2
    const int64_t b_bytes = B->nnzb * (int64_t)tile_bytes;
    int npanels = (int)((b_bytes + (int64_t)tn.panel_mb * (1 << 20) - 1) / ((int64_t)tn.panel_mb * (1 << 20)));
    if (npanels < 1) npanels = 1;
    if (npanels > 64) npanels = 64;
    if (npanels > B->nbc) npanels = B->nbc > 0 ? B->nbc : 1;
    const int panel_w = (B->nbc + npanels - 1) / npanels;
    // chunk size: at least a few output tiles' worth of products, so that a chunk is rarely a
    // single over-long tile next to empty neighbours (dense chain levels fold 100+ pairs per tile)
    int Q = tn.Q;
    if (!tn.q_forced && sym->nnzb_c > 0) {
        // ten tiles' worth, at most 2048 (chain top product, 105 pairs per tile: Q = 420 -> 6.55 ms,
        // 1024 -> 5.98 ms, 2048 -> 6.1 ms, 4096 -> 7.2 ms)
        const int64_t avg10 = 10 * (sym->nprod / sym->nnzb_c);
        if (avg10 > Q) Q = (int)(avg10 > 2048 ? 2048 : avg10);
    }
    if (!tn.q_forced && tile_bytes >= 2048) Q = Q / 4 > 8 ? Q / 4 : 8; // big tiles: a step is 8-64x the work
    cs->nseg = (int64_t)nbr * npanels;
    const int64_t max_chunks64 = sym->nprod / Q + cs->nseg + 1;
    if (max_chunks64 > 0x7fffff00LL) BSP_FAIL(ctx, BSP_ERR_LIMIT, "numeric pass: too many chunks");
    cs->max_chunks = (int)max_chunks64;
    cs->seg_ts = cs->seg_nch = cs->seg_off = nullptr;
    cs->chunk_tiles = nullptr;
    bsp_scratch sc(ctx);
    BSP_TRY(sc.alloc(&cs->seg_ts, (size_t)cs->nseg + 1));
    BSP_TRY(sc.alloc(&cs->seg_nch, (size_t)cs->nseg + 1));
    BSP_TRY(sc.alloc(&cs->seg_off, (size_t)cs->nseg + 1));
    BSP_TRY(sc.alloc(&cs->chunk_tiles, (size_t)cs->max_chunks));
    seg_bounds_kernel<<<(unsigned)((cs->nseg + 1 + 255) / 256), 256, 0, ctx->stream>>>(
        nbr, npanels, panel_w, c_rowptr, sym->c_colidx, sym->pair_ptr, Q, cs->seg_ts, cs->seg_nch);
    BSP_LAUNCH_CHECK(ctx);
    BSP_TRY(bsp_scan_i32(ctx, cs->seg_nch, cs->seg_off, cs->nseg + 1, nullptr));
    // panels off (the default): chunks in descending size class; with panels the panel-major order is the point
    const bool by_size = npanels == 1 && tn.sort_chunks;
    unsigned char *key = nullptr;
    int2 *unsorted = nullptr;
    int32_t *hist = nullptr;
    if (by_size) {
        BSP_TRY(sc.alloc(&key, (size_t)cs->max_chunks));
        BSP_TRY(sc.alloc(&unsorted, (size_t)cs->max_chunks));
        BSP_TRY(sc.alloc(&hist, (size_t)2 * CHUNK_CLASSES));
        BSP_CUDA(ctx, cudaMemsetAsync(hist, 0, 2 * CHUNK_CLASSES * sizeof(int32_t), ctx->stream));
    }
    chunk_fill_kernel<<<(unsigned)((cs->max_chunks + 255) / 256), 256, 0, ctx->stream>>>(
        nbr, npanels, panel_w, c_rowptr, sym->c_colidx, sym->pair_ptr, Q, cs->seg_ts, cs->seg_off,
        by_size ? unsorted : cs->chunk_tiles, key);
    BSP_LAUNCH_CHECK(ctx);
    if (by_size) {
        const unsigned g = (unsigned)((cs->max_chunks + CSORT_THREADS - 1) / CSORT_THREADS);
        chunk_hist_kernel<<<g, CSORT_THREADS, 0, ctx->stream>>>(key, cs->seg_off + cs->nseg, hist);
        BSP_LAUNCH_CHECK(ctx);
        chunk_scatter_kernel<<<g, CSORT_THREADS, 0, ctx->stream>>>(key, unsorted, cs->seg_off + cs->nseg, hist, hist + CHUNK_CLASSES,
                                                                   cs->chunk_tiles);
        BSP_LAUNCH_CHECK(ctx);
    }
    bsp_trace_mark(ctx, "num_schedule");
    sc.keep(cs->seg_ts); sc.keep(cs->seg_nch); sc.keep(cs->seg_off); sc.keep(cs->chunk_tiles); // released by release_schedule
    return BSP_OK;
}

void release_schedule(bsp_ctx *ctx, ChunkSchedule *cs)
{
    bsp_release(ctx, cs->seg_ts); bsp_release(ctx, cs->seg_nch); bsp_release(ctx, cs->seg_off); bsp_release(ctx, cs->chunk_tiles);
}

template <int K, int STAGES, int WARPS, int MINB, int MACV>
int launch_rows_cfg(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, int nbr,
                    const int32_t *c_rowptr, uint64_t *c_vals, const NumericTuning &tn)
{
    using Cfg = RowsCfg<K>;
    const int groups_per_cta = WARPS * Cfg::GROUPS;
    ChunkSchedule cs;
    BSP_TRY(build_schedule(ctx, B, sym, nbr, c_rowptr, Cfg::TILE_BYTES, tn, &cs));
    const size_t smem = (size_t)WARPS * Cfg::WARP_BYTES(STAGES);
    auto kern = numeric_rows_kernel<K, STAGES, WARPS, MINB, MACV>;
    BSP_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (cs.max_chunks + groups_per_cta - 1) / groups_per_cta;
    kern<<<grid, WARPS * 32, smem, ctx->stream>>>(A->vals, B->vals, sym->pair_ptr, sym->pairs, cs.chunk_tiles,
                                                  cs.seg_off + cs.nseg, c_vals, (unsigned long long *)ctx->d_counters);
    BSP_LAUNCH_CHECK(ctx);
    release_schedule(ctx, &cs);
    return BSP_OK;
}

template <int K>
int launch_bigtile(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, int nbr,
                   const int32_t *c_rowptr, uint64_t *c_vals)
{
    using Cfg = BigCfg<K>;
    constexpr int WARPS = 4;
    const NumericTuning tn = numeric_tuning();
    ChunkSchedule cs;
    BSP_TRY(build_schedule(ctx, B, sym, nbr, c_rowptr, Cfg::TILE_BYTES, tn, &cs));
    const size_t smem = (size_t)WARPS * 2 * Cfg::STAGE_BYTES;
    auto kern = tn.macv == 0 ? numeric_bigtile_kernel<K, WARPS, 0> : numeric_bigtile_kernel<K, WARPS, 3>;
    BSP_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int groups_per_cta = WARPS * Cfg::GROUPS;
    const int grid = (cs.max_chunks + groups_per_cta - 1) / groups_per_cta;
    kern<<<grid, WARPS * 32, smem, ctx->stream>>>(A->vals, B->vals, sym->pair_ptr, sym->pairs, cs.chunk_tiles,
                                                  cs.seg_off + cs.nseg, c_vals, (unsigned long long *)ctx->d_counters);
    BSP_LAUNCH_CHECK(ctx);
    release_schedule(ctx, &cs);
    return BSP_OK;
}

template <int K>
int launch_rows(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, int nbr,
                const int32_t *c_rowptr, uint64_t *c_vals)
{
    const NumericTuning tn = numeric_tuning();
    // resident CTAs per SM the register allocation aims at: 3 (12 warps).  Measured on configs[1]: a 4-CTA build
    // (126 registers) runs 16 % slower, the kernel is bound by integer instruction issue, not by latency.
    constexpr int MINB = K == 8 ? 1 : 3;
    if (tn.macv == 0) return launch_rows_cfg<K, 2, 4, MINB, 0>(ctx, A, B, sym, nbr, c_rowptr, c_vals, tn);
    return launch_rows_cfg<K, 2, 4, MINB, 3>(ctx, A, B, sym, nbr, c_rowptr, c_vals, tn);
}

} // namespace

int bsp_numeric(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym,
                uint64_t *c_vals, int64_t *exact_refolds, int32_t *path)
{
    if (exact_refolds) *exact_refolds = 0;
    if (path) *path = 0;
    if (sym->nnzb_c == 0) return BSP_OK;
    const int k = A->k;
    if (!ctx->force_generic && (k == 4 || k == 8 || k == 16 || k == 32)) {
        BSP_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, sizeof(unsigned long long), ctx->stream));
        if (k == 8) BSP_TRY(launch_rows<8>(ctx, A, B, sym, A->nbr, sym->c_rowptr, c_vals));
        else if (k == 4) BSP_TRY(launch_rows<4>(ctx, A, B, sym, A->nbr, sym->c_rowptr, c_vals));
        else if (k == 16) BSP_TRY(launch_bigtile<16>(ctx, A, B, sym, A->nbr, sym->c_rowptr, c_vals));
        else BSP_TRY(launch_bigtile<32>(ctx, A, B, sym, A->nbr, sym->c_rowptr, c_vals));
        if (path) *path = k;
        if (exact_refolds) {
            // the caller reads the counter at ctx->pinned + 64 after its event sync (bsp_capi.cu)
            bsp_publish_list pl = {};
            pl.n = 1;
            pl.src[0] = ctx->d_counters; pl.offset[0] = 64; pl.words[0] = 2;
            BSP_TRY(bsp_publish(ctx, pl));
        }
        return BSP_OK;
    }
    const int E = k * k;
    const int slots = E <= GEN_THREADS ? GEN_THREADS / E : 1;
    int64_t grid = (sym->nnzb_c + slots - 1) / slots;
    const int64_t max_grid = (int64_t)ctx->sm_count * 64;
    if (grid > max_grid) grid = max_grid;
    numeric_generic_kernel<<<(unsigned)grid, GEN_THREADS, 0, ctx->stream>>>(A->vals, B->vals, sym->pair_ptr,
                                                                             sym->pairs, sym->nnzb_c, k, c_vals);
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}
