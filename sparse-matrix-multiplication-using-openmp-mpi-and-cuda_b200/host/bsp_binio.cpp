// bsp_binio.cpp -- binary container for block-sparse matrices.
//
// No counterpart in the reference, which reads and writes decimal text only (cu:345-383,
// cu:595-608): the result of BASELINE.json configs[1] is 35-50 GB as text and 19.5 GB here.
// Layout (little-endian, no padding between the sections):
//   64-byte header: "BSPMAT01", rows i32, cols i32, k i32, esize i32 (8 = uint64, 4 = float32),
//                   nnzb i64, checksum u64 (bsp_mat_checksum of the content), 24 reserved zero bytes
//   brow[nnzb] i32, bcol[nnzb] i32 -- tiles in (brow, bcol) order, as the text writer emits them
//   vals[nnzb * k*k] of esize bytes, tile-major, tiles row-major
// The reader rebuilds the matrix on the device and recomputes the checksum there; a mismatch is
// BSP_ERR_IO.  Files are read and written in 16 MiB pieces through two pinned staging buffers, so
// the copy to or from the GPU of one piece overlaps the file I/O of the next.
#include <cstdio>
#include <cstring>
#include <vector>

#include "../csrc/bsp_internal.h"

namespace {

struct bin_header {
    char magic[8];
    int32_t rows, cols, k, esize;
    int64_t nnzb;
    uint64_t checksum;
    char reserved[24];
};
static_assert(sizeof(bin_header) == 64, "header is 64 bytes");

const char MAGIC[8] = {'B', 'S', 'P', 'M', 'A', 'T', '0', '1'};
constexpr size_t PIECE = 16u << 20;

// the two staging pieces live in the context (pinned allocation costs ~1 ms per MiB, far more than
// moving a small matrix), allocated by the first binary read or write and freed with the context
struct pinned_pair {
    char **buf;
    cudaEvent_t *done;
    bsp_ctx *ctx;
    explicit pinned_pair(bsp_ctx *c) : buf(c->io_buf), done(c->io_ev), ctx(c) {}
    bool init()
    {
        for (int i = 0; i < 2; i++) {
            if (!buf[i] && cudaMallocHost((void **)&buf[i], PIECE) != cudaSuccess) { buf[i] = nullptr; cudaGetLastError(); return false; }
            if (!done[i] && cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming) != cudaSuccess) { done[i] = nullptr; return false; }
        }
        return true;
    }
};

// device -> file, double-buffered
int stream_out(bsp_ctx *ctx, pinned_pair &pp, FILE *f, const void *dev, size_t bytes, const char *path)
{
    size_t off = 0;
    int slot = 0;
    size_t pending[2] = {0, 0};
    auto flush = [&](int sl) -> int { // wait for the download into buffer sl, then write it
        if (!pending[sl]) return BSP_OK;
        BSP_CUDA(ctx, cudaEventSynchronize(pp.done[sl]));
        if (fwrite(pp.buf[sl], 1, pending[sl], f) != pending[sl]) BSP_FAIL(ctx, BSP_ERR_IO, "short write to %s", path);
        pending[sl] = 0;
        return BSP_OK;
    };
    while (off < bytes) {
        const size_t n = bytes - off < PIECE ? bytes - off : PIECE;
        BSP_CUDA(ctx, cudaMemcpyAsync(pp.buf[slot], (const char *)dev + off, n, cudaMemcpyDeviceToHost, ctx->stream));
        BSP_CUDA(ctx, cudaEventRecord(pp.done[slot], ctx->stream));
        pending[slot] = n;
        off += n;
        BSP_TRY(flush(slot ^ 1)); // the previous piece goes to the file while this one is in flight
        slot ^= 1;
    }
    BSP_TRY(flush(slot));
    BSP_TRY(flush(slot ^ 1));
    return BSP_OK;
}

// file -> device, double-buffered
int stream_in(bsp_ctx *ctx, pinned_pair &pp, FILE *f, void *dev, size_t bytes, const char *path)
{
    size_t off = 0;
    int slot = 0;
    bool used[2] = {false, false};
    while (off < bytes) {
        const size_t n = bytes - off < PIECE ? bytes - off : PIECE;
        if (used[slot]) BSP_CUDA(ctx, cudaEventSynchronize(pp.done[slot])); // the upload from this buffer is over
        if (fread(pp.buf[slot], 1, n, f) != n) BSP_FAIL(ctx, BSP_ERR_IO, "%s: truncated", path);
        BSP_CUDA(ctx, cudaMemcpyAsync((char *)dev + off, pp.buf[slot], n, cudaMemcpyHostToDevice, ctx->stream));
        BSP_CUDA(ctx, cudaEventRecord(pp.done[slot], ctx->stream));
        used[slot] = true;
        off += n;
        slot ^= 1;
    }
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BSP_OK;
}

struct file_closer {
    FILE *f;
    ~file_closer() { if (f) fclose(f); }
};

} // namespace

extern "C" int bsp_mat_write_bin(bsp_ctx *ctx, const bsp_mat *m, const char *path)
{
    if (!ctx || !m || !path) return BSP_ERR_ARG;
    if (m->device != ctx->device) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_write_bin: matrix lives on device %d, context on %d", m->device, ctx->device);
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    bin_header h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, MAGIC, 8);
    h.rows = m->rows; h.cols = m->cols; h.k = m->k; h.esize = m->esize; h.nnzb = m->nnzb;
    BSP_TRY(bsp_mat_checksum(ctx, m, &h.checksum));
    file_closer fc{fopen(path, "wb")};
    if (!fc.f) BSP_FAIL(ctx, BSP_ERR_IO, "cannot create %s", path);
    if (fwrite(&h, sizeof(h), 1, fc.f) != 1) BSP_FAIL(ctx, BSP_ERR_IO, "short write to %s", path);
    if (m->nnzb > 0) {
        pinned_pair pp(ctx);
        if (!pp.init()) BSP_FAIL(ctx, BSP_ERR_NOMEM, "bsp_mat_write_bin: pinned staging buffers");
        bsp_scratch sc(ctx);
        int32_t *d_brow = nullptr;
        BSP_TRY(sc.alloc(&d_brow, (size_t)m->nnzb));
        BSP_TRY(bsp_expand_rows(ctx, m, d_brow));
        BSP_TRY(stream_out(ctx, pp, fc.f, d_brow, (size_t)m->nnzb * 4, path));
        BSP_TRY(stream_out(ctx, pp, fc.f, m->colidx, (size_t)m->nnzb * 4, path));
        BSP_TRY(stream_out(ctx, pp, fc.f, m->vals, (size_t)m->nnzb * m->k * m->k * m->esize, path));
    }
    FILE *f = fc.f;
    fc.f = nullptr;
    if (fclose(f) != 0) BSP_FAIL(ctx, BSP_ERR_IO, "short write to %s", path);
    return BSP_OK;
}

extern "C" int bsp_mat_read_bin(bsp_ctx *ctx, const char *path, bsp_mat **out)
{
    if (!ctx || !path || !out) return BSP_ERR_ARG;
    *out = nullptr;
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    file_closer fc{fopen(path, "rb")};
    if (!fc.f) BSP_FAIL(ctx, BSP_ERR_IO, "cannot open %s", path);
    bin_header h;
    if (fread(&h, sizeof(h), 1, fc.f) != 1 || memcmp(h.magic, MAGIC, 8) != 0)
        BSP_FAIL(ctx, BSP_ERR_IO, "%s: not a BSPMAT01 file", path);
    if (h.k < 1 || (h.esize != 8 && h.esize != 4) || h.nnzb < 0 || h.nnzb > 0x7fffffffLL)
        BSP_FAIL(ctx, BSP_ERR_IO, "%s: bad header (k=%d esize=%d nnzb=%lld)", path, h.k, h.esize, (long long)h.nnzb);
    const size_t tb = (size_t)h.k * h.k * h.esize;
    bsp_scratch sc(ctx);
    int32_t *d_r = nullptr, *d_c = nullptr;
    char *d_v = nullptr;
    BSP_TRY(sc.alloc(&d_r, (size_t)h.nnzb));
    BSP_TRY(sc.alloc(&d_c, (size_t)h.nnzb));
    BSP_TRY(sc.alloc(&d_v, (size_t)h.nnzb * tb));
    int32_t max_row = -1, max_col = -1;
    if (h.nnzb > 0) {
        pinned_pair pp(ctx);
        if (!pp.init()) BSP_FAIL(ctx, BSP_ERR_NOMEM, "bsp_mat_read_bin: pinned staging buffers");
        // coordinates first: they are validated on the host (8 bytes per tile) before anything is built
        std::vector<int32_t> rc((size_t)h.nnzb * 2);
        if (fread(rc.data(), 4, rc.size(), fc.f) != rc.size()) BSP_FAIL(ctx, BSP_ERR_IO, "%s: truncated", path);
        for (int64_t i = 0; i < h.nnzb; i++) {
            const int32_t r = rc[(size_t)i], c = rc[(size_t)h.nnzb + i];
            if (r < 0 || c < 0) BSP_FAIL(ctx, BSP_ERR_IO, "%s: negative block coordinate at tile %lld", path, (long long)i);
            if (r > max_row) max_row = r;
            if (c > max_col) max_col = c;
        }
        BSP_CUDA(ctx, cudaMemcpyAsync(d_r, rc.data(), (size_t)h.nnzb * 4, cudaMemcpyHostToDevice, ctx->stream));
        BSP_CUDA(ctx, cudaMemcpyAsync(d_c, rc.data() + h.nnzb, (size_t)h.nnzb * 4, cudaMemcpyHostToDevice, ctx->stream));
        BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        BSP_TRY(stream_in(ctx, pp, fc.f, d_v, (size_t)h.nnzb * tb, path));
    }
    bsp_mat *m = nullptr;
    int rc = bsp_ingest_coo(ctx, h.rows, h.cols, h.k, h.esize, h.nnzb, d_r, d_c, d_v, max_row, max_col, &m);
    uint64_t sum = 0;
    if (rc == BSP_OK) rc = bsp_mat_checksum(ctx, m, &sum);
    if (rc == BSP_OK && sum != h.checksum) {
        snprintf(ctx->err, sizeof(ctx->err), "%s: checksum mismatch (file %016llx, content %016llx)", path,
                 (unsigned long long)h.checksum, (unsigned long long)sum);
        rc = BSP_ERR_IO;
    }
    if (rc != BSP_OK) { bsp_mat_free(ctx, m); return rc; }
    *out = m;
    return BSP_OK;
}
