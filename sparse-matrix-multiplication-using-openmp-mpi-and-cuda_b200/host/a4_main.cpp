// a4 -- drop-in for the reference's command line (cu:402-682):
//
//     a4 <folder> [--ranks P] [--devices D] [--binary] [--checksum] [--no-output]
//
// Reads <folder>/size ("N k": number of matrices and tile edge, cu:411-419) and
// <folder>/matrix1..N, multiplies the chain M1*M2*...*MN and writes the product to ./matrix in
// the current directory (cu:595-608).  Prints the reference's "multiplying a b" lines
// (cu:301) and a final "time taken X seconds" (cu:679).  Exit code 1 when <folder>/size cannot
// be opened (cu:413-417).
//
// --ranks P reproduces the association tree of `mpirun -np P ./a4` (the tree is part of the
// answer): the chain is cut into P contiguous sub-chains exactly as cu:438-456
// (N/P each, the last one takes the remainder; fewer matrices than ranks -> one sub-chain),
// every sub-chain is reduced by the pairwise tree, the P partial products are reduced by the
// same tree (cu:569-571), and all-zero tiles are dropped from the final product (cu:577-592).
// Sub-chain r runs on CUDA device r mod D on its own host thread; partial products move
// device-to-device into device 0, which does the merge.  Default P = 1, D = all visible.
//
// Additions without a counterpart in the reference (off by default, the default run is the
// reference's): --binary reads <folder>/matrixI.bin where it exists (else the text file) and
// writes ./matrix.bin instead of ./matrix; --checksum prints "checksum <16 hex digits>" of the
// final product (computed on the device); --no-output skips the output file (with --checksum:
// compare a 20 GB product without writing it).
//
// Host code only: everything that touches the GPU goes through the C ABI of libbspgemm.so.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "bspgemm.h"

namespace {

struct Rank {
    int start = 0, end = -1, device = 0;
    bsp_ctx *ctx = nullptr;
    std::vector<bsp_mat *> mats;
    bsp_mat *partial = nullptr;
    std::string log;
    int rc = 0;
    std::string err;
};

bool file_exists(const std::string &p)
{
    FILE *f = fopen(p.c_str(), "rb");
    if (f) fclose(f);
    return f != nullptr;
}

void run_rank(Rank *r, const std::string &folder, int k, bool binary)
{
    r->rc = bsp_ctx_create(r->device, nullptr, &r->ctx);
    if (r->rc) { r->err = r->ctx ? bsp_last_error(r->ctx) : "context"; return; }
    const int n = r->end - r->start + 1;
    r->mats.assign((size_t)n, nullptr);
    // The reference parses its files with one OpenMP task each (cu:334-339).  Here the text files of
    // the sub-chain go through the pipelined reader: host threads parse into pinned buffers while this
    // thread uploads and builds the finished ones.
    std::vector<std::string> text_paths;
    std::vector<size_t> text_slot;
    for (int i = 0; i < n && !r->rc; i++) {
        const std::string path = folder + "/matrix" + std::to_string(r->start + i + 1); // 1-based, cu:338-344
        if (binary && file_exists(path + ".bin")) {
            r->rc = bsp_mat_read_bin(r->ctx, (path + ".bin").c_str(), &r->mats[(size_t)i]);
            if (r->rc) r->err = bsp_last_error(r->ctx);
        } else {
            text_paths.push_back(path);
            text_slot.push_back((size_t)i);
        }
    }
    if (!r->rc && !text_paths.empty()) {
        std::vector<const char *> cp;
        for (auto &p : text_paths) cp.push_back(p.c_str());
        std::vector<bsp_mat *> got(cp.size(), nullptr);
        r->rc = bsp_mat_read_text_many(r->ctx, cp.data(), (int32_t)cp.size(), k, got.data());
        if (r->rc) r->err = bsp_last_error(r->ctx);
        else for (size_t j = 0; j < got.size(); j++) r->mats[text_slot[j]] = got[j];
    }
    if (r->rc) return;
    std::vector<char> log((size_t)n * 64 + 64);
    r->rc = bsp_chain(r->ctx, r->mats.data(), n, &r->partial, nullptr, log.data(), log.size(), r->start);
    if (r->rc) { r->err = bsp_last_error(r->ctx); return; }
    r->log = log.data();
}

} // namespace

int main(int argc, char **argv)
{
    auto t0 = std::chrono::high_resolution_clock::now();
    if (argc < 2) {
        fprintf(stderr, "usage: %s <folder> [--ranks P] [--devices D] [--binary] [--checksum] [--no-output]\n", argv[0]);
        return 2;
    }
    std::string folder = argv[1];
    int P = getenv("BSP_RANKS") ? atoi(getenv("BSP_RANKS")) : 1;
    int D = 0;
    bool binary = false, want_checksum = false, no_output = false;
    for (int i = 2; i < argc; i++) {
        if ((!strcmp(argv[i], "--ranks") || !strcmp(argv[i], "--gpus")) && i + 1 < argc) P = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--devices") && i + 1 < argc) D = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--binary")) binary = true;
        else if (!strcmp(argv[i], "--checksum")) want_checksum = true;
        else if (!strcmp(argv[i], "--no-output")) no_output = true;
        else { fprintf(stderr, "a4: unknown argument %s\n", argv[i]); return 2; }
    }
    if (P < 1) P = 1;

    int N = 0, k = 0;
    {
        FILE *f = fopen((folder + "/size").c_str(), "r");
        if (!f) {
            fprintf(stderr, "Cannot open size file!\n"); // cu:415
            return 1;
        }
        if (fscanf(f, "%d %d", &N, &k) != 2) { fclose(f); fprintf(stderr, "Cannot parse size file!\n"); return 1; }
        fclose(f);
    }
    if (N < 1 || k < 1) { fprintf(stderr, "size file: N=%d k=%d\n", N, k); return 1; }

    bsp_ctx *probe = nullptr;
    if (bsp_ctx_create(0, nullptr, &probe) != BSP_OK) {
        fprintf(stderr, "a4: %s\n", probe ? bsp_last_error(probe) : "cannot create context");
        bsp_ctx_destroy(probe);
        return 3;
    }
    if (D < 1) {
        D = 1;
        for (int d = 1; d < 64; d++) {
            bsp_ctx *c = nullptr;
            int rc = bsp_ctx_create(d, nullptr, &c);
            bsp_ctx_destroy(c);
            if (rc != BSP_OK) break;
            D = d + 1;
        }
    }

    // sub-chains, cu:438-456 / cu:612-616
    const int per = N / P;
    std::vector<Rank> ranks;
    if (per == 0) {
        ranks.resize(1);
        ranks[0].start = 0; ranks[0].end = N - 1;
    } else {
        ranks.resize((size_t)P);
        for (int r = 0; r < P; r++) {
            ranks[(size_t)r].start = r * per;
            ranks[(size_t)r].end = (r == P - 1) ? N - 1 : (r + 1) * per - 1;
        }
    }
    for (size_t r = 0; r < ranks.size(); r++) ranks[r].device = (int)(r % (size_t)D);
    {
        std::vector<std::thread> th;
        for (auto &r : ranks) th.emplace_back(run_rank, &r, folder, k, binary);
        for (auto &t : th) t.join();
    }
    int rc = 0;
    for (auto &r : ranks) {
        fputs(r.log.c_str(), stdout);
        if (r.rc) { fprintf(stderr, "a4: sub-chain %d..%d: %s\n", r.start, r.end, r.err.c_str()); rc = 3; }
    }

    bsp_mat *result = nullptr;
    if (!rc) {
        if (ranks.size() == 1) {
            result = ranks[0].partial;
            ranks[0].partial = nullptr;
            bsp_ctx_destroy(probe);
            probe = ranks[0].ctx; // the result lives in that context
            ranks[0].ctx = nullptr;
            for (auto *m : ranks[0].mats) bsp_mat_free(probe, m);
            ranks[0].mats.clear();
        } else {
            // gather the partial products on device 0 (replaces MPI_Send/Recv, cu:471-552)
            std::vector<bsp_mat *> parts;
            for (auto &r : ranks) {
                bsp_mat *m = nullptr;
                int e = bsp_mat_copy_to_ctx(r.ctx, r.partial, probe, &m);
                if (e) { fprintf(stderr, "a4: gather: %s\n", bsp_last_error(probe)); rc = 3; break; }
                parts.push_back(m);
            }
            if (!rc) {
                bsp_ctx_sync(probe);
                std::vector<char> log(parts.size() * 64 + 64);
                int e = bsp_chain(probe, parts.data(), (int)parts.size(), &result, nullptr, log.data(), log.size(), 0);
                if (e) { fprintf(stderr, "a4: merge: %s\n", bsp_last_error(probe)); rc = 3; }
                else fputs(log.data(), stdout);
            }
            for (auto *m : parts) bsp_mat_free(probe, m);
        }
    }
    if (!rc) {
        int e = bsp_prune_zero(probe, result);
        if (!e && want_checksum) {
            uint64_t sum = 0;
            e = bsp_mat_checksum(probe, result, &sum);
            if (!e) printf("checksum %016llx\n", (unsigned long long)sum);
        }
        if (!e && !no_output) e = binary ? bsp_mat_write_bin(probe, result, "matrix.bin") : bsp_mat_write_text(probe, result, "matrix");
        if (e) {
            fprintf(stderr, "a4: output: %s\n", bsp_last_error(probe));
            rc = 3;
        }
    }
    bsp_mat_free(probe, result);
    for (auto &r : ranks) {
        for (auto *m : r.mats) bsp_mat_free(r.ctx, m);
        bsp_mat_free(r.ctx, r.partial);
        bsp_ctx_destroy(r.ctx);
    }
    bsp_ctx_destroy(probe);
    auto t1 = std::chrono::high_resolution_clock::now();
    printf("time taken %g seconds\n", std::chrono::duration<double>(t1 - t0).count()); // cu:679
    return rc;
}
