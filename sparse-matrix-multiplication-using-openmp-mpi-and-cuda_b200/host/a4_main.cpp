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
// Sub-chain r runs on CUDA device r mod D on its own host thread.  With P <= D every rank has a device
// of its own: the ranks form an NCCL communicator and bsp_chain_distributed reduces the whole chain
// (ncclSend / ncclRecv over NVLink, every merge product sharded over the ranks of its subtree).  With
// P > D the partial products move device-to-device into device 0 and every large product of the merge
// is sharded by block rows over the D devices.  Same tree, same bits either way.  Default P = 1,
// D = all visible.
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
    bsp_comm *comm = nullptr; // one NCCL rank per device when every sub-chain has a device of its own
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

// nccl_id != nullptr: this thread is NCCL rank `rank` of `nranks` (one device each) and the whole chain is
// reduced by bsp_chain_distributed, which replaces MPI_Send / MPI_Recv of the partial products and the
// rank-0 merge (cu:471-571); every rank must reach the collective calls, so a rank that failed to load
// its files still takes part with an empty sub-chain and reports its error afterwards.
void run_rank(Rank *r, const std::string &folder, int k, bool binary, const void *nccl_id, int rank, int nranks)
{
    r->rc = bsp_ctx_create(r->device, nullptr, &r->ctx);
    if (r->rc) { r->err = r->ctx ? bsp_last_error(r->ctx) : "context"; if (!nccl_id) return; }
    if (nccl_id && r->ctx) {
        int crc = bsp_comm_create(r->ctx, nccl_id, nranks, rank, &r->comm);
        if (crc && !r->rc) { r->rc = crc; r->err = bsp_last_error(r->ctx); }
    }
    if (r->rc) return;
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
    std::vector<char> log((size_t)n * 64 + 64);
    if (r->comm) {
        // collective: a rank whose files failed passes an empty sub-chain so that the others do not hang
        const int load_rc = r->rc;
        const int nl = load_rc ? 0 : n;
        int crc = bsp_chain_distributed(r->ctx, r->comm, r->mats.data(), nl, r->start, &r->partial, nullptr, log.data(), log.size());
        if (!load_rc && crc) { r->rc = crc; r->err = bsp_last_error(r->ctx); }
        if (!r->rc) r->log = log.data();
        return;
    }
    if (r->rc) return;
    r->rc = bsp_chain(r->ctx, r->mats.data(), n, &r->partial, nullptr, log.data(), log.size(), r->start);
    if (r->rc) { r->err = bsp_last_error(r->ctx); return; }
    r->log = log.data();
}

// One product of the merge, sharded over the devices: device d multiplies the block rows
// [bounds[d], bounds[d+1]) of L (balanced by tiles per row) by R.  Rows are independent and every
// output element keeps its accumulation order, so the concatenated panels equal L*R bit for bit.
// devs[0] is the context that holds L and R and receives the product.  Each context is used by one
// thread at a time: the copies to and from the other devices are made by the calling thread, only the
// panel products run side by side.
int sharded_product(const std::vector<bsp_ctx *> &devs, const bsp_mat *L, const bsp_mat *R, bsp_mat **out, std::string *err)
{
    bsp_ctx *root = devs[0];
    const int D = (int)devs.size();
    int64_t nl = 0, nr = 0;
    bsp_mat_info(L, nullptr, nullptr, nullptr, &nl);
    bsp_mat_info(R, nullptr, nullptr, nullptr, &nr);
    int64_t shard_min = 200000; // tiles of L and R together; BSP_A4_SHARD_MIN lowers it for tests
    if (const char *e = getenv("BSP_A4_SHARD_MIN")) shard_min = atoll(e);
    if (D == 1 || nl + nr < shard_min) { // small product: moving it would cost more than it saves
        int rc = bsp_spgemm(root, L, R, out, nullptr);
        if (rc) *err = bsp_last_error(root);
        return rc;
    }
    int32_t nbr = 0;
    bsp_mat_device_view(L, &nbr, nullptr, nullptr, nullptr, nullptr);
    std::vector<int32_t> brow((size_t)nl);
    int rc = bsp_mat_to_coo(root, L, brow.data(), nullptr, nullptr);
    if (rc) { *err = bsp_last_error(root); return rc; }
    std::vector<int32_t> bounds((size_t)D + 1, nbr);
    bounds[0] = 0;
    for (int d = 1; d < D; d++) { // first row whose first tile is at or beyond d/D of all tiles (brow is sorted)
        const int64_t want = nl * d / D;
        bounds[(size_t)d] = want < nl ? brow[(size_t)want] : nbr;
        if (bounds[(size_t)d] < bounds[(size_t)d - 1]) bounds[(size_t)d] = bounds[(size_t)d - 1];
    }
    std::vector<bsp_mat *> Ld((size_t)D, nullptr), Rd((size_t)D, nullptr), panel((size_t)D, nullptr), home((size_t)D, nullptr);
    std::vector<int> prc((size_t)D, 0);
    for (int d = 1; d < D && !rc; d++) {
        rc = bsp_mat_copy_to_ctx(root, L, devs[(size_t)d], &Ld[(size_t)d]);
        if (!rc) rc = bsp_mat_copy_to_ctx(root, R, devs[(size_t)d], &Rd[(size_t)d]);
        if (rc) *err = bsp_last_error(devs[(size_t)d]);
    }
    if (!rc) {
        std::vector<std::thread> th;
        for (int d = 0; d < D; d++)
            th.emplace_back([&, d]() {
                prc[(size_t)d] = bsp_spgemm_panel(devs[(size_t)d], d ? Ld[(size_t)d] : L, d ? Rd[(size_t)d] : R, bounds[(size_t)d],
                                                  bounds[(size_t)d + 1], &panel[(size_t)d], nullptr);
            });
        for (auto &t : th) t.join();
        for (int d = 0; d < D && !rc; d++)
            if (prc[(size_t)d]) { rc = prc[(size_t)d]; *err = bsp_last_error(devs[(size_t)d]); }
    }
    home[0] = panel[0];
    for (int d = 1; d < D && !rc; d++) {
        rc = bsp_mat_copy_to_ctx(devs[(size_t)d], panel[(size_t)d], root, &home[(size_t)d]);
        if (rc) *err = bsp_last_error(root);
    }
    if (!rc) {
        rc = bsp_mat_concat_panels(root, home.data(), D, out);
        if (rc) *err = bsp_last_error(root);
    }
    for (int d = 0; d < D; d++) {
        bsp_mat_free(devs[(size_t)d], Ld[(size_t)d]);
        bsp_mat_free(devs[(size_t)d], Rd[(size_t)d]);
        bsp_mat_free(devs[(size_t)d], panel[(size_t)d]);
        if (d) bsp_mat_free(root, home[(size_t)d]);
    }
    return rc;
}

// The reference's rank 0 reduces the P partial products with helper2 (cu:569-571): adjacent pairs,
// odd tail carried (cu:297-320), one "multiplying a b" line per product (cu:301).  Same tree here,
// every product through sharded_product.  `parts` are owned by the caller.
int merge_partials(const std::vector<bsp_ctx *> &devs, const std::vector<bsp_mat *> &parts, bsp_mat **out, std::string *log,
                   std::string *err)
{
    bsp_ctx *root = devs[0];
    std::vector<bsp_mat *> cur(parts);
    std::vector<char> owned(parts.size(), 0);
    int rc = 0;
    while (cur.size() > 1 && !rc) {
        const size_t len = cur.size();
        std::vector<bsp_mat *> nxt;
        std::vector<char> nxt_owned;
        for (size_t i = 0; i + 1 < len; i += 2) *log += "multiplying " + std::to_string(i) + " " + std::to_string(i + 1) + "\n";
        for (size_t i = 0; i + 1 < len && !rc; i += 2) {
            bsp_mat *c = nullptr;
            rc = sharded_product(devs, cur[i], cur[i + 1], &c, err);
            if (rc) { bsp_mat_free(root, c); break; }
            nxt.push_back(c);
            nxt_owned.push_back(1);
            if (owned[i]) bsp_mat_free(root, cur[i]);
            if (owned[i + 1]) bsp_mat_free(root, cur[i + 1]);
            cur[i] = cur[i + 1] = nullptr;
        }
        if (!rc && (len & 1)) {
            nxt.push_back(cur[len - 1]);
            nxt_owned.push_back(owned[len - 1]);
            cur[len - 1] = nullptr;
        }
        if (rc) {
            for (size_t i = 0; i < nxt.size(); i++) if (nxt_owned[i]) bsp_mat_free(root, nxt[i]);
            for (size_t i = 0; i < cur.size(); i++) if (cur[i] && owned[i]) bsp_mat_free(root, cur[i]);
            return rc;
        }
        cur.swap(nxt);
        owned.swap(nxt_owned);
    }
    if (owned[0]) { *out = cur[0]; return 0; }
    // a single partial product: helper2 leaves it as it is; hand back a copy the caller may free
    rc = bsp_mat_copy_to_ctx(root, cur[0], root, out);
    if (rc) *err = bsp_last_error(root);
    return rc;
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
    // every sub-chain on a device of its own: the ranks form an NCCL communicator and the merge runs
    // distributed over them (BSP_A4_NO_NCCL=1 keeps the gather-on-device-0 path below)
    unsigned char nccl_id[BSP_COMM_ID_BYTES];
    const bool use_nccl = ranks.size() > 1 && ranks.size() <= (size_t)D && !getenv("BSP_A4_NO_NCCL") &&
                          bsp_comm_unique_id(nccl_id, sizeof(nccl_id)) == BSP_OK;
    {
        std::vector<std::thread> th;
        for (size_t r = 0; r < ranks.size(); r++)
            th.emplace_back(run_rank, &ranks[r], folder, k, binary, use_nccl ? (const void *)nccl_id : nullptr, (int)r, (int)ranks.size());
        for (auto &t : th) t.join();
    }
    int rc = 0;
    for (auto &r : ranks) {
        fputs(r.log.c_str(), stdout);
        if (r.rc) { fprintf(stderr, "a4: sub-chain %d..%d: %s\n", r.start, r.end, r.err.c_str()); rc = 3; }
    }

    bsp_mat *result = nullptr;
    if (!rc && use_nccl) {
        // bsp_chain_distributed left the product on rank 0; the reference's rank 0 prints the merge log (cu:301)
        for (size_t n = ranks.size(); n > 1; n = n / 2 + (n & 1))
            for (size_t i = 0; i + 1 < n; i += 2) printf("multiplying %zu %zu\n", i, i + 1);
        result = ranks[0].partial;
        ranks[0].partial = nullptr;
        bsp_ctx_destroy(probe);
        probe = ranks[0].ctx; // the result lives in that context
        if (ranks[0].comm) { bsp_comm_destroy(probe, ranks[0].comm); ranks[0].comm = nullptr; }
        ranks[0].ctx = nullptr;
        for (auto *m : ranks[0].mats) bsp_mat_free(probe, m);
        ranks[0].mats.clear();
    } else if (!rc) {
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
                // one context per device for the sharded merge: device 0 = probe, device d = the context of
                // the first sub-chain that ran there (its thread has finished)
                std::vector<bsp_ctx *> devs{probe};
                for (int d = 1; d < D; d++)
                    for (auto &r : ranks)
                        if (r.device == d) { devs.push_back(r.ctx); break; }
                std::string log, err;
                int e = merge_partials(devs, parts, &result, &log, &err);
                if (e) { fprintf(stderr, "a4: merge: %s\n", err.c_str()); rc = 3; }
                else fputs(log.c_str(), stdout);
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
        if (r.comm) bsp_comm_destroy(r.ctx, r.comm);
        bsp_ctx_destroy(r.ctx);
    }
    bsp_ctx_destroy(probe);
    auto t1 = std::chrono::high_resolution_clock::now();
    printf("time taken %g seconds\n", std::chrono::duration<double>(t1 - t0).count()); // cu:679
    return rc;
}
