/*
 * example_sharded.cpp -- output-channel sharding of one batch across the GPUs
 * of a box from plain C++: no Python, no NCCL, only the C ABI.
 *
 *   example_sharded model.gcm WORLD BATCH [NDEV]
 *
 * The program forks WORLD processes ("ranks", one per GPU: rank r drives
 * device r % NDEV; with NDEV = 1 all ranks share GPU 0, which is how the
 * test-suite runs it on a one-GPU box).  Every rank loads the same material
 * (DGCM file with the garbler-side input encoding), takes shard (r, WORLD),
 * prepares the peer-memory exchange (dashb_set_shard_lanes), hands the CUDA
 * IPC handles of its gather buffers to the other ranks through files in a
 * temporary directory, evaluates the batch (dashb_evaluate_sharded) and
 * decodes it.  Rank 0 additionally evaluates the batch unsharded and checks
 * that the sharded labels are the same, label for label.
 *
 * The reference has no counterpart (one GPU, one inference per call:
 * garbling/garbled_circuit_interface.h:603-612); this is the C++ face of
 * SURVEY.md section 8(e)(2).  Build: g++ -std=c++17 -Iinclude
 * integration/example_sharded.cpp -Ldash_b200 -ldash_b200 -o example_sharded
 */
#include <sys/stat.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "dash_b200.h"

#define CHECK(call)                                                                   \
    do {                                                                              \
        if ((call) != DASHB_OK) {                                                     \
            fprintf(stderr, "rank %d: %s failed: %s\n", rank, #call, dashb_last_error()); \
            return 1;                                                                 \
        }                                                                             \
    } while (0)

static bool write_file(const std::string& path, const void* p, size_t n) {
    const std::string tmp = path + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return false;
    const bool ok = fwrite(p, 1, n, f) == n;
    fclose(f);
    return ok && rename(tmp.c_str(), path.c_str()) == 0;
}

static bool read_file_wait(const std::string& path, void* p, size_t n) {
    for (int tries = 0; tries < 3000; ++tries) { /* up to 30 s */
        FILE* f = fopen(path.c_str(), "rb");
        if (f) {
            const bool ok = fread(p, 1, n, f) == n;
            fclose(f);
            if (ok) return true;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    return false;
}

static int run_rank(int rank, int world, int ndev, const char* gcm, uint64_t batch, const std::string& dir) {
    CHECK(dashb_init(rank % ndev));
    dashb_circuit* c = nullptr;
    CHECK(dashb_circuit_load(&c, gcm));
    CHECK(dashb_set_shard(c, rank, world));
    CHECK(dashb_set_shard_lanes(c, 2));
    CHECK(dashb_finalize(c, batch));
    const uint64_t n_in = dashb_input_size(c), n_out = dashb_output_size(c);
    const int k = dashb_crt_base_size(c);

    /* the batch: the same synthetic values on every rank */
    std::vector<int64_t> x(batch * n_in);
    uint64_t st = 12345;
    for (auto& v : x) {
        st = st * 6364136223846793005ull + 1442695040888963407ull;
        v = (int64_t)((st >> 33) % 256);
    }
    CHECK(dashb_garble_inputs(c, x.data(), batch));

    /* hand the IPC handles round: lane l of rank r -> <dir>/h_<r>_<l> */
    const int lanes = dashb_shard_lanes_used(c);
    for (int l = 0; l < lanes; ++l) {
        unsigned char h[192];
        CHECK(dashb_shard_export(c, l, h));
        if (!write_file(dir + "/h_" + std::to_string(rank) + "_" + std::to_string(l), h, sizeof(h))) return 1;
    }
    for (int r = 0; r < world; ++r) {
        if (r == rank) continue;
        for (int l = 0; l < lanes; ++l) {
            unsigned char h[192];
            if (!read_file_wait(dir + "/h_" + std::to_string(r) + "_" + std::to_string(l), h, sizeof(h))) {
                fprintf(stderr, "rank %d: no handles from rank %d\n", rank, r);
                return 1;
            }
            CHECK(dashb_shard_import(c, l, r, h));
        }
    }

    std::vector<std::vector<int16_t>> out(k);
    std::vector<int16_t*> outp(k);
    for (int j = 0; j < k; ++j) {
        out[j].resize(batch * n_out * (size_t)dashb_nr_comps(dashb_output_modulus(c, j)));
        outp[j] = out[j].data();
    }
    for (int rep = 0; rep < 3; ++rep) {
        CHECK(dashb_garble_inputs(c, x.data(), batch));
        CHECK(dashb_evaluate_sharded(c));
    }
    CHECK(dashb_move_outputs(c, outp.data()));
    for (int l = 0; l < lanes; ++l) {
        uint32_t timed_out = 0;
        CHECK(dashb_shard_status(c, l, &timed_out));
        if (timed_out) {
            fprintf(stderr, "rank %d: a peer never delivered its slice (lane %d)\n", rank, l);
            return 1;
        }
    }
    std::vector<int64_t> values(batch * n_out);
    std::vector<uint8_t> ok(batch);
    CHECK(dashb_decode_outputs(c, values.data(), ok.data()));
    for (uint64_t b = 0; b < batch; ++b)
        if (!ok[b]) {
            fprintf(stderr, "rank %d: inference %llu does not decode\n", rank, (unsigned long long)b);
            return 1;
        }
    /* every rank holds the whole result; rank 0 compares it with the unsharded evaluation */
    if (rank == 0) {
        dashb_circuit* whole = nullptr;
        CHECK(dashb_circuit_load(&whole, gcm));
        CHECK(dashb_finalize(whole, batch));
        CHECK(dashb_garble_inputs(whole, x.data(), batch));
        CHECK(dashb_evaluate_device(whole));
        std::vector<std::vector<int16_t>> ref(k);
        std::vector<int16_t*> refp(k);
        for (int j = 0; j < k; ++j) {
            ref[j].resize(out[j].size());
            refp[j] = ref[j].data();
        }
        CHECK(dashb_move_outputs(whole, refp.data()));
        for (int j = 0; j < k; ++j)
            if (ref[j] != out[j]) {
                fprintf(stderr, "rank 0: sharded and unsharded labels differ (modulus %d)\n", j);
                return 1;
            }
        dashb_circuit_destroy(whole);
        printf("sharded == unsharded: %d ranks, batch %llu, %llu outputs, first values %lld %lld\n", world,
               (unsigned long long)batch, (unsigned long long)n_out, (long long)values[0], (long long)values[n_out > 1 ? 1 : 0]);
        fflush(stdout); /* the rank leaves through _exit() */
    }
    /* nobody unmaps a peer's buffers while that peer may still be pushing */
    write_file(dir + "/done_" + std::to_string(rank), "x", 1);
    for (int r = 0; r < world; ++r) {
        char ch;
        if (!read_file_wait(dir + "/done_" + std::to_string(r), &ch, 1)) return 1;
    }
    dashb_circuit_destroy(c);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s model.gcm WORLD BATCH [NDEV]\n", argv[0]);
        return 2;
    }
    const int world = atoi(argv[2]);
    const uint64_t batch = strtoull(argv[3], nullptr, 10);
    const int ndev = argc > 4 ? atoi(argv[4]) : world;
    if (world < 1 || world > 16 || batch < 1 || ndev < 1) return 2;
    char tmpl[] = "/tmp/dashb_shard_XXXXXX";
    if (!mkdtemp(tmpl)) return 1;
    const std::string dir = tmpl;
    std::vector<pid_t> pids;
    for (int r = 0; r < world; ++r) { /* fork BEFORE any CUDA call: every rank gets a context of its own */
        const pid_t pid = fork();
        if (pid == 0) _exit(run_rank(r, world, ndev, argv[1], batch, dir));
        pids.push_back(pid);
    }
    int bad = 0;
    for (pid_t pid : pids) {
        int st = 0;
        waitpid(pid, &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) ++bad;
    }
    const std::string rm = "rm -rf " + dir;
    if (system(rm.c_str()) != 0) bad += 0;
    return bad ? 1 : 0;
}
