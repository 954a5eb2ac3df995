// File-system rendezvous for the one-time exchange of CUDA IPC handles between the ranks of a node.
//
// The reference bootstraps through MPI (mpi4py.MPI.COMM_WORLD, part3_fedavg_overlap_mpi_gpu.py:149-151);
// here the hot paths need no communication library at all, only this set-up step: every rank publishes a
// small blob (its two 64-byte IPC handles) and collects everybody else's.  C callers that do not run under
// torch.distributed use this instead of all_gather_object: each rank writes <dir>/<tag>.<rank> atomically
// (write to a temp name, then rename) and polls for the others.  Works for any processes that share a
// directory (tmpfs, /dev/shm); no sockets, no daemon.
#include <errno.h>
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "crossscale.h"

namespace {
thread_local std::string g_rdv_err;

bool read_all(const std::string& path, void* dst, size_t n) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    const size_t got = fread(dst, 1, n, f);
    fclose(f);
    return got == n;
}
}  // namespace

extern "C" {

const char* cs_rendezvous_last_error(void) { return g_rdv_err.c_str(); }

// All-gather of `blob_bytes` per rank through directory `dir`.  `out` receives world * blob_bytes, ordered by
// rank.  `tag` distinguishes successive exchanges in one directory.  Returns 0, or non-zero after `timeout_s`.
int cs_rendezvous_allgather(const char* dir, const char* tag, int rank, int world, const void* blob, size_t blob_bytes,
                            void* out, double timeout_s) {
    if (!dir || !tag || !blob || !out || world < 1 || rank < 0 || rank >= world) { g_rdv_err = "rendezvous: bad arguments"; return 1; }
    mkdir(dir, 0777);
    const std::string base = std::string(dir) + "/" + tag + ".";
    const std::string mine = base + std::to_string(rank);
    const std::string tmp = mine + ".tmp." + std::to_string((long)getpid());
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) { g_rdv_err = "rendezvous: cannot write " + tmp + ": " + strerror(errno); return 2; }
    const bool ok = fwrite(blob, 1, blob_bytes, f) == blob_bytes;
    fclose(f);
    if (!ok || rename(tmp.c_str(), mine.c_str()) != 0) { g_rdv_err = "rendezvous: cannot publish " + mine; return 2; }
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<bool> have(world, false);
    int missing = world;
    while (missing > 0) {
        for (int r = 0; r < world; ++r) {
            if (have[r]) continue;
            struct stat st;
            const std::string p = base + std::to_string(r);
            if (stat(p.c_str(), &st) == 0 && (size_t)st.st_size == blob_bytes &&
                read_all(p, (char*)out + (size_t)r * blob_bytes, blob_bytes)) {
                have[r] = true; --missing;
            }
        }
        if (missing == 0) break;
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) {
            g_rdv_err = "rendezvous: timed out waiting for " + std::to_string(missing) + " rank(s) in " + dir;
            return 3;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(2));
    }
    return 0;
}

// Remove this rank's file once everybody has read it (call after a later barrier / at shutdown).
int cs_rendezvous_cleanup(const char* dir, const char* tag, int rank) {
    const std::string mine = std::string(dir) + "/" + tag + "." + std::to_string(rank);
    return unlink(mine.c_str()) == 0 ? 0 : 1;
}

}  // extern "C"
