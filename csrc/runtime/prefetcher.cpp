// Pinned-ring shard prefetcher (native threads).
//
// Semantics of the reference's experimental LABL loader (Module_1/labl_loader(EXPERIMENTAL).py:
// LABLShardedReader :7-28, PinnedRing :30-36, LABLPrefetcher :38-136): a background thread walks the
// shards sequentially through mmap, optionally z-scores every window (float64 statistics), fills
// caller-provided pinned slots batch by batch and hands (slot, n_filled, fill_ms) to the consumer,
// who does ONE coalesced H2D copy per batch and recycles the slot.  Batches never span shards.
// Differences: no Python in the fill loop (the GIL is never taken), optional bf16 output so the
// H2D copy is half the bytes, labels from .lbl sidecars, optional looping, and the exact-multiple
// end-of-shard case does not emit an empty batch (reference bug).
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "crossscale.h"

namespace {

thread_local std::string g_err;

struct Mapped {
    int fd = -1; void* base = nullptr; size_t size = 0; int64_t n = 0, l = 0;
    std::vector<uint8_t> labels;
    bool open(const std::string& path, std::string& err) {
        fd = ::open(path.c_str(), O_RDONLY);
        if (fd < 0) { err = "cannot open " + path; return false; }
        struct stat st;
        if (fstat(fd, &st) != 0 || st.st_size < 16) { err = "bad shard " + path; close(); return false; }
        size = (size_t)st.st_size;
        base = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (base == MAP_FAILED) { base = nullptr; err = "mmap failed " + path; close(); return false; }
        madvise(base, size, MADV_SEQUENTIAL);
        memcpy(&n, base, 8); memcpy(&l, (char*)base + 8, 8);
        // overflow-safe: compare n against the number of windows the payload can hold (a corrupt header must not wrap)
        if (n < 0 || l <= 0 || (uint64_t)l > (size - 16) / 4 || (uint64_t)n > (size - 16) / 4 / (uint64_t)l) { err = "shard header/payload mismatch " + path; close(); return false; }
        // optional label sidecar: <stem>.lbl = [int64 N][int64 classes][N uint8]
        std::string lp = path;
        size_t dot = lp.rfind('.');
        if (dot != std::string::npos) lp = lp.substr(0, dot);
        lp += ".lbl";
        FILE* f = fopen(lp.c_str(), "rb");
        if (f) {
            int64_t hdr[2];
            if (fread(hdr, 8, 2, f) == 2 && hdr[0] == n) {
                labels.resize((size_t)n);
                if (fread(labels.data(), 1, (size_t)n, f) != (size_t)n) labels.clear();
            }
            fclose(f);
        }
        return true;
    }
    const float* row(int64_t i) const { return (const float*)((const char*)base + 16) + i * l; }
    void close() {
        if (base) munmap(base, size);
        if (fd >= 0) ::close(fd);
        base = nullptr; fd = -1; labels.clear();
    }
};

inline uint16_t f32_to_bf16(float f) {
    uint32_t u; memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);   // NaN
    u += 0x7fffu + ((u >> 16) & 1u);                                            // round to nearest even
    return (uint16_t)(u >> 16);
}

struct Full { int slot; int n; double ms; };

}  // namespace

struct cs_prefetcher {
    std::vector<std::string> paths;
    int batch = 0, slots = 0, L = 0;
    bool normalize = false, bf16 = false, loop = false;
    std::vector<void*> slot_ptr;
    std::vector<int64_t*> label_ptr;
    std::mutex mu;
    std::condition_variable cv_free, cv_full;
    std::deque<int> q_free;
    std::deque<Full> q_full;
    std::atomic<bool> stop{false};
    bool done = false, failed = false;
    std::string error;
    std::thread th;

    void fill_row(const float* src, void* slot, int r) const {
        float mean = 0.0f, inv = 1.0f;
        if (normalize) {
            double s = 0.0, ss = 0.0;
            for (int i = 0; i < L; ++i) { double v = src[i]; s += v; ss += v * v; }
            const double m = s / L;
            double var = ss / L - m * m; if (var < 0) var = 0;
            mean = (float)m; inv = (float)(1.0 / (std::sqrt(var) + 1e-8));
        }
        if (bf16) {
            uint16_t* dst = (uint16_t*)slot + (size_t)r * L;
            if (normalize) for (int i = 0; i < L; ++i) dst[i] = f32_to_bf16((src[i] - mean) * inv);
            else for (int i = 0; i < L; ++i) dst[i] = f32_to_bf16(src[i]);
        } else {
            float* dst = (float*)slot + (size_t)r * L;
            if (normalize) for (int i = 0; i < L; ++i) dst[i] = (src[i] - mean) * inv;
            else memcpy(dst, src, (size_t)L * 4);
        }
    }

    void push(Full f) { { std::lock_guard<std::mutex> g(mu); q_full.push_back(f); } cv_full.notify_one(); }

    void run() {
        size_t shard = 0;
        Mapped m; bool have = false; int64_t off = 0;
        for (;;) {
            int slot;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_free.wait(lk, [&] { return stop.load() || !q_free.empty(); });
                if (stop.load()) break;
                slot = q_free.front(); q_free.pop_front();
            }
            // position on a shard that still has windows (skips empty shards, handles exact multiples)
            bool end = false;
            while (!have || off >= m.n) {
                if (have) { m.close(); have = false; ++shard; off = 0; }
                if (shard >= paths.size()) { if (loop && !paths.empty()) shard = 0; else { end = true; break; } }
                std::string err;
                if (!m.open(paths[shard], err)) { std::lock_guard<std::mutex> g(mu); failed = true; error = err; end = true; break; }
                if (m.l != L) { m.close(); std::lock_guard<std::mutex> g(mu); failed = true; error = "window length differs in " + paths[shard]; end = true; break; }
                have = true; off = 0;
            }
            if (end) { { std::lock_guard<std::mutex> g(mu); done = true; q_free.push_back(slot); } cv_full.notify_all(); break; }
            const auto t0 = std::chrono::steady_clock::now();
            const int n = (int)std::min<int64_t>(batch, m.n - off);
            for (int r = 0; r < n; ++r) fill_row(m.row(off + r), slot_ptr[slot], r);
            if (!label_ptr.empty() && label_ptr[slot]) {
                for (int r = 0; r < n; ++r) label_ptr[slot][r] = m.labels.empty() ? 0 : (int64_t)m.labels[(size_t)(off + r)];
            }
            off += n;
            const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            push(Full{slot, n, ms});
        }
        if (have) m.close();
        { std::lock_guard<std::mutex> g(mu); done = true; }
        cv_full.notify_all();
    }
};

extern "C" {

const char* cs_rt_last_error(void) { return g_err.c_str(); }

cs_prefetcher_t* cs_prefetcher_create(const char* const* shard_paths, int num_shards, int batch, int num_slots,
                                      void* const* slot_ptrs, int64_t* const* label_slots, int normalize,
                                      int out_bf16, int loop) {
    if (num_shards <= 0 || batch <= 0 || num_slots <= 0 || !slot_ptrs) { g_err = "prefetcher: bad arguments"; return nullptr; }
    cs_prefetcher* p = new cs_prefetcher();
    for (int i = 0; i < num_shards; ++i) p->paths.emplace_back(shard_paths[i]);
    Mapped probe; std::string err;
    if (!probe.open(p->paths[0], err)) { g_err = err; delete p; return nullptr; }
    p->L = (int)probe.l; probe.close();
    p->batch = batch; p->slots = num_slots; p->normalize = normalize != 0; p->bf16 = out_bf16 != 0; p->loop = loop != 0;
    for (int i = 0; i < num_slots; ++i) {
        p->slot_ptr.push_back(slot_ptrs[i]);
        if (label_slots) p->label_ptr.push_back(label_slots[i]);
        p->q_free.push_back(i);
    }
    return p;
}

int cs_prefetcher_window_len(cs_prefetcher_t* p) { return p ? p->L : 0; }

int cs_prefetcher_start(cs_prefetcher_t* p) {
    if (!p || p->th.joinable()) { g_err = "prefetcher: already started"; return 1; }
    p->th = std::thread([p] { p->run(); });
    return 0;
}

int cs_prefetcher_next(cs_prefetcher_t* p, int* slot, int* n_filled, double* fill_ms) {
    std::unique_lock<std::mutex> lk(p->mu);
    p->cv_full.wait(lk, [&] { return !p->q_full.empty() || p->done || p->stop.load(); });
    if (!p->q_full.empty()) {
        Full f = p->q_full.front(); p->q_full.pop_front();
        *slot = f.slot; *n_filled = f.n; *fill_ms = f.ms;
        return 1;
    }
    if (p->failed) { g_err = p->error; return -1; }
    return 0;
}

int cs_prefetcher_recycle(cs_prefetcher_t* p, int slot) {
    if (slot < 0 || slot >= p->slots) { g_err = "prefetcher: bad slot"; return 1; }
    { std::lock_guard<std::mutex> g(p->mu); p->q_free.push_back(slot); }
    p->cv_free.notify_one();
    return 0;
}

void cs_prefetcher_destroy(cs_prefetcher_t* p) {
    if (!p) return;
    {   // set the flag under the mutex the worker's predicates are checked under: a store + notify that lands between the
        // worker's predicate check and its wait would otherwise be lost and join() below would hang forever
        std::lock_guard<std::mutex> g(p->mu);
        p->stop.store(true);
    }
    p->cv_free.notify_all(); p->cv_full.notify_all();
    if (p->th.joinable()) p->th.join();
    delete p;
}

}  // extern "C"
