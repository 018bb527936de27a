// One-sided windows over POSIX shared memory for the ranks of one node (host runtime; no GPU involved).
//
// Role: what MPI RMA windows are to the reference's MPI engine (reference DGraph/distributed/mpi/MPIBackendEngine.py:33-193 --
// Win.Create / Fence / Get per row / Accumulate(SUM) per row / Free), as a persistent runtime instead of a window per call:
//   * every rank owns one control block (fixed size) and one data segment (grown on demand, generation-numbered);
//   * peers map each other's segments once and re-map lazily when the owner has published a new generation;
//   * a fence is an epoch flag per rank (store-release of my epoch, acquire-spin until every peer's epoch has caught up) with a
//     bounded wait, so that a rank that died is reported instead of hanging the job;
//   * gets are row memcpys out of the owner's segment, accumulates are row-locked vector adds into it (one lock byte per row), both run over a small
//     thread pool for large requests.  Nothing is allocated or mapped between the two fences of an access epoch.
//
// The Python side is dgraph_b200/parallel/shm_engine.py.  Exported with C linkage; every call returns 0 on success.
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#define DGBH_API extern "C" __attribute__((visibility("default")))

namespace {

enum : int {
    OK = 0,
    ERR_TIMEOUT = 1,    // a peer did not arrive at the fence / did not create its segment in time
    ERR_BOUNDS = 2,     // a row lies outside the owner's segment
    ERR_SYS = 3,        // shm_open / ftruncate / mmap failed (errno is printed)
    ERR_ARG = 4,        // bad rank, null pointer, ...
    ERR_ABORTED = 5,    // another rank reported a failure
};

struct alignas(64) Control {
    std::atomic<uint64_t> magic;  // set last by the owner: the block is initialised
    std::atomic<uint64_t> epoch;  // fences completed by the owner (arrival counter)
    std::atomic<uint64_t> gen;    // generation of the owner's data segment
    std::atomic<uint64_t> bytes;  // capacity of that generation
    std::atomic<uint32_t> abort;  // set by anyone who gave up waiting
    std::atomic<int64_t> lock_off;  // byte offset of the owner's per-row lock bytes for the coming accumulate epoch (-1: none)
};
constexpr uint64_t MAGIC = 0x4447425f57494e31ull;  // "DGB_WIN1"

struct Peer {
    Control* ctl = nullptr;
    char* data = nullptr;
    uint64_t gen = ~0ull;
    uint64_t bytes = 0;
};

struct Window {
    std::string job;
    int rank = 0, world = 1;
    uint64_t epoch = 0;
    int threads = 1;
    std::vector<Peer> peers;
};

std::string ctl_name(const std::string& job, int r) { return "/" + job + "_c" + std::to_string(r); }
std::string dat_name(const std::string& job, int r, uint64_t gen) {
    return "/" + job + "_d" + std::to_string(r) + "_" + std::to_string(gen);
}

double now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

void relax(unsigned& spins) {
    if (++spins < 256) {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    } else if (spins < 4096) {
        sched_yield();
    } else {
        timespec ts{0, 50000};  // 50 us: ranks may share cores on a small box
        nanosleep(&ts, nullptr);
    }
}

void* map_segment(const std::string& name, uint64_t bytes, bool create) {
    int fd = shm_open(name.c_str(), create ? (O_CREAT | O_EXCL | O_RDWR) : O_RDWR, 0600);
    if (fd < 0) return nullptr;
    if (create && ftruncate(fd, (off_t)bytes) != 0) {
        std::fprintf(stderr, "dgraph_b200 shm window: ftruncate(%s, %llu): %s\n", name.c_str(), (unsigned long long)bytes, std::strerror(errno));
        close(fd);
        shm_unlink(name.c_str());
        return nullptr;
    }
    if (!create) {  // the owner may still be between shm_open and ftruncate
        struct stat st;
        if (fstat(fd, &st) != 0 || (uint64_t)st.st_size < bytes) {
            close(fd);
            return nullptr;
        }
    }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    return p == MAP_FAILED ? nullptr : p;
}

uint64_t round_up(uint64_t b) {
    const uint64_t page = 1ull << 16;
    return (b + page - 1) / page * page;
}

// Map (or re-map) the data segment of peer p if its owner has published a newer generation.  Only called right after a fence
// or before the first one: the acquire load of the owner's epoch ordered its gen / bytes stores before our loads.
int refresh_peer(Window* w, int p) {
    Peer& pe = w->peers[p];
    const uint64_t gen = pe.ctl->gen.load(std::memory_order_acquire);
    if (gen == pe.gen) return OK;
    const uint64_t bytes = pe.ctl->bytes.load(std::memory_order_acquire);
    void* m = map_segment(dat_name(w->job, p, gen), bytes, false);
    if (!m) {
        std::fprintf(stderr, "dgraph_b200 shm window: rank %d cannot map generation %llu of rank %d: %s\n", w->rank, (unsigned long long)gen, p,
                     std::strerror(errno));
        return ERR_SYS;
    }
    if (pe.data) munmap(pe.data, pe.bytes);
    pe.data = static_cast<char*>(m);
    pe.gen = gen;
    pe.bytes = bytes;
    return OK;
}

template <class Fn>
void parallel_rows(const Window* w, int64_t n, int64_t bytes_per_row, Fn fn) {
    int t = w->threads;
    if (t <= 1 || n * bytes_per_row < (4ll << 20) || n < 2 * t) {
        fn(0, n);
        return;
    }
    std::vector<std::thread> pool;
    pool.reserve(t - 1);
    const int64_t chunk = (n + t - 1) / t;
    for (int i = 1; i < t; ++i) {
        const int64_t lo = i * chunk, hi = std::min<int64_t>(n, lo + chunk);
        if (lo < hi) pool.emplace_back([=] { fn(lo, hi); });
    }
    fn(0, std::min<int64_t>(n, chunk));
    for (auto& th : pool) th.join();
}

inline void atomic_add_f32(float* addr, float v) {
    auto* a = reinterpret_cast<uint32_t*>(addr);
    uint32_t old = __atomic_load_n(a, __ATOMIC_RELAXED), want;
    do {
        float f;
        std::memcpy(&f, &old, 4);
        f += v;
        std::memcpy(&want, &f, 4);
    } while (!__atomic_compare_exchange_n(a, &old, want, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
}

inline void atomic_add_f64(double* addr, double v) {
    auto* a = reinterpret_cast<uint64_t*>(addr);
    uint64_t old = __atomic_load_n(a, __ATOMIC_RELAXED), want;
    do {
        double f;
        std::memcpy(&f, &old, 8);
        f += v;
        std::memcpy(&want, &f, 8);
    } while (!__atomic_compare_exchange_n(a, &old, want, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
}

}  // namespace

// Create this rank's control block and first data segment.  Collective in spirit (every rank calls it once), but it does not
// wait for anyone: dgbh_win_connect does.
DGBH_API void* dgbh_win_open(const char* job, int rank, int world, uint64_t bytes, int threads) {
    if (!job || rank < 0 || world < 1 || rank >= world) return nullptr;
    auto* w = new Window;
    w->job = job;
    w->rank = rank;
    w->world = world;
    w->threads = threads > 0 ? threads : 1;
    w->peers.resize(world);
    bytes = round_up(bytes ? bytes : 1);
    shm_unlink(ctl_name(w->job, rank).c_str());  // leftovers of a crashed job with the same name
    auto* ctl = static_cast<Control*>(map_segment(ctl_name(w->job, rank), sizeof(Control), true));
    if (!ctl) {
        std::fprintf(stderr, "dgraph_b200 shm window: cannot create %s: %s\n", ctl_name(w->job, rank).c_str(), std::strerror(errno));
        delete w;
        return nullptr;
    }
    shm_unlink(dat_name(w->job, rank, 0).c_str());
    void* d = map_segment(dat_name(w->job, rank, 0), bytes, true);
    if (!d) {
        std::fprintf(stderr, "dgraph_b200 shm window: cannot create a %llu-byte segment: %s\n", (unsigned long long)bytes, std::strerror(errno));
        munmap(ctl, sizeof(Control));
        shm_unlink(ctl_name(w->job, rank).c_str());
        delete w;
        return nullptr;
    }
    new (ctl) Control;
    ctl->epoch.store(0, std::memory_order_relaxed);
    ctl->gen.store(0, std::memory_order_relaxed);
    ctl->bytes.store(bytes, std::memory_order_relaxed);
    ctl->abort.store(0, std::memory_order_relaxed);
    ctl->lock_off.store(-1, std::memory_order_relaxed);
    ctl->magic.store(MAGIC, std::memory_order_release);
    Peer& me = w->peers[rank];
    me.ctl = ctl;
    me.data = static_cast<char*>(d);
    me.gen = 0;
    me.bytes = bytes;
    return w;
}

// Map every peer's control block and current data segment; waits (bounded) for peers that have not created theirs yet.
DGBH_API int dgbh_win_connect(void* h, double timeout_ms) {
    auto* w = static_cast<Window*>(h);
    if (!w) return ERR_ARG;
    const double t0 = now_ms();
    for (int p = 0; p < w->world; ++p) {
        if (p == w->rank) continue;
        Peer& pe = w->peers[p];
        unsigned spins = 0;
        while (!pe.ctl) {
            auto* c = static_cast<Control*>(map_segment(ctl_name(w->job, p), sizeof(Control), false));
            if (c && c->magic.load(std::memory_order_acquire) == MAGIC) {
                pe.ctl = c;
                break;
            }
            if (c) munmap(c, sizeof(Control));
            if (now_ms() - t0 > timeout_ms) return ERR_TIMEOUT;
            relax(spins);
        }
        if (int rc = refresh_peer(w, p)) return rc;
    }
    return OK;
}

DGBH_API void* dgbh_win_base(void* h) { return h ? static_cast<Window*>(h)->peers[static_cast<Window*>(h)->rank].data : nullptr; }
DGBH_API uint64_t dgbh_win_bytes(void* h) { return h ? static_cast<Window*>(h)->peers[static_cast<Window*>(h)->rank].bytes : 0; }
DGBH_API uint64_t dgbh_win_epoch(void* h) { return h ? static_cast<Window*>(h)->epoch : 0; }

// Replace this rank's data segment by a larger one (contents are NOT carried over).  Only legal outside an access epoch, i.e.
// when no peer can be reading or writing the segment: after the closing fence of the previous operation and before the
// opening fence of the next.  Peers pick the new generation up at their next fence.
DGBH_API int dgbh_win_grow(void* h, uint64_t bytes) {
    auto* w = static_cast<Window*>(h);
    if (!w) return ERR_ARG;
    Peer& me = w->peers[w->rank];
    bytes = round_up(bytes);
    if (bytes <= me.bytes) return OK;
    const uint64_t gen = me.gen + 1;
    shm_unlink(dat_name(w->job, w->rank, gen).c_str());
    void* d = map_segment(dat_name(w->job, w->rank, gen), bytes, true);
    if (!d) {
        std::fprintf(stderr, "dgraph_b200 shm window: cannot grow to %llu bytes: %s\n", (unsigned long long)bytes, std::strerror(errno));
        return ERR_SYS;
    }
    munmap(me.data, me.bytes);
    shm_unlink(dat_name(w->job, w->rank, me.gen).c_str());  // peers that still map it keep their mapping until they refresh
    me.data = static_cast<char*>(d);
    me.bytes = bytes;
    me.gen = gen;
    me.ctl->bytes.store(bytes, std::memory_order_release);
    me.ctl->gen.store(gen, std::memory_order_release);
    return OK;
}

// Epoch fence: everything this rank wrote (into its own segment or a peer's) before the call is visible to every rank after it,
// and every peer has finished the accesses it issued before its own call.  Bounded: ERR_TIMEOUT names the first missing rank on
// stderr and raises the abort flag so that the others stop waiting too.
DGBH_API int dgbh_win_fence(void* h, double timeout_ms) {
    auto* w = static_cast<Window*>(h);
    if (!w) return ERR_ARG;
    const uint64_t e = ++w->epoch;
    w->peers[w->rank].ctl->epoch.store(e, std::memory_order_release);
    const double t0 = now_ms();
    for (int p = 0; p < w->world; ++p) {
        if (p == w->rank) continue;
        unsigned spins = 0;
        while (w->peers[p].ctl->epoch.load(std::memory_order_acquire) < e) {
            if (w->peers[p].ctl->abort.load(std::memory_order_relaxed) || w->peers[w->rank].ctl->abort.load(std::memory_order_relaxed))
                return ERR_ABORTED;
            if ((spins & 1023) == 1023 && now_ms() - t0 > timeout_ms) {
                std::fprintf(stderr, "dgraph_b200 shm window: rank %d waited %.0f ms for rank %d at fence %llu\n", w->rank, timeout_ms, p,
                             (unsigned long long)e);
                for (int q = 0; q < w->world; ++q) w->peers[q].ctl->abort.store(1, std::memory_order_relaxed);
                return ERR_TIMEOUT;
            }
            relax(spins);
        }
    }
    for (int p = 0; p < w->world; ++p)
        if (p != w->rank)
            if (int rc = refresh_peer(w, p)) return rc;
    return OK;
}

// dst[dst_row[i]] = segment(owner[i])[base_off + row[i] * row_bytes ...]   (dst_row == nullptr: i)
DGBH_API int dgbh_win_get_rows(void* h, const int64_t* row, const int32_t* owner, const int64_t* dst_row, int64_t n, int64_t row_bytes,
                               int64_t base_off, void* dst) {
    auto* w = static_cast<Window*>(h);
    if (!w || !row || !owner || (!dst && n > 0) || row_bytes <= 0 || base_off < 0) return ERR_ARG;
    std::atomic<int> rc{OK};
    parallel_rows(w, n, row_bytes, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            const int p = owner[i];
            if (p < 0 || p >= w->world) { rc = ERR_ARG; return; }
            const Peer& pe = w->peers[p];
            const int64_t off = base_off + row[i] * row_bytes;
            if (row[i] < 0 || (uint64_t)(off + row_bytes) > pe.bytes) { rc = ERR_BOUNDS; return; }
            std::memcpy(static_cast<char*>(dst) + (dst_row ? dst_row[i] : i) * row_bytes, pe.data + off, (size_t)row_bytes);
        }
    });
    return rc.load();
}

// segment(peer)[dst_off ...] = src[0 .. nbytes)   -- the contiguous one-sided write of a halo `put`
DGBH_API int dgbh_win_put(void* h, int peer, int64_t dst_off, const void* src, int64_t nbytes) {
    auto* w = static_cast<Window*>(h);
    if (!w || peer < 0 || peer >= w->world || dst_off < 0 || nbytes < 0 || (!src && nbytes > 0)) return ERR_ARG;
    const Peer& pe = w->peers[peer];
    if ((uint64_t)(dst_off + nbytes) > pe.bytes) return ERR_BOUNDS;
    char* d = pe.data + dst_off;
    const char* s = static_cast<const char*>(src);
    parallel_rows(w, (nbytes + 4095) / 4096, 4096, [&](int64_t lo, int64_t hi) {
        const int64_t a = lo * 4096, b = std::min<int64_t>(nbytes, hi * 4096);
        if (a < b) std::memcpy(d + a, s + a, (size_t)(b - a));
    });
    return OK;
}

// segment(owner[i])[base_off + row[i] * F ...] += src[i * F ...]   (elem_bytes 4: float, 8: double); concurrent adds from any
// number of ranks and threads to the same row are safe.  An owner that published a lock array (dgbh_win_set_lock_offset, before the
// epoch's opening fence; one zeroed byte per row at segment[lock_off + row]) has its rows locked with one atomic exchange, added with
// plain vector adds and released with one store: two atomics per row instead of one compare-and-swap loop per element.  Owners
// without a lock array (-1) get the compare-and-swap loop.
DGBH_API int dgbh_win_accumulate_rows(void* h, const int64_t* row, const int32_t* owner, int64_t n, int64_t F, int elem_bytes, int64_t base_off,
                                      const void* src) {
    auto* w = static_cast<Window*>(h);
    if (!w || !row || !owner || (!src && n > 0) || F <= 0 || base_off < 0 || (elem_bytes != 4 && elem_bytes != 8) || base_off % elem_bytes)
        return ERR_ARG;
    const int64_t row_bytes = F * elem_bytes;
    std::atomic<int> rc{OK};
    parallel_rows(w, n, row_bytes, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            const int p = owner[i];
            if (p < 0 || p >= w->world) { rc = ERR_ARG; return; }
            const Peer& pe = w->peers[p];
            const int64_t off = base_off + row[i] * row_bytes;
            if (row[i] < 0 || (uint64_t)(off + row_bytes) > pe.bytes) { rc = ERR_BOUNDS; return; }
            const int64_t lock_off = pe.ctl->lock_off.load(std::memory_order_relaxed);  // ordered by the fence's acquire
            if (lock_off >= 0) {
                if ((uint64_t)(lock_off + row[i] + 1) > pe.bytes) { rc = ERR_BOUNDS; return; }
                auto* lock = reinterpret_cast<unsigned char*>(pe.data + lock_off + row[i]);
                unsigned spins = 0;
                double t0 = 0;
                while (__atomic_exchange_n(lock, (unsigned char)1, __ATOMIC_ACQUIRE)) {
                    if ((++spins & 0xFFFF) == 0) {  // a holder that died would otherwise stall everybody for ever
                        if (t0 == 0) t0 = now_ms();
                        if (pe.ctl->abort.load(std::memory_order_relaxed)) { rc = ERR_ABORTED; return; }
                        if (now_ms() - t0 > 120000.0) { rc = ERR_TIMEOUT; return; }
                    }
#if defined(__x86_64__) || defined(__i386__)
                    __builtin_ia32_pause();
#endif
                }
                if (elem_bytes == 4) {
                    auto* d = reinterpret_cast<float*>(pe.data + off);
                    const float* s = static_cast<const float*>(src) + i * F;
                    for (int64_t c = 0; c < F; ++c) d[c] += s[c];
                } else {
                    auto* d = reinterpret_cast<double*>(pe.data + off);
                    const double* s = static_cast<const double*>(src) + i * F;
                    for (int64_t c = 0; c < F; ++c) d[c] += s[c];
                }
                __atomic_store_n(lock, (unsigned char)0, __ATOMIC_RELEASE);
            } else if (elem_bytes == 4) {
                auto* d = reinterpret_cast<float*>(pe.data + off);
                const float* s = static_cast<const float*>(src) + i * F;
                for (int64_t c = 0; c < F; ++c) atomic_add_f32(d + c, s[c]);
            } else {
                auto* d = reinterpret_cast<double*>(pe.data + off);
                const double* s = static_cast<const double*>(src) + i * F;
                for (int64_t c = 0; c < F; ++c) atomic_add_f64(d + c, s[c]);
            }
        }
    });
    return rc.load();
}

// Publish (or withdraw, -1) the offset of this rank's lock bytes; takes effect for peers at the next fence.
DGBH_API int dgbh_win_set_lock_offset(void* h, int64_t off) {
    auto* w = static_cast<Window*>(h);
    if (!w) return ERR_ARG;
    w->peers[w->rank].ctl->lock_off.store(off, std::memory_order_relaxed);
    return OK;
}

DGBH_API int dgbh_win_aborted(void* h) {
    auto* w = static_cast<Window*>(h);
    return w ? (int)w->peers[w->rank].ctl->abort.load(std::memory_order_relaxed) : 0;
}

DGBH_API void dgbh_win_close(void* h) {
    auto* w = static_cast<Window*>(h);
    if (!w) return;
    for (int p = 0; p < w->world; ++p) {
        Peer& pe = w->peers[p];
        if (pe.data) munmap(pe.data, pe.bytes);
        if (p == w->rank) {
            shm_unlink(dat_name(w->job, p, pe.gen).c_str());
            shm_unlink(ctl_name(w->job, p).c_str());
        }
        if (pe.ctl) munmap(pe.ctl, sizeof(Control));
    }
    delete w;
}
