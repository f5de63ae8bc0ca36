// stats_exchange.cu — the path's only inter-GPU exchange: the end-of-step statistics, summed over ranks for every
// generation (worlds are independent units, modules/fimo_worlds/src/World.zig:26,45-52, so nothing
// else crosses GPUs).  One process per GPU.  Two transports, see include/fimo_worlds_b200.h:
//
//   PEERS  the publishing CTA of the statistics kernel stores the generation's words into every rank's inbox over
//          NVLink (cudaIpc-mapped peer HBM), as self-validating {tag, half word} u64s (kernels_common.cuh:
//          stats_finalize).  Nothing is launched, nothing runs beside the persistent world-step grid.
//   NCCL   blocks of `bucket` generations of the statistics ring are all-reduced in place on a side stream
//          (ncclAllReduce / ncclUint64 / ncclSum), enqueued here from the launcher thread; events order the ring
//          block's reuse.  libnccl is dlopen()ed at bind time.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h> // types only; the library is resolved at run time

#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "kernels_common.cuh"

namespace fb::dev {
namespace {

constexpr uint32_t kNcclBlocks = 4;            // ring = kNcclBlocks x bucket generations
constexpr size_t kInboxRankWords = 16;         // u64 per (slot, rank): 8 statistics words as {tag, half} pairs
static_assert(sizeof(cudaIpcMemHandle_t) == FIMO_B200_STATS_PEER_HANDLE_BYTES, "cudaIpcMemHandle_t");
static_assert(sizeof(ncclUniqueId) == FIMO_B200_STATS_NCCL_ID_BYTES, "ncclUniqueId");

bool cuda_ok(cudaError_t err, const char *what) {
    if (err == cudaSuccess) return true;
    std::fprintf(stderr, "fimo_worlds_b200: stats exchange: %s: %s\n", what, cudaGetErrorString(err));
    (void)cudaGetLastError();
    return false;
}

struct Nccl { // the five entry points the exchange uses
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};
Nccl &nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char *name : {"libnccl.so.2", "libnccl.so"})
            if ((n.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL))) break;
        if (!n.lib) {
            std::fprintf(stderr, "fimo_worlds_b200: stats exchange: libnccl.so.2 not found (%s)\n", dlerror());
            return;
        }
        n.GetUniqueId = (decltype(n.GetUniqueId))dlsym(n.lib, "ncclGetUniqueId");
        n.CommInitRank = (decltype(n.CommInitRank))dlsym(n.lib, "ncclCommInitRank");
        n.AllReduce = (decltype(n.AllReduce))dlsym(n.lib, "ncclAllReduce");
        n.CommDestroy = (decltype(n.CommDestroy))dlsym(n.lib, "ncclCommDestroy");
        n.GetErrorString = (decltype(n.GetErrorString))dlsym(n.lib, "ncclGetErrorString");
        n.ok = n.GetUniqueId && n.CommInitRank && n.AllReduce && n.CommDestroy && n.GetErrorString;
    });
    return n;
}
bool nccl_ok(ncclResult_t r, const char *what) {
    if (r == ncclSuccess) return true;
    std::fprintf(stderr, "fimo_worlds_b200: stats exchange: %s: %s\n", what, nccl().GetErrorString ? nccl().GetErrorString(r) : "?");
    return false;
}

struct Guard {
    int previous = -1;
    explicit Guard(int device) {
        if (cudaGetDevice(&previous) != cudaSuccess) previous = -1, (void)cudaGetLastError();
        (void)cudaSetDevice(device);
    }
    ~Guard() {
        if (previous >= 0) (void)cudaSetDevice(previous);
    }
};

} // namespace

struct StatsExchange {
    int device = 0;
    cudaStream_t side = nullptr; // collectives (NCCL) and read-backs; never the group's stream
    std::mutex m;
    // PEERS
    unsigned long long *inbox = nullptr; // [slots][kStatsMaxRanks][16]
    uint32_t slots = 0, rank = 0, nranks = 0;
    PeerTable *table = nullptr;          // HBM copy the kernels read
    std::vector<void *> opened;          // peers' inboxes (cudaIpcOpenMemHandle)
    // NCCL
    ncclComm_t comm = nullptr;
    unsigned long long *ring = nullptr; // [kNcclBlocks * bucket][8]
    uint32_t bucket = 0;
    cudaEvent_t written[kNcclBlocks] = {}, reduced[kNcclBlocks] = {};
    uint64_t launched = 0, flushed = 0; // generations enqueued / all-reduced so far
    cudaStream_t last_stream = nullptr;
    unsigned long long *host = nullptr; // pinned read-back buffer
};

StatsExchange *stats_exchange_create(int device) {
    Guard guard(device);
    StatsExchange *x = new StatsExchange();
    x->device = device;
    if (!cuda_ok(cudaStreamCreateWithFlags(&x->side, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !cuda_ok(cudaHostAlloc((void **)&x->host, kStatsMaxRanks * kInboxRankWords * 8, cudaHostAllocDefault), "cudaHostAlloc")) {
        stats_exchange_destroy(x);
        return nullptr;
    }
    return x;
}

void stats_exchange_destroy(StatsExchange *x) {
    if (!x) return;
    Guard guard(x->device);
    if (x->side) cudaStreamSynchronize(x->side);
    for (void *p : x->opened) cudaIpcCloseMemHandle(p);
    if (x->comm && nccl().ok) nccl().CommDestroy(x->comm);
    for (uint32_t i = 0; i < kNcclBlocks; i++) {
        if (x->written[i]) cudaEventDestroy(x->written[i]);
        if (x->reduced[i]) cudaEventDestroy(x->reduced[i]);
    }
    if (x->inbox) cudaFree(x->inbox);
    if (x->table) cudaFree(x->table);
    if (x->ring) cudaFree(x->ring);
    if (x->host) cudaFreeHost(x->host);
    if (x->side) cudaStreamDestroy(x->side);
    (void)cudaGetLastError();
    delete x;
}

bool stats_exchange_peers_open(StatsExchange *x, uint32_t slots, void *handle_out) {
    Guard guard(x->device);
    std::lock_guard<std::mutex> lock(x->m);
    if (x->inbox || slots == 0) return false;
    const size_t bytes = (size_t)slots * kStatsMaxRanks * kInboxRankWords * 8;
    if (!cuda_ok(cudaMalloc((void **)&x->inbox, bytes), "cudaMalloc (inbox)") || !cuda_ok(cudaMemset(x->inbox, 0, bytes), "cudaMemset")) return false;
    x->slots = slots;
    cudaIpcMemHandle_t h;
    if (!cuda_ok(cudaIpcGetMemHandle(&h, x->inbox), "cudaIpcGetMemHandle")) return false;
    std::memcpy(handle_out, &h, sizeof h);
    return true;
}

bool stats_exchange_peers_bind(StatsExchange *x, const void *handles, uint32_t rank, uint32_t nranks) {
    Guard guard(x->device);
    std::lock_guard<std::mutex> lock(x->m);
    if (!x->inbox || x->table || nranks == 0 || nranks > (uint32_t)kStatsMaxRanks || rank >= nranks) return false;
    PeerTable t{};
    t.nranks = nranks, t.rank = rank, t.slots = x->slots;
    for (uint32_t r = 0; r < nranks; r++) {
        if (r == rank) {
            t.inbox[r] = x->inbox;
            continue;
        }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, (const char *)handles + (size_t)r * sizeof h, sizeof h);
        void *p = nullptr;
        if (!cuda_ok(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle")) return false;
        x->opened.push_back(p);
        t.inbox[r] = (unsigned long long *)p;
    }
    if (!cuda_ok(cudaMalloc((void **)&x->table, sizeof t), "cudaMalloc (peer table)") ||
        !cuda_ok(cudaMemcpy(x->table, &t, sizeof t, cudaMemcpyHostToDevice), "cudaMemcpy (peer table)"))
        return false;
    x->rank = rank, x->nranks = nranks;
    return true;
}

const void *stats_exchange_peer_table(StatsExchange *x) { return x ? x->table : nullptr; }

bool stats_nccl_unique_id(void *out) {
    if (!nccl().ok) return false;
    ncclUniqueId id;
    if (!nccl_ok(nccl().GetUniqueId(&id), "ncclGetUniqueId")) return false;
    std::memcpy(out, &id, sizeof id);
    return true;
}

bool stats_exchange_nccl_bind(StatsExchange *x, const void *unique_id, uint32_t rank, uint32_t nranks, uint32_t bucket) {
    if (!nccl().ok || x->comm || bucket == 0 || rank >= nranks) return false;
    Guard guard(x->device);
    std::lock_guard<std::mutex> lock(x->m);
    ncclUniqueId id;
    std::memcpy(&id, unique_id, sizeof id);
    if (!nccl_ok(nccl().CommInitRank(&x->comm, (int)nranks, id, (int)rank), "ncclCommInitRank")) return false;
    const size_t bytes = (size_t)kNcclBlocks * bucket * FIMO_B200_STAT_WORDS * 8;
    if (!cuda_ok(cudaMalloc((void **)&x->ring, bytes), "cudaMalloc (ring)") || !cuda_ok(cudaMemset(x->ring, 0, bytes), "cudaMemset")) return false;
    for (uint32_t i = 0; i < kNcclBlocks; i++)
        if (!cuda_ok(cudaEventCreateWithFlags(&x->written[i], cudaEventDisableTiming), "cudaEventCreate") ||
            !cuda_ok(cudaEventCreateWithFlags(&x->reduced[i], cudaEventDisableTiming), "cudaEventCreate"))
            return false;
    x->bucket = bucket;
    if (!x->nranks) x->rank = rank, x->nranks = nranks;
    // Warm the exact collective (same count, type, stream) once: the first all-reduce of a communicator sets up
    // channels and proxy threads, which must not land inside anybody's timed window.
    if (!nccl_ok(nccl().AllReduce(x->ring, x->ring, (size_t)bucket * FIMO_B200_STAT_WORDS, ncclUint64, ncclSum, x->comm, x->side), "ncclAllReduce (warm-up)"))
        return false;
    return cuda_ok(cudaStreamSynchronize(x->side), "cudaStreamSynchronize") && cuda_ok(cudaMemset(x->ring, 0, bytes), "cudaMemset");
}

void *stats_exchange_ring(StatsExchange *x, uint32_t *slots) {
    if (!x || !x->ring) return nullptr;
    *slots = kNcclBlocks * x->bucket;
    return x->ring;
}

namespace {
// m held.  All-reduce ring rows of generations [flushed, upto) — they lie in one block — on the side stream.
bool reduce_upto(StatsExchange *x, uint64_t upto) {
    if (upto <= x->flushed) return true;
    const uint32_t slots = kNcclBlocks * x->bucket;
    const uint32_t blk = (uint32_t)(((upto - 1) / x->bucket) % kNcclBlocks);
    const uint64_t lo = x->flushed % slots, n = upto - x->flushed;
    unsigned long long *rows = x->ring + lo * FIMO_B200_STAT_WORDS;
    bool ok = cuda_ok(cudaEventRecord(x->written[blk], x->last_stream), "cudaEventRecord") &&
              cuda_ok(cudaStreamWaitEvent(x->side, x->written[blk], 0), "cudaStreamWaitEvent") &&
              nccl_ok(nccl().AllReduce(rows, rows, (size_t)n * FIMO_B200_STAT_WORDS, ncclUint64, ncclSum, x->comm, x->side), "ncclAllReduce") &&
              cuda_ok(cudaEventRecord(x->reduced[blk], x->side), "cudaEventRecord");
    x->flushed = upto;
    return ok;
}
} // namespace

// A launch of n generations starting at generation `first` overwrites the rows of generations [first - slots,
// first + n - slots).  All of those must already have had their all-reduce enqueued (generation < flushed, and
// flushed > first - bucket) and before_launch must find the `reduced` event of every block the launch enters already
// recorded: both hold exactly when n <= slots - bucket.
uint32_t stats_exchange_max_gens(StatsExchange *x) {
    if (!x) return UINT32_MAX;
    std::lock_guard<std::mutex> lock(x->m);
    return x->comm ? (kNcclBlocks - 1) * x->bucket : UINT32_MAX;
}

bool stats_exchange_before_launch(StatsExchange *x, void *stream, uint32_t n) {
    if (!x || !x->comm) return true;
    Guard guard(x->device);
    std::lock_guard<std::mutex> lock(x->m);
    const uint32_t slots = kNcclBlocks * x->bucket;
    bool ok = true;
    for (uint64_t g = x->launched; g < x->launched + n; g++) // a block about to be overwritten: its collective must be done
        if (g % x->bucket == 0 && g >= slots) ok = cuda_ok(cudaStreamWaitEvent((cudaStream_t)stream, x->reduced[(g / x->bucket) % kNcclBlocks], 0), "cudaStreamWaitEvent") && ok;
    return ok;
}

bool stats_exchange_after_launch(StatsExchange *x, void *stream, uint32_t n) {
    if (!x) return true;
    std::lock_guard<std::mutex> lock(x->m);
    const uint64_t first = x->launched;
    x->launched += n;
    x->last_stream = (cudaStream_t)stream;
    if (!x->comm) return true;
    Guard guard(x->device);
    bool ok = true;
    for (uint64_t g = first; g < first + n; g++)
        if ((g + 1) % x->bucket == 0) ok = reduce_upto(x, g + 1) && ok;
    return ok;
}

bool stats_exchange_read(StatsExchange *x, uint64_t generation, uint32_t mode, uint64_t *out) {
    if (!x) return false;
    Guard guard(x->device);
    std::lock_guard<std::mutex> lock(x->m);
    if (mode == FIMO_B200_STATS_MODE_NCCL) {
        if (!x->comm || generation >= x->launched) return false;
        if (!reduce_upto(x, x->launched)) return false; // the partial block (collective: every rank is here)
        const uint32_t slots = kNcclBlocks * x->bucket;
        if (x->launched - generation > slots - x->bucket) return false; // already overwritten
        if (!cuda_ok(cudaMemcpyAsync(x->host, x->ring + (generation % slots) * FIMO_B200_STAT_WORDS, FIMO_B200_STAT_WORDS * 8, cudaMemcpyDeviceToHost, x->side),
                     "cudaMemcpyAsync") ||
            !cuda_ok(cudaStreamSynchronize(x->side), "cudaStreamSynchronize"))
            return false;
        for (int i = 0; i < FIMO_B200_STAT_WORDS; i++) out[i] = x->host[i];
        return true;
    }
    if (mode != FIMO_B200_STATS_MODE_PEERS || !x->table) return false;
    const unsigned long long tag = (generation + 1) & 0xFFFFFFFFull;
    const unsigned long long *slot = x->inbox + (generation % x->slots) * kStatsMaxRanks * kInboxRankWords;
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(10);
    for (;;) {
        if (!cuda_ok(cudaMemcpyAsync(x->host, slot, (size_t)x->nranks * kInboxRankWords * 8, cudaMemcpyDeviceToHost, x->side), "cudaMemcpyAsync") ||
            !cuda_ok(cudaStreamSynchronize(x->side), "cudaStreamSynchronize"))
            return false;
        bool complete = true, overwritten = false;
        for (size_t i = 0; i < (size_t)x->nranks * kInboxRankWords; i++) {
            const unsigned long long t = x->host[i] >> 32;
            if (t != tag) complete = false;
            if (t > tag) overwritten = true; // a later generation already took the slot
        }
        if (complete) break;
        if (overwritten || std::chrono::steady_clock::now() > deadline) {
            std::fprintf(stderr, "fimo_worlds_b200: stats exchange: generation %llu %s\n", (unsigned long long)generation,
                         overwritten ? "has been overwritten in the inbox" : "did not arrive from every rank within 10 s");
            return false;
        }
        std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
    for (int i = 0; i < FIMO_B200_STAT_WORDS; i++) {
        uint64_t sum = 0;
        for (uint32_t r = 0; r < x->nranks; r++) {
            const unsigned long long *w = x->host + (size_t)r * kInboxRankWords + 2 * i;
            sum += (w[0] & 0xFFFFFFFFull) | ((w[1] & 0xFFFFFFFFull) << 32);
        }
        out[i] = sum;
    }
    return true;
}

} // namespace fb::dev
