// device.cu — CUDA side of the scheduler: HBM resource store, the per-group executor that turns the
// compiled system DAG into a CUDA graph, and the completion thread that turns "generation finished on
// the device" into the reference's host-side protocol (generation counter, timeline semaphore, fence).
//
// Replaces, for device systems, the reference's per-system dispatch through the pool's event loop
// (modules/fimo_tasks/src/Pool.zig:660-765, CommandBuffer.zig:150-201,390-411): dependencies between
// systems become edges of a CUDA graph, so one generation is ONE graph launch and no host thread is
// touched between systems.  Generations of a group serialise by stream order (SystemGroup.zig:430-452
// does it with a timeline semaphore).
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <deque>
#include <mutex>
#include <string>
#include <thread>

#include "device.h"

namespace fb::dev {
namespace {

thread_local std::string t_cuda_error;

bool ok(cudaError_t err, const char *what) {
    if (err == cudaSuccess) return true;
    t_cuda_error = std::string(what) + ": " + cudaGetErrorString(err);
    std::fprintf(stderr, "fimo_worlds_b200: %s\n", t_cuda_error.c_str());
    (void)cudaGetLastError();
    return false;
}

struct DeviceGuard { // the calling thread may belong to anyone: restore its device
    int previous = -1;
    bool valid = false;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&previous) != cudaSuccess) {
            (void)cudaGetLastError();
            previous = -1;
        }
        valid = cudaSetDevice(device) == cudaSuccess;
        if (!valid) (void)cudaGetLastError();
    }
    ~DeviceGuard() {
        if (previous >= 0 && valid) cudaSetDevice(previous);
    }
};

// ---- completion thread ---------------------------------------------------------------------------
struct Completion {
    cudaEvent_t start, end;
    Executor *ex;
    void (*done)(void *);
    void *user;
    bool generation = true; // false: a stream-ordered transfer (not part of the per-generation timers)
    cudaEvent_t landed = nullptr; // a generation with result copies behind it: completes when they have landed
    uint64_t seq = 0;             // position in the executor's queue (Executor::posted_seq)
    uint32_t batch = 1;           // generations the start / end pair brackets (executor_launch_batch)
    bool owns_events = true;      // the last generation of a batch recycles the shared events
};

struct Completer {
    std::mutex m;
    std::condition_variable cv;
    std::deque<Completion> pending;
    bool started = false;
    void run();
    void post(const Completion &c) {
        std::lock_guard<std::mutex> g(m);
        pending.push_back(c);
        if (!started) {
            started = true;
            std::thread([this] { run(); }).detach();
        }
        cv.notify_one();
    }
};
Completer &completer() {
    static Completer *c = new Completer(); // leaked on purpose: the thread may outlive static destructors
    return *c;
}

} // namespace

struct Executor {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::vector<cudaStream_t> lanes; // side streams for DAG fan-out
    std::vector<cudaEvent_t> node_events;
    cudaEvent_t fork_event = nullptr;
    std::vector<Node> nodes;
    FimoB200_LaunchInfo info{};
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    bool use_graph = false;
    uint32_t kernels = 0;
    std::atomic<uint64_t> kernels_launched{0}; // kernel launches of generations enqueued so far (copies not counted)

    std::mutex m; // event pool, in-flight count, timing
    std::condition_variable idle_cv;
    std::vector<cudaEvent_t> free_events;
    size_t in_flight = 0;
    double ms_sum = 0.0;
    uint64_t ms_count = 0;
    cudaEvent_t span_anchor = nullptr; // start of the first generation since the timers were read
    double span_ms = 0.0;              // anchor -> end of the latest generation (includes inter-generation gaps)
    std::atomic<double> last_ms{0.0};  // device time of the generation being retired

    // Staged uploads (executor_launch): the next generation's inputs cross PCIe on `copy_stream` into one of two HBM
    // staging slots while the previous generation still runs; in stream order only a device-to-device copy remains.
    cudaStream_t copy_stream = nullptr;
    void *staging[2] = {nullptr, nullptr};
    size_t staging_bytes = 0;
    cudaEvent_t staged[2] = {nullptr, nullptr}, slot_free[2] = {nullptr, nullptr};
    uint64_t staged_count = 0;
    // Packed cull read-back (executor_launch with a Readback): two HBM staging slots, drained by the copy engine on
    // `readback_stream` while the next generation runs; `rb_drained[s]` fences the reuse of slot s.
    cudaStream_t readback_stream = nullptr;
    void *rb_staging[2] = {nullptr, nullptr};
    size_t rb_bytes = 0;
    cudaEvent_t rb_packed[2] = {nullptr, nullptr}, rb_drained[2] = {nullptr, nullptr};
    uint64_t rb_count = 0;
    // Shadow mode (the executor is one fused world-step node that can write its cull outputs twice, Node::launch_shadow):
    // generation g leaves a second copy of mask / list / count in shadow set g & 1, the pack reads THAT copy on
    // `readback_stream` beside generation g + 1 (which overwrites the resources) and `rb_ready[s]` orders it behind its
    // generation; generation g + 2 waits for `rb_packed[s]` before it overwrites the set.
    CullShadow rb_shadow[2] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
    size_t rb_shadow_words[3] = {0, 0, 0};
    cudaEvent_t rb_ready[2] = {nullptr, nullptr};
    // Host memory that queued device -> host copies have yet to write.  An upload that reads such memory (a result fed
    // back as the next step's input without waiting) must stay in stream order behind that copy: it is not staged.
    struct HostRange {
        const char *lo, *hi;
        uint64_t seq;
    };
    std::vector<HostRange> host_writes; // guarded by m
    uint64_t posted_seq = 0, retired_seq = 0;
    // m held.  Registers `writes` under a new sequence number; returns whether any of `reads` overlaps a pending write.
    bool post(const Copy *reads, size_t n_reads, const Copy *writes, size_t n_writes, uint64_t *seq) {
        host_writes.erase(std::remove_if(host_writes.begin(), host_writes.end(), [&](const HostRange &r) { return r.seq <= retired_seq; }),
                          host_writes.end());
        bool overlap = false;
        for (size_t i = 0; i < n_reads && !overlap; i++) {
            const char *lo = (const char *)reads[i].src, *hi = lo + reads[i].bytes;
            for (const HostRange &r : host_writes)
                if (lo < r.hi && r.lo < hi) overlap = true;
        }
        *seq = ++posted_seq;
        for (size_t i = 0; i < n_writes; i++) host_writes.push_back(HostRange{(const char *)writes[i].dst, (const char *)writes[i].dst + writes[i].bytes, *seq});
        return overlap;
    }

    cudaEvent_t take_event() {
        {
            std::lock_guard<std::mutex> g(m);
            if (!free_events.empty()) {
                cudaEvent_t e = free_events.back();
                free_events.pop_back();
                return e;
            }
        }
        cudaEvent_t e = nullptr;
        ok(cudaEventCreate(&e), "cudaEventCreate");
        return e;
    }
};

namespace {

void Completer::run() {
    for (;;) {
        Completion c;
        {
            std::unique_lock<std::mutex> g(m);
            cv.wait(g, [&] { return !pending.empty(); });
            c = pending.front();
            pending.pop_front();
        }
        cudaSetDevice(c.ex->device);
        ok(cudaEventSynchronize(c.landed ? c.landed : c.end), "cudaEventSynchronize");
        float ms = 0.0f;
        const bool timed = c.generation && cudaEventElapsedTime(&ms, c.start, c.end) == cudaSuccess;
        if (c.generation && !timed) (void)cudaGetLastError();
        ms /= (float)c.batch; // a multi-generation launch: every generation gets its share
        if (timed) c.ex->last_ms.store((double)ms, std::memory_order_relaxed);
        if (c.done) c.done(c.user);
        Executor *ex = c.ex;
        std::lock_guard<std::mutex> g(ex->m);
        if (timed) {
            ex->ms_sum += ms;
            ex->ms_count++;
        }
        if (!c.owns_events) {
            // an earlier generation of a batch: the events belong to the batch's last generation
        } else if (c.generation) {
            if (!ex->span_anchor) ex->span_anchor = c.start;
            else ex->free_events.push_back(c.start);
            float span = 0.0f;
            if (cudaEventElapsedTime(&span, ex->span_anchor, c.end) == cudaSuccess) ex->span_ms = span;
            else (void)cudaGetLastError();
        } else if (c.start) {
            ex->free_events.push_back(c.start);
        }
        if (c.owns_events) ex->free_events.push_back(c.end);
        if (c.landed) ex->free_events.push_back(c.landed);
        if (c.seq > ex->retired_seq) ex->retired_seq = c.seq;
        if (--ex->in_flight == 0) ex->idle_cv.notify_all();
    }
}

// Replays the DAG on ex->stream: node i runs on a lane stream, waits for its dependencies' events,
// records its own; the origin stream joins every lane at the end.  Used both live (no graph) and
// under stream capture (to build the graph).
bool replay(Executor *ex, bool capturing) {
    const size_t n = ex->nodes.size();
    while (ex->node_events.size() < n) {
        cudaEvent_t e = nullptr;
        if (!ok(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate")) return false;
        ex->node_events.push_back(e);
    }
    // lane assignment: a node continues the lane of its first dependency if nobody else has yet
    std::vector<int> lane(n, -1);
    std::vector<char> lane_continued(n, 0);
    int lanes_used = 1;
    for (size_t i = 0; i < n; i++) {
        int pick = -1;
        for (uint32_t d : ex->nodes[i].deps)
            if (!lane_continued[d]) {
                pick = lane[d];
                lane_continued[d] = 1;
                break;
            }
        if (pick < 0) pick = (i == 0) ? 0 : lanes_used++;
        lane[i] = pick;
    }
    while ((int)ex->lanes.size() < lanes_used - 1) {
        cudaStream_t s = nullptr;
        if (!ok(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
        ex->lanes.push_back(s);
    }
    auto stream_of = [&](int l) { return l == 0 ? ex->stream : ex->lanes[(size_t)l - 1]; };
    if (lanes_used > 1) {
        if (!ex->fork_event && !ok(cudaEventCreateWithFlags(&ex->fork_event, cudaEventDisableTiming), "cudaEventCreate"))
            return false;
        if (!ok(cudaEventRecord(ex->fork_event, ex->stream), "cudaEventRecord")) return false;
    }
    std::vector<char> lane_started(lanes_used, 0);
    lane_started[0] = 1;
    FimoB200_LaunchInfo info = ex->info;
    info.flags = capturing ? FIMO_B200_LAUNCH_FLAG_CAPTURING : 0u;
    for (size_t i = 0; i < n; i++) {
        cudaStream_t s = stream_of(lane[i]);
        if (!lane_started[lane[i]]) { // fork the lane from the origin so it belongs to this generation
            if (!ok(cudaStreamWaitEvent(s, ex->fork_event, 0), "cudaStreamWaitEvent")) return false;
            lane_started[lane[i]] = 1;
        }
        for (uint32_t d : ex->nodes[i].deps)
            if (lane[d] != lane[i] && !ok(cudaStreamWaitEvent(s, ex->node_events[d], 0), "cudaStreamWaitEvent"))
                return false;
        info.stream = s;
        if (ex->nodes[i].launch(&info) != FIMO_STATUS_OK) return false;
        if (!ok(cudaEventRecord(ex->node_events[i], s), "cudaEventRecord")) return false;
    }
    // join: the last node of every side lane
    std::vector<int> last_on_lane(lanes_used, -1);
    for (size_t i = 0; i < n; i++) last_on_lane[lane[i]] = (int)i;
    for (int l = 1; l < lanes_used; l++)
        if (last_on_lane[l] >= 0 &&
            !ok(cudaStreamWaitEvent(ex->stream, ex->node_events[(size_t)last_on_lane[l]], 0), "cudaStreamWaitEvent"))
            return false;
    return true;
}

void drop_graph(Executor *ex) {
    if (ex->graph_exec) cudaGraphExecDestroy(ex->graph_exec);
    if (ex->graph) cudaGraphDestroy(ex->graph);
    ex->graph_exec = nullptr;
    ex->graph = nullptr;
}

} // namespace

const char *last_cuda_error() { return t_cuda_error.c_str(); }

int resolve_device(int device) {
    if (device >= 0) return device;
    int current = 0;
    if (cudaGetDevice(&current) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    return current;
}

bool device_ok(int device) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) {
        t_cuda_error = "no CUDA device: the B200 path has no CPU fallback";
        (void)cudaGetLastError();
        return false;
    }
    device = resolve_device(device);
    if (device >= count) {
        t_cuda_error = "CUDA device ordinal out of range";
        return false;
    }
    return true;
}

void *hbm_alloc(int device, size_t bytes) {
    DeviceGuard guard(device);
    void *p = nullptr;
    if (!ok(cudaMalloc(&p, bytes ? bytes : 4), "cudaMalloc")) return nullptr;
    if (!ok(cudaMemset(p, 0, bytes ? bytes : 4), "cudaMemset")) {
        cudaFree(p);
        return nullptr;
    }
    return p;
}

void hbm_free(int device, void *ptr) {
    if (!ptr) return;
    DeviceGuard guard(device);
    ok(cudaFree(ptr), "cudaFree");
}

bool hbm_upload(int device, void *dst, const void *src, size_t bytes) {
    DeviceGuard guard(device);
    return ok(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice), "cudaMemcpy H2D");
}

bool hbm_download(int device, void *dst, const void *src, size_t bytes) {
    DeviceGuard guard(device);
    return ok(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost), "cudaMemcpy D2H");
}

Executor *executor_create(int device) {
    DeviceGuard guard(device);
    Executor *ex = new Executor();
    ex->device = device;
    if (!ok(cudaStreamCreateWithFlags(&ex->stream, cudaStreamNonBlocking), "cudaStreamCreate")) {
        delete ex;
        return nullptr;
    }
    // Without the copy stream uploads simply take the in-order path.
    int least = 0, greatest = 0;
    cudaDeviceGetStreamPriorityRange(&least, &greatest);
    bool staged = cudaStreamCreateWithPriority(&ex->copy_stream, cudaStreamNonBlocking, greatest) == cudaSuccess;
    for (int i = 0; i < 2 && staged; i++)
        staged = cudaEventCreateWithFlags(&ex->staged[i], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&ex->slot_free[i], cudaEventDisableTiming) == cudaSuccess;
    if (!staged) {
        cudaGetLastError();
        if (ex->copy_stream) cudaStreamDestroy(ex->copy_stream);
        ex->copy_stream = nullptr;
    }
    return ex;
}

void executor_wait_idle(Executor *ex) {
    std::unique_lock<std::mutex> g(ex->m);
    ex->idle_cv.wait(g, [&] { return ex->in_flight == 0; });
}

void executor_destroy(Executor *ex) {
    if (!ex) return;
    executor_wait_idle(ex);
    DeviceGuard guard(ex->device);
    cudaStreamSynchronize(ex->stream);
    drop_graph(ex);
    for (cudaEvent_t e : ex->free_events) cudaEventDestroy(e);
    for (cudaEvent_t e : ex->node_events) cudaEventDestroy(e);
    if (ex->fork_event) cudaEventDestroy(ex->fork_event);
    for (cudaStream_t s : ex->lanes) cudaStreamDestroy(s);
    if (ex->copy_stream) {
        cudaStreamSynchronize(ex->copy_stream);
        cudaStreamDestroy(ex->copy_stream);
    }
    if (ex->readback_stream) {
        cudaStreamSynchronize(ex->readback_stream);
        cudaStreamDestroy(ex->readback_stream);
    }
    for (int i = 0; i < 2; i++) {
        if (ex->staged[i]) cudaEventDestroy(ex->staged[i]);
        if (ex->slot_free[i]) cudaEventDestroy(ex->slot_free[i]);
        if (ex->staging[i]) cudaFree(ex->staging[i]);
        if (ex->rb_packed[i]) cudaEventDestroy(ex->rb_packed[i]);
        if (ex->rb_drained[i]) cudaEventDestroy(ex->rb_drained[i]);
        if (ex->rb_staging[i]) cudaFree(ex->rb_staging[i]);
        if (ex->rb_ready[i]) cudaEventDestroy(ex->rb_ready[i]);
        for (uint32_t *q : {ex->rb_shadow[i].mask, ex->rb_shadow[i].list, ex->rb_shadow[i].count})
            if (q) cudaFree(q);
    }
    cudaStreamDestroy(ex->stream);
    delete ex;
}

bool executor_compile(Executor *ex, std::vector<Node> nodes, const FimoB200_LaunchInfo &info, bool use_graph) {
    executor_wait_idle(ex); // never rebuild under a generation that is still running
    DeviceGuard guard(ex->device);
    drop_graph(ex);
    ex->nodes = std::move(nodes);
    ex->info = info;
    ex->info.device = ex->device;
    ex->use_graph = false;
    ex->kernels = 0;
    for (const Node &n : ex->nodes) ex->kernels += n.kernels;
    if (!use_graph || ex->nodes.empty()) return true;
    if (!ok(cudaStreamBeginCapture(ex->stream, cudaStreamCaptureModeRelaxed), "cudaStreamBeginCapture")) return false;
    const bool replayed = replay(ex, true);
    cudaGraph_t graph = nullptr;
    const bool ended = ok(cudaStreamEndCapture(ex->stream, &graph), "cudaStreamEndCapture");
    if (!replayed || !ended || !graph) {
        if (graph) cudaGraphDestroy(graph);
        return false;
    }
    ex->graph = graph;
    if (!ok(cudaGraphInstantiate(&ex->graph_exec, ex->graph, 0), "cudaGraphInstantiate")) return false;
    ok(cudaGraphUpload(ex->graph_exec, ex->stream), "cudaGraphUpload");
    ex->use_graph = true;
    return true;
}

// ---- small host <-> HBM copies as ONE kernel --------------------------------------------------------------------
// A copy-engine transfer costs ~8-10 us of stream time whatever its size; the per-step inputs / results of
// fimo_b200_group_schedule_io are a handful of small resources (frusta, dt, counts, statistics).  When every host
// buffer is page-locked (so the GPU can address it, UVA) the whole batch is moved by one kernel launch that
// reads / writes host memory over PCIe directly: 128-bit accesses, all requests in flight at once.
namespace {
constexpr int kIoMaxCopies = 8;
constexpr size_t kIoMaxBytes = 8u << 20; // beyond this the copy engine wins
struct IoBatch {
    void *dst[kIoMaxCopies];
    const void *src[kIoMaxCopies];
    unsigned long long bytes[kIoMaxCopies];
};
__global__ void __launch_bounds__(128) k_io_copy(const IoBatch b) {
    const int c = blockIdx.y;
    const unsigned long long bytes = b.bytes[c];
    const size_t stride = (size_t)gridDim.x * blockDim.x, first = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if ((((uintptr_t)b.dst[c] | (uintptr_t)b.src[c] | bytes) & 15u) == 0) {
        uint4 *dst = (uint4 *)b.dst[c];
        const uint4 *src = (const uint4 *)b.src[c];
        for (size_t i = first; i < bytes / 16; i += stride) dst[i] = src[i];
    } else {
        uint32_t *dst = (uint32_t *)b.dst[c];
        const uint32_t *src = (const uint32_t *)b.src[c];
        for (size_t i = first; i < bytes / 4; i += stride) dst[i] = src[i];
    }
}

// device-visible alias of a page-locked host pointer, or nullptr (pageable memory: the copy engine must stage it)
void *device_alias(const void *host) {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, host) != cudaSuccess) {
        (void)cudaGetLastError();
        return nullptr;
    }
    return attr.type == cudaMemoryTypeHost ? attr.devicePointer : nullptr;
}

bool launch_io_batch(cudaStream_t stream, const IoBatch &batch, size_t n, size_t largest) {
    // 128-thread CTAs at <= 32 registers fit beside the resident world-step CTAs, so the staged PCIe leg can overlap them
    const unsigned blocks = (unsigned)std::min<size_t>(std::max<size_t>((largest / 16 + 127) / 128, 1), 296);
    k_io_copy<<<dim3(blocks, (unsigned)n), 128, 0, stream>>>(batch);
    return ok(cudaGetLastError(), "k_io_copy");
}

// `ex` != nullptr (uploads only): stage through the executor's HBM slots on its copy stream, see Executor::staging.
bool enqueue_copies(cudaStream_t stream, const Copy *copies, size_t n, bool to_device, Executor *ex = nullptr) {
    if (n == 0) return true;
    bool by_kernel = n <= (size_t)kIoMaxCopies && !std::getenv("FIMO_B200_IO_MEMCPY");
    IoBatch batch{};
    size_t total = 0, largest = 0, padded = 0;
    for (size_t i = 0; i < n && by_kernel; i++) {
        void *alias = device_alias(to_device ? copies[i].src : copies[i].dst);
        if (!alias || copies[i].bytes % 4) by_kernel = false;
        batch.dst[i] = to_device ? copies[i].dst : alias;
        batch.src[i] = to_device ? alias : copies[i].src;
        batch.bytes[i] = copies[i].bytes;
        total += copies[i].bytes, largest = std::max(largest, copies[i].bytes);
        padded += (copies[i].bytes + 15 + 255) & ~(size_t)255;
    }
    if (by_kernel && total <= kIoMaxBytes && ex && to_device && ex->copy_stream && !std::getenv("FIMO_B200_IO_NO_STAGING")) {
        if (!ex->staging[0] && padded <= kIoMaxBytes) { // slots are made once; a later, larger batch takes the direct path
            const size_t want = std::max<size_t>(padded, 1u << 20);
            if (cudaMalloc(&ex->staging[0], want) == cudaSuccess && cudaMalloc(&ex->staging[1], want) == cudaSuccess) ex->staging_bytes = want;
            else cudaGetLastError();
        }
        if (padded <= ex->staging_bytes) {
            const int slot = (int)(ex->staged_count & 1);
            IoBatch in = batch, out = batch;
            size_t at = 0;
            for (size_t i = 0; i < n; i++) {
                // keep the slot address congruent to the resource pointer modulo 16 so that the device-to-device
                // leg stays on the 128-bit path
                char *p = (char *)ex->staging[slot] + at + ((uintptr_t)batch.dst[i] & 15u);
                in.dst[i] = p, out.src[i] = p;
                at += (copies[i].bytes + 15 + 255) & ~(size_t)255;
            }
            bool queued = true;
            if (ex->staged_count >= 2) queued = ok(cudaStreamWaitEvent(ex->copy_stream, ex->slot_free[slot], 0), "cudaStreamWaitEvent");
            // The PCIe leg goes to the copy engine: it runs beside the resident world-step CTAs whatever they occupy
            // (a copy kernel on the side stream waits for free SM resources and hides nothing: measured).
            for (size_t i = 0; i < n && queued; i++)
                queued = ok(cudaMemcpyAsync(in.dst[i], copies[i].src, copies[i].bytes, cudaMemcpyHostToDevice, ex->copy_stream),
                            "cudaMemcpyAsync H2D (staging)");
            queued = queued && ok(cudaEventRecord(ex->staged[slot], ex->copy_stream), "cudaEventRecord");
            queued = queued && ok(cudaStreamWaitEvent(stream, ex->staged[slot], 0), "cudaStreamWaitEvent");
            queued = queued && launch_io_batch(stream, out, n, largest);
            queued = queued && ok(cudaEventRecord(ex->slot_free[slot], stream), "cudaEventRecord");
            ex->staged_count++;
            return queued;
        }
    }
    if (by_kernel && total <= kIoMaxBytes) return launch_io_batch(stream, batch, n, largest);
    for (size_t i = 0; i < n; i++)
        if (!ok(cudaMemcpyAsync(copies[i].dst, copies[i].src, copies[i].bytes, to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, stream),
                to_device ? "cudaMemcpyAsync H2D" : "cudaMemcpyAsync D2H"))
            return false;
    return true;
}
} // namespace

// ---- packed cull read-back -------------------------------------------------------------------------------------
// Host / staging layout, u32 words: header[8] = {total visible, n_worlds, mask_words, list capacity (entries), truncated,
// generation lo, generation hi, entities}; counts[n_worlds]; offsets[n_worlds] (exclusive prefix of counts);
// mask[n_worlds][mask_words]; list[capacity] = the worlds' ordered visible lists back to back.
namespace {
constexpr int kScanThreads = 1024;
__global__ void __launch_bounds__(kScanThreads) k_readback_scan(const uint32_t *__restrict__ count, uint32_t W, uint32_t mask_words, uint32_t entities,
                                                                uint32_t capacity, unsigned long long generation, uint32_t *__restrict__ out) {
    __shared__ uint32_t s_warp[kScanThreads / 32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t *counts = out + 8, *offsets = out + 8 + W;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < W; base += kScanThreads) {
        const uint32_t w = base + tid;
        const uint32_t c = w < W ? count[w] : 0u;
        uint32_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t before = s_carry;
        for (int i = 0; i < warp; i++) before += s_warp[i];
        if (w < W) counts[w] = c, offsets[w] = before + incl - c;
        __syncthreads();
        if (tid == kScanThreads - 1) s_carry = before + incl;
        __syncthreads();
    }
    if (tid == 0) {
        const uint32_t total = s_carry;
        out[0] = total, out[1] = W, out[2] = mask_words, out[3] = capacity, out[4] = total > capacity ? 1u : 0u;
        out[5] = (uint32_t)(generation & 0xFFFFFFFFull), out[6] = (uint32_t)(generation >> 32), out[7] = entities;
    }
}
__global__ void __launch_bounds__(128) k_readback_pack(const uint32_t *__restrict__ mask, const uint32_t *__restrict__ list, uint32_t W, uint32_t mask_words,
                                                       uint32_t entities, uint32_t capacity, uint32_t *__restrict__ out) {
    const uint32_t w = blockIdx.x;
    const uint32_t n = out[8 + w], at = out[8 + W + w];
    uint32_t *dst_mask = out + 8 + 2ull * W + (uint64_t)w * mask_words;
    uint32_t *dst_list = out + 8 + 2ull * W + (uint64_t)W * mask_words;
    const uint32_t *src_mask = mask + (uint64_t)w * mask_words, *src_list = list + (uint64_t)w * entities;
    for (uint32_t i = threadIdx.x; i < mask_words; i += blockDim.x) dst_mask[i] = src_mask[i];
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
        if (at + i < capacity) dst_list[at + i] = src_list[i];
}

// Batches of up to kFusedPackWorlds worlds: ONE kernel.  CTA w sums the counts of the worlds before it itself (at most
// 64 coalesced L2 loads per thread) instead of waiting for a single-CTA scan kernel in front — one launch boundary and
// one serial 1-CTA kernel fewer per step on the e2e path.
constexpr uint32_t kFusedPackWorlds = 8192;
__global__ void __launch_bounds__(128) k_readback_scan_pack(const uint32_t *__restrict__ count, const uint32_t *__restrict__ mask,
                                                            const uint32_t *__restrict__ list, uint32_t W, uint32_t mask_words, uint32_t entities,
                                                            uint32_t capacity, unsigned long long generation, uint32_t *__restrict__ out) {
    __shared__ uint32_t s_warp[4];
    const uint32_t w = blockIdx.x, tid = threadIdx.x;
    uint32_t mine = 0;
    for (uint32_t i = tid; i < w; i += 128) mine += __ldcg(count + i);
    mine = __reduce_add_sync(0xFFFFFFFFu, mine);
    if ((tid & 31) == 0) s_warp[tid >> 5] = mine;
    __syncthreads();
    const uint32_t at = s_warp[0] + s_warp[1] + s_warp[2] + s_warp[3], n = __ldcg(count + w);
    if (tid == 0) {
        out[8 + w] = n, out[8 + W + w] = at;
        if (w == W - 1) {
            const uint32_t total = at + n;
            out[0] = total, out[1] = W, out[2] = mask_words, out[3] = capacity, out[4] = total > capacity ? 1u : 0u;
            out[5] = (uint32_t)(generation & 0xFFFFFFFFull), out[6] = (uint32_t)(generation >> 32), out[7] = entities;
        }
    }
    uint32_t *dst_mask = out + 8 + 2ull * W + (uint64_t)w * mask_words;
    uint32_t *dst_list = out + 8 + 2ull * W + (uint64_t)W * mask_words;
    const uint32_t *src_mask = mask + (uint64_t)w * mask_words, *src_list = list + (uint64_t)w * entities;
    for (uint32_t i = tid; i < mask_words; i += 128) dst_mask[i] = src_mask[i];
    for (uint32_t i = tid; i < n; i += 128)
        if (at + i < capacity) dst_list[at + i] = src_list[i];
}

// Before the generation is enqueued: stream, events, staging slots (and, in shadow mode, the shadow sets) exist; returns
// the slot this step uses.  Shadow mode: the generation about to be launched overwrites shadow set `slot`, which the
// pack of two steps ago read on the read-back stream.
bool readback_prepare(Executor *ex, const Readback &rb, bool shadow, int *slot_out) {
    const uint64_t fixed_words = 8 + 2 * rb.n_worlds + rb.n_worlds * rb.mask_words;
    if (rb.host_bytes < fixed_words * 4 || rb.n_worlds > 0xFFFFFFFFull) return false;
    const uint64_t capacity = std::min<uint64_t>(rb.host_bytes / 4 - fixed_words, rb.n_worlds * rb.entities);
    const size_t bytes = (size_t)(fixed_words + capacity) * 4;
    if (!ex->readback_stream) {
        if (!ok(cudaStreamCreateWithFlags(&ex->readback_stream, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
        for (int i = 0; i < 2; i++)
            if (!ok(cudaEventCreateWithFlags(&ex->rb_packed[i], cudaEventDisableTiming), "cudaEventCreate") ||
                !ok(cudaEventCreateWithFlags(&ex->rb_drained[i], cudaEventDisableTiming), "cudaEventCreate") ||
                !ok(cudaEventCreateWithFlags(&ex->rb_ready[i], cudaEventDisableTiming), "cudaEventCreate"))
                return false;
    }
    if (bytes > ex->rb_bytes) { // (re)size the slots: cold path, drains what is in flight
        if (!ok(cudaStreamSynchronize(ex->readback_stream), "cudaStreamSynchronize")) return false;
        for (int i = 0; i < 2; i++) {
            if (ex->rb_staging[i]) cudaFree(ex->rb_staging[i]);
            ex->rb_staging[i] = nullptr;
            if (!ok(cudaMalloc(&ex->rb_staging[i], bytes), "cudaMalloc (read-back staging)")) return false;
        }
        ex->rb_bytes = bytes, ex->rb_count = 0;
    }
    if (shadow) {
        const size_t want[3] = {(size_t)(rb.n_worlds * rb.mask_words), (size_t)(rb.n_worlds * rb.entities), (size_t)rb.n_worlds};
        if (want[0] > ex->rb_shadow_words[0] || want[1] > ex->rb_shadow_words[1] || want[2] > ex->rb_shadow_words[2]) { // cold path
            if (!ok(cudaStreamSynchronize(ex->readback_stream), "cudaStreamSynchronize") || !ok(cudaStreamSynchronize(ex->stream), "cudaStreamSynchronize")) return false;
            for (int i = 0; i < 2; i++) {
                uint32_t **set[3] = {&ex->rb_shadow[i].mask, &ex->rb_shadow[i].list, &ex->rb_shadow[i].count};
                for (int k = 0; k < 3; k++) {
                    if (*set[k]) cudaFree(*set[k]);
                    *set[k] = nullptr;
                    if (!ok(cudaMalloc((void **)set[k], std::max<size_t>(want[k], 1) * 4), "cudaMalloc (read-back shadow)")) return false;
                }
            }
            for (int k = 0; k < 3; k++) ex->rb_shadow_words[k] = want[k];
            ex->rb_count = 0;
        }
    }
    const int slot = (int)(ex->rb_count & 1);
    // the staging slot (in-stream pack) or the shadow set (pack beside the next generation) of two steps ago is free again
    if (ex->rb_count >= 2 && !ok(cudaStreamWaitEvent(ex->stream, shadow ? ex->rb_packed[slot] : ex->rb_drained[slot], 0), "cudaStreamWaitEvent")) return false;
    *slot_out = slot;
    return true;
}

// Behind the generation: scan + pack into the staging slot, one copy-engine transfer to the host, `landed`.
bool enqueue_readback(Executor *ex, const Readback &rb, cudaEvent_t landed, int slot, bool shadow) {
    const uint64_t fixed_words = 8 + 2 * rb.n_worlds + rb.n_worlds * rb.mask_words;
    const uint64_t capacity = std::min<uint64_t>(rb.host_bytes / 4 - fixed_words, rb.n_worlds * rb.entities);
    const size_t bytes = (size_t)(fixed_words + capacity) * 4;
    uint32_t *out = (uint32_t *)ex->rb_staging[slot];
    cudaStream_t pack_stream = shadow ? ex->readback_stream : ex->stream;
    const uint32_t *count = shadow ? ex->rb_shadow[slot].count : rb.count, *mask = shadow ? ex->rb_shadow[slot].mask : rb.mask,
                   *list = shadow ? ex->rb_shadow[slot].list : rb.list;
    bool queued = true;
    if (shadow) // the pack waits for its generation only; the staging slot's previous transfer precedes it on the same stream
        queued = ok(cudaEventRecord(ex->rb_ready[slot], ex->stream), "cudaEventRecord") &&
                 ok(cudaStreamWaitEvent(ex->readback_stream, ex->rb_ready[slot], 0), "cudaStreamWaitEvent");
    if (rb.n_worlds <= kFusedPackWorlds && !std::getenv("FIMO_B200_READBACK_TWO_KERNELS")) {
        k_readback_scan_pack<<<(unsigned)rb.n_worlds, 128, 0, pack_stream>>>(count, mask, list, (uint32_t)rb.n_worlds, (uint32_t)rb.mask_words,
                                                                              (uint32_t)rb.entities, (uint32_t)capacity, rb.generation, out);
    } else {
        k_readback_scan<<<1, kScanThreads, 0, pack_stream>>>(count, (uint32_t)rb.n_worlds, (uint32_t)rb.mask_words, (uint32_t)rb.entities, (uint32_t)capacity,
                                                             rb.generation, out);
        k_readback_pack<<<(unsigned)rb.n_worlds, 128, 0, pack_stream>>>(mask, list, (uint32_t)rb.n_worlds, (uint32_t)rb.mask_words, (uint32_t)rb.entities,
                                                                         (uint32_t)capacity, out);
    }
    queued = ok(cudaGetLastError(), "k_readback_pack") && queued;
    queued = queued && ok(cudaEventRecord(ex->rb_packed[slot], pack_stream), "cudaEventRecord");
    if (!shadow) queued = queued && ok(cudaStreamWaitEvent(ex->readback_stream, ex->rb_packed[slot], 0), "cudaStreamWaitEvent");
    queued = queued && ok(cudaMemcpyAsync(rb.host, out, bytes, cudaMemcpyDeviceToHost, ex->readback_stream), "cudaMemcpyAsync D2H (read-back)");
    queued = queued && ok(cudaEventRecord(ex->rb_drained[slot], ex->readback_stream), "cudaEventRecord");
    queued = queued && ok(cudaEventRecord(landed, ex->readback_stream), "cudaEventRecord");
    ex->rb_count++;
    return queued;
}
} // namespace

bool executor_launch(Executor *ex, void (*done)(void *), void *user, const Copy *uploads, size_t n_up, const Copy *downloads,
                     size_t n_down, const Readback *readback) {
    DeviceGuard guard(ex->device);
    Completion c{ex->take_event(), ex->take_event(), ex, done, user};
    if (!c.start || !c.end) return false;
    if ((n_down || readback) && !(c.landed = ex->take_event())) return false;
    bool busy, feedback;
    {
        std::lock_guard<std::mutex> g(ex->m);
        busy = ex->in_flight++ > 0;
        std::vector<Copy> writes(downloads, downloads + n_down); // host memory this step will write
        if (readback) writes.push_back(Copy{readback->host, nullptr, readback->host_bytes});
        feedback = ex->post(uploads, n_up, writes.data(), writes.size(), &c.seq);
    }
    // staging pays only when an earlier generation is still running for the PCIe leg to hide behind, and is only
    // correct when no queued result copy has yet to write the memory this upload reads
    bool launched = enqueue_copies(ex->stream, uploads, n_up, true, busy && !feedback ? ex : nullptr); // this generation's inputs, in stream order in front of it
    // read-back: the pack runs beside the NEXT generation when the executor is one fused world-step node that can leave a
    // second copy of its cull outputs (FIMO_B200_READBACK_IN_STREAM=1: pack in stream order, A/B)
    const bool shadow = readback && ex->nodes.size() == 1 && ex->nodes[0].launch_shadow && !std::getenv("FIMO_B200_READBACK_IN_STREAM");
    int rb_slot = 0;
    if (readback) launched = readback_prepare(ex, *readback, shadow, &rb_slot) && launched;
    launched = ok(cudaEventRecord(c.start, ex->stream), "cudaEventRecord") && launched;
    if (launched) {
        if (shadow) {
            FimoB200_LaunchInfo info = ex->info;
            info.stream = ex->stream, info.flags = 0u; // a plain launch: the shadow set alternates from step to step
            launched = ex->nodes[0].launch_shadow(&info, &ex->rb_shadow[rb_slot]) == FIMO_STATUS_OK;
        } else if (ex->use_graph) launched = ok(cudaGraphLaunch(ex->graph_exec, ex->stream), "cudaGraphLaunch");
        else launched = replay(ex, false);
    }
    launched = ok(cudaEventRecord(c.end, ex->stream), "cudaEventRecord") && launched;
    ex->kernels_launched.fetch_add(ex->kernels, std::memory_order_relaxed);
    if (launched) launched = enqueue_copies(ex->stream, downloads, n_down, false); // its results, behind it
    if (readback) launched = launched && enqueue_readback(ex, *readback, c.landed, rb_slot, shadow);
    if (c.landed && (!readback || !launched)) launched = ok(cudaEventRecord(c.landed, ex->stream), "cudaEventRecord") && launched;
    if (!launched) c.done = nullptr; // the caller reports the failure itself
    completer().post(c);
    return launched;
}

uint32_t executor_max_gens(Executor *ex) {
    return ex->nodes.size() == 1 && ex->nodes[0].launch_gens ? std::max<uint32_t>(ex->nodes[0].max_gens, 1) : 1;
}

bool executor_launch_batch(Executor *ex, uint32_t gens, void (*done)(void *), void *const *users) {
    if (gens == 0 || gens > executor_max_gens(ex)) return false;
    DeviceGuard guard(ex->device);
    Completion c{ex->take_event(), ex->take_event(), ex, done, nullptr};
    if (!c.start || !c.end) return false;
    c.batch = gens;
    {
        std::lock_guard<std::mutex> g(ex->m);
        ex->in_flight += gens;
        c.seq = ++ex->posted_seq;
    }
    FimoB200_LaunchInfo info = ex->info;
    info.stream = ex->stream, info.flags = 0u;
    bool launched = ok(cudaEventRecord(c.start, ex->stream), "cudaEventRecord");
    launched = launched && ex->nodes[0].launch_gens(&info, gens) == FIMO_STATUS_OK; // a plain launch: the generation count is a kernel argument
    launched = ok(cudaEventRecord(c.end, ex->stream), "cudaEventRecord") && launched;
    ex->kernels_launched.fetch_add(ex->nodes[0].kernels_gens, std::memory_order_relaxed); // one launch for all `gens` generations
    for (uint32_t i = 0; i < gens; i++) {
        Completion one = c;
        one.user = users[i];
        one.owns_events = i + 1 == gens;
        if (!launched) one.done = nullptr; // the caller reports the failure itself
        completer().post(one);
    }
    return launched;
}

bool executor_transfer(Executor *ex, void *dst, const void *src, size_t bytes, bool to_device, void (*done)(void *), void *user) {
    DeviceGuard guard(ex->device);
    Completion c{nullptr, ex->take_event(), ex, done, user, false};
    if (!c.end) return false;
    {
        std::lock_guard<std::mutex> g(ex->m);
        ex->in_flight++;
        const Copy copy{dst, src, bytes};
        ex->post(nullptr, 0, &copy, to_device ? 0 : 1, &c.seq);
    }
    bool queued = ok(cudaMemcpyAsync(dst, src, bytes, to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, ex->stream),
                     "cudaMemcpyAsync");
    queued = ok(cudaEventRecord(c.end, ex->stream), "cudaEventRecord") && queued;
    if (!queued) c.done = nullptr;
    completer().post(c);
    return queued;
}

bool enqueue_host_node(void *stream, const Copy *d2h, size_t n_d2h, void (*fn)(void *), void *user, const Copy *h2d, size_t n_h2d) {
    cudaStream_t s = (cudaStream_t)stream;
    for (size_t i = 0; i < n_d2h; i++)
        if (!ok(cudaMemcpyAsync(d2h[i].dst, d2h[i].src, d2h[i].bytes, cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync D2H (mirror)")) return false;
    if (!ok(cudaLaunchHostFunc(s, fn, user), "cudaLaunchHostFunc")) return false;
    for (size_t i = 0; i < n_h2d; i++)
        if (!ok(cudaMemcpyAsync(h2d[i].dst, h2d[i].src, h2d[i].bytes, cudaMemcpyHostToDevice, s), "cudaMemcpyAsync H2D (mirror)")) return false;
    return true;
}

void *host_alloc_pinned(size_t bytes) {
    void *p = nullptr;
    if (!ok(cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault), "cudaHostAlloc")) return nullptr;
    return p;
}
void host_free_pinned(void *p) {
    if (p) ok(cudaFreeHost(p), "cudaFreeHost");
}

void *executor_stream(Executor *ex) { return ex->stream; }
bool executor_uses_graph(Executor *ex) { return ex->use_graph; }
uint32_t executor_kernels_per_generation(Executor *ex) { return ex->kernels; }
uint64_t executor_kernels_launched(Executor *ex) { return ex->kernels_launched.load(std::memory_order_relaxed); }

double executor_device_ms(Executor *ex, uint64_t *generations) {
    std::lock_guard<std::mutex> g(ex->m);
    if (generations) *generations = ex->ms_count;
    const double avg = ex->ms_count ? ex->ms_sum / (double)ex->ms_count : 0.0;
    ex->ms_sum = 0.0;
    ex->ms_count = 0;
    return avg;
}

double executor_last_generation_ms(Executor *ex) { return ex->last_ms.load(std::memory_order_relaxed); }

double executor_span_ms(Executor *ex) {
    std::lock_guard<std::mutex> g(ex->m);
    const double span = ex->span_ms;
    ex->span_ms = 0.0;
    if (ex->span_anchor) ex->free_events.push_back(ex->span_anchor);
    ex->span_anchor = nullptr;
    return span;
}

} // namespace fb::dev
