// device.h — the seam between the host-side scheduler (worlds.cpp) and the CUDA side
// (device.cu, kernels.cu).  Plain C++: worlds.cpp never sees a CUDA header.
#pragma once

#include <cstddef>
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "fimo_worlds_b200.h"

namespace fb::dev {

// ---- HBM resource store -----------------------------------------------------------------------
// One allocation per (resource, batch of worlds): dense 4-byte words [n_worlds][components][count].
// All calls throw nothing and return false / nullptr after recording the CUDA error text.
bool device_ok(int device); // a usable CUDA device with that ordinal exists (-1 = current)
int resolve_device(int device);
void *hbm_alloc(int device, size_t bytes); // zero-filled
void hbm_free(int device, void *ptr);
bool hbm_upload(int device, void *dst, const void *src, size_t bytes);
bool hbm_download(int device, void *dst, const void *src, size_t bytes);
const char *last_cuda_error();

// ---- executor ------------------------------------------------------------------------------------
// A compiled system group: the DAG of device nodes, instantiated as a CUDA graph (or replayed with
// stream fork/join when graphs are disabled), launched once per generation on the group's stream.
struct Node {
    // Enqueue the node's kernels on info->stream (FimoB200_device_entry with bound arguments).
    std::function<FimoStatus(const FimoB200_LaunchInfo *)> launch;
    std::vector<uint32_t> deps; // indices of nodes that must finish first
    uint32_t kernels = 1;       // kernel launches this node issues (for ExecutorInfo / gpu_launches)
    // Optional: enqueue `gens` consecutive generations of this node in ONE launch (the fused world-step kernel when a
    // world belongs to one CTA).  Used when the node is the whole executor and the backlog holds several generations
    // with nothing between them (executor_launch_batch).
    std::function<FimoStatus(const FimoB200_LaunchInfo *, uint32_t gens)> launch_gens;
    uint32_t max_gens = 1;
    uint32_t kernels_gens = 1; // kernel launches of one launch_gens call
    // Optional: one generation that ALSO writes the cull outputs (mask words, ordered list, count: the layout of the
    // resources) to a shadow set — read-back generations, so that the pack of generation g can run beside generation g + 1.
    std::function<FimoStatus(const FimoB200_LaunchInfo *, const struct CullShadow *)> launch_shadow;
};
struct CullShadow {
    uint32_t *mask, *list, *count; // HBM: [n_worlds][mask_words], [n_worlds][entities], [n_worlds]
};

struct Executor;
Executor *executor_create(int device);
void executor_destroy(Executor *ex); // waits until idle
bool executor_compile(Executor *ex, std::vector<Node> nodes, const FimoB200_LaunchInfo &info, bool use_graph);
// Enqueue one generation.  `done(user)` runs on the completion thread once the device finished it
// (never inside a CUDA callback, so it may take locks and wake futex waiters).
// `uploads` are copied host -> HBM in front of the generation and `downloads` HBM -> host behind it, all in stream
// order; with downloads, `done` runs once they have landed.
struct Copy {
    void *dst;
    const void *src;
    size_t bytes;
};
// `readback` (fimo_b200_group_schedule_readback): behind the generation, one small pack of the cull outputs of every
// world — counts, exclusive offsets, mask words and the first `count` entries of each ordered visible list, back to
// back — into an HBM staging slot in stream order, then ONE copy-engine transfer of `host_bytes` to `host` on the
// executor's read-back stream, beside the next generation; `done` runs once it has landed.
struct Readback {
    const uint32_t *mask, *list, *count; // HBM: [n_worlds][mask_words], [n_worlds][entities], [n_worlds]
    uint64_t n_worlds, entities, mask_words, generation;
    void *host;
    size_t host_bytes;
};
bool executor_launch(Executor *ex, void (*done)(void *), void *user, const Copy *uploads = nullptr, size_t n_up = 0,
                     const Copy *downloads = nullptr, size_t n_down = 0, const Readback *readback = nullptr);
// `gens` (<= executor_max_gens) consecutive generations in one launch; `done(users[i])` runs once per generation, in
// order, when the launch has finished (generation numbers, fences and semaphores stay per generation).
uint32_t executor_max_gens(Executor *ex);
bool executor_launch_batch(Executor *ex, uint32_t gens, void (*done)(void *), void *const *users);
// Enqueue a stream-ordered copy between host memory and HBM behind everything launched so far.
bool executor_transfer(Executor *ex, void *dst, const void *src, size_t bytes, bool to_device, void (*done)(void *), void *user);
// A host system inside a device group: `d2h` copies (HBM -> pinned mirrors) in front, `fn(user)` as a host node
// (cudaLaunchHostFunc — also under stream capture, where it becomes a host node of the CUDA graph), `h2d` copies
// (mirrors -> HBM) behind, all in stream order on `stream`.  `fn` must not call the CUDA API.
bool enqueue_host_node(void *stream, const Copy *d2h, size_t n_d2h, void (*fn)(void *), void *user, const Copy *h2d, size_t n_h2d);
void *host_alloc_pinned(size_t bytes);
void host_free_pinned(void *p);
void executor_wait_idle(Executor *ex);
void *executor_stream(Executor *ex);
bool executor_uses_graph(Executor *ex);
uint32_t executor_kernels_per_generation(Executor *ex);
uint64_t executor_kernels_launched(Executor *ex); // kernel launches of all generations enqueued so far
double executor_device_ms(Executor *ex, uint64_t *generations);
double executor_last_generation_ms(Executor *ex); // device time of the generation whose `done` is running (completion thread)
double executor_span_ms(Executor *ex);

// ---- multi-GPU statistics exchange (stats_exchange.cu) -------------------------------------------
// See include/fimo_worlds_b200.h "Multi-GPU".  `before` / `after_launch` are called by the launcher thread around
// the enqueue of `n` generations on the group's stream (NCCL mode: ring-block reuse and the bucketed all-reduce).
struct StatsExchange;
StatsExchange *stats_exchange_create(int device);
void stats_exchange_destroy(StatsExchange *x);
bool stats_exchange_peers_open(StatsExchange *x, uint32_t slots, void *handle_out);
bool stats_exchange_peers_bind(StatsExchange *x, const void *handles, uint32_t rank, uint32_t nranks);
const void *stats_exchange_peer_table(StatsExchange *x); // HBM table the statistics kernel reads, or nullptr
bool stats_nccl_unique_id(void *out);
bool stats_exchange_nccl_bind(StatsExchange *x, const void *unique_id, uint32_t rank, uint32_t nranks, uint32_t bucket);
void *stats_exchange_ring(StatsExchange *x, uint32_t *slots); // NCCL mode: the ring the kernel snapshots into
// Generations ONE launch may step with this exchange bound: a launch writes its ring rows before any of its bucket
// all-reduces can be enqueued, so in NCCL mode it must not reach around the ring into rows that are still waiting for
// theirs — (blocks - 1) x bucket generations.  Unbounded (UINT32_MAX) otherwise.
uint32_t stats_exchange_max_gens(StatsExchange *x);
bool stats_exchange_before_launch(StatsExchange *x, void *stream, uint32_t n);
bool stats_exchange_after_launch(StatsExchange *x, void *stream, uint32_t n);
bool stats_exchange_read(StatsExchange *x, uint64_t generation, uint32_t mode, uint64_t *out);

// ---- built-in systems (kernels.cu) -------------------------------------------------------------
// device_entry of FIMO_B200_BUILTIN_{INTEGRATE,PROPAGATE,CULL}.
FimoStatus builtin_integrate(void *system, const FimoB200_ResourceView *const *exclusive,
                             const FimoB200_ResourceView *const *shared, const FimoB200_LaunchInfo *info);
FimoStatus builtin_propagate(void *system, const FimoB200_ResourceView *const *exclusive,
                             const FimoB200_ResourceView *const *shared, const FimoB200_LaunchInfo *info);
FimoStatus builtin_cull(void *system, const FimoB200_ResourceView *const *exclusive,
                        const FimoB200_ResourceView *const *shared, const FimoB200_LaunchInfo *info);
#define FB_DECLARE_ENTRY(fn)                                                                                       \
    FimoStatus fn(void *system, const FimoB200_ResourceView *const *exclusive, const FimoB200_ResourceView *const *shared, \
                  const FimoB200_LaunchInfo *info)
// systems_extra.cu: the systems of the 8- and 16-system DAGs
FB_DECLARE_ENTRY(builtin_forces);
FB_DECLARE_ENTRY(builtin_vel_clamp);
FB_DECLARE_ENTRY(builtin_angvel_damp);
FB_DECLARE_ENTRY(builtin_integrate_ang);
FB_DECLARE_ENTRY(builtin_world_aabb);
FB_DECLARE_ENTRY(builtin_prev_copy);
FB_DECLARE_ENTRY(builtin_lod);
FB_DECLARE_ENTRY(builtin_sleep_flags);
FB_DECLARE_ENTRY(builtin_cull_waabb);
FB_DECLARE_ENTRY(builtin_bounds_reduce);
FB_DECLARE_ENTRY(builtin_vis_bounds);
FB_DECLARE_ENTRY(builtin_group_stats);
#undef FB_DECLARE_ENTRY
uint32_t builtin_propagate_kernels(const FimoB200_ResourceView *world, const FimoB200_ResourceView *level_start);

constexpr size_t kBuiltinScratchBytes = 2048; // >= kStatScratchBytes (kernels_common.cuh)
// State of a built-in system instance (what `system_create` builds into the system blob).
struct BuiltinState {
    void *scratch; // HBM: statistics accumulators + completion ticket (cull only)
    int device;
    void *stats_ring = nullptr; // FimoB200_GroupDeviceExt::stats_ring, if this is the group's first cull system
    uint32_t stats_ring_slots = 0;
    const void *stats_peers = nullptr; // PeerTable in HBM (multi-GPU statistics exchange), same condition
};

// The fused world-step: integrate -> propagate -> cull of one batch in ONE kernel; every live-in
// resource is read once and every live-out resource written once (SURVEY.md §8d, B3 bytes).
struct WorldStepViews {
    const FimoB200_ResourceView *pos, *vel, *dt;                            // integrate
    const FimoB200_ResourceView *world, *rot, *scale, *parent, *level_start; // propagate
    const FimoB200_ResourceView *vis_mask, *vis_list, *vis_count, *stats;    // cull (exclusive)
    const FimoB200_ResourceView *aabb_c, *aabb_h, *frustum;                  // cull (shared)
    // set 1 — the fused 8-system DAG (forces, [vel_clamp], integrate_lin, integrate_ang, propagate, world_aabb,
    // cull on waabb, bounds_reduce, group_stats) — additionally uses:
    uint32_t set = 0;        // 0: 3-system chain; 1: 8-system DAG; 2: 15 systems of the 16-system DAG
    bool vel_clamp = false;  // vel_clamp of the 16-system DAG sits between forces and integrate_lin
    const FimoB200_ResourceView *gravity = nullptr, *angvel = nullptr, *vmax = nullptr, *waabb = nullptr, *wbounds = nullptr,
                                *gstats = nullptr;
    // set 2 — set 1 with vel_clamp, plus angvel_damp, sleep_flags, lod, cull2, vis_bounds, group_stats2 of the 16-system DAG:
    const FimoB200_ResourceView *damp = nullptr, *sleep_eps = nullptr, *campos = nullptr, *lod_dist2 = nullptr, *frustum2 = nullptr,
                                *sleep = nullptr, *lod = nullptr, *vis_mask2 = nullptr, *vis_list2 = nullptr, *vis_count2 = nullptr,
                                *stats2 = nullptr, *vbounds = nullptr, *gstats2 = nullptr;
    const FimoB200_ResourceView *prev_world = nullptr; // set 2, optional: prev_copy (X: prev_world, S: world) is a member too
    void *scratch2 = nullptr; // cull2's statistics accumulator block (its BuiltinState)
    void *stats_ring = nullptr; // FimoB200_GroupDeviceExt::stats_ring (per-generation snapshots of `stats`)
    uint32_t stats_ring_slots = 0;
    const void *stats_peers = nullptr; // PeerTable in HBM (fimo_b200_group_stats_peers_bind)
};
// How the fused kernel is launched for a given batch; made once when the executor is compiled (it reads
// level_start from HBM to size the shared-memory parent window), then reused by every launch.
struct WorldStepPlan {
    int threads = 0;            // entities per chunk = compute threads per CTA (128/256/512); 0 = v1 kernel
    uint32_t blocks_per_sm = 0; // resident CTAs per SM the plan is sized for
    uint32_t stages = 0;        // ring depth in chunks
    uint32_t lvl_cap = 0;       // parent matrices kept in shared memory: entities [0, lvl_cap)
    uint32_t max_parents = 0;   // max over worlds of level_start[levels-1] when the plan was made
    uint32_t team = 1;          // CTAs that step one world together (> 1: large worlds / few worlds)
    uint32_t teams = 0;         // worlds in flight = grid / team
    uint32_t kernels = 1;       // launches per generation (team mode adds the list-compaction kernel, a tail grid one launch)
    uint32_t max_gens = 1;      // generations ONE launch may step (> 1: one-CTA-per-world plans with enough chunks per CTA)
    uint32_t block_worlds = 0x40000000u; // worlds a CTA steps through all generations of a launch before it moves on (PipeArgs)
    // One CTA per world: the persistent grid steps worlds [0, tail_first) in whole rounds, a second grid launched
    // beside it steps [tail_first, n_worlds) one world per CTA as slots free up (tail_first == n_worlds: no tail).
    uint32_t grid = 0, tail_first = 0;
    uint32_t grid_all = 0;      // the persistent grid over ALL worlds (multi-generation launches do without the tail grid)
    std::shared_ptr<void> tail;     // side stream + fork / join events of the tail grid
    size_t smem = 0;            // dynamic shared memory per CTA
    std::shared_ptr<void> rows;     // HBM: [n_worlds][E] 48-byte parent rows (parents outside the window)
    std::shared_ptr<void> lvl_done; // HBM: [n_worlds][32] level-completion counters (team mode)
    std::shared_ptr<void> bounds;   // HBM: [6 + 6 n_worlds] bounds accumulators as float keys (set 1)
    std::shared_ptr<void> bounds2;  // HBM: [6] batch bounds of the visible entities (set 2)
    uint32_t set = 0;               // copied from the views
};
bool world_step_plan(const WorldStepViews &v, int device, WorldStepPlan &plan); // false: not fusable
FimoStatus world_step_launch(const WorldStepViews &v, const WorldStepPlan &plan, void *cull_scratch,
                             const FimoB200_LaunchInfo *info, uint32_t gens = 1, const CullShadow *shadow = nullptr);
bool world_step_supports_shadow(const WorldStepPlan &plan); // one-CTA-per-world pipelined plans
const char *world_step_kernel_name(const WorldStepPlan &plan); // which kernel world_step_launch picks
std::string world_step_describe(const WorldStepPlan &plan, uint64_t n_worlds); // ... and with which geometry
// v1: one CTA per world, whole world's matrices in shared memory, loads through registers (kernels.cu)
bool world_step_supported_v1(const WorldStepViews &v);
FimoStatus world_step_launch_v1(const WorldStepViews &v, void *cull_scratch, const FimoB200_LaunchInfo *info);

} // namespace fb::dev
