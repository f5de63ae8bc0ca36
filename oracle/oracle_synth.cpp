// oracle_synth.cpp — the synthetic systems as ordinary CPU `system_run` callbacks, registered
// through the unchanged descriptors exactly as a user module would on the reference's CPU path
// (systems.h:51-114), plus the timing loop of the CPU baseline.  TEST INFRASTRUCTURE ONLY.
#include <chrono>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "oracle_internal.h"
#include "systems_ref.h"

namespace {

using View = FimoB200_ResourceView;

template <typename T> T *world_ptr(const View *v, uint64_t w) {
    return (T *)((char *)v->host_ptr + w * v->world_stride);
}

bool create_noop(const void *, FimoWorldsMeta_SystemContext *, void *) { return true; }

// X: pos   S: vel, dt
void run_integrate(void *, void *const *exclusive, void *const *shared, FimoWorldsMeta_Fence *fence) {
    const View *pos = (const View *)exclusive[0];
    const View *vel = (const View *)shared[0];
    const View *dt = (const View *)shared[1];
    for (uint64_t w = 0; w < pos->n_worlds; w++)
        fimo_ref_integrate(world_ptr<float>(pos, w), world_ptr<const float>(vel, w), *world_ptr<const float>(dt, w),
                           pos->count);
    oracle::fence_signal(fence); // the system must signal its fence (systems.h:103-113)
}

// X: world   S: pos, rot, scale, parent, level_start
void run_propagate(void *, void *const *exclusive, void *const *shared, FimoWorldsMeta_Fence *fence) {
    const View *world = (const View *)exclusive[0];
    const View *pos = (const View *)shared[0];
    const View *rot = (const View *)shared[1];
    const View *scale = (const View *)shared[2];
    const View *parent = (const View *)shared[3];
    for (uint64_t w = 0; w < world->n_worlds; w++)
        if (fimo_ref_propagate(world_ptr<float>(world, w), world_ptr<const float>(pos, w), world_ptr<const float>(rot, w),
                               world_ptr<const float>(scale, w), world_ptr<const uint32_t>(parent, w), world->count) != 0)
            oracle::panic("hierarchy is not sorted by level");
    oracle::fence_signal(fence);
}

// X: vis_mask, vis_list, vis_count, stats   S: world, aabb_c, aabb_h, frustum
void run_cull(void *, void *const *exclusive, void *const *shared, FimoWorldsMeta_Fence *fence) {
    const View *mask = (const View *)exclusive[0];
    const View *list = (const View *)exclusive[1];
    const View *count = (const View *)exclusive[2];
    const View *stats = (const View *)exclusive[3];
    const View *world = (const View *)shared[0];
    const View *aabb_c = (const View *)shared[1];
    const View *aabb_h = (const View *)shared[2];
    const View *frustum = (const View *)shared[3];
    uint64_t *total = (uint64_t *)stats->host_ptr;
    uint64_t sum[FIMO_B200_STAT_WORDS] = {0};
    for (uint64_t w = 0; w < world->n_worlds; w++) {
        uint64_t one[FIMO_B200_STAT_WORDS] = {0};
        fimo_ref_cull(world_ptr<uint32_t>(mask, w), world_ptr<uint32_t>(list, w), world_ptr<uint32_t>(count, w), one,
                      world_ptr<const float>(world, w), world_ptr<const float>(aabb_c, w),
                      world_ptr<const float>(aabb_h, w), world_ptr<const float>(frustum, w), world->count);
        for (int i = 0; i < FIMO_B200_STAT_WORDS; i++) sum[i] += one[i];
    }
    total[FIMO_B200_STAT_VISIBLE] = sum[FIMO_B200_STAT_VISIBLE];
    total[FIMO_B200_STAT_ENTITIES] = sum[FIMO_B200_STAT_ENTITIES];
    total[FIMO_B200_STAT_MASK_SUM] = sum[FIMO_B200_STAT_MASK_SUM];
    total[FIMO_B200_STAT_WORLDS] = sum[FIMO_B200_STAT_WORLDS];
    total[FIMO_B200_STAT_STEPS] += sum[FIMO_B200_STAT_STEPS];
    oracle::fence_signal(fence);
}

// ---- the 8- and 16-system DAGs -----------------------------------------------------------------------------
#define FOR_WORLDS(v) for (uint64_t w = 0; w < (v)->n_worlds; w++)

// X: vel   S: gravity, dt
void run_forces(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *vel = (const View *)x[0], *g = (const View *)s[0], *dt = (const View *)s[1];
    FOR_WORLDS(vel) fimo_ref_forces(world_ptr<float>(vel, w), world_ptr<const float>(g, w), *world_ptr<const float>(dt, w), vel->count);
    oracle::fence_signal(fence);
}
// X: vel   S: vmax
void run_vel_clamp(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *vel = (const View *)x[0], *vmax = (const View *)s[0];
    FOR_WORLDS(vel) fimo_ref_vel_clamp(world_ptr<float>(vel, w), *world_ptr<const float>(vmax, w), vel->count);
    oracle::fence_signal(fence);
}
// X: angvel   S: damp
void run_angvel_damp(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *av = (const View *)x[0], *damp = (const View *)s[0];
    FOR_WORLDS(av) fimo_ref_angvel_damp(world_ptr<float>(av, w), *world_ptr<const float>(damp, w), av->count);
    oracle::fence_signal(fence);
}
// X: rot   S: angvel, dt
void run_integrate_ang(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *rot = (const View *)x[0], *av = (const View *)s[0], *dt = (const View *)s[1];
    FOR_WORLDS(rot) fimo_ref_integrate_ang(world_ptr<float>(rot, w), world_ptr<const float>(av, w), *world_ptr<const float>(dt, w), rot->count);
    oracle::fence_signal(fence);
}
// X: waabb   S: world, aabb_c, aabb_h
void run_world_aabb(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *wa = (const View *)x[0], *world = (const View *)s[0], *c = (const View *)s[1], *h = (const View *)s[2];
    FOR_WORLDS(wa) fimo_ref_world_aabb(world_ptr<float>(wa, w), world_ptr<const float>(world, w), world_ptr<const float>(c, w),
                                       world_ptr<const float>(h, w), wa->count);
    oracle::fence_signal(fence);
}
// X: vis_mask, vis_list, vis_count, stats   S: waabb, frustum
void run_cull_waabb(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *mask = (const View *)x[0], *list = (const View *)x[1], *count = (const View *)x[2], *stats = (const View *)x[3];
    const View *wa = (const View *)s[0], *frustum = (const View *)s[1];
    uint64_t *total = (uint64_t *)stats->host_ptr;
    uint64_t sum[FIMO_B200_STAT_WORDS] = {0};
    FOR_WORLDS(wa) {
        uint64_t one[FIMO_B200_STAT_WORDS] = {0};
        fimo_ref_cull_waabb(world_ptr<uint32_t>(mask, w), world_ptr<uint32_t>(list, w), world_ptr<uint32_t>(count, w), one,
                            world_ptr<const float>(wa, w), world_ptr<const float>(frustum, w), wa->count);
        for (int i = 0; i < FIMO_B200_STAT_WORDS; i++) sum[i] += one[i];
    }
    total[FIMO_B200_STAT_VISIBLE] = sum[FIMO_B200_STAT_VISIBLE];
    total[FIMO_B200_STAT_ENTITIES] = sum[FIMO_B200_STAT_ENTITIES];
    total[FIMO_B200_STAT_MASK_SUM] = sum[FIMO_B200_STAT_MASK_SUM];
    total[FIMO_B200_STAT_WORLDS] = sum[FIMO_B200_STAT_WORLDS];
    total[FIMO_B200_STAT_STEPS] += sum[FIMO_B200_STAT_STEPS];
    oracle::fence_signal(fence);
}
// X: wbounds   S: waabb
void run_bounds_reduce(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *b = (const View *)x[0], *wa = (const View *)s[0];
    FOR_WORLDS(wa) fimo_ref_bounds(world_ptr<float>(b, w), world_ptr<const float>(wa, w), nullptr, wa->count);
    oracle::fence_signal(fence);
}
// X: vbounds   S: waabb, vis_mask
void run_vis_bounds(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *b = (const View *)x[0], *wa = (const View *)s[0], *mask = (const View *)s[1];
    FOR_WORLDS(wa) fimo_ref_bounds(world_ptr<float>(b, w), world_ptr<const float>(wa, w), world_ptr<const uint32_t>(mask, w), wa->count);
    oracle::fence_signal(fence);
}
// X: gstats   S: vis_count, wbounds
void run_group_stats(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *gs = (const View *)x[0], *count = (const View *)s[0], *b = (const View *)s[1];
    uint32_t *out = (uint32_t *)gs->host_ptr;
    uint64_t visible = 0;
    float lo[3] = {FIMO_B200_BOUNDS_EMPTY_MIN, FIMO_B200_BOUNDS_EMPTY_MIN, FIMO_B200_BOUNDS_EMPTY_MIN};
    float hi[3] = {-FIMO_B200_BOUNDS_EMPTY_MIN, -FIMO_B200_BOUNDS_EMPTY_MIN, -FIMO_B200_BOUNDS_EMPTY_MIN};
    FOR_WORLDS(count) {
        visible += *world_ptr<const uint32_t>(count, w);
        const float *wb = world_ptr<const float>(b, w);
        for (int r = 0; r < 3; r++) lo[r] = std::fmin(lo[r], wb[r]), hi[r] = std::fmax(hi[r], wb[3 + r]);
    }
    for (int i = 0; i < FIMO_B200_GSTAT_WORDS; i++) out[i] = 0;
    const uint64_t worlds = count->n_worlds;
    std::memcpy(out + FIMO_B200_GSTAT_VISIBLE_LO, &visible, 8);
    std::memcpy(out + FIMO_B200_GSTAT_WORLDS_LO, &worlds, 8);
    std::memcpy(out + FIMO_B200_GSTAT_BOUNDS, lo, 12);
    std::memcpy(out + FIMO_B200_GSTAT_BOUNDS + 3, hi, 12);
    oracle::fence_signal(fence);
}
// X: prev_world   S: world
void run_prev_copy(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *prev = (const View *)x[0], *world = (const View *)s[0];
    std::memcpy(prev->host_ptr, world->host_ptr, world->world_stride * world->n_worlds);
    oracle::fence_signal(fence);
}
// X: lod   S: waabb, campos, lod_dist2
void run_lod(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *lod = (const View *)x[0], *wa = (const View *)s[0], *cam = (const View *)s[1], *d2 = (const View *)s[2];
    FOR_WORLDS(wa) fimo_ref_lod(world_ptr<uint32_t>(lod, w), world_ptr<const float>(wa, w), world_ptr<const float>(cam, w),
                                world_ptr<const float>(d2, w), wa->count);
    oracle::fence_signal(fence);
}
// X: sleep   S: vel, sleep_eps
void run_sleep_flags(void *, void *const *x, void *const *s, FimoWorldsMeta_Fence *fence) {
    const View *sleep = (const View *)x[0], *vel = (const View *)s[0], *eps = (const View *)s[1];
    FOR_WORLDS(vel) fimo_ref_sleep_flags(world_ptr<uint32_t>(sleep, w), world_ptr<const float>(vel, w), *world_ptr<const float>(eps, w), vel->count);
    oracle::fence_signal(fence);
}

} // namespace

extern "C" {

// CPU counterpart of fimo_b200_builtin_system: fills create/run for a built-in system id.
FIMO_B200_EXPORT FimoStatus fimo_oracle_builtin_system(uint32_t builtin_id, FimoWorldsMeta_SystemDescriptor *desc) {
    desc->system_size = 0;
    desc->system_alignment = 1;
    desc->system_create = &create_noop;
    desc->system_destroy = nullptr;
    switch (builtin_id) {
        case FIMO_B200_BUILTIN_INTEGRATE: desc->system_run = &run_integrate; break;
        case FIMO_B200_BUILTIN_PROPAGATE: desc->system_run = &run_propagate; break;
        case FIMO_B200_BUILTIN_CULL: desc->system_run = &run_cull; break;
        case FIMO_B200_BUILTIN_FORCES: desc->system_run = &run_forces; break;
        case FIMO_B200_BUILTIN_INTEGRATE_ANG: desc->system_run = &run_integrate_ang; break;
        case FIMO_B200_BUILTIN_WORLD_AABB: desc->system_run = &run_world_aabb; break;
        case FIMO_B200_BUILTIN_CULL_WAABB: desc->system_run = &run_cull_waabb; break;
        case FIMO_B200_BUILTIN_BOUNDS_REDUCE: desc->system_run = &run_bounds_reduce; break;
        case FIMO_B200_BUILTIN_GROUP_STATS: desc->system_run = &run_group_stats; break;
        case FIMO_B200_BUILTIN_PREV_COPY: desc->system_run = &run_prev_copy; break;
        case FIMO_B200_BUILTIN_LOD: desc->system_run = &run_lod; break;
        case FIMO_B200_BUILTIN_VEL_CLAMP: desc->system_run = &run_vel_clamp; break;
        case FIMO_B200_BUILTIN_SLEEP_FLAGS: desc->system_run = &run_sleep_flags; break;
        case FIMO_B200_BUILTIN_ANGVEL_DAMP: desc->system_run = &run_angvel_damp; break;
        case FIMO_B200_BUILTIN_VIS_BOUNDS: desc->system_run = &run_vis_bounds; break;
        default: oracle::set_last_error("NotFound"); return FIMO_STATUS_ERROR;
    }
    return FIMO_STATUS_OK;
}

// CPU baseline timing loop: `steps` times { schedule every group; wait every fence }.
// Returns wall seconds (steady_clock) for all steps.
FIMO_B200_EXPORT double fimo_oracle_step_groups(FimoWorldsMeta_SystemGroup *const *groups, size_t n, size_t steps) {
    std::vector<FimoWorldsMeta_Fence> fences(n);
    const auto t0 = std::chrono::steady_clock::now();
    for (size_t s = 0; s < steps; s++) {
        for (size_t i = 0; i < n; i++) {
            FimoWorldsMeta_fence_init(&fences[i]);
            if (fimo_worlds_system_group_schedule(groups[i], nullptr, 0, &fences[i]) != FIMO_STATUS_OK)
                oracle::panic("schedule failed");
        }
        for (size_t i = 0; i < n; i++) oracle::fence_wait(&fences[i]);
    }
    const auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// The same work the way the reference lets callers drive it: `schedule` is asynchronous and generations of one
// group queue behind its timeline semaphore (SystemGroup.zig:430-452), so every group gets all `steps` generations
// up front and only the last one of each carries a fence — no barrier across worlds between steps.
FIMO_B200_EXPORT double fimo_oracle_step_groups_queued(FimoWorldsMeta_SystemGroup *const *groups, size_t n, size_t steps) {
    std::vector<FimoWorldsMeta_Fence> fences(n);
    const auto t0 = std::chrono::steady_clock::now();
    for (size_t i = 0; i < n; i++) FimoWorldsMeta_fence_init(&fences[i]);
    for (size_t s = 0; s < steps; s++)
        for (size_t i = 0; i < n; i++)
            if (fimo_worlds_system_group_schedule(groups[i], nullptr, 0, s + 1 == steps ? &fences[i] : nullptr) != FIMO_STATUS_OK)
                oracle::panic("schedule failed");
    for (size_t i = 0; i < n; i++) oracle::fence_wait(&fences[i]);
    const auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// Raw single-thread math loop without any scheduling: the floor of what one host core can do,
// reported next to the scheduled number so the dispatch overhead is visible.
FIMO_B200_EXPORT unsigned fimo_oracle_hardware_threads(void) { return std::thread::hardware_concurrency(); }

} // extern "C"
