// oracle_synth.cpp — the synthetic systems as ordinary CPU `system_run` callbacks, registered
// through the unchanged descriptors exactly as a user module would on the reference's CPU path
// (systems.h:51-114), plus the timing loop of the CPU baseline.  TEST INFRASTRUCTURE ONLY.
#include <chrono>
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

// Raw single-thread math loop without any scheduling: the floor of what one host core can do,
// reported next to the scheduled number so the dispatch overhead is visible.
FIMO_B200_EXPORT unsigned fimo_oracle_hardware_threads(void) { return std::thread::hardware_concurrency(); }

} // extern "C"
