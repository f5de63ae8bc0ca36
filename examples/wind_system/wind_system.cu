// wind_system.cu — a THIRD-PARTY system for a fimo_worlds system group, built into its own shared library.
//
// It is registered exactly as the reference wants a system registered — one FimoWorldsMeta_SystemDescriptor with
// label, resource handles, dependencies, factory, system_create / system_destroy / system_run
// (pkgs/fimo_worlds_meta/include/fimo_worlds_meta/systems.h:51-114) — plus, chained on the descriptor's reserved
// `next` pointer (systems.h:52-53), a FimoB200_SystemDeviceExt whose `device_entry` enqueues the system's kernel for
// every world of the batch.  builtin_id is 0: the library knows nothing about this system; it orders it against the
// other systems by its dependencies and resource conflicts alone and launches it on the stream (or captures it into
// the generation's CUDA graph).  Without the extension (reference / CPU oracle) `system_run` does the same arithmetic
// on the host, one IEEE-754 binary32 rounding per operation on both sides, so the two paths agree bit for bit.
//
//   wind      X: vel (entity, 3)      S: pos (entity, 3), wind (world, 4: direction xyz, shear), dt (world, 1)
//             vel += gust * wind.xyz * (1 + shear * pos.y) * dt
//
// `gust` comes from the factory (copied at registration, Universe.zig:311-320) into the system's state.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "fimo_worlds_b200.h"

namespace {

struct WindFactory {
    float gust;
};
struct WindState {
    float gust;
    uint64_t device_launches, host_runs; // how often each path ran (the tests read them through wind_counters)
};
WindState *g_last_state = nullptr;
FimoTasksMeta_futex_wake g_futex_wake = nullptr;

bool wind_create(const void *factory, FimoWorldsMeta_SystemContext *, void *system) {
    WindState *st = (WindState *)system;
    st->gust = ((const WindFactory *)factory)->gust;
    st->device_launches = st->host_runs = 0;
    g_last_state = st;
    return true;
}
void wind_destroy(void *system) {
    if (g_last_state == system) g_last_state = nullptr;
}

// one entity: every line is one rounding
__host__ __device__ inline void wind_entity(float &vx, float &vy, float &vz, float py, const float *wind, float gust, float dt) {
#ifdef __CUDA_ARCH__
    const float k = __fmul_rn(__fmul_rn(gust, __fmaf_rn(wind[3], py, 1.0f)), dt);
    vx = __fmaf_rn(wind[0], k, vx), vy = __fmaf_rn(wind[1], k, vy), vz = __fmaf_rn(wind[2], k, vz);
#else
    const float k = (gust * fmaf(wind[3], py, 1.0f)) * dt; // -ffp-contract=off: two separate multiplies
    vx = fmaf(wind[0], k, vx), vy = fmaf(wind[1], k, vy), vz = fmaf(wind[2], k, vz);
#endif
}

// CPU path: the views' host_ptr IS the data (reference / oracle), [n_worlds][components][count] words
void wind_run(void *system, void *const *exclusive, void *const *shared, FimoWorldsMeta_Fence *fence) {
    WindState *st = (WindState *)system;
    const FimoB200_ResourceView *vel = (const FimoB200_ResourceView *)exclusive[0];
    const FimoB200_ResourceView *pos = (const FimoB200_ResourceView *)shared[0], *wind = (const FimoB200_ResourceView *)shared[1],
                                *dt = (const FimoB200_ResourceView *)shared[2];
    const uint64_t E = vel->count;
    for (uint64_t w = 0; w < vel->n_worlds; w++) {
        float *v = (float *)((char *)vel->host_ptr + w * vel->world_stride);
        const float *p = (const float *)((const char *)pos->host_ptr + w * pos->world_stride);
        const float *wd = (const float *)((const char *)wind->host_ptr + w * wind->world_stride);
        const float step = *(const float *)((const char *)dt->host_ptr + w * dt->world_stride);
        for (uint64_t e = 0; e < E; e++) wind_entity(v[e], v[E + e], v[2 * E + e], p[E + e], wd, st->gust, step);
    }
    st->host_runs++;
    // the system signals its fence (systems.h:103-113); nobody waits on it before system_run returns
    if (g_futex_wake) FimoWorldsMeta_fence_signal(fence, g_futex_wake);
    else __atomic_store_n(&fence->state, FIMO_WORLDS_META_FENCE_SIGNALED, __ATOMIC_RELEASE);
}

__global__ void __launch_bounds__(256) k_wind(float *__restrict__ vel, const float *__restrict__ pos, const float *__restrict__ wind,
                                              const float *__restrict__ dt, uint32_t E, uint64_t vel_world, uint64_t pos_world, uint64_t wind_world,
                                              uint64_t dt_world, float gust) {
    const uint64_t w = blockIdx.y;
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    float *v = vel + w * vel_world;
    float vx = v[e], vy = v[E + e], vz = v[2ull * E + e];
    wind_entity(vx, vy, vz, pos[w * pos_world + E + e], wind + w * wind_world, gust, dt[w * dt_world]);
    v[e] = vx, v[E + e] = vy, v[2ull * E + e] = vz;
}

// Device path: same argument order as system_run, views instead of blobs, the stream instead of the fence.
FimoStatus wind_device_entry(void *system, const FimoB200_ResourceView *const *exclusive, const FimoB200_ResourceView *const *shared,
                             const FimoB200_LaunchInfo *info) {
    WindState *st = (WindState *)system;
    const FimoB200_ResourceView *vel = exclusive[0], *pos = shared[0], *wind = shared[1], *dt = shared[2];
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    if (!vel->device_ptr || !pos->device_ptr || !wind->device_ptr || !dt->device_ptr || vel->components != 3 || pos->components != 3 ||
        wind->components != 4 || vel->count != E || pos->count != E || W > 65535)
        return FIMO_STATUS_ERROR;
    k_wind<<<dim3((unsigned)((E + 255) / 256), (unsigned)W), 256, 0, (cudaStream_t)info->stream>>>(
            (float *)vel->device_ptr, (const float *)pos->device_ptr, (const float *)wind->device_ptr, (const float *)dt->device_ptr, (uint32_t)E,
            vel->world_stride / 4, pos->world_stride / 4, wind->world_stride / 4, dt->world_stride / 4, st->gust);
    if (!(info->flags & FIMO_B200_LAUNCH_FLAG_CAPTURING)) st->device_launches++; // under capture the entry runs once, the graph replays it
    return cudaGetLastError() == cudaSuccess ? FIMO_STATUS_OK : FIMO_STATUS_ERROR;
}

WindFactory g_factory{1.0f};

} // namespace

extern "C" {

// Fills everything of the descriptor that belongs to the system itself; the caller adds label, resource handles and
// dependencies (they are handles of ITS universe).  `ext` != NULL: chain the device extension (B200 library);
// NULL: a plain CPU system (reference / oracle).
__attribute__((visibility("default"))) void wind_fill_descriptor(FimoWorldsMeta_SystemDescriptor *desc, FimoB200_SystemDeviceExt *ext, float gust) {
    g_factory.gust = gust;
    desc->factory = &g_factory, desc->factory_size = sizeof g_factory, desc->factory_alignment = alignof(WindFactory), desc->factory_destroy = nullptr;
    desc->system_size = sizeof(WindState), desc->system_alignment = alignof(WindState);
    desc->system_create = &wind_create, desc->system_destroy = &wind_destroy, desc->system_run = &wind_run;
    desc->next = nullptr;
    if (ext) {
        std::memset(ext, 0, sizeof *ext);
        ext->header.s_type = FIMO_B200_STYPE_SYSTEM_DEVICE;
        ext->device_entry = &wind_device_entry;
        ext->builtin_id = FIMO_B200_BUILTIN_NONE;
        desc->next = ext;
    }
}
__attribute__((visibility("default"))) void wind_bind_futex_wake(FimoTasksMeta_futex_wake wake) { g_futex_wake = wake; }
__attribute__((visibility("default"))) void wind_counters(uint64_t *device_launches, uint64_t *host_runs) {
    *device_launches = g_last_state ? g_last_state->device_launches : 0;
    *host_runs = g_last_state ? g_last_state->host_runs : 0;
}

} // extern "C"
