// kernels.cu — sm_100a kernels of the built-in synthetic systems and of the fused
// world-step.  HBM-bound integer/float streaming work: no tensor cores anywhere on this path.
//
// The reference contains no entity / transform / cull arithmetic (a "system" there is an opaque
// callback, pkgs/fimo_worlds_meta/include/fimo_worlds_meta/systems.h:103-113), so the arithmetic
// is defined by this build (kernels_common.cuh): every float operation is one IEEE-754 binary32
// operation — __fmul_rn / __fadd_rn / __fsub_rn, or __fmaf_rn exactly where the definition spells a
// fused multiply-add — in a fixed order, never contracted or reassociated by the compiler, so that
// results are bit-identical to a scalar CPU evaluation of the same expressions with fmaf().
//
// Data layout: every resource is dense 4-byte words [n_worlds][components][count];
// component plane k of world w starts at word  w*world_words + k*count.



#include "kernels_common.cuh"

namespace fb::dev {
namespace {

// Cull of one world by one CTA.  `world_at(i, e)` reads component i of entity e's world transform.
// Writes the bit mask (warp ballot), the ASCENDING visible-index list (warp popc prefix + block
// scan — never unordered atomics) and the count.  Returns (count, mask_sum) on thread 0.
template <int T, typename WorldFn>
__device__ __forceinline__ void cull_world(const StepArgs &a, uint64_t w, const float *s_frustum, uint32_t *s_warp,
                                           WorldFn world_at, unsigned long long &visible_out,
                                           unsigned long long &mask_sum_out) {
    const uint32_t E = a.entities;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = T / 32;
    const float *__restrict__ ac = fptr(a.aabb_c, w);
    const float *__restrict__ ah = fptr(a.aabb_h, w);
    uint32_t *__restrict__ mask = uptr(a.vis_mask, w);
    uint32_t *__restrict__ list = uptr(a.vis_list, w);
    uint32_t running = 0;
    unsigned long long mask_sum = 0; // lane 0 of every warp accumulates its own words
    for (uint32_t base = 0; base < E; base += T) {
        const uint32_t e = base + tid;
        bool vis = false;
        if (e < E) {
            float W[12];
#pragma unroll
            for (int i = 0; i < 12; i++) W[i] = world_at(i, e);
            vis = aabb_visible(W, ac[e], ac[E + e], ac[2 * E + e], ah[e], ah[E + e], ah[2 * E + e], s_frustum);
        }
        const unsigned ballot = __ballot_sync(kFull, vis);
        if (lane == 0) {
            if ((e >> 5) < a.mask_words) mask[e >> 5] = ballot;
            mask_sum += ballot;
            s_warp[warp] = __popc(ballot);
        }
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int i = 0; i < kWarps; i++) {
            const uint32_t c = s_warp[i];
            before += (i < warp) ? c : 0u;
            total += c;
        }
        if (vis) list[running + before + __popc(ballot & ((1u << lane) - 1u))] = e;
        running += total;
        __syncthreads();
    }
    // block sum of the per-warp mask sums
    if (lane == 0) {
        s_warp[warp] = (uint32_t)(mask_sum & 0xFFFFFFFFull);
        s_warp[kWarps + warp] = (uint32_t)(mask_sum >> 32);
    }
    __syncthreads();
    if (tid == 0) {
        unsigned long long total = 0;
        for (int i = 0; i < kWarps; i++) total += ((unsigned long long)s_warp[kWarps + i] << 32) | s_warp[i];
        *uptr(a.vis_count, w) = running;
        visible_out = running;
        mask_sum_out = total;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// per-system kernels (the CUDA-graph-of-kernels executor)
// ---------------------------------------------------------------------------------------------

// integrate: pos += vel * dt, 128-bit loads/stores. grid.x = chunks_per_world * n_worlds.
__global__ void __launch_bounds__(256) k_integrate_v4(Plane pos, Plane vel, Plane dt, uint32_t vec_per_world,
                                                      uint32_t chunks_per_world) {
    const uint64_t w = blockIdx.x / chunks_per_world;
    const uint32_t i = (blockIdx.x % chunks_per_world) * 256 + threadIdx.x;
    if (i >= vec_per_world) return;
    const float step = *fptr(dt, w);
    float4 *p = (float4 *)fptr(pos, w) + i;
    const float4 v = __ldg((const float4 *)fptr(vel, w) + i);
    float4 x = *p;
    x.x = fma1(v.x, step, x.x);
    x.y = fma1(v.y, step, x.y);
    x.z = fma1(v.z, step, x.z);
    x.w = fma1(v.w, step, x.w);
    *p = x;
}
__global__ void __launch_bounds__(256) k_integrate_s(Plane pos, Plane vel, Plane dt, uint32_t words_per_world,
                                                     uint32_t chunks_per_world) {
    const uint64_t w = blockIdx.x / chunks_per_world;
    const uint32_t i = (blockIdx.x % chunks_per_world) * 256 + threadIdx.x;
    if (i >= words_per_world) return;
    const float step = *fptr(dt, w);
    float *p = fptr(pos, w) + i;
    *p = fma1(fptr(vel, w)[i], step, *p);
}

// propagate, one CTA per world: the world's 3x4 matrices are staged in shared memory ([12][E]
// planes, conflict-free), levels are walked in order with one __syncthreads per level.
template <int T, bool kIntegrate, bool kCull>
__global__ void __launch_bounds__(T) k_world_cta(StepArgs a) {
    extern __shared__ float s_world[]; // [12][E]
    __shared__ __align__(16) float s_frustum[24];
    __shared__ uint32_t s_levels[kMaxLevels + 1];
    __shared__ uint32_t s_warp[2 * (T / 32)];
    const uint32_t E = a.entities;
    const int tid = threadIdx.x;
    unsigned long long cta_visible = 0, cta_mask_sum = 0, cta_worlds = 0;
    for (uint64_t w = blockIdx.x; w < a.n_worlds; w += gridDim.x) {
        if (tid <= (int)a.levels) s_levels[tid] = uptr(a.level_start, w)[tid];
        if (kCull && tid < 24) s_frustum[tid] = fptr(a.frustum, w)[tid];
        __syncthreads();
        float *__restrict__ pos = fptr(a.pos, w);
        const float *__restrict__ vel = kIntegrate ? fptr(a.vel, w) : nullptr;
        const float dt = kIntegrate ? *fptr(a.dt, w) : 0.0f;
        const float *__restrict__ rot = fptr(a.rot, w);
        const float *__restrict__ scale = fptr(a.scale, w);
        const uint32_t *__restrict__ parent = uptr(a.parent, w);
        float *__restrict__ world = fptr(a.world, w);
        for (uint32_t l = 0; l < a.levels; l++) {
            const uint32_t end = s_levels[l + 1];
            for (uint32_t e = s_levels[l] + tid; e < end; e += T) {
                propagate_entity<kIntegrate>(
                        e, E, pos, vel, dt, rot, scale, parent, [&](int i, uint32_t p) { return s_world[i * E + p]; },
                        [&](int i, float v) {
                            s_world[i * E + e] = v;
                            world[(uint64_t)i * E + e] = v;
                        });
            }
            __syncthreads();
        }
        if (kCull) {
            unsigned long long visible = 0, mask_sum = 0;
            cull_world<T>(a, w, s_frustum, s_warp, [&](int i, uint32_t e) { return s_world[i * E + e]; }, visible, mask_sum);
            cta_visible += visible;
            cta_mask_sum += mask_sum;
            cta_worlds += 1;
        }
    }
    if (kCull && tid == 0) publish_stats(a, cta_visible, cta_mask_sum, cta_worlds);
}

// propagate for worlds too large for one CTA's shared memory: one launch per hierarchy level,
// parents are read back from HBM/L2 (written by the previous launch).
__global__ void __launch_bounds__(256) k_propagate_level(StepArgs a, uint32_t level, uint32_t chunks_per_world) {
    const uint64_t w = blockIdx.x / chunks_per_world;
    const uint32_t *ls = uptr(a.level_start, w);
    const uint32_t e = ls[level] + (blockIdx.x % chunks_per_world) * 256 + threadIdx.x;
    if (e >= ls[level + 1]) return;
    const uint32_t E = a.entities;
    float *world = fptr(a.world, w);
    propagate_entity<false>(
            e, E, fptr(a.pos, w), nullptr, 0.0f, fptr(a.rot, w), fptr(a.scale, w), uptr(a.parent, w),
            [&](int i, uint32_t p) { return world[(uint64_t)i * E + p]; },
            [&](int i, float v) { world[(uint64_t)i * E + e] = v; });
}

// cull, one CTA per world, transforms read from HBM.
template <int T> __global__ void __launch_bounds__(T) k_cull_cta(StepArgs a) {
    __shared__ __align__(16) float s_frustum[24];
    __shared__ uint32_t s_warp[2 * (T / 32)];
    const uint32_t E = a.entities;
    unsigned long long cta_visible = 0, cta_mask_sum = 0, cta_worlds = 0;
    for (uint64_t w = blockIdx.x; w < a.n_worlds; w += gridDim.x) {
        if (threadIdx.x < 24) s_frustum[threadIdx.x] = fptr(a.frustum, w)[threadIdx.x];
        __syncthreads();
        const float *__restrict__ world = fptr(a.world, w);
        unsigned long long visible = 0, mask_sum = 0;
        cull_world<T>(a, w, s_frustum, s_warp, [&](int i, uint32_t e) { return world[(uint64_t)i * E + e]; }, visible,
                      mask_sum);
        cta_visible += visible;
        cta_mask_sum += mask_sum;
        cta_worlds += 1;
    }
    if (threadIdx.x == 0) publish_stats(a, cta_visible, cta_mask_sum, cta_worlds);
}

} // namespace

FimoStatus builtin_integrate(void *, const FimoB200_ResourceView *const *exclusive, const FimoB200_ResourceView *const *shared,
                             const FimoB200_LaunchInfo *info) {
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    try {
        const Plane pos = plane_of(exclusive[0], FIMO_B200_SCOPE_ENTITY, 3, E, W, "pos");
        const Plane vel = plane_of(shared[0], FIMO_B200_SCOPE_ENTITY, 3, E, W, "vel");
        const Plane dt = plane_of(shared[1], FIMO_B200_SCOPE_WORLD, 1, 1, W, "dt");
        cudaStream_t stream = (cudaStream_t)info->stream;
        const uint64_t words = 3 * E;
        if (words % 4 == 0 && ((uintptr_t)pos.base % 16) == 0 && ((uintptr_t)vel.base % 16) == 0) {
            const uint32_t vec = (uint32_t)(words / 4), chunks = (vec + 255) / 256;
            k_integrate_v4<<<(unsigned)(chunks * W), 256, 0, stream>>>(pos, vel, dt, vec, chunks);
        } else {
            const uint32_t chunks = (uint32_t)((words + 255) / 256);
            k_integrate_s<<<(unsigned)(chunks * W), 256, 0, stream>>>(pos, vel, dt, (uint32_t)words, chunks);
        }
        return check_launch("integrate");
    } catch (const Fail &f) {
        return fail(f.what);
    }
}

uint32_t builtin_propagate_kernels(const FimoB200_ResourceView *world, const FimoB200_ResourceView *level_start) {
    if (!world || !level_start) return 1;
    return 48 * world->count <= kMaxDynamicSmem ? 1u : level_start->components - 1;
}

FimoStatus builtin_propagate(void *, const FimoB200_ResourceView *const *exclusive, const FimoB200_ResourceView *const *shared,
                             const FimoB200_LaunchInfo *info) {
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    try {
        StepArgs a{};
        a.n_worlds = (uint32_t)W, a.entities = (uint32_t)E;
        fill_propagate(a, exclusive[0], shared[0], shared[1], shared[2], shared[3], shared[4], W, E);
        cudaStream_t stream = (cudaStream_t)info->stream;
        const size_t smem = 48 * E;
        if (smem <= kMaxDynamicSmem) {
            auto kernel = k_world_cta<256, false, false>;
            if (!allow_smem(kernel, smem)) return check_launch("propagate (shared memory opt-in)");
            kernel<<<(unsigned)W, 256, smem, stream>>>(a);
        } else {
            const uint32_t chunks = (uint32_t)((E + 255) / 256); // upper bound of any level's size
            for (uint32_t l = 0; l < a.levels; l++)
                k_propagate_level<<<(unsigned)(chunks * W), 256, 0, stream>>>(a, l, chunks);
        }
        return check_launch("propagate");
    } catch (const Fail &f) {
        return fail(f.what);
    }
}

FimoStatus builtin_cull(void *system, const FimoB200_ResourceView *const *exclusive, const FimoB200_ResourceView *const *shared,
                        const FimoB200_LaunchInfo *info) {
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0) return FIMO_STATUS_OK; // empty worlds (E == 0) still publish their counts and the batch statistics
    try {
        StepArgs a{};
        a.n_worlds = (uint32_t)W, a.entities = (uint32_t)E;
        fill_cull(a, exclusive[0], exclusive[1], exclusive[2], exclusive[3], shared[0], shared[1], shared[2], shared[3],
                  ((BuiltinState *)system)->scratch, W, E);
        a.ring = (unsigned long long *)((BuiltinState *)system)->stats_ring, a.ring_slots = ((BuiltinState *)system)->stats_ring_slots;
        a.peers = (const PeerTable *)((BuiltinState *)system)->stats_peers;
        cudaStream_t stream = (cudaStream_t)info->stream;
        const unsigned grid = (unsigned)(W < 8ull * sm_count(info->device) ? W : 8ull * sm_count(info->device));
        if (E <= 2048) k_cull_cta<256><<<grid, 256, 0, stream>>>(a);
        else k_cull_cta<1024><<<grid, 1024, 0, stream>>>(a);
        return check_launch("cull");
    } catch (const Fail &f) {
        return fail(f.what);
    }
}

bool world_step_supported_v1(const WorldStepViews &v) {
    return v.world && 48 * v.world->count <= kMaxDynamicSmem && v.world->count > 0;
}

FimoStatus world_step_launch_v1(const WorldStepViews &v, void *cull_scratch, const FimoB200_LaunchInfo *info) {
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    try {
        StepArgs a{};
        a.n_worlds = (uint32_t)W, a.entities = (uint32_t)E;
        fill_propagate(a, v.world, v.pos, v.rot, v.scale, v.parent, v.level_start, W, E);
        fill_cull(a, v.vis_mask, v.vis_list, v.vis_count, v.stats, v.world, v.aabb_c, v.aabb_h, v.frustum, cull_scratch, W, E);
        a.ring = (unsigned long long *)v.stats_ring, a.ring_slots = v.stats_ring_slots, a.peers = (const PeerTable *)v.stats_peers;
        a.vel = plane_of(v.vel, FIMO_B200_SCOPE_ENTITY, 3, E, W, "vel");
        a.dt = plane_of(v.dt, FIMO_B200_SCOPE_WORLD, 1, 1, W, "dt");
        cudaStream_t stream = (cudaStream_t)info->stream;
        const size_t smem = 48 * E;
        auto kernel = k_world_cta<256, true, true>;
        if (!allow_smem(kernel, smem)) return check_launch("world_step (shared memory opt-in)");
        kernel<<<(unsigned)W, 256, smem, stream>>>(a);
        return check_launch("world_step");
    } catch (const Fail &f) {
        return fail(f.what);
    }
}

} // namespace fb::dev
