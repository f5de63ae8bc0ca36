// kernels.cu — sm_100a kernels of the built-in synthetic systems and of the fused
// world-step.  HBM-bound integer/float streaming work: no tensor cores anywhere on this path.
//
// The reference contains no entity / transform / cull arithmetic (a "system" there is an opaque
// callback, pkgs/fimo_worlds_meta/include/fimo_worlds_meta/systems.h:103-113), so the arithmetic
// below is defined by this build: every float operation is one IEEE-754 binary32 operation
// (__fmul_rn/__fadd_rn/__fsub_rn, never contracted into an FMA), in a fixed order, so that results
// are bit-identical to a scalar CPU evaluation of the same expressions.
//
// Data layout: every resource is dense 4-byte words [n_worlds][components][count];
// component plane k of world w starts at word  w*world_words + k*count.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

#include "device.h"

namespace fb::dev {
namespace {

constexpr uint32_t kRoot = FIMO_B200_PARENT_ROOT;
constexpr int kMaxLevels = 32;
constexpr unsigned kFull = 0xFFFFFFFFu;

// ---------------------------------------------------------------------------------------------
// math (each line is one rounding)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }

// local = T(pos) * R(quat x,y,z,w) * S(uniform s): 3x4 row-major affine
__device__ __forceinline__ void local_trs(float (&L)[12], float px, float py, float pz, float x, float y, float z, float w,
                                          float s) {
    const float xx = mul(x, x), yy = mul(y, y), zz = mul(z, z);
    const float xy = mul(x, y), xz = mul(x, z), yz = mul(y, z);
    const float wx = mul(w, x), wy = mul(w, y), wz = mul(w, z);
    const float r00 = sub(1.0f, mul(2.0f, add(yy, zz))), r01 = mul(2.0f, sub(xy, wz)), r02 = mul(2.0f, add(xz, wy));
    const float r10 = mul(2.0f, add(xy, wz)), r11 = sub(1.0f, mul(2.0f, add(xx, zz))), r12 = mul(2.0f, sub(yz, wx));
    const float r20 = mul(2.0f, sub(xz, wy)), r21 = mul(2.0f, add(yz, wx)), r22 = sub(1.0f, mul(2.0f, add(xx, yy)));
    L[0] = mul(r00, s), L[1] = mul(r01, s), L[2] = mul(r02, s), L[3] = px;
    L[4] = mul(r10, s), L[5] = mul(r11, s), L[6] = mul(r12, s), L[7] = py;
    L[8] = mul(r20, s), L[9] = mul(r21, s), L[10] = mul(r22, s), L[11] = pz;
}

// W = P o L  (both 3x4 affine)
__device__ __forceinline__ void compose(float (&W)[12], const float (&P)[12], const float (&L)[12]) {
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const float a = P[r * 4 + 0], b = P[r * 4 + 1], c = P[r * 4 + 2], t = P[r * 4 + 3];
#pragma unroll
        for (int col = 0; col < 3; col++)
            W[r * 4 + col] = add(add(mul(a, L[col]), mul(b, L[4 + col])), mul(c, L[8 + col]));
        W[r * 4 + 3] = add(add(add(mul(a, L[3]), mul(b, L[7])), mul(c, L[11])), t);
    }
}

// world-space AABB of (c,h) under W, tested against 6 planes (nx,ny,nz,d), inside = n.x + d >= 0
__device__ __forceinline__ bool aabb_visible(const float (&W)[12], float c0, float c1, float c2, float h0, float h1,
                                             float h2, const float *__restrict__ fr /* 24 floats, shared memory */) {
    float wc[3], wh[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        wc[r] = add(add(add(mul(W[r * 4 + 0], c0), mul(W[r * 4 + 1], c1)), mul(W[r * 4 + 2], c2)), W[r * 4 + 3]);
        wh[r] = add(add(mul(fabsf(W[r * 4 + 0]), h0), mul(fabsf(W[r * 4 + 1]), h1)), mul(fabsf(W[r * 4 + 2]), h2));
    }
    bool visible = true;
#pragma unroll
    for (int p = 0; p < 6; p++) {
        const float nx = fr[p * 4 + 0], ny = fr[p * 4 + 1], nz = fr[p * 4 + 2], d = fr[p * 4 + 3];
        const float dist = add(add(add(mul(nx, wc[0]), mul(ny, wc[1])), mul(nz, wc[2])), d);
        const float rad = add(add(mul(fabsf(nx), wh[0]), mul(fabsf(ny), wh[1])), mul(fabsf(nz), wh[2]));
        if (!(add(dist, rad) >= 0.0f)) visible = false;
    }
    return visible;
}

// ---------------------------------------------------------------------------------------------
// argument blocks (word strides; one block per kernel, passed by value)
// ---------------------------------------------------------------------------------------------
struct Plane { // a resource of the batch
    uint32_t *base;
    uint64_t world_words;
};
__device__ __forceinline__ float *fptr(const Plane &p, uint64_t w) { return (float *)(p.base + w * p.world_words); }
__device__ __forceinline__ uint32_t *uptr(const Plane &p, uint64_t w) { return p.base + w * p.world_words; }

struct StepArgs {
    Plane pos, vel, dt;                             // integrate
    Plane world, rot, scale, parent, level_start;   // propagate
    Plane vis_mask, vis_list, vis_count;            // cull outputs
    Plane aabb_c, aabb_h, frustum;                  // cull inputs
    unsigned long long *stats;                      // FIMO_B200_STAT_* words of the batch
    unsigned long long *scratch;                    // [0..3] accumulators, [4] ticket
    uint32_t n_worlds, entities, levels, mask_words;
};

// One entity of one hierarchy level: integrate (optional), build the local transform, compose with
// the parent's world transform (read through `parent_world(i, p)`), hand the result to `emit(i, v)`.
template <bool kIntegrate, typename ParentFn, typename EmitFn>
__device__ __forceinline__ void propagate_entity(uint32_t e, uint32_t E, float *__restrict__ pos, const float *__restrict__ vel,
                                                 float dt, const float *__restrict__ rot, const float *__restrict__ scale,
                                                 const uint32_t *__restrict__ parent, ParentFn parent_world, EmitFn emit) {
    float px = pos[e], py = pos[E + e], pz = pos[2 * E + e];
    if (kIntegrate) {
        px = add(px, mul(vel[e], dt));
        py = add(py, mul(vel[E + e], dt));
        pz = add(pz, mul(vel[2 * E + e], dt));
        pos[e] = px, pos[E + e] = py, pos[2 * E + e] = pz;
    }
    float L[12], W[12];
    local_trs(L, px, py, pz, rot[e], rot[E + e], rot[2 * E + e], rot[3 * E + e], scale[e]);
    const uint32_t p = parent[e];
    if (p == kRoot) {
#pragma unroll
        for (int i = 0; i < 12; i++) W[i] = L[i];
    } else {
        float P[12];
#pragma unroll
        for (int i = 0; i < 12; i++) P[i] = parent_world(i, p);
        compose(W, P, L);
    }
#pragma unroll
    for (int i = 0; i < 12; i++) emit(i, W[i]);
}

// Folds one CTA's statistics into the batch accumulators; the last CTA of the grid publishes them
// and re-arms the scratch block (so no memset node is needed between generations).
__device__ __forceinline__ void publish_stats(const StepArgs &a, unsigned long long visible, unsigned long long mask_sum,
                                              unsigned long long worlds) {
    atomicAdd(&a.scratch[0], visible);
    atomicAdd(&a.scratch[1], mask_sum);
    atomicAdd(&a.scratch[2], worlds);
    __threadfence();
    const unsigned long long ticket = atomicAdd(&a.scratch[4], 1ull);
    if (ticket == (unsigned long long)gridDim.x - 1) {
        __threadfence();
        const unsigned long long v = atomicExch(&a.scratch[0], 0ull);
        const unsigned long long m = atomicExch(&a.scratch[1], 0ull);
        const unsigned long long n = atomicExch(&a.scratch[2], 0ull);
        a.stats[FIMO_B200_STAT_VISIBLE] = v;
        a.stats[FIMO_B200_STAT_ENTITIES] = n * (unsigned long long)a.entities;
        a.stats[FIMO_B200_STAT_MASK_SUM] = m;
        a.stats[FIMO_B200_STAT_WORLDS] = n;
        a.stats[FIMO_B200_STAT_STEPS] += n;
        atomicExch(&a.scratch[4], 0ull);
    }
}

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
    x.x = add(x.x, mul(v.x, step));
    x.y = add(x.y, mul(v.y, step));
    x.z = add(x.z, mul(v.z, step));
    x.w = add(x.w, mul(v.w, step));
    *p = x;
}
__global__ void __launch_bounds__(256) k_integrate_s(Plane pos, Plane vel, Plane dt, uint32_t words_per_world,
                                                     uint32_t chunks_per_world) {
    const uint64_t w = blockIdx.x / chunks_per_world;
    const uint32_t i = (blockIdx.x % chunks_per_world) * 256 + threadIdx.x;
    if (i >= words_per_world) return;
    const float step = *fptr(dt, w);
    float *p = fptr(pos, w) + i;
    *p = add(*p, mul(fptr(vel, w)[i], step));
}

// propagate, one CTA per world: the world's 3x4 matrices are staged in shared memory ([12][E]
// planes, conflict-free), levels are walked in order with one __syncthreads per level.
template <int T, bool kIntegrate, bool kCull>
__global__ void __launch_bounds__(T) k_world_cta(StepArgs a) {
    extern __shared__ float s_world[]; // [12][E]
    __shared__ float s_frustum[24];
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
    __shared__ float s_frustum[24];
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

// ---------------------------------------------------------------------------------------------
// host side: argument checking and launches
// ---------------------------------------------------------------------------------------------
struct Fail {
    const char *what;
};

Plane plane_of(const FimoB200_ResourceView *v, uint32_t scope, uint32_t components, uint64_t count, uint64_t n_worlds,
               const char *name) {
    if (!v || !v->device_ptr) throw Fail{name};
    if (v->elem_bytes != 4 || v->scope != scope || (components && v->components != components) || v->count != count ||
        v->n_worlds != n_worlds || v->comp_stride != 4 * count || v->world_stride != 4ull * v->components * count)
        throw Fail{name};
    return Plane{(uint32_t *)v->device_ptr, v->world_stride / 4};
}

FimoStatus fail(const char *what) {
    std::fprintf(stderr, "fimo_worlds_b200: built-in system: resource `%s` does not have the expected shape\n", what);
    return FIMO_STATUS_ERROR;
}

int sm_count(int device) {
    static int cached[64] = {0};
    if (device < 0 || device >= 64) device = 0;
    if (!cached[device]) cudaDeviceGetAttribute(&cached[device], cudaDevAttrMultiProcessorCount, device);
    return cached[device] ? cached[device] : 148;
}

constexpr size_t kMaxDynamicSmem = 200 * 1024; // of the 227 KB a CTA may use on sm_100a

template <typename K> bool allow_smem(K kernel, size_t bytes) {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) == cudaSuccess;
}

FimoStatus check_launch(const char *name) {
    const cudaError_t err = cudaGetLastError();
    if (err == cudaSuccess) return FIMO_STATUS_OK;
    std::fprintf(stderr, "fimo_worlds_b200: launch of %s failed: %s\n", name, cudaGetErrorString(err));
    return FIMO_STATUS_ERROR;
}

void fill_propagate(StepArgs &a, const FimoB200_ResourceView *world, const FimoB200_ResourceView *pos,
                    const FimoB200_ResourceView *rot, const FimoB200_ResourceView *scale,
                    const FimoB200_ResourceView *parent, const FimoB200_ResourceView *level_start, uint64_t W, uint64_t E) {
    a.world = plane_of(world, FIMO_B200_SCOPE_ENTITY, 12, E, W, "world");
    a.pos = plane_of(pos, FIMO_B200_SCOPE_ENTITY, 3, E, W, "pos");
    a.rot = plane_of(rot, FIMO_B200_SCOPE_ENTITY, 4, E, W, "rot");
    a.scale = plane_of(scale, FIMO_B200_SCOPE_ENTITY, 1, E, W, "scale");
    a.parent = plane_of(parent, FIMO_B200_SCOPE_ENTITY, 1, E, W, "parent");
    a.level_start = plane_of(level_start, FIMO_B200_SCOPE_WORLD, 0, 1, W, "level_start");
    if (level_start->components < 2 || level_start->components > kMaxLevels + 1) throw Fail{"level_start"};
    a.levels = level_start->components - 1;
}

void fill_cull(StepArgs &a, const FimoB200_ResourceView *mask, const FimoB200_ResourceView *list,
               const FimoB200_ResourceView *count, const FimoB200_ResourceView *stats, const FimoB200_ResourceView *world,
               const FimoB200_ResourceView *aabb_c, const FimoB200_ResourceView *aabb_h,
               const FimoB200_ResourceView *frustum, void *scratch, uint64_t W, uint64_t E) {
    a.vis_mask = plane_of(mask, FIMO_B200_SCOPE_ENTITY_BITS, 1, (E + 31) / 32, W, "vis_mask");
    a.vis_list = plane_of(list, FIMO_B200_SCOPE_ENTITY, 1, E, W, "vis_list");
    a.vis_count = plane_of(count, FIMO_B200_SCOPE_WORLD, 1, 1, W, "vis_count");
    a.world = plane_of(world, FIMO_B200_SCOPE_ENTITY, 12, E, W, "world");
    a.aabb_c = plane_of(aabb_c, FIMO_B200_SCOPE_ENTITY, 3, E, W, "aabb_c");
    a.aabb_h = plane_of(aabb_h, FIMO_B200_SCOPE_ENTITY, 3, E, W, "aabb_h");
    a.frustum = plane_of(frustum, FIMO_B200_SCOPE_WORLD, 24, 1, W, "frustum");
    if (!stats || !stats->device_ptr || stats->scope != FIMO_B200_SCOPE_BATCH || stats->components != 2 * FIMO_B200_STAT_WORDS)
        throw Fail{"stats"};
    if (!scratch) throw Fail{"cull scratch"};
    a.stats = (unsigned long long *)stats->device_ptr;
    a.scratch = (unsigned long long *)scratch;
    a.mask_words = (uint32_t)((E + 31) / 32);
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
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    try {
        StepArgs a{};
        a.n_worlds = (uint32_t)W, a.entities = (uint32_t)E;
        fill_cull(a, exclusive[0], exclusive[1], exclusive[2], exclusive[3], shared[0], shared[1], shared[2], shared[3],
                  ((BuiltinState *)system)->scratch, W, E);
        cudaStream_t stream = (cudaStream_t)info->stream;
        const unsigned grid = (unsigned)(W < 8ull * sm_count(info->device) ? W : 8ull * sm_count(info->device));
        if (E <= 2048) k_cull_cta<256><<<grid, 256, 0, stream>>>(a);
        else k_cull_cta<1024><<<grid, 1024, 0, stream>>>(a);
        return check_launch("cull");
    } catch (const Fail &f) {
        return fail(f.what);
    }
}

bool world_step_supported(const WorldStepViews &v) {
    return v.world && 48 * v.world->count <= kMaxDynamicSmem && v.world->count > 0;
}

FimoStatus world_step_launch(const WorldStepViews &v, void *cull_scratch, const FimoB200_LaunchInfo *info) {
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    try {
        StepArgs a{};
        a.n_worlds = (uint32_t)W, a.entities = (uint32_t)E;
        fill_propagate(a, v.world, v.pos, v.rot, v.scale, v.parent, v.level_start, W, E);
        fill_cull(a, v.vis_mask, v.vis_list, v.vis_count, v.stats, v.world, v.aabb_c, v.aabb_h, v.frustum, cull_scratch, W, E);
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
