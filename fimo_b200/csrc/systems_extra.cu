// systems_extra.cu — sm_100a kernels of the built-in systems that make up the 8-system fan-out / fan-in DAG and
// the 16-system DAG with mixed read/write conflicts (BASELINE.json configs[2] and configs[4]).
// Each system is one node of the group's CUDA graph (device.cu); the DAG compiler (worlds.cpp) orders them by
// explicit dependencies and resource conflicts exactly like SystemGroup.zig:160-338.  All of them are HBM
// streaming passes over [world][component][entity] planes.  Arithmetic: one IEEE binary32 rounding per written
// operation, identical to oracle/systems_ref.c (the definition).
#include "kernels_common.cuh"

namespace fb::dev {
namespace {

constexpr int kT = 256;
constexpr float kEmptyMin = FIMO_B200_BOUNDS_EMPTY_MIN;

// grid.x = chunks_per_world * n_worlds; one entity (or word) per thread
struct Tile {
    uint64_t w;
    uint32_t i;
};
__device__ __forceinline__ Tile tile_of(uint32_t chunks_per_world) {
    return Tile{blockIdx.x / chunks_per_world, (blockIdx.x % chunks_per_world) * kT + threadIdx.x};
}

// forces: vel[k] = fma(g[k], dt, vel[k])       words = 3E per world
__global__ void __launch_bounds__(kT) k_forces(Plane vel, Plane gravity, Plane dt, uint32_t E, uint32_t chunks) {
    const Tile t = tile_of(chunks);
    if (t.i >= 3 * E) return;
    float *v = fptr(vel, t.w) + t.i;
    *v = fma1(fptr(gravity, t.w)[t.i / E], *fptr(dt, t.w), *v);
}

// vel_clamp: v = fminf(fmaxf(v, -vmax), vmax)
__global__ void __launch_bounds__(kT) k_vel_clamp(Plane vel, Plane vmax, uint32_t E, uint32_t chunks) {
    const Tile t = tile_of(chunks);
    if (t.i >= 3 * E) return;
    float *v = fptr(vel, t.w) + t.i;
    const float m = *fptr(vmax, t.w);
    *v = fminf(fmaxf(*v, -m), m);
}

// angvel_damp: w = w * damp
__global__ void __launch_bounds__(kT) k_angvel_damp(Plane angvel, Plane damp, uint32_t E, uint32_t chunks) {
    const Tile t = tile_of(chunks);
    if (t.i >= 3 * E) return;
    float *v = fptr(angvel, t.w) + t.i;
    *v = mul(*v, *fptr(damp, t.w));
}

// prev_copy: prev_world = world            words = 12E per world, 128-bit when aligned
__global__ void __launch_bounds__(kT) k_copy_words(uint32_t *__restrict__ dst, const uint32_t *__restrict__ src, uint64_t words) {
    const uint64_t i = (uint64_t)blockIdx.x * kT + threadIdx.x;
    if (i < words) dst[i] = src[i];
}
__global__ void __launch_bounds__(kT) k_copy_vec4(uint4 *__restrict__ dst, const uint4 *__restrict__ src, uint64_t vecs) {
    const uint64_t i = (uint64_t)blockIdx.x * kT + threadIdx.x;
    if (i < vecs) dst[i] = src[i];
}

// integrate_ang: q += (dt/2) (omega (x) q); q /= |q|      (IEEE sqrt and division: __fsqrt_rn, __fdiv_rn)
__global__ void __launch_bounds__(kT) k_integrate_ang(Plane rot, Plane angvel, Plane dt, uint32_t E, uint32_t chunks) {
    const Tile t = tile_of(chunks);
    if (t.i >= E) return;
    const uint32_t e = t.i;
    float *q = fptr(rot, t.w);
    const float *a = fptr(angvel, t.w);
    const float h = mul(0.5f, *fptr(dt, t.w));
    float x = q[e], y = q[E + e], z = q[2 * E + e], w = q[3 * E + e];
    integrate_quat(x, y, z, w, a[e], a[E + e], a[2 * E + e], h);
    q[e] = x, q[E + e] = y, q[2 * E + e] = z, q[3 * E + e] = w;
}

// world_aabb: c' = W c + t, h' = |W| h
__global__ void __launch_bounds__(kT) k_world_aabb(Plane waabb, Plane world, Plane aabb_c, Plane aabb_h, uint32_t E, uint32_t chunks) {
    const Tile t = tile_of(chunks);
    if (t.i >= E) return;
    const uint32_t e = t.i;
    const float *Wp = fptr(world, t.w);
    float W[12];
#pragma unroll
    for (int i = 0; i < 12; i++) W[i] = Wp[(uint64_t)i * E + e];
    const float *c = fptr(aabb_c, t.w), *h = fptr(aabb_h, t.w);
    const float c0 = c[e], c1 = c[E + e], c2 = c[2 * E + e], h0 = h[e], h1 = h[E + e], h2 = h[2 * E + e];
    float *out = fptr(waabb, t.w);
#pragma unroll
    for (int r = 0; r < 3; r++) {
        out[(uint64_t)r * E + e] = fma1(W[r * 4 + 2], c2, fma1(W[r * 4 + 1], c1, fma1(W[r * 4 + 0], c0, W[r * 4 + 3])));
        out[(uint64_t)(3 + r) * E + e] = fma1(fabsf(W[r * 4 + 2]), h2, fma1(fabsf(W[r * 4 + 1]), h1, mul(fabsf(W[r * 4 + 0]), h0)));
    }
}

// lod = #{i : |c' - cam|^2 >= lod_dist2[i]}
__global__ void __launch_bounds__(kT) k_lod(Plane lod, Plane waabb, Plane campos, Plane dist2, uint32_t E, uint32_t chunks) {
    const Tile t = tile_of(chunks);
    if (t.i >= E) return;
    const uint32_t e = t.i;
    const float *wa = fptr(waabb, t.w), *cam = fptr(campos, t.w), *d = fptr(dist2, t.w);
    const float dx = sub(wa[e], cam[0]), dy = sub(wa[E + e], cam[1]), dz = sub(wa[2 * E + e], cam[2]);
    const float d2 = fma1(dz, dz, fma1(dy, dy, mul(dx, dx)));
    uptr(lod, t.w)[e] = (uint32_t)(d2 >= d[0]) + (uint32_t)(d2 >= d[1]) + (uint32_t)(d2 >= d[2]);
}

// sleep bit = |vx| + |vy| + |vz| < eps (one mask word per warp)
__global__ void __launch_bounds__(kT) k_sleep_flags(Plane sleep, Plane vel, Plane eps, uint32_t E, uint32_t words, uint32_t chunks) {
    const Tile t = tile_of(chunks);
    const uint32_t e = t.i;
    bool bit = false;
    if (e < E) {
        const float *v = fptr(vel, t.w);
        bit = add(add(fabsf(v[e]), fabsf(v[E + e])), fabsf(v[2 * E + e])) < *fptr(eps, t.w);
    }
    const unsigned ballot = __ballot_sync(kFull, bit);
    if ((threadIdx.x & 31) == 0 && (e >> 5) < words) uptr(sleep, t.w)[e >> 5] = ballot;
}

// Cull on precomputed world AABBs, one CTA per world (persistent): same mask / ordered list / count / stats
// contract as k_cull_cta in kernels.cu.
template <int T> __global__ void __launch_bounds__(T) k_cull_waabb(StepArgs a, Plane waabb) {
    __shared__ __align__(16) float s_frustum[24];
    __shared__ uint32_t s_warp[2 * (T / 32)];
    const uint32_t E = a.entities;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = T / 32;
    unsigned long long cta_visible = 0, cta_mask_sum = 0, cta_worlds = 0;
    for (uint64_t w = blockIdx.x; w < a.n_worlds; w += gridDim.x) {
        if (tid < 24) s_frustum[tid] = fptr(a.frustum, w)[tid];
        __syncthreads();
        const float *__restrict__ wa = fptr(waabb, w);
        uint32_t *__restrict__ mask = uptr(a.vis_mask, w);
        uint32_t *__restrict__ list = uptr(a.vis_list, w);
        uint32_t running = 0;
        unsigned long long mask_sum = 0;
        for (uint32_t base = 0; base < E; base += T) {
            const uint32_t e = base + tid;
            bool vis = false;
            if (e < E) {
                const float wc[3] = {wa[e], wa[E + e], wa[2 * E + e]};
                const float wh[3] = {wa[3 * (uint64_t)E + e], wa[4 * (uint64_t)E + e], wa[5 * (uint64_t)E + e]};
                vis = planes_visible(wc, wh, s_frustum);
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
        if (lane == 0) {
            s_warp[warp] = (uint32_t)(mask_sum & 0xFFFFFFFFull);
            s_warp[kWarps + warp] = (uint32_t)(mask_sum >> 32);
        }
        __syncthreads();
        if (tid == 0) {
            unsigned long long total = 0;
            for (int i = 0; i < kWarps; i++) total += ((unsigned long long)s_warp[kWarps + i] << 32) | s_warp[i];
            *uptr(a.vis_count, w) = running;
            cta_visible += running, cta_mask_sum += total, cta_worlds += 1;
        }
        __syncthreads();
    }
    if (tid == 0) publish_stats(a, cta_visible, cta_mask_sum, cta_worlds);
}

// bounds of a world's AABBs (all entities, or only those whose bit is set in `mask`): min / max are exact and
// order-independent, so any reduction tree gives the oracle's bits.  One CTA per world.
template <bool kMasked> __global__ void __launch_bounds__(kT) k_bounds(Plane bounds, Plane waabb, Plane mask, uint32_t E) {
    __shared__ float s_red[6][kT / 32];
    const uint64_t w = blockIdx.x;
    const float *wa = fptr(waabb, w);
    const uint32_t *m = kMasked ? uptr(mask, w) : nullptr;
    float lo[3] = {kEmptyMin, kEmptyMin, kEmptyMin}, hi[3] = {-kEmptyMin, -kEmptyMin, -kEmptyMin};
    for (uint32_t e = threadIdx.x; e < E; e += kT) {
        if (kMasked && !((m[e >> 5] >> (e & 31)) & 1u)) continue;
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const float c = wa[(uint64_t)r * E + e], h = wa[(uint64_t)(3 + r) * E + e];
            lo[r] = fminf(lo[r], sub(c, h));
            hi[r] = fmaxf(hi[r], add(c, h));
        }
    }
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            lo[r] = fminf(lo[r], __shfl_xor_sync(kFull, lo[r], d));
            hi[r] = fmaxf(hi[r], __shfl_xor_sync(kFull, hi[r], d));
        }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int r = 0; r < 3; r++) s_red[r][threadIdx.x >> 5] = lo[r], s_red[3 + r][threadIdx.x >> 5] = hi[r];
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = s_red[threadIdx.x][0];
        for (int i = 1; i < kT / 32; i++) v = threadIdx.x < 3 ? fminf(v, s_red[threadIdx.x][i]) : fmaxf(v, s_red[threadIdx.x][i]);
        fptr(bounds, w)[threadIdx.x] = v;
    }
}

// group statistics: u64 sum of vis_count, world count, bounds of all worlds' bounds.  One CTA.
__global__ void __launch_bounds__(kT) k_group_stats(uint32_t *gstats, Plane vis_count, Plane wbounds, uint32_t n_worlds) {
    __shared__ float s_red[6][kT / 32];
    __shared__ unsigned long long s_vis;
    if (threadIdx.x == 0) s_vis = 0;
    __syncthreads();
    unsigned long long vis = 0;
    float lo[3] = {kEmptyMin, kEmptyMin, kEmptyMin}, hi[3] = {-kEmptyMin, -kEmptyMin, -kEmptyMin};
    for (uint64_t w = threadIdx.x; w < n_worlds; w += kT) {
        vis += *uptr(vis_count, w);
        const float *b = fptr(wbounds, w);
#pragma unroll
        for (int r = 0; r < 3; r++) lo[r] = fminf(lo[r], b[r]), hi[r] = fmaxf(hi[r], b[3 + r]);
    }
    atomicAdd(&s_vis, vis); // integer: exact in any order
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            lo[r] = fminf(lo[r], __shfl_xor_sync(kFull, lo[r], d));
            hi[r] = fmaxf(hi[r], __shfl_xor_sync(kFull, hi[r], d));
        }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int r = 0; r < 3; r++) s_red[r][threadIdx.x >> 5] = lo[r], s_red[3 + r][threadIdx.x >> 5] = hi[r];
    __syncthreads();
    if (threadIdx.x < FIMO_B200_GSTAT_WORDS) {
        uint32_t out = 0;
        const int i = threadIdx.x;
        if (i == FIMO_B200_GSTAT_VISIBLE_LO) out = (uint32_t)(s_vis & 0xFFFFFFFFull);
        if (i == FIMO_B200_GSTAT_VISIBLE_LO + 1) out = (uint32_t)(s_vis >> 32);
        if (i == FIMO_B200_GSTAT_WORLDS_LO) out = n_worlds;
        if (i >= FIMO_B200_GSTAT_BOUNDS && i < FIMO_B200_GSTAT_BOUNDS + 6) {
            const int k = i - FIMO_B200_GSTAT_BOUNDS;
            float v = s_red[k][0];
            for (int j = 1; j < kT / 32; j++) v = k < 3 ? fminf(v, s_red[k][j]) : fmaxf(v, s_red[k][j]);
            out = __float_as_uint(v);
        }
        gstats[i] = out;
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------
// device entry points (FimoB200_device_entry): argument order = the descriptor's resource order
// ---------------------------------------------------------------------------------------------
namespace {
struct Dims {
    uint64_t W, E;
    cudaStream_t stream;
};
bool dims_of(const FimoB200_LaunchInfo *info, Dims &d) {
    d = Dims{info->n_worlds, info->entities_per_world, (cudaStream_t)info->stream};
    return d.W != 0 && d.E != 0;
}
uint32_t chunks_for(uint64_t items) { return (uint32_t)((items + kT - 1) / kT); }
Plane ent(const FimoB200_ResourceView *v, uint32_t comps, const Dims &d, const char *name) {
    return plane_of(v, FIMO_B200_SCOPE_ENTITY, comps, d.E, d.W, name);
}
Plane per_world(const FimoB200_ResourceView *v, uint32_t comps, const Dims &d, const char *name) {
    return plane_of(v, FIMO_B200_SCOPE_WORLD, comps, 1, d.W, name);
}
Plane bits(const FimoB200_ResourceView *v, const Dims &d, const char *name) {
    return plane_of(v, FIMO_B200_SCOPE_ENTITY_BITS, 1, (d.E + 31) / 32, d.W, name);
}
template <typename F> FimoStatus guarded(const char *name, F body) {
    try {
        body();
        return check_launch(name);
    } catch (const Fail &f) {
        return fail(f.what);
    }
}
} // namespace

#define ENTRY(fn) \
    FimoStatus fn(void *system, const FimoB200_ResourceView *const *x, const FimoB200_ResourceView *const *s, const FimoB200_LaunchInfo *info)

ENTRY(builtin_forces) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("forces", [&] {
        const uint32_t c = chunks_for(3 * d.E);
        k_forces<<<(unsigned)(c * d.W), kT, 0, d.stream>>>(ent(x[0], 3, d, "vel"), per_world(s[0], 3, d, "gravity"), per_world(s[1], 1, d, "dt"), (uint32_t)d.E, c);
    });
}
ENTRY(builtin_vel_clamp) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("vel_clamp", [&] {
        const uint32_t c = chunks_for(3 * d.E);
        k_vel_clamp<<<(unsigned)(c * d.W), kT, 0, d.stream>>>(ent(x[0], 3, d, "vel"), per_world(s[0], 1, d, "vmax"), (uint32_t)d.E, c);
    });
}
ENTRY(builtin_angvel_damp) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("angvel_damp", [&] {
        const uint32_t c = chunks_for(3 * d.E);
        k_angvel_damp<<<(unsigned)(c * d.W), kT, 0, d.stream>>>(ent(x[0], 3, d, "angvel"), per_world(s[0], 1, d, "damp"), (uint32_t)d.E, c);
    });
}
ENTRY(builtin_integrate_ang) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("integrate_ang", [&] {
        const uint32_t c = chunks_for(d.E);
        k_integrate_ang<<<(unsigned)(c * d.W), kT, 0, d.stream>>>(ent(x[0], 4, d, "rot"), ent(s[0], 3, d, "angvel"), per_world(s[1], 1, d, "dt"), (uint32_t)d.E, c);
    });
}
ENTRY(builtin_world_aabb) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("world_aabb", [&] {
        const uint32_t c = chunks_for(d.E);
        k_world_aabb<<<(unsigned)(c * d.W), kT, 0, d.stream>>>(ent(x[0], 6, d, "waabb"), ent(s[0], 12, d, "world"), ent(s[1], 3, d, "aabb_c"),
                                                              ent(s[2], 3, d, "aabb_h"), (uint32_t)d.E, c);
    });
}
ENTRY(builtin_prev_copy) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("prev_copy", [&] {
        const Plane dst = ent(x[0], 12, d, "prev_world"), src = ent(s[0], 12, d, "world");
        const uint64_t words = 12 * d.E * d.W;
        if (words % 4 == 0 && (uintptr_t)dst.base % 16 == 0 && (uintptr_t)src.base % 16 == 0)
            k_copy_vec4<<<(unsigned)((words / 4 + kT - 1) / kT), kT, 0, d.stream>>>((uint4 *)dst.base, (const uint4 *)src.base, words / 4);
        else
            k_copy_words<<<(unsigned)((words + kT - 1) / kT), kT, 0, d.stream>>>(dst.base, src.base, words);
    });
}
ENTRY(builtin_lod) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("lod", [&] {
        const uint32_t c = chunks_for(d.E);
        k_lod<<<(unsigned)(c * d.W), kT, 0, d.stream>>>(ent(x[0], 1, d, "lod"), ent(s[0], 6, d, "waabb"), per_world(s[1], 3, d, "campos"),
                                                       per_world(s[2], 3, d, "lod_dist2"), (uint32_t)d.E, c);
    });
}
ENTRY(builtin_sleep_flags) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("sleep_flags", [&] {
        const uint32_t c = chunks_for(d.E);
        k_sleep_flags<<<(unsigned)(c * d.W), kT, 0, d.stream>>>(bits(x[0], d, "sleep"), ent(s[0], 3, d, "vel"), per_world(s[1], 1, d, "sleep_eps"),
                                                               (uint32_t)d.E, (uint32_t)((d.E + 31) / 32), c);
    });
}
ENTRY(builtin_cull_waabb) {
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("cull_waabb", [&] {
        StepArgs a{};
        a.n_worlds = (uint32_t)d.W, a.entities = (uint32_t)d.E;
        a.vis_mask = bits(x[0], d, "vis_mask");
        a.vis_list = ent(x[1], 1, d, "vis_list");
        a.vis_count = per_world(x[2], 1, d, "vis_count");
        if (!x[3] || !x[3]->device_ptr || x[3]->scope != FIMO_B200_SCOPE_BATCH || x[3]->components != 2 * FIMO_B200_STAT_WORDS) throw Fail{"stats"};
        a.stats = (unsigned long long *)x[3]->device_ptr;
        a.scratch = (unsigned long long *)((BuiltinState *)system)->scratch;
        a.ring = (unsigned long long *)((BuiltinState *)system)->stats_ring, a.ring_slots = ((BuiltinState *)system)->stats_ring_slots;
        a.peers = (const PeerTable *)((BuiltinState *)system)->stats_peers;
        if (!a.scratch) throw Fail{"cull scratch"};
        a.frustum = per_world(s[1], 24, d, "frustum");
        a.mask_words = (uint32_t)((d.E + 31) / 32);
        const Plane waabb = ent(s[0], 6, d, "waabb");
        const unsigned grid = (unsigned)std::min<uint64_t>(d.W, 8ull * sm_count(info->device));
        if (d.E <= 2048) k_cull_waabb<256><<<grid, 256, 0, d.stream>>>(a, waabb);
        else k_cull_waabb<1024><<<grid, 1024, 0, d.stream>>>(a, waabb);
    });
}
ENTRY(builtin_bounds_reduce) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("bounds_reduce", [&] {
        k_bounds<false><<<(unsigned)d.W, kT, 0, d.stream>>>(per_world(x[0], 6, d, "wbounds"), ent(s[0], 6, d, "waabb"), Plane{nullptr, 0}, (uint32_t)d.E);
    });
}
ENTRY(builtin_vis_bounds) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("vis_bounds", [&] {
        k_bounds<true><<<(unsigned)d.W, kT, 0, d.stream>>>(per_world(x[0], 6, d, "vbounds"), ent(s[0], 6, d, "waabb"), bits(s[1], d, "vis_mask"), (uint32_t)d.E);
    });
}
ENTRY(builtin_group_stats) {
    (void)system;
    Dims d;
    if (!dims_of(info, d)) return FIMO_STATUS_OK;
    return guarded("group_stats", [&] {
        if (!x[0] || !x[0]->device_ptr || x[0]->scope != FIMO_B200_SCOPE_BATCH || x[0]->components != FIMO_B200_GSTAT_WORDS) throw Fail{"gstats"};
        k_group_stats<<<1, kT, 0, d.stream>>>((uint32_t *)x[0]->device_ptr, per_world(s[0], 1, d, "vis_count"), per_world(s[1], 6, d, "wbounds"), (uint32_t)d.W);
    });
}
#undef ENTRY

} // namespace fb::dev
