// kernels_common.cuh — math, argument blocks and host-side launch helpers shared by kernels.cu
// (per-system kernels, v1 world-step) and world_step.cu (TMA-pipelined world-step).
#pragma once
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
__device__ __forceinline__ float fma1(float a, float b, float c) { return __fmaf_rn(a, b, c); } // one rounding

// local = T(pos) * R(quat x,y,z,w) * S(uniform s): 3x4 row-major affine.  33 roundings.
__device__ __forceinline__ void local_trs(float (&L)[12], float px, float py, float pz, float x, float y, float z, float w,
                                          float s) {
    const float x2 = add(x, x), y2 = add(y, y), z2 = add(z, z);
    const float xx = mul(x, x2), yy = mul(y, y2), zz = mul(z, z2);
    const float xy = mul(x, y2), xz = mul(x, z2), yz = mul(y, z2);
    const float wx = mul(w, x2), wy = mul(w, y2), wz = mul(w, z2);
    L[0] = fma1(-add(yy, zz), s, s), L[1] = mul(sub(xy, wz), s), L[2] = mul(add(xz, wy), s), L[3] = px;
    L[4] = mul(add(xy, wz), s), L[5] = fma1(-add(xx, zz), s, s), L[6] = mul(sub(yz, wx), s), L[7] = py;
    L[8] = mul(sub(xz, wy), s), L[9] = mul(add(yz, wx), s), L[10] = fma1(-add(xx, yy), s, s), L[11] = pz;
}

// W = P o L  (both 3x4 affine): 36 roundings
__device__ __forceinline__ void compose(float (&W)[12], const float (&P)[12], const float (&L)[12]) {
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const float a = P[r * 4 + 0], b = P[r * 4 + 1], c = P[r * 4 + 2], t = P[r * 4 + 3];
#pragma unroll
        for (int col = 0; col < 3; col++) W[r * 4 + col] = fma1(c, L[8 + col], fma1(b, L[4 + col], mul(a, L[col])));
        W[r * 4 + 3] = fma1(c, L[11], fma1(b, L[7], fma1(a, L[3], t)));
    }
}

// world-space AABB of (c,h) under W: centre wc = W c + t, half extent wh = |W| h.  18 roundings.
__device__ __forceinline__ void world_aabb_of(float (&wc)[3], float (&wh)[3], const float (&W)[12], float c0, float c1, float c2,
                                              float h0, float h1, float h2) {
#pragma unroll
    for (int r = 0; r < 3; r++) {
        wc[r] = fma1(W[r * 4 + 2], c2, fma1(W[r * 4 + 1], c1, fma1(W[r * 4 + 0], c0, W[r * 4 + 3])));
        wh[r] = fma1(fabsf(W[r * 4 + 2]), h2, fma1(fabsf(W[r * 4 + 1]), h1, mul(fabsf(W[r * 4 + 0]), h0)));
    }
}

// AABB (wc, wh) against 6 planes (nx,ny,nz,d), inside = n.x + d >= 0.  36 roundings.
__device__ __forceinline__ bool planes_visible(const float (&wc)[3], const float (&wh)[3],
                                               const float *__restrict__ fr /* 24 floats, shared memory, 16-byte aligned */) {
    bool visible = true;
#pragma unroll
    for (int p = 0; p < 6; p++) {
        const float4 pl = *(const float4 *)(fr + p * 4);
        float sd = fma1(pl.z, wc[2], fma1(pl.y, wc[1], fma1(pl.x, wc[0], pl.w)));
        sd = fma1(fabsf(pl.z), wh[2], fma1(fabsf(pl.y), wh[1], fma1(fabsf(pl.x), wh[0], sd)));
        if (!(sd >= 0.0f)) visible = false;
    }
    return visible;
}

__device__ __forceinline__ bool aabb_visible(const float (&W)[12], float c0, float c1, float c2, float h0, float h1,
                                             float h2, const float *__restrict__ fr) {
    float wc[3], wh[3];
    world_aabb_of(wc, wh, W, c0, c1, c2, h0, h1, h2);
    return planes_visible(wc, wh, fr);
}

// q += (dt/2) (omega (x) q); q /= |q|   (IEEE sqrt and division).  h = 0.5 * dt, already rounded.
__device__ __forceinline__ void integrate_quat(float &x, float &y, float &z, float &w, float ax, float ay, float az, float h) {
    const float dx = fma1(ay, z, fma1(-az, y, mul(ax, w)));
    const float dy = fma1(az, x, fma1(-ax, z, mul(ay, w)));
    const float dz = fma1(ax, y, fma1(-ay, x, mul(az, w)));
    const float dw = -fma1(az, z, fma1(ay, y, mul(ax, x)));
    x = fma1(dx, h, x), y = fma1(dy, h, y), z = fma1(dz, h, z), w = fma1(dw, h, w);
    const float len = __fsqrt_rn(fma1(w, w, fma1(z, z, fma1(y, y, mul(x, x)))));
    x = __fdiv_rn(x, len), y = __fdiv_rn(y, len), z = __fdiv_rn(z, len), w = __fdiv_rn(w, len);
}

// order-preserving int encoding of a finite float (for atomicMin / atomicMax on bounds)
__device__ __forceinline__ int float_key(float f) {
    const int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7FFFFFFF); }

// ---------------------------------------------------------------------------------------------
// argument blocks (word strides; one block per kernel, passed by value)
// ---------------------------------------------------------------------------------------------
struct Plane { // a resource of the batch
    uint32_t *base;
    uint64_t world_words;
};
__device__ __forceinline__ float *fptr(const Plane &p, uint64_t w) { return (float *)(p.base + w * p.world_words); }
__device__ __forceinline__ uint32_t *uptr(const Plane &p, uint64_t w) { return p.base + w * p.world_words; }

// Multi-GPU statistics exchange over peer memory (fimo_b200_group_stats_peers_bind): where the publishing CTA stores
// a generation's statistics — every rank's inbox, [slots][kStatsMaxRanks][16] u64 in that rank's HBM, mapped here
// through cudaIpc (NVLink).  Each u64 is {generation tag : 32, half of a statistics word : 32}: an aligned 8-byte
// store arrives whole, so a reader that finds all 16 tags of (slot, rank) equal to the generation's tag has the
// words — no fence, no flag (the LL protocol of NCCL, used here for 64 bytes per generation).
constexpr int kStatsMaxRanks = FIMO_B200_STATS_MAX_RANKS;
struct __align__(16) PeerTable {
    uint32_t nranks, rank, slots, pad;
    unsigned long long *inbox[kStatsMaxRanks];
};
__device__ __forceinline__ void st_relaxed_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

struct StepArgs {
    Plane pos, vel, dt;                             // integrate
    Plane world, rot, scale, parent, level_start;   // propagate
    Plane vis_mask, vis_list, vis_count;            // cull outputs
    Plane aabb_c, aabb_h, frustum;                  // cull inputs
    unsigned long long *stats;                      // FIMO_B200_STAT_* words of the batch
    unsigned long long *scratch;                    // per-generation accumulator blocks + counters (kStatScratchBytes, see stats_contribute)
    unsigned long long *ring;                       // per-generation snapshots of `stats` ([ring_slots][8]) or nullptr
    uint32_t ring_slots;
    const PeerTable *peers;                         // multi-GPU statistics exchange over peer memory, or nullptr
    // the fused 8-system DAG (world_step.cu, kSet = 1) additionally uses:
    Plane angvel, gravity, vmax, waabb, wbounds;
    uint32_t *gstats;     // FIMO_B200_GSTAT_* words of the batch
    int *bounds_keys;     // [kStatSlots][6] batch bounds as float_key accumulators (one set per generation of a launch), then [n_worlds][6] per-world ones (team mode)
    // the fused 16-system DAG (kSet = 2) additionally uses:
    Plane damp, sleep_eps, campos, lod_dist2, frustum2;     // per-world constants
    Plane sleep, lod, vis_mask2, vis_list2, vis_count2, vbounds, prev_world;
    unsigned long long *stats2, *scratch2; // cull2's statistics and its accumulator block
    uint32_t *gstats2;                     // group_stats2
    int *bounds_keys2;                     // [kStatSlots][6] batch bounds of the visible entities (float keys)
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
        px = fma1(vel[e], dt, px);
        py = fma1(vel[E + e], dt, py);
        pz = fma1(vel[2 * E + e], dt, pz);
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
// Returns true on the last CTA (the one that published).  kGroupStats: the fused 8-system DAG also publishes
// BUILTIN_GROUP_STATS' words — total visible, worlds, and the batch bounds every CTA merged into
// a.bounds_keys[0..5] (float keys) before calling — and re-arms those keys.
struct StatsSink { // where one cull's batch statistics go (the DAGs have up to two culls)
    unsigned long long *stats, *scratch;
    uint32_t *gstats; // group statistics fed by this cull (nullptr: none)
    int *bounds_keys; // batch bounds accumulators that go with gstats
    uint32_t entities;
    unsigned long long *ring = nullptr; // FimoB200_GroupDeviceExt::stats_ring
    uint32_t ring_slots = 0;
    const PeerTable *peers = nullptr;   // fimo_b200_group_stats_peers_bind
};
// Accumulator blocks in `scratch`, one per generation a launch steps (a CTA of a multi-generation launch is in any
// generation at any time, so each generation gets its own sums): slot s at scratch[8 s ..]: [0] visible, [1] mask
// sum, [2] worlds; scratch[4] (slot 0's spare word) is the launch's completion ticket.  Behind them
// scratch[kStatGlobals] counts the generations published since the system was created (ring / inbox row index).
constexpr int kStatSlots = 16;
constexpr int kStatGlobals = kStatSlots * 8;
constexpr size_t kStatScratchBytes = (kStatGlobals + 8) * 8;
static_assert(kStatScratchBytes <= kBuiltinScratchBytes, "built-in scratch block too small");
// The three steps of publishing: every CTA adds its sums (fire-and-forget reductions), draws a ticket behind a fence,
// and the CTA that drew the last one turns the accumulators into the published words and re-arms them.
__device__ __forceinline__ void stats_contribute(const StatsSink &a, uint32_t slot, unsigned long long visible, unsigned long long mask_sum,
                                                 unsigned long long worlds) {
    atomicAdd(&a.scratch[slot * 8 + 0], visible);
    atomicAdd(&a.scratch[slot * 8 + 1], mask_sum);
    atomicAdd(&a.scratch[slot * 8 + 2], worlds);
}
__device__ __forceinline__ bool stats_last_ticket(const StatsSink &a, uint32_t slot, unsigned long long n_ctas) {
    __threadfence();
    const unsigned long long ticket = atomicAdd(&a.scratch[slot * 8 + 4], 1ull);
    if (ticket != n_ctas - 1) return false;
    __threadfence();
    return true;
}
// The words of generation `gen` into every rank's inbox (our own included), tagged with the generation.
__device__ __forceinline__ void stats_store_peers(const StatsSink &a, const uint4 &peers /* nranks, rank, slots */, unsigned long long gen,
                                                  const unsigned long long (&words)[5], const unsigned long long (&rest)[FIMO_B200_STAT_WORDS - 5]) {
    const unsigned long long tag = ((gen + 1) & 0xFFFFFFFFull) << 32;
    const unsigned long long at = ((gen % peers.z) * kStatsMaxRanks + peers.y) * 16ull;
    for (uint32_t r = 0; r < peers.x; r++) {
        unsigned long long *dst = a.peers->inbox[r] + at;
        for (int i = 0; i < FIMO_B200_STAT_WORDS; i++) {
            const unsigned long long w = i < 5 ? words[i] : rest[i - 5];
            st_relaxed_sys(dst + 2 * i, tag | (w & 0xFFFFFFFFull));
            st_relaxed_sys(dst + 2 * i + 1, tag | (w >> 32));
        }
    }
}
template <bool kGroupStats>
__device__ __forceinline__ void stats_finalize(const StatsSink &a, uint32_t slot) {
    // This runs on the kernel's critical path (the last CTA, after everyone else has left): the loads that do not
    // depend on the accumulators go first so that their latency overlaps the exchanges, and the ring slot is written
    // from registers instead of reading `stats` back.
    const unsigned long long steps_before = a.stats[FIMO_B200_STAT_STEPS];
    unsigned long long gen = 0, rest[FIMO_B200_STAT_WORDS - 5] = {};
    uint4 peers = make_uint4(0u, 0u, 1u, 0u); // nranks, rank, slots
    if (a.peers) peers = *(const uint4 *)a.peers;
    if (a.ring || a.peers) {
        gen = a.scratch[kStatGlobals];
        for (int i = 5; i < FIMO_B200_STAT_WORDS; i++) rest[i - 5] = a.stats[i];
    }
    const unsigned long long v = atomicExch(&a.scratch[slot * 8 + 0], 0ull);
    const unsigned long long m = atomicExch(&a.scratch[slot * 8 + 1], 0ull);
    const unsigned long long n = atomicExch(&a.scratch[slot * 8 + 2], 0ull);
    unsigned long long words[5];
    words[FIMO_B200_STAT_VISIBLE] = v;
    words[FIMO_B200_STAT_ENTITIES] = n * (unsigned long long)a.entities;
    words[FIMO_B200_STAT_MASK_SUM] = m;
    words[FIMO_B200_STAT_WORLDS] = n;
    words[FIMO_B200_STAT_STEPS] = steps_before + n;
    for (int i = 0; i < 5; i++) a.stats[i] = words[i];
    if (a.ring || a.peers) a.scratch[kStatGlobals] = gen + 1;
    if (a.ring) { // the snapshot of this generation: slot = generations published so far % slots
        unsigned long long *row = a.ring + (gen % a.ring_slots) * FIMO_B200_STAT_WORDS;
        for (int i = 0; i < 5; i++) row[i] = words[i];
        for (int i = 5; i < FIMO_B200_STAT_WORDS; i++) row[i] = rest[i - 5];
    }
    if (a.peers) stats_store_peers(a, peers, gen, words, rest);
    if (kGroupStats) {
        uint32_t *g = a.gstats;
        for (int i = 0; i < FIMO_B200_GSTAT_WORDS; i++) g[i] = 0u;
        g[FIMO_B200_GSTAT_VISIBLE_LO] = (uint32_t)(v & 0xFFFFFFFFull), g[FIMO_B200_GSTAT_VISIBLE_LO + 1] = (uint32_t)(v >> 32);
        g[FIMO_B200_GSTAT_WORLDS_LO] = (uint32_t)(n & 0xFFFFFFFFull), g[FIMO_B200_GSTAT_WORLDS_LO + 1] = (uint32_t)(n >> 32);
        for (int i = 0; i < 6; i++) {
            const int empty = i < 3 ? float_key(FIMO_B200_BOUNDS_EMPTY_MIN) : float_key(-FIMO_B200_BOUNDS_EMPTY_MIN);
            g[FIMO_B200_GSTAT_BOUNDS + i] = __float_as_uint(key_float(atomicExch(a.bounds_keys + slot * 6 + i, empty)));
        }
    }
}
template <bool kGroupStats>
__device__ __forceinline__ bool publish_stats_to(const StatsSink &a, unsigned long long visible, unsigned long long mask_sum,
                                                 unsigned long long worlds, unsigned long long n_ctas) {
    stats_contribute(a, 0, visible, mask_sum, worlds);
    if (!stats_last_ticket(a, 0, n_ctas)) return false;
    stats_finalize<kGroupStats>(a, 0);
    atomicExch(&a.scratch[4], 0ull); // re-arm the ticket
    return true;
}
// A launch that stepped `gens` generations (world_step.cu): ONE warp publishes them all.  Lane g adds the CTA's sums of
// generation g to that generation's accumulators; one fence and ONE ticket per CTA and launch; the warp of the CTA
// that drew the last ticket turns every generation's accumulators into its published words — `stats` (and the group
// statistics) end up holding the last generation's, every generation gets its own row in the snapshot ring and in the
// peers' inboxes, the step counter advances by every generation's worlds.  kPair: two culls (the 16-system set) share
// the fence and the ticket (the first sink's).
template <bool kGroupStats>
__device__ __forceinline__ void stats_finalize_launch(const StatsSink &a, uint32_t gens, int lane) {
    const bool mine = (uint32_t)lane < gens;
    const unsigned long long steps_before = a.stats[FIMO_B200_STAT_STEPS];
    unsigned long long gen0 = 0, rest[FIMO_B200_STAT_WORDS - 5] = {};
    uint4 peers = make_uint4(0u, 0u, 1u, 0u);
    if (a.peers) peers = *(const uint4 *)a.peers;
    if (a.ring || a.peers) {
        gen0 = a.scratch[kStatGlobals];
        for (int i = 5; i < FIMO_B200_STAT_WORDS; i++) rest[i - 5] = a.stats[i];
    }
    const unsigned long long v = mine ? atomicExch(&a.scratch[lane * 8 + 0], 0ull) : 0ull;
    const unsigned long long m = mine ? atomicExch(&a.scratch[lane * 8 + 1], 0ull) : 0ull;
    const unsigned long long n = mine ? atomicExch(&a.scratch[lane * 8 + 2], 0ull) : 0ull;
    unsigned long long steps = n; // inclusive prefix over the generations: the step counter after generation `lane`
#pragma unroll
    for (int d = 1; d < kStatSlots; d <<= 1) {
        const unsigned long long up = __shfl_up_sync(kFull, steps, d);
        if (lane >= d) steps += up;
    }
    unsigned long long words[5];
    words[FIMO_B200_STAT_VISIBLE] = v;
    words[FIMO_B200_STAT_ENTITIES] = n * (unsigned long long)a.entities;
    words[FIMO_B200_STAT_MASK_SUM] = m;
    words[FIMO_B200_STAT_WORLDS] = n;
    words[FIMO_B200_STAT_STEPS] = steps_before + steps;
    const bool last = (uint32_t)lane + 1 == gens;
    if (last) {
        for (int i = 0; i < 5; i++) a.stats[i] = words[i];
        if (a.ring || a.peers) a.scratch[kStatGlobals] = gen0 + gens;
    }
    // a ring / inbox with fewer rows than the launch has generations keeps the latest ones
    if (mine && a.ring && (uint32_t)lane + a.ring_slots >= gens) {
        unsigned long long *row = a.ring + ((gen0 + lane) % a.ring_slots) * FIMO_B200_STAT_WORDS;
        for (int i = 0; i < 5; i++) row[i] = words[i];
        for (int i = 5; i < FIMO_B200_STAT_WORDS; i++) row[i] = rest[i - 5];
    }
    if (mine && a.peers && (uint32_t)lane + peers.z >= gens) stats_store_peers(a, peers, gen0 + lane, words, rest);
    if (kGroupStats && mine) {
        int keys[6];
        for (int i = 0; i < 6; i++) {
            const int empty = i < 3 ? float_key(FIMO_B200_BOUNDS_EMPTY_MIN) : float_key(-FIMO_B200_BOUNDS_EMPTY_MIN);
            keys[i] = atomicExch(a.bounds_keys + lane * 6 + i, empty); // every generation's accumulators are re-armed
        }
        if (last) {
            uint32_t *g = a.gstats;
            for (int i = 0; i < FIMO_B200_GSTAT_WORDS; i++) g[i] = 0u;
            g[FIMO_B200_GSTAT_VISIBLE_LO] = (uint32_t)(v & 0xFFFFFFFFull), g[FIMO_B200_GSTAT_VISIBLE_LO + 1] = (uint32_t)(v >> 32);
            g[FIMO_B200_GSTAT_WORLDS_LO] = (uint32_t)(n & 0xFFFFFFFFull), g[FIMO_B200_GSTAT_WORLDS_LO + 1] = (uint32_t)(n >> 32);
            for (int i = 0; i < 6; i++) g[FIMO_B200_GSTAT_BOUNDS + i] = __float_as_uint(key_float(keys[i]));
        }
    }
}
template <bool kGroupStats, bool kPair>
__device__ __forceinline__ bool publish_stats_launch(const StatsSink &a, const StatsSink &b, const unsigned long long *vis_a, const unsigned long long *mask_a,
                                                     const unsigned long long *vis_b, const unsigned long long *mask_b, uint32_t gens,
                                                     unsigned long long worlds, unsigned long long n_ctas, int lane) {
    if ((uint32_t)lane < gens) {
        stats_contribute(a, lane, vis_a[lane], mask_a[lane], worlds);
        if (kPair) stats_contribute(b, lane, vis_b[lane], mask_b[lane], worlds);
    }
    __threadfence();
    __syncwarp();
    unsigned long long ticket = 0;
    if (lane == 0) ticket = atomicAdd(&a.scratch[4], 1ull);
    ticket = __shfl_sync(kFull, ticket, 0);
    if (ticket != n_ctas - 1) return false;
    __threadfence();
    stats_finalize_launch<kGroupStats>(a, gens, lane);
    if (kPair) stats_finalize_launch<kGroupStats>(b, gens, lane);
    __syncwarp();
    if (lane == 0) atomicExch(&a.scratch[4], 0ull); // re-arm the ticket
    return true;
}
template <bool kGroupStats = false>
__device__ __forceinline__ bool publish_stats(const StepArgs &a, unsigned long long visible, unsigned long long mask_sum,
                                              unsigned long long worlds, unsigned long long n_ctas) {
    return publish_stats_to<kGroupStats>(StatsSink{a.stats, a.scratch, a.gstats, a.bounds_keys, a.entities, a.ring, a.ring_slots, a.peers}, visible,
                                         mask_sum, worlds, n_ctas);
}
__device__ __forceinline__ bool publish_stats(const StepArgs &a, unsigned long long visible, unsigned long long mask_sum,
                                              unsigned long long worlds) {
    return publish_stats<false>(a, visible, mask_sum, worlds, (unsigned long long)gridDim.x);
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
} // namespace fb::dev
