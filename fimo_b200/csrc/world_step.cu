// world_step.cu — the fused world-step as a persistent, TMA-fed pipeline (sm_100a).
//
// What it replaces: one generation of a system group, which the reference dispatches as one fiber task per system and
// world through the pool's event loop (modules/fimo_worlds/src/SystemGroup.zig:454-487,
// modules/fimo_tasks/src/CommandBuffer.zig:150-201).  Here ONE kernel steps every world of the batch for a whole
// fusable system set (SetTraits below: the 3-system chain, the 8-system DAG, 15 systems of the 16-system DAG); each
// live-in word is read from HBM once and each live-out word written once.
//
// Structure (HBM-bound streaming; no tensor cores):
//   * persistent, warp-specialised CTAs: T/32 compute warps + 1 TMA warp; (T, CTAs/SM) in {(128,4), (256,3), (256,2),
//     (512,1)} chosen per batch by world_step_plan; worlds strided over CTAs, or — few / large worlds — over CTA teams;
//   * a world is cut into chunks of T consecutive entities (entities are sorted by hierarchy level);
//   * the input planes of a chunk are brought into a shared-memory ring by 1-D bulk async copies (cp.async.bulk, the
//     TMA engine, SASS UBLKCP) that signal an mbarrier with complete_tx; the ring is three chunks deep and runs across
//     world boundaries, so HBM latency is hidden by the copy engine, not by occupancy;
//   * one thread owns one entity of the chunk and does all the systems' arithmetic in registers.  Matrices of
//     entities that can still be parents live in a shared-memory window of 48-byte rows; parents beyond the window
//     (and all parents in team mode) travel through a 48-byte-row buffer that stays L2-resident; the levels present
//     in a chunk are walked in order with a named barrier (compute warps only) between them;
//   * cull compaction: __ballot_sync gives the 32-entity mask word, __popc prefix inside the warp and a dp4a scan over
//     the warps' count bytes give each visible entity its slot in the ASCENDING index list — never unordered atomics;
//   * outputs overwrite the thread's own (dead) column of the ring stage and leave through bulk stores issued by the
//     TMA warp; the last CTA publishes the batch statistics (and the per-generation snapshot ring).
#include <algorithm>
#include <cstdlib>
#include <string>
#include <vector>

#include "kernels_common.cuh"

namespace fb::dev {
namespace {

// System sets this kernel fuses.  kSet 0: integrate -> propagate -> cull (3-system chain).
// kSet 1: the 8-system DAG forces -> [vel_clamp] -> integrate_lin || integrate_ang -> propagate -> world_aabb ->
//         cull || bounds_reduce -> group_stats (vel_clamp of the 16-system DAG is a runtime option).
// Input planes of a chunk: pos 0-2, vel 3-5, rot 6-9, scale 10, parent 11, aabb_c 12-14, aabb_h 15-17, [angvel 18-20].
// Output planes left in the same stage: world 0-11, pos 12-14, [vel 15-17, rot 18-21].
template <int kSet> struct SetTraits;
template <> struct SetTraits<0> {
    static constexpr int kIn = 18, kOut = 15, kStage = 18;
};
template <> struct SetTraits<1> {
    static constexpr int kIn = 21, kOut = 22, kStage = 22;
};
// kSet 2: set 1 with vel_clamp plus six more systems of the 16-system DAG that only touch data the kernel already
// holds: angvel_damp (angvel *= damp, written back: planes 22-24), sleep_flags (bit per entity from the clamped
// velocity), lod (u32 per entity from the world AABB centre), cull2 (second frustum: mask2 / list2 / count2 /
// stats2), vis_bounds (bounds over the VISIBLE entities of cull) and group_stats2 — and, as a runtime option
// (kPipePrevCopy), prev_copy: every thread copies its entity's OLD world matrix (12 coalesced plane loads through L2,
// 12 plane stores) before the chunk's new matrices leave the stage, i.e. before the R/W conflict the reference
// resolves by ordering prev_copy in front of propagate (SystemGroup.zig:264-336) can arise.
template <> struct SetTraits<2> {
    static constexpr int kIn = 21, kOut = 25, kStage = 25;
};

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// try_wait with a suspend-time hint: the thread sleeps in hardware until the phase completes (or the hint
// expires) instead of spinning through issue slots the compute warps need.
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity), "r"(0x989680u)
            : "memory");
    return done != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// 1-D bulk async copy global -> shared, completion reported to `bar` in bytes (TMA engine).
__device__ __forceinline__ void bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}

// L2 eviction priorities (createpolicy): team mode streams 3.3 GB per step through the 126 MB L2 while the 48-byte
// parent rows (58 MB, re-read a few tens of microseconds after they are written) must stay in it — the streamed chunk
// inputs and outputs are marked evict_first, the rows evict_last.
__device__ __forceinline__ uint64_t l2_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void bulk_load_hint(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar)), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ void bulk_store_hint(void *dst, const void *src, uint32_t bytes, uint64_t policy) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst), "r"(smem_addr(src)), "r"(bytes), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ void st_row_hint(float4 *dst, const float4 &v, uint64_t policy) {
    asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(dst), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(policy) : "memory");
}
__device__ __forceinline__ float4 ld_row_hint(const float4 *src, uint64_t policy) {
    float4 v;
    asm volatile("ld.global.cg.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(src), "l"(policy) : "memory");
    return v;
}

// 1-D bulk async copy shared -> global (TMA engine), tracked by the issuing thread's bulk async-group.
__device__ __forceinline__ void bulk_store(void *dst, const void *src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_addr(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// all but the two most recent bulk groups have COMPLETED (their global writes are visible to this thread)
__device__ __forceinline__ void bulk_wait_all_but_2() { asm volatile("cp.async.bulk.wait_group 2;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// generic-proxy writes to shared memory become visible to the async proxy (the TMA store reads them)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
template <int kThreads> __device__ __forceinline__ void consumer_sync() { // the T compute threads only (barrier 1)
    asm volatile("bar.sync 1, %0;" ::"n"(kThreads) : "memory");
}


// Debug build switch (make TIMING=1): per-CTA clock64 accounting of where the chunk loop's time goes, printed every 16th
// launch when FIMO_B200_TIMING=1 and the group runs without a CUDA graph.  Compiled out by default.
#ifdef FIMO_B200_TIMING_BUILD
constexpr bool kTimingBuild = true;
#else
constexpr bool kTimingBuild = false;
#endif

struct PipeArgs {
    StepArgs a;
    uint32_t stages;   // ring depth (chunks)
    uint32_t lvl_cap;  // parent matrices kept in shared memory: entities [0, lvl_cap) of a world
    uint32_t chunks;   // chunks per world = ceil(E / T)
    // Generations this launch steps (1..kStatSlots).  A world belongs to one CTA, so generation g + 1 of a world needs
    // nothing but that CTA's own generation g of it.  A CTA therefore walks its worlds in BLOCKS of `block_worlds`
    // consecutive worlds and steps a block through ALL generations before it moves on (block, generation, world,
    // chunk): the block's state — written by generation g, read back by generation g + 1 a few microseconds later —
    // stays in L2, so HBM sees a world's inputs once and its final outputs once per launch instead of once per
    // generation.  The ring runs straight across world, generation and block boundaries; a chunk's loads are ordered
    // behind the stores of its previous generation by a bulk-group wait (the block holds at least stages + 2 chunks).
    // Every generation still writes every output and accumulates its own statistics.  One-CTA-per-world plans only.
    uint32_t gens;
    uint32_t block_worlds;
    uint32_t team;     // CTAs that step one world together (1 = a world belongs to one CTA)
    uint32_t teams;    // gridDim.x / team
    // The worlds this launch steps: [world_first, world_end).  One generation of a one-CTA-per-world plan is two
    // launches — a persistent grid over whole rounds of worlds and a tail grid of one world per CTA (launch_pipe) —
    // whose CTAs share one completion ticket: `stat_ctas` CTAs in all.
    uint32_t world_first, world_end, stat_ctas, trace_base;
    // Parent matrices that do not fit the shared-memory window: [n_worlds][E] rows of 48 bytes in HBM, written
    // by whoever computes the parent and read back through L2 with three 128-bit loads.  nullptr if unused.
    float4 *rows;
    uint32_t flags; // kPipe*
    // Team mode: entities of (world, level) whose matrices are visible in `rows` — the dataflow counters the
    // CTAs of a team synchronise on ([n_worlds][kMaxLevels], zeroed by the compaction kernel).
    uint32_t *lvl_done;
    // Read-back generations (fimo_b200_group_schedule_readback): a second copy of the cull outputs — mask words, ordered list,
    // count; the layout of the resources — goes to this shadow set, so that the pack of generation g can run beside generation
    // g + 1 (which overwrites the resources themselves).  nullptr: none.  One-CTA-per-world plans only.
    uint32_t *rb_mask, *rb_list, *rb_count;
    uint32_t poll_ns; // team mode: pause between two polls of a level counter (FIMO_B200_POLL_NS, default 64)
    // FIMO_B200_TIMING=1 (debug, plain stream launches only): thread 0 of every CTA accumulates clock64 ticks spent
    // [0] waiting for chunk inputs, [1] waiting for a level to complete, [2] in the whole chunk loop, [3] chunks,
    // [4] prefetch section (team mode), [5] thread 0 waiting for its parent row, [6] chunks whose rows were prefetched,
    // [7] the longest chunk loop of any CTA.
    unsigned long long *timing;
};

constexpr uint32_t kTimingCtas = 4096;      // timing build: CTAs whose (smid, loop start ns, loop end ns, chunks, entry ns, exit ns) are recorded
constexpr uint32_t kTimingRow = 6;
constexpr uint32_t kPipeVelClamp = 8u;         // kSet 1: vel_clamp sits between forces and integrate_lin
constexpr uint32_t kPipePrevCopy = 16u;        // kSet 2: prev_copy is a member too (prev_world = the matrices this generation overwrites)
constexpr uint32_t kPipePublishPerChunk = 2u; // FIMO_B200_PUBLISH=chunk: announce level progress after every chunk
constexpr uint32_t kPipePrefetchBlocking = 4u; // FIMO_B200_PREFETCH=2: wait for the next chunk's inputs before this chunk's math
constexpr uint32_t kPipeNoPrefetch = 1u; // FIMO_B200_PREFETCH=0: parent rows are loaded where they are used
constexpr uint32_t kParSlots = 4; // team mode: depth of the parent-plane look-ahead ring
constexpr uint32_t kPipeNoL2Hints = 32u; // FIMO_B200_L2_HINTS=0: team mode without L2 eviction priorities (A/B runs)

// Per-world constants, double-buffered in shared memory (world k uses buffer k & 1): the compute threads
// fetch world k+1's words into a register while world k is being stepped and publish them before the last
// barrier of world k, so a world boundary costs no extra barrier and no exposed HBM latency.
struct WorldConst {
    float frustum[24];              // 6 planes (nx, ny, nz, d); 16-byte aligned for LDS.128
    uint32_t ls[kMaxLevels + 1];    // level_start
    float dt;
    float gravity[3], vmax;         // kSet 1
    uint32_t pad[2];
    float frustum2[24];             // kSet 2 (16-byte aligned)
    float damp, sleep_eps, campos[3], lod_dist2[3];
};
static_assert(sizeof(WorldConst) % 16 == 0, "WorldConst");

// Source plane p of world w (floats / u32), see SetTraits.
__device__ __forceinline__ const uint32_t *plane_src(const StepArgs &a, int p, uint64_t w, uint32_t E) {
    if (p < 3) return uptr(a.pos, w) + (uint64_t)p * E;
    if (p < 6) return uptr(a.vel, w) + (uint64_t)(p - 3) * E;
    if (p < 10) return uptr(a.rot, w) + (uint64_t)(p - 6) * E;
    if (p == 10) return uptr(a.scale, w);
    if (p == 11) return uptr(a.parent, w);
    if (p < 15) return uptr(a.aabb_c, w) + (uint64_t)(p - 12) * E;
    if (p < 18) return uptr(a.aabb_h, w) + (uint64_t)(p - 15) * E;
    return uptr(a.angvel, w) + (uint64_t)(p - 18) * E;
}

// Work of one CTA: worlds world_first + team_id, + teams, ... (below world_end) and of each world the chunks rank, rank + team, ...
struct Walk {
    uint32_t team_id, rank, my_worlds, my_chunks, total, first_world;
    __device__ __forceinline__ Walk(const PipeArgs &pa) {
        team_id = blockIdx.x / pa.team;
        rank = blockIdx.x % pa.team;
        first_world = pa.world_first + team_id;
        my_worlds = pa.world_end > first_world ? (pa.world_end - first_world + pa.teams - 1) / pa.teams : 0;
        my_chunks = pa.chunks > rank ? (pa.chunks - rank + pa.team - 1) / pa.team : 0;
        total = my_worlds * my_chunks;
    }
};

// Position of a CTA in its work: world k (of my_worlds), chunk c, generation g, inside the block of worlds [b0, b1).
// Order: block, generation, world, chunk.  The TMA thread runs two cursors (loads, and stores `stages` behind), the
// compute threads one; all three advance identically.
struct Cursor {
    uint32_t k, c, g, b0, b1;
    __device__ __forceinline__ uint32_t block_end(const PipeArgs &pa, const Walk &wk, uint32_t from) const {
        return wk.my_worlds - from >= 2 * pa.block_worlds ? from + pa.block_worlds : wk.my_worlds; // a short last block joins its neighbour
    }
    __device__ __forceinline__ Cursor(const PipeArgs &pa, const Walk &wk) : k(0), c(wk.rank), g(0), b0(0), b1(0) { b1 = block_end(pa, wk, 0); }
    // the world visited after world k's visit ends (k itself again in the next generation if the block is one world); my_worlds: none
    __device__ __forceinline__ uint32_t next_world(const PipeArgs &pa, const Walk &wk) const {
        if (k + 1 < b1) return k + 1;
        if (g + 1 < pa.gens) return b0;
        return b1; // first world of the next block, or my_worlds
    }
    __device__ __forceinline__ void advance(const PipeArgs &pa, const Walk &wk) {
        if ((c += pa.team) < pa.chunks) return;
        c = wk.rank;
        if (++k < b1) return;
        if (++g < pa.gens) {
            k = b0;
            return;
        }
        g = 0, b0 = b1, b1 = block_end(pa, wk, b0);
    }
};

// The TMA warp (one elected thread): loads chunk inputs into the ring and stores the finished chunk's
// outputs, which the compute threads leave in the same ring stage (world matrices over planes 0-11, new
// positions over 12-14).
template <int T, int kSet, bool kHint>
__device__ __forceinline__ void tma_warp(const PipeArgs &pa, const Walk &wk, float *ring, uint64_t *full, uint64_t *done, uint32_t *s_par,
                                         uint64_t *pfull) {
    constexpr int kIn = SetTraits<kSet>::kIn, kStage = SetTraits<kSet>::kStage;
    const StepArgs &a = pa.a;
    const uint64_t stream_policy = kHint ? l2_evict_first() : 0ull; // kHint (team mode): streamed data leaves L2 first
    auto bulk_store = [&](void *dst, const void *src, uint32_t bytes) {
        if (kHint && !(pa.flags & kPipeNoL2Hints)) bulk_store_hint(dst, src, bytes, stream_policy);
        else fb::dev::bulk_store(dst, src, bytes);
    };
    auto bulk_load = [&](void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
        if (kHint && !(pa.flags & kPipeNoL2Hints)) bulk_load_hint(dst, src, bytes, bar, stream_policy);
        else fb::dev::bulk_load(dst, src, bytes, bar);
    };
    const uint32_t E = a.entities, N = pa.stages, all = wk.total * pa.gens;
    uint32_t s = 0, round = 0;
    Cursor ld(pa, wk), st_(pa, wk); // the chunk being loaded; the chunk being stored (N behind)
    auto store_chunk = [&](uint32_t stage, uint32_t wk_, uint32_t wc) {
        const uint64_t w = wk.first_world + (uint64_t)wk_ * pa.teams;
        const uint32_t first = wc * T, bytes = min((uint32_t)T, E - first) * 4u;
        const float *st = ring + (size_t)stage * kStage * T;
        float *world = fptr(a.world, w) + first;
        float *pos = fptr(a.pos, w) + first;
#pragma unroll
        for (int i = 0; i < 12; i++) bulk_store(world + (uint64_t)i * E, st + i * T, bytes);
#pragma unroll
        for (int i = 0; i < 3; i++) bulk_store(pos + (uint64_t)i * E, st + (12 + i) * T, bytes);
        if (kSet >= 1) {
            float *vel = fptr(a.vel, w) + first;
            float *rot = fptr(a.rot, w) + first;
#pragma unroll
            for (int i = 0; i < 3; i++) bulk_store(vel + (uint64_t)i * E, st + (15 + i) * T, bytes);
#pragma unroll
            for (int i = 0; i < 4; i++) bulk_store(rot + (uint64_t)i * E, st + (18 + i) * T, bytes);
        }
        if (kSet == 2) {
            float *angvel = fptr(a.angvel, w) + first;
#pragma unroll
            for (int i = 0; i < 3; i++) bulk_store(angvel + (uint64_t)i * E, st + (22 + i) * T, bytes);
        }
        bulk_commit();
    };
    for (uint32_t it = 0; it < all; it++) {
        if (it >= N) { // the stage still holds chunk it-N: wait for its outputs, drain them, then reuse it
            mbar_wait(&done[s], (round - 1) & 1u);
            store_chunk(s, st_.k, st_.c);
            st_.advance(pa, wk);
            bulk_wait_read_all();
        }
        // A later generation reads what the previous one stored: the bulk group of this chunk one generation ago must
        // have COMPLETED.  That group lies a whole block (>= N + 2 chunks, world_step_plan) back and groups up to
        // it - N are committed, so "all but the two most recent" covers it.
        if (ld.g > 0) bulk_wait_all_but_2();
        const uint64_t w = wk.first_world + (uint64_t)ld.k * pa.teams;
        const uint32_t first = ld.c * T, bytes = min((uint32_t)T, E - first) * 4u; // multiple of 16 (E % 4 == 0)
        float *st = ring + (size_t)s * kStage * T;
        mbar_expect_tx(&full[s], bytes * kIn);
#pragma unroll
        for (int p = 0; p < kIn; p++) bulk_load(st + p * T, plane_src(a, p, w, E) + first, bytes, &full[s]);
        ld.advance(pa, wk);
        if (kHint && it + 1 < all) { // team mode: parent indices of chunk it + 1, a whole ring turn before its stage is loaded
            // slot it % kParSlots last held chunk it - 3's indices, read while chunk it - 4 was computed: long done (done[s] of it - N)
            const uint64_t nw = wk.first_world + (uint64_t)ld.k * pa.teams;
            const uint32_t nfirst = ld.c * T, nbytes = min((uint32_t)T, E - nfirst) * 4u, slot = it % kParSlots; // use it / kParSlots of that slot
            mbar_expect_tx(&pfull[slot], nbytes);
            fb::dev::bulk_load(s_par + (size_t)slot * T, uptr(a.parent, nw) + nfirst, nbytes, &pfull[slot]);
        }
        if (++s == N) s = 0, round++;
    }
    // drain the last min(N, all) chunks
    const uint32_t tail = all < N ? all : N;
    uint32_t ds = all < N ? 0 : s, dround = all < N ? 0 : round - 1;
    for (uint32_t j = 0; j < tail; j++) {
        mbar_wait(&done[ds], dround & 1u);
        store_chunk(ds, st_.k, st_.c);
        st_.advance(pa, wk);
        if (++ds == N) ds = 0, dround++;
    }
    bulk_wait_all();
    if (kHint && all > 1) { // the compute warps need not have waited for the last parent-plane loads: none may land after the CTA is gone
        for (uint32_t j = all - 1 > kParSlots ? all - 1 - kParSlots : 0; j + 1 < all; j++) mbar_wait(&pfull[j % kParSlots], (j / kParSlots) & 1u);
    }
}

// non-blocking probe of an mbarrier phase
__device__ __forceinline__ bool mbar_test(uint64_t *bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity)
            : "memory");
    return done != 0;
}
// 16-byte asynchronous copy global (L2) -> shared, no registers in between (LDGSTS)
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16_hint(void *dst, const void *src, uint64_t policy) {
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(smem_addr(dst)), "l"(src), "l"(policy) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all_but_latest() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(uint32_t *p, uint32_t v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Warp-specialised: warps [0, T/32) compute (one entity of the chunk per thread), the last warp is the TMA warp.
// kTeam = false: a world belongs to one CTA; mask, ordered list, count and statistics leave this kernel.
// kTeam = true : `team` CTAs step one world together (large worlds, few worlds); chunks are dealt round-robin,
//                parents travel through `rows` in L2, and a CTA starts level l of a chunk once the (world, l-1)
//                counter says every entity of level l-1 is visible — dataflow, no team-wide barrier.  Only the
//                mask leaves this kernel; k_compact_lists derives list, count and statistics from it.
// kShadow: the cull outputs are written twice (PipeArgs::rb_*).  A variant of its own: the three extra predicated stores cost
// the plain kernel 3 % at C2 when they were a run-time option.
template <int T, int kMinBlocks, bool kTeam, int kSet, bool kShadow = false>
__global__ void __launch_bounds__(T + 32, kMinBlocks) k_world_step_pipe(const PipeArgs pa) {
    constexpr int kStage = SetTraits<kSet>::kStage;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const StepArgs &a = pa.a;
    const uint32_t E = a.entities, N = pa.stages, cap = pa.lvl_cap, nc = pa.chunks, G = pa.team;
    float *ring = (float *)smem_raw;                              // [N][kStage][T] planes
    float4 *s_win = (float4 *)(ring + (size_t)N * kStage * T);    // [cap][3]: parent window, one 48-byte matrix per entity
    uint64_t *full = (uint64_t *)(s_win + (size_t)3 * cap);       // [N] inputs have landed
    uint64_t *done = full + N;                                    // [N] outputs are in the stage
    float4 *s_pre = (float4 *)(done + N);                         // team mode: [2][T][3] parent rows prefetched from L2
    // team mode: the parent-index plane of the chunk AFTER the one being loaded, one ring slot ahead of the stage ring
    // ([kParSlots][T] + kParSlots mbarriers), so that the row prefetch for chunk i + 1 never waits for that chunk's stage
    uint32_t *s_par = (uint32_t *)(s_pre + (size_t)2 * T * 3);
    uint64_t *pfull = (uint64_t *)(s_par + (size_t)kParSlots * T);
    __shared__ __align__(16) WorldConst s_c[2];
    __shared__ __align__(16) uint32_t s_cnt[2][4]; // per-warp visible counts of a chunk, one byte per warp
    // this CTA's sums per generation of the launch (one-CTA-per-world plans): visible entities (thread 0), mask words
    // (lane 0 of every warp), batch bounds as float keys (threads 0-5); published by warp 0 when the CTA is done
    constexpr int kGens = kTeam ? 1 : kStatSlots;
    __shared__ unsigned long long s_gen_vis[kGens], s_gen_mask[kGens];
    __shared__ unsigned long long s_gen_vis2[kSet == 2 ? kGens : 1], s_gen_mask2[kSet == 2 ? kGens : 1];
    __shared__ int s_gen_bounds[kSet >= 1 ? kGens : 1][6], s_gen_vbounds[kSet == 2 ? kGens : 1][6];
    __shared__ int s_bounds[2][6]; // kSet >= 1: bounds of the current world as float keys (min xyz, max xyz), by world parity
    __shared__ int s_vbounds[2][6]; // kSet 2: the same over the entities `cull` found visible (vis_bounds)
    __shared__ __align__(16) uint32_t s_cnt2[2][4]; // kSet 2: per-warp counts of cull2
    static_assert(!(kTeam && kSet == 2), "the 16-system set is fused for one-CTA-per-world plans only");

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long tm_entry = 0;
    if (kTimingBuild && pa.timing && tid == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tm_entry));
    constexpr int kWarps = T / 32;
    static_assert(kWarps <= 16, "one count byte per warp in s_cnt");
    const int kEmptyKey = tid % 6 < 3 ? float_key(FIMO_B200_BOUNDS_EMPTY_MIN) : float_key(-FIMO_B200_BOUNDS_EMPTY_MIN);
    if (kSet >= 1 && tid < 12) s_bounds[tid / 6][tid % 6] = kEmptyKey;
    if (kSet == 2 && tid < 12) s_vbounds[tid / 6][tid % 6] = kEmptyKey;
    if (!kTeam && tid < kGens) {
        s_gen_vis[tid] = 0, s_gen_mask[tid] = 0;
        if (kSet == 2) s_gen_vis2[tid] = 0, s_gen_mask2[tid] = 0;
    }
    if (!kTeam && kSet >= 1 && tid < kGens * 6) {
        s_gen_bounds[tid / 6][tid % 6] = kEmptyKey;
        if (kSet == 2) s_gen_vbounds[tid / 6][tid % 6] = kEmptyKey;
    }
    const uint32_t levels = a.levels;
    const Walk wk(pa);
    const uint32_t my_worlds = wk.my_worlds, total = wk.total * (kTeam ? 1u : pa.gens); // chunks of every generation of the launch

    if (tid == 0) {
        for (uint32_t s = 0; s < N; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&done[s], kWarps);
        }
        if (kTeam)
            for (uint32_t s = 0; s < kParSlots; s++) mbar_init(&pfull[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == kWarps) {
        if (lane == 0) tma_warp<T, kSet, kTeam>(pa, wk, ring, full, done, kTeam ? s_par : nullptr, kTeam ? pfull : nullptr);
        return;
    }

    // ----------------------------------- compute warps ------------------------------------------------
    // which word of a world's constants this thread carries: frustum (24), dt (1), level_start (levels+1)
    auto const_word = [&](uint64_t w) -> uint32_t {
        if (tid < 24) return uptr(a.frustum, w)[tid];
        if (tid == 24) return uptr(a.dt, w)[0];
        if (kSet >= 1 && tid >= 25 && tid < 28) return uptr(a.gravity, w)[tid - 25];
        if (kSet >= 1 && tid == 28 && (pa.flags & kPipeVelClamp)) return uptr(a.vmax, w)[0];
        if (tid >= 32 && tid <= 32 + (int)levels) return uptr(a.level_start, w)[tid - 32];
        if (kSet == 2) { // threads 65..96 (T >= 128)
            if (tid >= 65 && tid < 89) return uptr(a.frustum2, w)[tid - 65];
            if (tid == 89) return uptr(a.damp, w)[0];
            if (tid == 90) return uptr(a.sleep_eps, w)[0];
            if (tid >= 91 && tid < 94) return uptr(a.campos, w)[tid - 91];
            if (tid >= 94 && tid < 97) return uptr(a.lod_dist2, w)[tid - 94];
        }
        return 0u;
    };
    auto publish_const = [&](WorldConst &c, uint32_t v) {
        if (tid < 24) c.frustum[tid] = __uint_as_float(v);
        if (tid == 24) c.dt = __uint_as_float(v);
        if (kSet >= 1 && tid >= 25 && tid < 28) c.gravity[tid - 25] = __uint_as_float(v);
        if (kSet >= 1 && tid == 28) c.vmax = __uint_as_float(v);
        if (tid >= 32 && tid <= 32 + (int)levels) c.ls[tid - 32] = v;
        if (kSet == 2) {
            if (tid >= 65 && tid < 89) c.frustum2[tid - 65] = __uint_as_float(v);
            if (tid == 89) c.damp = __uint_as_float(v);
            if (tid == 90) c.sleep_eps = __uint_as_float(v);
            if (tid >= 91 && tid < 94) c.campos[tid - 91] = __uint_as_float(v);
            if (tid >= 94 && tid < 97) c.lod_dist2[tid - 94] = __uint_as_float(v);
        }
    };
    if (total) publish_const(s_c[0], const_word(wk.first_world));
    consumer_sync<T>();

    uint32_t next_const = 0;
    unsigned long long warp_mask_sum = 0; // lane 0 of each warp: mask words of the current world visit (kSet 2 adds them to shared memory per chunk)
    uint32_t running = 0;                                   // visible entities of the current world so far
    uint32_t running2 = 0;
    uint32_t s = 0, parity = 0, l_lo = 0;
    uint32_t kk = 0;  // world visits begun so far: parity of the double-buffered per-world state
    Cursor cur(pa, wk);
    // team mode: entities of level pend_level this CTA has finished but not yet announced.  Dependants only
    // care about a level being COMPLETE, so a CTA announces its share once, when it moves past the level.
    int pend_level = -1;
    uint32_t pend_count = 0;
    auto flush_pending = [&](uint32_t *counters) {
        if (pend_level < 0) return;
        consumer_sync<T>(); // every thread's rows of that level are written
        if (tid == 0) red_release_gpu_add(counters + pend_level, pend_count);
        pend_level = -1, pend_count = 0;
    };
    bool pre_have = false; // team mode: this thread's parent row of the current chunk is (being) prefetched into s_pre
    int ready_upto = -1; // team mode: levels [0, ready_upto] of the current world are known to be complete
    uint32_t *mask_out = nullptr, *list_out = nullptr;
    float4 *rows = nullptr;
    uint32_t *lvl_done = nullptr, *lvl_prev = nullptr;
    int guard_upto = -1; // team mode: rows of levels [0, guard_upto] of the PREVIOUS world are known to have no readers left
    const uint64_t row_policy = kTeam ? l2_evict_last() : 0ull;
    const bool l2_hints = kTeam && !(pa.flags & kPipeNoL2Hints);
    float *waabb_out = nullptr;
    uint32_t *mask2_out = nullptr, *list2_out = nullptr, *sleep_out = nullptr, *lod_out = nullptr;
    const float *world_old = nullptr; // kSet 2 + prev_copy
    float *prev_out = nullptr;

    long long tm_input = 0, tm_level = 0, tm_pre = 0, tm_gather = 0, tm_prefetched = 0;
    const long long tm_begin = (kTimingBuild && pa.timing && tid == 0) ? clock64() : 0;
    unsigned long long tm_wall0 = 0;
    if (kTimingBuild && pa.timing && tid == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tm_wall0));
    for (uint32_t it = 0; it < total; it++) {
        const WorldConst &wc = s_c[kk & 1];
        const uint32_t k = cur.k, c = cur.c, gen = cur.g;
        const uint64_t w = wk.first_world + (uint64_t)k * pa.teams;
        const uint32_t next_k = cur.next_world(pa, wk);
        const bool more_worlds = next_k < my_worlds; // this CTA visits another world (or this one again, a generation on) afterwards
        if (c == wk.rank) { // a world visit begins
            if (kTeam && it != 0) consumer_sync<T>(); // its constants were published during the previous world
            if (more_worlds) next_const = const_word(wk.first_world + (uint64_t)next_k * pa.teams);
            running = 0, running2 = 0, l_lo = 0, ready_upto = -1;
            mask_out = uptr(a.vis_mask, w);
            list_out = uptr(a.vis_list, w);
            // team mode: ONE rows buffer per team, reused by every world the team steps, so the rows' footprint
            // (teams x parents x 48 B) fits L2 and a dirty row line is overwritten in place instead of written back
            rows = pa.rows ? pa.rows + (kTeam ? (uint64_t)wk.team_id : w) * (uint64_t)E * 3 : nullptr;
            if (kTeam) {
                lvl_prev = lvl_done; // the world this team stepped before (nullptr: none): its readers guard row reuse
                lvl_done = pa.lvl_done + w * kMaxLevels;
                guard_upto = -1;
            }
            if (kSet >= 1) waabb_out = fptr(a.waabb, w);
            if (kSet == 2) {
                mask2_out = uptr(a.vis_mask2, w), list2_out = uptr(a.vis_list2, w);
                sleep_out = uptr(a.sleep, w), lod_out = uptr(a.lod, w);
                if (pa.flags & kPipePrevCopy) world_old = fptr(a.world, w), prev_out = fptr(a.prev_world, w);
            }
        }
        const uint32_t first = c * T;
        const uint32_t e = first + tid;
        const bool valid = e < E;
        const uint32_t last = min(E, first + T) - 1;
        // hierarchy levels present in this chunk (uniform), and the level of this thread's entity
        while (l_lo + 1 < levels && first >= wc.ls[l_lo + 1]) l_lo++;
        uint32_t l_hi = l_lo;
        while (l_hi + 1 < levels && last >= wc.ls[l_hi + 1]) l_hi++;
        uint32_t l_me = l_lo;
        while (l_me < l_hi && e >= wc.ls[l_me + 1]) l_me++;
        const bool may_be_parent = l_me + 1 < levels;
        if (kTeam && pend_level >= 0 && (int)l_lo > pend_level) flush_pending(lvl_done); // moved past that level

        const long long t_in0 = (kTimingBuild && pa.timing && tid == 0) ? clock64() : 0;
        mbar_wait(&full[s], parity);
        if (kTimingBuild && pa.timing && tid == 0) tm_input += clock64() - t_in0;
        float *st = ring + (size_t)s * kStage * T + tid;
        float lo[3] = {FIMO_B200_BOUNDS_EMPTY_MIN, FIMO_B200_BOUNDS_EMPTY_MIN, FIMO_B200_BOUNDS_EMPTY_MIN};
        float hi[3] = {-FIMO_B200_BOUNDS_EMPTY_MIN, -FIMO_B200_BOUNDS_EMPTY_MIN, -FIMO_B200_BOUNDS_EMPTY_MIN};
        bool vis = false, vis2 = false, asleep = false;
        bool pre_next = false;
        const long long t_pre0 = (kTimingBuild && kTeam && pa.timing && tid == 0) ? clock64() : 0;
        if (kTeam) {
            // Parent rows come from L2 (~1 us under load): fetch the NEXT chunk's rows now, asynchronously and
            // without registers (cp.async), so that this chunk's arithmetic covers the latency.  Safe when this
            // chunk and the next lie inside one level nl whose parent level is already known to be complete.
            if (c + G < nc && !(pa.flags & kPipeNoPrefetch)) {
                const uint32_t nfirst = (c + G) * T, nlast = min(E, nfirst + T) - 1;
                uint32_t nl = l_hi;
                while (nl + 1 < levels && nfirst >= wc.ls[nl + 1]) nl++;
                const bool single = !(nl + 1 < levels && nlast >= wc.ls[nl + 1]);
                const uint32_t slot = it % kParSlots, spar = (it / kParSlots) & 1u; // loaded by the TMA thread's iteration `it`
                if (single && nl >= 1 && l_lo == nl && ready_upto >= (int)nl - 1 &&
                    ((pa.flags & kPipePrefetchBlocking) ? (mbar_wait(&pfull[slot], spar), true) : mbar_test(&pfull[slot], spar))) {
                    const uint32_t np = s_par[(size_t)slot * T + tid];
                    if (nfirst + tid < E && np != kRoot) {
                        const float4 *src = rows + (uint64_t)np * 3;
                        float4 *dst = s_pre + ((size_t)((it + 1) & 1) * T + tid) * 3;
                        if (l2_hints) cp_async16_hint(dst, src, row_policy), cp_async16_hint(dst + 1, src + 1, row_policy), cp_async16_hint(dst + 2, src + 2, row_policy);
                        else cp_async16(dst, src), cp_async16(dst + 1, src + 1), cp_async16(dst + 2, src + 2);
                    }
                    pre_next = true;
                }
            }
            cp_async_commit(); // one group per iteration, possibly empty
            if (kTimingBuild && pa.timing && tid == 0) tm_pre += clock64() - t_pre0, tm_prefetched += pre_next ? 1 : 0;
        }
        for (uint32_t l = l_lo; l <= l_hi; l++) {
            // Row reuse across worlds: this chunk overwrites the team's rows of level l, which level l + 1 of the
            // PREVIOUS world read.  That level must be complete (it almost always is: the wait returns at once).
            const bool guard = kTeam && lvl_prev && l + 1 < levels && guard_upto < (int)l;
            if (kTeam && ((l >= 1 && ready_upto < (int)l - 1) || guard)) {
                const bool wait_level = l >= 1 && ready_upto < (int)l - 1;
                if (wait_level) flush_pending(lvl_done); // our own share of level l-1 (same chunk) counts too
                if (tid == 0) {
                    const long long t_lv0 = (kTimingBuild && pa.timing) ? clock64() : 0;
                    if (guard) {
                        const WorldConst &pc = s_c[(kk + 1) & 1]; // the previous world's constants are still in the other buffer
                        const uint32_t need_prev = pc.ls[l + 2] - pc.ls[l + 1];
                        while (ld_acquire_gpu(lvl_prev + (l + 1)) < need_prev) __nanosleep(64);
                    }
                    // every entity of level l-1 — in whichever CTA's chunk — must be visible in `rows`
                    if (wait_level) {
                        const uint32_t need = wc.ls[l] - wc.ls[l - 1];
                        while (ld_acquire_gpu(lvl_done + (l - 1)) < need)
                            if (pa.poll_ns) __nanosleep(pa.poll_ns);
                    }
                    if (kTimingBuild && pa.timing) tm_level += clock64() - t_lv0;
                }
                consumer_sync<T>();
                if (wait_level) ready_upto = (int)l - 1;
                if (guard) guard_upto = (int)l;
            }
            if (valid && l_me == l) {
                const float dt = wc.dt;
                float vx = st[3 * T], vy = st[4 * T], vz = st[5 * T];
                float rx = st[6 * T], ry = st[7 * T], rz = st[8 * T], rw = st[9 * T];
                if (kSet >= 1) {
                    vx = fma1(wc.gravity[0], dt, vx), vy = fma1(wc.gravity[1], dt, vy), vz = fma1(wc.gravity[2], dt, vz); // forces
                    if (kSet == 2 || (pa.flags & kPipeVelClamp)) {
                        const float m = wc.vmax;
                        vx = fminf(fmaxf(vx, -m), m), vy = fminf(fmaxf(vy, -m), m), vz = fminf(fmaxf(vz, -m), m);
                    }
                    float ax = st[18 * T], ay = st[19 * T], az = st[20 * T];
                    if (kSet == 2) {
                        ax = mul(ax, wc.damp), ay = mul(ay, wc.damp), az = mul(az, wc.damp); // angvel_damp
                        st[22 * T] = ax, st[23 * T] = ay, st[24 * T] = az;
                        asleep = add(add(fabsf(vx), fabsf(vy)), fabsf(vz)) < wc.sleep_eps;       // sleep_flags
                    }
                    integrate_quat(rx, ry, rz, rw, ax, ay, az, mul(0.5f, dt)); // integrate_ang
                }
                const float px = fma1(vx, dt, st[0 * T]); // integrate (linear)
                const float py = fma1(vy, dt, st[1 * T]);
                const float pz = fma1(vz, dt, st[2 * T]);
                float L[12], Wm[12];
                local_trs(L, px, py, pz, rx, ry, rz, rw, st[10 * T]);
                const uint32_t p = __float_as_uint(st[11 * T]);
                if (p == kRoot) {
#pragma unroll
                    for (int i = 0; i < 12; i++) Wm[i] = L[i];
                } else {
                    float4 r0, r1, r2;
                    if (!kTeam && p < cap) {
                        r0 = s_win[3 * p], r1 = s_win[3 * p + 1], r2 = s_win[3 * p + 2];
                    } else if (kTeam && pre_have) {
                        const long long t_g0 = (kTimingBuild && pa.timing && tid == 0) ? clock64() : 0;
                        cp_async_wait_all_but_latest(); // the group committed one iteration ago
                        const float4 *row = s_pre + ((size_t)(it & 1) * T + tid) * 3;
                        r0 = row[0], r1 = row[1], r2 = row[2];
                        if (kTimingBuild && pa.timing && tid == 0) tm_gather += clock64() - t_g0;
                    } else { // beyond the shared-memory window: the parent's row comes back through L2
                        const float4 *row = rows + (uint64_t)p * 3;
                        const long long t_g0 = (kTimingBuild && pa.timing && tid == 0) ? clock64() : 0;
                        if (l2_hints) r0 = ld_row_hint(row, row_policy), r1 = ld_row_hint(row + 1, row_policy), r2 = ld_row_hint(row + 2, row_policy);
                        else r0 = __ldcg(row), r1 = __ldcg(row + 1), r2 = __ldcg(row + 2);
                        if (kTimingBuild && pa.timing && tid == 0) tm_gather += (long long)(r0.x != r1.x) * 0 + clock64() - t_g0;
                    }
                    const float P[12] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
                    compose(Wm, P, L);
                }
                float wcn[3], whn[3];
                world_aabb_of(wcn, whn, Wm, st[12 * T], st[13 * T], st[14 * T], st[15 * T], st[16 * T], st[17 * T]);
                vis = planes_visible(wcn, whn, wc.frustum);
                if (kSet >= 1) { // world_aabb leaves through plain coalesced stores; bounds_reduce folds it
#pragma unroll
                    for (int r = 0; r < 3; r++) {
                        waabb_out[(uint64_t)r * E + e] = wcn[r];
                        waabb_out[(uint64_t)(3 + r) * E + e] = whn[r];
                        lo[r] = sub(wcn[r], whn[r]), hi[r] = add(wcn[r], whn[r]);
                    }
                }
                if (kSet == 2) {
                    vis2 = planes_visible(wcn, whn, wc.frustum2); // cull2
                    const float dx = sub(wcn[0], wc.campos[0]), dy = sub(wcn[1], wc.campos[1]), dz = sub(wcn[2], wc.campos[2]);
                    const float d2 = fma1(dz, dz, fma1(dy, dy, mul(dx, dx)));
                    lod_out[e] = (uint32_t)(d2 >= wc.lod_dist2[0]) + (uint32_t)(d2 >= wc.lod_dist2[1]) + (uint32_t)(d2 >= wc.lod_dist2[2]); // lod
                }
                if (may_be_parent) { // 48-byte rows: three 128-bit stores, conflict-free in shared memory
                    float4 *row = (!kTeam && e < cap) ? s_win + 3 * e : rows + (uint64_t)e * 3;
                    if (l2_hints) {
                        st_row_hint(row, make_float4(Wm[0], Wm[1], Wm[2], Wm[3]), row_policy);
                        st_row_hint(row + 1, make_float4(Wm[4], Wm[5], Wm[6], Wm[7]), row_policy);
                        st_row_hint(row + 2, make_float4(Wm[8], Wm[9], Wm[10], Wm[11]), row_policy);
                    } else {
                        row[0] = make_float4(Wm[0], Wm[1], Wm[2], Wm[3]);
                        row[1] = make_float4(Wm[4], Wm[5], Wm[6], Wm[7]);
                        row[2] = make_float4(Wm[8], Wm[9], Wm[10], Wm[11]);
                    }
                }
                // outputs replace this thread's own (now dead) column of the stage; the TMA warp stores them
#pragma unroll
                for (int i = 0; i < 12; i++) st[i * T] = Wm[i];
                st[12 * T] = px, st[13 * T] = py, st[14 * T] = pz;
                if (kSet >= 1) {
                    st[15 * T] = vx, st[16 * T] = vy, st[17 * T] = vz;
                    st[18 * T] = rx, st[19 * T] = ry, st[20 * T] = rz, st[21 * T] = rw;
                }
            }
            if (kTeam) {
                { // this chunk's share of level l is done (rows written, or — last level — rows read)
                    if (pend_level != (int)l) flush_pending(lvl_done);
                    pend_level = (int)l;
                    pend_count += min(last, wc.ls[l + 1] - 1) - max(first, wc.ls[l]) + 1;
                    if ((pa.flags & kPipePublishPerChunk) && l + 1 < levels) flush_pending(lvl_done);
                }
            } else if (l < l_hi) {
                consumer_sync<T>(); // the next level of this chunk reads what this one wrote
            }
        }
        // prev_copy (kSet 2): this entity's matrix of the previous generation.  The twelve loads are issued HERE — the
        // matrix registers are dead, the chunk's new matrices sit in the stage — and consumed only after the cull
        // compaction below, which covers their L2 latency (they were the kernel's top stall: 34 % long_scoreboard).  The
        // bulk store that overwrites `world` cannot pass them: a warp arrives on done[s] only after its prev_copy stores.
        float old_w[kSet == 2 ? 12 : 1];
        const bool copy_prev = kSet == 2 && (pa.flags & kPipePrevCopy) && valid;
        if (kSet == 2 && copy_prev) {
#pragma unroll
            for (int i = 0; i < 12; i++) old_w[i] = __ldcg(world_old + (uint64_t)i * E + e);
        }
        fence_async_smem();
        const unsigned ballot = __ballot_sync(kFull, vis);
        const bool world_ends = c + G >= nc;
        if (kSet >= 1) { // min / max are exact in any order: warp REDUX on order-preserving keys, then shared atomics
            int key[6], vkey[6];
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const int klo = float_key(lo[r]), khi = float_key(hi[r]);
                key[r] = __reduce_min_sync(kFull, klo);
                key[3 + r] = __reduce_max_sync(kFull, khi);
                if (kSet == 2) { // vis_bounds: the same over the entities cull found visible
                    vkey[r] = __reduce_min_sync(kFull, vis ? klo : float_key(FIMO_B200_BOUNDS_EMPTY_MIN));
                    vkey[3 + r] = __reduce_max_sync(kFull, vis ? khi : float_key(-FIMO_B200_BOUNDS_EMPTY_MIN));
                }
            }
            if (lane < 6) {
                int mine = key[0], vmine = kSet == 2 ? vkey[0] : 0;
#pragma unroll
                for (int r = 1; r < 6; r++) {
                    mine = lane == r ? key[r] : mine;
                    if (kSet == 2) vmine = lane == r ? vkey[r] : vmine;
                }
                if (lane < 3) atomicMin(&s_bounds[kk & 1][lane], mine);
                else atomicMax(&s_bounds[kk & 1][lane], mine);
                if (kSet == 2) {
                    if (lane < 3) atomicMin(&s_vbounds[kk & 1][lane], vmine);
                    else atomicMax(&s_vbounds[kk & 1][lane], vmine);
                }
            }
        }
        unsigned ballot2 = 0, ballot_sleep = 0;
        if (kSet == 2) ballot2 = __ballot_sync(kFull, vis2), ballot_sleep = __ballot_sync(kFull, asleep);
        if (kTeam) {
            if (lane == 0) {
                mbar_arrive(&done[s]); // this warp's outputs are in the stage
                if ((e >> 5) < a.mask_words) mask_out[e >> 5] = ballot;
            }
            if (world_ends) flush_pending(lvl_done);
            if (world_ends && more_worlds) publish_const(s_c[(kk + 1) & 1], next_const);
            if (kSet >= 1 && world_ends) { // this CTA's share of the world's bounds joins the team's
                consumer_sync<T>();
                if (tid < 6) {
                    int *dst = a.bounds_keys + kStatSlots * 6 + w * 6 + tid;
                    if (tid < 3) atomicMin(dst, s_bounds[kk & 1][tid]);
                    else atomicMax(dst, s_bounds[kk & 1][tid]);
                    s_bounds[kk & 1][tid] = kEmptyKey;
                }
            }
        } else {
            // mask word + ordered compaction
            uint32_t *cnt = s_cnt[it & 1];
            if (lane == 0) {
                if (kSet != 2) mbar_arrive(&done[s]);
                if ((e >> 5) < a.mask_words) {
                    mask_out[e >> 5] = ballot;
                    if (kShadow) pa.rb_mask[w * a.vis_mask.world_words + (e >> 5)] = ballot;
                }
                if (kSet == 2) atomicAdd(&s_gen_mask[gen], (unsigned long long)ballot); // set 2 runs at the register cap: no per-thread 64-bit accumulators
                else warp_mask_sum += ballot;
                ((uint8_t *)cnt)[warp] = (uint8_t)__popc(ballot);
                if (kSet == 2) {
                    if ((e >> 5) < a.mask_words) mask2_out[e >> 5] = ballot2, sleep_out[e >> 5] = ballot_sleep;
                    atomicAdd(&s_gen_mask2[gen], (unsigned long long)ballot2);
                    ((uint8_t *)s_cnt2[it & 1])[warp] = (uint8_t)__popc(ballot2);
                }
            }
            if (world_ends && more_worlds) publish_const(s_c[(kk + 1) & 1], next_const);
            consumer_sync<T>();
            // the warps' counts are bytes: sums over "warps before me" and "all warps" are two dp4a per word
            uint32_t before = 0, chunk_total = 0;
#pragma unroll
            for (int j = 0; j < (kWarps + 3) / 4; j++) {
                const uint32_t word = cnt[j];
                const int nb = warp - 4 * j; // bytes of this word that belong to earlier warps
                const uint32_t below = nb >= 4 ? 0xFFFFFFFFu : (nb <= 0 ? 0u : (1u << (8 * nb)) - 1u);
                before = __dp4a(word & below, 0x01010101u, before);
                chunk_total = __dp4a(word, 0x01010101u, chunk_total);
            }
            if (vis) {
                const uint32_t at = running + before + __popc(ballot & ((1u << lane) - 1u));
                list_out[at] = e;
                if (kShadow) pa.rb_list[w * a.vis_list.world_words + at] = e;
            }
            running += chunk_total;
            if (kSet == 2) { // cull2's ordered list, same scan on its own count bytes
                const uint32_t *cnt2 = s_cnt2[it & 1];
                uint32_t before2 = 0, chunk_total2 = 0;
#pragma unroll
                for (int j = 0; j < (kWarps + 3) / 4; j++) {
                    const uint32_t word = cnt2[j];
                    const int nb = warp - 4 * j;
                    const uint32_t below = nb >= 4 ? 0xFFFFFFFFu : (nb <= 0 ? 0u : (1u << (8 * nb)) - 1u);
                    before2 = __dp4a(word & below, 0x01010101u, before2);
                    chunk_total2 = __dp4a(word, 0x01010101u, chunk_total2);
                }
                if (vis2) list2_out[running2 + before2 + __popc(ballot2 & ((1u << lane) - 1u))] = e;
                running2 += chunk_total2;
                if (copy_prev) {
#pragma unroll
                    for (int i = 0; i < 12; i++) prev_out[(uint64_t)i * E + e] = old_w[i];
                }
                __syncwarp(); // every lane's old matrix has been consumed: the stage (and `world`) may be drained
                if (lane == 0) mbar_arrive(&done[s]);
            }
        }
        if (world_ends) {
            if (!kTeam && tid == 0) {
                uptr(a.vis_count, w)[0] = running;
                if (kShadow) pa.rb_count[w * a.vis_count.world_words] = running;
                s_gen_vis[gen] += running;
                if (kSet == 2) uptr(a.vis_count2, w)[0] = running2, s_gen_vis2[gen] += running2;
            }
            if (!kTeam && lane == 0) { // mask words of this world visit join its generation's sum
                if (kSet != 2) atomicAdd(&s_gen_mask[gen], warp_mask_sum), warp_mask_sum = 0;
            }
            if (kSet >= 1 && !kTeam && tid < 6) { // after this chunk's barrier: every warp's keys are in
                const int key = s_bounds[kk & 1][tid];
                fptr(a.wbounds, w)[tid] = key_float(key);
                s_gen_bounds[gen][tid] = tid < 3 ? min(s_gen_bounds[gen][tid], key) : max(s_gen_bounds[gen][tid], key);
                s_bounds[kk & 1][tid] = kEmptyKey;
                if (kSet == 2) {
                    const int vkey = s_vbounds[kk & 1][tid];
                    fptr(a.vbounds, w)[tid] = key_float(vkey);
                    s_gen_vbounds[gen][tid] = tid < 3 ? min(s_gen_vbounds[gen][tid], vkey) : max(s_gen_vbounds[gen][tid], vkey);
                    s_vbounds[kk & 1][tid] = kEmptyKey;
                }
            }
            kk++;
        }
        cur.advance(pa, wk);
        if (++s == N) s = 0, parity ^= 1u;
        pre_have = pre_next;
    }

    if (kTimingBuild && pa.timing && tid == 0) {
        atomicAdd(pa.timing + 0, (unsigned long long)tm_input);
        atomicAdd(pa.timing + 1, (unsigned long long)tm_level);
        atomicAdd(pa.timing + 2, (unsigned long long)(clock64() - tm_begin));
        atomicAdd(pa.timing + 3, (unsigned long long)total);
        atomicAdd(pa.timing + 4, (unsigned long long)tm_pre);
        atomicAdd(pa.timing + 5, (unsigned long long)tm_gather);
        atomicAdd(pa.timing + 6, (unsigned long long)tm_prefetched);
        atomicMax(pa.timing + 7, (unsigned long long)(clock64() - tm_begin));
        if (pa.trace_base + blockIdx.x < kTimingCtas) { // per-CTA trace of the latest launch: where it ran, when, for how long
            unsigned long long wall1;
            unsigned smid;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(wall1));
            asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
            unsigned long long *row = pa.timing + 8 + kTimingRow * (size_t)(pa.trace_base + blockIdx.x);
            row[0] = smid, row[1] = tm_wall0, row[2] = wall1, row[3] = (unsigned long long)total, row[4] = tm_entry;
        }
    }
    if (!kTeam) { // publish every generation of the launch: batch bounds first, then warp 0 with one ticket for all of them
        consumer_sync<T>(); // the per-generation sums in shared memory are complete
        if (kSet >= 1 && tid < 6 && my_worlds) {
            for (uint32_t g = 0; g < pa.gens; g++) {
                if (tid < 3) atomicMin(a.bounds_keys + g * 6 + tid, s_gen_bounds[g][tid]);
                else atomicMax(a.bounds_keys + g * 6 + tid, s_gen_bounds[g][tid]);
                if (kSet == 2) {
                    if (tid < 3) atomicMin(a.bounds_keys2 + g * 6 + tid, s_gen_vbounds[g][tid]);
                    else atomicMax(a.bounds_keys2 + g * 6 + tid, s_gen_vbounds[g][tid]);
                }
            }
        }
        if (warp == 0) {
            const StatsSink sink{a.stats, a.scratch, a.gstats, a.bounds_keys, a.entities, a.ring, a.ring_slots, a.peers};
            if (kSet == 2)
                publish_stats_launch<true, true>(sink, StatsSink{a.stats2, a.scratch2, a.gstats2, a.bounds_keys2, a.entities}, s_gen_vis, s_gen_mask,
                                                 s_gen_vis2, s_gen_mask2, pa.gens, my_worlds, (unsigned long long)pa.stat_ctas, lane);
            else publish_stats_launch<kSet >= 1, false>(sink, sink, s_gen_vis, s_gen_mask, s_gen_vis, s_gen_mask, pa.gens, my_worlds,
                                                        (unsigned long long)pa.stat_ctas, lane);
        }
    }
    if (kTimingBuild && !kTeam && pa.timing && tid == 0 && pa.trace_base + blockIdx.x < kTimingCtas) { // statistics published: thread 0 leaves
        unsigned long long wall2;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(wall2));
        pa.timing[8 + kTimingRow * (size_t)(pa.trace_base + blockIdx.x) + 5] = wall2;
    }
}

// Team mode, second kernel: the ordered visible list, the per-world count and the batch statistics, derived
// from the mask alone (1 bit per entity in, 4 f_vis bytes per entity out — < 1 % of the step's traffic).
// One CTA per (segment of 256 mask words = 8192 entities, world); a segment's base offset is the popcount of
// the world's mask words before it (re-read through L2).  Segment 0 also re-arms the world's level counters.
constexpr int kCompactThreads = 256;
template <bool kGroupStats>
__global__ void __launch_bounds__(kCompactThreads) k_compact_lists(const StepArgs a, uint32_t *lvl_done_all) {
    __shared__ uint32_t s_warp[2][kCompactThreads / 32];
    __shared__ unsigned long long s_mask_sum;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t seg = blockIdx.x;
    const uint64_t w = blockIdx.y;
    const uint32_t *__restrict__ mask = uptr(a.vis_mask, w);
    uint32_t *__restrict__ list = uptr(a.vis_list, w);
    if (tid == 0) s_mask_sum = 0;
    if (seg == 0 && tid < kMaxLevels && lvl_done_all) lvl_done_all[w * kMaxLevels + tid] = 0;
    if (kGroupStats && seg == 0 && tid < 6) { // the team's bounds of this world: publish, fold into the batch's, re-arm
        const int empty = tid < 3 ? float_key(FIMO_B200_BOUNDS_EMPTY_MIN) : float_key(-FIMO_B200_BOUNDS_EMPTY_MIN);
        const int key = atomicExch(a.bounds_keys + kStatSlots * 6 + w * 6 + tid, empty);
        fptr(a.wbounds, w)[tid] = key_float(key);
        if (tid < 3) atomicMin(a.bounds_keys + tid, key);
        else atomicMax(a.bounds_keys + tid, key);
    }

    uint32_t before = 0; // visible entities of this world in earlier segments (this thread's share)
    for (uint32_t i = tid; i < seg * kCompactThreads; i += kCompactThreads) before += __popc(__ldcg(mask + i));
    const uint32_t wi = seg * kCompactThreads + tid;
    uint32_t word = wi < a.mask_words ? __ldcg(mask + wi) : 0u;
    const uint32_t mine = __popc(word);
    // warp-inclusive scan of `mine`, warp sum of `before`
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += v;
    }
    const uint32_t before_warp = __reduce_add_sync(kFull, before);
    if (lane == 31) s_warp[0][warp] = incl;
    if (lane == 0) s_warp[1][warp] = before_warp;
    __syncthreads();
    uint32_t all_before = 0, earlier = 0, seg_total = 0;
#pragma unroll
    for (int i = 0; i < kCompactThreads / 32; i++) {
        all_before += s_warp[1][i];
        earlier += i < warp ? s_warp[0][i] : 0u;
        seg_total += s_warp[0][i];
    }
    uint32_t off = all_before + earlier + incl - mine;
    const unsigned long long word64 = word;
    while (word) {
        const uint32_t b = __ffs(word) - 1;
        list[off++] = wi * 32u + b;
        word &= word - 1;
    }
    const unsigned long long warp_sum = __reduce_add_sync(kFull, (uint32_t)(word64 & 0xFFFFu)) +
                                        ((unsigned long long)__reduce_add_sync(kFull, (uint32_t)(word64 >> 16)) << 16);
    if (lane == 0) atomicAdd(&s_mask_sum, warp_sum);
    __syncthreads();
    if (tid == 0) {
        if (seg == gridDim.x - 1) uptr(a.vis_count, w)[0] = all_before + seg_total;
        publish_stats<kGroupStats>(a, seg_total, s_mask_sum, seg == 0 ? 1ull : 0ull, (unsigned long long)gridDim.x * gridDim.y);
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
namespace {

bool aligned16(const FimoB200_ResourceView *v) { return v && ((uintptr_t)v->device_ptr % 16) == 0; }

// Side stream and fork / join events of a plan's tail grid (WorldStepPlan::tail).
struct TailLaunch {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    ~TailLaunch() {
        if (stream) cudaStreamDestroy(stream);
        if (fork) cudaEventDestroy(fork);
        if (join) cudaEventDestroy(join);
    }
};

constexpr size_t kSmemPerSm = 227 * 1024;  // usable by CTAs on sm_100a
constexpr size_t kSmemPerCtaReserve = 1024; // the hardware reserves 1 KB per resident CTA
constexpr size_t static_smem(uint32_t set) { return set == 2 ? 2304 : set == 1 ? 1536 : 1152; } // WorldConst x2, counts, per-generation sums and bounds keys (see -Xptxas -v)

size_t ring_bytes(uint32_t set, int T, uint32_t stages) {
    return (size_t)stages * (set == 2 ? SetTraits<2>::kStage : set == 1 ? SetTraits<1>::kStage : SetTraits<0>::kStage) * T * 4;
}

// Largest index of an entity that can be a parent, over all worlds: max level_start[levels-1].  Read once
// from HBM when the executor is compiled (cold path) to size the shared-memory window.  If the host later
// uploads a different hierarchy the kernel stays correct — parents outside the window travel through the
// `rows` buffer, which always covers every entity — only the sizing is stale.
uint64_t max_parents(const WorldStepViews &v, int device) {
    const FimoB200_ResourceView *ls = v.level_start;
    const uint64_t E = v.world->count;
    if (!ls || !ls->device_ptr || ls->components < 2) return E;
    const size_t words = (size_t)ls->n_worlds * ls->components;
    std::vector<uint32_t> host(words);
    if (!hbm_download(device, host.data(), ls->device_ptr, words * 4)) return E;
    uint64_t m = 0;
    for (uint64_t w = 0; w < ls->n_worlds; w++) m = std::max<uint64_t>(m, host[w * ls->components + ls->components - 2]);
    return std::min(m, E);
}

bool make_plan(WorldStepPlan &p, int T, uint32_t blocks, uint32_t stages, uint64_t want_cap, uint64_t E) {
    const size_t budget = kSmemPerSm / blocks - kSmemPerCtaReserve - static_smem(p.set);
    const size_t fixed = ring_bytes(p.set, T, stages) + (size_t)stages * 16;
    if (fixed > budget) return false;
    uint64_t cap = std::min<uint64_t>({want_cap, E, (budget - fixed) / 48});
    p.threads = T, p.blocks_per_sm = blocks, p.stages = stages, p.lvl_cap = (uint32_t)cap;
    p.smem = ring_bytes(p.set, T, stages) + (size_t)cap * 48 + (size_t)stages * 16;
    return true;
}

// How many CTAs step one world together.  Few worlds (W below the CTA slots of the GPU) leave slots idle and
// serialise each world inside one CTA, so the slots are dealt out as `teams` x `team`: the split that keeps
// the most slots busy with an even number of worlds per team, ties to the smaller team.
void choose_team(WorldStepPlan &p, uint64_t W, uint64_t chunks, uint64_t slots) {
    p.team = 1, p.teams = 0;
    if (chunks < 2 || W * 2 > slots) return;
    double best = 0.0;
    for (uint64_t per = 1; per <= 8 && per <= W; per++) {
        const uint64_t teams = (W + per - 1) / per;
        const uint64_t team = std::min<uint64_t>(slots / teams, chunks);
        if (team < 1) continue;
        const double used = (double)(teams * team) / (double)slots * ((double)W / (double)(teams * per));
        if (used > best + 1e-9) best = used, p.team = (uint32_t)team, p.teams = (uint32_t)teams;
    }
    if (p.team <= 1) p.team = 1, p.teams = 0;
}

std::shared_ptr<void> hbm_shared(int device, size_t bytes) {
    void *ptr = hbm_alloc(device, bytes);
    if (!ptr) return nullptr;
    return std::shared_ptr<void>(ptr, [device](void *q) { hbm_free(device, q); });
}

} // namespace

bool world_step_plan(const WorldStepViews &v, int device, WorldStepPlan &plan) {
    plan = WorldStepPlan{};
    if (!v.world || !v.level_start) return false;
    const uint64_t E = v.world->count, W = v.world->n_worlds;
    if (E == 0 || E > 0x7FFFFFFFull) return false;
    bool pipe = E % 4 == 0; // the pipelined kernel needs 16-byte aligned plane rows (bulk copies)
    for (const FimoB200_ResourceView *r : {v.pos, v.vel, v.rot, v.scale, v.parent, v.aabb_c, v.aabb_h, v.world})
        if (!aligned16(r)) pipe = false;
    if (!pipe) { // ragged worlds: v1 (one CTA per world, loads through registers)
        if (v.set != 0 || !world_step_supported_v1(v)) return false; // v1 knows the 3-system chain only
        plan.threads = 0;
        return true;
    }
    const uint64_t parents = max_parents(v, device);
    plan.max_parents = (uint32_t)parents;
    plan.set = v.set;
    if (v.set == 2 && (!v.vel_clamp || !v.damp || !v.sleep_eps || !v.campos || !v.lod_dist2 || !v.frustum2 || !v.sleep || !v.lod ||
                       !v.vis_mask2 || !v.vis_list2 || !v.vis_count2 || !v.stats2 || !v.vbounds || !v.gstats2 || !v.scratch2))
        return false;
    if (v.set >= 1) {
        if (!v.gravity || !v.angvel || !v.waabb || !v.wbounds || !v.gstats || (v.vel_clamp && !v.vmax) || !aligned16(v.angvel)) return false;
        // bounds accumulators: batch [0..5] then per world, armed "empty" (min slots +FLT_MAX, max slots -FLT_MAX as keys)
        std::vector<int> keys(kStatSlots * 6 + 6 * W);
        const int32_t kMinKey = 0x7F7FFFFF, kMaxKey = (int32_t)(0xFF7FFFFFu ^ 0x7FFFFFFFu);
        for (size_t i = 0; i < keys.size(); i++) keys[i] = i % 6 < 3 ? kMinKey : kMaxKey;
        plan.bounds = hbm_shared(device, keys.size() * 4);
        if (!plan.bounds || !hbm_upload(device, plan.bounds.get(), keys.data(), keys.size() * 4)) return false;
        if (v.set == 2) {
            plan.bounds2 = hbm_shared(device, kStatSlots * 6 * 4);
            if (!plan.bounds2 || !hbm_upload(device, plan.bounds2.get(), keys.data(), kStatSlots * 6 * 4)) return false;
        }
    }
    bool planned = false;
    // FIMO_B200_PIPE=T,blocks,stages[,cap] overrides the heuristic (tuning sweeps; any value is correct)
    if (const char *env = std::getenv("FIMO_B200_PIPE")) {
        int T = 0;
        unsigned blocks = 0, stages = 0;
        unsigned long long cap = parents;
        if (std::sscanf(env, "%d,%u,%u,%llu", &T, &blocks, &stages, &cap) >= 3 && (T == 128 || T == 256 || T == 512) &&
            blocks >= 1 && blocks <= 6 && stages >= 2 && stages <= 16 && (v.set == 0 || blocks <= 4 || (v.set == 1 && blocks == 5 && T == 128)) &&
            make_plan(plan, T, blocks, stages, cap, E))
            planned = true;
        else
            std::fprintf(stderr, "fimo_worlds_b200: ignoring FIMO_B200_PIPE=%s\n", env);
    }
    // Heuristic (measured on B200, profiles/r1h_sweep.txt): parent matrices gathered from shared memory are what
    // keeps the kernel on the HBM roofline, so the plan that holds the largest share of the parents in its window
    // wins; among equals, more and smaller CTAs (they hide the level barriers best).  A 3-deep ring beats deeper
    // ones on every shape measured.  Parents outside the window travel through `rows`.
    double window_coverage = 1.0;
    if (!planned) {
        static const struct { int T; uint32_t blocks, stages; } kCandidates[] = {{128, 4, 3}, {256, 3, 3}, {256, 2, 3}, {512, 1, 3}};
        // ... but not at any price: a window that holds 90 % of the best coverage on offer with more CTAs per SM beats the
        // full window with fewer (4096-entity worlds: (256, 2) at 0.92 coverage 0.968 of the HBM peak, (512, 1) at 1.0
        // coverage 0.921; profiles/r2f_sweep.txt).
        double best = -1.0, chosen = -1.0;
        for (const auto &cand : kCandidates) {
            WorldStepPlan trial = plan;
            if (!make_plan(trial, cand.T, cand.blocks, cand.stages, parents, E)) continue;
            best = std::max(best, parents ? std::min(1.0, (double)trial.lvl_cap / (double)parents) : 1.0);
        }
        if (best < 0.0) return false;
        for (const auto &cand : kCandidates) { // in order of decreasing CTAs per SM
            WorldStepPlan trial = plan;
            if (!make_plan(trial, cand.T, cand.blocks, cand.stages, parents, E)) continue;
            const double coverage = parents ? std::min(1.0, (double)trial.lvl_cap / (double)parents) : 1.0;
            if (coverage >= 0.9 * best) {
                chosen = coverage, plan = trial;
                break;
            }
        }
        window_coverage = chosen;
    }
    // Few worlds (or a forced team size): CTA teams, with the geometry that suits them (no window needed).
    const char *team_var = std::getenv("FIMO_B200_TEAM"); // force the team size (1 = one CTA per world)
    bool team_env = team_var != nullptr;
    uint64_t forced = team_env ? std::strtoull(team_var, nullptr, 10) : 0;
    // Many LARGE worlds: with the window holding under 30 % of the parents a lone CTA gathers most rows from L2 with
    // nothing to overlap them (592 x 64k entities: 0.24 of the HBM peak); teams of 16 prefetch the rows one chunk ahead
    // and keep far fewer worlds' rows live in L2 (0.52, profiles/r1m_two_world_interleave_rejected.txt).
    if (!team_env && !planned && v.set != 2 && window_coverage < 0.3 && (E + 127) / 128 >= 32 && W * 2 > 4ull * sm_count(device))
        team_env = true, forced = 16;
    if (!planned && v.set != 2 && (!team_env || forced > 1)) {
        WorldStepPlan trial = plan;
        make_plan(trial, 128, 4, 3, 0, E);
        choose_team(trial, W, (E + 127) / 128, 4ull * sm_count(device));
        if (trial.team > 1 || forced > 1) plan = trial;
    }
    const uint64_t chunks = (E + plan.threads - 1) / plan.threads;
    const uint64_t slots = (uint64_t)plan.blocks_per_sm * sm_count(device);
    if (v.set == 2) { // the 16-system set runs one CTA per world only; few large worlds fall back to set 1 + teams
        if (!team_env) choose_team(plan, W, chunks, slots);
        else plan.team = (uint32_t)std::max<uint64_t>(forced, 1);
        if (plan.team > 1) return false;
        plan.team = 1, plan.teams = 0;
    } else if (!team_env) choose_team(plan, W, chunks, slots);
    else if (forced <= 1) plan.team = 1, plan.teams = 0;
    else plan.team = (uint32_t)std::min<uint64_t>({forced, chunks, slots}), plan.teams = (uint32_t)std::min<uint64_t>(W, slots / plan.team);
    if (plan.team <= 1) plan.team = 1, plan.teams = 0;
    if (plan.team > 1) { // no window: every parent row goes through L2
        plan.lvl_cap = 0;
        plan.stages = std::min<uint32_t>(plan.stages, kParSlots - 1); // the parent look-ahead ring runs one slot ahead of the stage ring
        plan.smem = ring_bytes(plan.set, plan.threads, plan.stages) + (size_t)plan.stages * 16 + (size_t)2 * plan.threads * 48 +
                    (size_t)kParSlots * plan.threads * 4 + (size_t)kParSlots * 8; // + parent-plane look-ahead ring
        plan.kernels = 2;
        plan.lvl_done = hbm_shared(device, (size_t)W * kMaxLevels * 4);
        if (!plan.lvl_done) return false;
    }
    if (plan.team == 1) {
        // Persistent grid with an even share of worlds per CTA: a ragged tail costs more than idle CTA slots on an
        // HBM-bound kernel.
        const uint64_t per_cta = std::max<uint64_t>((W + slots - 1) / slots, 1);
        plan.grid = plan.grid_all = (uint32_t)((W + per_cta - 1) / per_cta), plan.tail_first = (uint32_t)W;
        // CTAs do not all run at the same speed.  On the 8-system set whole groups of SMs (by %smid, the same ones every
        // launch) step their chunks 5-16 % slower than the rest, so with an even static share the kernel ends with its
        // slowest CTA, 1.23x the mean (profiles/r1zc_cta_trace.txt).  There the last round of worlds is not assigned
        // ahead: it is a second grid of one-world CTAs launched beside the persistent one, and the hardware block
        // scheduler hands them to whichever SM frees a slot first (4096 worlds 0.80 -> 0.84 of the HBM peak, 8192
        // worlds 0.82 -> 0.88).  World ids stay arithmetic on blockIdx and kernel parameters in both grids (uniform
        // datapath in the TMA thread; an atomic world queue inside the kernel is not, r1z_world_queue_experiment.txt).
        // The 3-system chain and the 16-system set show no slow SM groups — only a ~10 % spread between the CTAs of one
        // SM — and a one-world CTA's prologue and statistics epilogue cost them more than the tail returns (-1.5 % /
        // -0.5 %), so they keep one grid.  FIMO_B200_TAIL=0 disables, =n moves about n rounds to the tail for any set.
        uint64_t tail_rounds = plan.set == 1 ? 1 : 0;
        if (const char *env = std::getenv("FIMO_B200_TAIL")) tail_rounds = std::strtoull(env, nullptr, 10);
        const uint64_t rounds = W / slots; // whole rounds the batch fills
        if (tail_rounds && rounds >= 2) { // single-round batches: measured slower (one-world CTAs everywhere)
            // the remainder always goes to the tail; whole rounds join it until it holds at least tail_rounds - 1/2 rounds
            uint64_t fixed = rounds;
            while (fixed > 1 && (W - fixed * slots) * 2 < (2 * tail_rounds - 1) * slots) fixed--;
            if (W - fixed * slots > 0) {
                auto tail = std::make_shared<TailLaunch>();
                int previous = -1; // the calling thread may belong to anyone: create on the plan's device, restore
                if (cudaGetDevice(&previous) != cudaSuccess) previous = -1, (void)cudaGetLastError();
                (void)cudaSetDevice(resolve_device(device));
                if (cudaStreamCreateWithFlags(&tail->stream, cudaStreamNonBlocking) == cudaSuccess &&
                    cudaEventCreateWithFlags(&tail->fork, cudaEventDisableTiming) == cudaSuccess &&
                    cudaEventCreateWithFlags(&tail->join, cudaEventDisableTiming) == cudaSuccess) {
                    plan.grid = (uint32_t)slots, plan.tail_first = (uint32_t)(fixed * slots), plan.kernels = 2;
                    plan.tail = tail;
                } else cudaGetLastError();
                if (previous >= 0) (void)cudaSetDevice(previous);
            }
        }
    }
    // Multi-generation launches (PipeArgs::gens): every CTA — the tail grid's one-world CTAs included — must have at
    // least stages + 2 chunks per generation, so that a chunk's previous-generation stores are covered by the TMA
    // thread's "all but the two most recent bulk groups" wait.
    plan.max_gens = 1, plan.block_worlds = 0x40000000u; // one block: plain world order
    if (plan.team == 1 && plan.grid_all) { // multi-generation launches always use the one persistent grid (no tail grid)
        const uint64_t min_worlds = W / plan.grid_all;
        const uint64_t block = ((uint64_t)plan.stages + 2 + chunks - 1) / chunks; // fewest worlds that hold stages + 2 chunks
        if (min_worlds >= block) plan.max_gens = kStatSlots, plan.block_worlds = (uint32_t)block;
        // FIMO_B200_BLOCK=n: at least n worlds per block (A/B runs; 0 = all of a CTA's worlds, i.e. generation-major order)
        if (const char *env = std::getenv("FIMO_B200_BLOCK")) {
            const uint64_t want = std::strtoull(env, nullptr, 10);
            plan.block_worlds = want == 0 ? 0x40000000u : (uint32_t)std::min<uint64_t>(std::max<uint64_t>(want, block), 0x40000000u);
        }
    }
    if (const char *env = std::getenv("FIMO_B200_GENS")) plan.max_gens = std::min<uint32_t>(plan.max_gens, (uint32_t)std::max(1, std::atoi(env)));
    // parent rows beyond the window: per world; team mode: per TEAM (reused world after world, L2-resident)
    plan.rows = hbm_shared(device, (size_t)(plan.team > 1 ? plan.teams : W) * E * 48);
    return plan.rows != nullptr;
}

const char *world_step_kernel_name(const WorldStepPlan &plan) {
    if (plan.threads == 0) return "k_world_cta<256,true,true> (one CTA per world)";
    if (plan.team > 1) return "k_world_step_pipe<team> + k_compact_lists (TMA-fed persistent world-step, CTA teams per world)";
    return "k_world_step_pipe (TMA-fed persistent world-step)";
}

std::string world_step_describe(const WorldStepPlan &plan, uint64_t n_worlds) {
    char buf[320];
    if (plan.threads == 0) return world_step_kernel_name(plan);
    if (plan.team > 1)
        std::snprintf(buf, sizeof buf, "k_world_step_pipe<T=%d, %u CTAs/SM, team, set %u>: %u teams x %u CTAs, %u-deep TMA ring, parents through L2 rows; + k_compact_lists",
                      plan.threads, plan.blocks_per_sm, plan.set, plan.teams, plan.team, plan.stages);
    else if (plan.tail_first < n_worlds)
        std::snprintf(buf, sizeof buf, "k_world_step_pipe<T=%d, %u CTAs/SM, set %u>: persistent grid %u over worlds [0, %u) + tail grid of %llu one-world CTAs, "
                      "%u-deep TMA ring, window %u of %u parents",
                      plan.threads, plan.blocks_per_sm, plan.set, plan.grid, plan.tail_first, (unsigned long long)(n_worlds - plan.tail_first), plan.stages,
                      plan.lvl_cap, plan.max_parents);
    else
        std::snprintf(buf, sizeof buf, "k_world_step_pipe<T=%d, %u CTAs/SM, set %u>: persistent grid %u x %llu worlds per CTA, %u-deep TMA ring, window %u of %u parents",
                      plan.threads, plan.blocks_per_sm, plan.set, plan.grid, (unsigned long long)((n_worlds + plan.grid - 1) / (plan.grid ? plan.grid : 1)),
                      plan.stages, plan.lvl_cap, plan.max_parents);
    return buf;
}

namespace {
template <int T, int kBlocks, int kSet>
FimoStatus launch_pipe(PipeArgs &pa, const WorldStepPlan &plan, uint64_t W, const FimoB200_LaunchInfo *info) {
    cudaStream_t stream = (cudaStream_t)info->stream;
    if (pa.gens < 1 || pa.gens > plan.max_gens) return fail("world_step plan (generations per launch)");
    if (plan.team > 1) {
        if constexpr (kSet == 2) return fail("world_step plan (teams cannot run the 16-system set)");
        else {
        auto kernel = k_world_step_pipe<T, kBlocks, true, kSet>;
        if (!allow_smem(kernel, plan.smem)) return check_launch("world_step (shared memory opt-in)");
        pa.team = plan.team, pa.teams = plan.teams;
        pa.world_first = 0, pa.world_end = (uint32_t)W, pa.stat_ctas = plan.team * plan.teams, pa.trace_base = 0;
        kernel<<<plan.team * plan.teams, T + 32, plan.smem, stream>>>(pa);
        if (check_launch("world_step") != FIMO_STATUS_OK) return FIMO_STATUS_ERROR;
        const unsigned segments = (pa.a.mask_words + kCompactThreads - 1) / kCompactThreads;
        k_compact_lists<kSet == 1><<<dim3(segments, (unsigned)W), kCompactThreads, 0, stream>>>(pa.a, pa.lvl_done);
        return check_launch("compact_lists");
        }
    }
    const bool shadow = pa.rb_mask && pa.rb_list && pa.rb_count; // read-back generation: second copy of the cull outputs
    auto kernel = shadow ? k_world_step_pipe<T, kBlocks, false, kSet, true> : k_world_step_pipe<T, kBlocks, false, kSet, false>;
    if (!allow_smem(kernel, plan.smem)) return check_launch("world_step (shared memory opt-in)");
    // The tail grid evens out ONE generation's slowest CTAs; a multi-generation launch amortises that tail over its
    // generations anyway (dag8 4096 x 1024: 0.974 with, 0.975 without), so it uses the one persistent grid.
    const bool use_tail = plan.tail_first < W && pa.gens == 1;
    const unsigned grid = use_tail ? plan.grid : plan.grid_all, tail_ctas = use_tail ? (unsigned)(W - plan.tail_first) : 0u;
    pa.team = 1, pa.teams = grid;
    pa.world_first = 0, pa.world_end = use_tail ? plan.tail_first : (uint32_t)W, pa.stat_ctas = grid + tail_ctas, pa.trace_base = 0;
    if (!tail_ctas) {
        kernel<<<grid, T + 32, plan.smem, stream>>>(pa);
        return check_launch("world_step");
    }
    // fork: the persistent grid first (it takes every slot), the tail grid beside it on the plan's side stream; join.
    // Under stream capture the two become parallel kernel nodes of the generation's graph.
    const TailLaunch *tail = (const TailLaunch *)plan.tail.get();
    if (cudaEventRecord(tail->fork, stream) != cudaSuccess) return check_launch("world_step (fork)");
    kernel<<<grid, T + 32, plan.smem, stream>>>(pa);
    if (check_launch("world_step") != FIMO_STATUS_OK) return FIMO_STATUS_ERROR;
    if (cudaStreamWaitEvent(tail->stream, tail->fork, 0) != cudaSuccess) return check_launch("world_step (fork)");
    PipeArgs pt = pa;
    pt.teams = tail_ctas, pt.world_first = plan.tail_first, pt.world_end = (uint32_t)W, pt.trace_base = grid;
    kernel<<<tail_ctas, T + 32, plan.smem, tail->stream>>>(pt);
    if (check_launch("world_step (tail)") != FIMO_STATUS_OK) return FIMO_STATUS_ERROR;
    if (cudaEventRecord(tail->join, tail->stream) != cudaSuccess || cudaStreamWaitEvent(stream, tail->join, 0) != cudaSuccess)
        return check_launch("world_step (join)");
    return FIMO_STATUS_OK;
}
} // namespace

bool world_step_supports_shadow(const WorldStepPlan &plan) { return plan.threads != 0 && plan.team == 1; }

FimoStatus world_step_launch(const WorldStepViews &v, const WorldStepPlan &plan, void *cull_scratch,
                             const FimoB200_LaunchInfo *info, uint32_t gens, const CullShadow *shadow) {
    if (shadow && (!world_step_supports_shadow(plan) || gens != 1)) return fail("world_step plan (shadow cull outputs)");
    if (plan.threads == 0) return gens == 1 ? world_step_launch_v1(v, cull_scratch, info) : FIMO_STATUS_ERROR;
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    try {
        PipeArgs pa{};
        StepArgs &a = pa.a;
        a.n_worlds = (uint32_t)W, a.entities = (uint32_t)E;
        fill_propagate(a, v.world, v.pos, v.rot, v.scale, v.parent, v.level_start, W, E);
        fill_cull(a, v.vis_mask, v.vis_list, v.vis_count, v.stats, v.world, v.aabb_c, v.aabb_h, v.frustum, cull_scratch, W, E);
        a.ring = (unsigned long long *)v.stats_ring, a.ring_slots = v.stats_ring_slots, a.peers = (const PeerTable *)v.stats_peers;
        a.vel = plane_of(v.vel, FIMO_B200_SCOPE_ENTITY, 3, E, W, "vel");
        a.dt = plane_of(v.dt, FIMO_B200_SCOPE_WORLD, 1, 1, W, "dt");
        pa.stages = plan.stages;
        pa.gens = gens;
        pa.block_worlds = plan.block_worlds;
        pa.lvl_cap = plan.lvl_cap;
        pa.chunks = (uint32_t)((E + plan.threads - 1) / plan.threads);
        pa.rows = (float4 *)plan.rows.get();
        if (shadow) pa.rb_mask = shadow->mask, pa.rb_list = shadow->list, pa.rb_count = shadow->count;
        pa.lvl_done = (uint32_t *)plan.lvl_done.get();
        static unsigned long long *timing_buf = nullptr; // debug only (FIMO_B200_TIMING=1 with plain stream launches)
        static int timing_launches = 0;
        const bool timing = kTimingBuild && std::getenv("FIMO_B200_TIMING") != nullptr && !(info->flags & FIMO_B200_LAUNCH_FLAG_CAPTURING);
        if (timing && !timing_buf) timing_buf = (unsigned long long *)hbm_alloc(info->device, (8 + kTimingRow * (size_t)kTimingCtas) * sizeof(unsigned long long));
        pa.timing = timing ? timing_buf : nullptr;
        pa.flags = kPipePrefetchBlocking; // FIMO_B200_PREFETCH=0 / 1 (only if already landed) / 2 (default)
        if (const char *env = std::getenv("FIMO_B200_PREFETCH"))
            pa.flags = (pa.flags & ~kPipePrefetchBlocking) | (env[0] == '0' ? kPipeNoPrefetch : (env[0] == '2' ? kPipePrefetchBlocking : 0u));
        pa.poll_ns = 64;
        if (const char *env = std::getenv("FIMO_B200_POLL_NS")) pa.poll_ns = (uint32_t)std::min(std::max(std::atoi(env), 0), 100000);
        if (const char *env = std::getenv("FIMO_B200_L2_HINTS"))
            if (env[0] == '0') pa.flags |= kPipeNoL2Hints;
        if (const char *env = std::getenv("FIMO_B200_PUBLISH"))
            if (env[0] == 'c') pa.flags |= kPipePublishPerChunk;
        if (plan.set >= 1) {
            a.angvel = plane_of(v.angvel, FIMO_B200_SCOPE_ENTITY, 3, E, W, "angvel");
            a.gravity = plane_of(v.gravity, FIMO_B200_SCOPE_WORLD, 3, 1, W, "gravity");
            a.waabb = plane_of(v.waabb, FIMO_B200_SCOPE_ENTITY, 6, E, W, "waabb");
            a.wbounds = plane_of(v.wbounds, FIMO_B200_SCOPE_WORLD, 6, 1, W, "wbounds");
            if (v.vel_clamp) a.vmax = plane_of(v.vmax, FIMO_B200_SCOPE_WORLD, 1, 1, W, "vmax"), pa.flags |= kPipeVelClamp;
            if (!v.gstats->device_ptr || v.gstats->scope != FIMO_B200_SCOPE_BATCH || v.gstats->components != FIMO_B200_GSTAT_WORDS) throw Fail{"gstats"};
            a.gstats = (uint32_t *)v.gstats->device_ptr;
            a.bounds_keys = (int *)plan.bounds.get();
        }
        if (plan.set == 2) {
            const uint64_t words = (E + 31) / 32;
            a.damp = plane_of(v.damp, FIMO_B200_SCOPE_WORLD, 1, 1, W, "damp");
            a.sleep_eps = plane_of(v.sleep_eps, FIMO_B200_SCOPE_WORLD, 1, 1, W, "sleep_eps");
            a.campos = plane_of(v.campos, FIMO_B200_SCOPE_WORLD, 3, 1, W, "campos");
            a.lod_dist2 = plane_of(v.lod_dist2, FIMO_B200_SCOPE_WORLD, 3, 1, W, "lod_dist2");
            a.frustum2 = plane_of(v.frustum2, FIMO_B200_SCOPE_WORLD, 24, 1, W, "frustum2");
            a.sleep = plane_of(v.sleep, FIMO_B200_SCOPE_ENTITY_BITS, 1, words, W, "sleep");
            a.lod = plane_of(v.lod, FIMO_B200_SCOPE_ENTITY, 1, E, W, "lod");
            a.vis_mask2 = plane_of(v.vis_mask2, FIMO_B200_SCOPE_ENTITY_BITS, 1, words, W, "vis_mask2");
            a.vis_list2 = plane_of(v.vis_list2, FIMO_B200_SCOPE_ENTITY, 1, E, W, "vis_list2");
            a.vis_count2 = plane_of(v.vis_count2, FIMO_B200_SCOPE_WORLD, 1, 1, W, "vis_count2");
            a.vbounds = plane_of(v.vbounds, FIMO_B200_SCOPE_WORLD, 6, 1, W, "vbounds");
            if (!v.stats2->device_ptr || v.stats2->scope != FIMO_B200_SCOPE_BATCH || v.stats2->components != 2 * FIMO_B200_STAT_WORDS) throw Fail{"stats2"};
            if (!v.gstats2->device_ptr || v.gstats2->scope != FIMO_B200_SCOPE_BATCH || v.gstats2->components != FIMO_B200_GSTAT_WORDS) throw Fail{"gstats2"};
            a.stats2 = (unsigned long long *)v.stats2->device_ptr;
            a.scratch2 = (unsigned long long *)v.scratch2;
            a.gstats2 = (uint32_t *)v.gstats2->device_ptr;
            a.bounds_keys2 = (int *)plan.bounds2.get();
            if (v.prev_world) a.prev_world = plane_of(v.prev_world, FIMO_B200_SCOPE_ENTITY, 12, E, W, "prev_world"), pa.flags |= kPipePrevCopy;
        }
        const int key = plan.threads * 100 + (int)plan.blocks_per_sm * 10 + (int)plan.set;
        struct TimingReport { // prints the accumulated ticks every 16th launch, after the launch below has been enqueued
            bool on;
            int device;
            cudaStream_t stream;
            ~TimingReport() {
                if (!on || (++timing_launches % 16) != 0) return;
                cudaStreamSynchronize(stream);
                unsigned long long t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                hbm_download(device, t, timing_buf, sizeof t);
                if (const char *path = std::getenv("FIMO_B200_TIMING_DUMP")) { // per-CTA trace of the latest launch as CSV
                    std::vector<unsigned long long> rows(kTimingRow * (size_t)kTimingCtas);
                    hbm_download(device, rows.data(), timing_buf + 8, rows.size() * sizeof(unsigned long long));
                    if (FILE *f = std::fopen(path, "w")) {
                        std::fprintf(f, "cta,smid,start_ns,end_ns,chunks,entry_ns,exit_ns\n");
                        for (unsigned i = 0; i < grid_ctas && i < kTimingCtas; i++) {
                            const unsigned long long *r = rows.data() + kTimingRow * (size_t)i;
                            std::fprintf(f, "%u,%llu,%llu,%llu,%llu,%llu,%llu\n", i, r[0], r[1], r[2], r[3], r[4], r[5]);
                        }
                        std::fclose(f);
                    }
                }
                cudaMemsetAsync(timing_buf, 0, sizeof t, stream);
                const double loop = (double)(t[2] ? t[2] : 1), chunks = (double)(t[3] ? t[3] : 1);
                std::fprintf(stderr,
                             "fimo_worlds_b200 timing (16 launches): of the chunk loop: input wait %.1f%%, level wait %.1f%%, prefetch section %.1f%%, "
                             "row wait (thread 0) %.1f%%; %.0f ticks per chunk, %.0f%% of chunks prefetched; longest CTA loop %.0f ticks = %.2fx the mean\n",
                             100.0 * t[0] / loop, 100.0 * t[1] / loop, 100.0 * t[4] / loop, 100.0 * t[5] / loop, loop / chunks, 100.0 * t[6] / chunks,
                             (double)t[7], (double)t[7] * 16.0 * (double)grid_ctas / loop);
            }
            unsigned grid_ctas;
        } report{timing, info->device, (cudaStream_t)info->stream,
                 plan.team > 1 ? plan.team * plan.teams : plan.grid + (unsigned)(W - plan.tail_first)};
        switch (key) {
        case 12840: case 12830: case 12820: case 12810: return launch_pipe<128, 4, 0>(pa, plan, W, info);
        case 12850: return launch_pipe<128, 5, 0>(pa, plan, W, info);
        case 12860: return launch_pipe<128, 6, 0>(pa, plan, W, info);
        case 25610: case 25620: return launch_pipe<256, 2, 0>(pa, plan, W, info);
        case 25630: return launch_pipe<256, 3, 0>(pa, plan, W, info);
        case 51210: return launch_pipe<512, 1, 0>(pa, plan, W, info);
        case 12841: case 12831: case 12821: case 12811: return launch_pipe<128, 4, 1>(pa, plan, W, info);
        case 12851: return launch_pipe<128, 5, 1>(pa, plan, W, info);
        case 25611: case 25621: return launch_pipe<256, 2, 1>(pa, plan, W, info);
        case 25631: return launch_pipe<256, 3, 1>(pa, plan, W, info);
        case 51211: return launch_pipe<512, 1, 1>(pa, plan, W, info);
        case 12842: case 12832: case 12822: case 12812: return launch_pipe<128, 4, 2>(pa, plan, W, info);
        case 25612: case 25622: return launch_pipe<256, 2, 2>(pa, plan, W, info);
        case 25632: return launch_pipe<256, 3, 2>(pa, plan, W, info);
        case 51212: return launch_pipe<512, 1, 2>(pa, plan, W, info);
        default: return fail("world_step plan");
        }
    } catch (const Fail &f) {
        return fail(f.what);
    }
}

} // namespace fb::dev
