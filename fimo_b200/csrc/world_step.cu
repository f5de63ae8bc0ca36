// world_step.cu — the fused world-step as a persistent, TMA-fed pipeline (sm_100a).
//
// What it replaces: one generation of the 3-system chain integrate -> transform propagate -> AABB
// frustum cull, which the reference dispatches as three fiber tasks per world through the pool's event
// loop (modules/fimo_worlds/src/SystemGroup.zig:454-487, modules/fimo_tasks/src/CommandBuffer.zig:150-201).
// Here ONE kernel steps every world of the batch; each live-in word is read from HBM once and each
// live-out word written once (B3 = 132.125 + 4 f_vis bytes per entity-step).
//
// Structure (HBM-bound streaming; no tensor cores):
//   * persistent CTAs (2 per SM with 256 threads, or 1 with 512), worlds strided over CTAs;
//   * a world is cut into chunks of T consecutive entities (entities are sorted by hierarchy level);
//   * the 18 input planes of a chunk (pos3 vel3 rot4 scale parent aabb_c3 aabb_h3 = 72 B/entity) are
//     brought into a shared-memory ring by 1-D bulk async copies (cp.async.bulk, the TMA engine,
//     SASS UBLKCP) that signal an mbarrier with complete_tx; the ring is several chunks deep and runs
//     across world boundaries, so HBM latency is hidden by the copy engine, not by occupancy;
//   * one thread owns one entity of the chunk and does integrate, local TRS, composition with the
//     parent's world matrix, the 6-plane AABB test — all in registers.  Only matrices of entities that
//     can still be parents (all levels but the last) are kept in shared memory ([12][cap] planes);
//     the levels present in a chunk are walked in order with a __syncthreads between them;
//   * cull compaction: __ballot_sync gives the 32-entity mask word, __popc prefix inside the warp and a
//     scan over the warps' counts give each visible entity its slot in the ASCENDING index list;
//   * outputs go straight from registers to HBM as coalesced 128-byte rows per plane.
#include "kernels_common.cuh"

namespace fb::dev {
namespace {

constexpr int kPlanes = 18; // pos 0-2, vel 3-5, rot 6-9, scale 10, parent 11, aabb_c 12-14, aabb_h 15-17

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity)
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

// 1-D bulk async copy shared -> global (TMA engine), tracked by the issuing thread's bulk async-group.
__device__ __forceinline__ void bulk_store(void *dst, const void *src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_addr(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// generic-proxy writes to shared memory become visible to the async proxy (the TMA store reads them)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
template <int kThreads> __device__ __forceinline__ void consumer_sync() { // the T compute threads only (barrier 1)
    asm volatile("bar.sync 1, %0;" ::"n"(kThreads) : "memory");
}

struct PipeArgs {
    StepArgs a;
    uint32_t stages;   // ring depth (chunks)
    uint32_t lvl_cap;  // parent matrices kept in shared memory: entities [0, lvl_cap) of a world
    uint32_t chunks;   // chunks per world = ceil(E / T)
};

// Source plane p of world w (floats / u32), see kPlanes.
__device__ __forceinline__ const uint32_t *plane_src(const StepArgs &a, int p, uint64_t w, uint32_t E) {
    if (p < 3) return uptr(a.pos, w) + (uint64_t)p * E;
    if (p < 6) return uptr(a.vel, w) + (uint64_t)(p - 3) * E;
    if (p < 10) return uptr(a.rot, w) + (uint64_t)(p - 6) * E;
    if (p == 10) return uptr(a.scale, w);
    if (p == 11) return uptr(a.parent, w);
    if (p < 15) return uptr(a.aabb_c, w) + (uint64_t)(p - 12) * E;
    return uptr(a.aabb_h, w) + (uint64_t)(p - 15) * E;
}

// Warp-specialised: warps [0, T/32) compute (one entity of the chunk per thread), the last warp drives the
// TMA engine — it loads chunk inputs into the ring and stores the finished chunk's outputs, which the
// compute threads leave in the same ring stage (world matrices over planes 0-11, new positions over 12-14).
template <int T, int kMinBlocks>
__global__ void __launch_bounds__(T + 32, kMinBlocks) k_world_step_pipe(const PipeArgs pa) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const StepArgs &a = pa.a;
    const uint32_t E = a.entities, N = pa.stages, cap = pa.lvl_cap, nc = pa.chunks;
    float *ring = (float *)smem_raw;                        // [N][kPlanes][T]
    float *s_lvl = ring + (size_t)N * kPlanes * T;          // [12][cap]
    uint64_t *full = (uint64_t *)(s_lvl + (size_t)12 * cap); // [N] inputs have landed     (cap is even => 8-byte aligned)
    uint64_t *done = full + N;                               // [N] outputs are in the stage
    __shared__ __align__(16) float s_frustum[24];
    __shared__ uint32_t s_ls[kMaxLevels + 1];
    __shared__ float s_dt;
    __shared__ __align__(16) uint32_t s_cnt[2][T / 32];
    __shared__ unsigned long long s_sum;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = T / 32;
    const uint32_t levels = a.levels;
    const uint32_t my_worlds = a.n_worlds > blockIdx.x ? (a.n_worlds - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const uint32_t total = my_worlds * nc; // chunk iterations of this CTA

    if (tid == 0) {
        for (uint32_t s = 0; s < N; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&done[s], kWarps);
        }
        s_sum = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == kWarps) {
        // ------------------------------- TMA warp -------------------------------------------------
        if (lane != 0) return;
        uint32_t k = 0, c = 0, s = 0, round = 0;       // position of the chunk being loaded
        uint32_t sk = 0, sc = 0;                        // position of the chunk being stored (N behind)
        auto store_chunk = [&](uint32_t stage, uint32_t wk, uint32_t wc) {
            const uint64_t w = blockIdx.x + (uint64_t)wk * gridDim.x;
            const uint32_t first = wc * T, bytes = min((uint32_t)T, E - first) * 4u;
            const float *st = ring + (size_t)stage * kPlanes * T;
            float *world = fptr(a.world, w) + first;
            float *pos = fptr(a.pos, w) + first;
#pragma unroll
            for (int i = 0; i < 12; i++) bulk_store(world + (uint64_t)i * E, st + i * T, bytes);
#pragma unroll
            for (int i = 0; i < 3; i++) bulk_store(pos + (uint64_t)i * E, st + (12 + i) * T, bytes);
            bulk_commit();
        };
        for (uint32_t it = 0; it < total; it++) {
            if (it >= N) { // the stage still holds chunk it-N: wait for its outputs, drain them, then reuse it
                mbar_wait(&done[s], (round - 1) & 1u);
                store_chunk(s, sk, sc);
                if (++sc == nc) sc = 0, sk++;
                bulk_wait_read_all();
            }
            const uint64_t w = blockIdx.x + (uint64_t)k * gridDim.x;
            const uint32_t first = c * T, bytes = min((uint32_t)T, E - first) * 4u; // multiple of 16 (E % 4 == 0)
            float *st = ring + (size_t)s * kPlanes * T;
            mbar_expect_tx(&full[s], bytes * kPlanes);
#pragma unroll
            for (int p = 0; p < kPlanes; p++) bulk_load(st + p * T, plane_src(a, p, w, E) + first, bytes, &full[s]);
            if (++c == nc) c = 0, k++;
            if (++s == N) s = 0, round++;
        }
        // drain the last min(N, total) chunks
        const uint32_t tail = total < N ? total : N;
        uint32_t ds = total < N ? 0 : s, dround = total < N ? 0 : round - 1;
        for (uint32_t j = 0; j < tail; j++) {
            mbar_wait(&done[ds], dround & 1u);
            store_chunk(ds, sk, sc);
            if (++sc == nc) sc = 0, sk++;
            if (++ds == N) ds = 0, dround++;
        }
        bulk_wait_all();
        return;
    }

    // ----------------------------------- compute warps ------------------------------------------------
    // per-world constants travel one world ahead in a register: frustum (24), dt (1), level_start (levels+1)
    auto const_word = [&](uint64_t w) -> uint32_t {
        if (tid < 24) return uptr(a.frustum, w)[tid];
        if (tid == 24) return uptr(a.dt, w)[0];
        if (tid >= 32 && tid <= 32 + (int)levels) return uptr(a.level_start, w)[tid - 32];
        return 0u;
    };
    uint32_t next_const = my_worlds ? const_word(blockIdx.x) : 0u;
    unsigned long long cta_visible = 0, warp_mask_sum = 0; // thread 0 / lane 0 of each warp
    uint32_t running = 0;                                   // visible entities of the current world so far
    uint32_t k = 0, c = 0, s = 0, parity = 0;
    uint64_t w = blockIdx.x;
    uint32_t *mask_out = nullptr, *list_out = nullptr;
    float *world_out = nullptr;

    for (uint32_t it = 0; it < total; it++) {
        if (c == 0) { // a new world begins: publish its constants, fetch the next world's
            if (tid < 24) s_frustum[tid] = __uint_as_float(next_const);
            if (tid == 24) s_dt = __uint_as_float(next_const);
            if (tid >= 32 && tid <= 32 + (int)levels) s_ls[tid - 32] = next_const;
            if (k + 1 < my_worlds) next_const = const_word(w + gridDim.x);
            running = 0;
            mask_out = uptr(a.vis_mask, w);
            list_out = uptr(a.vis_list, w);
            world_out = fptr(a.world, w);
            consumer_sync<T>();
        }
        const uint32_t first = c * T;
        const uint32_t e = first + tid;
        const bool valid = e < E;
        const uint32_t last = min(E, first + T) - 1;
        // hierarchy levels present in this chunk, and the level of this thread's entity
        uint32_t l_lo = 0, l_hi = 0, l_me = 0;
        for (uint32_t l = 1; l < levels; l++) {
            const uint32_t start = s_ls[l];
            l_lo += (first >= start);
            l_hi += (last >= start);
            l_me += (e >= start);
        }

        mbar_wait(&full[s], parity);
        float *st = ring + (size_t)s * kPlanes * T + tid;
        bool vis = false;
        for (uint32_t l = l_lo; l <= l_hi; l++) {
            if (valid && l_me == l) {
                const float dt = s_dt;
                const float px = fma1(st[3 * T], dt, st[0 * T]); // integrate
                const float py = fma1(st[4 * T], dt, st[1 * T]);
                const float pz = fma1(st[5 * T], dt, st[2 * T]);
                float L[12], Wm[12];
                local_trs(L, px, py, pz, st[6 * T], st[7 * T], st[8 * T], st[9 * T], st[10 * T]);
                const uint32_t p = __float_as_uint(st[11 * T]);
                if (p == kRoot) {
#pragma unroll
                    for (int i = 0; i < 12; i++) Wm[i] = L[i];
                } else {
                    float P[12];
                    if (p < cap) {
#pragma unroll
                        for (int i = 0; i < 12; i++) P[i] = s_lvl[i * cap + p];
                    } else { // beyond the shared-memory window: written to HBM by this CTA below, read back through L2
#pragma unroll
                        for (int i = 0; i < 12; i++) P[i] = __ldcg(world_out + (uint64_t)i * E + p);
                    }
                    compose(Wm, P, L);
                }
                vis = aabb_visible(Wm, st[12 * T], st[13 * T], st[14 * T], st[15 * T], st[16 * T], st[17 * T], s_frustum);
                if (l + 1 < levels) { // may be somebody's parent
                    if (e < cap) {
#pragma unroll
                        for (int i = 0; i < 12; i++) s_lvl[i * cap + e] = Wm[i];
                    } else {
#pragma unroll
                        for (int i = 0; i < 12; i++) world_out[(uint64_t)i * E + e] = Wm[i];
                    }
                }
                // outputs replace this thread's own (now dead) column of the stage; the TMA warp stores them
#pragma unroll
                for (int i = 0; i < 12; i++) st[i * T] = Wm[i];
                st[12 * T] = px, st[13 * T] = py, st[14 * T] = pz;
            }
            if (l < l_hi) consumer_sync<T>(); // the next level of this chunk reads what this one wrote
        }
        fence_async_smem();
        // mask word + ordered compaction
        const unsigned ballot = __ballot_sync(kFull, vis);
        uint32_t *cnt = s_cnt[it & 1];
        if (lane == 0) {
            mbar_arrive(&done[s]); // this warp's outputs are in the stage
            if ((e >> 5) < a.mask_words) mask_out[e >> 5] = ballot;
            warp_mask_sum += ballot;
            cnt[warp] = __popc(ballot);
        }
        consumer_sync<T>();
        uint32_t before = 0, chunk_total = 0;
#pragma unroll
        for (int i = 0; i < kWarps; i++) {
            const uint32_t n = cnt[i];
            before += (i < warp) ? n : 0u;
            chunk_total += n;
        }
        if (vis) list_out[running + before + __popc(ballot & ((1u << lane) - 1u))] = e;
        running += chunk_total;
        if (++c == nc) {
            if (tid == 0) {
                uptr(a.vis_count, w)[0] = running;
                cta_visible += running;
            }
            c = 0, k++, w += gridDim.x;
        }
        if (++s == N) s = 0, parity ^= 1u;
    }

    // fold the per-warp mask sums, publish the batch statistics
    if (lane == 0) atomicAdd(&s_sum, warp_mask_sum);
    consumer_sync<T>();
    if (tid == 0) publish_stats(a, cta_visible, s_sum, my_worlds);
}

} // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
namespace {

struct PipeConfig {
    int threads = 0;      // 256 (2 CTAs / SM) or 512 (1 CTA / SM); 0 = not applicable
    uint32_t stages = 0;
    uint32_t lvl_cap = 0;
    size_t smem = 0;
};

bool aligned16(const FimoB200_ResourceView *v) { return v && ((uintptr_t)v->device_ptr % 16) == 0; }

// The pipelined kernel needs 16-byte aligned plane rows (bulk copies): E % 4 == 0 and aligned bases.
PipeConfig pipe_config(const WorldStepViews &v) {
    PipeConfig cfg;
    if (!v.world || !v.level_start) return cfg;
    const uint64_t E = v.world->count;
    if (E == 0 || E % 4 != 0 || E > 0x7FFFFFFFull) return cfg;
    for (const FimoB200_ResourceView *r : {v.pos, v.vel, v.rot, v.scale, v.parent, v.aabb_c, v.aabb_h})
        if (!aligned16(r)) return cfg;
    // Entities that can be parents are [0, level_start[last]); the host does not read level_start (it
    // lives in HBM), so the window is sized for the worst case the budget allows and the kernel falls
    // back to L2 reads beyond it.
    const size_t kBudgetTwo = 110 * 1024, kBudgetOne = 220 * 1024; // per CTA, of the 227 KB / SM
    auto fit = [&](int T, size_t budget, uint32_t stages) {
        PipeConfig c;
        const size_t ring = (size_t)stages * kPlanes * T * 4 + stages * 16 + 1024 /* static + reserve */;
        if (ring >= budget) return c;
        uint64_t cap = (budget - ring) / 48;
        if (cap > E) cap = E;
        cap &= ~1ull;
        c.threads = T, c.stages = stages, c.lvl_cap = (uint32_t)cap;
        c.smem = (size_t)stages * kPlanes * T * 4 + (size_t)cap * 48 + stages * 16;
        return c;
    };
    if (E <= 2048) return fit(256, kBudgetTwo, 4);
    return fit(512, kBudgetOne, 4);
}

} // namespace

bool world_step_supported(const WorldStepViews &v) { return pipe_config(v).threads != 0 || world_step_supported_v1(v); }

const char *world_step_kernel_name(const WorldStepViews &v) {
    const PipeConfig cfg = pipe_config(v);
    if (cfg.threads == 256) return "k_world_step_pipe<256,2> (TMA-fed persistent world-step)";
    if (cfg.threads == 512) return "k_world_step_pipe<512,1> (TMA-fed persistent world-step)";
    return "k_world_cta<256,true,true> (one CTA per world)";
}

FimoStatus world_step_launch(const WorldStepViews &v, void *cull_scratch, const FimoB200_LaunchInfo *info) {
    const PipeConfig cfg = pipe_config(v);
    if (cfg.threads == 0) return world_step_launch_v1(v, cull_scratch, info);
    const uint64_t W = info->n_worlds, E = info->entities_per_world;
    if (W == 0 || E == 0) return FIMO_STATUS_OK;
    try {
        PipeArgs pa{};
        StepArgs &a = pa.a;
        a.n_worlds = (uint32_t)W, a.entities = (uint32_t)E;
        fill_propagate(a, v.world, v.pos, v.rot, v.scale, v.parent, v.level_start, W, E);
        fill_cull(a, v.vis_mask, v.vis_list, v.vis_count, v.stats, v.world, v.aabb_c, v.aabb_h, v.frustum, cull_scratch, W, E);
        a.vel = plane_of(v.vel, FIMO_B200_SCOPE_ENTITY, 3, E, W, "vel");
        a.dt = plane_of(v.dt, FIMO_B200_SCOPE_WORLD, 1, 1, W, "dt");
        pa.stages = cfg.stages;
        pa.lvl_cap = cfg.lvl_cap;
        pa.chunks = (uint32_t)((E + cfg.threads - 1) / cfg.threads);
        cudaStream_t stream = (cudaStream_t)info->stream;
        const int sms = sm_count(info->device);
        if (cfg.threads == 256) {
            auto kernel = k_world_step_pipe<256, 2>;
            if (!allow_smem(kernel, cfg.smem)) return check_launch("world_step (shared memory opt-in)");
            const unsigned grid = (unsigned)(W < 2ull * sms ? W : 2ull * sms);
            kernel<<<grid, 256 + 32, cfg.smem, stream>>>(pa);
        } else {
            auto kernel = k_world_step_pipe<512, 1>;
            if (!allow_smem(kernel, cfg.smem)) return check_launch("world_step (shared memory opt-in)");
            const unsigned grid = (unsigned)(W < (uint64_t)sms ? W : (uint64_t)sms);
            kernel<<<grid, 512 + 32, cfg.smem, stream>>>(pa);
        }
        return check_launch("world_step");
    } catch (const Fail &f) {
        return fail(f.what);
    }
}

} // namespace fb::dev
