// cull_pack_static.cu — fimo_b200_cull_pack_static: the packed cull read-back for callers that must not allocate.
//
// fimo_b200_group_schedule_readback (device.cu: k_readback_scan_pack) packs the cull outputs into HBM staging slots the
// executor cudaMalloc's on first use and re-allocates when a step asks for more.  A frame loop that cannot afford the
// device-wide synchronisation of cudaMalloc / cudaFree, an engine that handed all of HBM to its own pool at start-up,
// or a host that bans driver allocations after initialisation cannot use that.  This file is the same pack with its
// staging area in a fixed array of the device image: nothing here calls cudaMalloc, cudaFree or any allocating
// library routine.  The reference has nothing like it (no device, no read-back).
//
// Everything that touches the array lives in THIS translation unit on purpose: the array has internal linkage, and in
// host code its name is only a handle for cudaGetSymbolAddress — handing `g_pack_area` itself to a kernel or a copy
// would pass a host-side shadow address, not HBM.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <mutex>

#include "rt.h"

namespace {

constexpr uint32_t kAreaWords = FIMO_B200_PACK_STATIC_MAX_BYTES / 4;
constexpr uint32_t kHeader = FIMO_B200_READBACK_HEADER_WORDS;
constexpr int kMaxDevices = 64;

// The staging area: FIMO_B200_PACK_STATIC_MAX_BYTES of HBM on every device that loads this image, until exit.
// Zero at load; afterwards it holds whatever the last call on that device packed.
static __device__ __align__(256) uint32_t g_pack_area[kAreaWords];

// One CTA per world.  CTA w adds up the counts of the worlds in front of it itself (n_worlds <= 8192: at most 64
// loads per thread, all L2 hits after the first CTA) — no scan kernel in front, no second pass.  `area` is the device
// address of g_pack_area; the list part [list_at, list_at + capacity) was zeroed in stream order before this launch.
__global__ void __launch_bounds__(128) k_cull_pack_static(const uint32_t *__restrict__ count, const uint32_t *__restrict__ mask,
                                                          const uint32_t *__restrict__ list, uint32_t n_worlds, uint32_t mask_words,
                                                          uint32_t entities, uint32_t capacity, unsigned long long generation,
                                                          uint32_t *__restrict__ area) {
    __shared__ uint32_t s_part[4];
    const uint32_t w = blockIdx.x, tid = threadIdx.x;
    if (n_worlds == 0) { // an empty batch still gets its header
        if (tid == 0) {
            area[0] = 0u, area[1] = 0u, area[2] = mask_words, area[3] = capacity, area[4] = 0u;
            area[5] = (uint32_t)(generation & 0xFFFFFFFFull), area[6] = (uint32_t)(generation >> 32), area[7] = entities;
        }
        return;
    }
    uint32_t before = 0;
    for (uint32_t i = tid; i < w; i += 128) before += min(__ldcg(count + i), entities); // a count can never exceed the row it indexes
    before = __reduce_add_sync(0xFFFFFFFFu, before);
    if ((tid & 31) == 0) s_part[tid >> 5] = before;
    __syncthreads();
    const uint32_t at = s_part[0] + s_part[1] + s_part[2] + s_part[3];
    const uint32_t n = min(__ldcg(count + w), entities);
    uint32_t *counts = area + kHeader, *offsets = counts + n_worlds;
    uint32_t *dst_mask = offsets + n_worlds + (uint64_t)w * mask_words;
    uint32_t *dst_list = offsets + n_worlds + (uint64_t)n_worlds * mask_words;
    if (tid == 0) {
        counts[w] = n, offsets[w] = at;
        if (w + 1 == n_worlds) {
            const uint32_t total = at + n;
            area[0] = total, area[1] = n_worlds, area[2] = mask_words, area[3] = capacity, area[4] = total > capacity ? 1u : 0u;
            area[5] = (uint32_t)(generation & 0xFFFFFFFFull), area[6] = (uint32_t)(generation >> 32), area[7] = entities;
        }
    }
    const uint32_t *src_mask = mask + (uint64_t)w * mask_words, *src_list = list + (uint64_t)w * entities;
    for (uint32_t i = tid; i < mask_words; i += 128) dst_mask[i] = src_mask[i];
    for (uint32_t i = tid; i < n; i += 128)
        if (at + i < capacity) dst_list[at + i] = src_list[i];
}

FimoStatus device_error(const char *what, cudaError_t err) {
    std::fprintf(stderr, "fimo_worlds_b200: fimo_b200_cull_pack_static: %s: %s\n", what, cudaGetErrorString(err));
    (void)cudaGetLastError();
    fb::set_last_error("DeviceError");
    return FIMO_STATUS_ERROR;
}

// Device address of g_pack_area on `device`, which must be current.  Looked up once per device: the first call on a
// device asks the runtime (cudaGetSymbolAddress — under lazy module loading this is also what loads the image and
// gives the array its HBM), later calls read the table.
std::mutex g_lookup_mutex;
uint32_t *g_area_of[kMaxDevices] = {};
cudaError_t area_on_current_device(int device, uint32_t **area) {
    std::lock_guard<std::mutex> lock(g_lookup_mutex);
    if (!g_area_of[device]) {
        uint32_t *ptr = nullptr;
        const cudaError_t err = cudaGetSymbolAddress((void **)&ptr, g_pack_area);
        if (err != cudaSuccess) return err;
        g_area_of[device] = ptr;
    }
    *area = g_area_of[device];
    return cudaSuccess;
}

} // namespace

extern "C" FimoStatus fimo_b200_cull_pack_static(const uint32_t *vis_count, const uint32_t *vis_mask, const uint32_t *vis_list,
                                                 uint64_t n_worlds, uint64_t entities, uint64_t mask_words, uint64_t generation,
                                                 void *out, uint64_t out_bytes, void *stream) {
    if (!out || (n_worlds && (!vis_count || (mask_words && !vis_mask) || (entities && !vis_list))) || entities > 0xFFFFFFFFull ||
        mask_words > 0xFFFFFFFFull) {
        fb::set_last_error("InvalidArgument");
        return FIMO_STATUS_ERROR;
    }
    // Sizes first, before anything is queued.  With at most 8192 worlds and 32-bit row lengths none of the products
    // below can overflow 64 bits.
    const auto too_large = [&](uint64_t packed_words) {
        std::fprintf(stderr,
                     "fimo_worlds_b200: fimo_b200_cull_pack_static: %llu worlds / %llu packed bytes do not fit the fixed staging area "
                     "(at most %u worlds and %u bytes per call): pass a smaller out_bytes or pack the batch in slices of worlds\n",
                     (unsigned long long)n_worlds, (unsigned long long)packed_words * 4ull, FIMO_B200_PACK_STATIC_MAX_WORLDS,
                     FIMO_B200_PACK_STATIC_MAX_BYTES);
        fb::set_last_error("TooLarge");
        return FIMO_STATUS_ERROR;
    };
    if (n_worlds > FIMO_B200_PACK_STATIC_MAX_WORLDS) return too_large(0);
    const uint64_t fixed_words = kHeader + 2 * n_worlds + n_worlds * mask_words;
    if (fixed_words > kAreaWords) return too_large(fixed_words);
    if (out_bytes / 4 < fixed_words) { // counts, offsets and masks are never cut: `out` must hold at least those
        fb::set_last_error("InvalidArgument");
        return FIMO_STATUS_ERROR;
    }
    const uint64_t room = out_bytes / 4 - fixed_words, all = n_worlds * entities;
    const uint64_t capacity = room < all ? room : all;
    if (capacity > kAreaWords - fixed_words) return too_large(fixed_words + capacity);
    int device = -1;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return device_error("cudaGetDevice", err);
    if (device < 0 || device >= kMaxDevices) {
        fb::set_last_error("InvalidArgument");
        return FIMO_STATUS_ERROR;
    }
    uint32_t *area = nullptr;
    if ((err = area_on_current_device(device, &area)) != cudaSuccess) return device_error("cudaGetSymbolAddress", err);

    cudaStream_t s = (cudaStream_t)stream;
    // The array still holds the previous call's lists.  The copy below sends `capacity` entries whatever the total
    // turns out to be, so the entries this call does not write must not be somebody else's: zero them first.
    if (capacity && (err = cudaMemsetAsync(area + fixed_words, 0, capacity * 4, s)) != cudaSuccess) return device_error("cudaMemsetAsync", err);
    k_cull_pack_static<<<(unsigned)(n_worlds ? n_worlds : 1), 128, 0, s>>>(vis_count, vis_mask, vis_list, (uint32_t)n_worlds, (uint32_t)mask_words,
                                                                         (uint32_t)entities, (uint32_t)capacity, generation, area);
    if ((err = cudaGetLastError()) != cudaSuccess) return device_error("launch of k_cull_pack_static", err);
    if ((err = cudaMemcpyAsync(out, area, (fixed_words + capacity) * 4, cudaMemcpyDefault, s)) != cudaSuccess) return device_error("cudaMemcpyAsync", err);
    return FIMO_STATUS_OK;
}
