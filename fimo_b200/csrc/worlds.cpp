// worlds.cpp — universe, world and system group of the B200-native `fimo-worlds`, and its 32 exports.
//
// Reference behaviour restated here (modules/fimo_worlds/src):
//   Universe.zig:131-381   resource / system registry: validation, link maps, cycle rejection
//   World.zig:39-216       resource instances, per-resource reader/writer lock, ordered multi-lock
//   SystemGroup.zig:58-158 membership with transitive weak inclusion and refcounts
//   SystemGroup.zig:160-338 DAG compile (longest-path rank, conflict edges)
//   SystemGroup.zig:430-487 schedule / run: one generation at a time, all group resources locked
//   SystemContext.zig:42-110 per-system instance, run trampoline, allocator strategies
//   FimoWorlds.zig:108-447 C-ABI shims and error reporting
// What is different by design:
//   * a world created with FimoB200_WorldBatchExt stands for N independent worlds on one GPU;
//     resources registered with FimoB200_ResourceDeviceExt live in HBM as SoA buffers and their blob
//     is a FimoB200_ResourceView;
//   * the compile step emits a true DAG (dev::Node list) instead of an in-order command buffer with
//     wait entries; device systems become a CUDA graph, a generation is one graph launch;
//   * generations of a device group are pipelined on the group's stream: `schedule` returns after the
//     enqueue, the completion thread performs the end-of-generation protocol;
//   * host systems still run on the pool through a command buffer, exactly as in the reference.
#include <algorithm>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <deque>
#include <memory>
#include <thread>
#include <unordered_map>
#include <unordered_set>

#include "device.h"
#include "rt.h"

namespace fb {
namespace {

std::string label_of(const char *label, size_t len) { return label && len ? std::string(label, len) : std::string("<unlabelled>"); }

void *aligned_new(size_t size, size_t alignment) {
    if (alignment < sizeof(void *)) alignment = sizeof(void *);
    void *p = nullptr;
    if (posix_memalign(&p, alignment, size ? size : 1) != 0) return nullptr;
    return p;
}

template <typename T> const T *find_ext(const void *next, uint32_t s_type) {
    for (const FimoB200_ExtHeader *h = (const FimoB200_ExtHeader *)next; h; h = (const FimoB200_ExtHeader *)h->next)
        if (h->s_type == s_type) return (const T *)h;
    return nullptr;
}

// Bulk-freed allocations (std.heap.ArenaAllocator stand-in) and tracked allocations
// (heap/TracingAllocator.zig: leftovers are released with the owner).
struct Bump {
    std::mutex m;
    std::vector<void *> blocks;
    void *alloc(size_t size, size_t alignment) {
        void *p = aligned_new(size, alignment);
        if (!p) return nullptr;
        std::lock_guard<std::mutex> g(m);
        blocks.push_back(p);
        return p;
    }
    void reset() {
        std::lock_guard<std::mutex> g(m);
        for (void *p : blocks) free(p);
        blocks.clear();
    }
    ~Bump() { reset(); }
};
struct Owned {
    std::mutex m;
    std::unordered_set<void *> live;
    void *alloc(size_t size, size_t alignment) {
        void *p = aligned_new(size, alignment);
        if (!p) return nullptr;
        std::lock_guard<std::mutex> g(m);
        live.insert(p);
        return p;
    }
    void release(void *p) {
        if (!p) return;
        {
            std::lock_guard<std::mutex> g(m);
            if (!live.erase(p)) return;
        }
        free(p);
    }
    ~Owned() {
        for (void *p : live) free(p);
    }
};

// ---------------------------------------------------------------------------------------------
// registry
// ---------------------------------------------------------------------------------------------
struct ResourceType { // Universe.zig:57-62 + the device extension
    std::string label;
    size_t size = 0, alignment = 1;
    std::atomic<size_t> references{0};
    bool on_device = false;
    uint32_t scope = 0, components = 0;
};

struct Edge { // Universe.zig:96-101
    bool implicit, weak, ignore_deferred;
};

struct SystemType;
using EdgeList = std::vector<std::pair<SystemType *, Edge>>; // insertion ordered, small

Edge *edge_to(EdgeList &l, SystemType *s) {
    for (auto &e : l)
        if (e.first == s) return &e.second;
    return nullptr;
}
void edge_erase(EdgeList &l, SystemType *s) {
    for (size_t i = 0; i < l.size(); i++)
        if (l[i].first == s) {
            l.erase(l.begin() + (ptrdiff_t)i);
            return;
        }
}

struct SystemType { // Universe.zig:64-106
    std::string label;
    std::vector<ResourceType *> exclusive, shared;
    std::atomic<size_t> references{0};
    RwLock links_lock;
    size_t declared_before = 0, declared_after = 0;
    EdgeList before; // systems that run BEFORE this one (this one depends on them)
    EdgeList after;  // systems that run AFTER this one
    void *factory = nullptr;
    void (*factory_destroy)(const void *) = nullptr;
    size_t system_size = 0, system_alignment = 1;
    bool (*system_create)(const void *, FimoWorldsMeta_SystemContext *, void *) = nullptr;
    void (*system_destroy)(void *) = nullptr;
    void (*system_run)(void *, void *const *, void *const *, FimoWorldsMeta_Fence *) = nullptr;
    FimoB200_device_entry device_entry = nullptr; // set => the system runs on the device
    uint32_t builtin_id = FIMO_B200_BUILTIN_NONE;
};

struct Registry {
    RwLock lock;
    std::unordered_set<ResourceType *> resources;
    std::unordered_set<SystemType *> systems;
    size_t worlds = 0;
};
Registry g_registry;

ResourceType *register_resource(const FimoWorldsMeta_ResourceDescriptor *d, const char **err) {
    // Universe.zig:131-150; `next` is where the reference reserves room for extensions (resources.h:17-18)
    auto r = std::make_unique<ResourceType>();
    r->label = label_of(d->label, d->label_len);
    r->size = d->size;
    r->alignment = d->alignment ? d->alignment : 1;
    if (const auto *ext = find_ext<FimoB200_ResourceDeviceExt>(d->next, FIMO_B200_STYPE_RESOURCE_DEVICE)) {
        if (d->size != sizeof(FimoB200_ResourceView) || ext->elem_bytes != 4 || ext->components == 0 ||
            ext->scope > FIMO_B200_SCOPE_BATCH)
            return *err = "InvalidDeviceResource", nullptr;
        r->on_device = true;
        r->scope = ext->scope;
        r->components = ext->components;
    }
    g_registry.lock.lock_write();
    ResourceType *raw = r.release();
    g_registry.resources.insert(raw);
    g_registry.lock.unlock_write();
    return raw;
}

void unregister_resource(ResourceType *r) {
    // Universe.zig:152-163
    g_registry.lock.lock_write();
    if (!g_registry.resources.erase(r)) panic("invalid resource");
    if (r->references.load(std::memory_order_acquire) != 0) panic("resource in use");
    g_registry.lock.unlock_write();
    delete r;
}

template <typename T> bool repeats_after(T *const *items, size_t from, size_t len, T *v) {
    for (size_t i = from; i < len; i++)
        if (items[i] == v) return true;
    return false;
}

SystemType *register_system(const FimoWorldsMeta_SystemDescriptor *d, const char **err) {
    // Universe.zig:165-334
    struct FactoryOwner { // the factory is ours from the first line on (errdefer, :166-169)
        const FimoWorldsMeta_SystemDescriptor *d;
        bool keep = false;
        ~FactoryOwner() {
            if (!keep && d->factory_destroy) d->factory_destroy(d->factory_size ? d->factory : nullptr);
        }
    } factory_owner{d};
    g_registry.lock.lock_write();
    struct Unlock {
        ~Unlock() { g_registry.lock.unlock_write(); }
    } unlock;

    ResourceType *const *ex = (ResourceType *const *)d->exclusive_handles;
    ResourceType *const *sh = (ResourceType *const *)d->shared_handles;
    const size_t n_ex = d->exclusive_handles_len, n_sh = d->shared_handles_len;
    for (size_t i = 0; i < n_ex; i++) { // :174-178
        if (!g_registry.resources.count(ex[i])) return *err = "NotFound", nullptr;
        if (repeats_after(ex, i + 1, n_ex, ex[i])) return *err = "Deadlock", nullptr;
    }
    for (size_t i = 0; i < n_sh; i++) { // :179-185
        if (!g_registry.resources.count(sh[i])) return *err = "NotFound", nullptr;
        if (repeats_after(ex, 0, n_ex, sh[i])) return *err = "Deadlock", nullptr;
        if (repeats_after(sh, i + 1, n_sh, sh[i])) return *err = "Duplicate", nullptr;
    }
    auto edge_of = [](const FimoWorldsMeta_SystemDependency &dep) {
        return Edge{false, (dep.flags & FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_WEAK) != 0,
                    (dep.flags & FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_IGNORE_DEFERRED) != 0};
    };
    EdgeList before, after;
    for (size_t i = 0; i < d->before_len; i++) { // :195-206
        SystemType *dep = (SystemType *)d->before[i].system;
        if (!g_registry.systems.count(dep)) return *err = "NotFound", nullptr;
        if (edge_to(before, dep)) return *err = "Duplicate", nullptr;
        before.emplace_back(dep, edge_of(d->before[i]));
    }
    for (size_t i = 0; i < d->after_len; i++) { // :207-219
        SystemType *dep = (SystemType *)d->after[i].system;
        if (!g_registry.systems.count(dep)) return *err = "NotFound", nullptr;
        if (edge_to(before, dep)) return *err = "Deadlock", nullptr;
        if (edge_to(after, dep)) return *err = "Duplicate", nullptr;
        after.emplace_back(dep, edge_of(d->after[i]));
    }
    { // :221-240 — following explicit successor links from our successors must never reach a predecessor
        std::vector<SystemType *> frontier, seen;
        for (auto &e : after) frontier.push_back(e.first), seen.push_back(e.first);
        while (!frontier.empty()) {
            SystemType *s = frontier.back();
            frontier.pop_back();
            if (edge_to(before, s)) return *err = "Deadlock", nullptr;
            s->links_lock.lock_read();
            for (auto &e : s->after)
                if (!e.second.implicit && std::find(seen.begin(), seen.end(), e.first) == seen.end()) {
                    seen.push_back(e.first);
                    frontier.push_back(e.first);
                }
            s->links_lock.unlock_read();
        }
    }

    SystemType *sys = new SystemType();
    sys->label = label_of(d->label, d->label_len);
    sys->exclusive.assign(ex, ex + n_ex);
    sys->shared.assign(sh, sh + n_sh);
    sys->declared_before = before.size();
    sys->declared_after = after.size();
    sys->before = before;
    sys->after = after;
    sys->factory_destroy = d->factory_destroy;
    sys->system_size = d->system_size;
    sys->system_alignment = d->system_alignment ? d->system_alignment : 1;
    sys->system_create = d->system_create;
    sys->system_destroy = d->system_destroy;
    sys->system_run = d->system_run;
    if (const auto *ext = find_ext<FimoB200_SystemDeviceExt>(d->next, FIMO_B200_STYPE_SYSTEM_DEVICE)) {
        sys->device_entry = ext->device_entry;
        sys->builtin_id = ext->builtin_id;
    }
    for (auto &e : before) { // implicit reverse links, :284-309
        e.first->links_lock.lock_write();
        e.first->after.emplace_back(sys, Edge{true, e.second.weak, e.second.ignore_deferred});
        e.first->links_lock.unlock_write();
    }
    for (auto &e : after) {
        e.first->links_lock.lock_write();
        e.first->before.emplace_back(sys, Edge{true, e.second.weak, e.second.ignore_deferred});
        e.first->links_lock.unlock_write();
    }
    if (d->factory_size != 0) { // the factory bytes are copied, :311-320
        sys->factory = aligned_new(d->factory_size, d->factory_alignment);
        std::memcpy(sys->factory, d->factory, d->factory_size);
    }
    g_registry.systems.insert(sys);
    for (ResourceType *r : sys->exclusive) r->references.fetch_add(1, std::memory_order_relaxed); // :323-330
    for (ResourceType *r : sys->shared) r->references.fetch_add(1, std::memory_order_relaxed);
    factory_owner.keep = true;
    return sys;
}

void unregister_system(SystemType *sys) {
    // Universe.zig:336-381
    g_registry.lock.lock_write();
    if (!g_registry.systems.erase(sys)) panic("invalid system");
    if (sys->before.size() != sys->declared_before || sys->after.size() != sys->declared_after) panic("system referenced");
    if (sys->references.load(std::memory_order_acquire) != 0) panic("system in use");
    if (sys->factory_destroy && sys->factory) sys->factory_destroy(sys->factory);
    for (ResourceType *r : sys->exclusive) r->references.fetch_sub(1, std::memory_order_relaxed);
    for (ResourceType *r : sys->shared) r->references.fetch_sub(1, std::memory_order_relaxed);
    for (auto &e : sys->before) {
        if (e.second.implicit) panic("system referenced");
        e.first->links_lock.lock_write();
        edge_erase(e.first->after, sys);
        e.first->links_lock.unlock_write();
    }
    for (auto &e : sys->after) {
        if (e.second.implicit) panic("system referenced");
        e.first->links_lock.lock_write();
        edge_erase(e.first->before, sys);
        e.first->links_lock.unlock_write();
    }
    g_registry.lock.unlock_write();
    free(sys->factory);
    delete sys;
}

// ---------------------------------------------------------------------------------------------
// world: resource store
// ---------------------------------------------------------------------------------------------
struct Instance { // World.zig:28-32
    RwLock lock;
    void *blob = nullptr; // host blob; for a device resource: the FimoB200_ResourceView
    std::atomic<size_t> references{0};
    size_t hbm_bytes = 0;
    void *mirror = nullptr; // pinned host copy of a device resource, made when a host system of a device group uses it
};

struct World {
    RwLock lock;
    std::string label;
    FimoTasksMeta_Pool executor{};
    bool inherited_executor = false;
    Owned allocator;
    std::atomic<size_t> groups{0};
    std::unordered_map<ResourceType *, Instance *> resources;
    std::atomic<uint64_t> resource_epoch{0}; // bumped on add/remove: compiled executors hold blob pointers
    // batch extension
    uint64_t n_worlds = 1, entities = 0;
    int device = -1; // >= 0: the batch lives on that GPU
};

World *world_create(const FimoWorldsMeta_WorldDescriptor *d, const char **err) {
    // World.zig:39-71
    auto w = std::make_unique<World>();
    w->label = label_of(d->label, d->label_len);
    if (const auto *ext = find_ext<FimoB200_WorldBatchExt>(d->next, FIMO_B200_STYPE_WORLD_BATCH)) {
        if (!dev::device_ok(ext->device)) return *err = "NoDevice", nullptr; // no CPU fallback
        w->n_worlds = ext->n_worlds;
        w->entities = ext->entities_per_world;
        w->device = dev::resolve_device(ext->device);
    }
    if (d->pool) {
        w->executor = *d->pool;
        w->executor.vtable->acquire(w->executor.data);
        w->inherited_executor = true;
    } else { // one pool per world (World.zig:45-52); a GPU batch needs only a few host threads
        const std::string l = "world `" + w->label + "` executor";
        w->executor = pool_handle(pool_create(l.data(), l.size(), w->device >= 0 ? 2 : 0));
    }
    g_registry.lock.lock_write();
    g_registry.worlds++;
    g_registry.lock.unlock_write();
    return w.release();
}

void world_destroy(World *w) {
    // World.zig:71-82
    if (w->groups.load(std::memory_order_acquire) != 0 || !w->resources.empty()) panic("world not empty");
    if (!w->inherited_executor) w->executor.vtable->request_close(w->executor.data);
    w->executor.vtable->release(w->executor.data);
    g_registry.lock.lock_write();
    g_registry.worlds--;
    g_registry.lock.unlock_write();
    delete w;
}

uint64_t scope_count(const World *w, uint32_t scope) {
    switch (scope) {
        case FIMO_B200_SCOPE_ENTITY: return w->entities;
        case FIMO_B200_SCOPE_ENTITY_BITS: return (w->entities + 31) / 32;
        default: return 1;
    }
}

const char *world_add_resource(World *w, ResourceType *type, const void *value) {
    // World.zig:84-112.  Host resource: the blob is copied.  Device resource: the blob is a view; the
    // data is uploaded into (or adopted as) an HBM buffer laid out [n_worlds][components][count].
    g_registry.lock.lock_read();
    if (!g_registry.resources.count(type)) panic("invalid resource");
    type->references.fetch_add(1, std::memory_order_relaxed);
    g_registry.lock.unlock_read();

    auto inst = std::make_unique<Instance>();
    inst->blob = aligned_new(type->size, type->alignment);
    std::memcpy(inst->blob, value, type->size);
    const char *err = nullptr;
    bool owns_hbm = false;
    if (type->on_device) {
        FimoB200_ResourceView *view = (FimoB200_ResourceView *)inst->blob;
        if (w->device < 0) err = "NoDevice"; // a device resource needs a batch world
        else {
            const void *host_src = view->host_ptr;
            view->n_worlds = type->scope == FIMO_B200_SCOPE_BATCH ? 1 : w->n_worlds;
            view->count = scope_count(w, type->scope);
            view->components = type->components;
            view->elem_bytes = 4;
            view->comp_stride = 4 * view->count;
            view->world_stride = 4ull * view->components * view->count;
            view->scope = type->scope;
            view->host_ptr = nullptr;
            inst->hbm_bytes = view->world_stride * view->n_worlds;
            if (view->device_ptr) {
                view->flags |= FIMO_B200_VIEW_FLAG_EXTERNAL; // adopted: the caller keeps ownership
            } else {
                view->flags &= ~FIMO_B200_VIEW_FLAG_EXTERNAL;
                view->device_ptr = dev::hbm_alloc(w->device, inst->hbm_bytes);
                owns_hbm = view->device_ptr != nullptr;
                if (!view->device_ptr) err = "OutOfDeviceMemory";
            }
            if (!err && host_src && inst->hbm_bytes &&
                !dev::hbm_upload(w->device, view->device_ptr, host_src, inst->hbm_bytes))
                err = "DeviceError";
        }
    }
    if (!err) {
        w->lock.lock_write();
        if (w->resources.count(type)) err = "Duplicate"; // World.zig:109
        else {
            w->resources.emplace(type, inst.get());
            w->resource_epoch.fetch_add(1, std::memory_order_release);
        }
        w->lock.unlock_write();
    }
    if (err) {
        if (owns_hbm) dev::hbm_free(w->device, ((FimoB200_ResourceView *)inst->blob)->device_ptr);
        free(inst->blob);
        type->references.fetch_sub(1, std::memory_order_relaxed);
        return err;
    }
    inst.release();
    return nullptr;
}

const char *world_remove_resource(World *w, ResourceType *type, void *value) {
    // World.zig:114-137: the blob is copied back out; a device resource is first drained to
    // `value->host_ptr` (if given) — this is also the checkpoint path.
    w->lock.lock_write();
    auto it = w->resources.find(type);
    if (it == w->resources.end()) panic("invalid resource");
    Instance *inst = it->second;
    if (inst->references.load(std::memory_order_acquire) != 0) {
        w->lock.unlock_write();
        return "InUse";
    }
    w->resources.erase(it);
    w->resource_epoch.fetch_add(1, std::memory_order_release);
    w->lock.unlock_write();
    const char *err = nullptr;
    if (type->on_device) {
        FimoB200_ResourceView *view = (FimoB200_ResourceView *)inst->blob;
        void *host_dst = ((FimoB200_ResourceView *)value)->host_ptr;
        if (host_dst && inst->hbm_bytes && !dev::hbm_download(w->device, host_dst, view->device_ptr, inst->hbm_bytes))
            err = "DeviceError";
        if (!(view->flags & FIMO_B200_VIEW_FLAG_EXTERNAL)) dev::hbm_free(w->device, view->device_ptr);
        if (inst->mirror) dev::host_free_pinned(inst->mirror);
        view->host_ptr = host_dst;
    }
    std::memcpy(value, inst->blob, type->size);
    free(inst->blob);
    delete inst;
    type->references.fetch_sub(1, std::memory_order_relaxed);
    return err;
}

struct LockRequest {
    size_t slot;
    ResourceType *type;
    Instance *inst;
    bool exclusive;
};

std::vector<LockRequest> collect_locks(World *w, ResourceType *const *ex, size_t n_ex, ResourceType *const *sh, size_t n_sh) {
    // World.zig:145-189: validate and pin, then order by handle address (worlds.h:67).
    std::vector<LockRequest> reqs;
    reqs.reserve(n_ex + n_sh);
    w->lock.lock_read();
    for (size_t i = 0; i < n_ex; i++) {
        if (repeats_after(ex, i + 1, n_ex, ex[i])) panic("deadlock");
        auto it = w->resources.find(ex[i]);
        if (it == w->resources.end()) panic("invalid resource");
        it->second->references.fetch_add(1, std::memory_order_relaxed);
        reqs.push_back({i, ex[i], it->second, true});
    }
    for (size_t i = 0; i < n_sh; i++) {
        if (repeats_after(sh, i + 1, n_sh, sh[i])) panic("duplicate");
        if (repeats_after(ex, 0, n_ex, sh[i])) panic("deadlock");
        auto it = w->resources.find(sh[i]);
        if (it == w->resources.end()) panic("invalid resource");
        it->second->references.fetch_add(1, std::memory_order_relaxed);
        reqs.push_back({n_ex + i, sh[i], it->second, false});
    }
    w->lock.unlock_read();
    std::stable_sort(reqs.begin(), reqs.end(),
                     [](const LockRequest &a, const LockRequest &b) { return (uintptr_t)a.type < (uintptr_t)b.type; });
    return reqs;
}

void world_lock_resources(World *w, ResourceType *const *ex, size_t n_ex, ResourceType *const *sh, size_t n_sh, void **out) {
    // World.zig:190-198
    for (const LockRequest &r : collect_locks(w, ex, n_ex, sh, n_sh)) {
        if (r.exclusive) r.inst->lock.lock_write();
        else r.inst->lock.lock_read();
        out[r.slot] = r.inst->blob;
    }
}

// All-or-nothing variant for the non-blocking schedule path.
bool world_try_lock_resources(World *w, ResourceType *const *ex, size_t n_ex, void **out) {
    std::vector<LockRequest> reqs = collect_locks(w, ex, n_ex, nullptr, 0);
    for (size_t i = 0; i < reqs.size(); i++) {
        if (reqs[i].inst->lock.try_lock_write()) {
            out[reqs[i].slot] = reqs[i].inst->blob;
            continue;
        }
        for (size_t j = 0; j < i; j++) reqs[j].inst->lock.unlock_write();
        for (const LockRequest &r : reqs) r.inst->references.fetch_sub(1, std::memory_order_relaxed);
        return false;
    }
    return true;
}

void world_unlock_resource(World *w, ResourceType *type, bool exclusive) {
    // World.zig:200-216
    w->lock.lock_read();
    auto it = w->resources.find(type);
    if (it == w->resources.end()) panic("invalid resource");
    Instance *inst = it->second;
    w->lock.unlock_read();
    inst->references.fetch_sub(1, std::memory_order_relaxed);
    if (exclusive) inst->lock.unlock_write();
    else inst->lock.unlock_read();
}

// ---------------------------------------------------------------------------------------------
// system group
// ---------------------------------------------------------------------------------------------
struct Group;

struct Member { // SystemContext.zig:8-40
    Group *group = nullptr;
    SystemType *type = nullptr;
    bool weak = false;
    size_t dependants = 0; // reference_count in the reference
    void *state = nullptr;
    bool merge_deferred = false;
    FimoWorldsMeta_Fence deferred_fence{FIMO_WORLDS_META_FENCE_UNSIGNALED};
    std::vector<Member *> deferred_deps;
    std::vector<void *> blobs; // exclusive first, then shared (SystemGroup.zig:345-350)
    Bump transient;
    Owned persistent;
    std::vector<FimoWorldsMeta_Fence *> removal_waiters;
    FimoTasksMeta_Task task{};
    size_t rank = 0, order = 0;
};

Member *member_create(Group *g, SystemType *type, bool weak) {
    // SystemContext.zig:42-74
    Member *m = new Member();
    m->group = g;
    m->type = type;
    m->weak = weak;
    m->state = aligned_new(type->system_size, type->system_alignment);
    type->references.fetch_add(1, std::memory_order_relaxed);
    if (!type->system_create(type->factory, (FimoWorldsMeta_SystemContext *)m, m->state)) {
        type->references.fetch_sub(1, std::memory_order_relaxed);
        free(m->state);
        delete m;
        return nullptr;
    }
    return m;
}

void member_destroy(Member *m) {
    // SystemContext.zig:76-89
    if (m->type->system_destroy) m->type->system_destroy(m->state);
    m->type->references.fetch_sub(1, std::memory_order_relaxed);
    for (FimoWorldsMeta_Fence *f : m->removal_waiters) fence_signal(f);
    free(m->state);
    delete m;
}

void member_run_on_host(Member *m) {
    // SystemContext.zig:91-100
    FimoWorldsMeta_fence_reset(&m->deferred_fence);
    for (Member *dep : m->deferred_deps) fence_wait(&dep->deferred_fence);
    void *const *ex = m->blobs.data();
    void *const *sh = m->blobs.data() + m->type->exclusive.size();
    m->type->system_run(m->state, ex, sh, &m->deferred_fence);
    m->transient.reset();
    if (m->merge_deferred) fence_wait(&m->deferred_fence);
}

void member_task_start(FimoTasksMeta_Task *task) { // TaskInfo.taskStart, SystemGroup.zig:203-207
    member_run_on_host((Member *)((char *)task - offsetof(Member, task)));
}

struct Job { // one scheduled generation (Job.zig:216-264 wraps SystemGroup.run the same way)
    Group *group;
    size_t generation;
    std::vector<FimoWorldsMeta_Fence *> wait_on;
    FimoWorldsMeta_Fence *signal;
    // A stream-ordered transfer of one group resource instead of a generation (fimo_b200_group_transfer):
    // queued and ordered exactly like a generation, but it does not consume a generation number.
    ResourceType *xfer_resource = nullptr;
    void *xfer_host = nullptr;
    bool xfer_to_device = false;
    // fimo_b200_group_schedule_io: a generation with its own input copies in front and result copies behind
    std::vector<std::pair<ResourceType *, void *>> io_up, io_down;
    // fimo_b200_group_schedule_readback: the packed cull outputs of this generation go to host memory
    bool has_readback = false;
    FimoB200_CullReadback readback{};
};

struct Group {
    std::string label;
    World *world = nullptr;
    FimoTasksMeta_Pool executor{};
    std::atomic<size_t> generation{0};
    uint32_t ext_flags = 0;
    FimoB200_post_step post_step = nullptr;
    void *post_step_user = nullptr;
    void *stats_ring = nullptr; // FimoB200_GroupDeviceExt::stats_ring
    uint32_t stats_ring_slots = 0;
    dev::StatsExchange *stats_x = nullptr; // multi-GPU statistics exchange (fimo_b200_group_stats_*_bind)

    std::mutex m; // the graph mutex (SystemGroup.zig:40-52)
    bool dirty = true;
    size_t next_generation = 0;
    FimoWorldsMeta_TimelineSemaphore schedule_semaphore{0};
    std::vector<std::pair<SystemType *, Member *>> members; // insertion ordered
    std::vector<Member *> retired;                          // deinit_list
    std::vector<Member *> order;                            // compiled: rank-sorted members
    std::vector<ResourceType *> resources;                  // union of the members' resources
    std::vector<void *> resource_blobs;
    std::vector<FimoTasksMeta_CommandBufferEntry> entries; // host path (SystemGroup.zig:264-336)
    std::string systems_label;
    Bump single_generation;
    Bump multi_generation[4];

    // device path
    bool has_device_systems = false, has_host_systems = false;
    dev::Executor *executor_dev = nullptr;
    FimoB200_ExecutorInfo info{};
    std::string plan_text; // what the DAG compiler produced, in words (fimo_b200_group_executor_describe)
    size_t in_flight = 0; // device generations enqueued and not yet completed (locks are held while > 0)
    // Multi-generation launches: when the executor is one fused world-step node that can step several generations per
    // launch (dev::executor_max_gens) and two launches are already on the stream, further plain generations wait here
    // and leave as ONE launch when a launch retires — fewer launch boundaries, same results, generation numbers and
    // fences still per generation (SystemGroup.zig:430-452,483-486).
    std::deque<Job *> batch_pending;
    size_t launches_in_flight = 0;
    uint64_t compiled_epoch = 0; // world->resource_epoch the executor was compiled against

    // deferred path: generations that could not be started from inside `schedule`
    std::deque<Job *> backlog;
    std::condition_variable backlog_cv;
    std::thread dispatcher;
    bool dispatcher_running = false, dispatcher_busy = false, stop_dispatcher = false;

    Member *find(SystemType *t) {
        for (auto &p : members)
            if (p.first == t) return p.second;
        return nullptr;
    }
    void erase(SystemType *t) {
        for (size_t i = 0; i < members.size(); i++)
            if (members[i].first == t) {
                members.erase(members.begin() + (ptrdiff_t)i);
                return;
            }
    }
    bool generation_running() const {
        return !FimoWorldsMeta_timeline_semaphore_is_signaled(&schedule_semaphore, (uint64_t)next_generation);
    }
};

void drain_retired(Group *g) { // SystemGroup.zig:54-59
    for (Member *m : g->retired) member_destroy(m);
    g->retired.clear();
}

void graph_remove(Group *g, SystemType *type, FimoWorldsMeta_Fence *fence, bool may_defer);

Member *graph_add(Group *g, SystemType *type, bool weak, const char **err) {
    // SystemGroup.zig:61-115: strong dependencies come along as weak, refcounted members
    if (Member *existing = g->find(type)) {
        if (!weak) existing->weak = false;
        return existing;
    }
    type->links_lock.lock_read();
    std::vector<Member *> pinned;
    auto undo = [&] {
        for (Member *m : pinned)
            if (--m->dependants == 0 && m->weak) graph_remove(g, m->type, nullptr, false);
        type->links_lock.unlock_read();
    };
    for (EdgeList *links : {&type->before, &type->after})
        for (auto &e : *links) {
            if (e.second.implicit || e.second.weak) continue;
            Member *dep = graph_add(g, e.first, true, err);
            if (!dep) return undo(), nullptr;
            dep->dependants++;
            pinned.push_back(dep);
        }
    Member *m = member_create(g, type, weak);
    if (!m) {
        *err = "InitError"; // SystemContext.zig:71
        return undo(), nullptr;
    }
    g->members.emplace_back(type, m);
    g->dirty = true;
    type->links_lock.unlock_read();
    return m;
}

void graph_remove(Group *g, SystemType *type, FimoWorldsMeta_Fence *fence, bool may_defer) {
    // SystemGroup.zig:117-158
    Member *first = g->find(type);
    if (!first) return;
    g->erase(type);
    std::vector<Member *> worklist{first};
    for (size_t i = 0; i < worklist.size(); i++) {
        Member *m = worklist[i];
        m->type->links_lock.lock_read();
        for (EdgeList *links : {&m->type->before, &m->type->after})
            for (auto &e : *links) {
                if (e.second.implicit || e.second.weak) continue;
                Member *dep = g->find(e.first);
                if (!dep) continue;
                if (--dep->dependants == 0 && dep->weak) {
                    worklist.push_back(dep);
                    g->erase(e.first);
                }
            }
        m->type->links_lock.unlock_read();
    }
    for (size_t i = worklist.size(); i-- > 0;) {
        Member *m = worklist[i];
        if (may_defer && g->generation_running()) { // :146-150 destroy at the next recompile
            if (m->type == type && fence) m->removal_waiters.push_back(fence);
            g->retired.push_back(m);
        } else {
            drain_retired(g);
            const bool is_target = m->type == type;
            member_destroy(m);
            if (is_target && fence) fence_signal(fence);
        }
    }
    g->dirty = true;
}

size_t assign_rank(Group *g, Member *m, std::vector<Member *> &emitted) {
    // SystemGroup.zig:214-246: rank = 1 + max(rank of predecessors that are members)
    if (m->order != (size_t)-1) return m->rank;
    size_t rank = 0;
    m->order = (size_t)-2; // in progress (cycles were rejected at registration)
    m->type->links_lock.lock_read();
    for (auto &e : m->type->before)
        if (Member *dep = g->find(e.first)) rank = std::max(rank, assign_rank(g, dep, emitted) + 1);
    m->type->links_lock.unlock_read();
    m->rank = rank;
    m->order = emitted.size();
    emitted.push_back(m);
    return rank;
}

// ---- device executor construction -----------------------------------------------------------------
const FimoB200_ResourceView *view_of(Group *g, ResourceType *t) {
    for (size_t i = 0; i < g->resources.size(); i++)
        if (g->resources[i] == t) return (const FimoB200_ResourceView *)g->resource_blobs[i];
    return nullptr;
}

FimoB200_LaunchInfo launch_info(Group *g) {
    FimoB200_LaunchInfo info{};
    info.n_worlds = g->world->n_worlds;
    info.entities_per_world = g->world->entities;
    info.device = g->world->device;
    return info;
}

uint64_t algorithmic_bytes(Group *g) {
    // every live-in resource read once + every live-out resource written once.  A
    // resource is live-in if some system reads it (shared, or exclusive = read-modify-write for the
    // built-ins that update in place) before any system of the generation fully overwrites it.
    // Derived from the registered resource tables: shared => read; exclusive => written; exclusive
    // of a built-in that updates in place (integrate's pos) => read as well.
    std::unordered_set<ResourceType *> read, written;
    for (Member *m : g->order) {
        for (ResourceType *r : m->type->shared)
            if (!written.count(r)) read.insert(r);
        for (ResourceType *r : m->type->exclusive) {
            const uint32_t id = m->type->builtin_id; // built-ins that update their exclusive resource in place
            const bool in_place = id == FIMO_B200_BUILTIN_INTEGRATE || id == FIMO_B200_BUILTIN_FORCES || id == FIMO_B200_BUILTIN_VEL_CLAMP ||
                                  id == FIMO_B200_BUILTIN_ANGVEL_DAMP || id == FIMO_B200_BUILTIN_INTEGRATE_ANG;
            if (in_place && !written.count(r)) read.insert(r);
            written.insert(r);
        }
    }
    auto bytes = [&](ResourceType *r) -> uint64_t {
        if (!r->on_device || r->scope == FIMO_B200_SCOPE_BATCH || r->scope == FIMO_B200_SCOPE_WORLD) return 0; // < 0.1 %
        const FimoB200_ResourceView *v = view_of(g, r);
        return v ? v->world_stride * v->n_worlds : 0;
    };
    uint64_t total = 0;
    for (ResourceType *r : read) total += bytes(r);
    for (ResourceType *r : written) total += bytes(r);
    return total;
}

// A host system inside a device group.  It runs where the reference runs every system — as a task on
// the group's pool (SystemContext.run, SystemContext.zig:91-100) — started from a host node of the generation's CUDA
// graph, so it is ordered with the device systems by the same DAG edges.  Device resources it uses appear to it as
// `FimoB200_ResourceView::host_ptr`: a pinned mirror, refreshed (HBM -> host) in front of the system and, for the
// resources it holds exclusively, written back (host -> HBM) behind it.
void member_run_from_stream(void *user) {
    Member *m = (Member *)user;
    Group *g = m->group;
    FimoTasksMeta_CommandBufferEntry entry{};
    entry.type = FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK;
    entry.payload.enqueue_task = &m->task;
    FimoTasksMeta_CommandBuffer buffer{};
    buffer.label = m->type->label.data();
    buffer.label_len = m->type->label.size();
    buffer.entries = &entry;
    buffer.entries_len = 1;
    FimoTasksMeta_CommandBufferHandle handle{};
    if (g->executor.vtable->enqueue_buffer(g->executor.data, &buffer, &handle) != FIMO_STATUS_OK)
        panic("enqueue of a host system of a device group failed");
    handle.vtable->wait_on(handle.data);
    handle.vtable->release(handle.data);
}

void *ensure_mirror(World *w, ResourceType *type) {
    w->lock.lock_read();
    auto it = w->resources.find(type);
    Instance *inst = it == w->resources.end() ? nullptr : it->second;
    w->lock.unlock_read();
    if (!inst || !type->on_device) return nullptr;
    FimoB200_ResourceView *view = (FimoB200_ResourceView *)inst->blob;
    if (!inst->mirror) {
        inst->mirror = dev::host_alloc_pinned(inst->hbm_bytes);
        view->host_ptr = inst->mirror; // the group holds the resource exclusively whenever a system looks at the view
    }
    return inst->mirror;
}

bool host_node(Group *g, Member *m, dev::Node &node, const char **err) {
    std::vector<dev::Copy> in, out;
    auto add = [&](ResourceType *r, bool exclusive) {
        if (!r->on_device) return true;
        const FimoB200_ResourceView *view = view_of(g, r);
        void *mirror = ensure_mirror(g->world, r);
        if (!view || !view->device_ptr || !mirror) return false;
        const size_t bytes = view->world_stride * view->n_worlds;
        in.push_back(dev::Copy{mirror, view->device_ptr, bytes});
        if (exclusive) out.push_back(dev::Copy{view->device_ptr, mirror, bytes});
        return true;
    };
    for (ResourceType *r : m->type->exclusive)
        if (!add(r, true)) return *err = "OutOfMemory", false;
    for (ResourceType *r : m->type->shared)
        if (!add(r, false)) return *err = "OutOfMemory", false;
    node.launch = [m, in, out](const FimoB200_LaunchInfo *info) {
        return dev::enqueue_host_node(info->stream, in.data(), in.size(), &member_run_from_stream, m, out.data(), out.size())
                       ? FIMO_STATUS_OK
                       : FIMO_STATUS_ERROR;
    };
    node.kernels = 0;
    return true;
}

bool build_device_executor(Group *g, const char **err) {
    // Emits the true DAG: node j depends on node i (i earlier in rank order) if j names i as an
    // explicit predecessor or they conflict on a resource (W/W, R/W, W/R) — the same ordering
    // SystemGroup.zig:274-336 enforces with wait entries, minus the false serialisation of an
    // in-order command buffer.
    const size_t n = g->order.size();
    std::vector<std::vector<uint32_t>> deps(n);
    uint64_t edges = 0;
    for (size_t j = 0; j < n; j++) {
        SystemType *tj = g->order[j]->type;
        for (size_t i = 0; i < j; i++) {
            SystemType *ti = g->order[i]->type;
            bool ordered = false;
            tj->links_lock.lock_read();
            ordered = edge_to(tj->before, ti) != nullptr;
            tj->links_lock.unlock_read();
            auto touches = [](const std::vector<ResourceType *> &l, ResourceType *r) { return std::find(l.begin(), l.end(), r) != l.end(); };
            for (ResourceType *r : tj->exclusive)
                if (touches(ti->exclusive, r) || touches(ti->shared, r)) ordered = true;
            for (ResourceType *r : tj->shared)
                if (touches(ti->exclusive, r)) ordered = true;
            if (ordered) deps[j].push_back((uint32_t)i), edges++;
        }
    }
    // transitive reduction keeps the graph small (an edge i->j is dropped if some k with i->..->k->j exists)
    std::vector<std::vector<char>> reach(n, std::vector<char>(n, 0));
    for (size_t j = 0; j < n; j++)
        for (uint32_t i : deps[j]) {
            reach[j][i] = 1;
            for (size_t k = 0; k < n; k++)
                if (reach[i][k]) reach[j][k] = 1;
        }
    for (size_t j = 0; j < n; j++) {
        std::vector<uint32_t> kept;
        for (uint32_t i : deps[j]) {
            bool implied = false;
            for (uint32_t k : deps[j])
                if (k != i && reach[k][i]) implied = true;
            if (!implied) kept.push_back(i);
        }
        deps[j].swap(kept);
    }

    // Fusion.  Two patterns collapse into the world-step kernel (world_step.cu):
    //   set 0: integrate -> propagate -> cull (the whole group);
    //   set 1: forces -> [vel_clamp] -> integrate_lin || integrate_ang -> propagate -> world_aabb -> cull_waabb ||
    //          bounds_reduce -> group_stats, wired through the same resources.  Other systems of the group stay
    //          separate graph nodes around the fused one, provided the fused set is convex in the DAG (nothing
    //          outside it sits between two of its members).
    auto by_id = [&](uint32_t id, auto pred) -> Member * {
        for (Member *m : g->order)
            if (m->type->builtin_id == id && pred(m)) return m;
        return nullptr;
    };
    auto any = [](Member *) { return true; };
    auto arity = [](Member *m, size_t x, size_t s) { return m && m->type->exclusive.size() == x && m->type->shared.size() == s; };
    std::vector<Member *> fused, fused16; // fused16: the larger set (set 2), tried first
    dev::WorldStepViews views{}, views16{};
    dev::WorldStepPlan plan{};
    Member *mc = nullptr; // the cull member: owner of the statistics scratch block
    if (!(g->ext_flags & FIMO_B200_GROUP_FLAG_NO_FUSION)) {
        Member *mi = by_id(FIMO_B200_BUILTIN_INTEGRATE, any), *mp = by_id(FIMO_B200_BUILTIN_PROPAGATE, any);
        if (arity(mi, 1, 2) && arity(mp, 1, 5) && mp->type->shared[0] == mi->type->exclusive[0] && mi->order < mp->order) {
            ResourceType *pos = mi->type->exclusive[0], *vel = mi->type->shared[0], *dt = mi->type->shared[1];
            ResourceType *world = mp->type->exclusive[0], *rot = mp->type->shared[1];
            views.pos = view_of(g, pos), views.vel = view_of(g, vel), views.dt = view_of(g, dt);
            views.world = view_of(g, world), views.rot = view_of(g, rot);
            views.scale = view_of(g, mp->type->shared[2]), views.parent = view_of(g, mp->type->shared[3]);
            views.level_start = view_of(g, mp->type->shared[4]);
            Member *c3 = by_id(FIMO_B200_BUILTIN_CULL, [&](Member *m) { return arity(m, 4, 4) && m->type->shared[0] == world; });
            if (c3 && mp->order < c3->order) {
                mc = c3;
                views.aabb_c = view_of(g, mc->type->shared[1]), views.aabb_h = view_of(g, mc->type->shared[2]);
                views.frustum = view_of(g, mc->type->shared[3]);
                fused = {mi, mp, mc};
            } else {
                Member *mf = by_id(FIMO_B200_BUILTIN_FORCES, [&](Member *m) { return arity(m, 1, 2) && m->type->exclusive[0] == vel && m->type->shared[1] == dt; });
                Member *ma = by_id(FIMO_B200_BUILTIN_INTEGRATE_ANG, [&](Member *m) { return arity(m, 1, 2) && m->type->exclusive[0] == rot && m->type->shared[1] == dt; });
                Member *mw = by_id(FIMO_B200_BUILTIN_WORLD_AABB, [&](Member *m) { return arity(m, 1, 3) && m->type->shared[0] == world; });
                ResourceType *waabb = mw ? mw->type->exclusive[0] : nullptr;
                Member *mb = by_id(FIMO_B200_BUILTIN_BOUNDS_REDUCE, [&](Member *m) { return arity(m, 1, 1) && m->type->shared[0] == waabb; });
                Member *ms = nullptr, *c8 = nullptr;
                for (Member *cand : g->order) { // the cull whose count feeds a group_stats together with bounds_reduce's output
                    if (cand->type->builtin_id != FIMO_B200_BUILTIN_CULL_WAABB || !arity(cand, 4, 2) || cand->type->shared[0] != waabb || !mb) continue;
                    ms = by_id(FIMO_B200_BUILTIN_GROUP_STATS, [&](Member *m) {
                        return arity(m, 1, 2) && m->type->shared[0] == cand->type->exclusive[2] && m->type->shared[1] == mb->type->exclusive[0];
                    });
                    if (ms) {
                        c8 = cand;
                        break;
                    }
                }
                Member *mv = by_id(FIMO_B200_BUILTIN_VEL_CLAMP, [&](Member *m) { return arity(m, 1, 1) && m->type->exclusive[0] == vel; });
                const bool ordered = mf && ma && mw && mb && ms && c8 && mf->order < mi->order && ma->order < mp->order &&
                                     mp->order < mw->order && mw->order < c8->order && mw->order < mb->order && c8->order < ms->order &&
                                     mb->order < ms->order && (!mv || (mf->order < mv->order && mv->order < mi->order));
                if (ordered) {
                    mc = c8;
                    views.set = 1;
                    views.gravity = view_of(g, mf->type->shared[0]);
                    views.angvel = view_of(g, ma->type->shared[0]);
                    views.waabb = view_of(g, waabb);
                    views.aabb_c = view_of(g, mw->type->shared[1]), views.aabb_h = view_of(g, mw->type->shared[2]);
                    views.frustum = view_of(g, mc->type->shared[1]);
                    views.wbounds = view_of(g, mb->type->exclusive[0]);
                    views.gstats = view_of(g, ms->type->exclusive[0]);
                    if (mv) views.vel_clamp = true, views.vmax = view_of(g, mv->type->shared[0]);
                    fused = {mf, mi, ma, mp, mw, mc, mb, ms};
                    if (mv) fused.push_back(mv);
                    // set 2: six more systems of the 16-system DAG that only touch data the kernel already holds
                    ResourceType *angvel = ma->type->shared[0], *mask = mc->type->exclusive[0];
                    Member *md = by_id(FIMO_B200_BUILTIN_ANGVEL_DAMP, [&](Member *m) { return arity(m, 1, 1) && m->type->exclusive[0] == angvel; });
                    Member *msl = by_id(FIMO_B200_BUILTIN_SLEEP_FLAGS, [&](Member *m) { return arity(m, 1, 2) && m->type->shared[0] == vel; });
                    Member *ml = by_id(FIMO_B200_BUILTIN_LOD, [&](Member *m) { return arity(m, 1, 3) && m->type->shared[0] == waabb; });
                    Member *c2 = by_id(FIMO_B200_BUILTIN_CULL_WAABB, [&](Member *m) { return m != c8 && arity(m, 4, 2) && m->type->shared[0] == waabb; });
                    Member *mvb = by_id(FIMO_B200_BUILTIN_VIS_BOUNDS, [&](Member *m) {
                        return arity(m, 1, 2) && m->type->shared[0] == waabb && m->type->shared[1] == mask;
                    });
                    Member *ms2 = (c2 && mvb) ? by_id(FIMO_B200_BUILTIN_GROUP_STATS, [&](Member *m) {
                        return m != ms && arity(m, 1, 2) && m->type->shared[0] == c2->type->exclusive[2] && m->type->shared[1] == mvb->type->exclusive[0];
                    }) : nullptr;
                    if (mv && md && msl && ml && c2 && mvb && ms2 && md->order < ma->order && mv->order < msl->order && mw->order < ml->order &&
                        mw->order < c2->order && c8->order < mvb->order && c2->order < ms2->order && mvb->order < ms2->order) {
                        views16 = views;
                        views16.set = 2;
                        views16.damp = view_of(g, md->type->shared[0]);
                        views16.sleep = view_of(g, msl->type->exclusive[0]), views16.sleep_eps = view_of(g, msl->type->shared[1]);
                        views16.lod = view_of(g, ml->type->exclusive[0]);
                        views16.campos = view_of(g, ml->type->shared[1]), views16.lod_dist2 = view_of(g, ml->type->shared[2]);
                        views16.vis_mask2 = view_of(g, c2->type->exclusive[0]), views16.vis_list2 = view_of(g, c2->type->exclusive[1]);
                        views16.vis_count2 = view_of(g, c2->type->exclusive[2]), views16.stats2 = view_of(g, c2->type->exclusive[3]);
                        views16.frustum2 = view_of(g, c2->type->shared[1]);
                        views16.vbounds = view_of(g, mvb->type->exclusive[0]);
                        views16.gstats2 = view_of(g, ms2->type->exclusive[0]);
                        views16.scratch2 = ((dev::BuiltinState *)c2->state)->scratch;
                        fused16 = fused;
                        for (Member *m : {md, msl, ml, c2, mvb, ms2}) fused16.push_back(m);
                        // prev_copy (X: prev_world, S: world), ordered in front of propagate by the R/W conflict on `world`:
                        // inside the kernel every thread copies its entity's old matrix before the chunk's new ones are stored
                        Member *mpc = by_id(FIMO_B200_BUILTIN_PREV_COPY, [&](Member *m) { return arity(m, 1, 1) && m->type->shared[0] == world; });
                        if (mpc && mpc->order < mp->order && !std::getenv("FIMO_B200_NO_PREV_FUSION")) {
                            views16.prev_world = view_of(g, mpc->type->exclusive[0]);
                            fused16.push_back(mpc);
                        }
                    }
                }
            }
            if (!fused.empty()) {
                views.vis_mask = view_of(g, mc->type->exclusive[0]), views.vis_list = view_of(g, mc->type->exclusive[1]);
                views.vis_count = view_of(g, mc->type->exclusive[2]), views.stats = view_of(g, mc->type->exclusive[3]);
                views16.vis_mask = views.vis_mask, views16.vis_list = views.vis_list, views16.vis_count = views.vis_count, views16.stats = views.stats;
            }
        }
    }
    // convexity: an outside node may be before the fused set or after it, never both
    std::vector<char> in_set, pred, succ;
    auto convex = [&](std::vector<Member *> &set) {
        in_set.assign(n, 0), pred.assign(n, 0), succ.assign(n, 0);
        for (Member *m : set) in_set[m->order] = 1;
        for (size_t x = 0; x < n && !set.empty(); x++) {
            if (in_set[x]) continue;
            for (size_t f = 0; f < n; f++) {
                if (!in_set[f]) continue;
                if (reach[f][x]) pred[x] = 1;
                if (reach[x][f]) succ[x] = 1;
            }
            if (pred[x] && succ[x]) set.clear();
        }
        return !set.empty();
    };
    // per-generation statistics snapshots (FimoB200_GroupDeviceExt::stats_ring): written by whichever kernel publishes
    // the statistics of the group's FIRST cull system — the fused world-step if that system is fused, its own kernel otherwise
    {
        Member *first_cull = nullptr;
        for (Member *m : g->order) {
            const uint32_t id = m->type->builtin_id;
            if (!m->type->device_entry || (id != FIMO_B200_BUILTIN_CULL && id != FIMO_B200_BUILTIN_CULL_WAABB)) continue;
            dev::BuiltinState *st = (dev::BuiltinState *)m->state;
            st->stats_ring = first_cull ? nullptr : g->stats_ring;
            st->stats_ring_slots = first_cull ? 0 : g->stats_ring_slots;
            st->stats_peers = first_cull ? nullptr : dev::stats_exchange_peer_table(g->stats_x);
            if (!first_cull) first_cull = m;
        }
        if (mc && mc == first_cull) {
            views.stats_ring = views16.stats_ring = g->stats_ring;
            views.stats_ring_slots = views16.stats_ring_slots = g->stats_ring_slots;
            views.stats_peers = views16.stats_peers = dev::stats_exchange_peer_table(g->stats_x);
        }
    }
    bool fuse = false;
    if (!fused16.empty() && !std::getenv("FIMO_B200_NO_SET2") && convex(fused16) && dev::world_step_plan(views16, g->world->device, plan))
        fuse = true, fused = fused16, views = views16;
    if (!fuse) fuse = convex(fused) && dev::world_step_plan(views, g->world->device, plan);

    std::vector<dev::Node> nodes;
    bool nodes_ok = true;
    auto system_node = [&](Member *m) {
        dev::Node node;
        if (!m->type->device_entry) { // a host system of this device group
            if (!host_node(g, m, node, err)) nodes_ok = false;
            return node;
        }
        const size_t n_ex = m->type->exclusive.size();
        node.launch = [m, n_ex](const FimoB200_LaunchInfo *info) {
            const FimoB200_ResourceView *const *blobs = (const FimoB200_ResourceView *const *)m->blobs.data();
            return m->type->device_entry(m->state, blobs, blobs + n_ex, info);
        };
        node.kernels = 1;
        if (m->type->builtin_id == FIMO_B200_BUILTIN_PROPAGATE)
            node.kernels = dev::builtin_propagate_kernels(view_of(g, m->type->exclusive[0]), view_of(g, m->type->shared[4]));
        return node;
    };
    if (fuse) {
        // new order: everything that is not after the fused set (original order), the fused node, the rest
        std::vector<int> index(n, -1);
        for (size_t x = 0; x < n; x++)
            if (!in_set[x] && !succ[x]) index[x] = (int)nodes.size(), nodes.push_back(system_node(g->order[x]));
        const int fused_index = (int)nodes.size();
        {
            dev::Node node;
            void *scratch = ((dev::BuiltinState *)mc->state)->scratch;
            node.launch = [views, plan, scratch](const FimoB200_LaunchInfo *info) { return dev::world_step_launch(views, plan, scratch, info); };
            node.launch_gens = [views, plan, scratch](const FimoB200_LaunchInfo *info, uint32_t gens) {
                return dev::world_step_launch(views, plan, scratch, info, gens);
            };
            if (dev::world_step_supports_shadow(plan))
                node.launch_shadow = [views, plan, scratch](const FimoB200_LaunchInfo *info, const dev::CullShadow *shadow) {
                    return dev::world_step_launch(views, plan, scratch, info, 1, shadow);
                };
            node.max_gens = plan.max_gens;
            node.kernels = plan.kernels;
            nodes.push_back(std::move(node));
        }
        for (size_t x = 0; x < n; x++) {
            if (in_set[x]) index[x] = fused_index;
            else if (succ[x]) index[x] = (int)nodes.size(), nodes.push_back(system_node(g->order[x]));
        }
        edges = 0;
        for (size_t x = 0; x < n; x++)
            for (uint32_t d : deps[x]) {
                const int from = index[d], to = index[x];
                if (from == to) continue;
                auto &list = nodes[(size_t)to].deps;
                if (std::find(list.begin(), list.end(), (uint32_t)from) == list.end()) list.push_back((uint32_t)from), edges++;
            }
    } else {
        for (size_t j = 0; j < n; j++) {
            dev::Node node = system_node(g->order[j]);
            node.deps = deps[j];
            nodes.push_back(std::move(node));
        }
    }
    if (!nodes_ok) return false;
    const size_t nodes_count = nodes.size();
    if (!g->executor_dev) g->executor_dev = dev::executor_create(g->world->device);
    if (!g->executor_dev) return *err = "DeviceError", false;
    g->info = FimoB200_ExecutorInfo{};
    g->info.systems = (uint32_t)n;
    g->info.device_systems = 0;
    for (Member *m : g->order) g->info.device_systems += m->type->device_entry ? 1u : 0u;
    g->info.nodes = (uint32_t)nodes.size();
    g->info.fused_world_step = fuse ? 1u : 0u;
    g->info.edges = edges;
    g->info.algorithmic_bytes_per_step = algorithmic_bytes(g);
    const bool use_graph = !(g->ext_flags & FIMO_B200_GROUP_FLAG_NO_GRAPH);
    if (!dev::executor_compile(g->executor_dev, std::move(nodes), launch_info(g), use_graph)) return *err = "DeviceError", false;
    g->info.uses_cuda_graph = dev::executor_uses_graph(g->executor_dev) ? 1u : 0u;
    g->info.kernel_launches_per_generation = dev::executor_kernels_per_generation(g->executor_dev);
    g->info.max_generations_per_launch = dev::executor_max_gens(g->executor_dev);
    g->plan_text = fuse ? dev::world_step_describe(plan, g->world->n_worlds) : "one kernel per system";
    if (fuse && nodes_count == 1 && plan.max_gens > 1)
        g->plan_text += "; up to " + std::to_string(plan.max_gens) + " queued generations per launch";
    if (g->info.nodes > 1) g->plan_text += " (" + std::to_string(g->info.nodes) + " graph nodes)";
    return true;
}

// The resource blobs of a batch world do not move while they are in the world, so the compiled
// executor may keep pointers to them; they are looked up here without taking the resource locks.
bool snapshot_blobs(Group *g) {
    g->resource_blobs.assign(g->resources.size(), nullptr);
    g->world->lock.lock_read();
    bool complete = true;
    for (size_t i = 0; i < g->resources.size(); i++) {
        auto it = g->world->resources.find(g->resources[i]);
        if (it == g->world->resources.end()) complete = false;
        else g->resource_blobs[i] = it->second->blob;
    }
    g->world->lock.unlock_read();
    return complete;
}

void scatter_blobs(Group *g) { // SystemGroup.zig:345-350
    auto blob = [&](ResourceType *t) -> void * {
        for (size_t i = 0; i < g->resources.size(); i++)
            if (g->resources[i] == t) return g->resource_blobs[i];
        return nullptr;
    };
    for (Member *m : g->order) {
        size_t i = 0;
        for (ResourceType *r : m->type->exclusive) m->blobs[i++] = blob(r);
        for (ResourceType *r : m->type->shared) m->blobs[i++] = blob(r);
    }
}

// SystemGroup.zig:160-338.  Called with g->m held.
bool compile(Group *g, const char **err) {
    if (!g->dirty) return true;
    if (g->executor_dev) dev::executor_wait_idle(g->executor_dev);
    drain_retired(g);
    g->resources.clear();
    g->has_device_systems = g->has_host_systems = false;
    for (auto &p : g->members) {
        Member *m = p.second;
        for (ResourceType *r : m->type->exclusive)
            if (std::find(g->resources.begin(), g->resources.end(), r) == g->resources.end()) g->resources.push_back(r);
        for (ResourceType *r : m->type->shared)
            if (std::find(g->resources.begin(), g->resources.end(), r) == g->resources.end()) g->resources.push_back(r);
        m->type->links_lock.lock_read();
        m->merge_deferred = true; // :178-181
        for (auto &e : m->type->after)
            if (e.second.ignore_deferred) m->merge_deferred = false;
        m->deferred_deps.clear(); // :184-191
        for (auto &e : m->type->before)
            if (!e.second.ignore_deferred)
                if (Member *dep = g->find(e.first))
                    if (!(m->type->device_entry || dep->type->device_entry)) // device work is ordered by the stream, not by fences
                        m->deferred_deps.push_back(dep);
        m->type->links_lock.unlock_read();
        m->blobs.assign(m->type->exclusive.size() + m->type->shared.size(), nullptr);
        m->order = (size_t)-1;
        (m->type->device_entry ? g->has_device_systems : g->has_host_systems) = true;
    }
    std::vector<Member *> emitted;
    for (auto &p : g->members) assign_rank(g, p.second, emitted);
    // The reference sorts unstably (:262) => order inside a rank is unspecified there; here the
    // tie-break is emission order (= add_systems order), identical on every run and every GPU.
    std::stable_sort(emitted.begin(), emitted.end(), [](Member *a, Member *b) { return a->rank < b->rank; });
    for (size_t i = 0; i < emitted.size(); i++) emitted[i]->order = i;
    g->order.swap(emitted);

    // host path: the in-order command buffer with wait entries (:264-336)
    g->entries.clear();
    std::unordered_map<SystemType *, size_t> issued_at;
    struct Holders {
        bool exclusive = false;
        std::vector<SystemType *> systems;
    };
    std::unordered_map<ResourceType *, Holders> holders;
    auto wait_for = [&](SystemType *s) {
        auto it = issued_at.find(s);
        if (it == issued_at.end()) return;
        FimoTasksMeta_CommandBufferEntry e{};
        e.type = FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_INDIRECT;
        e.payload.wait_on_command_indirect = g->entries.size() - it->second;
        issued_at.erase(it);
        g->entries.push_back(e);
    };
    for (Member *m : g->order) {
        SystemType *t = m->type;
        t->links_lock.lock_read();
        for (auto &e : t->before) wait_for(e.first);
        t->links_lock.unlock_read();
        for (ResourceType *r : t->exclusive) {
            Holders &h = holders[r];
            for (SystemType *s : h.systems) wait_for(s);
            h.exclusive = true;
            h.systems.assign(1, t);
        }
        for (ResourceType *r : t->shared) {
            Holders &h = holders[r];
            if (h.exclusive) {
                for (SystemType *s : h.systems) wait_for(s);
                h.systems.clear();
            }
            h.exclusive = false;
            h.systems.push_back(t);
        }
        issued_at[t] = g->entries.size();
        m->task = FimoTasksMeta_Task{};
        m->task.label = t->label.data();
        m->task.label_len = t->label.size();
        m->task.on_start = &member_task_start;
        FimoTasksMeta_CommandBufferEntry e{};
        e.type = FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK;
        e.payload.enqueue_task = &m->task;
        g->entries.push_back(e);
    }

    g->info = FimoB200_ExecutorInfo{};
    g->info.systems = (uint32_t)g->order.size();
    if (g->has_device_systems) {
        if (g->world->device < 0) return *err = "NoDevice", false;
        g->compiled_epoch = g->world->resource_epoch.load(std::memory_order_acquire);
        if (!snapshot_blobs(g)) return *err = "NotFound", false;
        scatter_blobs(g);
        if (!build_device_executor(g, err)) return false;
    }
    g->dirty = false;
    return true;
}

// ---- generation protocol ----------------------------------------------------------------------------
void end_generation_locked(Group *g) { // SystemGroup.zig:483-486 (g->m held)
    const size_t gen = g->generation.fetch_add(1, std::memory_order_acq_rel) + 1;
    g->single_generation.reset();
    g->multi_generation[gen % 4].reset(); // the documented 4-ring, not the `& 4` of MultiGenerationAllocator.zig:33
}

void unlock_group_resources(Group *g) { // :477-481
    for (ResourceType *r : g->resources) world_unlock_resource(g->world, r, true);
}

struct DeviceDone {
    Group *group;
    FimoWorldsMeta_Fence *signal;
    uint64_t generation = 0;
    int64_t enqueue_ns = 0;
    bool ends_launch = true; // the last generation of its launch
};
void launch_pending_batches(Group *g);
void dispatcher_main(Group *g);

std::atomic<FimoB200_trace_sink> g_trace_sink{nullptr};
std::atomic<void *> g_trace_user{nullptr};
bool tracing_wanted() { return g_trace_sink.load(std::memory_order_relaxed) || trace_bridge_active(); }
int64_t monotonic_ns() {
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

void device_generation_done(void *user) {
    // Runs on the completion thread once the device finished the generation: the second half of
    // SystemGroup.run (:477-486) and the Job cleanup (Job.zig:234-247).
    std::unique_ptr<DeviceDone> done((DeviceDone *)user);
    Group *g = done->group;
    FimoB200_trace_sink sink = g_trace_sink.load(std::memory_order_acquire);
    if (sink || trace_bridge_active()) { // before the group may be destroyed
        FimoB200_TraceEvent ev{};
        ev.group_label = g->label.data(), ev.group_label_len = g->label.size();
        ev.generation = done->generation;
        ev.device_ms = dev::executor_last_generation_ms(g->executor_dev);
        ev.host_enqueue_ns = done->enqueue_ns, ev.host_complete_ns = monotonic_ns();
        ev.kernels = g->info.kernel_launches_per_generation, ev.fused_world_step = g->info.fused_world_step;
        if (sink) sink(g_trace_user.load(std::memory_order_acquire), &ev);
        trace_bridge_generation(g, ev); // fimo tracing subscribers (tracing.cpp)
    }
    size_t gen;
    {
        std::lock_guard<std::mutex> lock(g->m);
        if (done->ends_launch) {
            g->launches_in_flight--;
            if (!g->batch_pending.empty()) launch_pending_batches(g); // keeps the stream fed: one launch is still running
        }
        if (--g->in_flight == 0) unlock_group_resources(g);
        end_generation_locked(g);
        gen = g->generation.load(std::memory_order_acquire);
    }
    FimoWorldsMeta_Fence *signal = done->signal;
    FimoWorldsMeta_timeline_semaphore_signal(&g->schedule_semaphore, (uint64_t)gen, g_futex_wake);
    if (signal) fence_signal(signal); // the group may be destroyed by the waiter from here on
}

void device_transfer_done(void *user) {
    std::unique_ptr<DeviceDone> done((DeviceDone *)user);
    Group *g = done->group;
    {
        std::lock_guard<std::mutex> lock(g->m);
        if (--g->in_flight == 0) unlock_group_resources(g);
    }
    if (done->signal) fence_signal(done->signal);
}

// Enqueues one device generation.  g->m held.  `blocking`: may wait for the resource locks.
// Returns 0 = started, 1 = would block, -1 = failed (*err set).
int start_device_generation(Group *g, Job *job, bool blocking, const char **err) {
    if (g->compiled_epoch != g->world->resource_epoch.load(std::memory_order_acquire)) g->dirty = true;
    if (g->dirty) {
        if (g->in_flight != 0 && !blocking) return 1; // a recompile has to wait for the stream to drain
        if (!compile(g, err)) return -1;
    }
    if (g->in_flight == 0) {
        // all group resources are held exclusively while generations are in flight (:340-351)
        std::vector<void *> blobs(g->resources.size());
        if (blocking) world_lock_resources(g->world, g->resources.data(), g->resources.size(), nullptr, 0, blobs.data());
        else if (!world_try_lock_resources(g->world, g->resources.data(), g->resources.size(), blobs.data())) return 1;
    }
    if (job->xfer_resource) {
        const FimoB200_ResourceView *view = view_of(g, job->xfer_resource);
        if (!view || !view->device_ptr) {
            if (g->in_flight == 0) unlock_group_resources(g);
            *err = "NotFound";
            return -1;
        }
        g->in_flight++;
        DeviceDone *done = new DeviceDone{g, job->signal};
        const size_t bytes = view->world_stride * view->n_worlds;
        void *dst = job->xfer_to_device ? view->device_ptr : job->xfer_host;
        const void *src = job->xfer_to_device ? job->xfer_host : view->device_ptr;
        if (!dev::executor_transfer(g->executor_dev, dst, src, bytes, job->xfer_to_device, &device_transfer_done, done)) {
            delete done;
            if (--g->in_flight == 0) unlock_group_resources(g);
            *err = "DeviceError";
            return -1;
        }
        return 0;
    }
    std::vector<dev::Copy> ups, downs;
    for (int dir = 0; dir < 2; dir++)
        for (auto &io : dir == 0 ? job->io_up : job->io_down) {
            const FimoB200_ResourceView *view = view_of(g, io.first);
            if (!view || !view->device_ptr) {
                if (g->in_flight == 0) unlock_group_resources(g);
                *err = "NotFound";
                return -1;
            }
            const size_t bytes = view->world_stride * view->n_worlds;
            if (dir == 0) ups.push_back(dev::Copy{view->device_ptr, io.second, bytes});
            else downs.push_back(dev::Copy{io.second, view->device_ptr, bytes});
        }
    dev::Readback rb{};
    if (job->has_readback) {
        const FimoB200_ResourceView *mask = view_of(g, (ResourceType *)job->readback.vis_mask), *list = view_of(g, (ResourceType *)job->readback.vis_list),
                                    *count = view_of(g, (ResourceType *)job->readback.vis_count);
        const uint64_t W = g->world->n_worlds, E = g->world->entities;
        if (!mask || !list || !count || !mask->device_ptr || !list->device_ptr || !count->device_ptr || mask->scope != FIMO_B200_SCOPE_ENTITY_BITS ||
            mask->components != 1 || list->scope != FIMO_B200_SCOPE_ENTITY || list->components != 1 || count->scope != FIMO_B200_SCOPE_WORLD ||
            count->components != 1 || mask->n_worlds != W || list->n_worlds != W || count->n_worlds != W) {
            if (g->in_flight == 0) unlock_group_resources(g);
            *err = "NotFound";
            return -1;
        }
        rb = dev::Readback{(const uint32_t *)mask->device_ptr, (const uint32_t *)list->device_ptr, (const uint32_t *)count->device_ptr, W, E,
                           (E + 31) / 32, (uint64_t)job->generation, job->readback.host, (size_t)job->readback.host_bytes};
    }
    g->in_flight++;
    g->launches_in_flight++;
    DeviceDone *done = new DeviceDone{g, job->signal, (uint64_t)job->generation, tracing_wanted() ? monotonic_ns() : 0};
    if (g->stats_x) dev::stats_exchange_before_launch(g->stats_x, dev::executor_stream(g->executor_dev), 1);
    if (!dev::executor_launch(g->executor_dev, &device_generation_done, done, ups.data(), ups.size(), downs.data(), downs.size(),
                              job->has_readback ? &rb : nullptr)) {
        // The completion entry still drains through the completion thread (without `done`); undo here.
        delete done;
        g->launches_in_flight--;
        if (--g->in_flight == 0) unlock_group_resources(g);
        *err = "DeviceError";
        return -1;
    }
    if (g->stats_x) dev::stats_exchange_after_launch(g->stats_x, dev::executor_stream(g->executor_dev), 1);
    if (g->post_step) g->post_step(g->post_step_user, dev::executor_stream(g->executor_dev), (uint64_t)job->generation);
    return 0;
}

// A generation that may share a launch with its neighbours: nothing but the generation itself.
bool batchable(const Group *g, const Job *job) {
    return !job->xfer_resource && job->io_up.empty() && job->io_down.empty() && !job->has_readback && job->wait_on.empty() && !g->post_step;
}

// g->m held, resources locked (in_flight > 0).  Launches everything in batch_pending, up to max_gens generations per
// launch.  Called from the completion thread when a launch retires and from `schedule` when something that must
// stay behind the pending generations arrives.
void launch_pending_batches(Group *g) {
    if (g->dirty || g->compiled_epoch != g->world->resource_epoch.load(std::memory_order_acquire)) {
        // a recompile has to drain the stream, which this thread may be needed for: hand the jobs to the dispatcher
        while (!g->batch_pending.empty()) g->backlog.push_front(g->batch_pending.back()), g->batch_pending.pop_back();
        if (!g->dispatcher_running) {
            if (g->dispatcher.joinable()) g->dispatcher.join();
            g->dispatcher_running = true, g->stop_dispatcher = false;
            g->dispatcher = std::thread(dispatcher_main, g);
        }
        g->backlog_cv.notify_one();
        return;
    }
    // with a NCCL statistics ring bound, a launch must not wrap the ring into rows whose all-reduce is still to come
    const uint32_t max_gens = std::min(dev::executor_max_gens(g->executor_dev), dev::stats_exchange_max_gens(g->stats_x));
    const int64_t now = tracing_wanted() ? monotonic_ns() : 0;
    while (!g->batch_pending.empty()) {
        const uint32_t n = (uint32_t)std::min<size_t>(g->batch_pending.size(), max_gens);
        std::vector<void *> users(n);
        for (uint32_t i = 0; i < n; i++) {
            Job *job = g->batch_pending[i];
            DeviceDone *done = new DeviceDone{g, job->signal, (uint64_t)job->generation, now};
            done->ends_launch = i + 1 == n;
            users[i] = done;
        }
        g->in_flight += n;
        g->launches_in_flight++;
        void *stream = dev::executor_stream(g->executor_dev);
        if (g->stats_x) dev::stats_exchange_before_launch(g->stats_x, stream, n);
        if (!dev::executor_launch_batch(g->executor_dev, n, &device_generation_done, users.data())) panic("device generations failed to start");
        if (g->stats_x) dev::stats_exchange_after_launch(g->stats_x, stream, n);
        for (uint32_t i = 0; i < n; i++) delete g->batch_pending[i];
        g->batch_pending.erase(g->batch_pending.begin(), g->batch_pending.begin() + n);
    }
}

bool fences_signaled(const Job *job) {
    for (FimoWorldsMeta_Fence *f : job->wait_on)
        if (!FimoWorldsMeta_fence_is_signaled(f)) return false;
    return true;
}

// One generation of host systems: SystemGroup.run (:454-487) executed by the dispatcher thread,
// which plays the role of the group job's fiber.
void run_host_generation(Group *g, Job *job) {
    const char *err = nullptr;
    {
        std::lock_guard<std::mutex> lock(g->m);
        if (!compile(g, &err)) panic("system group compile failed");
        g->resource_blobs.assign(g->resources.size(), nullptr);
        world_lock_resources(g->world, g->resources.data(), g->resources.size(), nullptr, 0, g->resource_blobs.data());
        scatter_blobs(g);
    }
    FimoTasksMeta_CommandBuffer buffer{};
    buffer.label = g->systems_label.data();
    buffer.label_len = g->systems_label.size();
    buffer.entries = g->entries.data();
    buffer.entries_len = g->entries.size();
    FimoTasksMeta_CommandBufferHandle handle{};
    const int64_t t_enqueue = trace_bridge_active() ? monotonic_ns() : 0;
    if (g->executor.vtable->enqueue_buffer(g->executor.data, &buffer, &handle) != FIMO_STATUS_OK)
        panic("enqueue of the systems buffer failed");
    handle.vtable->wait_on(handle.data);
    handle.vtable->release(handle.data);
    if (t_enqueue) { // a host generation is a span too: the systems buffer between enqueue and completion (kernels = 0)
        FimoB200_TraceEvent ev{};
        ev.group_label = g->label.data(), ev.group_label_len = g->label.size();
        ev.generation = (uint64_t)job->generation;
        ev.host_enqueue_ns = t_enqueue, ev.host_complete_ns = monotonic_ns();
        ev.device_ms = (double)(ev.host_complete_ns - t_enqueue) / 1.0e6;
        trace_bridge_generation(g, ev);
    }
    size_t gen;
    {
        std::lock_guard<std::mutex> lock(g->m);
        unlock_group_resources(g);
        end_generation_locked(g);
        gen = g->generation.load(std::memory_order_acquire);
    }
    FimoWorldsMeta_Fence *signal = job->signal;
    FimoWorldsMeta_timeline_semaphore_signal(&g->schedule_semaphore, (uint64_t)gen, g_futex_wake);
    if (signal) fence_signal(signal);
}

void dispatcher_main(Group *g) {
    std::unique_lock<std::mutex> lock(g->m);
    for (;;) {
        g->backlog_cv.wait(lock, [&] { return g->stop_dispatcher || !g->backlog.empty(); });
        if (g->backlog.empty()) break; // stop requested and nothing left
        Job *job = g->backlog.front();
        g->dispatcher_busy = true;
        lock.unlock();
        for (FimoWorldsMeta_Fence *f : job->wait_on) fence_wait(f); // Job.zig:224-227
        // generations serialise: host generations by running here one at a time, device generations
        // by stream order (Job.zig:228-232 waits on the timeline semaphore for the same effect)
        const char *err = nullptr;
        bool device, stale;
        {
            std::lock_guard<std::mutex> relock(g->m);
            stale = g->dirty || g->compiled_epoch != g->world->resource_epoch.load(std::memory_order_acquire);
        }
        // A recompile must not overlap generations that are still on the stream, and the completion
        // thread needs g->m to retire them: drain first, without holding the graph mutex.
        if (stale) FimoWorldsMeta_timeline_semaphore_wait(&g->schedule_semaphore, (uint64_t)job->generation, g_futex_wait);
        {
            std::lock_guard<std::mutex> relock(g->m);
            if (stale) g->dirty = true;
            if (g->dirty && !compile(g, &err)) panic("system group compile failed");
            device = g->has_device_systems;
        }
        if (device) {
            std::lock_guard<std::mutex> relock(g->m);
            if (start_device_generation(g, job, true, &err) != 0) panic("device generation failed to start");
        } else if (job->xfer_resource) {
            panic("transfer queued on a group without device systems");
        } else {
            FimoWorldsMeta_timeline_semaphore_wait(&g->schedule_semaphore, (uint64_t)job->generation, g_futex_wait);
            run_host_generation(g, job);
        }
        lock.lock();
        g->backlog.pop_front();
        g->dispatcher_busy = false;
        delete job;
    }
    g->dispatcher_running = false;
}

struct Transfer {
    ResourceType *resource;
    void *host;
    bool to_device;
};

struct StepIo {
    const FimoB200_StepTransfer *uploads;
    size_t n_up;
    const FimoB200_StepTransfer *downloads;
    size_t n_down;
    const FimoB200_CullReadback *readback = nullptr;
};

const char *group_schedule(Group *g, FimoWorldsMeta_Fence *const *wait_on, size_t n, FimoWorldsMeta_Fence *signal,
                           const Transfer *xfer = nullptr, const StepIo *io = nullptr) {
    // SystemGroup.zig:430-452
    std::unique_lock<std::mutex> lock(g->m);
    Job *job = new Job{g, g->next_generation, std::vector<FimoWorldsMeta_Fence *>(wait_on, wait_on + n), signal};
    if (io) {
        for (size_t i = 0; i < io->n_up; i++) job->io_up.emplace_back((ResourceType *)io->uploads[i].resource, io->uploads[i].host);
        for (size_t i = 0; i < io->n_down; i++) job->io_down.emplace_back((ResourceType *)io->downloads[i].resource, io->downloads[i].host);
        if (io->readback) job->has_readback = true, job->readback = *io->readback;
    }
    const size_t consumed = xfer ? 0 : 1; // a transfer rides between generations, it is not one
    if (xfer) job->xfer_resource = xfer->resource, job->xfer_host = xfer->host, job->xfer_to_device = xfer->to_device;
    g->next_generation += consumed;
    const char *err = nullptr;
    // fast path: a device group with nothing queued ahead starts the generation right here
    if (g->backlog.empty() && !g->dispatcher_busy && fences_signaled(job)) {
        bool device = g->has_device_systems && !g->dirty;
        if (g->dirty && g->in_flight == 0) {
            bool any_device = false;
            for (auto &p : g->members) any_device |= p.second->type->device_entry != nullptr;
            if (any_device) {
                if (!compile(g, &err)) {
                    g->next_generation -= consumed;
                    delete job;
                    return err;
                }
                device = g->has_device_systems;
            }
        }
        // FIMO_B200_GENS_HOLD=n (profiling aid: under ncu every launch is synchronous, so nothing ever queues up): hold
        // plain generations until n are pending or one carries a fence, then launch them together
        static const size_t hold = [] { const char *e = std::getenv("FIMO_B200_GENS_HOLD"); return e ? (size_t)std::strtoull(e, nullptr, 10) : 0; }();
        if (hold > 1 && device && dev::executor_max_gens(g->executor_dev) > 1 && batchable(g, job) &&
            g->compiled_epoch == g->world->resource_epoch.load(std::memory_order_acquire)) {
            g->batch_pending.push_back(job);
            if (g->batch_pending.size() >= std::min<size_t>(hold, dev::executor_max_gens(g->executor_dev)) || job->signal) {
                if (g->in_flight == 0) {
                    std::vector<void *> blobs(g->resources.size());
                    world_lock_resources(g->world, g->resources.data(), g->resources.size(), nullptr, 0, blobs.data());
                }
                launch_pending_batches(g);
            }
            return nullptr;
        }
        if (device && dev::executor_max_gens(g->executor_dev) > 1 && g->in_flight != 0 &&
            g->compiled_epoch == g->world->resource_epoch.load(std::memory_order_acquire)) {
            if (batchable(g, job) && (g->launches_in_flight >= 2 || !g->batch_pending.empty())) {
                g->batch_pending.push_back(job); // leaves with its neighbours when a launch retires
                return nullptr;
            }
            if (!g->batch_pending.empty()) launch_pending_batches(g); // this job must stay behind them
        }
        if (device && g->backlog.empty()) {
            const int rc = start_device_generation(g, job, false, &err);
            if (rc == 0) {
                delete job;
                return nullptr;
            }
            if (rc < 0) {
                g->next_generation -= consumed;
                delete job;
                return err;
            }
        }
    }
    while (!g->batch_pending.empty()) g->backlog.push_back(g->batch_pending.front()), g->batch_pending.pop_front(); // earlier generations first
    g->backlog.push_back(job);
    if (!g->dispatcher_running) {
        if (g->dispatcher.joinable()) g->dispatcher.join();
        g->dispatcher_running = true;
        g->stop_dispatcher = false;
        g->dispatcher = std::thread(dispatcher_main, g);
    }
    g->backlog_cv.notify_one();
    return nullptr;
}

Group *group_create(const FimoWorldsMeta_SystemGroupDescriptor *d) {
    // SystemGroup.zig:359-378
    Group *g = new Group();
    g->label = label_of(d->label, d->label_len);
    g->world = (World *)d->world;
    g->executor = d->pool ? *d->pool : g->world->executor;
    g->executor.vtable->acquire(g->executor.data);
    g->world->groups.fetch_add(1, std::memory_order_relaxed);
    if (const auto *ext = find_ext<FimoB200_GroupDeviceExt>(d->next, FIMO_B200_STYPE_GROUP_DEVICE)) {
        g->ext_flags = ext->flags;
        g->post_step = ext->post_step;
        g->post_step_user = ext->post_step_user;
        if ((ext->flags & FIMO_B200_GROUP_FLAG_STATS_RING) && ext->stats_ring && ext->stats_ring_slots)
            g->stats_ring = ext->stats_ring, g->stats_ring_slots = ext->stats_ring_slots;
    }
    char buf[64];
    std::snprintf(buf, sizeof buf, "%p systems", (void *)g); // :465
    g->systems_label = buf;
    return g;
}

void group_destroy(Group *g) {
    // SystemGroup.zig:380-397
    {
        std::unique_lock<std::mutex> lock(g->m);
        if (g->generation_running()) panic("system group still running");
        if (!g->members.empty()) panic("system group not empty");
        drain_retired(g);
        g->stop_dispatcher = true;
        g->backlog_cv.notify_all();
    }
    if (g->dispatcher.joinable()) g->dispatcher.join();
    trace_bridge_group_destroyed(g); // its call stack ends (tracing.cpp)
    if (g->executor_dev) dev::executor_destroy(g->executor_dev);
    dev::stats_exchange_destroy(g->stats_x);
    g->world->groups.fetch_sub(1, std::memory_order_relaxed);
    g->executor.vtable->release(g->executor.data);
    delete g;
}

const char *group_add_systems(Group *g, SystemType *const *handles, size_t n) {
    // SystemGroup.zig:399-418
    std::lock_guard<std::mutex> lock(g->m);
    g_registry.lock.lock_read();
    struct Unlock {
        ~Unlock() { g_registry.lock.unlock_read(); }
    } unlock;
    for (size_t i = 0; i < n; i++) {
        if (!g_registry.systems.count(handles[i])) panic("invalid system");
        if (Member *m = g->find(handles[i]))
            if (!m->weak) return "Duplicate";
    }
    for (size_t i = 0; i < n; i++) {
        const char *err = "InitError";
        if (!graph_add(g, handles[i], false, &err)) {
            for (size_t j = 0; j < i; j++) graph_remove(g, handles[j], nullptr, false);
            return err;
        }
    }
    return nullptr;
}

void group_remove_system(Group *g, SystemType *handle, FimoWorldsMeta_Fence *fence) {
    // SystemGroup.zig:420-428
    std::lock_guard<std::mutex> lock(g->m);
    Member *m = g->find(handle);
    if (!m || m->weak) panic("invalid system");
    graph_remove(g, handle, fence, true);
}

// ---- built-in systems: host-side glue ---------------------------------------------------------------
bool builtin_create(const void *factory, FimoWorldsMeta_SystemContext *context, void *system) {
    (void)factory;
    Member *m = (Member *)context;
    dev::BuiltinState *state = (dev::BuiltinState *)system;
    state->device = m->group->world->device;
    state->scratch = nullptr;
    state->stats_ring = nullptr;
    state->stats_ring_slots = 0;
    state->stats_peers = nullptr;
    if (state->device < 0) return false; // built-in systems exist on the device only
    if (m->type->builtin_id == FIMO_B200_BUILTIN_CULL || m->type->builtin_id == FIMO_B200_BUILTIN_CULL_WAABB) {
        state->scratch = dev::hbm_alloc(state->device, dev::kBuiltinScratchBytes);
        if (!state->scratch) return false;
    }
    return true;
}
void builtin_destroy(void *system) {
    dev::BuiltinState *state = (dev::BuiltinState *)system;
    if (state->scratch) dev::hbm_free(state->device, state->scratch);
}
void builtin_run_stub(void *, void *const *, void *const *, FimoWorldsMeta_Fence *) {
    panic("a built-in device system was dispatched to the CPU: there is no CPU fallback");
}

} // namespace
} // namespace fb

// ---------------------------------------------------------------------------------------------
// C ABI (FimoWorlds.zig:108-447)
// ---------------------------------------------------------------------------------------------
using namespace fb;

static FimoStatus report(const char *name) { // ctx.setResult(...); return .err  (FimoWorlds.zig:117-120)
    set_last_error(name);
    return FIMO_STATUS_ERROR;
}

extern "C" {

FimoStatus fimo_worlds_resource_register(const FimoWorldsMeta_ResourceDescriptor *resource, FimoWorldsMeta_ResourceHandle *handle) {
    const char *err = nullptr;
    ResourceType *r = register_resource(resource, &err);
    if (!r) return report(err);
    *handle = (FimoWorldsMeta_ResourceHandle)r;
    return FIMO_STATUS_OK;
}
void fimo_worlds_resource_unregister(FimoWorldsMeta_ResourceHandle handle) { unregister_resource((ResourceType *)handle); }

FimoStatus fimo_worlds_system_register(const FimoWorldsMeta_SystemDescriptor *system, FimoWorldsMeta_SystemHandle *handle) {
    const char *err = nullptr;
    SystemType *s = register_system(system, &err);
    if (!s) return report(err);
    *handle = (FimoWorldsMeta_SystemHandle)s;
    return FIMO_STATUS_OK;
}
void fimo_worlds_system_unregister(FimoWorldsMeta_SystemHandle handle) { unregister_system((SystemType *)handle); }

FimoStatus fimo_worlds_system_group_create(const FimoWorldsMeta_SystemGroupDescriptor *descriptor, FimoWorldsMeta_SystemGroup **group) {
    *group = (FimoWorldsMeta_SystemGroup *)group_create(descriptor);
    return FIMO_STATUS_OK;
}
void fimo_worlds_system_group_destroy(FimoWorldsMeta_SystemGroup *group) { group_destroy((Group *)group); }
FimoWorldsMeta_World *fimo_worlds_system_group_get_world(FimoWorldsMeta_SystemGroup *group) {
    return (FimoWorldsMeta_World *)((Group *)group)->world;
}
const char *fimo_worlds_system_group_get_label(FimoWorldsMeta_SystemGroup *group, FimoUSize *len) {
    Group *g = (Group *)group;
    *len = g->label.size();
    return g->label.data();
}
FimoTasksMeta_Pool fimo_worlds_system_group_get_pool(FimoWorldsMeta_SystemGroup *group) {
    Group *g = (Group *)group;
    g->executor.vtable->acquire(g->executor.data); // a new reference (FimoWorlds.zig:207-210)
    return g->executor;
}
FimoStatus fimo_worlds_system_group_add_systems(FimoWorldsMeta_SystemGroup *group, const FimoWorldsMeta_SystemHandle *systems,
                                                FimoUSize systems_len) {
    const char *err = group_add_systems((Group *)group, (SystemType *const *)systems, systems_len);
    return err ? report(err) : FIMO_STATUS_OK;
}
void fimo_worlds_system_group_remove_system(FimoWorldsMeta_SystemGroup *group, FimoWorldsMeta_SystemHandle handle,
                                            FimoWorldsMeta_Fence *signal) {
    group_remove_system((Group *)group, (SystemType *)handle, signal);
}
FimoStatus fimo_worlds_system_group_schedule(FimoWorldsMeta_SystemGroup *group, FimoWorldsMeta_Fence *const *wait_on,
                                             FimoUSize wait_on_len, FimoWorldsMeta_Fence *signal) {
    const char *err = group_schedule((Group *)group, wait_on, wait_on_len, signal);
    return err ? report(err) : FIMO_STATUS_OK;
}

FimoWorldsMeta_SystemGroup *fimo_worlds_system_context_get_group(FimoWorldsMeta_SystemContext *context) {
    return (FimoWorldsMeta_SystemGroup *)((Member *)context)->group;
}
FimoUSize fimo_worlds_system_context_get_generation(FimoWorldsMeta_SystemContext *context) {
    return ((Member *)context)->group->generation.load(std::memory_order_acquire);
}

static void *strategy_alloc(Member *m, FimoWorldsMeta_SystemAllocatorStrategy strategy, size_t size, size_t alignment) {
    // SystemContext.zig:102-110
    switch (strategy) {
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_TRANSIENT: return m->transient.alloc(size, alignment);
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SINGLE_GENERATION: return m->group->single_generation.alloc(size, alignment);
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_MULTI_GENERATION:
            return m->group->multi_generation[m->group->generation.load() % 4].alloc(size, alignment);
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT: return m->persistent.alloc(size, alignment);
        default: panic("unknown allocator strategy");
    }
}
void *fimo_worlds_system_context_allocator_alloc(FimoWorldsMeta_SystemContext *context, FimoWorldsMeta_SystemAllocatorStrategy strategy,
                                                 FimoUSize size, FimoUSize alignment, FimoUSize) {
    return strategy_alloc((Member *)context, strategy, size, alignment);
}
bool fimo_worlds_system_context_allocator_resize(FimoWorldsMeta_SystemContext *, FimoWorldsMeta_SystemAllocatorStrategy, void *,
                                                 FimoUSize size, FimoUSize, FimoUSize new_size, FimoUSize) {
    return new_size <= size;
}
void *fimo_worlds_system_context_allocator_remap(FimoWorldsMeta_SystemContext *context, FimoWorldsMeta_SystemAllocatorStrategy strategy,
                                                 void *ptr, FimoUSize size, FimoUSize alignment, FimoUSize new_size, FimoUSize) {
    if (new_size <= size) return ptr;
    Member *m = (Member *)context;
    void *p = strategy_alloc(m, strategy, new_size, alignment);
    if (p && ptr) std::memcpy(p, ptr, size);
    if (p && strategy == FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT) m->persistent.release(ptr);
    return p;
}
void fimo_worlds_system_context_allocator_free(FimoWorldsMeta_SystemContext *context, FimoWorldsMeta_SystemAllocatorStrategy strategy,
                                               void *ptr, FimoUSize, FimoUSize, FimoUSize) {
    if (strategy == FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT) ((Member *)context)->persistent.release(ptr);
}

FimoStatus fimo_worlds_world_create(const FimoWorldsMeta_WorldDescriptor *descriptor, FimoWorldsMeta_World **world) {
    const char *err = nullptr;
    World *w = world_create(descriptor, &err);
    if (!w) return report(err);
    *world = (FimoWorldsMeta_World *)w;
    return FIMO_STATUS_OK;
}
void fimo_worlds_world_destroy(FimoWorldsMeta_World *world) { world_destroy((World *)world); }
const char *fimo_worlds_world_get_label(FimoWorldsMeta_World *world, FimoUSize *len) {
    World *w = (World *)world;
    *len = w->label.size();
    return w->label.data();
}
FimoTasksMeta_Pool fimo_worlds_world_get_pool(FimoWorldsMeta_World *world) {
    World *w = (World *)world;
    w->executor.vtable->acquire(w->executor.data); // a new reference (FimoWorlds.zig:338-341)
    return w->executor;
}
bool fimo_worlds_world_has_resource(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle) {
    World *w = (World *)world;
    w->lock.lock_read();
    const bool has = w->resources.count((ResourceType *)handle) != 0;
    w->lock.unlock_read();
    return has;
}
FimoStatus fimo_worlds_world_add_resource(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle, const void *value) {
    const char *err = world_add_resource((World *)world, (ResourceType *)handle, value);
    return err ? report(err) : FIMO_STATUS_OK;
}
FimoStatus fimo_worlds_world_remove_resource(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle, void *value) {
    const char *err = world_remove_resource((World *)world, (ResourceType *)handle, value);
    return err ? report(err) : FIMO_STATUS_OK;
}
void fimo_worlds_world_lock_resources(FimoWorldsMeta_World *world, const FimoWorldsMeta_ResourceHandle *exclusive_handles,
                                      FimoUSize exclusive_handles_len, const FimoWorldsMeta_ResourceHandle *shared_handles,
                                      FimoUSize shared_handles_len, void **out_resources) {
    world_lock_resources((World *)world, (ResourceType *const *)exclusive_handles, exclusive_handles_len,
                         (ResourceType *const *)shared_handles, shared_handles_len, out_resources);
}
void fimo_worlds_world_unlock_resource_exclusive(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle) {
    world_unlock_resource((World *)world, (ResourceType *)handle, true);
}
void fimo_worlds_world_unlock_resource_shared(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle) {
    world_unlock_resource((World *)world, (ResourceType *)handle, false);
}
void *fimo_worlds_world_allocator_alloc(FimoWorldsMeta_World *world, FimoUSize size, FimoUSize alignment, FimoUSize) {
    return ((World *)world)->allocator.alloc(size, alignment);
}
bool fimo_worlds_world_allocator_resize(FimoWorldsMeta_World *, void *, FimoUSize size, FimoUSize, FimoUSize new_size, FimoUSize) {
    return new_size <= size;
}
void *fimo_worlds_world_allocator_remap(FimoWorldsMeta_World *world, void *ptr, FimoUSize size, FimoUSize alignment,
                                        FimoUSize new_size, FimoUSize) {
    if (new_size <= size) return ptr;
    World *w = (World *)world;
    void *p = w->allocator.alloc(new_size, alignment);
    if (p && ptr) std::memcpy(p, ptr, size);
    if (p) w->allocator.release(ptr);
    return p;
}
void fimo_worlds_world_allocator_free(FimoWorldsMeta_World *world, void *ptr, FimoUSize, FimoUSize, FimoUSize) {
    ((World *)world)->allocator.release(ptr);
}

#define FIMO_B200_ROW(n) {#n, "fimo-worlds", (void (*)(void)) & fimo_worlds_##n}
static const FimoB200_Export g_export_table[] = {
        // `all_symbols`, pkgs/fimo_worlds_meta/src/symbols.zig:22-58
        FIMO_B200_ROW(resource_register), FIMO_B200_ROW(resource_unregister), FIMO_B200_ROW(system_register),
        FIMO_B200_ROW(system_unregister), FIMO_B200_ROW(system_group_create), FIMO_B200_ROW(system_group_destroy),
        FIMO_B200_ROW(system_group_get_world), FIMO_B200_ROW(system_group_get_label), FIMO_B200_ROW(system_group_get_pool),
        FIMO_B200_ROW(system_group_add_systems), FIMO_B200_ROW(system_group_remove_system), FIMO_B200_ROW(system_group_schedule),
        FIMO_B200_ROW(system_context_get_group), FIMO_B200_ROW(system_context_get_generation),
        FIMO_B200_ROW(system_context_allocator_alloc), FIMO_B200_ROW(system_context_allocator_resize),
        FIMO_B200_ROW(system_context_allocator_remap), FIMO_B200_ROW(system_context_allocator_free), FIMO_B200_ROW(world_create),
        FIMO_B200_ROW(world_destroy), FIMO_B200_ROW(world_get_label), FIMO_B200_ROW(world_get_pool), FIMO_B200_ROW(world_has_resource),
        FIMO_B200_ROW(world_add_resource), FIMO_B200_ROW(world_remove_resource), FIMO_B200_ROW(world_lock_resources),
        FIMO_B200_ROW(world_unlock_resource_exclusive), FIMO_B200_ROW(world_unlock_resource_shared),
        FIMO_B200_ROW(world_allocator_alloc), FIMO_B200_ROW(world_allocator_resize), FIMO_B200_ROW(world_allocator_remap),
        FIMO_B200_ROW(world_allocator_free),
};
const FimoB200_Export *fimo_worlds_b200_exports(FimoUSize *count) {
    *count = sizeof(g_export_table) / sizeof(g_export_table[0]);
    return g_export_table;
}

// ---- device-side helpers --------------------------------------------------------------------------
FimoStatus fimo_b200_builtin_system(uint32_t builtin_id, FimoWorldsMeta_SystemDescriptor *desc, FimoB200_SystemDeviceExt *ext) {
    FimoB200_device_entry entry = nullptr;
    switch (builtin_id) {
        case FIMO_B200_BUILTIN_INTEGRATE: entry = &dev::builtin_integrate; break;
        case FIMO_B200_BUILTIN_PROPAGATE: entry = &dev::builtin_propagate; break;
        case FIMO_B200_BUILTIN_CULL: entry = &dev::builtin_cull; break;
        case FIMO_B200_BUILTIN_FORCES: entry = &dev::builtin_forces; break;
        case FIMO_B200_BUILTIN_INTEGRATE_ANG: entry = &dev::builtin_integrate_ang; break;
        case FIMO_B200_BUILTIN_WORLD_AABB: entry = &dev::builtin_world_aabb; break;
        case FIMO_B200_BUILTIN_CULL_WAABB: entry = &dev::builtin_cull_waabb; break;
        case FIMO_B200_BUILTIN_BOUNDS_REDUCE: entry = &dev::builtin_bounds_reduce; break;
        case FIMO_B200_BUILTIN_GROUP_STATS: entry = &dev::builtin_group_stats; break;
        case FIMO_B200_BUILTIN_PREV_COPY: entry = &dev::builtin_prev_copy; break;
        case FIMO_B200_BUILTIN_LOD: entry = &dev::builtin_lod; break;
        case FIMO_B200_BUILTIN_VEL_CLAMP: entry = &dev::builtin_vel_clamp; break;
        case FIMO_B200_BUILTIN_SLEEP_FLAGS: entry = &dev::builtin_sleep_flags; break;
        case FIMO_B200_BUILTIN_ANGVEL_DAMP: entry = &dev::builtin_angvel_damp; break;
        case FIMO_B200_BUILTIN_VIS_BOUNDS: entry = &dev::builtin_vis_bounds; break;
        default: return report("NotFound");
    }
    *ext = FimoB200_SystemDeviceExt{};
    ext->header.s_type = FIMO_B200_STYPE_SYSTEM_DEVICE;
    ext->header.next = desc->next;
    ext->device_entry = entry;
    ext->builtin_id = builtin_id;
    desc->next = ext;
    desc->system_size = sizeof(dev::BuiltinState);
    desc->system_alignment = alignof(dev::BuiltinState);
    desc->system_create = &builtin_create;
    desc->system_destroy = &builtin_destroy;
    desc->system_run = &builtin_run_stub;
    return FIMO_STATUS_OK;
}

FimoStatus fimo_b200_resource_upload(const FimoB200_ResourceView *view, const void *host_src) {
    if (!view || !view->device_ptr || !host_src) return report("InvalidArgument");
    // the device is implied by the pointer; cudaMemcpy resolves it through UVA
    return dev::hbm_upload(dev::resolve_device(-1), view->device_ptr, host_src, view->world_stride * view->n_worlds)
                   ? FIMO_STATUS_OK
                   : report("DeviceError");
}
FimoStatus fimo_b200_resource_download(const FimoB200_ResourceView *view, void *host_dst) {
    if (!view || !view->device_ptr || !host_dst) return report("InvalidArgument");
    return dev::hbm_download(dev::resolve_device(-1), host_dst, view->device_ptr, view->world_stride * view->n_worlds)
                   ? FIMO_STATUS_OK
                   : report("DeviceError");
}

FimoStatus fimo_b200_group_transfer(FimoWorldsMeta_SystemGroup *group, FimoWorldsMeta_ResourceHandle handle, void *host,
                                    uint32_t direction, FimoWorldsMeta_Fence *signal) {
    Group *g = (Group *)group;
    if (!g || !handle || !host || direction > FIMO_B200_TRANSFER_DEVICE_TO_HOST) return report("InvalidArgument");
    {
        std::lock_guard<std::mutex> lock(g->m);
        const char *err = nullptr;
        if (g->dirty && g->in_flight == 0 && g->backlog.empty() && !g->dispatcher_busy && !compile(g, &err)) return report(err);
        if (!g->has_device_systems) return report("NoDevice");
        if (std::find(g->resources.begin(), g->resources.end(), (ResourceType *)handle) == g->resources.end())
            return report("NotFound"); // only resources the group's systems use ride on its stream
    }
    const Transfer xfer{(ResourceType *)handle, host, direction == FIMO_B200_TRANSFER_HOST_TO_DEVICE};
    if (const char *err = group_schedule(g, nullptr, 0, signal, &xfer)) return report(err);
    return FIMO_STATUS_OK;
}
FimoStatus fimo_b200_group_schedule_io(FimoWorldsMeta_SystemGroup *group, const FimoB200_StepTransfer *uploads, FimoUSize n_uploads,
                                       const FimoB200_StepTransfer *downloads, FimoUSize n_downloads,
                                       FimoWorldsMeta_Fence *const *wait_on, FimoUSize n_wait, FimoWorldsMeta_Fence *signal) {
    Group *g = (Group *)group;
    if ((n_uploads && !uploads) || (n_downloads && !downloads)) return report("InvalidArgument");
    for (size_t i = 0; i < n_uploads; i++)
        if (!uploads[i].resource || !uploads[i].host) return report("InvalidArgument");
    for (size_t i = 0; i < n_downloads; i++)
        if (!downloads[i].resource || !downloads[i].host) return report("InvalidArgument");
    {
        std::lock_guard<std::mutex> lock(g->m);
        const char *err = nullptr;
        if (g->dirty && g->in_flight == 0 && g->backlog.empty() && !g->dispatcher_busy && !compile(g, &err)) return report(err); // never under a running generation
        if (!g->has_device_systems) return report("NoDevice");
        auto used = [&](FimoWorldsMeta_ResourceHandle h) {
            return std::find(g->resources.begin(), g->resources.end(), (ResourceType *)h) != g->resources.end();
        };
        for (size_t i = 0; i < n_uploads; i++)
            if (!used(uploads[i].resource)) return report("NotFound");
        for (size_t i = 0; i < n_downloads; i++)
            if (!used(downloads[i].resource)) return report("NotFound");
    }
    const StepIo io{uploads, n_uploads, downloads, n_downloads};
    if (const char *err = group_schedule(g, wait_on, n_wait, signal, nullptr, &io)) return report(err);
    return FIMO_STATUS_OK;
}
FimoStatus fimo_b200_group_schedule_readback(FimoWorldsMeta_SystemGroup *group, const FimoB200_StepTransfer *uploads, FimoUSize n_uploads,
                                             const FimoB200_CullReadback *readback, FimoWorldsMeta_Fence *const *wait_on, FimoUSize n_wait,
                                             FimoWorldsMeta_Fence *signal) {
    Group *g = (Group *)group;
    if ((n_uploads && !uploads) || !readback || !readback->host || !readback->vis_mask || !readback->vis_list || !readback->vis_count)
        return report("InvalidArgument");
    for (size_t i = 0; i < n_uploads; i++)
        if (!uploads[i].resource || !uploads[i].host) return report("InvalidArgument");
    {
        std::lock_guard<std::mutex> lock(g->m);
        const char *err = nullptr;
        if (g->dirty && g->in_flight == 0 && g->backlog.empty() && !g->dispatcher_busy && !compile(g, &err)) return report(err);
        if (!g->has_device_systems) return report("NoDevice");
        auto used = [&](FimoWorldsMeta_ResourceHandle h) {
            return std::find(g->resources.begin(), g->resources.end(), (ResourceType *)h) != g->resources.end();
        };
        for (size_t i = 0; i < n_uploads; i++)
            if (!used(uploads[i].resource)) return report("NotFound");
        if (!used(readback->vis_mask) || !used(readback->vis_list) || !used(readback->vis_count)) return report("NotFound");
        const uint64_t W = g->world->n_worlds, words = (g->world->entities + 31) / 32;
        if (readback->host_bytes < 4 * (8 + 2 * W + W * words)) return report("InvalidArgument"); // header + counts + offsets + masks must fit
    }
    const StepIo io{uploads, n_uploads, nullptr, 0, readback};
    if (const char *err = group_schedule(g, wait_on, n_wait, signal, nullptr, &io)) return report(err);
    return FIMO_STATUS_OK;
}
void fimo_worlds_b200_set_trace_sink(FimoB200_trace_sink sink, void *user) {
    g_trace_user.store(user, std::memory_order_release);
    g_trace_sink.store(sink, std::memory_order_release);
}
void *fimo_b200_host_alloc(FimoUSize bytes) { return dev::host_alloc_pinned(bytes); }
void fimo_b200_host_free(void *ptr) { dev::host_free_pinned(ptr); }

FimoStatus fimo_b200_group_executor_info(FimoWorldsMeta_SystemGroup *group, FimoB200_ExecutorInfo *info) {
    Group *g = (Group *)group;
    std::lock_guard<std::mutex> lock(g->m);
    const char *err = nullptr;
    if (g->dirty && g->in_flight == 0 && g->backlog.empty() && !g->dispatcher_busy && !compile(g, &err)) return report(err); // never under a running generation
    *info = g->info;
    return FIMO_STATUS_OK;
}
// ---- multi-GPU statistics exchange ------------------------------------------------------------------------------
static dev::StatsExchange *stats_exchange_of(Group *g) { // g->m held
    if (!g->stats_x && g->world->device >= 0) g->stats_x = dev::stats_exchange_create(g->world->device);
    return g->stats_x;
}
FimoStatus fimo_b200_group_stats_peers_open(FimoWorldsMeta_SystemGroup *group, uint32_t slots, void *handle_out) {
    Group *g = (Group *)group;
    std::lock_guard<std::mutex> lock(g->m);
    dev::StatsExchange *x = stats_exchange_of(g);
    if (!x) return report("NoDevice");
    return dev::stats_exchange_peers_open(x, slots, handle_out) ? FIMO_STATUS_OK : report("DeviceError");
}
FimoStatus fimo_b200_group_stats_peers_bind(FimoWorldsMeta_SystemGroup *group, const void *handles, uint32_t rank, uint32_t nranks) {
    Group *g = (Group *)group;
    std::lock_guard<std::mutex> lock(g->m);
    if (!g->stats_x) return report("NotFound"); // peers_open first
    if (g->in_flight != 0 || !g->backlog.empty()) return report("InUse");
    if (!dev::stats_exchange_peers_bind(g->stats_x, handles, rank, nranks)) return report("DeviceError");
    g->dirty = true; // the executor picks the peer table up when it is (re)compiled
    return FIMO_STATUS_OK;
}
FimoStatus fimo_b200_stats_nccl_unique_id(void *id_out) { return dev::stats_nccl_unique_id(id_out) ? FIMO_STATUS_OK : report("DeviceError"); }
FimoStatus fimo_b200_group_stats_nccl_bind(FimoWorldsMeta_SystemGroup *group, const void *unique_id, uint32_t rank, uint32_t nranks,
                                           uint32_t bucket) {
    Group *g = (Group *)group;
    std::lock_guard<std::mutex> lock(g->m);
    dev::StatsExchange *x = stats_exchange_of(g);
    if (!x) return report("NoDevice");
    if (g->in_flight != 0 || !g->backlog.empty()) return report("InUse");
    if (g->stats_ring) return report("Duplicate"); // the caller already supplied its own ring
    if (!dev::stats_exchange_nccl_bind(x, unique_id, rank, nranks, bucket)) return report("DeviceError");
    g->stats_ring = dev::stats_exchange_ring(x, &g->stats_ring_slots);
    g->dirty = true;
    return FIMO_STATUS_OK;
}
FimoStatus fimo_b200_group_global_stats(FimoWorldsMeta_SystemGroup *group, uint64_t generation, uint32_t mode, uint64_t *out_words) {
    Group *g = (Group *)group;
    dev::StatsExchange *x;
    {
        std::lock_guard<std::mutex> lock(g->m);
        x = g->stats_x;
    }
    if (!x) return report("NotFound");
    return dev::stats_exchange_read(x, generation, mode, out_words) ? FIMO_STATUS_OK : report("DeviceError");
}

FimoUSize fimo_b200_group_executor_describe(FimoWorldsMeta_SystemGroup *group, char *buffer, FimoUSize buffer_len) {
    Group *g = (Group *)group;
    std::lock_guard<std::mutex> lock(g->m);
    if (buffer && buffer_len) {
        const size_t n = std::min<size_t>(g->plan_text.size(), buffer_len - 1);
        std::memcpy(buffer, g->plan_text.data(), n);
        buffer[n] = 0;
    }
    return g->plan_text.size();
}

uint64_t fimo_b200_group_kernels_launched(FimoWorldsMeta_SystemGroup *group) {
    Group *g = (Group *)group;
    std::lock_guard<std::mutex> lock(g->m);
    return g->executor_dev ? dev::executor_kernels_launched(g->executor_dev) : 0;
}

void *fimo_b200_group_stream(FimoWorldsMeta_SystemGroup *group) {
    Group *g = (Group *)group;
    std::lock_guard<std::mutex> lock(g->m);
    return g->executor_dev ? dev::executor_stream(g->executor_dev) : nullptr;
}
void fimo_b200_group_wait_idle(FimoWorldsMeta_SystemGroup *group) {
    Group *g = (Group *)group;
    FimoWorldsMeta_TimelineSemaphore *sem = &g->schedule_semaphore;
    size_t target;
    {
        std::lock_guard<std::mutex> lock(g->m);
        target = g->next_generation;
    }
    FimoWorldsMeta_timeline_semaphore_wait(sem, (uint64_t)target, g_futex_wait);
}
double fimo_b200_group_device_ms(FimoWorldsMeta_SystemGroup *group, uint64_t *generations) {
    Group *g = (Group *)group;
    dev::Executor *ex;
    {
        std::lock_guard<std::mutex> lock(g->m);
        ex = g->executor_dev;
    }
    if (!ex) {
        if (generations) *generations = 0;
        return 0.0;
    }
    dev::executor_wait_idle(ex);
    return dev::executor_device_ms(ex, generations);
}
double fimo_b200_group_span_ms(FimoWorldsMeta_SystemGroup *group) {
    Group *g = (Group *)group;
    dev::Executor *ex;
    {
        std::lock_guard<std::mutex> lock(g->m);
        ex = g->executor_dev;
    }
    if (!ex) return 0.0;
    dev::executor_wait_idle(ex);
    return dev::executor_span_ms(ex);
}
}
