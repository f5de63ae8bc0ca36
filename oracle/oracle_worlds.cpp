// oracle_worlds.cpp — CPU restatement of modules/fimo_worlds (TEST INFRASTRUCTURE ONLY).
//
// Universe  <- modules/fimo_worlds/src/Universe.zig      (registry, validation, cycle check)
// World     <- modules/fimo_worlds/src/World.zig         (resource blobs, per-resource RwLock)
// Group     <- modules/fimo_worlds/src/SystemGroup.zig   (DAG -> command buffer, generation loop)
// Context   <- modules/fimo_worlds/src/SystemContext.zig (per-system instance, run trampoline)
// exports   <- modules/fimo_worlds/src/FimoWorlds.zig:108-447
//
// Deliberate, documented differences from the reference:
//  * Order of systems inside one DAG generation: the reference uses an unstable sort
//    (SystemGroup.zig:262) => unspecified; the oracle uses a stable sort, i.e. the
//    tie-break is toposort emission order (= add_systems order). The product uses the same rule.
//  * The group job does not block a worker while its systems run: `run` is split into
//    run_begin (task) / run_end (command buffer on_complete) because oracle tasks are OS
//    threads, not fibers (see oracle_pool.cpp). Observable ordering is unchanged: generation
//    g+1 starts after g ended; `signal` fires after the generation ended.
//  * toposort records the systems it has visited (markers.put below).  The reference's sort reads `markers` but
//    never writes it (SystemGroup.zig:214-246), so a system reached over two dependency paths would be appended
//    twice into a task list sized for one entry per system; the contract (every system once per generation) is
//    what is restated here.
//  * MultiGenerationAllocator follows its documentation (4-generation ring), not the `& 4`
//    typo of heap/MultiGenerationAllocator.zig:33.
#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <unordered_set>
#include <vector>

#include "oracle_internal.h"

namespace oracle {

// Insertion-ordered map with swap-remove, the two properties of std.AutoArrayHashMapUnmanaged
// that the reference's iteration orders depend on.
template <typename K, typename V> struct ArrayMap {
    std::vector<K> keys;
    std::vector<V> values;
    size_t count() const { return keys.size(); }
    ptrdiff_t index_of(const K &k) const {
        for (size_t i = 0; i < keys.size(); i++)
            if (keys[i] == k) return (ptrdiff_t)i;
        return -1;
    }
    bool contains(const K &k) const { return index_of(k) >= 0; }
    V *get(const K &k) {
        ptrdiff_t i = index_of(k);
        return i < 0 ? nullptr : &values[(size_t)i];
    }
    void put(const K &k, const V &v) {
        ptrdiff_t i = index_of(k);
        if (i >= 0) values[(size_t)i] = v;
        else {
            keys.push_back(k);
            values.push_back(v);
        }
    }
    bool swap_remove(const K &k) {
        ptrdiff_t i = index_of(k);
        if (i < 0) return false;
        keys[(size_t)i] = keys.back();
        values[(size_t)i] = values.back();
        keys.pop_back();
        values.pop_back();
        return true;
    }
    void clear() {
        keys.clear();
        values.clear();
    }
};

static void *aligned_new(size_t size, size_t alignment) {
    if (alignment < sizeof(void *)) alignment = sizeof(void *);
    void *p = nullptr;
    if (posix_memalign(&p, alignment, size ? size : 1) != 0) return nullptr;
    return p;
}

// Arena: stand-in for std.heap.ArenaAllocator (blocks freed together).
struct Arena {
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
    ~Arena() { reset(); }
};

// Tracked allocator: stand-in for heap/TracingAllocator.zig (frees leftovers on deinit).
struct Tracked {
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
            live.erase(p);
        }
        free(p);
    }
    void reset() {
        std::lock_guard<std::mutex> g(m);
        for (void *p : live) free(p);
        live.clear();
    }
    ~Tracked() { reset(); }
};

// ---------------------------------------------------------------------------------
// Universe (Universe.zig)
// ---------------------------------------------------------------------------------
struct Resource { // Universe.zig:57-62
    std::string label;
    size_t size;
    size_t alignment;
    std::atomic<size_t> references{0};
};

struct Link { // Universe.zig:96-101
    bool implicit;
    bool weak;
    bool ignore_deferred;
};

struct SystemContext;
struct System { // Universe.zig:64-106
    std::string label;
    std::vector<Resource *> exclusive_resources;
    std::vector<Resource *> shared_resources;
    std::atomic<size_t> references{0};
    RwLock rwlock;
    size_t static_before_count = 0;
    size_t static_after_count = 0;
    ArrayMap<System *, Link> before; // systems that run BEFORE this one
    ArrayMap<System *, Link> after;  // systems that run AFTER this one
    void *factory = nullptr;
    size_t factory_size = 0;
    void (*factory_destroy)(const void *) = nullptr;
    size_t system_size = 0;
    size_t system_alignment = 1;
    bool (*system_create)(const void *, FimoWorldsMeta_SystemContext *, void *) = nullptr;
    void (*system_destroy)(void *) = nullptr;
    void (*system_run)(void *, void *const *, void *const *, FimoWorldsMeta_Fence *) = nullptr;
};

struct Universe {
    RwLock rwlock;
    size_t num_worlds = 0;
    ArrayMap<Resource *, char> resources;
    ArrayMap<System *, char> systems;
};
static Universe g_universe;

struct Err {
    const char *name;
};

static Resource *universe_register_resource(const FimoWorldsMeta_ResourceDescriptor *d) {
    // Universe.zig:131-150
    g_universe.rwlock.lock_write();
    Resource *r = new Resource();
    r->label = d->label_len ? std::string(d->label, d->label_len) : std::string("<unlabelled>");
    r->size = d->size;
    r->alignment = d->alignment;
    g_universe.resources.put(r, 0);
    g_universe.rwlock.unlock_write();
    return r;
}

static void universe_unregister_resource(Resource *r) {
    // Universe.zig:152-163
    g_universe.rwlock.lock_write();
    if (!g_universe.resources.swap_remove(r)) panic("invalid resource");
    if (r->references.load(std::memory_order_acquire) != 0) panic("resource in use");
    delete r;
    g_universe.rwlock.unlock_write();
}

template <typename T> static bool contains_from(const T *arr, size_t from, size_t len, T v) {
    for (size_t i = from; i < len; i++)
        if (arr[i] == v) return true;
    return false;
}

// Universe.zig:165-334. Returns nullptr and sets *err on failure.
static System *universe_register_system(const FimoWorldsMeta_SystemDescriptor *d, const char **err) {
    struct FactoryGuard { // errdefer factory_deinit (Universe.zig:166-169)
        const FimoWorldsMeta_SystemDescriptor *d;
        bool armed = true;
        ~FactoryGuard() {
            if (armed && d->factory_destroy) d->factory_destroy(d->factory_size ? d->factory : nullptr);
        }
    } fguard{d};

    g_universe.rwlock.lock_write();
    struct Unlock {
        ~Unlock() { g_universe.rwlock.unlock_write(); }
    } unlock;

    Resource *const *excl = (Resource *const *)d->exclusive_handles;
    Resource *const *shar = (Resource *const *)d->shared_handles;
    const size_t ne = d->exclusive_handles_len, ns = d->shared_handles_len;
    for (size_t i = 0; i < ne; i++) { // :174-178
        if (!g_universe.resources.contains(excl[i])) return *err = "NotFound", nullptr;
        if (contains_from(excl, i + 1, ne, excl[i])) return *err = "Deadlock", nullptr;
    }
    for (size_t i = 0; i < ns; i++) { // :179-185
        if (!g_universe.resources.contains(shar[i])) return *err = "NotFound", nullptr;
        if (contains_from(excl, 0, ne, shar[i])) return *err = "Deadlock", nullptr;
        if (contains_from(shar, i + 1, ns, shar[i])) return *err = "Duplicate", nullptr;
    }

    ArrayMap<System *, Link> before, after;
    for (size_t i = 0; i < d->before_len; i++) { // :195-206
        System *dep = (System *)d->before[i].system;
        if (!g_universe.systems.contains(dep)) return *err = "NotFound", nullptr;
        if (before.contains(dep)) return *err = "Duplicate", nullptr;
        before.put(dep, Link{false, (d->before[i].flags & FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_WEAK) != 0,
                             (d->before[i].flags & FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_IGNORE_DEFERRED) != 0});
    }
    for (size_t i = 0; i < d->after_len; i++) { // :207-219
        System *dep = (System *)d->after[i].system;
        if (!g_universe.systems.contains(dep)) return *err = "NotFound", nullptr;
        if (before.contains(dep)) return *err = "Deadlock", nullptr;
        if (after.contains(dep)) return *err = "Duplicate", nullptr;
        after.put(dep, Link{false, (d->after[i].flags & FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_WEAK) != 0,
                            (d->after[i].flags & FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_IGNORE_DEFERRED) != 0});
    }

    { // cycle check, :221-240: walk explicit `after` links from our successors; reaching a predecessor is a cycle
        std::vector<System *> stack(after.keys);
        std::vector<System *> visited(after.keys);
        while (!stack.empty()) {
            System *s = stack.back();
            stack.pop_back();
            if (before.contains(s)) return *err = "Deadlock", nullptr;
            s->rwlock.lock_read();
            for (size_t i = 0; i < s->after.count(); i++) {
                System *n = s->after.keys[i];
                if (s->after.values[i].implicit) continue;
                if (std::find(visited.begin(), visited.end(), n) != visited.end()) continue;
                visited.push_back(n);
                stack.push_back(n);
            }
            s->rwlock.unlock_read();
        }
    }

    System *sys = new System();
    sys->label = d->label_len ? std::string(d->label, d->label_len) : std::string("<unlabelled>");
    sys->exclusive_resources.assign(excl, excl + ne);
    sys->shared_resources.assign(shar, shar + ns);
    sys->static_before_count = before.count();
    sys->static_after_count = after.count();
    sys->before = before;
    sys->after = after;
    sys->factory_destroy = d->factory_destroy;
    sys->system_size = d->system_size;
    sys->system_alignment = d->system_alignment ? d->system_alignment : 1;
    sys->system_create = d->system_create;
    sys->system_destroy = d->system_destroy;
    sys->system_run = d->system_run;

    // implicit reverse links, :284-309
    for (size_t i = 0; i < before.count(); i++) {
        System *s2 = before.keys[i];
        s2->rwlock.lock_write();
        s2->after.put(sys, Link{true, before.values[i].weak, before.values[i].ignore_deferred});
        s2->rwlock.unlock_write();
    }
    for (size_t i = 0; i < after.count(); i++) {
        System *s2 = after.keys[i];
        s2->rwlock.lock_write();
        s2->before.put(sys, Link{true, after.values[i].weak, after.values[i].ignore_deferred});
        s2->rwlock.unlock_write();
    }

    if (d->factory_size != 0) { // factory bytes are copied, :311-320
        sys->factory = aligned_new(d->factory_size, d->factory_alignment);
        std::memcpy(sys->factory, d->factory, d->factory_size);
        sys->factory_size = d->factory_size;
    }
    g_universe.systems.put(sys, 0);
    for (Resource *r : sys->exclusive_resources) r->references.fetch_add(1, std::memory_order_relaxed); // :323-330
    for (Resource *r : sys->shared_resources) r->references.fetch_add(1, std::memory_order_relaxed);
    fguard.armed = false;
    return sys;
}

static void universe_unregister_system(System *sys) {
    // Universe.zig:336-381
    g_universe.rwlock.lock_write();
    if (!g_universe.systems.swap_remove(sys)) panic("invalid system");
    if (sys->before.count() != sys->static_before_count) panic("system referenced");
    if (sys->after.count() != sys->static_after_count) panic("system referenced");
    if (sys->references.load(std::memory_order_acquire) != 0) panic("system in use");
    if (sys->factory_destroy && sys->factory) sys->factory_destroy(sys->factory);
    for (Resource *r : sys->exclusive_resources) r->references.fetch_sub(1, std::memory_order_relaxed);
    for (Resource *r : sys->shared_resources) r->references.fetch_sub(1, std::memory_order_relaxed);
    for (size_t i = 0; i < sys->before.count(); i++) {
        if (sys->before.values[i].implicit) panic("system referenced");
        System *s2 = sys->before.keys[i];
        s2->rwlock.lock_write();
        s2->after.swap_remove(sys);
        s2->rwlock.unlock_write();
    }
    for (size_t i = 0; i < sys->after.count(); i++) {
        if (sys->after.values[i].implicit) panic("system referenced");
        System *s2 = sys->after.keys[i];
        s2->rwlock.lock_write();
        s2->before.swap_remove(sys);
        s2->rwlock.unlock_write();
    }
    free(sys->factory);
    delete sys;
    g_universe.rwlock.unlock_write();
}

// ---------------------------------------------------------------------------------
// World (World.zig)
// ---------------------------------------------------------------------------------
struct ResourceValue { // World.zig:28-32
    RwLock rwlock;
    void *value_ptr = nullptr;
    std::atomic<size_t> references{0};
};

struct World {
    RwLock rwlock;
    std::string label;
    FimoTasksMeta_Pool executor{};
    bool inherited_executor = false;
    Tracked allocator;
    std::atomic<size_t> system_group_count{0};
    ArrayMap<Resource *, ResourceValue *> resources;
};

static World *world_init(const FimoWorldsMeta_WorldDescriptor *d) {
    // World.zig:39-71
    World *w = new World();
    w->label = d->label_len ? std::string(d->label, d->label_len) : std::string("<unlabelled>");
    if (d->pool) {
        w->executor = *d->pool;
        w->executor.vtable->acquire(w->executor.data);
        w->inherited_executor = true;
    } else {
        std::string l = "world `" + w->label + "` executor";
        w->executor = pool_handle(pool_create(l.data(), l.size(), 0)); // one pool per world, #CPU workers (:45-52)
        w->inherited_executor = false;
    }
    g_universe.rwlock.lock_write();
    g_universe.num_worlds++;
    g_universe.rwlock.unlock_write();
    return w;
}

static void world_deinit(World *w) {
    // World.zig:71-82
    if (w->system_group_count.load(std::memory_order_acquire) != 0) panic("world not empty");
    if (w->resources.count() != 0) panic("world not empty");
    if (!w->inherited_executor) w->executor.vtable->request_close(w->executor.data);
    w->executor.vtable->release(w->executor.data);
    g_universe.rwlock.lock_write();
    g_universe.num_worlds--;
    g_universe.rwlock.unlock_write();
    delete w;
}

static const char *world_add_resource(World *w, Resource *h, const void *value) {
    // World.zig:84-112: the blob is COPIED into a world-owned allocation.
    g_universe.rwlock.lock_read();
    if (!g_universe.resources.contains(h)) panic("invalid resource");
    h->references.fetch_add(1, std::memory_order_relaxed);
    g_universe.rwlock.unlock_read();

    ResourceValue *rv = new ResourceValue();
    rv->value_ptr = aligned_new(h->size, h->alignment);
    std::memcpy(rv->value_ptr, value, h->size);

    w->rwlock.lock_write();
    if (w->resources.contains(h)) {
        w->rwlock.unlock_write();
        free(rv->value_ptr);
        delete rv;
        h->references.fetch_sub(1, std::memory_order_relaxed);
        return "Duplicate";
    }
    w->resources.put(h, rv);
    w->rwlock.unlock_write();
    return nullptr;
}

static const char *world_remove_resource(World *w, Resource *h, void *value) {
    // World.zig:114-137: the blob is copied back out.
    w->rwlock.lock_write();
    ResourceValue **slot = w->resources.get(h);
    if (!slot) panic("invalid resource");
    ResourceValue *rv = *slot;
    if (rv->references.load(std::memory_order_acquire) != 0) {
        w->rwlock.unlock_write();
        return "InUse";
    }
    w->resources.swap_remove(h);
    w->rwlock.unlock_write();
    std::memcpy(value, rv->value_ptr, h->size);
    free(rv->value_ptr);
    delete rv;
    h->references.fetch_sub(1, std::memory_order_relaxed);
    return nullptr;
}

static void world_lock_resources(World *w, Resource *const *exclusive, size_t ne, Resource *const *shared, size_t ns,
                                 void **out) {
    // World.zig:145-198: validate, then lock in ascending handle-address order.
    struct Info {
        size_t index;
        Resource *handle;
        ResourceValue *resource;
        bool exclusive;
    };
    std::vector<Info> infos(ne + ns);
    w->rwlock.lock_read();
    for (size_t i = 0; i < ne; i++) {
        if (contains_from(exclusive, i + 1, ne, exclusive[i])) panic("deadlock");
        ResourceValue **rv = w->resources.get(exclusive[i]);
        if (!rv) panic("invalid resource");
        (*rv)->references.fetch_add(1, std::memory_order_relaxed);
        infos[i] = Info{i, exclusive[i], *rv, true};
    }
    for (size_t i = 0; i < ns; i++) {
        if (contains_from(shared, i + 1, ns, shared[i])) panic("duplicate");
        if (contains_from(exclusive, 0, ne, shared[i])) panic("deadlock");
        ResourceValue **rv = w->resources.get(shared[i]);
        if (!rv) panic("invalid resource");
        (*rv)->references.fetch_add(1, std::memory_order_relaxed);
        infos[ne + i] = Info{ne + i, shared[i], *rv, false};
    }
    w->rwlock.unlock_read();
    std::stable_sort(infos.begin(), infos.end(),
                     [](const Info &a, const Info &b) { return (uintptr_t)a.handle < (uintptr_t)b.handle; });
    for (const Info &info : infos) {
        if (info.exclusive) info.resource->rwlock.lock_write();
        else info.resource->rwlock.lock_read();
        out[info.index] = info.resource->value_ptr;
    }
}

static void world_unlock_resource(World *w, Resource *h, bool exclusive) {
    // World.zig:200-216
    w->rwlock.lock_read();
    ResourceValue **rv = w->resources.get(h);
    if (!rv) panic("invalid resource");
    ResourceValue *r = *rv;
    w->rwlock.unlock_read();
    r->references.fetch_sub(1, std::memory_order_relaxed);
    if (exclusive) r->rwlock.unlock_write();
    else r->rwlock.unlock_read();
}

// ---------------------------------------------------------------------------------
// SystemContext (SystemContext.zig)
// ---------------------------------------------------------------------------------
struct SystemGroup;

struct SystemContext {
    SystemGroup *group = nullptr;
    System *sys = nullptr;
    bool weak = false;
    size_t reference_count = 0;
    void *value_ptr = nullptr;
    bool merge_deferred = false;
    FimoWorldsMeta_Fence deferred_fence{FIMO_WORLDS_META_FENCE_UNSIGNALED};
    std::vector<SystemContext *> deferred_dep;
    std::vector<void *> resources;
    Arena arena;                                   // transient: reset after every run (:96)
    Tracked persistent;                            // system_persistent: freed with the context
    std::vector<FimoWorldsMeta_Fence *> waiters;   // fences to signal on deinit (:111-116)
};

static SystemContext *context_init(SystemGroup *group, System *sys, bool weak) {
    // SystemContext.zig:42-74
    SystemContext *ctx = new SystemContext();
    ctx->group = group;
    ctx->sys = sys;
    ctx->weak = weak;
    ctx->value_ptr = aligned_new(sys->system_size, sys->system_alignment);
    sys->references.fetch_add(1, std::memory_order_relaxed);
    if (!sys->system_create(sys->factory, (FimoWorldsMeta_SystemContext *)ctx, ctx->value_ptr)) {
        sys->references.fetch_sub(1, std::memory_order_relaxed);
        free(ctx->value_ptr);
        delete ctx;
        return nullptr; // error.InitError
    }
    return ctx;
}

static void context_deinit(SystemContext *ctx) {
    // SystemContext.zig:76-89
    if (ctx->sys->system_destroy) ctx->sys->system_destroy(ctx->value_ptr);
    ctx->sys->references.fetch_sub(1, std::memory_order_relaxed);
    for (FimoWorldsMeta_Fence *f : ctx->waiters) fence_signal(f);
    free(ctx->value_ptr);
    delete ctx;
}

static void context_run(SystemContext *ctx) {
    // SystemContext.zig:91-100
    FimoWorldsMeta_fence_reset(&ctx->deferred_fence);
    for (SystemContext *dep : ctx->deferred_dep) fence_wait(&dep->deferred_fence);
    void *const *exclusive = ctx->resources.data();
    void *const *shared = ctx->resources.data() + ctx->sys->exclusive_resources.size();
    ctx->sys->system_run(ctx->value_ptr, exclusive, shared, &ctx->deferred_fence);
    ctx->arena.reset();
    if (ctx->merge_deferred) fence_wait(&ctx->deferred_fence);
}

// ---------------------------------------------------------------------------------
// SystemGroup (SystemGroup.zig)
// ---------------------------------------------------------------------------------
struct TaskInfo { // SystemGroup.zig:196-247
    SystemContext *ctx;
    size_t generation;
    FimoTasksMeta_Task task;
};

struct SystemGroup {
    std::string label;
    World *world = nullptr;
    FimoTasksMeta_Pool executor{};
    std::atomic<size_t> generation{0};
    // Graph (SystemGroup.zig:40-52)
    std::mutex mutex;
    bool dirty = true;
    size_t next_generation = 0;
    FimoWorldsMeta_TimelineSemaphore schedule_semaphore{0};
    std::vector<FimoTasksMeta_CommandBufferEntry> entries;
    ArrayMap<Resource *, void *> resources;
    ArrayMap<System *, SystemContext *> systems;
    std::vector<SystemContext *> deinit_list;
    std::vector<TaskInfo> tasks;
    Arena single_allocator;      // heap/SingleGenerationAllocator.zig
    Arena multi_allocator[4];    // heap/MultiGenerationAllocator.zig (documented 4-ring)
    std::string systems_label;
};

static void graph_clear_deinit_list(SystemGroup *g) { // :54-59
    for (SystemContext *c : g->deinit_list) context_deinit(c);
    g->deinit_list.clear();
}

static void graph_remove_system(SystemGroup *g, System *sys, FimoWorldsMeta_Fence *fence, bool allow_deferred);

static SystemContext *graph_add_system(SystemGroup *g, System *sys, bool weak, const char **err) {
    // SystemGroup.zig:61-115
    if (SystemContext **existing = g->systems.get(sys)) {
        if (!weak) (*existing)->weak = false;
        return *existing;
    }
    sys->rwlock.lock_read();
    std::vector<SystemContext *> taken; // dependencies whose refcount we raised (errdefer, :70-92)
    auto rollback = [&]() {
        for (SystemContext *c : taken) {
            c->reference_count -= 1;
            if (c->reference_count == 0 && c->weak) graph_remove_system(g, c->sys, nullptr, false);
        }
        sys->rwlock.unlock_read();
    };
    for (int pass = 0; pass < 2; pass++) { // `before` then `after`, :94-107
        ArrayMap<System *, Link> &links = pass == 0 ? sys->before : sys->after;
        for (size_t i = 0; i < links.count(); i++) {
            if (links.values[i].implicit || links.values[i].weak) continue;
            SystemContext *c = graph_add_system(g, links.keys[i], true, err);
            if (!c) return rollback(), nullptr;
            c->reference_count += 1;
            taken.push_back(c);
        }
    }
    SystemContext *ctx = context_init(g, sys, weak);
    if (!ctx) {
        *err = "InitError";
        return rollback(), nullptr;
    }
    g->systems.put(sys, ctx);
    g->dirty = true;
    sys->rwlock.unlock_read();
    return ctx;
}

static void graph_remove_system(SystemGroup *g, System *sys, FimoWorldsMeta_Fence *fence, bool allow_deferred) {
    // SystemGroup.zig:117-158
    SystemContext **slot = g->systems.get(sys);
    if (!slot) return;
    SystemContext *first = *slot;
    g->systems.swap_remove(sys);

    std::vector<SystemContext *> queue{first}, deinit_stack;
    for (size_t qi = 0; qi < queue.size(); qi++) {
        SystemContext *ctx = queue[qi];
        ctx->sys->rwlock.lock_read();
        for (int pass = 0; pass < 2; pass++) {
            ArrayMap<System *, Link> &links = pass == 0 ? ctx->sys->before : ctx->sys->after;
            for (size_t i = 0; i < links.count(); i++) {
                if (links.values[i].implicit || links.values[i].weak) continue;
                SystemContext **dep = g->systems.get(links.keys[i]);
                if (!dep) continue;
                (*dep)->reference_count -= 1;
                if ((*dep)->reference_count == 0 && (*dep)->weak) {
                    queue.push_back(*dep);
                    g->systems.swap_remove(links.keys[i]);
                }
            }
        }
        ctx->sys->rwlock.unlock_read();
        deinit_stack.push_back(ctx);
    }
    while (!deinit_stack.empty()) {
        SystemContext *ctx = deinit_stack.back();
        deinit_stack.pop_back();
        const bool running =
                !FimoWorldsMeta_timeline_semaphore_is_signaled(&g->schedule_semaphore, (uint64_t)g->next_generation);
        if (allow_deferred && running) { // :146-150: a generation is in flight, destroy at the next recompile
            if (ctx->sys == sys && fence) ctx->waiters.push_back(fence);
            g->deinit_list.push_back(ctx);
        } else {
            graph_clear_deinit_list(g);
            const bool is_target = ctx->sys == sys;
            context_deinit(ctx);
            if (is_target && fence) fence_signal(fence);
        }
    }
    g->dirty = true;
}

static void task_start(FimoTasksMeta_Task *task) { // TaskInfo.taskStart, :203-207
    TaskInfo *info = (TaskInfo *)((char *)task - offsetof(TaskInfo, task));
    context_run(info->ctx);
}

static size_t toposort(SystemGroup *g, std::vector<TaskInfo> &tasks, ArrayMap<System *, size_t> &markers, System *handle,
                       SystemContext *ctx) {
    // SystemGroup.zig:214-246: generation = 1 + max(generation of predecessors present in the group)
    if (size_t *idx = markers.get(handle)) return *idx;
    size_t generation = 0;
    ctx->sys->rwlock.lock_read();
    for (size_t i = 0; i < ctx->sys->before.count(); i++) {
        SystemContext **dep = g->systems.get(ctx->sys->before.keys[i]);
        if (!dep) continue; // legal only for implicit / weak links (:226-229)
        size_t idx = toposort(g, tasks, markers, ctx->sys->before.keys[i], *dep);
        generation = std::max(generation, tasks[idx].generation + 1);
    }
    ctx->sys->rwlock.unlock_read();
    const size_t idx = tasks.size();
    TaskInfo info{};
    info.ctx = ctx;
    info.generation = generation;
    info.task.label = ctx->sys->label.data();
    info.task.label_len = ctx->sys->label.size();
    info.task.on_start = &task_start;
    tasks.push_back(info);
    markers.put(handle, idx);
    return idx;
}

static void graph_populate_buffers(SystemGroup *g) {
    // SystemGroup.zig:160-338
    if (!g->dirty) return;
    g->dirty = false;
    graph_clear_deinit_list(g);
    g->resources.clear();

    for (SystemContext *ctx : g->systems.values) {
        for (Resource *r : ctx->sys->exclusive_resources) g->resources.put(r, nullptr); // :171-172
        for (Resource *r : ctx->sys->shared_resources) g->resources.put(r, nullptr);
        ctx->sys->rwlock.lock_read();
        ctx->merge_deferred = true; // :178-181
        for (const Link &l : ctx->sys->after.values)
            if (l.ignore_deferred) ctx->merge_deferred = false;
        ctx->deferred_dep.clear(); // :184-191
        for (size_t i = 0; i < ctx->sys->before.count(); i++) {
            if (ctx->sys->before.values[i].ignore_deferred) continue;
            SystemContext **dep = g->systems.get(ctx->sys->before.keys[i]);
            if (!dep) continue;
            ctx->deferred_dep.push_back(*dep);
        }
        ctx->sys->rwlock.unlock_read();
        ctx->resources.assign(ctx->sys->exclusive_resources.size() + ctx->sys->shared_resources.size(), nullptr);
    }

    std::vector<TaskInfo> tasks;
    tasks.reserve(g->systems.count());
    ArrayMap<System *, size_t> markers;
    for (size_t i = 0; i < g->systems.count(); i++) toposort(g, tasks, markers, g->systems.keys[i], g->systems.values[i]);
    // :262 sorts unstably (order inside a generation unspecified); the oracle pins it with a stable sort.
    std::stable_sort(tasks.begin(), tasks.end(),
                     [](const TaskInfo &a, const TaskInfo &b) { return a.generation < b.generation; });
    g->tasks.swap(tasks);

    struct ResourceInfo {
        bool exclusive = false;
        std::vector<System *> referenced_by;
    };
    std::vector<ResourceInfo> infos(g->resources.count());
    ArrayMap<System *, size_t> running; // system -> index of its enqueue_task entry
    g->entries.clear();
    auto wait_on = [&](System *id) {
        size_t *entry_idx = running.get(id);
        if (!entry_idx) return;
        FimoTasksMeta_CommandBufferEntry e{};
        e.type = FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_INDIRECT;
        e.payload.wait_on_command_indirect = g->entries.size() - *entry_idx;
        running.swap_remove(id);
        g->entries.push_back(e);
    };
    for (TaskInfo &task : g->tasks) {
        System *sys = task.ctx->sys;
        sys->rwlock.lock_read();
        for (System *dep : sys->before.keys) wait_on(dep); // explicit predecessors, :279-288
        for (Resource *res : sys->exclusive_resources) {   // W after anything, :291-307
            ResourceInfo &info = infos[(size_t)g->resources.index_of(res)];
            for (System *id : info.referenced_by) wait_on(id);
            info.exclusive = true;
            info.referenced_by.clear();
            info.referenced_by.push_back(sys);
        }
        for (Resource *res : sys->shared_resources) { // R after W, :310-328
            ResourceInfo &info = infos[(size_t)g->resources.index_of(res)];
            if (info.exclusive) {
                for (System *id : info.referenced_by) wait_on(id);
                info.referenced_by.clear();
            }
            info.exclusive = false;
            info.referenced_by.push_back(sys);
        }
        sys->rwlock.unlock_read();
        running.put(sys, g->entries.size());
        FimoTasksMeta_CommandBufferEntry e{};
        e.type = FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK;
        e.payload.enqueue_task = &task.task;
        g->entries.push_back(e);
    }
}

static void graph_acquire_resources(SystemGroup *g) {
    // SystemGroup.zig:340-351: ALL group resources are locked exclusively for the generation.
    world_lock_resources(g->world, g->resources.keys.data(), g->resources.count(), nullptr, 0, g->resources.values.data());
    for (SystemContext *ctx : g->systems.values) {
        size_t i = 0;
        for (Resource *r : ctx->sys->exclusive_resources) ctx->resources[i++] = *g->resources.get(r);
        for (Resource *r : ctx->sys->shared_resources) ctx->resources[i++] = *g->resources.get(r);
    }
}

// ---- generation loop: schedule (SystemGroup.zig:430-452) + run (:454-487) + Job.go (Job.zig:216-264) ----
struct Job {
    SystemGroup *group;
    size_t generation;
    std::vector<FimoWorldsMeta_Fence *> fences; // Job.zig:221 dups the fence array
    FimoWorldsMeta_Fence *signal;
    std::atomic<int> pending{2};                 // two command buffers reference this block
    FimoTasksMeta_Task task{};                   // the group job's single task
    FimoTasksMeta_CommandBufferEntry job_entries[2]; // future.zig:91-188: [abort_on_error(true), enqueue_task]
    FimoTasksMeta_CommandBuffer job_buffer{};
    FimoTasksMeta_CommandBuffer systems_buffer{}; // SystemGroup.zig:464-472
};

static void job_release(Job *job) {
    if (job->pending.fetch_sub(1, std::memory_order_acq_rel) == 1) delete job;
}
static void job_buffer_deinit(FimoTasksMeta_CommandBuffer *cb) {
    job_release((Job *)((char *)cb - offsetof(Job, job_buffer)));
}
static void systems_buffer_deinit(FimoTasksMeta_CommandBuffer *cb) {
    job_release((Job *)((char *)cb - offsetof(Job, systems_buffer)));
}

static void group_run_end(FimoTasksMeta_CommandBuffer *cb) {
    // Second half of SystemGroup.run (:477-486), then Job cleanup (Job.zig:234-247).
    Job *job = (Job *)((char *)cb - offsetof(Job, systems_buffer));
    SystemGroup *g = job->group;
    for (size_t i = 0; i < g->resources.count(); i++) {
        g->resources.values[i] = nullptr;
        world_unlock_resource(g->world, g->resources.keys[i], true);
    }
    const size_t gen = g->generation.fetch_add(1, std::memory_order_acq_rel) + 1;
    g->single_allocator.reset();
    g->multi_allocator[gen % 4].reset();
    FimoWorldsMeta_Fence *signal = job->signal;
    FimoWorldsMeta_timeline_semaphore_signal(&g->schedule_semaphore, (uint64_t)gen, &futex_wake);
    if (signal) fence_signal(signal); // nothing of the group may be touched after this line
}

static void group_run_begin(FimoTasksMeta_Task *task) {
    // First half of SystemGroup.run (:454-473), executed as the job's task on a worker.
    Job *job = (Job *)((char *)task - offsetof(Job, task));
    SystemGroup *g = job->group;
    if (g->generation.load(std::memory_order_acquire) != job->generation) panic("generation mismatch");
    {
        std::lock_guard<std::mutex> lock(g->mutex);
        graph_populate_buffers(g);
        graph_acquire_resources(g);
    }
    job->systems_buffer.label = g->systems_label.data();
    job->systems_buffer.label_len = g->systems_label.size();
    job->systems_buffer.entries = g->entries.data();
    job->systems_buffer.entries_len = g->entries.size();
    job->systems_buffer.on_complete = &group_run_end;
    job->systems_buffer.on_abort = &group_run_end;
    job->systems_buffer.on_deinit = &systems_buffer_deinit;
    if (g->executor.vtable->enqueue_buffer(g->executor.data, &job->systems_buffer, nullptr) != FIMO_STATUS_OK)
        panic("enqueue of systems buffer failed");
}

static void job_start(Job *job) {
    SystemGroup *g = job->group;
    job->task.label = g->label.data();
    job->task.label_len = g->label.size();
    job->task.on_start = &group_run_begin;
    job->job_entries[0].type = FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR;
    job->job_entries[0].payload.abort_on_error = true;
    job->job_entries[1].type = FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK;
    job->job_entries[1].payload.enqueue_task = &job->task;
    job->job_buffer.label = g->label.data();
    job->job_buffer.label_len = g->label.size();
    job->job_buffer.entries = job->job_entries;
    job->job_buffer.entries_len = 2;
    job->job_buffer.on_deinit = &job_buffer_deinit;
    if (g->executor.vtable->enqueue_buffer(g->executor.data, &job->job_buffer, nullptr) != FIMO_STATUS_OK)
        panic("enqueue of group job failed");
}

// Continuation form of TimelineSemaphore.wait (Job.zig:101-138).
static void semaphore_then(FimoWorldsMeta_TimelineSemaphore *s, uint64_t value, std::function<void()> fn) {
    for (;;) {
        const uint64_t state = __atomic_load_n(&s->state, __ATOMIC_ACQUIRE);
        if (state >= value) {
            fn();
            return;
        }
        auto again = [s, value, fn]() { semaphore_then(s, value, fn); };
        if (futex_wait_async(&s->state, 8, state, (size_t)value, again)) return;
    }
}

static void job_wait_fences(Job *job, size_t i) {
    // Job.zig:224-232: wait all fences, then the semaphore (generations serialise), then run.
    if (i < job->fences.size()) {
        fence_then(job->fences[i], [job, i]() { job_wait_fences(job, i + 1); });
        return;
    }
    semaphore_then(&job->group->schedule_semaphore, (uint64_t)job->generation, [job]() { job_start(job); });
}

static void group_schedule(SystemGroup *g, FimoWorldsMeta_Fence *const *fences, size_t n, FimoWorldsMeta_Fence *signal) {
    Job *job = new Job();
    {
        std::lock_guard<std::mutex> lock(g->mutex);
        job->group = g;
        job->generation = g->next_generation;
        job->fences.assign(fences, fences + n);
        job->signal = signal;
        g->next_generation += 1;
    }
    job_wait_fences(job, 0);
}

static SystemGroup *group_init(const FimoWorldsMeta_SystemGroupDescriptor *d) {
    // SystemGroup.zig:359-378
    SystemGroup *g = new SystemGroup();
    g->label = d->label_len ? std::string(d->label, d->label_len) : std::string("<unlabelled>");
    g->world = (World *)d->world;
    g->executor = d->pool ? *d->pool : g->world->executor;
    g->executor.vtable->acquire(g->executor.data);
    g->world->system_group_count.fetch_add(1, std::memory_order_relaxed);
    char buf[64];
    std::snprintf(buf, sizeof buf, "%p systems", (void *)g); // :465
    g->systems_label = buf;
    return g;
}

static void group_deinit(SystemGroup *g) {
    // SystemGroup.zig:380-397
    {
        std::lock_guard<std::mutex> lock(g->mutex);
        if (!FimoWorldsMeta_timeline_semaphore_is_signaled(&g->schedule_semaphore, (uint64_t)g->next_generation))
            panic("system group still running");
        if (g->systems.count() != 0) panic("system group not empty");
        graph_clear_deinit_list(g);
    }
    g->world->system_group_count.fetch_sub(1, std::memory_order_relaxed);
    g->executor.vtable->release(g->executor.data);
    delete g;
}

static const char *group_add_systems(SystemGroup *g, System *const *handles, size_t n) {
    // SystemGroup.zig:399-418
    std::lock_guard<std::mutex> lock(g->mutex);
    g_universe.rwlock.lock_read();
    struct U {
        ~U() { g_universe.rwlock.unlock_read(); }
    } u;
    for (size_t i = 0; i < n; i++) {
        if (!g_universe.systems.contains(handles[i])) panic("invalid system");
        if (SystemContext **c = g->systems.get(handles[i]))
            if (!(*c)->weak) return "Duplicate";
    }
    for (size_t i = 0; i < n; i++) {
        const char *err = "InitError";
        if (!graph_add_system(g, handles[i], false, &err)) {
            for (size_t j = 0; j < i; j++) graph_remove_system(g, handles[j], nullptr, false);
            return err;
        }
    }
    return nullptr;
}

static void group_remove_system(SystemGroup *g, System *handle, FimoWorldsMeta_Fence *fence) {
    // SystemGroup.zig:420-428
    std::lock_guard<std::mutex> lock(g->mutex);
    SystemContext **c = g->systems.get(handle);
    if (!c) panic("invalid system");
    if ((*c)->weak) panic("invalid system");
    graph_remove_system(g, handle, fence, true);
}

} // namespace oracle

// ---------------------------------------------------------------------------------
// C ABI (FimoWorlds.zig:108-447)
// ---------------------------------------------------------------------------------
using namespace oracle;

static FimoStatus fail(const char *name) {
    set_last_error(name); // ctx.setResult(.initErr(.initError(err))); return .err;
    return FIMO_STATUS_ERROR;
}

extern "C" {

FimoStatus fimo_worlds_resource_register(const FimoWorldsMeta_ResourceDescriptor *resource,
                                         FimoWorldsMeta_ResourceHandle *handle) {
    *handle = (FimoWorldsMeta_ResourceHandle)universe_register_resource(resource);
    return FIMO_STATUS_OK;
}
void fimo_worlds_resource_unregister(FimoWorldsMeta_ResourceHandle handle) {
    universe_unregister_resource((Resource *)handle);
}
FimoStatus fimo_worlds_system_register(const FimoWorldsMeta_SystemDescriptor *system, FimoWorldsMeta_SystemHandle *handle) {
    const char *err = nullptr;
    System *s = universe_register_system(system, &err);
    if (!s) return fail(err);
    *handle = (FimoWorldsMeta_SystemHandle)s;
    return FIMO_STATUS_OK;
}
void fimo_worlds_system_unregister(FimoWorldsMeta_SystemHandle handle) { universe_unregister_system((System *)handle); }

FimoStatus fimo_worlds_system_group_create(const FimoWorldsMeta_SystemGroupDescriptor *descriptor,
                                           FimoWorldsMeta_SystemGroup **group) {
    *group = (FimoWorldsMeta_SystemGroup *)group_init(descriptor);
    return FIMO_STATUS_OK;
}
void fimo_worlds_system_group_destroy(FimoWorldsMeta_SystemGroup *group) { group_deinit((SystemGroup *)group); }
FimoWorldsMeta_World *fimo_worlds_system_group_get_world(FimoWorldsMeta_SystemGroup *group) {
    return (FimoWorldsMeta_World *)((SystemGroup *)group)->world;
}
const char *fimo_worlds_system_group_get_label(FimoWorldsMeta_SystemGroup *group, FimoUSize *len) {
    SystemGroup *g = (SystemGroup *)group;
    *len = g->label.size();
    return g->label.data();
}
FimoTasksMeta_Pool fimo_worlds_system_group_get_pool(FimoWorldsMeta_SystemGroup *group) {
    SystemGroup *g = (SystemGroup *)group;
    g->executor.vtable->acquire(g->executor.data); // returns a NEW reference (FimoWorlds.zig:207-210)
    return g->executor;
}
FimoStatus fimo_worlds_system_group_add_systems(FimoWorldsMeta_SystemGroup *group, const FimoWorldsMeta_SystemHandle *systems,
                                                FimoUSize systems_len) {
    const char *err = group_add_systems((SystemGroup *)group, (System *const *)systems, systems_len);
    return err ? fail(err) : FIMO_STATUS_OK;
}
void fimo_worlds_system_group_remove_system(FimoWorldsMeta_SystemGroup *group, FimoWorldsMeta_SystemHandle handle,
                                            FimoWorldsMeta_Fence *signal) {
    group_remove_system((SystemGroup *)group, (System *)handle, signal);
}
FimoStatus fimo_worlds_system_group_schedule(FimoWorldsMeta_SystemGroup *group, FimoWorldsMeta_Fence *const *wait_on,
                                             FimoUSize wait_on_len, FimoWorldsMeta_Fence *signal) {
    group_schedule((SystemGroup *)group, wait_on, wait_on_len, signal);
    return FIMO_STATUS_OK;
}

FimoWorldsMeta_SystemGroup *fimo_worlds_system_context_get_group(FimoWorldsMeta_SystemContext *context) {
    return (FimoWorldsMeta_SystemGroup *)((SystemContext *)context)->group;
}
FimoUSize fimo_worlds_system_context_get_generation(FimoWorldsMeta_SystemContext *context) {
    return ((SystemContext *)context)->group->generation.load(std::memory_order_acquire);
}

// SystemContext.zig:102-110: strategy -> allocator
static void *ctx_alloc(SystemContext *c, FimoWorldsMeta_SystemAllocatorStrategy strategy, size_t size, size_t alignment) {
    switch (strategy) {
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_TRANSIENT: return c->arena.alloc(size, alignment);
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SINGLE_GENERATION:
            return c->group->single_allocator.alloc(size, alignment);
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_MULTI_GENERATION:
            return c->group->multi_allocator[c->group->generation.load() % 4].alloc(size, alignment);
        case FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT: return c->persistent.alloc(size, alignment);
        default: panic("unknown strategy");
    }
}
void *fimo_worlds_system_context_allocator_alloc(FimoWorldsMeta_SystemContext *context,
                                                 FimoWorldsMeta_SystemAllocatorStrategy strategy, FimoUSize size,
                                                 FimoUSize alignment, FimoUSize) {
    return ctx_alloc((SystemContext *)context, strategy, size, alignment);
}
bool fimo_worlds_system_context_allocator_resize(FimoWorldsMeta_SystemContext *, FimoWorldsMeta_SystemAllocatorStrategy,
                                                 void *, FimoUSize size, FimoUSize, FimoUSize new_size, FimoUSize) {
    return new_size <= size; // shrinking in place always succeeds; growing never does
}
void *fimo_worlds_system_context_allocator_remap(FimoWorldsMeta_SystemContext *context,
                                                 FimoWorldsMeta_SystemAllocatorStrategy strategy, void *ptr, FimoUSize size,
                                                 FimoUSize alignment, FimoUSize new_size, FimoUSize) {
    if (new_size <= size) return ptr;
    void *p = ctx_alloc((SystemContext *)context, strategy, new_size, alignment);
    if (p && ptr) std::memcpy(p, ptr, size);
    if (p && strategy == FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT)
        ((SystemContext *)context)->persistent.release(ptr);
    return p;
}
void fimo_worlds_system_context_allocator_free(FimoWorldsMeta_SystemContext *context,
                                               FimoWorldsMeta_SystemAllocatorStrategy strategy, void *ptr, FimoUSize,
                                               FimoUSize, FimoUSize) {
    if (strategy == FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT)
        ((SystemContext *)context)->persistent.release(ptr);
    // arena strategies free in bulk
}

FimoStatus fimo_worlds_world_create(const FimoWorldsMeta_WorldDescriptor *descriptor, FimoWorldsMeta_World **world) {
    *world = (FimoWorldsMeta_World *)world_init(descriptor);
    return FIMO_STATUS_OK;
}
void fimo_worlds_world_destroy(FimoWorldsMeta_World *world) { world_deinit((World *)world); }
const char *fimo_worlds_world_get_label(FimoWorldsMeta_World *world, FimoUSize *len) {
    World *w = (World *)world;
    *len = w->label.size();
    return w->label.data();
}
FimoTasksMeta_Pool fimo_worlds_world_get_pool(FimoWorldsMeta_World *world) {
    World *w = (World *)world;
    w->executor.vtable->acquire(w->executor.data); // new reference (FimoWorlds.zig:338-341)
    return w->executor;
}
bool fimo_worlds_world_has_resource(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle) {
    World *w = (World *)world;
    w->rwlock.lock_read();
    bool r = w->resources.contains((Resource *)handle);
    w->rwlock.unlock_read();
    return r;
}
FimoStatus fimo_worlds_world_add_resource(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle,
                                          const void *value) {
    const char *err = world_add_resource((World *)world, (Resource *)handle, value);
    return err ? fail(err) : FIMO_STATUS_OK;
}
FimoStatus fimo_worlds_world_remove_resource(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle, void *value) {
    const char *err = world_remove_resource((World *)world, (Resource *)handle, value);
    return err ? fail(err) : FIMO_STATUS_OK;
}
void fimo_worlds_world_lock_resources(FimoWorldsMeta_World *world, const FimoWorldsMeta_ResourceHandle *exclusive_handles,
                                      FimoUSize exclusive_handles_len, const FimoWorldsMeta_ResourceHandle *shared_handles,
                                      FimoUSize shared_handles_len, void **out_resources) {
    world_lock_resources((World *)world, (Resource *const *)exclusive_handles, exclusive_handles_len,
                         (Resource *const *)shared_handles, shared_handles_len, out_resources);
}
void fimo_worlds_world_unlock_resource_exclusive(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle) {
    world_unlock_resource((World *)world, (Resource *)handle, true);
}
void fimo_worlds_world_unlock_resource_shared(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle) {
    world_unlock_resource((World *)world, (Resource *)handle, false);
}
void *fimo_worlds_world_allocator_alloc(FimoWorldsMeta_World *world, FimoUSize size, FimoUSize alignment, FimoUSize) {
    return ((World *)world)->allocator.alloc(size, alignment);
}
bool fimo_worlds_world_allocator_resize(FimoWorldsMeta_World *, void *, FimoUSize size, FimoUSize, FimoUSize new_size,
                                        FimoUSize) {
    return new_size <= size;
}
void *fimo_worlds_world_allocator_remap(FimoWorldsMeta_World *world, void *ptr, FimoUSize size, FimoUSize alignment,
                                        FimoUSize new_size, FimoUSize) {
    if (new_size <= size) return ptr;
    void *p = ((World *)world)->allocator.alloc(new_size, alignment);
    if (p && ptr) std::memcpy(p, ptr, size);
    if (p) ((World *)world)->allocator.release(ptr);
    return p;
}
void fimo_worlds_world_allocator_free(FimoWorldsMeta_World *world, void *ptr, FimoUSize, FimoUSize, FimoUSize) {
    ((World *)world)->allocator.release(ptr);
}

#define EXPORT_ROW(n) {#n, "fimo-worlds", (void (*)(void)) & fimo_worlds_##n}
static const FimoB200_Export g_exports[] = {
        // order of `all_symbols`, pkgs/fimo_worlds_meta/src/symbols.zig:22-58
        EXPORT_ROW(resource_register), EXPORT_ROW(resource_unregister), EXPORT_ROW(system_register),
        EXPORT_ROW(system_unregister), EXPORT_ROW(system_group_create), EXPORT_ROW(system_group_destroy),
        EXPORT_ROW(system_group_get_world), EXPORT_ROW(system_group_get_label), EXPORT_ROW(system_group_get_pool),
        EXPORT_ROW(system_group_add_systems), EXPORT_ROW(system_group_remove_system), EXPORT_ROW(system_group_schedule),
        EXPORT_ROW(system_context_get_group), EXPORT_ROW(system_context_get_generation),
        EXPORT_ROW(system_context_allocator_alloc), EXPORT_ROW(system_context_allocator_resize),
        EXPORT_ROW(system_context_allocator_remap), EXPORT_ROW(system_context_allocator_free), EXPORT_ROW(world_create),
        EXPORT_ROW(world_destroy), EXPORT_ROW(world_get_label), EXPORT_ROW(world_get_pool), EXPORT_ROW(world_has_resource),
        EXPORT_ROW(world_add_resource), EXPORT_ROW(world_remove_resource), EXPORT_ROW(world_lock_resources),
        EXPORT_ROW(world_unlock_resource_exclusive), EXPORT_ROW(world_unlock_resource_shared),
        EXPORT_ROW(world_allocator_alloc), EXPORT_ROW(world_allocator_resize), EXPORT_ROW(world_allocator_remap),
        EXPORT_ROW(world_allocator_free),
};
const FimoB200_Export *fimo_worlds_b200_exports(FimoUSize *count) {
    *count = sizeof(g_exports) / sizeof(g_exports[0]);
    return g_exports;
}

} // extern "C"
