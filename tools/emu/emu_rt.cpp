// emu_rt.cpp — scheduler of the SIMT emulator (see emu_cuda.h).  TEST INFRASTRUCTURE ONLY; x86-64 System V.
#include "emu_cuda.h"
#include <sys/mman.h>
#include <mutex>
#include <deque>

extern "C" void emu_switch(void** save_sp, void* load_sp);
asm(R"(
.text
.globl emu_switch
.type emu_switch,@function
emu_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size emu_switch,.-emu_switch
)");

namespace emu {

Lane* cur = nullptr;

enum State { RUNNABLE = 0, WAIT_COLL, WAIT_BAR, SLEEPING, DONE };
constexpr size_t STACK_BYTES = 96 * 1024;

struct Cta;
struct Warp { Lane lanes[32]; int n = 0; Cta* cta = nullptr; bool at_barrier = false; bool done = false; };
struct Cta {
    std::vector<Warp*> warps; std::vector<void*> shared; void* dyn = nullptr; Idx bidx{}; int live = 0;
};

static std::mutex g_mu;                       // one launch at a time (host helper threads call into the engine too)
static void* g_sched_sp = nullptr;
static const std::function<void()>* g_body = nullptr;
static uint64_t g_progress = 0;
static uint8_t* g_stacks = nullptr; static size_t g_nstacks = 0; static std::vector<int> g_free_stacks;
static int g_lane_order = -1;
static int g_strict_sites = 0;

static int env_int(const char* name, int dflt) { const char* e = getenv(name); return e && *e ? atoi(e) : dflt; }
int sm_count() { static int v = env_int("B2_EMU_SMS", 2); return v; }
int device_count() { return env_int("B2_EMU_DEVICES", 1); }
int& current_device() { static thread_local int d = 0; return d; }
double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
void* dev_alloc(size_t n) {                   // "device" memory starts out as garbage, like the real thing
    void* p = nullptr;
    if (posix_memalign(&p, 256, n ? n : 1)) return nullptr;
    memset(p, 0xDB, n);
    return p;
}
void note_progress() { ++g_progress; }

static void yield_to_scheduler() { Lane* me = cur; emu_switch(&me->sp, g_sched_sp); }

uint32_t collective(int op, uint32_t a, uint32_t b) {
    Lane* me = cur;
    me->op = op; me->a = a; me->b = b; me->site = __builtin_return_address(0); me->state = WAIT_COLL;
    yield_to_scheduler();
    return me->res;
}
void syncthreads() { cur->state = WAIT_BAR; yield_to_scheduler(); }
void nanosleep_yield() { cur->state = SLEEPING; yield_to_scheduler(); }

void* shared_alloc(size_t bytes) {
    Lane* me = cur; Cta* c = me->warp->cta;
    const unsigned k = me->shared_cursor++;
    if (k >= c->shared.size()) {
        void* p = nullptr;
        if (posix_memalign(&p, 128, bytes ? bytes : 1)) abort();
        memset(p, 0xCD, bytes);               // shared memory is not zeroed on the device either
        c->shared.push_back(p);
    }
    return c->shared[k];
}
void* dyn_shared() { return cur->warp->cta->dyn; }

static void fiber_main() {
    Lane* me = cur;
    (*g_body)();
    me->state = DONE;
    ++g_progress;
    yield_to_scheduler();
    abort();                                  // a finished fiber is never resumed
}

static void init_stacks(size_t want) {
    if (g_nstacks >= want) return;
    // (the pool only grows between launches: nothing is live here)
    if (g_stacks) munmap(g_stacks, g_nstacks * STACK_BYTES);
    g_stacks = (uint8_t*)mmap(nullptr, want * STACK_BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (g_stacks == MAP_FAILED) { perror("emu: mmap"); abort(); }
    g_nstacks = want;
    g_free_stacks.clear();
    for (int i = (int)want - 1; i >= 0; --i) g_free_stacks.push_back(i);
}

static void start_lane(Lane& l) {
    l.stack_slot = g_free_stacks.back(); g_free_stacks.pop_back();
    uint8_t* top = g_stacks + (size_t)(l.stack_slot + 1) * STACK_BYTES;
    uint64_t* sp = (uint64_t*)(top - 64);     // 6 callee-saved registers, the entry point, one slot of padding
    memset(sp, 0, 64);
    sp[6] = (uint64_t)(uintptr_t)&fiber_main;
    l.sp = sp;
    l.state = RUNNABLE;
}

static void resolve(Warp* w) {
    // every live lane waits at a collective: they must be the same kind (different call sites are tolerated for
    // __syncwarp only, as on the hardware)
    int op = 0; const void* site = nullptr; uint32_t members = 0;
    for (int i = 0; i < w->n; ++i) {
        Lane& l = w->lanes[i];
        if (l.state != WAIT_COLL) continue;
        if (!op) { op = l.op; site = l.site; }
        else if (l.op != op || (g_strict_sites && op != OP_SYNCWARP && l.site != site)) {   // (the compiler may duplicate one source-level call)
            fprintf(stderr, "emu: warp diverged at a collective: lane %d op %d site %p vs op %d site %p (block %u, warp of thread %u)\n",
                    i, l.op, l.site, op, site, l.bidx.x, w->lanes[0].tidx.x);
            abort();
        }
        members |= 1u << i;
    }
    auto val = [&](int i) { return w->lanes[i].a; };
    uint32_t agg = 0;
    switch (op) {
        case OP_BALLOT: case OP_ANY: for (int i = 0; i < w->n; ++i) if ((members >> i & 1u) && val(i)) agg |= 1u << i; break;
        case OP_ALL: agg = 1; for (int i = 0; i < w->n; ++i) if ((members >> i & 1u) && !val(i)) agg = 0; break;
        case OP_RED_OR: for (int i = 0; i < w->n; ++i) if (members >> i & 1u) agg |= val(i); break;
        case OP_RED_AND: agg = 0xFFFFFFFFu; for (int i = 0; i < w->n; ++i) if (members >> i & 1u) agg &= val(i); break;
        case OP_RED_ADD: for (int i = 0; i < w->n; ++i) if (members >> i & 1u) agg += val(i); break;
        case OP_RED_MIN: agg = 0xFFFFFFFFu; for (int i = 0; i < w->n; ++i) if ((members >> i & 1u) && val(i) < agg) agg = val(i); break;
        case OP_RED_MAX: for (int i = 0; i < w->n; ++i) if ((members >> i & 1u) && val(i) > agg) agg = val(i); break;
        default: break;
    }
    for (int i = 0; i < w->n; ++i) {
        Lane& l = w->lanes[i];
        if (!(members >> i & 1u)) continue;
        int src = i;
        switch (op) {
            case OP_SHFL: src = (int)(l.b & 31u); break;
            case OP_SHFL_UP: src = i - (int)l.b; if (src < 0) src = i; break;
            case OP_SHFL_DOWN: src = i + (int)l.b; if (src > 31) src = i; break;
            case OP_SHFL_XOR: src = i ^ (int)(l.b & 31u); break;
            default: break;
        }
        switch (op) {
            case OP_SHFL: case OP_SHFL_UP: case OP_SHFL_DOWN: case OP_SHFL_XOR:
                l.res = (src < w->n && (members >> src & 1u)) ? w->lanes[src].a : l.a; break;   // inactive source: undefined on the device
            case OP_ANY: l.res = agg != 0; break;
            case OP_SYNCWARP: l.res = 0; break;
            default: l.res = agg; break;
        }
    }
    for (int i = 0; i < w->n; ++i) if (members >> i & 1u) w->lanes[i].state = RUNNABLE;
    ++g_progress;
}

// runs the warp until every live lane is blocked on something another warp (or a later time slice) has to provide
static void run_warp(Warp* w) {
    if (w->done || w->at_barrier) return;
    for (;;) {
        bool ran = false;
        for (int k = 0; k < w->n; ++k) {
            int i = k;
            if (g_lane_order == 1) i = w->n - 1 - k;
            else if (g_lane_order == 2) i = (int)((k * 13u + (unsigned)g_progress * 7u) % (unsigned)w->n);   // 13 is coprime with 32
            Lane& l = w->lanes[i];
            if (l.state != RUNNABLE) continue;
            cur = &l;
            emu_switch(&g_sched_sp, l.sp);
            cur = nullptr;
            ran = true;
            if (l.state == DONE) { g_free_stacks.push_back(l.stack_slot); l.stack_slot = -1; }
        }
        int n_coll = 0, n_bar = 0, n_sleep = 0, n_live = 0, n_run = 0;
        for (int i = 0; i < w->n; ++i) {
            switch (w->lanes[i].state) {
                case WAIT_COLL: ++n_coll; ++n_live; break;
                case WAIT_BAR: ++n_bar; ++n_live; break;
                case SLEEPING: ++n_sleep; ++n_live; break;
                case RUNNABLE: ++n_run; ++n_live; break;
                default: break;
            }
        }
        if (n_live == 0) { w->done = true; return; }
        if (n_run) continue;                                   // (random order may have skipped one)
        if (n_coll == n_live) { resolve(w); continue; }
        if (n_sleep) {                                         // give the rest of the machine a turn, then wake them
            for (int i = 0; i < w->n; ++i) if (w->lanes[i].state == SLEEPING) w->lanes[i].state = RUNNABLE;
            return;
        }
        if (n_bar == n_live) { w->at_barrier = true; return; }
        if (!ran || (n_coll && n_bar)) {
            fprintf(stderr, "emu: warp stuck: %d lanes at a warp collective, %d at __syncthreads (block %u)\n", n_coll, n_bar, w->lanes[0].bidx.x);
            abort();
        }
    }
}

void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()>& body) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (g_lane_order < 0) {
        const char* e = getenv("B2_EMU_LANE_ORDER");
        g_lane_order = !e ? 0 : !strcmp(e, "reverse") ? 1 : !strcmp(e, "random") ? 2 : 0;
        g_strict_sites = env_int("B2_EMU_STRICT_SITES", 0);
    }
    const unsigned nthreads = block.x * block.y * block.z;
    const unsigned nctas = grid.x * grid.y * grid.z;
    if (!nthreads || !nctas) return;
    const unsigned resident_limit = (unsigned)env_int("B2_EMU_RESIDENT_THREADS", 8192);
    const unsigned max_resident_ctas = std::max(1u, resident_limit / nthreads);
    init_stacks((size_t)std::min(nctas, max_resident_ctas) * nthreads);
    g_body = &body;
    const unsigned nwarps = (nthreads + 31) / 32;
    std::deque<Cta*> resident;
    unsigned next_cta = 0;
    uint64_t idle_passes = 0, last_progress = g_progress;
    const uint64_t idle_limit = (uint64_t)env_int("B2_EMU_IDLE_PASSES", 2000000);
    while (next_cta < nctas || !resident.empty()) {
        while (next_cta < nctas && resident.size() < max_resident_ctas) {
            Cta* c = new Cta();
            c->bidx = Idx{ next_cta % grid.x, (next_cta / grid.x) % grid.y, next_cta / (grid.x * grid.y) };
            if (smem) { if (posix_memalign(&c->dyn, 128, smem)) abort(); memset(c->dyn, 0xCD, smem); }
            for (unsigned wi = 0; wi < nwarps; ++wi) {
                Warp* w = new Warp(); w->cta = c; w->n = (int)std::min(32u, nthreads - wi * 32u);
                for (int i = 0; i < w->n; ++i) {
                    Lane& l = w->lanes[i];
                    const unsigned t = wi * 32u + (unsigned)i;
                    l.lane = i; l.warp = w; l.shared_cursor = 0;
                    l.tidx = Idx{ t % block.x, (t / block.x) % block.y, t / (block.x * block.y) };
                    l.bidx = c->bidx; l.bdim = Idx{ block.x, block.y, block.z }; l.gdim = Idx{ grid.x, grid.y, grid.z };
                    start_lane(l);
                }
                c->warps.push_back(w);
            }
            c->live = (int)nwarps;
            resident.push_back(c);
            ++next_cta;
        }
        for (size_t ci = 0; ci < resident.size();) {
            Cta* c = resident[ci];
            int at_bar = 0, live = 0;
            for (Warp* w : c->warps) { run_warp(w); if (!w->done) { ++live; if (w->at_barrier) ++at_bar; } }
            if (live && at_bar == live) {                      // __syncthreads: everybody (still alive) has arrived
                for (Warp* w : c->warps) if (!w->done) {
                    w->at_barrier = false;
                    for (int i = 0; i < w->n; ++i) if (w->lanes[i].state == WAIT_BAR) w->lanes[i].state = RUNNABLE;
                }
                ++g_progress;
            }
            if (!live) {
                for (Warp* w : c->warps) delete w;
                for (void* p : c->shared) free(p);
                free(c->dyn);
                delete c;
                resident.erase(resident.begin() + (long)ci);
                ++g_progress;
            } else ++ci;
        }
        if (g_progress == last_progress) {
            if (++idle_passes > idle_limit) {
                fprintf(stderr, "emu: no progress in %llu scheduler passes — the kernel waits for something no resident warp will do (deadlock)\n",
                        (unsigned long long)idle_passes);
                for (Cta* c : resident) for (Warp* w : c->warps) if (!w->done) {
                    int st[5] = {0, 0, 0, 0, 0}; for (int i = 0; i < w->n; ++i) ++st[w->lanes[i].state];
                    fprintf(stderr, "  block %u warp %u: runnable %d, collective %d, barrier %d, sleeping %d, done %d\n", c->bidx.x, w->lanes[0].tidx.x / 32u, st[0], st[1], st[2], st[3], st[4]);
                }
                abort();
            }
        } else { idle_passes = 0; last_progress = g_progress; }
    }
    g_body = nullptr;
}

}  // namespace emu
