// lz4f_compat.cu — the LZ4F_* streaming entry points of include/lz4frame.h on top of the batch engine.
//
// These are the twelve symbols src/Codec/Compression/LZ4/Conduit.hsc binds through inline-c (list with
// line numbers in include/lz4frame.h).  The reference implements them as a per-16 KB CPU state machine
// (lz4/lib/lz4frame.c:600-1018 compress side, :1109-1840 decompress side); here they are a thin host layer
// that collects whole blocks in pinned staging buffers and hands them to the GPU a batch at a time
// (b2lz4_compress_blocks_host / b2lz4_decompress_blocks_host).  Nothing in this file compresses or
// decompresses: without a CUDA device the calls return ERROR_cudaFailure.
#include "b2_cuda.h"
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include <algorithm>
#include <chrono>
#include <future>
#include <string>
#include <mutex>
#include <vector>
#include <stdio.h>
#include "format.h"
#include "../../include/b200lz4.h"
#include "../../include/lz4frame.h"

using namespace b2;

namespace {

inline size_t ec(uint32_t e) { return (size_t)0 - (size_t)e; }

// uncompressed bytes handed to the GPU per call.  B2LZ4_BATCH_MB sets both directions (4..2048 MiB).  Defaults:
//   decompression 32 MiB — the GPU is done with a batch before the host has copied the next one in;
//   compression   12 MiB — measured through the unmodified conduit's call sequence (profiles/r2y_tier1_batch_sweep.txt):
//     `compress` runs as fast with 12 MiB batches as with 32 (4.9 vs 4.5-5.1 GB/s, the layer is bound by the caller's
//     copies), and LZ4F_compressBound() — two batches — stays at 24 MiB, so that a caller who allocates the bound for EVERY
//     input piece (compressYieldImmediately, Conduit.hsc:385-400) is served from the allocator's heap instead of a fresh
//     mapping per call: 4.9 GB/s instead of 2.3.
const char* batch_env() { const char* e = getenv("B2LZ4_BATCH_MB"); return (e && *e) ? e : nullptr; }   // empty = unset
size_t batch_mb(long dflt) {
    const char* e = batch_env();
    long mb = e ? atol(e) : dflt;
    if (mb < 4) mb = 4;
    if (mb > 2048) mb = 2048;
    return ((size_t)mb << 20) & ~(((size_t)4 << 20) - 1);
}
size_t batch_bytes() { static size_t v = 0; if (!v) v = batch_mb(32); return v; }
size_t cbatch_bytes() { static size_t v = 0; if (!v) v = batch_mb(12); return v; }
int device_index() { const char* e = getenv("B2LZ4_DEVICE"); return e ? atoi(e) : 0; }
// B2LZ4_DEVICES=0,1,2,3: the batches of this stream shard across these GPUs by block index (b2lz4_set_devices)
size_t apply_devices(b2lz4_ctx* gpu) {
    const char* e = getenv("B2LZ4_DEVICES");
    if (!e || !*e) return 0;
    int devs[16]; int n = 0;
    for (const char* p = e; *p && n < 16;) {
        char* q = nullptr; long v = strtol(p, &q, 10);
        if (q == p) break;
        devs[n++] = (int)v; p = q;
        while (*p == ',' || *p == ' ') ++p;
    }
    return n ? b2lz4_set_devices(gpu, devs, n) : 0;
}

// B2LZ4_TRACE=1: every context prints, when it is freed, where its wall-clock time went (host copies in / GPU batches /
// host copies out) — the Tier-1 layer is host-copy bound, this is how that was measured
bool trace_on() { static int v = -1; if (v < 0) { const char* e = getenv("B2LZ4_TRACE"); v = (e && *e && *e != '0') ? 1 : 0; } return v == 1; }
struct Trace {
    double t_in = 0, t_gpu = 0, t_out = 0, t_all = 0; uint64_t calls = 0, batches = 0;
    static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
    void report(const char* what) const {
        if (trace_on()) fprintf(stderr, "[b2lz4 %s] calls %llu, batches %llu: inside the library %.3f s (copy-in %.3f, gpu %.3f, copy-out %.3f)\n", what,
                                (unsigned long long)calls, (unsigned long long)batches, t_all, t_in, t_gpu, t_out);
    }
};
struct Span { double& acc; double t0; explicit Span(double& a) : acc(a), t0(trace_on() ? Trace::now() : 0) {} ~Span() { if (trace_on()) acc += Trace::now() - t0; } };

// Pinning host memory costs ~0.5 ms per MB (and unpinning as much), which is more than a short stream's whole GPU time and
// erratic under virtualisation: staging buffers of finished contexts are kept in a small process-wide pool and handed to
// the next context instead of going back to the driver (a program that runs one conduit after the other pins once).
struct PinnedPool {
    static constexpr int SLOTS = 32;
    static constexpr size_t MAX_TOTAL = (size_t)768 << 20;
    std::mutex mu; uint8_t* p[SLOTS] = {}; size_t cap[SLOTS] = {}; size_t total = 0;
    uint8_t* get(size_t n, size_t* got) {                      // smallest pooled buffer that fits without wasting more than 4x
        std::lock_guard<std::mutex> g(mu);
        int best = -1;
        for (int i = 0; i < SLOTS; ++i) if (p[i] && cap[i] >= n && cap[i] <= 4 * n + (1u << 20) && (best < 0 || cap[i] < cap[best])) best = i;
        if (best < 0) return nullptr;
        uint8_t* q = p[best]; *got = cap[best]; total -= cap[best]; p[best] = nullptr; cap[best] = 0;
        return q;
    }
    // keeps the buffer if there is room; large buffers are the expensive ones to pin again, so when the pool is full the
    // smallest ones give way to a larger newcomer (they are handed back to the driver)
    bool put(uint8_t* q, size_t n) {
        uint8_t* drop[SLOTS]; int nd = 0; bool kept = false;
        {
            std::lock_guard<std::mutex> g(mu);
            if (n <= MAX_TOTAL) for (;;) {
                int slot = -1, small = -1;
                for (int i = 0; i < SLOTS; ++i) {
                    if (!p[i]) { if (slot < 0) slot = i; }
                    else if (small < 0 || cap[i] < cap[small]) small = i;
                }
                if (slot >= 0 && total + n <= MAX_TOTAL) { p[slot] = q; cap[slot] = n; total += n; kept = true; break; }
                if (small < 0 || cap[small] >= n) break;          // nothing smaller to evict
                drop[nd++] = p[small]; total -= cap[small]; p[small] = nullptr; cap[small] = 0;
            }
        }
        for (int i = 0; i < nd; ++i) b2lz4_pinned_free(drop[i]);
        return kept;
    }
};
PinnedPool& pinned_pool() { static PinnedPool* pool = new PinnedPool(); return *pool; }   // never destroyed: outlives CUDA teardown order

struct Pinned {                              // growable pinned host buffer
    uint8_t* p = nullptr; size_t cap = 0;
    // buffers of known size are allocated exactly, growing ones in few, large steps
    bool ensure(size_t n, size_t keep = 0, bool exact = false) {
        if (n <= cap) return true;
        size_t want = exact ? n : 2 * n + 4096;
        uint8_t* q = pinned_pool().get(n, &want);
        if (!q) q = (uint8_t*)b2lz4_pinned_alloc(want);
        if (!q) return false;
        if (keep) memcpy(q, p, keep);
        release();
        p = q; cap = want;
        return true;
    }
    void release() { if (p && !pinned_pool().put(p, cap)) b2lz4_pinned_free(p); p = nullptr; cap = 0; }
};

// Engine contexts (streams, events, device workspaces, the lanes on other devices) are pooled the same way: creating one
// costs a dozen cudaMallocs, destroying one as many cudaFrees (device-wide synchronisations; measured between 2 ms and
// 0.6 s on the pool's virtualised boxes), and a program that runs one conduit after the other would pay both per stream.
// A context goes back to the pool only if it never saw a CUDA error; it is handed out again for the same device and the
// same B2LZ4_DEVICES setting, with its preferences and dictionary reset.
struct EnginePool {
    static constexpr int SLOTS = 4;
    std::mutex mu; b2lz4_ctx* c[SLOTS] = {}; int dev[SLOTS] = {}; std::string devs[SLOTS];
    b2lz4_ctx* get(int device, const std::string& dv) {
        std::lock_guard<std::mutex> g(mu);
        for (int i = 0; i < SLOTS; ++i) if (c[i] && dev[i] == device && devs[i] == dv) { b2lz4_ctx* x = c[i]; c[i] = nullptr; return x; }
        return nullptr;
    }
    bool put(b2lz4_ctx* x, int device, const std::string& dv) {
        std::lock_guard<std::mutex> g(mu);
        for (int i = 0; i < SLOTS; ++i) if (!c[i]) { c[i] = x; dev[i] = device; devs[i] = dv; return true; }
        return false;
    }
};
EnginePool& engine_pool() { static EnginePool* pool = new EnginePool(); return *pool; }   // never destroyed (as the pinned pool)
std::string devices_env() { const char* e = getenv("B2LZ4_DEVICES"); return e ? e : ""; }
bool engine_pool_on() { static int v = -1; if (v < 0) { const char* e = getenv("B2LZ4_CTX_POOL"); v = (e && *e == '0') ? 0 : 1; } return v == 1; }

struct Engine {                               // an engine context and where it came from
    b2lz4_ctx* gpu = nullptr; int device = 0; std::string devs;
    size_t acquire(const b2lz4_prefs* bp) {
        device = device_index(); devs = devices_env();
        if (engine_pool_on()) {
            b2lz4_ctx* x = engine_pool().get(device, devs);
            if (x) {
                size_t r = b2lz4_set_prefs(x, bp);
                if (!b2lz4_is_error(r)) r = b2lz4_set_dict(x, nullptr, 0, 0);
                if (!b2lz4_is_error(r)) { gpu = x; return 0; }
                b2lz4_destroy(x);
            }
        }
        size_t r = b2lz4_create(&gpu, device, bp);
        if (b2lz4_is_error(r)) { gpu = nullptr; return r; }
        r = apply_devices(gpu);
        if (b2lz4_is_error(r)) { b2lz4_destroy(gpu); gpu = nullptr; }
        return r;
    }
    void release() {
        if (!gpu) return;
        const char* e = b2lz4_last_cuda_error(gpu);
        if (!(engine_pool_on() && (!e || !*e) && engine_pool().put(gpu, device, devs))) b2lz4_destroy(gpu);
        gpu = nullptr;
    }
};

// LZ4F_default -> 256 KB blocks here, each parsed as ONE independent unit: such a frame needs no side information to be
// decoded block-parallel (by this library's decoder as by any other), and its ratio on log-like data is 1 % better than
// the reference's default (linked 64 KB blocks).  4 MB / 1 MB blocks stay available through blockSizeID.
constexpr unsigned DEFAULT_BLOCK_ID = 5u;
constexpr size_t HINT_MAX = (size_t)512 << 10;
uint32_t block_bytes_of(unsigned id) { return block_size_of(id ? id : DEFAULT_BLOCK_ID); }

// worst case bytes of the block region for n input bytes: every block stored + header word + checksum
size_t blocks_bound(size_t n, uint32_t bs, bool bc) { return n + ((n + bs - 1) / bs) * (4u + (bc ? 4u : 0u)); }

}  // namespace

// =========================================================================================== compression
struct LZ4F_cctx_s {
    b2lz4_ctx* gpu = nullptr; Engine eng;
    LZ4F_preferences_t prefs{};
    b2lz4_prefs bp{};
    uint32_t block_size = 0;
    size_t batch = 0;
    Pinned in;  size_t in_len = 0;
    Pinned in2, out;                                         // the batch in flight reads in2 and writes out
    std::future<size_t> job; bool flight = false;            // at most one batch is on the GPU while the caller fills the next
    bool begun = false, first = true;
    uint64_t total_in = 0;
    Trace tr;
};

static void drop_flight(LZ4F_cctx* c) { if (c->flight) { c->job.wait(); (void)c->job.get(); c->flight = false; } }

extern "C" {

unsigned LZ4F_isError(LZ4F_errorCode_t code) { return b2lz4_is_error(code); }
const char* LZ4F_getErrorName(LZ4F_errorCode_t code) { return b2lz4_error_name(code); }
unsigned LZ4F_getVersion(void) { return LZ4F_VERSION; }

LZ4F_errorCode_t LZ4F_createCompressionContext(LZ4F_cctx** cctxPtr, unsigned version) {
    if (!cctxPtr) return ec(E_GENERIC);
    (void)version;
    LZ4F_cctx* c = new (std::nothrow) LZ4F_cctx_s();
    if (!c) return ec(E_allocation_failed);
    *cctxPtr = c;
    return 0;
}

LZ4F_errorCode_t LZ4F_freeCompressionContext(LZ4F_cctx* c) {
    if (!c) return 0;                                                    // lz4frame.c:554
    drop_flight(c);
    c->tr.report("cctx");
    const double t0 = trace_on() ? Trace::now() : 0;
    c->eng.release(); c->gpu = nullptr;
    const double t1 = trace_on() ? Trace::now() : 0;
    c->in.release(); c->out.release(); c->in2.release();
    if (trace_on()) fprintf(stderr, "[b2lz4 cctx] free: engine context %.3f s, staging %.3f s\n", t1 - t0, Trace::now() - t1);
    delete c;
    return 0;
}

size_t LZ4F_compressBound(size_t srcSize, const LZ4F_preferences_t* p) {
    // What one LZ4F_compressUpdate (or LZ4F_flush / LZ4F_compressEnd) can emit at most: everything buffered so
    // far (< one batch) plus srcSize, every block stored, plus EndMark and content checksum (lz4frame.c:324-349
    // has the same shape with "one block" in place of "one batch").
    LZ4F_preferences_t z; memset(&z, 0, sizeof z);
    if (!p) p = &z;
    const unsigned id = (unsigned)p->frameInfo.blockSizeID;
    if (id != 0 && (id < 4 || id > 7)) return ec(E_maxBlockSize_invalid);
    const uint32_t bs = block_bytes_of(id);
    const size_t buffered = p->autoFlush ? 0 : 2 * cbatch_bytes() - 1;            // one batch on the GPU, one being staged
    return blocks_bound(buffered + srcSize, bs, p->frameInfo.blockChecksumFlag != 0) + 8;
}

}  // extern "C"

struct LZ4F_CDict_s { std::vector<uint8_t> bytes; };

// LZ4F_blockLinked is the ZERO value of blockMode, i.e. what every caller that never thought about it passes (the reference
// conduit included, Conduit.hsc:248-263).  Frames with independent blocks are what this codec decodes in parallel, and any
// decoder accepts them, so plain LZ4F_compressBegin writes independent blocks unless B2LZ4_BLOCK_MODE=honour asks for the
// field to be taken literally; LZ4F_compressBegin_usingCDict always takes it literally (a dictionary is only used by linked
// blocks here, and a caller that brings one wants it used).
static bool honour_block_mode() { static int v = -1; if (v < 0) { const char* e = getenv("B2LZ4_BLOCK_MODE"); v = (e && (*e == 'h' || *e == 'l')) ? 1 : 0; } return v == 1; }

static size_t begin_impl(LZ4F_cctx* c, void* dst, size_t cap, const LZ4F_preferences_t* p, const LZ4F_CDict_s* cdict, bool literal_mode);

extern "C" {

size_t LZ4F_compressBegin(LZ4F_cctx* c, void* dst, size_t cap, const LZ4F_preferences_t* p) {
    return begin_impl(c, dst, cap, p, nullptr, honour_block_mode());
}

LZ4F_CDict* LZ4F_createCDict(const void* dictBuffer, size_t dictSize) {            // lz4frame.c:478-512: keeps the last 64 KB
    if (!dictBuffer && dictSize) return nullptr;
    LZ4F_CDict_s* d = new (std::nothrow) LZ4F_CDict_s();
    if (!d) return nullptr;
    const size_t keep = std::min<size_t>(dictSize, 65536);
    d->bytes.assign((const uint8_t*)dictBuffer + (dictSize - keep), (const uint8_t*)dictBuffer + dictSize);
    return d;
}
void LZ4F_freeCDict(LZ4F_CDict* d) { delete d; }                                   // lz4frame.c:514-521

size_t LZ4F_compressBegin_usingCDict(LZ4F_cctx* c, void* dst, size_t cap, const LZ4F_CDict* cdict, const LZ4F_preferences_t* p) {
    return begin_impl(c, dst, cap, p, cdict, true);
}

}  // extern "C"

static size_t begin_impl(LZ4F_cctx* c, void* dst, size_t cap, const LZ4F_preferences_t* p, const LZ4F_CDict_s* cdict, bool literal_mode) {
    if (!c || !dst) return ec(E_GENERIC);
    if (cap < (size_t)LZ4F_HEADER_SIZE_MAX) return ec(E_dstMaxSize_tooSmall);       // lz4frame.c:613
    drop_flight(c);                                                              // a frame abandoned without compressEnd
    LZ4F_preferences_t z; memset(&z, 0, sizeof z);
    c->prefs = p ? *p : z;
    const unsigned id = (unsigned)c->prefs.frameInfo.blockSizeID;
    if (id != 0 && (id < 4 || id > 7)) return ec(E_maxBlockSize_invalid);
    b2lz4_default_prefs(&c->bp);
    c->bp.block_size_id = id ? id : DEFAULT_BLOCK_ID;
    c->bp.sub_block_log2 = c->bp.block_size_id == 5u ? 18u : 17u;     // 256 KB blocks are not cut into sub-blocks
    c->bp.block_checksum = c->prefs.frameInfo.blockChecksumFlag != 0;
    c->bp.content_checksum = c->prefs.frameInfo.contentChecksumFlag != 0;
    c->bp.content_size = c->prefs.frameInfo.contentSize != 0;
    c->bp.write_index = 0;
    c->bp.block_linked = (literal_mode && c->prefs.frameInfo.blockMode == LZ4F_blockLinked) ? 1u : 0u;
    c->block_size = block_bytes_of(id);
    c->batch = cbatch_bytes();
    if (!c->gpu) {
        size_t r = c->eng.acquire(&c->bp);
        if (b2lz4_is_error(r)) return r;
        c->gpu = c->eng.gpu;
    } else {
        size_t r = b2lz4_set_prefs(c->gpu, &c->bp);
        if (b2lz4_is_error(r)) return r;
    }
    c->in_len = 0; c->total_in = 0; c->first = true; c->begun = true;
    // the dictionary (used by linked blocks only) and the dictID of the preferences (written as given, lz4frame.c:692-695)
    const bool use_dict = cdict && c->bp.block_linked && !cdict->bytes.empty();
    size_t r = b2lz4_set_dict(c->gpu, use_dict ? cdict->bytes.data() : nullptr, use_dict ? cdict->bytes.size() : 0, c->prefs.frameInfo.dictID);
    if (b2lz4_is_error(r)) return r;
    return b2lz4_frame_header_write2(dst, cap, &c->bp, c->prefs.frameInfo.contentSize, c->prefs.frameInfo.dictID);
}

extern "C" {

// compresses the staged bytes [0, n) into dst; returns bytes written
// the batch in flight has to finish; its blocks go to dst.  Returns bytes written.
static size_t collect_flight(LZ4F_cctx* c, uint8_t* dst, size_t cap) {
    if (!c->flight) return 0;
    size_t r;
    { Span sp(c->tr.t_gpu); r = c->job.get(); }
    c->flight = false;
    if (b2lz4_is_error(r)) return r;
    if (r > cap) return ec(E_dstMaxSize_tooSmall);
    { Span sp(c->tr.t_out); memcpy(dst, c->out.p, r); }
    return r;
}

// hands the staged batch [0, n) to the GPU on a helper thread and makes the other staging buffer the one being filled:
// the caller keeps copying input while the GPU compresses (this layer's time is host copies for the most part, B2LZ4_TRACE)
static size_t launch_staged(LZ4F_cctx* c, size_t n) {
    const size_t ocap = blocks_bound(n, c->block_size, c->bp.block_checksum != 0) + 16;
    if (!c->out.ensure(ocap, 0, true) || !c->in2.ensure(c->batch + 16, 0, true)) return ec(E_allocation_failed);   // what the next batch needs
    std::swap(c->in, c->in2);
    ++c->tr.batches;
    b2lz4_ctx* gpu = c->gpu; const uint8_t* src = c->in2.p; uint8_t* dst = c->out.p;
    const unsigned flags = c->first ? B2LZ4_STREAM_FIRST : 0u;
    auto run = [=]() { return b2lz4_compress_blocks_host(gpu, src, n, dst, ocap, flags); };
    try { c->job = std::async(std::launch::async, run); }
    catch (...) { std::promise<size_t> pr; pr.set_value(run()); c->job = pr.get_future(); }      // no thread to be had: run it here
    c->flight = true; c->first = false;
    return 0;
}

// compresses the staged bytes [0, n) into dst now (autoFlush, flush, end); returns bytes written
static size_t emit_staged(LZ4F_cctx* c, size_t n, uint8_t* dst, size_t cap) {
    if (n == 0) return 0;
    Span sp(c->tr.t_gpu); ++c->tr.batches;
    size_t r = b2lz4_compress_blocks_host(c->gpu, c->in.p, n, dst, cap, c->first ? B2LZ4_STREAM_FIRST : 0u);
    if (b2lz4_is_error(r)) return r;
    c->first = false;
    return r;
}

size_t LZ4F_compressUpdate(LZ4F_cctx* c, void* dstBuffer, size_t cap, const void* src, size_t n,
                           const LZ4F_compressOptions_t* opt) {
    (void)opt;
    if (!c || !c->begun) return ec(E_GENERIC);                                       // lz4frame.c:836
    if (n && !src) return ec(E_srcPtr_wrong);
    ++c->tr.calls;
    Span whole(c->tr.t_all);
    uint8_t* dst = (uint8_t*)dstBuffer;
    const bool bc = c->bp.block_checksum != 0;
    // bytes this call may emit: a batch leaves when the NEXT one is full (it is compressed while that one is being
    // staged), so whole batches as before; everything when autoFlush is set
    const size_t have = c->in_len + n;
    const size_t go = c->prefs.autoFlush ? have : (have / c->batch) * c->batch;
    if (cap < blocks_bound(go, c->block_size, bc)) return ec(E_dstMaxSize_tooSmall); // lz4frame.c:837-838
    if (!dst && go) return ec(E_GENERIC);
    const uint8_t* s = (const uint8_t*)src;
    size_t written = 0, left = n;
    const size_t stage_cap = c->prefs.autoFlush ? std::max(have, c->batch) : c->batch;
    {   // short streams should not pay for pinning a whole batch: the staging buffer grows with what has arrived
        const size_t need_in = std::min(stage_cap, std::max<size_t>(have, (size_t)1 << 20));
        const bool full = 2 * need_in >= stage_cap;
        if (!c->in.ensure((full ? stage_cap : need_in) + 16, c->in_len, full)) return ec(E_allocation_failed);
    }
    while (left) {
        const size_t room = stage_cap - c->in_len;
        const size_t take = left < room ? left : room;
        { Span sp(c->tr.t_in); memcpy(c->in.p + c->in_len, s, take); }              // pinned staging (replaces tmpIn, lz4frame.c:846-868)
        c->in_len += take; s += take; left -= take;
        if (c->in_len == c->batch && !c->prefs.autoFlush) {
            size_t r = collect_flight(c, dst + written, cap - written);              // the batch before this one
            if (b2lz4_is_error(r)) return r;
            written += r;
            r = launch_staged(c, c->in_len);
            if (b2lz4_is_error(r)) return r;
            c->in_len = 0;
        }
    }
    if (c->prefs.autoFlush && c->in_len) {
        size_t r = emit_staged(c, c->in_len, dst + written, cap - written);
        if (b2lz4_is_error(r)) return r;
        written += r; c->in_len = 0;
    }
    c->total_in += n;
    return written;
}

size_t LZ4F_flush(LZ4F_cctx* c, void* dstBuffer, size_t cap, const LZ4F_compressOptions_t* opt) {
    (void)opt;
    if (!c || !c->begun) return ec(E_GENERIC);
    uint8_t* dst = (uint8_t*)dstBuffer;
    size_t w = collect_flight(c, dst, cap);                                          // the batch still on the GPU comes first
    if (b2lz4_is_error(w)) return w;
    if (c->in_len == 0) return w;                                                    // lz4frame.c:945
    if (cap - w < blocks_bound(c->in_len, c->block_size, c->bp.block_checksum != 0)) return ec(E_dstMaxSize_tooSmall);
    size_t r = emit_staged(c, c->in_len, dst + w, cap - w);
    if (b2lz4_is_error(r)) return r;
    c->in_len = 0;
    return w + r;
}

size_t LZ4F_compressEnd(LZ4F_cctx* c, void* dstBuffer, size_t cap, const LZ4F_compressOptions_t* opt) {
    if (!c || !c->begun) return ec(E_GENERIC);
    uint8_t* dst = (uint8_t*)dstBuffer;
    size_t w = LZ4F_flush(c, dst, cap, opt);
    if (b2lz4_is_error(w)) return w;
    if (cap - w < 8) return ec(E_dstMaxSize_tooSmall);                               // lz4frame.c:997
    if (c->first) {                                                                  // empty frame: still restart the checksum
        size_t r = b2lz4_compress_blocks_host(c->gpu, nullptr, 0, dst + w, cap - w, B2LZ4_STREAM_FIRST);
        if (b2lz4_is_error(r)) return r;
    }
    size_t e = b2lz4_frame_end(c->gpu, dst + w, cap - w);
    if (b2lz4_is_error(e)) return e;
    c->begun = false;                                                                // reusable after a new compressBegin (lz4frame.c:1009)
    if (c->prefs.frameInfo.contentSize && c->prefs.frameInfo.contentSize != c->total_in) return ec(E_frameSize_wrong);   // lz4frame.c:1012-1015
    return w + e;
}

}  // extern "C"

// =========================================================================================== decompression
// Input is collected in a pinned buffer until a batch of WHOLE blocks is there (or the EndMark arrives), the
// batch is decoded on the GPU into a pinned output buffer, and the output is handed out in the slices the
// caller asks for.  Stages mirror the reference's dStage machine (lz4frame.c:1037-1056) at block granularity.
namespace {
enum DStage { DS_HEADER, DS_SKIP_SIZE, DS_SKIP_DATA, DS_BLOCK_HEADER, DS_BLOCK_DATA, DS_SUFFIX, DS_DONE };
}

struct LZ4F_dctx_s {
    b2lz4_ctx* gpu = nullptr; Engine eng;
    DStage stage = DS_HEADER;
    uint8_t hdr[HEADER_MAX]; uint32_t hdr_len = 0;          // header / small fixed-size elements being collected
    b2lz4_frame_info info{};
    Pinned in;  size_t in_len = 0;                           // whole blocks (+ the one being collected)
    Pinned in2, out2;                                        // the batch in flight reads in2 and writes out2
    std::future<size_t> job; bool flight = false;            // at most one batch is on the GPU while the caller drains `out`
    bool check_sum = false; uint32_t stored_sum = 0;         // content checksum read from the stream, compared when all is decoded
    size_t complete_end = 0; uint32_t n_complete = 0;        // end of the last whole block in `in`, how many there are
    size_t need = 0;                                         // bytes still missing for the element being collected
    uint64_t skip_left = 0;
    Pinned out; size_t out_len = 0, out_pos = 0;
    uint64_t total_out = 0;
    bool first = true;
    size_t pre_acked = 0;                                    // input bytes already buffered but reported as unconsumed
    bool launch_pending = false;                             // a batch of whole blocks is staged and goes to the GPU as soon as no other is in flight
    size_t batch = 0;
    std::vector<uint8_t> dict;                               // LZ4F_decompress_usingDict: last 64 KB of the dictionary of the current frame
    Trace tr;
};

namespace {

void dctx_reset(LZ4F_dctx_s* d) {
    if (d->flight) { d->job.wait(); (void)d->job.get(); d->flight = false; }
    d->stage = DS_HEADER; d->hdr_len = 0; d->in_len = 0; d->complete_end = 0; d->n_complete = 0; d->need = 0;
    d->skip_left = 0; d->out_len = d->out_pos = 0; d->total_out = 0; d->first = true; d->pre_acked = 0;
    d->check_sum = false; d->stored_sum = 0; d->launch_pending = false;
    d->dict.clear();                                             // a dictionary belongs to one frame (LZ4F_decompress_usingDict registers it again)
}

bool job_done(LZ4F_dctx_s* d) { return d->job.wait_for(std::chrono::seconds(0)) == std::future_status::ready; }

// the batch in flight has to finish: its output becomes the one handed out (only called when `out` is drained)
size_t collect(LZ4F_dctx_s* d) {
    if (!d->flight) return 0;
    Span sp(d->tr.t_gpu);
    const size_t r = d->job.get();
    d->flight = false;
    if (b2lz4_is_error(r)) return r;
    std::swap(d->out, d->out2);
    d->out_len = r; d->out_pos = 0; d->total_out += r;
    return 0;
}

// Decode the whole blocks collected so far — on a helper thread, so that the GPU works on this batch while the caller
// drains the previous one and feeds the next (the calls of this layer are host copies for the most part, B2LZ4_TRACE).
// Only called when `out` is drained; the batch before this one, if still in flight, is collected first.
size_t decode_batch(LZ4F_dctx_s* d) {
    size_t r = collect(d);
    if (b2lz4_is_error(r)) return r;
    if (d->n_complete == 0) return 0;
    ++d->tr.batches;
    if (!d->gpu) {
        b2lz4_prefs bp; b2lz4_default_prefs(&bp); bp.use_index = 0;
        r = d->eng.acquire(&bp);
        if (b2lz4_is_error(r)) return r;
        d->gpu = d->eng.gpu;
    }
    const size_t cap = (size_t)d->n_complete * d->info.block_size;
    if (!d->out2.ensure(cap + 16, 0, true)) return ec(E_allocation_failed);
    // the bytes after the last whole block move to the other staging buffer, which becomes the one being filled
    const size_t tail = d->in_len - d->complete_end;
    if (!d->in2.ensure(std::max(tail + 8, d->in.cap / 2))) return ec(E_allocation_failed);
    memcpy(d->in2.p, d->in.p + d->complete_end, tail);
    std::swap(d->in, d->in2);
    if (d->first) {                                          // (also clears the dictionary of an earlier frame)
        r = b2lz4_set_dict(d->gpu, d->dict.empty() ? nullptr : d->dict.data(), d->dict.size(), 0);
        if (b2lz4_is_error(r)) return r;
    }
    b2lz4_ctx* gpu = d->gpu; const b2lz4_frame_info info = d->info;
    const uint8_t* src = d->in2.p; const size_t n = d->complete_end; uint8_t* dst = d->out2.p;
    const unsigned flags = d->first ? B2LZ4_STREAM_FIRST : 0u;
    auto run = [=]() { return b2lz4_decompress_blocks_host(gpu, &info, src, n, dst, cap, flags); };
    try { d->job = std::async(std::launch::async, run); }
    catch (...) { std::promise<size_t> pr; pr.set_value(run()); d->job = pr.get_future(); }      // no thread to be had: run it here
    d->flight = true; d->first = false;
    d->in_len = tail; d->complete_end = 0; d->n_complete = 0;
    return 0;
}

// header bytes complete -> parse; returns 0 or an error code
size_t header_done(LZ4F_dctx_s* d) {
    size_t r = b2lz4_frame_header_parse(d->hdr, d->hdr_len, &d->info);
    if (b2lz4_is_error(r)) return r;
    d->stage = DS_BLOCK_HEADER; d->hdr_len = 0; d->need = 4;
    // a large foreign block is one unit of the dataflow decoder (one CTA, ~0.5 GB/s on log-like data): what makes a batch
    // fast is the number of blocks in it, so batches of large blocks are sized by count (16 blocks = 64 MiB of 4 MB blocks;
    // more would cost more in pinned staging than it gains)
    d->batch = std::max(batch_bytes(), std::min<size_t>((size_t)16 * d->info.block_size, (size_t)64 << 20));
    if (batch_env()) d->batch = batch_bytes();
    return 0;
}

}  // namespace

extern "C" {

LZ4F_errorCode_t LZ4F_createDecompressionContext(LZ4F_dctx** dctxPtr, unsigned version) {
    if (!dctxPtr) return ec(E_GENERIC);
    (void)version;
    LZ4F_dctx* d = new (std::nothrow) LZ4F_dctx_s();
    if (!d) return ec(E_allocation_failed);
    dctx_reset(d);
    *dctxPtr = d;
    return 0;
}

LZ4F_errorCode_t LZ4F_freeDecompressionContext(LZ4F_dctx* d) {
    if (!d) return 0;                                                                // lz4frame.c:1081
    const size_t stage = (size_t)d->stage;                                          // the reference returns the stage, not an error
    if (d->flight) { d->job.wait(); (void)d->job.get(); d->flight = false; }
    d->tr.report("dctx");
    const double t0 = trace_on() ? Trace::now() : 0;
    d->eng.release(); d->gpu = nullptr;
    const double t1 = trace_on() ? Trace::now() : 0;
    d->in.release(); d->out.release(); d->in2.release(); d->out2.release();
    if (trace_on()) fprintf(stderr, "[b2lz4 dctx] free: engine context %.3f s, staging %.3f s\n", t1 - t0, Trace::now() - t1);
    delete d;
    return stage == DS_DONE ? 0 : stage;
}

void LZ4F_resetDecompressionContext(LZ4F_dctx* d) { if (d) dctx_reset(d); }

size_t LZ4F_getFrameInfo(LZ4F_dctx* d, LZ4F_frameInfo_t* fi, const void* src, size_t* srcSizePtr) {
    if (!d || !fi || !srcSizePtr) return ec(E_GENERIC);
    auto fill = [&]() {
        memset(fi, 0, sizeof *fi);
        fi->blockSizeID = (LZ4F_blockSizeID_t)d->info.block_size_id;
        fi->blockMode = d->info.block_independent ? LZ4F_blockIndependent : LZ4F_blockLinked;
        fi->contentChecksumFlag = d->info.content_checksum ? LZ4F_contentChecksumEnabled : LZ4F_noContentChecksum;
        fi->blockChecksumFlag = d->info.block_checksum ? LZ4F_blockChecksumEnabled : LZ4F_noBlockChecksum;
        fi->frameType = LZ4F_frame;
        fi->contentSize = d->info.content_size; fi->dictID = d->info.dict_id;
    };
    if (d->stage != DS_HEADER) {                                                     // header already decoded (lz4frame.c:1251-1257)
        if (d->stage == DS_SKIP_SIZE || d->stage == DS_SKIP_DATA) { *srcSizePtr = 0; return ec(E_frameType_unknown); }
        *srcSizePtr = 0; fill();
        return d->need ? d->need : 4;
    }
    if (d->hdr_len != 0) { *srcSizePtr = 0; return ec(E_frameDecoding_alreadyStarted); }   // lz4frame.c:1258-1262
    const uint8_t* s = (const uint8_t*)src;
    const size_t n = *srcSizePtr;
    *srcSizePtr = 0;
    if (!s || n < 5) return ec(E_frameHeader_incomplete);
    if ((ld_le32(s) & 0xFFFFFFF0u) == MAGIC_SKIP_LO) return ec(E_frameType_unknown);
    if (ld_le32(s) != MAGIC_FRAME) return ec(E_frameType_unknown);
    const size_t hs = header_size_from_flg(s[4]);
    if (n < hs) return ec(E_frameHeader_incomplete);                                 // lz4frame.c:1266-1269
    memcpy(d->hdr, s, hs); d->hdr_len = (uint32_t)hs;
    size_t r = header_done(d);
    if (b2lz4_is_error(r)) { d->hdr_len = 0; return r; }
    *srcSizePtr = hs; fill();
    return 4;                                                                        // next: a block header
}

// lz4frame.c:1842-1860: the dictionary is registered when the frame starts; it is copied (the reference uses it in place)
size_t LZ4F_decompress_usingDict(LZ4F_dctx* d, void* dstBuffer, size_t* dstSizePtr, const void* srcBuffer, size_t* srcSizePtr,
                                 const void* dict, size_t dictSize, const LZ4F_decompressOptions_t* opt) {
    if (!d) return ec(E_GENERIC);
    if (d->stage == DS_HEADER && d->hdr_len == 0 && d->first) {
        const size_t keep = std::min<size_t>(dict ? dictSize : 0, 65536);
        if (keep) d->dict.assign((const uint8_t*)dict + (dictSize - keep), (const uint8_t*)dict + dictSize); else d->dict.clear();
    }
    return LZ4F_decompress(d, dstBuffer, dstSizePtr, srcBuffer, srcSizePtr, opt);
}

size_t LZ4F_decompress(LZ4F_dctx* d, void* dstBuffer, size_t* dstSizePtr, const void* srcBuffer, size_t* srcSizePtr,
                       const LZ4F_decompressOptions_t* opt) {
    (void)opt;
    if (!d || !dstSizePtr || !srcSizePtr) return ec(E_GENERIC);
    const uint8_t* src = (const uint8_t*)srcBuffer;
    uint8_t* dst = (uint8_t*)dstBuffer;
    const size_t srcSize = src ? *srcSizePtr : 0, dstCap = dst ? *dstSizePtr : 0;
    size_t consumed = 0, produced = 0;
    ++d->tr.calls;
    Span whole(d->tr.t_all);
    *srcSizePtr = 0; *dstSizePtr = 0;
    auto fail = [&](size_t e) { dctx_reset(d); return e; };
    {   // bytes taken in an earlier call but handed back so that the caller comes again for the pending output
        const size_t sk = d->pre_acked < srcSize ? d->pre_acked : srcSize;
        consumed += sk; d->pre_acked -= sk;
    }
    bool skipped = false;
    for (;;) {
        if (skipped) break;
        if (d->out_pos < d->out_len) {                                               // hand out decoded bytes first
            const size_t k = std::min(dstCap - produced, d->out_len - d->out_pos);
            { Span sp(d->tr.t_out); memcpy(dst + produced, d->out.p + d->out_pos, k); }
            produced += k; d->out_pos += k;
        }
        const bool dst_full = d->out_pos < d->out_len;
        // `out` is drained: the batch in flight is taken over as soon as it is done, or when nothing else is left to do
        if (!dst_full && d->flight && (d->stage == DS_DONE || d->launch_pending || job_done(d))) {
            size_t r = collect(d);
            if (b2lz4_is_error(r)) return fail(r);
            continue;
        }
        // a staged batch goes to the GPU as soon as no other is in flight — also while the caller is still draining the output
        // of the one before it (the destination being full does not stop the INPUT side: the GPU then works while the caller copies)
        if (d->launch_pending && !d->flight) {
            d->launch_pending = false;
            size_t r = decode_batch(d);
            if (b2lz4_is_error(r)) return fail(r);
            continue;
        }
        if (dst_full && d->launch_pending) break;                                   // (a batch is in flight and the next one is staged: only draining helps)
        if (d->stage == DS_DONE) break;
        if (consumed == srcSize) break;
        const size_t avail = srcSize - consumed;
        switch (d->stage) {
        case DS_HEADER: {
            // 5 bytes tell the header size; then the rest of it (lz4frame.c:1395-1432)
            size_t want = d->hdr_len < 5 ? 5 : 0;
            if (!want) {
                const uint32_t magic = ld_le32(d->hdr);
                if ((magic & 0xFFFFFFF0u) == MAGIC_SKIP_LO) want = 8;                // skippable frame: magic + size
                else if (magic != MAGIC_FRAME) return fail(ec(E_frameType_unknown));
                else want = header_size_from_flg(d->hdr[4]);
            }
            const size_t k = std::min(avail, want - d->hdr_len);
            memcpy(d->hdr + d->hdr_len, src + consumed, k); d->hdr_len += (uint32_t)k; consumed += k;
            if (d->hdr_len >= 5 && (ld_le32(d->hdr) & 0xFFFFFFF0u) == MAGIC_SKIP_LO) {
                if (d->hdr_len == 8) { d->skip_left = ld_le32(d->hdr + 4); d->hdr_len = 0; d->stage = DS_SKIP_DATA; if (d->skip_left == 0) skipped = true; }
            } else if (d->hdr_len >= 5 && ld_le32(d->hdr) == MAGIC_FRAME && d->hdr_len == header_size_from_flg(d->hdr[4])) {
                size_t r = header_done(d);
                if (b2lz4_is_error(r)) return fail(r);
            }
            break;
        }
        case DS_SKIP_DATA: {                                                         // lz4frame.c:1756-1801
            const size_t k = (size_t)std::min<uint64_t>(avail, d->skip_left);
            consumed += k; d->skip_left -= k;
            if (d->skip_left == 0) skipped = true;                                   // frame boundary: reset and return 0 (lz4frame.c:1789-1799)
            break;
        }
        case DS_BLOCK_HEADER: {
            if (!d->in.ensure(d->in_len + 8, d->in_len)) return fail(ec(E_allocation_failed));
            const size_t k = std::min(avail, d->need);
            memcpy(d->in.p + d->in_len, src + consumed, k); d->in_len += k; consumed += k; d->need -= k;
            if (d->need) break;
            const uint32_t w = ld_le32(d->in.p + d->in_len - 4);
            const uint32_t cs = w & 0x7FFFFFFFu;
            if (cs == 0) {                                                           // EndMark (lz4frame.c:1485-1488)
                d->in_len -= 4;
                d->launch_pending = d->n_complete != 0;                              // what is staged is the frame's last batch
                d->stage = DS_SUFFIX; d->need = d->info.content_checksum ? 4 : 0; d->hdr_len = 0;
                if (d->need == 0) d->stage = DS_DONE;                                // size / checksum are checked when all is decoded
                break;
            }
            if (cs > d->info.block_size) return fail(ec(E_maxBlockSize_invalid));     // lz4frame.c:1489-1490
            d->need = (size_t)cs + (d->info.block_checksum ? 4u : 0u);
            d->stage = DS_BLOCK_DATA;
            break;
        }
        case DS_BLOCK_DATA: {
            if (!d->in.ensure(d->in_len + d->need + 8, d->in_len)) return fail(ec(E_allocation_failed));
            const size_t k = std::min(avail, d->need);
            { Span sp(d->tr.t_in); memcpy(d->in.p + d->in_len, src + consumed, k); }
            d->in_len += k; consumed += k; d->need -= k;
            if (d->need) break;
            d->complete_end = d->in_len; ++d->n_complete;
            d->stage = DS_BLOCK_HEADER; d->need = 4;
            // enough for the GPU: the batch's output bound or its compressed size reached the batch size
            if ((size_t)(d->n_complete + 1) * d->info.block_size > d->batch || d->complete_end >= d->batch) d->launch_pending = true;
            break;
        }
        case DS_SUFFIX: {                                                            // content checksum (lz4frame.c:1734-1745)
            const size_t k = std::min(avail, d->need);
            memcpy(d->hdr + d->hdr_len, src + consumed, k); d->hdr_len += (uint32_t)k; consumed += k; d->need -= k;
            if (d->need) break;
            d->stored_sum = ld_le32(d->hdr); d->check_sum = true;
            d->stage = DS_DONE;
            break;
        }
        default: break;
        }
    }
    if (skipped) {                                                                   // a skippable frame ends like any frame: context reset, 0
        *srcSizePtr = consumed; *dstSizePtr = produced;
        dctx_reset(d);
        return 0;
    }
    const bool pending = d->out_pos < d->out_len || ((d->flight || d->launch_pending) && d->stage == DS_DONE);
    if (d->stage == DS_DONE && !pending) {                                           // everything is decoded: the frame's own checks
        if (d->check_sum) {                                                          // content checksum (lz4frame.c:1734-1745)
            uint32_t h = 0x02CC5D05u;                                                // XXH32 of no bytes, seed 0 (frame without blocks)
            if (!d->first) { size_t r = b2lz4_content_checksum(d->gpu, &h); if (b2lz4_is_error(r)) return fail(r); }
            if (h != d->stored_sum) return fail(ec(E_contentChecksum_invalid));
        }
        if (d->info.has_content_size && d->info.content_size != d->total_out) return fail(ec(E_frameSize_wrong));   // lz4frame.c:1715-1716
    }
    if (pending && consumed == srcSize && consumed > 0) { --consumed; ++d->pre_acked; }   // make the caller come back for the rest
    *srcSizePtr = consumed; *dstSizePtr = produced;
    if (d->stage == DS_DONE && !pending) {                                           // frame complete (lz4frame.h:457-468: returns 0)
        const size_t keep = d->pre_acked;
        dctx_reset(d); d->pre_acked = keep;
        return 0;
    }
    // hint: the caller may offer this much room.  Capped: the conduit sizes its next destination by the hint and then copies
    // the bytes once more into the ByteString it yields (Conduit.hsc:665, 676) — a destination that stays in the CPU's cache
    // makes that second copy cheap, a 64 MiB one does not
    if (pending) return std::max<size_t>(1, std::min<size_t>(d->out_len - d->out_pos, HINT_MAX));
    return d->need ? d->need + (d->stage == DS_BLOCK_DATA ? 4 : 0) : 4;
}

}  // extern "C"
