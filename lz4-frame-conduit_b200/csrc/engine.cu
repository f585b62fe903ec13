// engine.cu — host engine + C ABI (include/b200lz4.h) of the B200 LZ4 frame codec.
//
// One context owns a CUDA device, two streams (codec + content checksum) and reusable device
// workspaces; every compress / decompress call is a short fixed sequence of kernel launches followed
// by ONE small device->host copy of the result struct (no per-block host round trips).
// Replaces the reference's LZ4F_cctx / LZ4F_dctx life cycle (lz4/lib/lz4frame.c:536-563, 1065-1090)
// and the per-16 KB FFI call pattern of Conduit.hsc:308-314, 585-596 with batch calls.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <stddef.h>
#include <new>
#include <algorithm>
#include "kernels.h"
#include "../../include/b200lz4.h"

using namespace b2;

namespace {

inline size_t err_code(uint32_t e) { return (size_t)0 - (size_t)e; }

const char* const kErrNames[B2LZ4_ERROR_maxCode + 1] = {      // same strings as lz4frame.h:503-524 for 0..19
    "OK_NoError", "ERROR_GENERIC", "ERROR_maxBlockSize_invalid", "ERROR_blockMode_invalid",
    "ERROR_contentChecksumFlag_invalid", "ERROR_compressionLevel_invalid", "ERROR_headerVersion_wrong",
    "ERROR_blockChecksum_invalid", "ERROR_reservedFlag_set", "ERROR_allocation_failed",
    "ERROR_srcSize_tooLarge", "ERROR_dstMaxSize_tooSmall", "ERROR_frameHeader_incomplete",
    "ERROR_frameType_unknown", "ERROR_frameSize_wrong", "ERROR_srcPtr_wrong", "ERROR_decompressionFailed",
    "ERROR_headerChecksum_invalid", "ERROR_contentChecksum_invalid", "ERROR_frameDecoding_alreadyStarted",
    "ERROR_cudaFailure", "ERROR_noEndMark", "ERROR_maxCode" };

struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = n + n / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct Small {                 // small device-side state, mirrored to pinned host memory after each call
    FrameIndex fi;
    DecodeResult dres;
    EncodeResult eres;
    XxhState xxh;
    uint32_t ticket_dec, ticket_enc, crc_bad, hash_out;
};

}  // namespace

struct b2lz4_ctx {
    int device = 0;
    b2lz4_prefs prefs{};
    cudaStream_t st = nullptr, st2 = nullptr;     // codec stream, checksum stream
    cudaEvent_t ev[8] = {};
    cudaEvent_t ev_in = nullptr, ev_sum = nullptr;
    Small* d_small = nullptr;
    Small* h_small = nullptr;                     // pinned
    DevBuf scratch, subs, places, bplaces, blocks, pieces, chains, outs, stage_in, stage_out, enc_tabs, ring;
    size_t ring_pos = 0; uint32_t ring_hist = 0;     // streaming linked-block decode: next output offset / valid history bytes
    b2lz4_stats stats{};
    char cuda_err[256] = {0};
    int sm_count = 148;
};

namespace {

#define CU(call)                                                                                   \
    do { cudaError_t e_ = (call);                                                                  \
         if (e_ != cudaSuccess) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "%s: %s (%s:%d)", #call, \
                                           cudaGetErrorString(e_), __FILE__, __LINE__);            \
                                  cudaGetLastError(); return err_code(E_cudaFailure); } } while (0)

void fill_desc(const b2lz4_prefs& p, FrameDesc& d, uint64_t content_size) {
    memset(&d, 0, sizeof d);
    d.block_size_id = p.block_size_id ? p.block_size_id : 7u;
    d.block_independent = 1;
    d.block_checksum = p.block_checksum ? 1u : 0u;
    d.content_checksum = p.content_checksum ? 1u : 0u;
    // like the reference, the field is only emitted for a non-zero size (lz4frame.c:679, 686-690)
    d.has_content_size = (p.content_size && content_size) ? 1u : 0u;
    d.content_size = content_size;
}

uint32_t sub_size_of(const b2lz4_prefs& p, uint32_t block_size) {
    uint32_t lg = p.sub_block_log2 ? p.sub_block_log2 : 18u;
    lg = std::min(22u, std::max(16u, lg));
    return std::min(block_size, 1u << lg);
}

bool prefs_valid(const b2lz4_prefs& p) {
    if (p.block_size_id != 0 && (p.block_size_id < 4 || p.block_size_id > 7)) return false;
    if (p.sub_block_log2 != 0 && (p.sub_block_log2 < 16 || p.sub_block_log2 > 22)) return false;
    return true;
}

size_t set_device(b2lz4_ctx* ctx) { CU(cudaSetDevice(ctx->device)); return 0; }

}  // namespace

// ====================================================================================== C ABI
extern "C" {

unsigned b2lz4_is_error(size_t code) { return code > err_code(B2LZ4_ERROR_maxCode) ? 1u : 0u; }

const char* b2lz4_error_name(size_t code) {
    if (!b2lz4_is_error(code)) return "Unspecified error code";
    size_t e = (size_t)0 - code;
    return e < (size_t)B2LZ4_ERROR_maxCode ? kErrNames[e] : "Unspecified error code";
}

const char* b2lz4_last_cuda_error(const b2lz4_ctx* ctx) { return ctx ? ctx->cuda_err : ""; }

void b2lz4_default_prefs(b2lz4_prefs* p) {
    memset(p, 0, sizeof *p);
    p->block_size_id = 7; p->sub_block_log2 = 18; p->use_index = 1;
    // write_index stays 0: the stock lz4 CLI cannot skip a skippable frame on a pipe (lz4io.c:1124-1133 uses fseek),
    // so plain frames are the drop-in default and the seekable index is opt-in.
}

int b2lz4_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

size_t b2lz4_compress_bound(size_t n, const b2lz4_prefs* prefs) {
    b2lz4_prefs p; if (prefs) p = *prefs; else b2lz4_default_prefs(&p);
    if (!prefs_valid(p)) return err_code(E_maxBlockSize_invalid);
    size_t bs = block_size_of(p.block_size_id ? p.block_size_id : 7u);
    size_t ss = sub_size_of(p, (uint32_t)bs);
    size_t nb = (n + bs - 1) / bs, ns = (n + ss - 1) / ss + nb;
    size_t frame = HEADER_MAX + nb * 8 + n + 4 + 4;                    // blocks are stored raw at worst
    size_t index = p.write_index ? 8 + sizeof(IndexHead) + nb * sizeof(IndexBlock) + ns * sizeof(IndexChain) + 8 : 0;
    return frame + index;
}

size_t b2lz4_frame_header_write(void* dst, size_t cap, const b2lz4_prefs* prefs, uint64_t content_size) {
    b2lz4_prefs p; if (prefs) p = *prefs; else b2lz4_default_prefs(&p);
    if (!prefs_valid(p)) return err_code(E_maxBlockSize_invalid);
    if (cap < (size_t)HEADER_MAX) return err_code(E_dstMaxSize_tooSmall);
    FrameDesc d; fill_desc(p, d, content_size);
    return header_write((uint8_t*)dst, d);
}

size_t b2lz4_frame_header_parse(const void* src, size_t n, b2lz4_frame_info* info) {
    FrameDesc d{};
    Err e = header_parse((const uint8_t*)src, n, d);
    if (e != OK) return err_code(e);
    if (info) {
        info->block_size_id = d.block_size_id; info->block_independent = d.block_independent;
        info->block_checksum = d.block_checksum; info->content_checksum = d.content_checksum;
        info->has_content_size = d.has_content_size; info->has_dict_id = d.has_dict_id;
        info->dict_id = d.dict_id; info->header_size = d.header_size;
        info->content_size = d.content_size; info->block_size = block_size_of(d.block_size_id);
    }
    return d.header_size;
}

size_t b2lz4_create(b2lz4_ctx** out, int device, const b2lz4_prefs* prefs) {
    if (!out) return err_code(E_GENERIC);
    *out = nullptr;
    b2lz4_ctx* ctx = new (std::nothrow) b2lz4_ctx();
    if (!ctx) return err_code(E_allocation_failed);
    ctx->device = device;
    if (prefs) ctx->prefs = *prefs; else b2lz4_default_prefs(&ctx->prefs);
    if (!prefs_valid(ctx->prefs)) { delete ctx; return err_code(E_maxBlockSize_invalid); }
    auto fail = [&](cudaError_t e, const char* what) -> size_t {
        fprintf(stderr, "b200lz4: %s failed: %s\n", what, cudaGetErrorString(e));
        cudaGetLastError();
        b2lz4_destroy(ctx);
        return err_code(E_cudaFailure);
    };
    cudaError_t e;
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(e, "cudaSetDevice");
    if ((e = cudaStreamCreateWithFlags(&ctx->st, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "cudaStreamCreate");
    if ((e = cudaStreamCreateWithFlags(&ctx->st2, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "cudaStreamCreate");
    for (auto& v : ctx->ev) if ((e = cudaEventCreate(&v)) != cudaSuccess) return fail(e, "cudaEventCreate");
    if ((e = cudaEventCreateWithFlags(&ctx->ev_in, cudaEventDisableTiming)) != cudaSuccess) return fail(e, "cudaEventCreate");
    if ((e = cudaEventCreateWithFlags(&ctx->ev_sum, cudaEventDisableTiming)) != cudaSuccess) return fail(e, "cudaEventCreate");
    if ((e = cudaMalloc(&ctx->d_small, sizeof(Small))) != cudaSuccess) return fail(e, "cudaMalloc");
    if ((e = cudaMemset(ctx->d_small, 0, sizeof(Small))) != cudaSuccess) return fail(e, "cudaMemset");
    if ((e = cudaMallocHost(&ctx->h_small, sizeof(Small))) != cudaSuccess) return fail(e, "cudaMallocHost");
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    if ((e = cudaFuncSetAttribute(find_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)find_emit_smem_bytes())) != cudaSuccess) return fail(e, "cudaFuncSetAttribute");
    *out = ctx;
    return 0;
}

size_t b2lz4_destroy(b2lz4_ctx* ctx) {
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    if (ctx->st) cudaStreamSynchronize(ctx->st);
    if (ctx->st2) cudaStreamSynchronize(ctx->st2);
    for (DevBuf* b : { &ctx->scratch, &ctx->subs, &ctx->places, &ctx->bplaces, &ctx->blocks, &ctx->pieces,
                       &ctx->chains, &ctx->outs, &ctx->stage_in, &ctx->stage_out, &ctx->enc_tabs, &ctx->ring }) b->release();
    if (ctx->d_small) cudaFree(ctx->d_small);
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    for (auto& v : ctx->ev) if (v) cudaEventDestroy(v);
    if (ctx->ev_in) cudaEventDestroy(ctx->ev_in);
    if (ctx->ev_sum) cudaEventDestroy(ctx->ev_sum);
    if (ctx->st) cudaStreamDestroy(ctx->st);
    if (ctx->st2) cudaStreamDestroy(ctx->st2);
    cudaGetLastError();
    delete ctx;
    return 0;
}

size_t b2lz4_set_prefs(b2lz4_ctx* ctx, const b2lz4_prefs* prefs) {
    if (!ctx || !prefs) return err_code(E_GENERIC);
    if (!prefs_valid(*prefs)) return err_code(E_maxBlockSize_invalid);
    ctx->prefs = *prefs;
    return 0;
}

size_t b2lz4_last_stats(const b2lz4_ctx* ctx, b2lz4_stats* out) {
    if (!ctx || !out) return err_code(E_GENERIC);
    *out = ctx->stats;
    return 0;
}

void* b2lz4_pinned_alloc(size_t n) {
    void* p = nullptr;
    if (cudaMallocHost(&p, n ? n : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void b2lz4_pinned_free(void* p) { if (p) cudaFreeHost(p); }

// -------------------------------------------------------------------------------------- compress
// Launch sequence on ctx->st: find_emit -> layout -> gather -> [block checksums] ; the content checksum
// runs on ctx->st2 over the input, concurrently, and its digest is stored once layout knows the offset.
enum : unsigned { CL_WHOLE_FRAME = 1u, CL_SUM_RESET = 2u, CL_SUM_UPDATE = 4u };
static size_t compress_launch(b2lz4_ctx* ctx, const uint8_t* d_src, size_t n, uint8_t* d_dst, size_t cap,
                              unsigned flags)
{
    const bool whole_frame = (flags & CL_WHOLE_FRAME) != 0;
    const b2lz4_prefs& pr = ctx->prefs;
    EncParams P{};
    P.n = n;
    fill_desc(pr, P.desc, n);
    P.block_size = block_size_of(P.desc.block_size_id);
    P.sub_size = sub_size_of(pr, P.block_size);
    P.subs_per_block = P.block_size / P.sub_size;
    uint64_t nb = (n + P.block_size - 1) / P.block_size;
    if (nb > 0x7FFFFFFFull / P.subs_per_block) return err_code(E_srcSize_tooLarge);
    P.n_blocks = (uint32_t)nb;
    if (nb) {
        uint64_t last = n - (nb - 1) * P.block_size;
        P.n_subs = (uint32_t)((nb - 1) * P.subs_per_block + (last + P.sub_size - 1) / P.sub_size);
    }
    P.scratch_stride = (P.sub_size + P.sub_size / 255u + 32u + 15u) & ~15u;
    P.block_checksum = P.desc.block_checksum;
    P.content_checksum = whole_frame ? P.desc.content_checksum : 0u;
    P.write_header = whole_frame ? 1u : 0u;
    P.write_index = (whole_frame && pr.write_index) ? 1u : 0u;

    CU(ctx->scratch.ensure((size_t)P.n_subs * P.scratch_stride));
    CU(ctx->subs.ensure((size_t)P.n_subs * sizeof(SubOut)));
    CU(ctx->places.ensure((size_t)P.n_subs * sizeof(SubPlace)));
    CU(ctx->bplaces.ensure((size_t)P.n_blocks * sizeof(BlockPlace)));

    Small* ds = ctx->d_small;
    uint32_t launches = 0;
    const bool sum = P.desc.content_checksum != 0 && (flags & (CL_SUM_RESET | CL_SUM_UPDATE)) != 0;
    if (sum) {                                   // serial XXH32 of the input on the second stream
        CU(cudaEventRecord(ctx->ev_in, ctx->st));
        CU(cudaStreamWaitEvent(ctx->st2, ctx->ev_in, 0));
        if (flags & CL_SUM_RESET) { xxh32_reset_kernel<<<1, 32, 0, ctx->st2>>>(&ds->xxh); ++launches; }
        if (flags & CL_SUM_UPDATE) { xxh32_stream_kernel<<<1, 256, 0, ctx->st2>>>(d_src, n, nullptr, nullptr, &ds->xxh); ++launches; }
        CU(cudaEventRecord(ctx->ev_sum, ctx->st2));
    }
    CU(cudaEventRecord(ctx->ev[0], ctx->st));
    if (P.n_subs) {
        uint32_t grid = std::min<uint32_t>((P.n_subs + ENC_WARPS - 1) / ENC_WARPS, (uint32_t)ctx->sm_count * (uint32_t)ENC_CTAS_PER_SM);
        CU(ctx->enc_tabs.ensure(find_emit_gtab_bytes(grid)));
        find_emit_kernel<<<grid, ENC_WARPS * 32, find_emit_smem_bytes(), ctx->st>>>(
            d_src, P, (uint8_t*)ctx->scratch.p, (SubOut*)ctx->subs.p, &ds->ticket_enc, (uint16_t*)ctx->enc_tabs.p);
        ++launches;
    }
    CU(cudaEventRecord(ctx->ev[1], ctx->st));
    layout_kernel<<<1, 1024, 0, ctx->st>>>(P, (const SubOut*)ctx->subs.p, (SubPlace*)ctx->places.p,
                                           (BlockPlace*)ctx->bplaces.p, d_dst, cap, &ds->eres, &ds->ticket_enc);
    ++launches;
    if (P.n_subs) {
        uint32_t grid = std::min<uint32_t>(P.n_subs, (uint32_t)ctx->sm_count * 8u);
        gather_kernel<<<grid, 256, 0, ctx->st>>>(d_src, P, (const uint8_t*)ctx->scratch.p, (const SubOut*)ctx->subs.p,
                                                 (const SubPlace*)ctx->places.p, (const BlockPlace*)ctx->bplaces.p,
                                                 d_dst, &ds->eres);
        ++launches;
        if (P.block_checksum) {
            block_checksums_kernel<<<(P.n_blocks + 127) / 128, 128, 0, ctx->st>>>(P, (const BlockPlace*)ctx->bplaces.p,
                                                                                 d_dst, &ds->eres);
            ++launches;
        }
    }
    if (sum) {
        CU(cudaStreamWaitEvent(ctx->st, ctx->ev_sum, 0));
        if (whole_frame) {
            xxh32_digest_kernel<<<1, 32, 0, ctx->st>>>(&ds->xxh, &ds->hash_out, d_dst, &ds->eres.content_checksum_off);
            ++launches;
        }
    }
    CU(cudaEventRecord(ctx->ev[2], ctx->st));
    CU(cudaMemcpyAsync(&ctx->h_small->eres, &ds->eres, sizeof(EncodeResult), cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaGetLastError());
    ctx->stats.kernel_launches = launches;
    ctx->stats.n_blocks = P.n_blocks;
    return 0;
}

static size_t compress_finish(b2lz4_ctx* ctx, size_t n) {
    CU(cudaStreamSynchronize(ctx->st));
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&b, ctx->ev[0], ctx->ev[2]);
    ctx->stats.ms_main = a; ctx->stats.ms_total = b; ctx->stats.ms_aux = b - a;
    const EncodeResult& r = ctx->h_small->eres;
    ctx->stats.n_chains = r.n_chains; ctx->stats.indexed = r.index_off ? 1u : 0u;
    ctx->stats.bytes_in = n; ctx->stats.bytes_out = r.frame_len; ctx->stats.retries = 0;
    if (r.err) return err_code(r.err);
    return (size_t)r.frame_len;
}

size_t b2lz4_compress_device(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap) {
    if (!ctx || (!d_src && n) || !d_dst) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    ctx->stats = b2lz4_stats{};
    r = compress_launch(ctx, (const uint8_t*)d_src, n, (uint8_t*)d_dst, cap, CL_WHOLE_FRAME | CL_SUM_RESET | CL_SUM_UPDATE);
    if (r) return r;
    return compress_finish(ctx, n);
}

// -------------------------------------------------------------------------------------- decompress
static size_t decompress_launch(b2lz4_ctx* ctx, const uint8_t* d_src, size_t n, uint8_t* d_dst, size_t cap,
                                uint32_t mode, uint32_t cap_tables, uint32_t hist)
{
    CU(ctx->blocks.ensure((size_t)cap_tables * sizeof(BlockRef)));
    CU(ctx->pieces.ensure((size_t)cap_tables * sizeof(Piece)));
    CU(ctx->chains.ensure((size_t)cap_tables * sizeof(Chain)));
    CU(ctx->outs.ensure((size_t)cap_tables * sizeof(PieceOut)));
    Small* ds = ctx->d_small;
    uint32_t launches = 0;
    CU(cudaMemsetAsync(&ds->crc_bad, 0xFF, sizeof(uint32_t), ctx->st));
    CU(cudaEventRecord(ctx->ev[0], ctx->st));
    frame_scan_kernel<<<1, 256, 0, ctx->st>>>(d_src, n, cap, mode, &ds->fi, (BlockRef*)ctx->blocks.p, cap_tables,
                                              (Piece*)ctx->pieces.p, cap_tables, (Chain*)ctx->chains.p);
    verify_block_checksums_kernel<<<std::max(1u, std::min(cap_tables / 128u, 1024u)), 128, 0, ctx->st>>>(
        d_src, &ds->fi, (const BlockRef*)ctx->blocks.p, &ds->crc_bad);
    CU(cudaEventRecord(ctx->ev[1], ctx->st));
    decode_chains_kernel<<<ctx->sm_count * (64 / DEC_WARPS), DEC_WARPS * 32, 0, ctx->st>>>(
        d_src, d_dst, cap, &ds->fi, (const BlockRef*)ctx->blocks.p, (const Piece*)ctx->pieces.p,
        (const Chain*)ctx->chains.p, (PieceOut*)ctx->outs.p, &ds->ticket_dec, hist);
    CU(cudaEventRecord(ctx->ev[2], ctx->st));
    decode_finalize_kernel<<<1, 256, 0, ctx->st>>>(&ds->fi, (const Piece*)ctx->pieces.p, (const PieceOut*)ctx->outs.p,
                                                   &ds->crc_bad, &ds->dres, &ds->ticket_dec);
    // content checksum over the output (lz4frame.c:1676-1677, 1734-1745); the kernels return at once
    // when the frame carries none
    if (!(mode & SCAN_BATCH)) {                   // a batch of a streamed frame keeps its running checksum (caller updates it)
        xxh32_reset_kernel<<<1, 32, 0, ctx->st>>>(&ds->xxh);
        xxh32_stream_kernel<<<1, 256, 0, ctx->st>>>(d_dst, 0, &ds->dres.total_out, &ds->fi.desc.content_checksum, &ds->xxh);
        content_check_kernel<<<1, 32, 0, ctx->st>>>(&ds->fi, &ds->xxh, &ds->dres);
        launches += 3;
    }
    launches += 4;
    CU(cudaEventRecord(ctx->ev[3], ctx->st));
    CU(cudaMemcpyAsync(&ctx->h_small->fi, &ds->fi, sizeof(FrameIndex) + sizeof(DecodeResult), cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaGetLastError());
    ctx->stats.kernel_launches += launches;
    return 0;
}

static_assert(offsetof(Small, dres) == sizeof(FrameIndex), "fi and dres must be contiguous for the single copy");

static size_t decompress_device_impl(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, size_t* consumed,
                                     uint32_t hist, uint32_t mode0 = 0u);
size_t b2lz4_decompress_device(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, size_t* consumed) {
    return decompress_device_impl(ctx, d_src, n, d_dst, cap, consumed, 0u);
}
static size_t decompress_device_impl(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, size_t* consumed,
                                     uint32_t hist, uint32_t mode0) {
    if (consumed) *consumed = 0;
    if (!ctx || !d_src || (!d_dst && cap)) return err_code(E_GENERIC);
    if (n < (size_t)HEADER_MIN) return err_code(E_frameHeader_incomplete);
    size_t r = set_device(ctx); if (r) return r;
    ctx->stats = b2lz4_stats{};
    static uint8_t* dummy = nullptr;
    uint8_t* dst = (uint8_t*)d_dst;
    if (!dst) { if (!dummy) CU(cudaMalloc(&dummy, 256)); dst = dummy; }
    uint32_t mode = mode0 | (ctx->prefs.use_index ? SCAN_USE_INDEX : 0u);
    uint32_t cap_tables = (uint32_t)std::min<uint64_t>(0x7FFFFFFFull, std::max<uint64_t>(4096, (n >> 12) + (cap >> 16) + 64));
    for (int attempt = 0; attempt < 4; ++attempt) {
        r = decompress_launch(ctx, (const uint8_t*)d_src, n, dst, cap, mode, cap_tables, hist);
        if (r) return r;
        CU(cudaStreamSynchronize(ctx->st));
        const FrameIndex& fi = ctx->h_small->fi;
        const DecodeResult& dr = ctx->h_small->dres;
        if (fi.err == E_allocation_failed && attempt < 3) {            // tables too small: grow and rerun
            cap_tables = std::max(fi.need_blocks, fi.need_pieces) + 16;
            ++ctx->stats.retries; continue;
        }
        if (fi.err == OK && !fi.layout_known && fi.n_chains > 1 && !(mode & SCAN_SINGLE_CHAIN) &&
            (dr.err != OK || !dr.layout_ok) && attempt < 3) {
            // the "block b starts at b * blockSize" layout did not hold (a middle block decoded short: flushed
            // stream) or an error may stem from it: redo with the blocks chained in order, which is exact
            mode = (mode & ~SCAN_USE_INDEX) | SCAN_SINGLE_CHAIN;
            ++ctx->stats.retries; continue;
        }
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, ctx->ev[0], ctx->ev[1]);
        cudaEventElapsedTime(&b, ctx->ev[1], ctx->ev[2]);
        cudaEventElapsedTime(&c, ctx->ev[0], ctx->ev[3]);
        ctx->stats.ms_main = b; ctx->stats.ms_total = c; ctx->stats.ms_aux = c - b;
        ctx->stats.n_blocks = fi.n_blocks; ctx->stats.n_chains = fi.n_chains; ctx->stats.indexed = fi.indexed;
        ctx->stats.bytes_in = n; ctx->stats.bytes_out = dr.total_out;
        if (dr.err) return err_code(dr.err);
        if (consumed) *consumed = (size_t)fi.consumed_with_index;
        return (size_t)dr.total_out;
    }
    return err_code(E_GENERIC);
}

// -------------------------------------------------------------------------------------- host buffers
size_t b2lz4_compress_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap) {
    if (!ctx || (!src && n) || !dst) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    ctx->stats = b2lz4_stats{};
    size_t bound = b2lz4_compress_bound(n, &ctx->prefs);
    CU(ctx->stage_in.ensure(n + 16));
    CU(ctx->stage_out.ensure(bound + 16));
    CU(cudaEventRecord(ctx->ev[4], ctx->st));
    if (n) CU(cudaMemcpyAsync(ctx->stage_in.p, src, n, cudaMemcpyHostToDevice, ctx->st));
    CU(cudaEventRecord(ctx->ev[5], ctx->st));
    r = compress_launch(ctx, (const uint8_t*)ctx->stage_in.p, n, (uint8_t*)ctx->stage_out.p, std::min(bound, cap), CL_WHOLE_FRAME | CL_SUM_RESET | CL_SUM_UPDATE);
    if (r) return r;
    r = compress_finish(ctx, n);
    if (b2lz4_is_error(r)) return r;
    CU(cudaEventRecord(ctx->ev[6], ctx->st));
    CU(cudaMemcpyAsync(dst, ctx->stage_out.p, r, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaEventRecord(ctx->ev[7], ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    cudaEventElapsedTime(&ctx->stats.ms_h2d, ctx->ev[4], ctx->ev[5]);
    cudaEventElapsedTime(&ctx->stats.ms_d2h, ctx->ev[6], ctx->ev[7]);
    return r;
}

size_t b2lz4_decompress_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, size_t* consumed) {
    if (consumed) *consumed = 0;
    if (!ctx || !src || (!dst && cap)) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    CU(ctx->stage_in.ensure(n + 16));
    CU(ctx->stage_out.ensure(cap + 16));
    cudaEvent_t e0 = ctx->ev[4], e1 = ctx->ev[5], e2 = ctx->ev[6], e3 = ctx->ev[7];
    CU(cudaEventRecord(e0, ctx->st));
    CU(cudaMemcpyAsync(ctx->stage_in.p, src, n, cudaMemcpyHostToDevice, ctx->st));
    CU(cudaEventRecord(e1, ctx->st));
    r = b2lz4_decompress_device(ctx, ctx->stage_in.p, n, ctx->stage_out.p, cap, consumed);
    if (b2lz4_is_error(r)) return r;
    CU(cudaEventRecord(e2, ctx->st));
    if (r) CU(cudaMemcpyAsync(dst, ctx->stage_out.p, r, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaEventRecord(e3, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    cudaEventElapsedTime(&ctx->stats.ms_h2d, e0, e1);
    cudaEventElapsedTime(&ctx->stats.ms_d2h, e2, e3);
    return r;
}

// -------------------------------------------------------------------------------------- streaming (batches of blocks)
size_t b2lz4_compress_blocks_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, unsigned flags) {
    if (!ctx || (!src && n) || !dst) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    ctx->stats = b2lz4_stats{};
    if (n == 0) {                                                   // nothing to emit, but the checksum may have to restart
        if ((flags & B2LZ4_STREAM_FIRST) && ctx->prefs.content_checksum) {
            xxh32_reset_kernel<<<1, 32, 0, ctx->st>>>(&ctx->d_small->xxh);
            CU(cudaStreamSynchronize(ctx->st));
        }
        return 0;
    }
    size_t bound = b2lz4_compress_bound(n, &ctx->prefs);
    CU(ctx->stage_in.ensure(n + 16));
    CU(ctx->stage_out.ensure(bound + 16));
    CU(cudaMemcpyAsync(ctx->stage_in.p, src, n, cudaMemcpyHostToDevice, ctx->st));
    r = compress_launch(ctx, (const uint8_t*)ctx->stage_in.p, n, (uint8_t*)ctx->stage_out.p, std::min(bound, cap),
                        CL_SUM_UPDATE | ((flags & B2LZ4_STREAM_FIRST) ? CL_SUM_RESET : 0u));
    if (r) return r;
    r = compress_finish(ctx, n);
    if (b2lz4_is_error(r)) return r;
    CU(cudaMemcpyAsync(dst, ctx->stage_out.p, r, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    return r;
}

size_t b2lz4_content_checksum(b2lz4_ctx* ctx, uint32_t* out) {
    if (!ctx || !out) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    Small* ds = ctx->d_small;
    CU(cudaStreamSynchronize(ctx->st2));
    xxh32_digest_kernel<<<1, 32, 0, ctx->st>>>(&ds->xxh, &ds->hash_out, nullptr, nullptr);
    CU(cudaMemcpyAsync(&ctx->h_small->hash_out, &ds->hash_out, 4, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    CU(cudaGetLastError());
    *out = ctx->h_small->hash_out;
    return 0;
}

size_t b2lz4_frame_end(b2lz4_ctx* ctx, void* dst, size_t cap) {
    if (!ctx || !dst) return err_code(E_GENERIC);
    const size_t need = 4u + (ctx->prefs.content_checksum ? 4u : 0u);
    if (cap < need) return err_code(E_dstMaxSize_tooSmall);
    uint8_t* d = (uint8_t*)dst;
    st_le32(d, 0u);                                                 // EndMark (lz4frame.c:998-1000)
    if (ctx->prefs.content_checksum) {
        uint32_t h = 0;
        size_t r = b2lz4_content_checksum(ctx, &h); if (r) return r;
        st_le32(d + 4, h);
    }
    return need;
}

// keeps `hist` bytes of earlier output directly in front of the place the next batch decodes to
static size_t ring_reserve(b2lz4_ctx* ctx, size_t cap) {
    const size_t H = 65536;
    const size_t want = 2 * H + 2 * cap + 256;
    if (ctx->ring.cap < want) {                                     // grow, carrying the history over
        DevBuf nb;
        CU(nb.ensure(want));
        if (ctx->ring_hist) CU(cudaMemcpyAsync((uint8_t*)nb.p + H - ctx->ring_hist, (uint8_t*)ctx->ring.p + ctx->ring_pos - ctx->ring_hist,
                                               ctx->ring_hist, cudaMemcpyDeviceToDevice, ctx->st));
        CU(cudaStreamSynchronize(ctx->st));
        ctx->ring.release(); ctx->ring = nb; ctx->ring_pos = H;
    } else if (ctx->ring_pos + cap + 16 > ctx->ring.cap) {          // wrap: history moves to the front (never overlaps: ring_pos >= 2H here)
        if (ctx->ring_hist) CU(cudaMemcpyAsync((uint8_t*)ctx->ring.p + H - ctx->ring_hist, (uint8_t*)ctx->ring.p + ctx->ring_pos - ctx->ring_hist,
                                               ctx->ring_hist, cudaMemcpyDeviceToDevice, ctx->st));
        ctx->ring_pos = H;
    }
    return 0;
}

size_t b2lz4_decompress_blocks_host(b2lz4_ctx* ctx, const b2lz4_frame_info* info, const void* src, size_t n,
                                    void* dst, size_t cap, unsigned flags) {
    if (!ctx || !info || (!src && n) || (!dst && cap)) return err_code(E_GENERIC);
    if (info->block_size_id < 4 || info->block_size_id > 7) return err_code(E_maxBlockSize_invalid);
    size_t r = set_device(ctx); if (r) return r;
    Small* ds = ctx->d_small;
    if (flags & B2LZ4_STREAM_FIRST) {
        ctx->ring_pos = 65536; ctx->ring_hist = 0;
        if (info->content_checksum) xxh32_reset_kernel<<<1, 32, 0, ctx->st>>>(&ds->xxh);
    }
    if (n == 0) { CU(cudaStreamSynchronize(ctx->st)); return 0; }
    // a frame of its own around the batch: header without optional fields, the blocks, an EndMark
    FrameDesc d{};
    d.block_size_id = info->block_size_id; d.block_independent = info->block_independent ? 1u : 0u;
    d.block_checksum = info->block_checksum ? 1u : 0u;
    uint8_t hdr[HEADER_MAX];
    const uint32_t hs = header_write(hdr, d);
    const uint32_t zero = 0;
    CU(ctx->stage_in.ensure(hs + n + 4 + 16));
    uint8_t* din = (uint8_t*)ctx->stage_in.p;
    CU(cudaMemcpyAsync(din, hdr, hs, cudaMemcpyHostToDevice, ctx->st));
    CU(cudaMemcpyAsync(din + hs, src, n, cudaMemcpyHostToDevice, ctx->st));
    CU(cudaMemcpyAsync(din + hs + n, &zero, 4, cudaMemcpyHostToDevice, ctx->st));
    uint8_t* dout;
    uint32_t hist = 0;
    const bool linked = !info->block_independent;
    if (linked) {
        r = ring_reserve(ctx, cap); if (r) return r;
        dout = (uint8_t*)ctx->ring.p + ctx->ring_pos; hist = ctx->ring_hist;
    } else {
        CU(ctx->stage_out.ensure(cap + 16));
        dout = (uint8_t*)ctx->stage_out.p;
    }
    const uint32_t keep_index = ctx->prefs.use_index;
    ctx->prefs.use_index = 0;                                       // a batch never carries an index frame
    size_t consumed = 0;
    r = decompress_device_impl(ctx, din, hs + n + 4, dout, cap, &consumed, hist, SCAN_BATCH);
    ctx->prefs.use_index = keep_index;
    if (b2lz4_is_error(r)) return r;
    if (consumed != hs + n + 4) return err_code(E_decompressionFailed);   // an EndMark inside the batch: not whole blocks
    if (info->content_checksum && r) xxh32_stream_kernel<<<1, 256, 0, ctx->st>>>(dout, r, nullptr, nullptr, &ds->xxh);
    if (r) CU(cudaMemcpyAsync(dst, dout, r, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    CU(cudaGetLastError());
    if (linked) {
        ctx->ring_pos += r;
        ctx->ring_hist = (uint32_t)std::min<size_t>(65536, (size_t)ctx->ring_hist + r);
    }
    return r;
}

// -------------------------------------------------------------------------------------- XXH32
size_t b2lz4_xxh32_device(b2lz4_ctx* ctx, const void* d_src, size_t n, uint32_t* out) {
    if (!ctx || !out || (!d_src && n)) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    Small* ds = ctx->d_small;
    xxh32_reset_kernel<<<1, 32, 0, ctx->st>>>(&ds->xxh);
    xxh32_stream_kernel<<<1, 256, 0, ctx->st>>>((const uint8_t*)d_src, n, nullptr, nullptr, &ds->xxh);
    xxh32_digest_kernel<<<1, 32, 0, ctx->st>>>(&ds->xxh, &ds->hash_out, nullptr, nullptr);
    CU(cudaMemcpyAsync(&ctx->h_small->hash_out, &ds->hash_out, 4, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    CU(cudaGetLastError());
    *out = ctx->h_small->hash_out;
    return 0;
}

size_t b2lz4_xxh32_host(b2lz4_ctx* ctx, const void* src, size_t n, uint32_t* out) {
    if (!ctx || !out || (!src && n)) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    CU(ctx->stage_in.ensure(n + 16));
    if (n) CU(cudaMemcpyAsync(ctx->stage_in.p, src, n, cudaMemcpyHostToDevice, ctx->st));
    return b2lz4_xxh32_device(ctx, ctx->stage_in.p, n, out);
}

}  // extern "C"
