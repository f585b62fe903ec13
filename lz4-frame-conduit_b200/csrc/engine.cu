// engine.cu — host engine + C ABI (include/b200lz4.h) of the B200 LZ4 frame codec.
//
// One context owns a CUDA device, two streams (codec + content checksum) and reusable device
// workspaces; every compress / decompress call is a short fixed sequence of kernel launches followed
// by ONE small device->host copy of the result struct (no per-block host round trips).
// Replaces the reference's LZ4F_cctx / LZ4F_dctx life cycle (lz4/lib/lz4frame.c:536-563, 1065-1090)
// and the per-16 KB FFI call pattern of Conduit.hsc:308-314, 585-596 with batch calls.
#include "b2_cuda.h"
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <stddef.h>
#include <new>
#include <algorithm>
#include <vector>
#include <future>
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
    uint32_t walk[8];          // decoder-side block walk: chains published so far, done flag, role ticket, -, walker position (lo, hi), -, -
    DfState dfst;              // dataflow decode: tickets, first failure, byte counts
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
    DevBuf scratch, subs, places, bplaces, blocks, pieces, chains, outs, stored, stage_in, stage_out, enc_tabs, ring, idx_frag, stage_in2, stage_out2;
    DevBuf df_grp, df_heads, df_ent, df_eseq, df_nent, df_usize, df_nseq, df_tkbase, df_tk;   // dataflow decode (decode_df.cu), allocated when a frame first needs them
    uint32_t df_groups = 0, df_ntickets = 0;      // their capacities
    size_t ring_pos = 0; uint32_t ring_hist = 0;     // streaming linked-block decode: next output offset / valid history bytes
    b2lz4_stats stats{};
    char cuda_err[256] = {0};
    int sm_count = 148;
    // host-buffer calls cut the stream into chunks and keep several in flight, each on a lane = a child context with
    // its own stream and workspaces, so that H2D copies, kernels and D2H copies of different chunks overlap
    b2lz4_ctx* lanes[32] = {};
    int n_lanes = 0;
    // multi-GPU host-buffer calls (b2lz4_set_devices): lane k lives on devs[k % n_devs]; chunks of whole blocks go to the
    // lanes round robin, so the blocks of one frame shard across the GPUs by block index and are reassembled in order on
    // the host.  No collective: blocks are independent.  The running content checksum then lives in pinned host memory
    // (every device reads and writes it in stream order).
    int devs[16] = {};
    int n_devs = 0;
    uint8_t* h_frag = nullptr; size_t h_frag_cap = 0;   // pinned: index fragments of the chunks
    // dictionary (LZ4F_createCDict / LZ4F_decompress_usingDict, lz4frame.c:478-521, 1842-1860): its last 64 KB, what a
    // block can reach (lz4_Frame_format.md:226-240).  Used by linked-block frames, on both sides.
    std::vector<uint8_t> dict; uint32_t dict_id = 0;
    std::vector<uint8_t> tail;                    // linked-block streaming compression: the last 64 KB of the batches so far
    // content checksum of host-buffer calls: the bytes lie in host memory anyway, and XXH32 is one serial recurrence per
    // frame (xxhash.c:269-275) that a host core runs ~3 x faster than the GPU's single hashing warp — so a helper thread hashes
    // them while the GPU works.  Device-resident calls keep the GPU kernel (xxh32.cu).
    XxhState hsum{}; bool hsum_on = false;
    DevBuf dict_stage;
};
constexpr size_t HIST_MAX = 65536;

namespace {

#define CU(call)                                                                                   \
    do { cudaError_t e_ = (call);                                                                  \
         if (e_ != cudaSuccess) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "%s: %s (%s:%d)", #call, \
                                           cudaGetErrorString(e_), __FILE__, __LINE__);            \
                                  cudaGetLastError(); return err_code(E_cudaFailure); } } while (0)

void fill_desc(const b2lz4_prefs& p, FrameDesc& d, uint64_t content_size) {
    memset(&d, 0, sizeof d);
    d.block_size_id = p.block_size_id ? p.block_size_id : 7u;
    d.block_independent = p.block_linked ? 0u : 1u;
    d.block_checksum = p.block_checksum ? 1u : 0u;
    d.content_checksum = p.content_checksum ? 1u : 0u;
    // like the reference, the field is only emitted for a non-zero size (lz4frame.c:679, 686-690)
    d.has_content_size = (p.content_size && content_size) ? 1u : 0u;
    d.content_size = content_size;
}

uint32_t sub_size_of(const b2lz4_prefs& p, uint32_t block_size) {
    uint32_t lg = p.sub_block_log2 ? p.sub_block_log2 : 17u;
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
    p->block_size_id = 5; p->sub_block_log2 = 18; p->use_index = 1;
    // write_index stays 0: the stock lz4 CLI cannot skip a skippable frame on a pipe (lz4io.c:1124-1133 uses fseek),
    // so plain frames are the drop-in default and the seekable index is opt-in.  A plain frame is decoded one warp per
    // block by any block-parallel decoder, this one included, so the default block is 256 KB (parsed as one unit): 4096
    // blocks per GiB keep the GPU busy (141 GB/s, against 18 GB/s for plain 4 MB blocks) at a ratio 1 % better than
    // 4 MB blocks cut into 128 KB sub-blocks.  The bench shape (4 MB blocks + index) is set explicitly by its callers.
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

size_t b2lz4_frame_header_write2(void* dst, size_t cap, const b2lz4_prefs* prefs, uint64_t content_size, uint32_t dict_id) {
    b2lz4_prefs p; if (prefs) p = *prefs; else b2lz4_default_prefs(&p);
    if (!prefs_valid(p)) return err_code(E_maxBlockSize_invalid);
    if (cap < (size_t)HEADER_MAX) return err_code(E_dstMaxSize_tooSmall);
    FrameDesc d; fill_desc(p, d, content_size);
    if (dict_id) { d.has_dict_id = 1; d.dict_id = dict_id; }
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
    if ((e = cudaFuncSetAttribute(df_links_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)df_links_smem_bytes())) != cudaSuccess) return fail(e, "cudaFuncSetAttribute");
    if ((e = cudaFuncSetAttribute(decode_df_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)decode_df_smem_bytes(32))) != cudaSuccess) return fail(e, "cudaFuncSetAttribute");
    *out = ctx;
    return 0;
}

size_t b2lz4_destroy(b2lz4_ctx* ctx) {
    if (!ctx) return 0;
    for (auto& l : ctx->lanes) { b2lz4_destroy(l); l = nullptr; }
    cudaSetDevice(ctx->device);
    if (ctx->st) cudaStreamSynchronize(ctx->st);
    if (ctx->st2) cudaStreamSynchronize(ctx->st2);
    for (DevBuf* b : { &ctx->scratch, &ctx->subs, &ctx->places, &ctx->bplaces, &ctx->blocks, &ctx->pieces,
                       &ctx->chains, &ctx->outs, &ctx->stored, &ctx->stage_in, &ctx->stage_out, &ctx->enc_tabs, &ctx->ring, &ctx->idx_frag, &ctx->stage_in2, &ctx->stage_out2,
                       &ctx->dict_stage, &ctx->df_grp, &ctx->df_heads, &ctx->df_ent, &ctx->df_eseq, &ctx->df_nent, &ctx->df_usize, &ctx->df_nseq, &ctx->df_tkbase, &ctx->df_tk }) b->release();
    if (ctx->d_small) cudaFree(ctx->d_small);
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    if (ctx->h_frag) cudaFreeHost(ctx->h_frag);
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

size_t b2lz4_set_devices(b2lz4_ctx* ctx, const int* devices, int n) {
    if (!ctx || n < 0 || n > 16 || (n && !devices)) return err_code(E_GENERIC);
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess) { cudaGetLastError(); return err_code(E_cudaFailure); }
    for (int i = 0; i < n; ++i) if (devices[i] < 0 || devices[i] >= have) return err_code(E_GENERIC);
    for (auto& l : ctx->lanes) { b2lz4_destroy(l); l = nullptr; }    // lanes are re-created on their new devices
    ctx->n_lanes = 0;
    for (int i = 0; i < n; ++i) ctx->devs[i] = devices[i];
    ctx->n_devs = n;
    cudaSetDevice(ctx->device);
    return 0;
}

size_t b2lz4_set_dict(b2lz4_ctx* ctx, const void* dict, size_t n, uint32_t dict_id) {
    if (!ctx || (n && !dict)) return err_code(E_GENERIC);
    const size_t keep = std::min(n, HIST_MAX);                      // only the last 64 KB can ever be referenced (lz4frame.c:505-509)
    ctx->dict.assign((const uint8_t*)dict + (n - keep), (const uint8_t*)dict + n);
    ctx->dict_id = dict_id;
    for (int k = 0; k < ctx->n_lanes; ++k) if (ctx->lanes[k]) { ctx->lanes[k]->dict = ctx->dict; ctx->lanes[k]->dict_id = dict_id; }
    return 0;
}

// tuning builds (-DB2_DF_PROF): cycle counters of the last dataflow decode; zeros otherwise.  Not part of include/b200lz4.h.
size_t b2lz4_debug_df_prof(b2lz4_ctx* ctx, unsigned long long* out8) {
    if (!ctx || !out8) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    CU(cudaStreamSynchronize(ctx->st));
    CU(cudaMemcpy(out8, ctx->d_small->dfst.prof, 12 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
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
enum : unsigned { CL_WHOLE_FRAME = 1u, CL_SUM_RESET = 2u, CL_SUM_UPDATE = 4u, CL_INDEX_FRAGMENT = 8u };
static size_t compress_launch(b2lz4_ctx* ctx, const uint8_t* d_src, size_t n, uint8_t* d_dst, size_t cap,
                              unsigned flags, uint32_t hist = 0)
{
    const bool whole_frame = (flags & CL_WHOLE_FRAME) != 0;
    const b2lz4_prefs& pr = ctx->prefs;
    EncParams P{};
    P.n = n;
    fill_desc(pr, P.desc, n);
    if (ctx->dict_id) { P.desc.has_dict_id = 1; P.desc.dict_id = ctx->dict_id; }
    P.linked = pr.block_linked ? 1u : 0u;
    P.hist = P.linked ? hist : 0u;
    P.block_size = block_size_of(P.desc.block_size_id);
    const uint32_t dom = sub_size_of(pr, P.block_size);             // match domain = what one decoder warp decodes
    P.sub_size = std::min<uint32_t>(dom, ENC_UNIT);                 // encoder unit
    P.dom_units = dom / P.sub_size;
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
    P.write_index = (whole_frame && pr.write_index && !P.linked) ? 1u : 0u;   // a linked frame has no independent entry points

    CU(ctx->scratch.ensure((size_t)P.n_subs * P.scratch_stride));
    CU(ctx->subs.ensure((size_t)P.n_subs * sizeof(SubOut)));
    CU(ctx->places.ensure((size_t)P.n_subs * sizeof(SubPlace)));
    CU(ctx->bplaces.ensure((size_t)P.n_blocks * sizeof(BlockPlace)));
    if (flags & CL_INDEX_FRAGMENT) CU(ctx->idx_frag.ensure((size_t)P.n_blocks * sizeof(IndexBlock) + (size_t)P.n_subs * sizeof(IndexChain) + 64));

    Small* ds = ctx->d_small;
    uint32_t launches = 0;
    const bool sum = P.desc.content_checksum != 0 && (flags & (CL_SUM_RESET | CL_SUM_UPDATE)) != 0;
    if (sum) {                                   // serial XXH32 of the input on the second stream
        CU(cudaEventRecord(ctx->ev_in, ctx->st));
        CU(cudaStreamWaitEvent(ctx->st2, ctx->ev_in, 0));
        if (flags & CL_SUM_RESET) { B2_LAUNCH(xxh32_reset_kernel, 1, 32, 0, ctx->st2, &ds->xxh); ++launches; }
        if (flags & CL_SUM_UPDATE) { B2_LAUNCH(xxh32_stream_kernel, 1, 256, 0, ctx->st2, d_src, n, nullptr, nullptr, &ds->xxh); ++launches; }
        CU(cudaEventRecord(ctx->ev_sum, ctx->st2));
    }
    CU(cudaEventRecord(ctx->ev[0], ctx->st));
    if (P.n_subs) {
        uint32_t grid = std::min<uint32_t>((P.n_subs + ENC_WARPS - 1) / ENC_WARPS, (uint32_t)ctx->sm_count * (uint32_t)ENC_CTAS_PER_SM);
        CU(ctx->enc_tabs.ensure(find_emit_gtab_bytes(grid)));
        B2_LAUNCH(find_emit_kernel, grid, ENC_WARPS * 32, find_emit_smem_bytes(), ctx->st, 
            d_src, P, (uint8_t*)ctx->scratch.p, (SubOut*)ctx->subs.p, &ds->ticket_enc, (uint16_t*)ctx->enc_tabs.p);
        ++launches;
    }
    CU(cudaEventRecord(ctx->ev[1], ctx->st));
    uint8_t* const frag = (flags & CL_INDEX_FRAGMENT) ? (uint8_t*)ctx->idx_frag.p : nullptr;
    if (P.n_blocks) {                             // a warp per block stitches its sub-blocks
        B2_LAUNCH(layout_blocks_kernel, std::min<uint32_t>((P.n_blocks + 7u) / 8u, (uint32_t)ctx->sm_count * 4u), 256, 0, ctx->st,
                  P, (const SubOut*)ctx->subs.p, (SubPlace*)ctx->places.p, (BlockPlace*)ctx->bplaces.p);
        ++launches;
    }
    B2_LAUNCH(layout_kernel, 1, LAYOUT_THREADS, 0, ctx->st, P, (BlockPlace*)ctx->bplaces.p, d_dst, cap, &ds->eres, &ds->ticket_enc, frag);
    ++launches;
    if (P.n_subs) {
        uint32_t grid = std::min<uint32_t>(P.n_subs, (uint32_t)ctx->sm_count * 8u);
        B2_LAUNCH(gather_kernel, grid, 256, 0, ctx->st, d_src, P, (const uint8_t*)ctx->scratch.p, (const SubOut*)ctx->subs.p,
                                                 (const SubPlace*)ctx->places.p, (const BlockPlace*)ctx->bplaces.p,
                                                 d_dst, &ds->eres, frag);
        ++launches;
        if (P.block_checksum) {
            B2_LAUNCH(block_checksums_kernel, P.block_size >= (256u << 10) ? P.n_blocks : std::min<uint32_t>((P.n_blocks + 7u) / 8u, (uint32_t)ctx->sm_count * 8u), 256, 0, ctx->st, P, (const BlockPlace*)ctx->bplaces.p,
                                                                                 d_dst, &ds->eres);
            ++launches;
        }
    }
    if (sum) {
        CU(cudaStreamWaitEvent(ctx->st, ctx->ev_sum, 0));
        if (whole_frame) {
            B2_LAUNCH(xxh32_digest_kernel, 1, 32, 0, ctx->st, &ds->xxh, &ds->hash_out, d_dst, &ds->eres.content_checksum_off);
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
    B2Range range("b2lz4_compress_device");
    size_t r = set_device(ctx); if (r) return r;
    ctx->stats = b2lz4_stats{};
    uint32_t hist = 0;
    if (ctx->prefs.block_linked && !ctx->dict.empty() && n) {
        // the dictionary has to lie in front of the stream: dictionary + stream into a staging buffer (dictionary frames are
        // small-message territory; a device-to-device copy is the simple way)
        hist = (uint32_t)ctx->dict.size();
        CU(ctx->dict_stage.ensure(HIST_MAX + n + 16));
        uint8_t* st = (uint8_t*)ctx->dict_stage.p;
        CU(cudaMemcpyAsync(st + HIST_MAX - hist, ctx->dict.data(), hist, cudaMemcpyHostToDevice, ctx->st));
        CU(cudaMemcpyAsync(st + HIST_MAX, d_src, n, cudaMemcpyDeviceToDevice, ctx->st));
        d_src = st + HIST_MAX;
    }
    r = compress_launch(ctx, (const uint8_t*)d_src, n, (uint8_t*)d_dst, cap, CL_WHOLE_FRAME | CL_SUM_RESET | CL_SUM_UPDATE, hist);
    if (r) return r;
    return compress_finish(ctx, n);
}

// -------------------------------------------------------------------------------------- decompress
// dataflow mode is on unless B2LZ4_DF=0 (then unindexed frames decode one warp per block, linked frames one warp per frame)
static int df_level() { static int v = -1; if (v < 0) { const char* e = getenv("B2LZ4_DF"); v = !e ? 1 : (*e == '0') ? 0 : (*e == '2') ? 2 : 1; } return v; }
static bool df_enabled() { return df_level() != 0; }
static bool df_few() { static int v = -1; if (v < 0) { const char* e = getenv("B2LZ4_DF_FEW"); v = (e && *e == '0') ? 0 : 1; } return v == 1; }
static uint32_t df_mode() { return df_level() == 0 ? 0u : df_level() == 2 ? (SCAN_DF | SCAN_DF_ALL) : (SCAN_DF | (df_few() ? SCAN_DF_FEW : 0u)); }

// warps per CTA of decode_df_kernel: B2LZ4_DF_WARPS, else by the size of the call (a guess at the number of units: the
// host does not know the block size of a device-resident frame)
static uint32_t df_cta_warps(const b2lz4_ctx* ctx, size_t n, size_t cap) {
    static int env = -1;
    if (env < 0) { const char* e = getenv("B2LZ4_DF_WARPS"); env = e ? atoi(e) : 0; }
    if (B2_DF_SMEMWIN) return 32u;                 // the data window leaves room for one CTA per SM only
    if (env == 8 || env == 16 || env == 32) return (uint32_t)env;
    (void)n;
    return cap >= ((size_t)2 * ctx->sm_count << 22) ? 16u : 32u;      // >= two 4 MB blocks per SM: many units
}

static size_t decompress_launch(b2lz4_ctx* ctx, const uint8_t* d_src, size_t n, uint8_t* d_dst, size_t cap,
                                uint32_t mode, uint32_t cap_tables, uint32_t hist)
{
    DfWork W{};
    if (mode & SCAN_DF) {
        // workspaces of the dataflow path: whatever the context already owns (a frame that needs more says so through
        // ERROR_allocation_failed + need_*, and the caller comes again); contexts made for foreign frames start with an estimate
        if (ctx->df_groups) {
            CU(ctx->df_grp.ensure(((size_t)cap_tables + 1) * sizeof(uint32_t)));
            CU(ctx->df_nent.ensure((size_t)cap_tables * sizeof(uint32_t)));
            CU(ctx->df_usize.ensure((size_t)cap_tables * sizeof(uint32_t)));
            CU(ctx->df_nseq.ensure((size_t)cap_tables * sizeof(uint32_t)));
            CU(ctx->df_tkbase.ensure(((size_t)cap_tables + 1) * sizeof(uint32_t)));
            CU(ctx->df_heads.ensure((size_t)ctx->df_groups * DF_G * sizeof(DfHead)));
            CU(ctx->df_ent.ensure((size_t)ctx->df_groups * sizeof(uint2)));
            CU(ctx->df_eseq.ensure((size_t)ctx->df_groups * sizeof(uint32_t)));
            CU(ctx->df_tk.ensure((size_t)ctx->df_ntickets * sizeof(uint2)));
        }
        W.grp_base = (uint32_t*)ctx->df_grp.p; W.heads = (DfHead*)ctx->df_heads.p; W.ent = (uint2*)ctx->df_ent.p;
        W.ent_seq = (uint32_t*)ctx->df_eseq.p;
        W.blk_nent = (uint32_t*)ctx->df_nent.p; W.blk_usize = (uint32_t*)ctx->df_usize.p; W.blk_nseq = (uint32_t*)ctx->df_nseq.p;
        W.tk_base = (uint32_t*)ctx->df_tkbase.p; W.tk = (uint2*)ctx->df_tk.p; W.st = &ctx->d_small->dfst;
        W.cap_groups = ctx->df_groups; W.cap_tickets = ctx->df_ntickets;
    }
    CU(ctx->blocks.ensure((size_t)cap_tables * sizeof(BlockRef)));
    CU(ctx->pieces.ensure((size_t)cap_tables * sizeof(Piece)));
    CU(ctx->chains.ensure((size_t)cap_tables * sizeof(Chain)));
    CU(ctx->outs.ensure((size_t)cap_tables * sizeof(PieceOut)));
    CU(ctx->stored.ensure((size_t)cap_tables * sizeof(uint32_t)));
    Small* ds = ctx->d_small;
    uint32_t launches = 0;
    CU(cudaMemsetAsync(&ds->crc_bad, 0xFF, sizeof(uint32_t), ctx->st));
    CU(cudaEventRecord(ctx->ev[0], ctx->st));
    B2_LAUNCH(frame_scan_kernel, 1, 256, 0, ctx->st, d_src, n, cap, mode, &ds->fi, (BlockRef*)ctx->blocks.p, cap_tables,
                                              (Piece*)ctx->pieces.p, cap_tables, (Chain*)ctx->chains.p, (uint32_t*)ctx->stored.p, ds->walk, W);
    CU(cudaEventRecord(ctx->ev[1], ctx->st));
    B2_LAUNCH(decode_chains_kernel, ctx->sm_count * (64 / DEC_WARPS), DEC_WARPS * 32, 0, ctx->st, 
        d_src, n, d_dst, cap, mode, &ds->fi, (BlockRef*)ctx->blocks.p, (Piece*)ctx->pieces.p, (Chain*)ctx->chains.p, cap_tables,
        (uint32_t*)ctx->stored.p, (PieceOut*)ctx->outs.p, &ds->ticket_dec, ds->walk, hist);
    B2_LAUNCH(copy_stored_kernel, ctx->sm_count * 8, 256, 0, ctx->st, d_src, d_dst, cap, &ds->fi, (const Piece*)ctx->pieces.p,
                                                             (const Chain*)ctx->chains.p, (const uint32_t*)ctx->stored.p);
    // every one of these returns at once unless the scan chose the dataflow path; without workspaces the scan cannot choose
    // it (it answers ERROR_allocation_failed + need_*), so a context that never met such a frame does not pay the launches
    const bool df_launch = (mode & SCAN_DF) && ctx->df_groups;
    if (df_launch) {
        const BlockRef* bl = (const BlockRef*)ctx->blocks.p;
        B2_LAUNCH(df_links_kernel, ctx->sm_count * 3, DF_WARPS * 32, df_links_smem_bytes(), ctx->st, d_src, n, &ds->fi, bl, W);
        B2_LAUNCH(df_stitch_kernel, std::max(1u, std::min((cap_tables + 127u) / 128u, 1024u)), 128, 0, ctx->st, d_src, &ds->fi, bl, W);
        B2_LAUNCH(df_layout_kernel, 1, 256, 0, ctx->st, &ds->fi, (BlockRef*)ctx->blocks.p, cap, W);
        B2_LAUNCH(df_tickets_kernel, ctx->sm_count * 16, 128, 0, ctx->st, d_src, &ds->fi, bl, W);
        // one CTA per unit (independent block / whole linked frame): 32 warps per CTA when there are few units (all the
        // parallelism is inside them), 16 warps and two CTAs per SM when there are many
        const uint32_t dfw = df_cta_warps(ctx, n, cap);
        B2_LAUNCH(decode_df_kernel, ctx->sm_count * (32u / dfw), dfw * 32u, decode_df_smem_bytes(dfw), ctx->st, d_src, n, d_dst, cap, &ds->fi, bl, W, hist);
        launches += 5;
    }
    B2_LAUNCH(verify_block_checksums_kernel, std::max(1u, std::min(cap_tables / 8u, (uint32_t)ctx->sm_count * 8u)), 256, 0, ctx->st, 
        d_src, &ds->fi, (const BlockRef*)ctx->blocks.p, &ds->crc_bad);
    CU(cudaEventRecord(ctx->ev[2], ctx->st));
    B2_LAUNCH(decode_finalize_kernel, 1, 256, 0, ctx->st, &ds->fi, (const Piece*)ctx->pieces.p, (const PieceOut*)ctx->outs.p,
                                                   &ds->crc_bad, &ds->dres, &ds->ticket_dec);
    if (mode & SCAN_DF) { B2_LAUNCH(df_finalize_kernel, 1, 32, 0, ctx->st, &ds->fi, &ds->crc_bad, &ds->dres, W); ++launches; }
    // content checksum over the output (lz4frame.c:1676-1677, 1734-1745); the kernels return at once
    // when the frame carries none
    if (!(mode & SCAN_BATCH)) {                   // a batch of a streamed frame keeps its running checksum (caller updates it)
        B2_LAUNCH(xxh32_reset_kernel, 1, 32, 0, ctx->st, &ds->xxh);
        B2_LAUNCH(xxh32_stream_kernel, 1, 256, 0, ctx->st, d_dst, 0, &ds->dres.total_out, &ds->fi.desc.content_checksum, &ds->xxh);
        B2_LAUNCH(content_check_kernel, 1, 32, 0, ctx->st, &ds->fi, &ds->xxh, &ds->dres);
        launches += 3;
    }
    launches += 5;
    CU(cudaEventRecord(ctx->ev[3], ctx->st));
    CU(cudaMemcpyAsync(&ctx->h_small->fi, &ds->fi, sizeof(FrameIndex) + sizeof(DecodeResult), cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaGetLastError());
    ctx->stats.kernel_launches += launches;
    return 0;
}

static_assert(offsetof(Small, dres) == sizeof(FrameIndex), "fi and dres must be contiguous for the single copy");

static size_t decompress_device_impl(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, size_t* consumed,
                                     uint32_t hist, uint32_t mode0 = 0u, bool launched0 = false);
static uint32_t tables_for(size_t n, size_t cap) {
    return (uint32_t)std::min<uint64_t>(0x7FFFFFFFull, std::max<uint64_t>(4096, (n >> 12) + (cap >> 16) + 64));
}
// a context that decodes foreign frames (no index expected) sizes the dataflow workspaces up front; others wait for a
// frame that asks for them
static void df_presize(b2lz4_ctx* ctx, size_t n, size_t cap) {
    if (!df_enabled()) return;
    // the chain tables take 8 bytes per compressed byte: they are only allocated once a frame has asked for them (the scan
    // reports its needs through ERROR_allocation_failed and the call is repeated), and then kept for the frames that follow
    if (!ctx->df_groups) return;
    const uint64_t g = n / DF_G + tables_for(n, cap) + 16, t = n / (3u * DF_BATCH) + 2ull * tables_for(n, cap) + 16;
    if (g > ctx->df_groups) ctx->df_groups = (uint32_t)std::min<uint64_t>(0x7FFFFFFFull, g + g / 8);
    if (t > ctx->df_ntickets) ctx->df_ntickets = (uint32_t)std::min<uint64_t>(0x7FFFFFFFull, t + t / 8);
}
// queues the first decode attempt on ctx->st without waiting; decompress_device_impl(..., launched0 = true) completes it
static size_t decompress_begin(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, uint32_t hist, uint32_t mode0) {
    ctx->stats = b2lz4_stats{};
    uint32_t mode = mode0 | (ctx->prefs.use_index ? SCAN_USE_INDEX : 0u) | SCAN_OVERLAP_WALK | df_mode();
    df_presize(ctx, n, cap);
    return decompress_launch(ctx, (const uint8_t*)d_src, n, (uint8_t*)d_dst, cap, mode, tables_for(n, cap), hist);
}
size_t b2lz4_decompress_device(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, size_t* consumed) {
    if (ctx && !ctx->dict.empty() && d_dst && cap) {
        // the dictionary has to lie in front of the output: decode into a staging buffer behind it, then copy
        size_t r = set_device(ctx); if (r) return r;
        const uint32_t hist = (uint32_t)ctx->dict.size();
        CU(ctx->dict_stage.ensure(HIST_MAX + cap + 16));
        uint8_t* st = (uint8_t*)ctx->dict_stage.p + HIST_MAX;
        CU(cudaMemcpyAsync(st - hist, ctx->dict.data(), hist, cudaMemcpyHostToDevice, ctx->st));
        r = decompress_device_impl(ctx, d_src, n, st, cap, consumed, hist);
        if (b2lz4_is_error(r)) return r;
        if (r) CU(cudaMemcpyAsync(d_dst, st, r, cudaMemcpyDeviceToDevice, ctx->st));
        CU(cudaStreamSynchronize(ctx->st));
        return r;
    }
    return decompress_device_impl(ctx, d_src, n, d_dst, cap, consumed, 0u);
}
static size_t decompress_device_impl(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, size_t* consumed,
                                     uint32_t hist, uint32_t mode0, bool launched0) {
    if (consumed) *consumed = 0;
    if (!ctx || !d_src || (!d_dst && cap)) return err_code(E_GENERIC);
    if (n < (size_t)HEADER_MIN) return err_code(E_frameHeader_incomplete);
    B2Range range("b2lz4 decompress (device buffers)");
    size_t r = set_device(ctx); if (r) return r;
    if (!launched0) ctx->stats = b2lz4_stats{};
    static uint8_t* dummy = nullptr;
    uint8_t* dst = (uint8_t*)d_dst;
    if (!dst) { if (!dummy) CU(cudaMalloc(&dummy, 256)); dst = dummy; }
    uint32_t mode = mode0 | (ctx->prefs.use_index ? SCAN_USE_INDEX : 0u) | SCAN_OVERLAP_WALK | df_mode();
    uint32_t cap_tables = tables_for(n, cap);
    if (!launched0) df_presize(ctx, n, cap);
    for (int attempt = 0; attempt < 4; ++attempt) {
        if (!(launched0 && attempt == 0)) {
            r = decompress_launch(ctx, (const uint8_t*)d_src, n, dst, cap, mode, cap_tables, hist);
            if (r) return r;
        }
        CU(cudaStreamSynchronize(ctx->st));
        const FrameIndex& fi = ctx->h_small->fi;
        const DecodeResult& dr = ctx->h_small->dres;
        if (fi.err == E_allocation_failed && attempt < 3) {            // tables too small: grow and rerun
            if (fi.df) {
                cap_tables = std::max(cap_tables, fi.need_blocks + 16);
                ctx->df_groups = std::max(ctx->df_groups, fi.need_groups + 16);
                ctx->df_ntickets = std::max(ctx->df_ntickets, fi.need_tickets + 16);
            } else cap_tables = std::max(fi.need_blocks, fi.need_pieces) + 16;
            ++ctx->stats.retries; continue;
        }
        if (fi.indexed && (mode & SCAN_USE_INDEX) && dr.err != OK && attempt < 3) {
            // the index is side information in a skippable frame behind the frame: whatever is wrong with it (entries that
            // pass the structural checks but point at the wrong bytes) must not make a valid frame fail — decode without it;
            // if the frame itself is bad, that run reports the real error
            mode &= ~SCAN_USE_INDEX;
            ++ctx->stats.retries; continue;
        }
        if (fi.err == OK && !fi.layout_known && fi.n_chains > 1 && !(mode & SCAN_SINGLE_CHAIN) &&
            (dr.err != OK || !dr.layout_ok) && attempt < 3) {
            // the "block b starts at b * blockSize" layout did not hold (a middle block decoded short: flushed
            // stream) or an error may stem from it: redo with the exact layout — the dataflow path computes every block's
            // decoded size before it copies a byte (without it: the blocks chained in order by one warp)
            mode = (mode & ~SCAN_USE_INDEX) | (df_enabled() ? (SCAN_DF | SCAN_DF_ALL) : SCAN_SINGLE_CHAIN);

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
// pipeline shape of the host-buffer calls: B2LZ4_LANES chunks in flight (default 4), B2LZ4_PIPE_MB per chunk (default: a
// sixteenth of the stream, 32..256 MiB)
static int pipe_lanes(const b2lz4_ctx* ctx) {
    const char* e = getenv("B2LZ4_LANES"); int v = e ? atoi(e) : 4;
    if (ctx->n_devs > 1 && !e) v = std::max(v, 2 * ctx->n_devs);      // at least two chunks in flight per GPU
    return std::min(32, std::max(1, v));
}
static size_t pipe_chunk(const b2lz4_ctx* ctx, size_t n, size_t block_size) {
    const char* e = getenv("B2LZ4_PIPE_MB");
    const size_t parts = std::max(16, 2 * ctx->n_devs);   // 1 GiB -> 16 x 64 MB: measured best with 16 KB encoder units (23.1 / 23.9 ms vs 25.3 / 23.8 at 128 MB)
    size_t c = e ? (size_t)std::max(1, atoi(e)) << 20 : std::min<size_t>(256u << 20, std::max<size_t>(32u << 20, n / parts));
    c = std::max(block_size, c / block_size * block_size);
    return c;
}
// Chunk plan of a host-buffer call: start offsets in `unit`s, the last entry being the total.  Uniform chunks of `base`
// units (a plan that starts and ends with quarter / half chunks was measured twice and bought 0 - 1 %: 23.2 vs 22.9 ms per GiB).
static std::vector<size_t> plan_chunks(const b2lz4_ctx* ctx, size_t total, size_t base) {
    (void)ctx;
    std::vector<size_t> v{0};
    base = std::max<size_t>(1, base);
    for (size_t at = 0; at < total;) v.push_back(at = std::min(total, at + base));
    return v;
}
static bool multi_dev(const b2lz4_ctx* ctx) { return ctx->n_devs > 1; }
static int lane_device(const b2lz4_ctx* ctx, int k) { return ctx->n_devs > 0 ? ctx->devs[k % ctx->n_devs] : ctx->device; }
static size_t ensure_lanes(b2lz4_ctx* ctx, int want) {
    for (int k = ctx->n_lanes; k < want; ++k) {
        size_t r = b2lz4_create(&ctx->lanes[k], lane_device(ctx, k), &ctx->prefs);
        if (b2lz4_is_error(r)) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "lane context creation failed"); return r; }
        ctx->n_lanes = k + 1;
    }
    for (int k = 0; k < ctx->n_lanes; ++k) ctx->lanes[k]->prefs = ctx->prefs;
    cudaSetDevice(ctx->device);
    return 0;
}
extern "C++" {
// ---- XXH32 streaming state on the host (XXH32_reset / XXH32_update / XXH32_digest, xxhash.c:437-554), same XxhState as the kernels
static void hsum_reset(XxhState& s) { s = XxhState{}; s.v[0] = XP1 + XP2; s.v[1] = XP2; s.v[2] = 0; s.v[3] = 0u - XP1; }
static void hsum_update(XxhState& s, const uint8_t* p, size_t n) {
    s.total_len += (uint32_t)n;
    s.large |= (uint32_t)((n >= 16) | (s.total_len >= 16));
    if (s.memsize + n < 16) { memcpy((uint8_t*)s.mem + s.memsize, p, n); s.memsize += (uint32_t)n; return; }
    if (s.memsize) {
        const uint32_t take = 16 - s.memsize;
        memcpy((uint8_t*)s.mem + s.memsize, p, take);
        for (int k = 0; k < 4; ++k) s.v[k] = xxh32_round(s.v[k], s.mem[k]);
        p += take; n -= take; s.memsize = 0;
    }
    uint32_t v1 = s.v[0], v2 = s.v[1], v3 = s.v[2], v4 = s.v[3];
    for (; n >= 16; p += 16, n -= 16) {
        uint32_t w[4]; memcpy(w, p, 16);
        v1 = xxh32_round(v1, w[0]); v2 = xxh32_round(v2, w[1]); v3 = xxh32_round(v3, w[2]); v4 = xxh32_round(v4, w[3]);
    }
    s.v[0] = v1; s.v[1] = v2; s.v[2] = v3; s.v[3] = v4;
    if (n) { memcpy(s.mem, p, n); s.memsize = (uint32_t)n; }
}
static uint32_t hsum_digest(const XxhState& s) {
    uint32_t h = s.large ? xxh32_merge(s.v[0], s.v[1], s.v[2], s.v[3]) : s.v[2] + XP5;
    h += s.total_len;
    return xxh32_tail(h, (const uint8_t*)s.mem, s.memsize);
}
// runs f on a helper thread (or right here when no thread is to be had)
template <typename F> static std::future<void> run_aside(F f) {
    try { return std::async(std::launch::async, f); }
    catch (...) { f(); std::promise<void> p; p.set_value(); return p.get_future(); }
}
}  // extern "C++"

#define LANE_CU(L, call)                                                                           \
    do { cudaError_t e_ = (call);                                                                  \
         if (e_ != cudaSuccess) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "%s: %s (%s:%d)", #call, \
                                           cudaGetErrorString(e_), __FILE__, __LINE__);            \
                                  cudaGetLastError(); return err_code(E_cudaFailure); } } while (0)

// One frame from a host buffer.  The stream is cut into chunks of whole blocks; chunk i goes to lane i mod L:
// H2D copy -> match finder / layout / gather on the lane's stream, no host wait; chunks are completed in order
// (their compressed size decides where the next one lands in dst) and copied back while later chunks still run.
// mode: HOSTC_FRAME = a whole frame (header, blocks, EndMark, checksum, index); HOSTC_BLOCKS(_FIRST) = only the blocks of
// one batch of a streamed frame (b2lz4_compress_blocks_host), the running content checksum continuing across calls
enum : unsigned { HOSTC_FRAME = 0u, HOSTC_BLOCKS = 1u, HOSTC_BLOCKS_FIRST = 2u };
static size_t compress_host_impl(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, unsigned hmode);
size_t b2lz4_compress_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap) {
    return compress_host_impl(ctx, src, n, dst, cap, HOSTC_FRAME);
}
static size_t compress_host_impl(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, unsigned hmode) {
    if (!ctx || (!src && n) || !dst) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    ctx->stats = b2lz4_stats{};
    const bool whole = hmode == HOSTC_FRAME;
    b2lz4_prefs pr = ctx->prefs;
    if (!whole) pr.write_index = 0;
    const size_t bs = block_size_of(pr.block_size_id ? pr.block_size_id : 7u);
    const size_t chunk = pipe_chunk(ctx, n, bs);
    const std::vector<size_t> plan = plan_chunks(ctx, (n + bs - 1) / bs, chunk / bs);      // in blocks
    const size_t nchunks = plan.size() - 1;
    const int L = (int)std::min<size_t>((size_t)pipe_lanes(ctx), std::max<size_t>(1, nchunks));
    r = ensure_lanes(ctx, L); if (r) return r;
    uint8_t* out = (uint8_t*)dst;
    if (pr.block_linked) pr.write_index = 0;
    size_t o = whole ? b2lz4_frame_header_write2(out, cap, &pr, n, ctx->dict_id) : 0;
    if (b2lz4_is_error(o)) return o;
    const bool sum = pr.content_checksum != 0, want_index = pr.write_index != 0;
    const bool linked = pr.block_linked != 0;
    // linked blocks: every chunk goes to its lane with the 64 KB of the stream in front of it (the input is all here, in host
    // memory); in front of the first chunk lies the dictionary (first batch of a frame) or the tail of the batches before
    const std::vector<uint8_t>& front = (hmode == HOSTC_BLOCKS) ? ctx->tail : ctx->dict;
    std::future<void> hashing;                             // content checksum: the input is hashed on a helper thread meanwhile
    if (sum) {
        if (hmode != HOSTC_BLOCKS) hsum_reset(ctx->hsum);
        ctx->hsum_on = true;
        XxhState* st = &ctx->hsum; const uint8_t* p8 = (const uint8_t*)src;
        hashing = run_aside([st, p8, n]() { hsum_update(*st, p8, n); });
    }
    struct Joiner { std::future<void>& f; ~Joiner() { if (f.valid()) f.wait(); } } joiner{hashing};   // never outlives the call: src is the caller's
    std::vector<uint8_t> ix_blocks, ix_chains;             // index entries gathered from the chunks
    struct Frag { size_t off; uint32_t nb, nc, chain_base; };
    std::vector<Frag> frags;
    size_t frag_used = 0;
    if (want_index) {                                      // pinned landing area for the chunks' index fragments
        const size_t ss = sub_size_of(pr, (uint32_t)bs);
        const size_t need = (n / bs + nchunks + 2) * sizeof(IndexBlock) + (n / ss + 2 * nchunks + n / bs + 2) * sizeof(IndexChain);
        if (need > ctx->h_frag_cap) {
            if (ctx->h_frag) cudaFreeHost(ctx->h_frag);
            ctx->h_frag = nullptr; ctx->h_frag_cap = 0;
            CU(cudaMallocHost(&ctx->h_frag, need + need / 4));
            ctx->h_frag_cap = need + need / 4;
        }
    }
    uint32_t chains_total = 0, blocks_total = 0, launches = 0;
    float ms_main = 0, ms_total = 0;
    cudaEvent_t t0 = ctx->ev[4], t1 = ctx->ev[5];
    CU(cudaEventRecord(t0, ctx->st));

    auto submit = [&](size_t i) -> size_t {
        B2Range range("b2lz4 compress: chunk H2D + kernels");
        b2lz4_ctx* ln = ctx->lanes[i % L];
        const size_t off = plan[i] * bs, len = std::min(plan[i + 1] * bs, n) - off;
        LANE_CU(ln, cudaSetDevice(ln->device));
        const size_t pad = linked ? HIST_MAX : 0;
        LANE_CU(ln, ln->stage_in.ensure(pad + len + 16));
        LANE_CU(ln, ln->stage_out.ensure(b2lz4_compress_bound(len, &pr) + 16));
        uint8_t* din = (uint8_t*)ln->stage_in.p + pad;
        uint32_t hist = 0;
        if (linked) {
            const size_t own = std::min(off, HIST_MAX);                           // bytes of THIS call's input in front of the chunk
            hist = (uint32_t)own;
            if (own < HIST_MAX && !front.empty()) {                               // ... and, in front of those, the dictionary / earlier batches
                const size_t more = std::min(front.size(), HIST_MAX - own);
                LANE_CU(ln, cudaMemcpyAsync(din - own - more, front.data() + (front.size() - more), more, cudaMemcpyHostToDevice, ln->st));
                hist += (uint32_t)more;
            }
            LANE_CU(ln, cudaMemcpyAsync(din - own, (const uint8_t*)src + off - own, own + len, cudaMemcpyHostToDevice, ln->st));
        } else
        LANE_CU(ln, cudaMemcpyAsync(din, (const uint8_t*)src + off, len, cudaMemcpyHostToDevice, ln->st));
        ln->stats = b2lz4_stats{};
        ln->dict_id = ctx->dict_id;
        size_t e = compress_launch(ln, din, len, (uint8_t*)ln->stage_out.p, ln->stage_out.cap,
                                   want_index ? CL_INDEX_FRAGMENT : 0u, hist);
        if (e) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "%s", ln->cuda_err); return e; }
        return 0;
    };
    auto complete = [&](size_t i) -> size_t {
        B2Range range("b2lz4 compress: chunk wait + D2H");
        b2lz4_ctx* ln = ctx->lanes[i % L];
        const size_t off = plan[i] * bs, len = std::min(plan[i + 1] * bs, n) - off;
        LANE_CU(ln, cudaSetDevice(ln->device));
        size_t c = compress_finish(ln, len);                                    // waits for the lane
        if (b2lz4_is_error(c)) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "%s", ln->cuda_err); return c; }
        if (o + c + 8 > cap) return err_code(E_dstMaxSize_tooSmall);
        LANE_CU(ln, cudaMemcpyAsync(out + o, ln->stage_out.p, c, cudaMemcpyDeviceToHost, ln->st));
        const EncodeResult& er = ln->h_small->eres;
        if (want_index) {                                                       // fragment: [nb x IndexBlock][n_chains x IndexChain], chunk-relative
            const size_t fb = er.n_blocks * sizeof(IndexBlock) + er.n_chains * sizeof(IndexChain);
            if (frag_used + fb > ctx->h_frag_cap) return err_code(E_GENERIC);
            LANE_CU(ln, cudaMemcpyAsync(ctx->h_frag + frag_used, ln->idx_frag.p, fb, cudaMemcpyDeviceToHost, ln->st));
            frags.push_back({frag_used, er.n_blocks, er.n_chains, chains_total});
            frag_used += fb;
        }
        chains_total += er.n_chains; blocks_total += er.n_blocks;
        launches += ln->stats.kernel_launches; ms_main += ln->stats.ms_main; ms_total += ln->stats.ms_total;
        o += c;
        return 0;
    };
    for (size_t i = 0; i < nchunks; ++i) {
        if (i >= (size_t)L) { r = complete(i - L); if (r) return r; }
        r = submit(i); if (r) return r;
    }
    for (size_t i = nchunks > (size_t)L ? nchunks - L : 0; i < nchunks; ++i) { r = complete(i); if (r) return r; }
    for (int k = 0; k < L; ++k) LANE_CU(ctx->lanes[k], cudaStreamSynchronize(ctx->lanes[k]->st));
    if (hashing.valid()) hashing.wait();
    CU(cudaSetDevice(ctx->device));
    if (!whole && linked) {                                                     // what the next batch's first blocks may reach back into
        std::vector<uint8_t> t;
        const size_t own = std::min(n, HIST_MAX);
        if (own < HIST_MAX && !front.empty()) { const size_t more = std::min(front.size(), HIST_MAX - own); t.assign(front.end() - (long)more, front.end()); }
        t.insert(t.end(), (const uint8_t*)src + (n - own), (const uint8_t*)src + n);
        ctx->tail.swap(t);
    }
    if (!whole) {
        ctx->stats.n_blocks = blocks_total; ctx->stats.n_chains = chains_total; ctx->stats.kernel_launches = launches;
        ctx->stats.ms_main = ms_main; ctx->stats.ms_total = ms_total; ctx->stats.bytes_in = n; ctx->stats.bytes_out = o;
        return o;
    }

    // EndMark, content checksum, index frame: a few bytes, written by the host (lz4frame.c:998-1007)
    if (o + 4 + (sum ? 4u : 0u) > cap) return err_code(E_dstMaxSize_tooSmall);
    st_le32(out + o, 0u); o += 4;
    if (sum) { uint32_t h = 0; r = b2lz4_content_checksum(ctx, &h); if (r) return r; st_le32(out + o, h); o += 4; ++launches; }
    if (want_index) {
        for (const Frag& f : frags) {                                           // rebase chain numbers, concatenate
            IndexBlock* ib = (IndexBlock*)(ctx->h_frag + f.off);
            for (uint32_t k = 0; k < f.nb; ++k) ib[k].first_chain += f.chain_base;
            const uint8_t* p = ctx->h_frag + f.off;
            ix_blocks.insert(ix_blocks.end(), p, p + f.nb * sizeof(IndexBlock));
            ix_chains.insert(ix_chains.end(), p + f.nb * sizeof(IndexBlock), p + f.nb * sizeof(IndexBlock) + f.nc * sizeof(IndexChain));
        }
        const uint32_t pb = (uint32_t)(sizeof(IndexHead) + ix_blocks.size() + ix_chains.size());
        if (o + 16 + pb > cap) return err_code(E_dstMaxSize_tooSmall);
        uint8_t* x = out + o;
        st_le32(x, IDX_SKIP_MAGIC); st_le32(x + 4, pb + 8u); x += 8;
        IndexHead h{}; h.magic = IDX_MAGIC; h.version = 1; h.block_size = (uint32_t)bs; h.n_blocks = blocks_total;
        h.n_chains = chains_total; h.reserved = 0; h.content_size = n;
        memcpy(x, &h, sizeof h); x += sizeof h;
        if (!ix_blocks.empty()) memcpy(x, ix_blocks.data(), ix_blocks.size());
        x += ix_blocks.size();
        if (!ix_chains.empty()) memcpy(x, ix_chains.data(), ix_chains.size());
        x += ix_chains.size();
        st_le32(x, pb); st_le32(x + 4, IDX_MAGIC);
        o += 16 + pb;
    }
    CU(cudaEventRecord(t1, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    ctx->stats.n_blocks = blocks_total; ctx->stats.n_chains = chains_total; ctx->stats.indexed = want_index ? 1u : 0u;
    ctx->stats.kernel_launches = launches; ctx->stats.ms_main = ms_main; ctx->stats.ms_total = ms_total;
    ctx->stats.bytes_in = n; ctx->stats.bytes_out = o;
    return o;
}

static size_t decompress_host_whole(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, size_t* consumed) {
    if (consumed) *consumed = 0;
    if (!ctx || !src || (!dst && cap)) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    CU(ctx->stage_in.ensure(n + 16));
    CU(ctx->stage_out.ensure(HIST_MAX + cap + 16));
    cudaEvent_t e0 = ctx->ev[4], e1 = ctx->ev[5], e2 = ctx->ev[6], e3 = ctx->ev[7];
    CU(cudaEventRecord(e0, ctx->st));
    CU(cudaMemcpyAsync(ctx->stage_in.p, src, n, cudaMemcpyHostToDevice, ctx->st));
    // a dictionary (LZ4F_decompress_usingDict, lz4frame.c:1842-1860) is history in front of the first block (linked frames)
    uint8_t* dout = (uint8_t*)ctx->stage_out.p + HIST_MAX;
    const uint32_t hist = (uint32_t)ctx->dict.size();
    if (hist) CU(cudaMemcpyAsync(dout - hist, ctx->dict.data(), hist, cudaMemcpyHostToDevice, ctx->st));
    CU(cudaEventRecord(e1, ctx->st));
    r = decompress_device_impl(ctx, ctx->stage_in.p, n, dout, cap, consumed, hist);
    if (b2lz4_is_error(r)) return r;
    CU(cudaEventRecord(e2, ctx->st));
    if (r) CU(cudaMemcpyAsync(dst, dout, r, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaEventRecord(e3, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    cudaEventElapsedTime(&ctx->stats.ms_h2d, e0, e1);
    cudaEventElapsedTime(&ctx->stats.ms_d2h, e2, e3);
    return r;
}

// what the host learns from walking the block header words of a frame that lies in host memory
struct HostFrame {
    FrameDesc desc{};
    uint32_t block_size = 0;
    struct Blk { size_t off; uint32_t word; };       // offset of the header word, the word itself
    std::vector<Blk> blocks;
    size_t end_blocks = 0;                           // offset of the EndMark
    size_t consumed = 0;                             // through EndMark (+ content checksum) (+ index frame when used)
    uint32_t stored_sum = 0;
    const uint8_t* ix_blocks = nullptr; const uint8_t* ix_chains = nullptr;   // validated index payload, or null
    uint32_t ix_nchains = 0;
};

// false = leave it to the device scan (which reports the precise error)
static bool host_scan(const uint8_t* src, size_t n, bool use_index, HostFrame& f) {
    if (header_parse(src, n, f.desc) != OK) return false;
    f.block_size = block_size_of(f.desc.block_size_id);
    const uint32_t bc = f.desc.block_checksum ? 4u : 0u;
    size_t ip = f.desc.header_size;
    for (;;) {
        if (n - ip < 4) return false;
        const uint32_t w = ld_le32(src + ip);
        const uint32_t cs = w & 0x7FFFFFFFu;
        if (cs == 0) break;
        if (cs > f.block_size || n - ip - 4 < (size_t)cs + bc) return false;
        f.blocks.push_back({ip, w});
        ip += 4 + (size_t)cs + bc;
    }
    f.end_blocks = ip; ip += 4;
    if (f.desc.content_checksum) { if (n - ip < 4) return false; f.stored_sum = ld_le32(src + ip); ip += 4; }
    f.consumed = ip;
    if (use_index && f.desc.block_independent && n - ip >= 16 + sizeof(IndexHead) && ld_le32(src + ip) == IDX_SKIP_MAGIC) {
        const uint32_t sz = ld_le32(src + ip + 4);
        if (sz >= 8 + sizeof(IndexHead) && n - ip - 8 >= sz) {
            const uint8_t* pl = src + ip + 8; const uint32_t pb = sz - 8;
            IndexHead h; memcpy(&h, pl, sizeof h);
            const uint64_t want = sizeof(IndexHead) + (uint64_t)h.n_blocks * sizeof(IndexBlock) + (uint64_t)h.n_chains * sizeof(IndexChain);
            if (h.magic == IDX_MAGIC && h.version == 1 && want == pb && h.block_size == f.block_size &&
                h.n_blocks == f.blocks.size() && ld_le32(pl + pb) == pb && ld_le32(pl + pb + 4) == IDX_MAGIC) {
                f.ix_blocks = pl + sizeof(IndexHead); f.ix_chains = f.ix_blocks + (size_t)h.n_blocks * sizeof(IndexBlock);
                f.ix_nchains = h.n_chains;
                f.consumed = ip + 8 + sz;
            }
        }
    }
    return true;
}

// Upper bound of the decoded size of the first frame in src[0, n): its blocks are counted by walking the 4-byte header
// words on the host (what LZ4F_decompress learns block by block, lz4frame.c:1470-1500), every block taken as full.  A
// truncated frame is bounded by the blocks that are there (decoding it reports ERROR_noEndMark).  Host only, no GPU.
size_t b2lz4_decompress_bound(const void* src_, size_t n) {
    if (!src_) return err_code(E_GENERIC);
    const uint8_t* src = (const uint8_t*)src_;
    FrameDesc d{};
    const Err e = header_parse(src, n, d);
    if (e != OK) return err_code(e);
    const size_t bs = block_size_of(d.block_size_id);
    const uint32_t bc = d.block_checksum ? 4u : 0u;
    size_t ip = d.header_size, nb = 0;
    while (n - ip >= 4) {
        const uint32_t cs = ld_le32(src + ip) & 0x7FFFFFFFu;
        if (cs == 0) break;
        ++nb;
        if (cs > bs || n - ip - 4 < (size_t)cs + bc) break;       // malformed / truncated: the decoder reports it
        ip += 4 + (size_t)cs + bc;
    }
    return nb * bs;
}

static size_t decompress_host_groups(b2lz4_ctx* ctx, const uint8_t* src, const HostFrame& f, size_t per_group,
                                     void* dst, size_t cap, unsigned hmode);
// One frame from a host buffer.  Independent-block frames are cut into groups of whole blocks on the host (a serial walk
// over the 4-byte header words); group g goes to lane g mod L as a small frame of its own — header, the blocks, an EndMark
// and the group's slice of the index — so that its H2D copy, decode and D2H copy overlap with the other lanes'.
size_t b2lz4_decompress_host(b2lz4_ctx* ctx, const void* src_, size_t n, void* dst, size_t cap, size_t* consumed) {
    if (consumed) *consumed = 0;
    if (!ctx || !src_ || (!dst && cap)) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    const uint8_t* src = (const uint8_t*)src_;
    HostFrame f;
    const bool ok = n >= (size_t)HEADER_MIN && host_scan(src, n, ctx->prefs.use_index != 0, f);
    const size_t per_group = ok && f.block_size ? std::max<size_t>(1, pipe_chunk(ctx, (size_t)f.blocks.size() * f.block_size, f.block_size) / f.block_size) : 0;
    if (ok && f.desc.block_independent && !ctx->dict.empty() && !f.blocks.empty()) {
        // independent blocks + dictionary: block by block (b2lz4_decompress_blocks_host), then the frame's own checks
        b2lz4_frame_info info{};
        info.block_size_id = f.desc.block_size_id; info.block_independent = 1; info.block_checksum = f.desc.block_checksum;
        info.content_checksum = f.desc.content_checksum; info.block_size = f.block_size;
        const size_t b0 = f.blocks[0].off;
        r = b2lz4_decompress_blocks_host(ctx, &info, src + b0, f.end_blocks - b0, dst, cap, B2LZ4_STREAM_FIRST);
        if (b2lz4_is_error(r)) return r;
        if (f.desc.content_checksum) { uint32_t h = 0; size_t e = b2lz4_content_checksum(ctx, &h); if (e) return e; if (h != f.stored_sum) return err_code(E_contentChecksum_invalid); }
        if (f.desc.has_content_size && f.desc.content_size != r) return err_code(E_frameSize_wrong);
        if (consumed) *consumed = f.consumed;
        return r;
    }
    if (!ok || !f.desc.block_independent || f.blocks.size() <= per_group)
        return decompress_host_whole(ctx, src_, n, dst, cap, consumed);           // linked frames, small frames, anything odd
    r = decompress_host_groups(ctx, src, f, per_group, dst, cap, HOSTC_FRAME);
    if (b2lz4_is_error(r) && f.ix_blocks) {                                          // a bad index must not fail a valid frame: once more without it
        HostFrame g;
        if (host_scan(src, n, false, g)) { g.consumed = f.consumed; r = decompress_host_groups(ctx, src, g, per_group, dst, cap, HOSTC_FRAME); }
    }
    if (!b2lz4_is_error(r) && consumed) *consumed = f.consumed;
    return r;
}
static size_t decompress_host_groups(b2lz4_ctx* ctx, const uint8_t* src, const HostFrame& f, size_t per_group,
                                     void* dst, size_t cap, unsigned hmode) {
    size_t r = 0;
    const bool whole = hmode == HOSTC_FRAME;
    ctx->stats = b2lz4_stats{};
    const size_t nb = f.blocks.size();
    const std::vector<size_t> plan = plan_chunks(ctx, nb, per_group);
    const size_t ngroups = plan.size() - 1;
    const int L = (int)std::min<size_t>((size_t)pipe_lanes(ctx), ngroups);
    r = ensure_lanes(ctx, L); if (r) return r;
    const uint32_t bc = f.desc.block_checksum ? 4u : 0u;
    const bool sum = f.desc.content_checksum != 0;
    // content checksum: every group's output is hashed on a helper thread, in order, once its copy to the host is done
    std::future<void> hashing;
    struct Joiner { std::future<void>& f; ~Joiner() { if (f.valid()) f.wait(); } } joiner{hashing};       // dst is the caller's
    if (sum) { if (hmode != HOSTC_BLOCKS) hsum_reset(ctx->hsum); ctx->hsum_on = true; }
    FrameDesc gd{}; gd.block_size_id = f.desc.block_size_id; gd.block_independent = 1; gd.block_checksum = f.desc.block_checksum;
    uint8_t ghdr[HEADER_MAX]; const uint32_t ghs = header_write(ghdr, gd);
    uint8_t* out = (uint8_t*)dst;
    size_t o = 0;
    uint32_t launches = 0, chains_total = 0; float ms_main = 0, ms_total = 0;
    struct Group { size_t src_len, frame_len, out_cap; };
    std::vector<Group> groups(ngroups);

    auto submit = [&](size_t g) -> size_t {
        B2Range range("b2lz4 decompress: group H2D + kernels");
        b2lz4_ctx* ln = ctx->lanes[g % L];
        const size_t b0 = plan[g], b1 = plan[g + 1];
        const size_t s0 = f.blocks[b0].off, s1 = b1 < nb ? f.blocks[b1].off : f.end_blocks;
        (void)bc;
        // the group's slice of the index, rebased: [skip magic, size][IndexHead][blocks][chains][size, magic]
        std::vector<uint8_t> ix;
        if (f.ix_blocks) {
            const IndexBlock* ib = (const IndexBlock*)f.ix_blocks;
            IndexBlock first; memcpy(&first, ib + b0, sizeof first);
            IndexBlock last; memcpy(&last, ib + (b1 - 1), sizeof last);
            const uint32_t c0 = first.first_chain, c1 = last.first_chain + last.n_chain;
            if (c1 < c0 || c1 > f.ix_nchains) return err_code(E_GENERIC);
            const uint32_t pb = (uint32_t)(sizeof(IndexHead) + (b1 - b0) * sizeof(IndexBlock) + (size_t)(c1 - c0) * sizeof(IndexChain));
            ix.resize(16 + pb);
            uint8_t* x = ix.data();
            st_le32(x, IDX_SKIP_MAGIC); st_le32(x + 4, pb + 8u); x += 8;
            IndexHead h{}; h.magic = IDX_MAGIC; h.version = 1; h.block_size = f.block_size; h.n_blocks = (uint32_t)(b1 - b0); h.n_chains = c1 - c0;
            uint64_t usum = 0;
            IndexBlock* ob = (IndexBlock*)(x + sizeof h);
            for (size_t k = b0; k < b1; ++k) { IndexBlock e; memcpy(&e, ib + k, sizeof e); usum += e.usize; e.first_chain -= c0; memcpy(ob + (k - b0), &e, sizeof e); }
            h.content_size = usum;
            memcpy(x, &h, sizeof h);
            memcpy(x + sizeof h + (b1 - b0) * sizeof(IndexBlock), f.ix_chains + (size_t)c0 * sizeof(IndexChain), (size_t)(c1 - c0) * sizeof(IndexChain));
            st_le32(x + pb, pb); st_le32(x + pb + 4, IDX_MAGIC);
        }
        Group& G = groups[g];
        G.src_len = s1 - s0; G.frame_len = ghs + G.src_len + 4 + ix.size(); G.out_cap = (b1 - b0) * (size_t)f.block_size;
        LANE_CU(ln, cudaSetDevice(ln->device));
        LANE_CU(ln, ln->stage_in.ensure(G.frame_len + 16));
        LANE_CU(ln, ln->stage_out.ensure(G.out_cap + 16));
        uint8_t* din = (uint8_t*)ln->stage_in.p;
        const uint32_t zero = 0;
        LANE_CU(ln, cudaMemcpyAsync(din, ghdr, ghs, cudaMemcpyHostToDevice, ln->st));
        LANE_CU(ln, cudaMemcpyAsync(din + ghs, src + s0, G.src_len, cudaMemcpyHostToDevice, ln->st));
        LANE_CU(ln, cudaMemcpyAsync(din + ghs + G.src_len, &zero, 4, cudaMemcpyHostToDevice, ln->st));
        if (!ix.empty()) LANE_CU(ln, cudaMemcpyAsync(din + ghs + G.src_len + 4, ix.data(), ix.size(), cudaMemcpyHostToDevice, ln->st));
        ln->prefs.use_index = f.ix_blocks ? 1u : 0u;
        size_t e = decompress_begin(ln, din, G.frame_len, ln->stage_out.p, G.out_cap, 0u, SCAN_BATCH);
        if (e) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "%s", ln->cuda_err); return e; }
        return 0;
    };
    auto complete = [&](size_t g) -> size_t {
        B2Range range("b2lz4 decompress: group wait + D2H");
        b2lz4_ctx* ln = ctx->lanes[g % L];
        const Group& G = groups[g];
        size_t used = 0;
        LANE_CU(ln, cudaSetDevice(ln->device));
        size_t c = decompress_device_impl(ln, ln->stage_in.p, G.frame_len, ln->stage_out.p, G.out_cap, &used, 0u, SCAN_BATCH, true);
        if (b2lz4_is_error(c)) { snprintf(ctx->cuda_err, sizeof ctx->cuda_err, "%s", ln->cuda_err); return c; }
        if (used != G.frame_len) return err_code(E_decompressionFailed);
        if (o + c > cap) return err_code(E_dstMaxSize_tooSmall);
        if (c) LANE_CU(ln, cudaMemcpyAsync(out + o, ln->stage_out.p, c, cudaMemcpyDeviceToHost, ln->st));
        if (sum && c) {
            LANE_CU(ln, cudaEventRecord(ln->ev_sum, ln->st));                       // the copy above has landed when this event has
            XxhState* st = &ctx->hsum; const uint8_t* p8 = out + o; const size_t len = c; cudaEvent_t ev = ln->ev_sum;
            std::shared_future<void> prev = hashing.valid() ? hashing.share() : std::shared_future<void>();
            hashing = run_aside([st, p8, len, ev, prev]() { if (prev.valid()) prev.wait(); cudaEventSynchronize(ev); hsum_update(*st, p8, len); });
        }
        launches += ln->stats.kernel_launches; ms_main += ln->stats.ms_main; ms_total += ln->stats.ms_total;
        chains_total += ln->stats.n_chains;
        o += c;
        return 0;
    };
    for (size_t g = 0; g < ngroups; ++g) {
        if (g >= (size_t)L) { r = complete(g - L); if (r) return r; }
        r = submit(g); if (r) return r;
    }
    for (size_t g = ngroups > (size_t)L ? ngroups - L : 0; g < ngroups; ++g) { r = complete(g); if (r) return r; }
    for (int k = 0; k < L; ++k) LANE_CU(ctx->lanes[k], cudaStreamSynchronize(ctx->lanes[k]->st));
    if (hashing.valid()) hashing.wait();
    CU(cudaSetDevice(ctx->device));
    if (sum && whole) {
        uint32_t h = 0; r = b2lz4_content_checksum(ctx, &h); if (r) return r;
        if (h != f.stored_sum) return err_code(E_contentChecksum_invalid);      // lz4frame.c:1734-1745
    }
    if (whole && f.desc.has_content_size && f.desc.content_size != o) return err_code(E_frameSize_wrong);   // lz4frame.c:1715-1716
    ctx->stats.n_blocks = (uint32_t)nb; ctx->stats.n_chains = chains_total; ctx->stats.indexed = f.ix_blocks ? 1u : 0u;
    ctx->stats.kernel_launches = launches; ctx->stats.ms_main = ms_main; ctx->stats.ms_total = ms_total;
    ctx->stats.bytes_in = f.end_blocks; ctx->stats.bytes_out = o;
    return o;
}

// -------------------------------------------------------------------------------------- streaming (batches of blocks)
size_t b2lz4_compress_blocks_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, unsigned flags) {
    if (!ctx || (!src && n) || !dst) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    ctx->stats = b2lz4_stats{};
    if (n == 0) {                                                   // nothing to emit, but the checksum may have to restart
        if ((flags & B2LZ4_STREAM_FIRST) && ctx->prefs.content_checksum) { hsum_reset(ctx->hsum); ctx->hsum_on = true; }
        return 0;
    }
    // the batch goes through the same chunk pipeline as a whole frame (lanes, and every device of b2lz4_set_devices)
    return compress_host_impl(ctx, src, n, dst, cap, (flags & B2LZ4_STREAM_FIRST) ? HOSTC_BLOCKS_FIRST : HOSTC_BLOCKS);
}

size_t b2lz4_content_checksum(b2lz4_ctx* ctx, uint32_t* out) {
    if (!ctx || !out) return err_code(E_GENERIC);
    size_t r = set_device(ctx); if (r) return r;
    if (ctx->hsum_on) { *out = hsum_digest(ctx->hsum); return 0; }   // host-buffer stream: hashed on the host (xxhash.c:529-554)
    Small* ds = ctx->d_small;
    CU(cudaStreamSynchronize(ctx->st2));
    B2_LAUNCH(xxh32_digest_kernel, 1, 32, 0, ctx->st, &ds->xxh, &ds->hash_out, nullptr, nullptr);
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
    const bool legacy = info->block_size_id == B2LZ4_LEGACY_BLOCK_ID;      // [u32 size][LZ4 block] units of a legacy frame
    if (!legacy && (info->block_size_id < 4 || info->block_size_id > 7)) return err_code(E_maxBlockSize_invalid);
    size_t r = set_device(ctx); if (r) return r;
    Small* ds = ctx->d_small;
    if (flags & B2LZ4_STREAM_FIRST) {
        ctx->ring_pos = 65536; ctx->ring_hist = 0;
        if (info->content_checksum) { B2_LAUNCH(xxh32_reset_kernel, 1, 32, 0, ctx->st, &ds->xxh); hsum_reset(ctx->hsum); }
    }
    // independent blocks without a dictionary: the output is hashed on the host (group pipeline and the small-batch path below
    // alike); linked / legacy / dictionary streams keep the running checksum on the device
    const bool host_sum = !legacy && info->block_independent && ctx->dict.empty();
    ctx->hsum_on = host_sum && info->content_checksum;
    if (n == 0) { CU(cudaStreamSynchronize(ctx->st)); return 0; }
    if (!legacy && info->block_independent && !ctx->dict.empty()) {
        // independent blocks WITH a dictionary (what `lz4 -D` writes for single-block inputs, LZ4F_compressFrame_usingCDict,
        // lz4frame.c:392-395): every block may reach into the dictionary but not into its neighbours.  Each block is decoded
        // on its own as "a linked frame of one block whose history is the dictionary" — dictionary frames are small-message
        // territory, one launch sequence per block is fine.
        const uint32_t bs = block_size_of(info->block_size_id), bcs = info->block_checksum ? 4u : 0u, dl = (uint32_t)ctx->dict.size();
        FrameDesc d{}; d.block_size_id = info->block_size_id; d.block_independent = 0; d.block_checksum = info->block_checksum ? 1u : 0u;
        uint8_t hdr[HEADER_MAX]; const uint32_t hs = header_write(hdr, d); const uint32_t zero = 0;
        CU(ctx->dict_stage.ensure(HIST_MAX + bs + 16));
        uint8_t* dout = (uint8_t*)ctx->dict_stage.p + HIST_MAX;
        CU(cudaMemcpyAsync(dout - dl, ctx->dict.data(), dl, cudaMemcpyHostToDevice, ctx->st));
        const uint8_t* s8 = (const uint8_t*)src; size_t ip = 0, o = 0;
        const uint32_t keep_index = ctx->prefs.use_index; ctx->prefs.use_index = 0;
        while (ip < n) {
            if (n - ip < 4) { ctx->prefs.use_index = keep_index; return err_code(E_decompressionFailed); }
            const uint32_t cs = ld_le32(s8 + ip) & 0x7FFFFFFFu;
            if (cs == 0 || cs > bs || n - ip - 4 < (size_t)cs + bcs) { ctx->prefs.use_index = keep_index; return err_code(cs > bs ? E_maxBlockSize_invalid : E_decompressionFailed); }
            const size_t one = 4 + (size_t)cs + bcs;
            CU(ctx->stage_in.ensure(hs + one + 4 + 16));
            uint8_t* din = (uint8_t*)ctx->stage_in.p;
            CU(cudaMemcpyAsync(din, hdr, hs, cudaMemcpyHostToDevice, ctx->st));
            CU(cudaMemcpyAsync(din + hs, s8 + ip, one, cudaMemcpyHostToDevice, ctx->st));
            CU(cudaMemcpyAsync(din + hs + one, &zero, 4, cudaMemcpyHostToDevice, ctx->st));
            size_t used = 0;
            r = decompress_device_impl(ctx, din, hs + one + 4, dout, bs, &used, dl, SCAN_BATCH);
            if (b2lz4_is_error(r)) { ctx->prefs.use_index = keep_index; return r; }
            if (o + r > cap) { ctx->prefs.use_index = keep_index; return err_code(E_dstMaxSize_tooSmall); }
            if (info->content_checksum && r) B2_LAUNCH(xxh32_stream_kernel, 1, 256, 0, ctx->st, dout, r, nullptr, nullptr, &ds->xxh);
            if (r) CU(cudaMemcpyAsync((uint8_t*)dst + o, dout, r, cudaMemcpyDeviceToHost, ctx->st));
            CU(cudaStreamSynchronize(ctx->st));
            o += r; ip += one;
        }
        ctx->prefs.use_index = keep_index;
        return o;
    }
    if (!legacy && info->block_independent) {
        // independent blocks: the same group pipeline as a whole frame (lanes; every device of b2lz4_set_devices)
        HostFrame f;
        f.desc.block_size_id = info->block_size_id; f.desc.block_independent = 1;
        f.desc.block_checksum = info->block_checksum ? 1u : 0u; f.desc.content_checksum = info->content_checksum ? 1u : 0u;
        f.block_size = block_size_of(info->block_size_id);
        const uint8_t* s8 = (const uint8_t*)src;
        const uint32_t bcs = info->block_checksum ? 4u : 0u;
        size_t ip = 0; bool good = true;
        while (ip < n) {
            if (n - ip < 4) { good = false; break; }
            const uint32_t w = ld_le32(s8 + ip), cs = w & 0x7FFFFFFFu;
            if (cs == 0 || cs > f.block_size || n - ip - 4 < (size_t)cs + bcs) { good = false; break; }
            f.blocks.push_back({ip, w});
            ip += 4 + (size_t)cs + bcs;
        }
        f.end_blocks = n;
        const size_t per_group = std::max<size_t>(1, pipe_chunk(ctx, f.blocks.size() * (size_t)f.block_size, f.block_size) / f.block_size);
        if (good && (f.blocks.size() > per_group || multi_dev(ctx))) {
            return decompress_host_groups(ctx, s8, f, per_group, dst, cap, (flags & B2LZ4_STREAM_FIRST) ? HOSTC_BLOCKS_FIRST : HOSTC_BLOCKS);
        }
    }
    // a frame of its own around the batch: header without optional fields, the blocks, an EndMark
    FrameDesc d{};
    d.block_size_id = legacy ? 7u : info->block_size_id; d.block_independent = (legacy || info->block_independent) ? 1u : 0u;
    d.block_checksum = (!legacy && info->block_checksum) ? 1u : 0u;
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
    const bool linked = !legacy && !info->block_independent;
    if (linked) {
        r = ring_reserve(ctx, cap); if (r) return r;
        if ((flags & B2LZ4_STREAM_FIRST) && !ctx->dict.empty()) {      // the dictionary is the frame's first history (lz4frame.c:1842-1860)
            ctx->ring_hist = (uint32_t)ctx->dict.size();
            CU(cudaMemcpyAsync((uint8_t*)ctx->ring.p + ctx->ring_pos - ctx->ring_hist, ctx->dict.data(), ctx->ring_hist, cudaMemcpyHostToDevice, ctx->st));
        }
        dout = (uint8_t*)ctx->ring.p + ctx->ring_pos; hist = ctx->ring_hist;
    } else {
        CU(ctx->stage_out.ensure(cap + 16));
        dout = (uint8_t*)ctx->stage_out.p;
    }
    const uint32_t keep_index = ctx->prefs.use_index;
    ctx->prefs.use_index = 0;                                       // a batch never carries an index frame
    size_t consumed = 0;
    r = decompress_device_impl(ctx, din, hs + n + 4, dout, cap, &consumed, hist, SCAN_BATCH | (legacy ? SCAN_LEGACY : 0u));
    ctx->prefs.use_index = keep_index;
    if (b2lz4_is_error(r)) return r;
    if (consumed != hs + n + 4) return err_code(E_decompressionFailed);   // an EndMark inside the batch: not whole blocks
    if (info->content_checksum && r && !host_sum) B2_LAUNCH(xxh32_stream_kernel, 1, 256, 0, ctx->st, dout, r, nullptr, nullptr, &ds->xxh);
    if (r) CU(cudaMemcpyAsync(dst, dout, r, cudaMemcpyDeviceToHost, ctx->st));
    CU(cudaStreamSynchronize(ctx->st));
    CU(cudaGetLastError());
    if (info->content_checksum && r && host_sum) hsum_update(ctx->hsum, (const uint8_t*)dst, r);
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
    B2_LAUNCH(xxh32_reset_kernel, 1, 32, 0, ctx->st, &ds->xxh);
    B2_LAUNCH(xxh32_stream_kernel, 1, 256, 0, ctx->st, (const uint8_t*)d_src, n, nullptr, nullptr, &ds->xxh);
    B2_LAUNCH(xxh32_digest_kernel, 1, 32, 0, ctx->st, &ds->xxh, &ds->hash_out, nullptr, nullptr);
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
