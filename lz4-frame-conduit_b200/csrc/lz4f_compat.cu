// lz4f_compat.cu — the LZ4F_* streaming entry points of include/lz4frame.h on top of the batch engine.
//
// These are the twelve symbols src/Codec/Compression/LZ4/Conduit.hsc binds through inline-c (list with
// line numbers in include/lz4frame.h).  The reference implements them as a per-16 KB CPU state machine
// (lz4/lib/lz4frame.c:600-1018 compress side, :1109-1840 decompress side); here they are a thin host layer
// that collects whole blocks in pinned staging buffers and hands them to the GPU a batch at a time
// (b2lz4_compress_blocks_host / b2lz4_decompress_blocks_host).  Nothing in this file compresses or
// decompresses: without a CUDA device the calls return ERROR_cudaFailure.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include <algorithm>
#include "format.h"
#include "../../include/b200lz4.h"
#include "../../include/lz4frame.h"

using namespace b2;

namespace {

inline size_t ec(uint32_t e) { return (size_t)0 - (size_t)e; }

size_t batch_bytes() {                       // B2LZ4_BATCH_MB: uncompressed bytes handed to the GPU per call (default 64 MiB)
    static size_t v = 0;
    if (!v) {
        const char* e = getenv("B2LZ4_BATCH_MB");
        long mb = e ? atol(e) : 64;
        if (mb < 4) mb = 4;
        if (mb > 2048) mb = 2048;
        v = ((size_t)mb << 20) & ~(((size_t)4 << 20) - 1);
    }
    return v;
}
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

struct Pinned {                              // growable pinned host buffer
    uint8_t* p = nullptr; size_t cap = 0;
    bool ensure(size_t n, size_t keep = 0) {
        if (n <= cap) return true;
        uint8_t* q = (uint8_t*)b2lz4_pinned_alloc(n + n / 4 + 4096);
        if (!q) return false;
        if (keep) memcpy(q, p, keep);
        b2lz4_pinned_free(p);
        p = q; cap = n + n / 4 + 4096;
        return true;
    }
    void release() { b2lz4_pinned_free(p); p = nullptr; cap = 0; }
};

// LZ4F_default -> 256 KB blocks here, each parsed as ONE independent unit: such a frame needs no side information to be
// decoded block-parallel (by this library's decoder as by any other), and its ratio on log-like data is 1 % better than
// the reference's default (linked 64 KB blocks).  4 MB / 1 MB blocks stay available through blockSizeID.
constexpr unsigned DEFAULT_BLOCK_ID = 5u;
uint32_t block_bytes_of(unsigned id) { return block_size_of(id ? id : DEFAULT_BLOCK_ID); }

// worst case bytes of the block region for n input bytes: every block stored + header word + checksum
size_t blocks_bound(size_t n, uint32_t bs, bool bc) { return n + ((n + bs - 1) / bs) * (4u + (bc ? 4u : 0u)); }

}  // namespace

// =========================================================================================== compression
struct LZ4F_cctx_s {
    b2lz4_ctx* gpu = nullptr;
    LZ4F_preferences_t prefs{};
    b2lz4_prefs bp{};
    uint32_t block_size = 0;
    size_t batch = 0;
    Pinned in;  size_t in_len = 0;
    Pinned out;
    bool begun = false, first = true;
    uint64_t total_in = 0;
};

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
    b2lz4_destroy(c->gpu);
    c->in.release(); c->out.release();
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
    const size_t buffered = p->autoFlush ? 0 : batch_bytes() - 1;
    return blocks_bound(buffered + srcSize, bs, p->frameInfo.blockChecksumFlag != 0) + 8;
}

size_t LZ4F_compressBegin(LZ4F_cctx* c, void* dst, size_t cap, const LZ4F_preferences_t* p) {
    if (!c || !dst) return ec(E_GENERIC);
    if (cap < (size_t)LZ4F_HEADER_SIZE_MAX) return ec(E_dstMaxSize_tooSmall);       // lz4frame.c:613
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
    c->block_size = block_bytes_of(id);
    c->batch = batch_bytes();
    if (!c->gpu) {
        size_t r = b2lz4_create(&c->gpu, device_index(), &c->bp);
        if (b2lz4_is_error(r)) return r;
        r = apply_devices(c->gpu);
        if (b2lz4_is_error(r)) return r;
    } else {
        size_t r = b2lz4_set_prefs(c->gpu, &c->bp);
        if (b2lz4_is_error(r)) return r;
    }
    c->in_len = 0; c->total_in = 0; c->first = true; c->begun = true;
    // frames always carry independent blocks; a dictID in the preferences is not supported (no dictionary is applied)
    return b2lz4_frame_header_write(dst, cap, &c->bp, c->prefs.frameInfo.contentSize);
}

// compresses the staged bytes [0, n) into dst; returns bytes written
static size_t emit_staged(LZ4F_cctx* c, size_t n, uint8_t* dst, size_t cap) {
    if (n == 0) return 0;
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
    uint8_t* dst = (uint8_t*)dstBuffer;
    const bool bc = c->bp.block_checksum != 0;
    // bytes this call will push through the GPU: whole batches, or everything when autoFlush is set
    const size_t have = c->in_len + n;
    const size_t go = c->prefs.autoFlush ? have : (have / c->batch) * c->batch;
    if (cap < blocks_bound(go, c->block_size, bc)) return ec(E_dstMaxSize_tooSmall); // lz4frame.c:837-838
    if (!dst && go) return ec(E_GENERIC);
    const uint8_t* s = (const uint8_t*)src;
    size_t written = 0, left = n;
    const size_t stage_cap = c->prefs.autoFlush ? std::max(have, c->batch) : c->batch;
    if (!c->in.ensure(stage_cap + 16, c->in_len)) return ec(E_allocation_failed);
    while (left) {
        const size_t room = stage_cap - c->in_len;
        const size_t take = left < room ? left : room;
        memcpy(c->in.p + c->in_len, s, take);                                        // pinned staging (replaces tmpIn, lz4frame.c:846-868)
        c->in_len += take; s += take; left -= take;
        if (c->in_len == c->batch && !c->prefs.autoFlush) {
            size_t r = emit_staged(c, c->in_len, dst + written, cap - written);
            if (b2lz4_is_error(r)) return r;
            written += r; c->in_len = 0;
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
    if (c->in_len == 0) return 0;                                                    // lz4frame.c:945
    if (cap < blocks_bound(c->in_len, c->block_size, c->bp.block_checksum != 0)) return ec(E_dstMaxSize_tooSmall);
    size_t r = emit_staged(c, c->in_len, (uint8_t*)dstBuffer, cap);
    if (b2lz4_is_error(r)) return r;
    c->in_len = 0;
    return r;
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
    b2lz4_ctx* gpu = nullptr;
    DStage stage = DS_HEADER;
    uint8_t hdr[HEADER_MAX]; uint32_t hdr_len = 0;          // header / small fixed-size elements being collected
    b2lz4_frame_info info{};
    Pinned in;  size_t in_len = 0;                           // whole blocks (+ the one being collected)
    size_t complete_end = 0; uint32_t n_complete = 0;        // end of the last whole block in `in`, how many there are
    size_t need = 0;                                         // bytes still missing for the element being collected
    uint64_t skip_left = 0;
    Pinned out; size_t out_len = 0, out_pos = 0;
    uint64_t total_out = 0;
    bool first = true;
    size_t pre_acked = 0;                                    // input bytes already buffered but reported as unconsumed
    size_t batch = 0;
};

namespace {

void dctx_reset(LZ4F_dctx_s* d) {
    d->stage = DS_HEADER; d->hdr_len = 0; d->in_len = 0; d->complete_end = 0; d->n_complete = 0; d->need = 0;
    d->skip_left = 0; d->out_len = d->out_pos = 0; d->total_out = 0; d->first = true; d->pre_acked = 0;
}

// decode the whole blocks collected so far
size_t decode_batch(LZ4F_dctx_s* d) {
    if (d->n_complete == 0) return 0;
    if (!d->gpu) {
        b2lz4_prefs bp; b2lz4_default_prefs(&bp); bp.use_index = 0;
        size_t r = b2lz4_create(&d->gpu, device_index(), &bp);
        if (b2lz4_is_error(r)) return r;
        r = apply_devices(d->gpu);
        if (b2lz4_is_error(r)) return r;
    }
    const size_t cap = (size_t)d->n_complete * d->info.block_size;
    if (!d->out.ensure(cap + 16)) return ec(E_allocation_failed);
    size_t r = b2lz4_decompress_blocks_host(d->gpu, &d->info, d->in.p, d->complete_end, d->out.p, cap,
                                            d->first ? B2LZ4_STREAM_FIRST : 0u);
    if (b2lz4_is_error(r)) return r;
    d->first = false;
    d->out_len = r; d->out_pos = 0; d->total_out += r;
    memmove(d->in.p, d->in.p + d->complete_end, d->in_len - d->complete_end);
    d->in_len -= d->complete_end; d->complete_end = 0; d->n_complete = 0;
    return 0;
}

// header bytes complete -> parse; returns 0 or an error code
size_t header_done(LZ4F_dctx_s* d) {
    size_t r = b2lz4_frame_header_parse(d->hdr, d->hdr_len, &d->info);
    if (b2lz4_is_error(r)) return r;
    d->stage = DS_BLOCK_HEADER; d->hdr_len = 0; d->need = 4;
    d->batch = batch_bytes();
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
    b2lz4_destroy(d->gpu);
    d->in.release(); d->out.release();
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

size_t LZ4F_decompress(LZ4F_dctx* d, void* dstBuffer, size_t* dstSizePtr, const void* srcBuffer, size_t* srcSizePtr,
                       const LZ4F_decompressOptions_t* opt) {
    (void)opt;
    if (!d || !dstSizePtr || !srcSizePtr) return ec(E_GENERIC);
    const uint8_t* src = (const uint8_t*)srcBuffer;
    uint8_t* dst = (uint8_t*)dstBuffer;
    const size_t srcSize = src ? *srcSizePtr : 0, dstCap = dst ? *dstSizePtr : 0;
    size_t consumed = 0, produced = 0;
    *srcSizePtr = 0; *dstSizePtr = 0;
    auto fail = [&](size_t e) { dctx_reset(d); return e; };
    {   // bytes taken in an earlier call but handed back so that the caller comes again for the pending output
        const size_t sk = d->pre_acked < srcSize ? d->pre_acked : srcSize;
        consumed += sk; d->pre_acked -= sk;
    }
    for (;;) {
        if (d->out_pos < d->out_len) {                                               // hand out decoded bytes first
            const size_t k = std::min(dstCap - produced, d->out_len - d->out_pos);
            memcpy(dst + produced, d->out.p + d->out_pos, k);
            produced += k; d->out_pos += k;
            if (d->out_pos < d->out_len) break;                                      // destination full
        }
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
                if (d->hdr_len == 8) { d->skip_left = ld_le32(d->hdr + 4); d->hdr_len = 0; d->stage = DS_SKIP_DATA; }
            } else if (d->hdr_len >= 5 && ld_le32(d->hdr) == MAGIC_FRAME && d->hdr_len == header_size_from_flg(d->hdr[4])) {
                size_t r = header_done(d);
                if (b2lz4_is_error(r)) return fail(r);
            }
            break;
        }
        case DS_SKIP_DATA: {                                                         // lz4frame.c:1756-1801
            const size_t k = (size_t)std::min<uint64_t>(avail, d->skip_left);
            consumed += k; d->skip_left -= k;
            if (d->skip_left == 0) d->stage = DS_HEADER;                             // a skippable frame is not "a frame": keep going
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
                size_t r = decode_batch(d);
                if (b2lz4_is_error(r)) return fail(r);
                d->stage = DS_SUFFIX; d->need = d->info.content_checksum ? 4 : 0; d->hdr_len = 0;
                if (d->need == 0) {
                    if (d->info.has_content_size && d->info.content_size != d->total_out) return fail(ec(E_frameSize_wrong));
                    d->stage = DS_DONE;
                }
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
            memcpy(d->in.p + d->in_len, src + consumed, k); d->in_len += k; consumed += k; d->need -= k;
            if (d->need) break;
            d->complete_end = d->in_len; ++d->n_complete;
            d->stage = DS_BLOCK_HEADER; d->need = 4;
            // enough for the GPU: the batch's output bound or its compressed size reached the batch size
            if ((size_t)(d->n_complete + 1) * d->info.block_size > d->batch || d->complete_end >= d->batch) {
                size_t r = decode_batch(d);
                if (b2lz4_is_error(r)) return fail(r);
            }
            break;
        }
        case DS_SUFFIX: {                                                            // content checksum (lz4frame.c:1734-1745)
            const size_t k = std::min(avail, d->need);
            memcpy(d->hdr + d->hdr_len, src + consumed, k); d->hdr_len += (uint32_t)k; consumed += k; d->need -= k;
            if (d->need) break;
            uint32_t h = 0;
            if (d->first) h = 0x02CC5D05u;                                           // XXH32 of no bytes, seed 0 (frame without blocks)
            else { size_t r = b2lz4_content_checksum(d->gpu, &h); if (b2lz4_is_error(r)) return fail(r); }
            if (h != ld_le32(d->hdr)) return fail(ec(E_contentChecksum_invalid));
            if (d->info.has_content_size && d->info.content_size != d->total_out) return fail(ec(E_frameSize_wrong));
            d->stage = DS_DONE;
            break;
        }
        default: break;
        }
    }
    const bool pending = d->out_pos < d->out_len;
    if (pending && consumed == srcSize && consumed > 0) { --consumed; ++d->pre_acked; }   // make the caller come back for the rest
    *srcSizePtr = consumed; *dstSizePtr = produced;
    if (d->stage == DS_DONE && !pending) {                                           // frame complete (lz4frame.h:457-468: returns 0)
        const size_t keep = d->pre_acked;
        dctx_reset(d); d->pre_acked = keep;
        return 0;
    }
    if (pending) return d->out_len - d->out_pos;                                     // hint: the caller may offer this much room
    return d->need ? d->need + (d->stage == DS_BLOCK_DATA ? 4 : 0) : 4;
}

}  // extern "C"
