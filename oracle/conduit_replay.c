/* oracle/conduit_replay.c — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * A C replay of the exact call sequences the Haskell conduits issue against the reference
 * lz4 frame library.  GHC is not available in this image, so this harness *is* "the reference
 * conduit on CPU" (minus Haskell allocation overhead): it is linked against the UNMODIFIED
 * reference sources (see oracle/Makefile, target `ref`) and only uses the public LZ4F_* API.
 *
 *   ref_conduit_compress    follows src/Codec/Compression/LZ4/Conduit.hsc:457-533
 *                           (compressWithOutBufferSize; `compress` = bufferSize 0, :336-337)
 *   ref_conduit_compress_yi follows Conduit.hsc:364-425 (compressYieldImmediately)
 *   ref_conduit_decompress  follows Conduit.hsc:598-701 (decompress)
 *   ref_frame_compress      generic LZ4F streaming compression with arbitrary preferences, one
 *                           LZ4F_compressUpdate per `updateChunk` bytes (what lz4io.c:660-690 does
 *                           with updateChunk == blockSize and autoFlush == 1)
 *
 * All functions return the number of bytes produced, or an LZ4F error code (test with
 * LZ4F_isError), or REPLAY_ERR_* for harness-level failures that the Haskell code reports with
 * throwString.
 *   ref_frame_compress_dict / ref_frame_decompress_dict   the dictionary entry points (LZ4F_createCDict,
 *                           LZ4F_compressBegin_usingCDict, LZ4F_decompress_usingDict; lz4frame.h:530-607)
 */
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#define LZ4F_STATIC_LINKING_ONLY          /* the reference declares the dictionary API in its static-linking section */
#include "lz4frame.h"
/* -DREPLAY_AGAINST_B200: the same replay compiled against include/lz4frame.h and linked to libb200lz4.so —
 * tests/ use it to drive the product's LZ4F_* layer with exactly the calls the Haskell conduit makes. */
#ifndef REPLAY_AGAINST_B200
#include "lz4.h"
#define XXH_STATIC_LINKING_ONLY
#include "xxhash.h"
#endif

#define REPLAY_ERR_DST_TOO_SMALL   ((size_t)-100)
#define REPLAY_ERR_HEADER_SHORT    ((size_t)-101)  /* Conduit.hsc:615-616 "not enough bytes for header" */
#define REPLAY_ERR_NO_ENDMARK      ((size_t)-102)  /* Conduit.hsc:689 "stream ended before EndMark" */
#define REPLAY_ERR_ALLOC           ((size_t)-103)

#define MINSZ(a, b) ((a) < (b) ? (a) : (b))

/* lz4DefaultPreferences, Conduit.hsc:248-263 (everything zero = linked, 64 KB, no checksums) */
static void default_prefs(LZ4F_preferences_t* p) { memset(p, 0, sizeof(*p)); }

typedef struct { uint8_t* p; size_t cap, len; int overflow; } sink_t;
static void sink_put(sink_t* s, const void* b, size_t n) {
    if (s->len + n > s->cap) { s->overflow = 1; return; }
    memcpy(s->p + s->len, b, n); s->len += n;
}

/* Conduit.hsc:457-533.  `inChunk` is the size of the upstream ByteStrings (0 = one ByteString). */
size_t ref_conduit_compress(const void* src_, size_t n, void* dst, size_t cap,
                            size_t inChunk, size_t bufferSize)
{
    const uint8_t* src = (const uint8_t*)src_;
    sink_t out = { (uint8_t*)dst, cap, 0, 0 };
    LZ4F_cctx* ctx = NULL;
    LZ4F_preferences_t prefs; default_prefs(&prefs);
    size_t r = LZ4F_createCompressionContext(&ctx, LZ4F_VERSION);            /* :194-203 */
    if (LZ4F_isError(r)) return r;

    const size_t bsInChunkSize = 16 * 1024;                                  /* :464 */
    const size_t compressBound = LZ4F_compressBound(LZ4F_HEADER_SIZE_MAX + bsInChunkSize, &prefs); /* :466 */
    const size_t outBufferSize = bufferSize > compressBound ? bufferSize : compressBound;           /* :467 */
    uint8_t* outBuf = (uint8_t*)malloc(outBufferSize);
    if (!outBuf) { LZ4F_freeCompressionContext(ctx); return REPLAY_ERR_ALLOC; }

    r = LZ4F_compressBegin(ctx, outBuf, outBufferSize, &prefs);              /* :475 */
    if (LZ4F_isError(r)) goto done;
    size_t remaining = outBufferSize - r;                                    /* :533 */

    if (inChunk == 0) inChunk = n ? n : 1;
    for (size_t pos = 0; pos < n; ) {                                        /* await -> Just bs */
        size_t bsLen = MINSZ(inChunk, n - pos);
        for (size_t off = 0; off < bsLen; ) {                                /* bsChunksOf 16384, :501 */
            size_t piece = MINSZ(bsInChunkSize, bsLen - off);
            if (remaining < compressBound) {                                 /* :506-512 */
                sink_put(&out, outBuf, outBufferSize - remaining);
                remaining = outBufferSize;
            }
            r = LZ4F_compressUpdate(ctx, outBuf + (outBufferSize - remaining), remaining,
                                    src + pos + off, piece, NULL);           /* :520-525 */
            if (LZ4F_isError(r)) goto done;
            remaining -= r;
            off += piece;
        }
        pos += bsLen;
    }
    {   /* await -> Nothing, :485-498 */
        size_t footerSize = LZ4F_compressBound(0, &prefs);
        if (remaining < footerSize) {
            sink_put(&out, outBuf, outBufferSize - remaining);
            remaining = outBufferSize;
        }
        r = LZ4F_compressEnd(ctx, outBuf + (outBufferSize - remaining), remaining, NULL);
        if (LZ4F_isError(r)) goto done;
        sink_put(&out, outBuf, outBufferSize - remaining + r);
        r = out.overflow ? REPLAY_ERR_DST_TOO_SMALL : out.len;
    }
done:
    free(outBuf);
    LZ4F_freeCompressionContext(ctx);
    return r;
}

/* Conduit.hsc:364-425 */
size_t ref_conduit_compress_yi(const void* src_, size_t n, void* dst, size_t cap, size_t inChunk)
{
    const uint8_t* src = (const uint8_t*)src_;
    sink_t out = { (uint8_t*)dst, cap, 0, 0 };
    LZ4F_cctx* ctx = NULL;
    LZ4F_preferences_t prefs; default_prefs(&prefs);
    size_t r = LZ4F_createCompressionContext(&ctx, LZ4F_VERSION);
    if (LZ4F_isError(r)) return r;
    uint8_t hdr[LZ4F_HEADER_SIZE_MAX];
    r = LZ4F_compressBegin(ctx, hdr, sizeof hdr, &prefs);                    /* :371-373 */
    if (LZ4F_isError(r)) goto done;
    sink_put(&out, hdr, r);
    if (inChunk == 0) inChunk = n ? n : 1;
    for (size_t pos = 0; pos < n; ) {                                        /* awaitForever, :379 */
        size_t bsLen = MINSZ(inChunk, n - pos);
        size_t size = LZ4F_compressBound(bsLen, &prefs);                     /* :385 */
        uint8_t* buf = (uint8_t*)malloc(size);                               /* allocaBytes, :400 */
        if (!buf) { r = REPLAY_ERR_ALLOC; goto done; }
        r = LZ4F_compressUpdate(ctx, buf, size, src + pos, bsLen, NULL);     /* :403 */
        if (!LZ4F_isError(r) && r != 0) sink_put(&out, buf, r);              /* :405-407 */
        free(buf);
        if (LZ4F_isError(r)) goto done;
        pos += bsLen;
    }
    {   size_t footerSize = LZ4F_compressBound(0, &prefs);                   /* :419 */
        uint8_t* buf = (uint8_t*)malloc(footerSize);
        if (!buf) { r = REPLAY_ERR_ALLOC; goto done; }
        r = LZ4F_compressEnd(ctx, buf, footerSize, NULL);                    /* :422 */
        if (!LZ4F_isError(r)) sink_put(&out, buf, r);
        free(buf);
        if (LZ4F_isError(r)) goto done;
        r = out.overflow ? REPLAY_ERR_DST_TOO_SMALL : out.len;
    }
done:
    LZ4F_freeCompressionContext(ctx);
    return r;
}

/* Conduit.hsc:598-701.  `inChunk` = size of upstream ByteStrings after the header bytes (0 = all).
 * `*consumed` receives how many source bytes the conduit consumed (it stops at the first EndMark
 * [+ content checksum]; trailing bytes stay upstream, :693-695). */
size_t ref_conduit_decompress(const void* src_, size_t n, void* dst, size_t cap,
                              size_t inChunk, size_t* consumed)
{
    const uint8_t* src = (const uint8_t*)src_;
    sink_t out = { (uint8_t*)dst, cap, 0, 0 };
    LZ4F_dctx* ctx = NULL;
    size_t r = LZ4F_createDecompressionContext(&ctx, LZ4F_VERSION);          /* :568 */
    if (LZ4F_isError(r)) return r;
    size_t pos = 0;
    uint8_t* dstBuf = NULL;

    if (n < 5) { r = REPLAY_ERR_HEADER_SHORT; goto done; }                   /* :614-616 */
    {   int contentSizeBit = (src[4] >> 3) & 1;                              /* :618-619 */
        size_t hdrLen = 5 + (contentSizeBit ? 2 + 8 : 2);                    /* :621-627 */
        if (hdrLen > n) hdrLen = n;                                          /* CB.take returns what exists */
        LZ4F_frameInfo_t info;
        size_t srcSize = hdrLen;
        r = LZ4F_getFrameInfo(ctx, &info, src, &srcSize);                    /* :629-632 */
        if (LZ4F_isError(r)) goto done;
        pos = hdrLen;
    }
    {   size_t hint = r;
        size_t dstBufSize = 16 * 1024;                                       /* :634-635 */
        dstBuf = (uint8_t*)malloc(dstBufSize);
        if (!dstBuf) { r = REPLAY_ERR_ALLOC; goto done; }
        if (inChunk == 0) inChunk = n > pos ? n - pos : 1;
        while (hint != 0) {                                                  /* loop, :686-695 */
            if (pos >= n) { r = REPLAY_ERR_NO_ENDMARK; goto done; }          /* :689 */
            size_t bsLen = MINSZ(inChunk, n - pos);
            size_t off = 0;
            while (1) {                                                      /* loopSingleBs, :661-684 */
                size_t outBufSize = hint > 16 * 1024 ? hint : 16 * 1024;     /* :665 */
                if (outBufSize > dstBufSize) {                               /* :650-659 */
                    uint8_t* nb = (uint8_t*)realloc(dstBuf, outBufSize);
                    if (!nb) { r = REPLAY_ERR_ALLOC; goto done; }
                    dstBuf = nb; dstBufSize = outBufSize;
                }
                size_t dstSz = outBufSize, srcSz = bsLen - off;
                r = LZ4F_decompress(ctx, dstBuf, &dstSz, src + pos + off, &srcSz, NULL); /* :673 */
                if (LZ4F_isError(r)) goto done;
                sink_put(&out, dstBuf, dstSz);                               /* :676-679 */
                hint = r;
                off += srcSz;
                if (off >= bsLen) break;                                     /* :682-684 */
            }
            pos += bsLen;
            /* Note: when the frame ends inside this ByteString the Haskell code has still taken the
             * whole ByteString from upstream; `consumed` reports the codec-level position instead. */
            if (hint == 0) { pos = pos - bsLen + off; }
        }
        r = out.overflow ? REPLAY_ERR_DST_TOO_SMALL : out.len;
    }
done:
    if (consumed) *consumed = pos;
    free(dstBuf);
    LZ4F_freeDecompressionContext(ctx);
    return r;
}

/* Generic LZ4F streaming compression (what the stock CLI does for multi-block inputs,
 * lz4io.c:660-690, when updateChunk == block size and autoFlush == 1). */
size_t ref_frame_compress(const void* src_, size_t n, void* dst_, size_t cap,
                          int blockSizeID, int blockIndependent, int contentChecksum,
                          int blockChecksum, int withContentSize, int autoFlush, size_t updateChunk)
{
    const uint8_t* src = (const uint8_t*)src_;
    uint8_t* dst = (uint8_t*)dst_;
    LZ4F_cctx* ctx = NULL;
    LZ4F_preferences_t prefs; memset(&prefs, 0, sizeof prefs);
    prefs.frameInfo.blockSizeID = (LZ4F_blockSizeID_t)blockSizeID;
    prefs.frameInfo.blockMode = blockIndependent ? LZ4F_blockIndependent : LZ4F_blockLinked;
    prefs.frameInfo.contentChecksumFlag = contentChecksum ? LZ4F_contentChecksumEnabled : LZ4F_noContentChecksum;
    prefs.frameInfo.blockChecksumFlag = blockChecksum ? LZ4F_blockChecksumEnabled : LZ4F_noBlockChecksum;
    prefs.frameInfo.contentSize = withContentSize ? (unsigned long long)n : 0;
    prefs.autoFlush = (unsigned)autoFlush;
    size_t r = LZ4F_createCompressionContext(&ctx, LZ4F_VERSION);
    if (LZ4F_isError(r)) return r;
    size_t o = 0;
    r = LZ4F_compressBegin(ctx, dst, cap, &prefs);
    if (LZ4F_isError(r)) goto done;
    o += r;
    if (updateChunk == 0) updateChunk = n ? n : 1;
    for (size_t pos = 0; pos < n; ) {
        size_t piece = MINSZ(updateChunk, n - pos);
        r = LZ4F_compressUpdate(ctx, dst + o, cap - o, src + pos, piece, NULL);
        if (LZ4F_isError(r)) goto done;
        o += r; pos += piece;
    }
    r = LZ4F_compressEnd(ctx, dst + o, cap - o, NULL);
    if (LZ4F_isError(r)) goto done;
    o += r; r = o;
done:
    LZ4F_freeCompressionContext(ctx);
    return r;
}

/* One-shot full-frame decode with a large destination (library fast path), for cross-checks. */
size_t ref_frame_decompress(const void* src_, size_t n, void* dst_, size_t cap, size_t* consumed)
{
    const uint8_t* src = (const uint8_t*)src_;
    uint8_t* dst = (uint8_t*)dst_;
    LZ4F_dctx* ctx = NULL;
    size_t r = LZ4F_createDecompressionContext(&ctx, LZ4F_VERSION);
    if (LZ4F_isError(r)) return r;
    size_t ip = 0, op = 0;
    while (1) {
        size_t dstSz = cap - op, srcSz = n - ip;
        r = LZ4F_decompress(ctx, dst + op, &dstSz, src + ip, &srcSz, NULL);
        if (LZ4F_isError(r)) break;
        ip += srcSz; op += dstSz;
        if (r == 0) { r = op; break; }
        if (srcSz == 0 && dstSz == 0) { r = (ip >= n) ? REPLAY_ERR_NO_ENDMARK : REPLAY_ERR_DST_TOO_SMALL; break; }
    }
    if (consumed) *consumed = ip;
    LZ4F_freeDecompressionContext(ctx);
    return r;
}

#ifndef REPLAY_AGAINST_B200
unsigned ref_xxh32(const void* p, size_t n, unsigned seed) { return XXH32(p, n, seed); }

/* streaming XXH32 over arbitrary splits (content-checksum behaviour, xxhash.c:437-554) */
unsigned ref_xxh32_stream(const void* p_, size_t n, unsigned seed, size_t piece)
{
    const uint8_t* p = (const uint8_t*)p_;
    XXH32_state_t st; XXH32_reset(&st, seed);
    if (piece == 0) piece = n ? n : 1;
    for (size_t o = 0; o < n; o += piece) XXH32_update(&st, p + o, MINSZ(piece, n - o));
    return XXH32_digest(&st);
}

#endif

size_t ref_sizeof_frameinfo(void)   { return sizeof(LZ4F_frameInfo_t); }
size_t ref_sizeof_preferences(void) { return sizeof(LZ4F_preferences_t); }


/* Streaming compression with a dictionary: LZ4F_createCDict + LZ4F_compressBegin_usingCDict (lz4frame.c:478-521, 594-702),
 * then LZ4F_compressUpdate per `updateChunk` bytes and LZ4F_compressEnd (what lz4/tests/frametest.c:520-662 exercises). */
size_t ref_frame_compress_dict(const void* src_, size_t n, void* dst_, size_t cap, const void* dict, size_t dictSize,
                               int blockSizeID, int blockIndependent, int contentChecksum, unsigned dictID, size_t updateChunk)
{
    const uint8_t* src = (const uint8_t*)src_;
    uint8_t* dst = (uint8_t*)dst_;
    LZ4F_cctx* ctx = NULL;
    LZ4F_CDict* cd = LZ4F_createCDict(dict, dictSize);
    if (!cd) return REPLAY_ERR_ALLOC;
    LZ4F_preferences_t prefs; memset(&prefs, 0, sizeof prefs);
    prefs.frameInfo.blockSizeID = (LZ4F_blockSizeID_t)blockSizeID;
    prefs.frameInfo.blockMode = blockIndependent ? LZ4F_blockIndependent : LZ4F_blockLinked;
    prefs.frameInfo.contentChecksumFlag = contentChecksum ? LZ4F_contentChecksumEnabled : LZ4F_noContentChecksum;
    prefs.frameInfo.dictID = dictID;
    size_t r = LZ4F_createCompressionContext(&ctx, LZ4F_VERSION);
    if (LZ4F_isError(r)) { LZ4F_freeCDict(cd); return r; }
    size_t o = 0;
    r = LZ4F_compressBegin_usingCDict(ctx, dst, cap, cd, &prefs);
    if (LZ4F_isError(r)) goto done;
    o += r;
    if (updateChunk == 0) updateChunk = n ? n : 1;
    for (size_t pos = 0; pos < n; ) {
        size_t piece = MINSZ(updateChunk, n - pos);
        r = LZ4F_compressUpdate(ctx, dst + o, cap - o, src + pos, piece, NULL);
        if (LZ4F_isError(r)) goto done;
        o += r; pos += piece;
    }
    r = LZ4F_compressEnd(ctx, dst + o, cap - o, NULL);
    if (LZ4F_isError(r)) goto done;
    o += r; r = o;
done:
    LZ4F_freeCompressionContext(ctx);
    LZ4F_freeCDict(cd);
    return r;
}

/* LZ4F_decompress_usingDict (lz4frame.c:1842-1860) in a loop over `chunk`-byte input slices. */
size_t ref_frame_decompress_dict(const void* src_, size_t n, void* dst_, size_t cap, const void* dict, size_t dictSize, size_t chunk)
{
    const uint8_t* src = (const uint8_t*)src_;
    uint8_t* dst = (uint8_t*)dst_;
    LZ4F_dctx* ctx = NULL;
    size_t r = LZ4F_createDecompressionContext(&ctx, LZ4F_VERSION);
    if (LZ4F_isError(r)) return r;
    size_t ip = 0, op = 0, hint = 1;
    if (chunk == 0) chunk = n ? n : 1;
    while (hint != 0) {
        if (ip == n) { r = REPLAY_ERR_NO_ENDMARK; goto done; }
        size_t ssz = MINSZ(chunk, n - ip), dsz = cap - op;
        hint = LZ4F_decompress_usingDict(ctx, dst + op, &dsz, src + ip, &ssz, dict, dictSize, NULL);
        if (LZ4F_isError(hint)) { r = hint; goto done; }
        ip += ssz; op += dsz;
        if (ssz == 0 && dsz == 0 && op == cap) { r = REPLAY_ERR_DST_TOO_SMALL; goto done; }
    }
    r = op;
done:
    LZ4F_freeDecompressionContext(ctx);
    return r;
}
