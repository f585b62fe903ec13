/* lz4frame.h — "Tier 1" C ABI of libb200lz4.so: the twelve LZ4F_* entry points (and the struct layouts
 * hsc2hs reads) that the UNMODIFIED reference binds through inline-c, served by the B200 codec.
 *
 * Reference binding sites (src/Codec/Compression/LZ4/Conduit.hsc unless noted):
 *   LZ4F_isError / LZ4F_getErrorName           :150-151        (lz4/lib/lz4frame.h:100-101)
 *   LZ4F_createCompressionContext              :199, :242      (lz4frame.h:247)
 *   LZ4F_freeCompressionContext                :181, :210      (lz4frame.h:248)
 *   LZ4F_compressBegin                         :292            (lz4frame.h:272-275)
 *   LZ4F_compressBound                         :302            (lz4frame.h:291)
 *   LZ4F_compressUpdate                        :311            (lz4frame.h:304-308)
 *   LZ4F_compressEnd                           :321            (lz4frame.h:332-335)
 *   LZ4F_createDecompressionContext            :568            (lz4frame.h:360)
 *   LZ4F_freeDecompressionContext              :545            (lz4frame.h:361)
 *   LZ4F_getFrameInfo                          :579            (lz4frame.h:424-426)
 *   LZ4F_decompress                            :591            (lz4frame.h:457-460)
 *   struct sizes / field offsets / enum values : CTypes.hsc:60-76, 90-95, 172-199, 211-229
 * LZ4F_flush, LZ4F_getVersion and LZ4F_resetDecompressionContext are provided as well (same header
 * section of the reference) because callers of the streaming API commonly pair them with the above.
 *
 * Behavioural contract kept from the reference:
 *   - every size_t result is a byte count or an error code; LZ4F_isError() tells which; names are the
 *     reference's (lz4frame.h:503-524) plus ERROR_cudaFailure / ERROR_noEndMark;
 *   - LZ4F_compressUpdate may buffer; the caller must offer LZ4F_compressBound(srcSize) bytes;
 *   - LZ4F_decompress consumes any slice of input, produces at most *dstSizePtr bytes, returns a hint
 *     (0 = frame complete, else "more input expected") and is resumable at every byte;
 *   - free(NULL) is a no-op; contexts are not thread-safe, distinct contexts are independent.
 * What differs (allowed by the format, lz4_Frame_format.md): frames are always written with independent
 * blocks, LZ4F_default block size means 256 KB, and the encoder works in batches (B2LZ4_BATCH_MB, default
 * 12 MiB for compression, one on the GPU while the next is staged), so output arrives in larger pieces, one batch late,
 * and LZ4F_compressBound(srcSize) = the stored size of (2 batches - 1 + srcSize) bytes + 8: about 24 MiB with the
 * default batch (the reference: one block + srcSize).  It is what ANY single call may have to emit — LZ4F_compressEnd
 * included, which the reference conduit serves from a buffer of LZ4F_compressBound(19 + 16384) bytes, Conduit.hsc:466-495 —
 * so it cannot depend on srcSize alone.  The default batch was chosen with a measurement: callers that allocate the bound
 * per input piece (compressYieldImmediately, Conduit.hsc:385-400) run at the speed of `compress` below 32 MiB.
 * LZ4F_decompress, details a caller may notice (none is used by Conduit.hsc, which only tests the result for 0, error, or
 * "anything else"):
 *   - while decoded bytes are waiting to be handed out, the return value is a hint for the NEXT DESTINATION size (at most
 *     512 KB), not the next source size as in the reference (lz4frame.h:457-468 allows any non-zero "hint");
 *   - when a call has taken all its input but still holds output, *srcSizePtr reports one byte fewer than was taken, so
 *     that a caller that loops "until the input is consumed" comes back for the output; the library remembers the byte,
 *     the caller must present it again (any correct caller does: it re-sends what was reported unconsumed);
 *   - a skippable frame ends like any frame: the context is reset and the call returns 0 (lz4frame.c:1789-1799).
 * There is no CPU codec behind these symbols: without a usable CUDA device they return ERROR_cudaFailure.
 */
#ifndef B200_LZ4FRAME_H
#define B200_LZ4FRAME_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define LZ4F_VERSION 100
#define LZ4F_HEADER_SIZE_MIN 7
#define LZ4F_HEADER_SIZE_MAX 19
#define LZ4F_BLOCK_HEADER_SIZE 4
#define LZ4F_BLOCK_CHECKSUM_SIZE 4
#define LZ4F_CONTENT_CHECKSUM_SIZE 4

typedef size_t LZ4F_errorCode_t;

typedef enum { LZ4F_default = 0, LZ4F_max64KB = 4, LZ4F_max256KB = 5, LZ4F_max1MB = 6, LZ4F_max4MB = 7 } LZ4F_blockSizeID_t;
typedef enum { LZ4F_blockLinked = 0, LZ4F_blockIndependent } LZ4F_blockMode_t;
typedef enum { LZ4F_noContentChecksum = 0, LZ4F_contentChecksumEnabled } LZ4F_contentChecksum_t;
typedef enum { LZ4F_noBlockChecksum = 0, LZ4F_blockChecksumEnabled } LZ4F_blockChecksum_t;
typedef enum { LZ4F_frame = 0, LZ4F_skippableFrame } LZ4F_frameType_t;

typedef struct {                               /* 32 bytes, same field order as the reference (lz4frame.h:169-177) */
    LZ4F_blockSizeID_t     blockSizeID;
    LZ4F_blockMode_t       blockMode;
    LZ4F_contentChecksum_t contentChecksumFlag;
    LZ4F_frameType_t       frameType;
    unsigned long long     contentSize;
    unsigned               dictID;
    LZ4F_blockChecksum_t   blockChecksumFlag;
} LZ4F_frameInfo_t;

typedef struct {                               /* 56 bytes (lz4frame.h:186-192) */
    LZ4F_frameInfo_t frameInfo;
    int      compressionLevel;                 /* ignored: one fast level */
    unsigned autoFlush;                        /* 1 = every LZ4F_compressUpdate compresses what it was given */
    unsigned favorDecSpeed;                    /* ignored */
    unsigned reserved[3];
} LZ4F_preferences_t;

typedef struct { unsigned stableSrc; unsigned reserved[3]; } LZ4F_compressOptions_t;
typedef struct { unsigned stableDst; unsigned reserved[3]; } LZ4F_decompressOptions_t;

typedef struct LZ4F_cctx_s LZ4F_cctx;
typedef LZ4F_cctx* LZ4F_compressionContext_t;
typedef struct LZ4F_dctx_s LZ4F_dctx;
typedef LZ4F_dctx* LZ4F_decompressionContext_t;
typedef struct LZ4F_CDict_s LZ4F_CDict;

unsigned    LZ4F_isError(LZ4F_errorCode_t code);
const char* LZ4F_getErrorName(LZ4F_errorCode_t code);
unsigned    LZ4F_getVersion(void);

LZ4F_errorCode_t LZ4F_createCompressionContext(LZ4F_cctx** cctxPtr, unsigned version);
LZ4F_errorCode_t LZ4F_freeCompressionContext(LZ4F_cctx* cctx);
size_t LZ4F_compressBegin(LZ4F_cctx* cctx, void* dstBuffer, size_t dstCapacity, const LZ4F_preferences_t* prefsPtr);
size_t LZ4F_compressBound(size_t srcSize, const LZ4F_preferences_t* prefsPtr);
size_t LZ4F_compressUpdate(LZ4F_cctx* cctx, void* dstBuffer, size_t dstCapacity,
                           const void* srcBuffer, size_t srcSize, const LZ4F_compressOptions_t* cOptPtr);
size_t LZ4F_flush(LZ4F_cctx* cctx, void* dstBuffer, size_t dstCapacity, const LZ4F_compressOptions_t* cOptPtr);
size_t LZ4F_compressEnd(LZ4F_cctx* cctx, void* dstBuffer, size_t dstCapacity, const LZ4F_compressOptions_t* cOptPtr);

LZ4F_errorCode_t LZ4F_createDecompressionContext(LZ4F_dctx** dctxPtr, unsigned version);
LZ4F_errorCode_t LZ4F_freeDecompressionContext(LZ4F_dctx* dctx);
size_t LZ4F_getFrameInfo(LZ4F_dctx* dctx, LZ4F_frameInfo_t* frameInfoPtr, const void* srcBuffer, size_t* srcSizePtr);
size_t LZ4F_decompress(LZ4F_dctx* dctx, void* dstBuffer, size_t* dstSizePtr,
                       const void* srcBuffer, size_t* srcSizePtr, const LZ4F_decompressOptions_t* dOptPtr);
void   LZ4F_resetDecompressionContext(LZ4F_dctx* dctx);

/* Dictionaries (the reference lists them as TODO, Conduit.hsc:21-23; lz4/lib/lz4frame.h:530-607, lz4frame.c:478-521, 594-702,
 * 1842-1860).  A dictionary is stream history in front of the first block: it is used by LINKED-block frames, which
 * LZ4F_compressBegin_usingCDict writes whenever prefs->frameInfo.blockMode says so (LZ4F_blockLinked is the zero value).
 * Plain LZ4F_compressBegin keeps writing independent blocks unless B2LZ4_BLOCK_MODE=honour is set in the environment. */
LZ4F_CDict* LZ4F_createCDict(const void* dictBuffer, size_t dictSize);
void        LZ4F_freeCDict(LZ4F_CDict* CDict);
size_t LZ4F_compressBegin_usingCDict(LZ4F_cctx* cctx, void* dstBuffer, size_t dstCapacity, const LZ4F_CDict* cdict,
                                     const LZ4F_preferences_t* prefsPtr);
size_t LZ4F_decompress_usingDict(LZ4F_dctx* dctx, void* dstBuffer, size_t* dstSizePtr, const void* srcBuffer, size_t* srcSizePtr,
                                 const void* dict, size_t dictSize, const LZ4F_decompressOptions_t* decompressOptionsPtr);

#ifdef __cplusplus
}
#endif
#endif
