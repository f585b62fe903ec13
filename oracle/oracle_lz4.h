/* oracle/oracle_lz4.h — CPU restatement of the reference hot path (TEST INFRASTRUCTURE ONLY).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.  The product
 * (lz4-frame-conduit_b200/) never links, imports or executes anything under oracle/.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against the unmodified
 * reference library built by `make -C oracle ref` (oracle/_ref/libref.so) — block compressor output
 * byte-for-byte, decoder output, XXH32 values, frame headers — and against the committed golden
 * fixtures in tests/golden/ (generated from the reference by tests/golden/make_golden.py).
 */
#ifndef ORACLE_LZ4_H
#define ORACLE_LZ4_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* error codes mirror LZ4F_LIST_ERRORS (lz4frame.h:503-524); returned as (size_t)-code */
enum {
    ORC_OK = 0, ORC_ERR_GENERIC, ORC_ERR_maxBlockSize_invalid, ORC_ERR_blockMode_invalid,
    ORC_ERR_contentChecksumFlag_invalid, ORC_ERR_compressionLevel_invalid, ORC_ERR_headerVersion_wrong,
    ORC_ERR_blockChecksum_invalid, ORC_ERR_reservedFlag_set, ORC_ERR_allocation_failed,
    ORC_ERR_srcSize_tooLarge, ORC_ERR_dstMaxSize_tooSmall, ORC_ERR_frameHeader_incomplete,
    ORC_ERR_frameType_unknown, ORC_ERR_frameSize_wrong, ORC_ERR_srcPtr_wrong, ORC_ERR_decompressionFailed,
    ORC_ERR_headerChecksum_invalid, ORC_ERR_contentChecksum_invalid, ORC_ERR_frameDecoding_alreadyStarted,
    ORC_ERR_maxCode
};
#define ORC_ERROR(c)   ((size_t)-(ptrdiff_t)(c))
static inline int orc_is_error(size_t r) { return r > ORC_ERROR(ORC_ERR_maxCode); }

typedef struct {
    int      block_size_id;       /* 0 (=64 KB) or 4..7 */
    int      block_independent;   /* 1 = independent blocks */
    int      content_checksum;
    int      block_checksum;
    int      has_content_size;
    uint64_t content_size;
    int      has_dict_id;
    uint32_t dict_id;
} orc_frame_desc;

/* xxhash.c:392-416 (one-shot), :437-554 (streaming) */
uint32_t orc_xxh32(const void* data, size_t len, uint32_t seed);
typedef struct { uint32_t total_len, large, v[4], mem[4], memsize; } orc_xxh32_state;
void     orc_xxh32_reset(orc_xxh32_state* s, uint32_t seed);
void     orc_xxh32_update(orc_xxh32_state* s, const void* data, size_t len);
uint32_t orc_xxh32_digest(const orc_xxh32_state* s);

/* lz4.c:789-1177 via :1211-1245 (independent block, level 0, limitedOutput).  Returns the compressed
 * size, or 0 when the result would not fit into `cap` bytes. */
int orc_block_compress(const uint8_t* src, int n, uint8_t* dst, int cap);

/* lz4.c:1653-2068 (endOnInputSize, full, noDict / withPrefix64k).  `prefix` = number of valid
 * history bytes located immediately before dst.  Returns decoded size or a negative value. */
int orc_block_decompress(const uint8_t* src, int csize, uint8_t* dst, int cap, size_t prefix);

size_t orc_block_size(int block_size_id);                                        /* lz4frame.c:278-287 */
size_t orc_frame_header_write(uint8_t* dst, size_t cap, const orc_frame_desc* d);/* lz4frame.c:663-701 */
size_t orc_frame_header_parse(const uint8_t* src, size_t n, orc_frame_desc* d);  /* lz4frame.c:1109-1198 */
size_t orc_frame_compress_bound(size_t n, const orc_frame_desc* d);
/* independent-block frame compression == LZ4F_compressBegin / per-block LZ4F_makeBlock / compressEnd */
size_t orc_frame_compress(const uint8_t* src, size_t n, uint8_t* dst, size_t cap, const orc_frame_desc* d);
/* full frame decode (independent and linked blocks, stored blocks, all checksums verified) */
size_t orc_frame_decompress(const uint8_t* src, size_t n, uint8_t* dst, size_t cap, size_t* consumed);

#ifdef __cplusplus
}
#endif
#endif
