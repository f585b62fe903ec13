/* b200lz4.h — C ABI of libb200lz4.so, the B200-native LZ4 frame codec behind the
 * Codec.Compression.LZ4.Conduit compress / decompress conduits.
 *
 * Plain pointers and sizes only.  Every function that returns size_t follows the reference's error
 * convention (lz4/lib/lz4frame.c:249-252): a value > (size_t)-B2LZ4_ERROR_maxCode is an error code;
 * test with b2lz4_is_error(), name with b2lz4_error_name().  Codes 1..19 are numbered like
 * LZ4F_LIST_ERRORS (lz4/lib/lz4frame.h:503-524).
 *
 * "Tier 2" (this header) is the batch interface the modified conduit calls; "Tier 1" (lz4frame.h in
 * this directory) is the 12 LZ4F_* symbols the unmodified Conduit.hsc links against.  There is no CPU
 * codec in this library: every compress/decompress entry point launches sm_100a kernels and fails
 * with ERROR_cudaFailure when no device is usable.
 */
#ifndef B200LZ4_H
#define B200LZ4_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B2LZ4_ERROR_cudaFailure 20
#define B2LZ4_ERROR_noEndMark   21   /* Conduit.hsc:689 "stream ended before EndMark" */
#define B2LZ4_ERROR_maxCode     22

typedef struct b2lz4_ctx b2lz4_ctx;

/* Encoder preferences.  Mirrors the fields of the reference's Preferences / FrameInfo records
 * (src/Codec/Compression/LZ4/CTypes.hsc:155-232; lz4/lib/lz4frame.h:169-192) that make sense for
 * an independent-block GPU encoder, plus two knobs of this implementation. */
typedef struct {
    uint32_t block_size_id;     /* 4..7 = 64 KB, 256 KB, 1 MB, 4 MB (lz4frame.h:115-124); 0 = 4 MB; b2lz4_default_prefs sets 5 */
    uint32_t block_checksum;    /* XXH32 after every block (lz4frame.h:186-192 blockChecksumFlag) */
    uint32_t content_checksum;  /* XXH32 of the whole content after the EndMark */
    uint32_t content_size;      /* 1 = store the content size in the frame header (one-shot calls only) */
    uint32_t sub_block_log2;    /* independent sub-block size inside a block: 16..22; 0 = 17 (128 KB); b2lz4_default_prefs sets 18 */
    uint32_t write_index;       /* 1 = append the B2IX skippable index frame ("seekable" frame: block table + sub-block
                                   entry points, lets the decoder run every sub-block in parallel).  Default 0: the
                                   stock lz4 CLI skips skippable frames in files but fails on pipes (lz4io.c:1124-1133). */
    uint32_t use_index;         /* decoder: 1 = trust-but-verify a trailing B2IX index frame */
    uint32_t block_linked;      /* 1 = linked blocks (LZ4F_blockLinked, the reference conduit's own mode, Conduit.hsc:248-263): matches
                                   reach up to 64 KB back across block borders — 64 KB blocks then compress like the reference's.
                                   Such frames carry no index (write_index is ignored); a dictionary (b2lz4_set_dict) is used by
                                   linked frames only. */
} b2lz4_prefs;

typedef struct {                /* parsed frame header (replaces LZ4F_getFrameInfo, lz4frame.c:1246-1281) */
    uint32_t block_size_id, block_independent, block_checksum, content_checksum;
    uint32_t has_content_size, has_dict_id, dict_id, header_size;
    uint64_t content_size;
    uint64_t block_size;
} b2lz4_frame_info;

typedef struct {                /* what the last compress/decompress call on a context did */
    uint32_t n_blocks, n_chains, indexed, kernel_launches;
    float    ms_total;          /* first launch to last kernel, CUDA events on the context's stream */
    float    ms_main;           /* the dominant kernel: match finder (compress) / sequence decoder (decompress) */
    float    ms_aux;            /* everything else on the stream: scan, layout, gather, checksums */
    float    ms_h2d, ms_d2h;    /* host-buffer entry points only */
    uint64_t bytes_in, bytes_out;
    uint32_t retries;           /* workspace growth / layout fallbacks taken */
    uint32_t pad;
} b2lz4_stats;

/* ---- error helpers (replace LZ4F_isError / LZ4F_getErrorName, lz4frame.c:249-259) */
unsigned    b2lz4_is_error(size_t code);
const char* b2lz4_error_name(size_t code);
const char* b2lz4_last_cuda_error(const b2lz4_ctx* ctx);

/* ---- context (replaces LZ4F_create/freeCompressionContext + Decompression, lz4frame.c:536-563, 1065-1090).
 * One context = one CUDA device + its streams + reusable workspaces.  Not thread-safe; independent
 * contexts may be used concurrently from any OS thread. */
size_t b2lz4_create(b2lz4_ctx** out, int device, const b2lz4_prefs* prefs /* NULL = defaults */);
size_t b2lz4_destroy(b2lz4_ctx* ctx);                       /* NULL is a no-op */
size_t b2lz4_set_prefs(b2lz4_ctx* ctx, const b2lz4_prefs* prefs);
void   b2lz4_default_prefs(b2lz4_prefs* prefs);
/* Multi-GPU host-buffer calls: b2lz4_compress_host / b2lz4_decompress_host cut the frame into chunks of
 * whole blocks and send them round robin to lanes on these devices (each with its own streams and PCIe link); the host
 * reassembles the output in block order.  No collective is involved — independent blocks have no exchange step.
 * n = 0 returns to the context's own device.  The frame bytes do not depend on the device list. */
size_t b2lz4_set_devices(b2lz4_ctx* ctx, const int* devices, int n);
/* Dictionary (replaces LZ4F_createCDict + LZ4F_compressBegin_usingCDict / LZ4F_decompress_usingDict, lz4/lib/lz4frame.c:478-521,
 * 594-702, 1842-1860): the bytes are treated as stream history in front of the first block of every frame that follows.  Only the
 * last 64 KB matter and are kept (copied).  The compressor uses it for linked-block frames (prefs.block_linked); the decompressor
 * (host-buffer entry points) also takes independent-block frames with a dictionary, which `lz4 -D` writes for one-block inputs
 * (lz4frame.c:392-395), one block at a time.  dict_id != 0 is
 * written into the frame header (lz4frame.c:692-695).  n = 0 removes the dictionary. */
size_t b2lz4_set_dict(b2lz4_ctx* ctx, const void* dict, size_t n, uint32_t dict_id);
int    b2lz4_device_count(void);

/* ---- sizes (replaces LZ4F_compressBound / LZ4F_compressFrameBound, lz4frame.c:324-349, 370-381) */
size_t b2lz4_compress_bound(size_t n, const b2lz4_prefs* prefs);

/* upper bound of the decoded size of the first frame in src[0, n): number of blocks (host walk over the block header
 * words, lz4frame.c:1470-1500) x block size.  What a caller without a content-size field sizes its destination by. */
size_t b2lz4_decompress_bound(const void* src, size_t n);

/* ---- host-only frame helpers: no GPU involved (replace lz4frame.c:663-701 and :1109-1198) */
size_t b2lz4_frame_header_write(void* dst, size_t cap, const b2lz4_prefs* prefs, uint64_t content_size);
size_t b2lz4_frame_header_write2(void* dst, size_t cap, const b2lz4_prefs* prefs, uint64_t content_size, uint32_t dict_id);
size_t b2lz4_frame_header_parse(const void* src, size_t n, b2lz4_frame_info* info);   /* returns header size */

/* ---- one LZ4 frame, buffers resident in device memory (config 2).  Return the frame size / the decoded
 * size.  `consumed` (may be NULL) receives the compressed bytes the frame occupied (incl. a trailing index). */
size_t b2lz4_compress_device(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap);
size_t b2lz4_decompress_device(b2lz4_ctx* ctx, const void* d_src, size_t n, void* d_dst, size_t cap, size_t* consumed);

/* ---- one LZ4 frame, buffers in host memory (pinned buffers from b2lz4_pinned_alloc make the copies fast) */
size_t b2lz4_compress_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap);
size_t b2lz4_decompress_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, size_t* consumed);
void*  b2lz4_pinned_alloc(size_t n);
void   b2lz4_pinned_free(void* p);

/* ---- streaming pieces: a conduit that cuts the stream into batches of whole blocks itself calls these
 * (they are what the LZ4F_* layer in lz4frame.h is built on).  Replaces the LZ4F_compressUpdate /
 * LZ4F_compressEnd / LZ4F_decompress inner loops (lz4frame.c:823-925, 984-1018, 1371-1840) with one call
 * per batch.  Prefs of the context decide block size / checksums. */
#define B2LZ4_LEGACY_BLOCK_ID 8u /* b2lz4_frame_info.block_size_id for the [u32 size][block] units of a legacy frame
                                   (magic 0x184C2102, 8 MB blocks, lz4/programs/lz4io.c:895-940): b2lz4_decompress_blocks_host only */
#define B2LZ4_STREAM_FIRST 1u   /* first batch of a frame: resets the running content checksum and the decode history */
/* n bytes -> [u32 size|stored][data][block checksum]... ; n must be a multiple of the block size except in the last batch */
size_t b2lz4_compress_blocks_host(b2lz4_ctx* ctx, const void* src, size_t n, void* dst, size_t cap, unsigned flags);
/* EndMark + content checksum (when enabled) of everything compressed since B2LZ4_STREAM_FIRST: 4 or 8 bytes */
size_t b2lz4_frame_end(b2lz4_ctx* ctx, void* dst, size_t cap);
/* n bytes of WHOLE blocks of a frame described by `info` (no frame header, no EndMark) -> decoded bytes.
 * Linked-block frames keep their 64 KB history on the device between calls. */
size_t b2lz4_decompress_blocks_host(b2lz4_ctx* ctx, const b2lz4_frame_info* info, const void* src, size_t n,
                                    void* dst, size_t cap, unsigned flags);
/* XXH32 of everything compressed / decoded since B2LZ4_STREAM_FIRST (content checksum, xxhash.c:529-554) */
size_t b2lz4_content_checksum(b2lz4_ctx* ctx, uint32_t* out);

/* ---- XXH32 of a device / host buffer, seed 0 (replaces XXH32, xxhash.c:392-416) */
size_t b2lz4_xxh32_device(b2lz4_ctx* ctx, const void* d_src, size_t n, uint32_t* out);
size_t b2lz4_xxh32_host(b2lz4_ctx* ctx, const void* src, size_t n, uint32_t* out);

/* ---- XXH32 (seed 0) of n_ranges byte ranges [offs[i], offs[i] + lens[i]) of the device buffer d_src[0, n_src) — the
 * frame layer's block checksums — by an entry point that ALLOCATES NOTHING: no context, no cudaMalloc / cudaFree /
 * cudaMallocHost, no stream or event creation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Descriptors and results pass through a file-scope __device__ array of 1 MiB + 256 B (csrc/noalloc.cu) that every
 * GPU which loads the library holds until the process exits; hence the limit B2LZ4_STATIC_MAX_RANGES per call, beyond
 * which the call fails with ERROR_srcSize_tooLarge (as does a range that leaves the buffer) and
 * b2lz4_static_last_error() says why.  offs / lens / out are HOST arrays of n_ranges entries; `stream` is a
 * cudaStream_t of `device` (NULL = the default stream) on which the work is queued; the call returns after it has
 * finished and leaves the caller's current device unchanged.  All calls on a device share the one array, which keeps the
 * last call's contents: calls of one process are serialised inside, but work queued against the array by other means
 * while a call runs would race. */
#define B2LZ4_STATIC_MAX_RANGES 65536u
size_t b2lz4_xxh32_ranges_static(int device, void* stream, const void* d_src, size_t n_src,
                                 const uint64_t* offs, const uint32_t* lens, uint32_t n_ranges, uint32_t* out);
const char* b2lz4_static_last_error(void);   /* text of this thread's last failure of the call above ("" if none) */

size_t b2lz4_last_stats(const b2lz4_ctx* ctx, b2lz4_stats* out);

#ifdef __cplusplus
}
#endif
#endif
