// device_types.h — plain structs shared by the host engine and the sm_100a kernels.
#pragma once
#include <stdint.h>
#include "format.h"

namespace b2 {

// ------------------------------------------------------------------ decode side
// One LZ4 frame block as found by the frame scan (lz4frame.c:1475-1500 walks the same words).
struct BlockRef {
    uint64_t src_off;     // offset of the block's data bytes inside the compressed buffer
    uint32_t csize;       // stored/compressed size in bytes (flag stripped)
    uint32_t stored;      // 1 = raw (uncompressed) block
    uint64_t dst_off;     // where the block's output starts (valid when the layout is known up front)
    uint32_t usize;       // decoded size if known up front (index present), else 0
    uint32_t pad;
};

// A piece = a run of whole LZ4 sequences decoded in order by one warp.
//   foreign frame: piece = block;   indexed frame (ours): piece = independent sub-block of a block.
enum : uint32_t { PIECE_STORED = 1u, PIECE_ENDS_BLOCK = 2u, PIECE_EXACT = 4u,
                  PIECE_BULK = 8u };   // stored block whose copy is done by copy_stored_kernel, not by the decoding warp
struct Piece {
    uint64_t src_off;     // first token of the piece (absolute offset in the compressed buffer)
    uint32_t src_len;     // compressed bytes of the piece
    uint32_t flags;
    uint32_t dst_len;     // expected decoded bytes (PIECE_EXACT) — else unused
    uint32_t block;       // owning block (for the block-level limits the reference's checks use)
    uint32_t dst_rel;     // offset of the piece's output inside its block's output (PIECE_EXACT)
    uint32_t pad;
};

// A chain = consecutive pieces decoded by ONE warp into contiguous output.
//   independent blocks: one chain per piece;  linked-block frame: one chain for the whole frame.
struct Chain {
    uint64_t dst_off;     // first output byte of the chain
    uint32_t first_piece;
    uint32_t n_pieces;
};

struct PieceOut {
    uint32_t out_len;     // decoded bytes
    uint32_t err;         // 0 = ok, else b2::Err
};

// Result of the frame scan; lives in device memory and is mirrored to the host once per call.
struct FrameIndex {
    uint32_t err;             // b2::Err of the scan
    uint32_t n_blocks;
    uint32_t n_pieces;
    uint32_t n_chains;
    uint32_t block_size;      // bytes
    uint32_t indexed;         // 1 = a valid B2IX index frame followed the frame and was used
    uint32_t layout_known;    // 1 = every block's dst_off is final before decoding
    uint32_t stored_content_checksum;
    uint64_t consumed;        // compressed bytes from frame start through EndMark (+ content checksum)
    uint64_t consumed_with_index;   // ... plus the trailing index frame when present
    uint64_t total_out;       // decoded size when known up front (indexed), else upper bound n_blocks*block_size
    uint32_t need_blocks;     // table sizes the scan wanted (for the retry after E_allocation_failed)
    uint32_t need_pieces;
    uint32_t n_stored;        // stored blocks listed for copy_stored_kernel
    uint32_t walk_pending;    // 1 = the serial walk over the block headers is left to the decoder kernel (2 = done there)
    uint32_t df;              // 1 = decoded by the dataflow kernels (decode_df.cu): no side information, any block mode
    uint32_t need_groups;     // dataflow workspace the scan wanted (groups of DF_G compressed bytes)
    uint32_t need_tickets;    // ... and tickets (upper bound from the compressed sizes, exact after df_layout_kernel)
    FrameDesc desc;
};

// ---- dataflow decode (decode_df.cu)
constexpr uint32_t DF_G = 1024u;            // compressed bytes per group = one chain table
constexpr uint32_t DF_LA = 96u;             // bytes behind a group its parse may look at (length runs longer than that: DF_SLOW)
constexpr uint32_t DF_RUN = 64u;            // 255-bytes of one length field followed inside the table pass
constexpr uint32_t DF_BATCH = 32u;          // sequences per ticket (one per lane of the warp that decodes it)
// B2_DF_SMEMWIN: the unit's most recent output lives in a circular DATA window in shared memory next to the bitmap; every byte
// is written there, a match reads its source there (a shared-memory round trip instead of one through L2 per link of the
// dependency chain), and the warp that retires a ticket writes its bytes to global memory in 16-byte rows.
#ifndef B2_DF_SMEMWIN
#define B2_DF_SMEMWIN 0
#endif
#if B2_DF_SMEMWIN
constexpr uint32_t DF_WIN_BYTES = 128u << 10;   // output window covered by the bitmap (16 KB of shared memory) and by the data window (128 KB)
#else
constexpr uint32_t DF_WIN_BYTES = 512u << 10;   // output window covered by the CTA's "final bytes" bitmap (64 KB of shared memory)
#endif
constexpr uint32_t DF_WIN_MASK = DF_WIN_BYTES / 32u - 1u;
constexpr uint32_t DF_HALF = DF_WIN_BYTES / 2u; // a ticket is handed out when its output ends within this much of the watermark
constexpr uint32_t DF_DATA_BYTES = B2_DF_SMEMWIN ? DF_WIN_BYTES : 0u, DF_DATA_MASK = DF_DATA_BYTES ? DF_DATA_BYTES - 1u : 0u;
#ifndef B2_DF_NAP_MIN
#define B2_DF_NAP_MIN 64
#endif
#ifndef B2_DF_NAP_MAX
#define B2_DF_NAP_MAX 1024
#endif
constexpr uint32_t DF_NAP_MIN = B2_DF_NAP_MIN, DF_NAP_MAX = B2_DF_NAP_MAX;   // back-off of a waiting warp, ns
constexpr uint32_t DF_RING = 128u;          // tickets between the lowest unfinished and the newest handed out (done flags)
struct DfHead { uint32_t exit, out; };      // chain from one position of a group: packed (kind, sequences, where it leaves), bytes produced
struct DfShared {                           // per-CTA dataflow state in shared memory (decode_df_kernel)
    uint32_t* bits;                         // DF_WIN_BYTES / 32 words, circular: bit of output byte p is (p mod DF_WIN_BYTES)
    volatile unsigned long long* final_below;   // every output byte below this buffer offset is final
    uint8_t* wdata;                         // B2_DF_SMEMWIN: DF_DATA_BYTES of shared memory, byte of buffer offset p at (p + wlo) mod DF_DATA_BYTES
    uint8_t* gdst;                          // ... the destination buffer (buffer offset 0)
    uint32_t wlo;                           // ... low 4 bits of gdst (16-byte rows of the window are 16-byte rows of global memory)
    uint64_t gcap;                          // ... its capacity (nothing is flushed beyond it)
};
struct DfState {
    unsigned long long bad_key;             // min over failures of (block << 8 | Err); all ones = none
    unsigned long long produced, total_out;
    uint32_t n_tickets, n_units, ticket_links, ticket_units, ticket_tk, pad;
    unsigned long long prof[12];             // -DB2_DF_PROF: cycles spent per phase of decode_df_kernel, summed over warps (tools/foreign_speed.py prints them)
};
struct DfWork {                             // device workspaces of the dataflow path (sized by the host from the scan's needs)
    uint32_t* grp_base;                     // [blocks + 1] first group of every block
    DfHead* heads;                          // [groups x DF_G]
    uint2* ent;                             // [groups] (first token, output offset) of the k-th group a block's chain enters
    uint32_t* ent_seq;                      // [groups] ... and the number of that token's sequence within the block
    uint32_t* blk_nent; uint32_t* blk_usize; uint32_t* blk_nseq;   // [blocks]
    uint32_t* tk_base;                      // [blocks + 1] first ticket of every block
    uint2* tk;                              // [cap_tickets] (token, output offset) of every DF_BATCH-th sequence of a block
    DfState* st;
    uint32_t cap_groups, cap_tickets;
};

// Decode summary written by the finalise kernel
struct DecodeResult {
    uint32_t err;             // first error in block order (b2::Err), 0 = ok
    uint32_t err_block;
    uint64_t total_out;
    uint32_t content_checksum;    // XXH32 of the output when the frame carries one (computed on device)
    uint32_t layout_ok;           // 0 = a non-final block decoded short while the layout assumed full blocks
};

// ------------------------------------------------------------------ encode side
struct EncParams {
    uint64_t n;               // uncompressed bytes in this call
    uint32_t block_size;
    uint32_t sub_size;        // encoder unit: parsed by one warp, stitched by layout/gather (divides block_size)
    uint32_t subs_per_block;
    uint32_t dom_units;       // units per match domain (= the independent sub-block one decoder warp decodes; index entries)
    uint32_t n_blocks;
    uint32_t n_subs;          // total sub-blocks
    uint32_t scratch_stride;  // bytes of scratch per sub-block
    uint32_t block_checksum;
    uint32_t content_checksum;
    uint32_t write_header;    // 1 = emit frame header (+ EndMark, checksum slot, index) around the blocks
    uint32_t write_index;     // 1 = append the B2IX skippable index frame
    uint32_t linked;          // 1 = linked blocks: a match may reach up to 65535 bytes back across block borders (lz4_Frame_format.md:155-163)
    uint32_t hist;            // bytes of earlier stream (or dictionary) that lie directly in front of src in device memory
    FrameDesc desc;
};

struct SubOut {               // written by the match finder, one per sub-block
    uint32_t out_len;         // bytes emitted into scratch
    uint32_t first_ll;        // literal length of the first emitted sequence
    uint32_t first_hdr;       // token + length bytes of that first sequence
    uint32_t tail;            // trailing literals NOT emitted (carried into the next sub-block)
    uint32_t emitted;         // 0 = no sequence at all (whole sub-block is carried literals)
    uint32_t n_seq;
};

struct SubPlace {             // written by the layout kernel
    uint32_t dst_rel;         // offset of this sub-block's bytes inside the block's compressed data
    uint32_t carry_in;        // literals inherited from the previous sub-block(s)
    uint32_t new_hdr;         // size of the re-encoded first token + length bytes
    uint32_t size;            // rank + 1 of this sub-block among the decoder entry points of its block (first emitting unit of a match domain), else 0
};

struct BlockPlace {
    uint64_t frame_off;       // offset of the block's 4-byte header word inside the output
    uint32_t csize;           // bytes of block data (== usize when stored)
    uint32_t stored;
    uint32_t usize;
    uint32_t n_chain;         // emitted sub-blocks (index entries)
    uint32_t first_chain;
    uint32_t pad;
};

struct EncodeResult {
    uint32_t err;
    uint32_t n_blocks;
    uint64_t frame_len;           // total bytes written (incl. index frame when requested)
    uint64_t blocks_len;          // bytes of the block region only
    uint64_t content_checksum_off;// where the u32 content checksum goes (0 = none)
    uint64_t index_off;           // where the index frame starts (0 = none)
    uint32_t n_chains;
    uint32_t pad;
};

// XXH32 streaming state kept in device memory for the pipelined content checksum (xxhash.c:437-554)
struct XxhState {
    uint32_t v[4];
    uint32_t mem[4];
    uint32_t memsize;
    uint32_t total_len;   // wraps mod 2^32 like the reference (xxhash.c:464)
    uint32_t large;
    uint32_t pad;
};

// ---- B2IX index frame (a skippable frame, lz4_Frame_format.md:298-335, appended after the LZ4 frame).
// payload: IndexHead | n_blocks * IndexBlock | n_chains * IndexChain | u32 payload_bytes | u32 IDX_MAGIC
constexpr uint32_t IDX_SKIP_MAGIC = 0x184D2A5Bu;
constexpr uint32_t IDX_MAGIC      = 0x58493242u;   // "B2IX"
struct IndexHead  { uint32_t magic, version, block_size, n_blocks, n_chains, reserved; uint64_t content_size; };
struct IndexBlock { uint32_t hdr_word, usize, first_chain, n_chain; };
struct IndexChain { uint32_t tok_off, out_off; };   // relative to the block's data / the block's output

}  // namespace b2
