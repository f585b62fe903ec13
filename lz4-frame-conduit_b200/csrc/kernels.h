// kernels.h — launch-side declarations of the sm_100a kernels (decode.cu, encode.cu, xxh32.cu).
#pragma once
#include "b2_cuda.h"
#include <stdint.h>
#include "device_types.h"

namespace b2 {

#ifndef B2_DEC_WARPS
#define B2_DEC_WARPS 8
#endif
constexpr int DEC_WARPS = B2_DEC_WARPS;          // warps per CTA in the sequence decoder
#ifndef B2_ENC_WARPS
#define B2_ENC_WARPS 8
#endif
#ifndef B2_ENC_SMEM_WARPS
#define B2_ENC_SMEM_WARPS 6
#endif
#ifndef B2_ENC_CTAS
#define B2_ENC_CTAS 4
#endif
constexpr int ENC_WARPS = B2_ENC_WARPS;          // warps per CTA in the match finder (one hash table each)
constexpr int ENC_SMEM_WARPS = B2_ENC_SMEM_WARPS;     // ... of which this many keep their table in shared memory; the others use a global (L1/L2) table
constexpr int ENC_CTAS_PER_SM = B2_ENC_CTAS;   // resident CTAs per SM the kernel is built for
#ifndef B2_ENC_HLOG
#define B2_ENC_HLOG 12
#endif
constexpr int ENC_HLOG  = B2_ENC_HLOG;  // 4096-entry u16 hash table per warp (8 KB), the reference's table size for large inputs
#ifndef B2_ENC_UNIT_LOG2
#define B2_ENC_UNIT_LOG2 14
#endif
constexpr uint32_t ENC_UNIT = 1u << B2_ENC_UNIT_LOG2;   // bytes one match-finder warp parses at a time (a domain = sub-block holds several)
constexpr int LAYOUT_THREADS = 256;   // layout_kernel's single CTA
#ifndef B2_DF_WARPS
#define B2_DF_WARPS 8
#endif
constexpr int DF_WARPS = B2_DF_WARPS;            // warps per CTA in df_links_kernel (9.3 KB of shared memory each)
enum : uint32_t { SCAN_USE_INDEX = 1u, SCAN_SINGLE_CHAIN = 2u, SCAN_BATCH = 4u,
                  SCAN_DF = 32u,              // frames without a usable index may go to the dataflow kernels (decode_df.cu)
                  SCAN_DF_ALL = 64u,          // ... whatever their block size (B2LZ4_DF=2; the exact-layout retry)
                  SCAN_DF_FEW = 128u,         // ... and calls of few blocks (one warp per block would leave most SMs with one warp or none)
                  SCAN_LEGACY = 8u,
                  SCAN_OVERLAP_WALK = 16u };  // plain independent-block frames: walk the block headers inside the decoder kernel   // blocks of a legacy frame (lz4_Frame_format.md:338-386): 8 MB, always compressed, no flag bit
constexpr uint32_t LEGACY_BLOCK = 8u << 20;
constexpr uint32_t DF_FEW_BLOCKS = 256u;     // SCAN_DF_FEW: at most this many blocks fit the destination

// ---- decode
__global__ void frame_scan_kernel(const uint8_t* src, uint64_t n, uint64_t dst_cap, uint32_t mode,
                                  FrameIndex* fi, BlockRef* blocks, uint32_t cap_blocks,
                                  Piece* pieces, uint32_t cap_pieces, Chain* chains, uint32_t* stored_list, uint32_t* walk, DfWork W);
__global__ void copy_stored_kernel(const uint8_t* src, uint8_t* dst, uint64_t dst_cap, const FrameIndex* fi,
                                   const Piece* pieces, const Chain* chains, const uint32_t* stored_list);
__global__ void decode_chains_kernel(const uint8_t* src, uint64_t n_src, uint8_t* dst, uint64_t dst_cap, uint32_t mode,
                                     FrameIndex* fi, BlockRef* blocks, Piece* pieces, Chain* chains, uint32_t cap_tables,
                                     uint32_t* stored_list, PieceOut* outs, uint32_t* ticket, uint32_t* walk, uint32_t hist);
__global__ void decode_finalize_kernel(const FrameIndex* fi, const Piece* pieces, const PieceOut* outs,
                                       const uint32_t* block_crc_bad, DecodeResult* res, uint32_t* ticket);
__global__ void verify_block_checksums_kernel(const uint8_t* src, const FrameIndex* fi,
                                              const BlockRef* blocks, uint32_t* first_bad);

// ---- decode, dataflow path
__global__ void df_links_kernel(const uint8_t* src, uint64_t n_src, const FrameIndex* fi, const BlockRef* blocks, DfWork W);
size_t df_links_smem_bytes();
__global__ void df_stitch_kernel(const uint8_t* src, const FrameIndex* fi, const BlockRef* blocks, DfWork W);
__global__ void df_layout_kernel(FrameIndex* fi, BlockRef* blocks, uint64_t dst_cap, DfWork W);
__global__ void df_tickets_kernel(const uint8_t* src, const FrameIndex* fi, const BlockRef* blocks, DfWork W);
size_t decode_df_smem_bytes(uint32_t warps);
__global__ void decode_df_kernel(const uint8_t* src, uint64_t n_src, uint8_t* dst, uint64_t dst_cap,
                                 const FrameIndex* fi, const BlockRef* blocks, DfWork W, uint32_t hist);
__global__ void df_finalize_kernel(const FrameIndex* fi, const uint32_t* block_crc_bad, DecodeResult* res, DfWork W);

// ---- encode
__global__ void find_emit_kernel(const uint8_t* src, EncParams P, uint8_t* scratch, SubOut* subs, uint32_t* ticket, uint16_t* gtab);
size_t find_emit_gtab_bytes(uint32_t grid);
__global__ void layout_blocks_kernel(EncParams P, const SubOut* subs, SubPlace* places, BlockPlace* bplaces);
__global__ void layout_kernel(EncParams P, BlockPlace* bplaces,
                              uint8_t* dst, uint64_t dst_cap, EncodeResult* res, uint32_t* ticket, uint8_t* idx_frag);
__global__ void gather_kernel(const uint8_t* src, EncParams P, const uint8_t* scratch, const SubOut* subs,
                              const SubPlace* places, const BlockPlace* bplaces, uint8_t* dst,
                              const EncodeResult* res, uint8_t* idx_frag);
__global__ void block_checksums_kernel(EncParams P, const BlockPlace* bplaces, uint8_t* dst, const EncodeResult* res);

// ---- XXH32
__global__ void xxh32_stream_kernel(const uint8_t* data, uint64_t n, const uint64_t* n_ptr,
                                    const uint32_t* enable_ptr, XxhState* st);                // update (1 CTA x 256)
__global__ void xxh32_digest_kernel(const XxhState* st, uint32_t* out, uint8_t* base, const uint64_t* off_ptr);
__global__ void content_check_kernel(const FrameIndex* fi, const XxhState* st, DecodeResult* res);
__global__ void xxh32_reset_kernel(XxhState* st);
__global__ void xxh32_many_kernel(const uint8_t* data, const uint64_t* offs, const uint32_t* lens,
                                  uint32_t n, uint32_t* out);                                 // one hash per range

size_t find_emit_smem_bytes();

}  // namespace b2
