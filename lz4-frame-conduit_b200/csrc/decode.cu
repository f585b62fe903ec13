// decode.cu — LZ4 frame decoding on sm_100a: frame scan, sequence decoder, result reduction.
//
// Replaces, on this path (citations relative to /root/reference/):
//   LZ4F_decompress block walk            lz4/lib/lz4frame.c:1371-1840   -> frame_scan_kernel
//   LZ4_decompress_generic (safe loop)    lz4/lib/lz4.c:1653-2068        -> decode_chains_kernel
//   block / content checksum verification lz4frame.c:1604-1611,1734-1745 -> xxh32 kernels (xxh32.cu)
// Written from the format documents (lz4_Block_format.md, lz4_Frame_format.md) and the behaviour of
// the reference decoder (its acceptance rules, lz4.c:1882-2035, are mirrored here).
#include "common.cuh"
#include "kernels.h"

namespace b2 {

// ------------------------------------------------------------------------------------------------
// frame scan: header parse, block walk (or index-frame load), piece/chain tables.  One CTA.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_le32_dev(const uint8_t* p) { return ld32u(p); }

// Validates the B2IX index frame that may follow the LZ4 frame and returns its payload pointer.
__device__ const uint8_t* find_index(const uint8_t* src, uint64_t n, uint64_t* payload_bytes) {
    if (n < 16 + sizeof(IndexHead) + HEADER_MIN) return nullptr;
    if (ld_le32_dev(src + n - 4) != IDX_MAGIC) return nullptr;
    uint64_t pb = ld_le32_dev(src + n - 8);                // bytes before the 8-byte trailer
    if (pb < sizeof(IndexHead) || pb + 16 + HEADER_MIN > n) return nullptr;
    const uint8_t* skip = src + n - 16 - pb;              // skippable header: magic, size
    if (ld_le32_dev(skip) != IDX_SKIP_MAGIC) return nullptr;
    if (ld_le32_dev(skip + 4) != pb + 8) return nullptr;
    const uint8_t* payload = skip + 8;
    if (ld_le32_dev(payload) != IDX_MAGIC || ld_le32_dev(payload + 4) != 1u) return nullptr;
    *payload_bytes = pb;
    return payload;
}

// Serial walk over the block header words of a frame (lz4frame.c:1470-1500) by ONE thread: fills the block / piece /
// chain tables and the frame summary `s`.  When `progress` is given (chain-per-block layout only) the number of chains
// published so far is stored there after every block, behind a fence, so that decoding warps of the same kernel can
// start on block b while the walk is still hopping through the headers that follow it.
// Dataflow variant of the walk: only the block table and the group numbering (decode_df.cu finds the pieces).
__device__ B2_NOINLINE void walk_blocks_df(const uint8_t* __restrict__ src, uint64_t n, uint32_t mode, FrameIndex& s,
                                           BlockRef* blocks, uint32_t cap_blocks, const DfWork& W)
{
    const uint32_t bc = s.desc.block_checksum ? 4u : 0u;
    const uint32_t cc = s.desc.content_checksum ? 4u : 0u;
    uint64_t ip = s.desc.header_size;
    uint32_t nb = 0, groups = 0, tickets = 0;
    Err err = OK;
    for (;;) {
        if (n - ip < 4) { err = E_noEndMark; break; }
        const uint32_t w = ld_le32_dev(src + ip); ip += 4;
        const uint32_t cs = w & 0x7FFFFFFFu;
        if (cs == 0) break;                                      // EndMark
        if (mode & SCAN_LEGACY) {
            if ((w >> 31) || cs > LEGACY_BLOCK + LEGACY_BLOCK / 255u + 16u) { err = E_maxBlockSize_invalid; break; }
        } else if (cs > s.block_size) { err = E_maxBlockSize_invalid; break; }
        if (n - ip < (uint64_t)cs + bc) { err = E_noEndMark; break; }
        const uint32_t stored = w >> 31;
        if (nb < cap_blocks && W.cap_groups) {
            BlockRef br; br.src_off = ip; br.csize = cs; br.stored = stored; br.dst_off = 0; br.usize = 0; br.pad = 0;
            blocks[nb] = br;
            W.grp_base[nb] = groups;
        }
        if (stored) tickets += 1u;
        else { groups += (cs + DF_G - 1u) / DF_G; tickets += cs / (3u * DF_BATCH) + 2u; }   // a sequence takes at least 3 bytes
        ++nb;
        ip += (uint64_t)cs + bc;
    }
    if (err == OK && cc) {
        if (n - ip < 4) err = E_noEndMark;
        else { s.stored_content_checksum = ld_le32_dev(src + ip); ip += 4; }
    }
    s.need_blocks = nb; s.need_groups = groups; s.need_tickets = tickets; s.need_pieces = 0;
    if (err == OK && (nb > cap_blocks || !W.cap_groups || groups > W.cap_groups || tickets > W.cap_tickets)) err = E_allocation_failed;
    if (err == OK) W.grp_base[nb] = groups;
    s.err = err;
    s.n_blocks = nb; s.n_pieces = 0; s.n_chains = 0; s.n_stored = 0;
    s.consumed = ip; s.consumed_with_index = ip;
    s.total_out = (uint64_t)nb * s.block_size;                   // upper bound until df_layout_kernel knows
    s.layout_known = 0; s.walk_pending = 0; s.df = 1;
    DfState st; st.bad_key = 0xFFFFFFFFFFFFFFFFull; st.produced = 0; st.total_out = 0; st.n_tickets = 0; st.n_units = 0;
    st.ticket_links = 0; st.ticket_units = 0; st.ticket_tk = 0; st.pad = 0;
    for (int i = 0; i < 12; ++i) st.prof[i] = 0;
    *W.st = st;
}

// Companion of the decoder-side walk: the walk is a chain of dependent loads, one block header per hop, each a DRAM round trip
// (0.8 us: 13 ms for the 16 K blocks of a GiB of 64 KB blocks, more than decoding them takes).  A second warp keeps the next
// megabyte of the stream in L2 ahead of the walker — every line of it, as a header can sit anywhere; the decoding warps read
// the same lines a moment later — so that a hop costs an L2 hit.  w[1] = walk done, w[4..5] = the walker's position.
__device__ B2_NOINLINE void walk_prefetch(const uint8_t* __restrict__ src, uint64_t n, volatile uint32_t* w, uint32_t lane)
{
    constexpr uint64_t AHEAD = 1ull << 20, ROUND = 64ull << 10;
    uint64_t pf = 0;
    for (;;) {
        uint32_t done = 0, lo = 0, hi = 0;
        if (lane == 0) { done = w[1]; lo = w[4]; hi = w[5]; }
        done = __shfl_sync(FULL_MASK, done, 0); lo = __shfl_sync(FULL_MASK, lo, 0); hi = __shfl_sync(FULL_MASK, hi, 0);
        if (done) break;
        uint64_t cur = (uint64_t)lo | ((uint64_t)hi << 32);
        if (cur > n) cur = n;                                   // (a torn read of the two halves: harmless)
        const uint64_t lim = cur + AHEAD < n ? cur + AHEAD : n;
        if (pf < cur) pf = cur;
        if (pf >= lim) { __nanosleep(200); continue; }
        const uint64_t stop = pf + ROUND < lim ? pf + ROUND : lim;
        for (uint64_t a = (pf & ~127ull) + 128ull * lane; a < stop; a += 4096ull) prefetch_l2(src + a);
        pf = stop;
    }
}

__device__ B2_NOINLINE void walk_blocks(const uint8_t* __restrict__ src, uint64_t n, uint32_t mode, FrameIndex& s,
                            BlockRef* blocks, uint32_t cap_blocks, Piece* pieces, uint32_t cap_pieces, Chain* chains,
                            uint32_t* stored_list, volatile uint32_t* progress)
{
    const uint32_t bc = s.desc.block_checksum ? 4u : 0u;
    const uint32_t cc = s.desc.content_checksum ? 4u : 0u;
    const bool per_block = s.desc.block_independent && !(mode & SCAN_SINGLE_CHAIN);
    uint64_t ip = s.desc.header_size;
    uint32_t nb = 0, nstored = 0;
    Err err = OK;
    for (;;) {
        if (n - ip < 4) { err = E_noEndMark; break; }
        uint32_t w = ld_le32_dev(src + ip); ip += 4;
        uint32_t cs = w & 0x7FFFFFFFu;
        if (cs == 0) break;                                      // EndMark
        if (mode & SCAN_LEGACY) {                                // no stored flag; a block may be as large as LZ4_compressBound(8 MB)
            if ((w >> 31) || cs > LEGACY_BLOCK + LEGACY_BLOCK / 255u + 16u) { err = E_maxBlockSize_invalid; break; }
        } else
        if (cs > s.block_size) { err = E_maxBlockSize_invalid; break; }
        if (n - ip < (uint64_t)cs + bc) { err = E_noEndMark; break; }
        if (nb < cap_blocks && nb < cap_pieces) {
            if (progress) {
                // decoder-side walk: one 16-byte raw entry per block (in the chain slot); the warp that decodes the block
                // expands it into the BlockRef / Piece / Chain the later kernels read (expand_walk_entry)
                *(uint4*)(chains + nb) = make_uint4((uint32_t)ip, (uint32_t)(ip >> 32), cs, w >> 31);
                if (w >> 31) stored_list[nstored++] = nb;
                if ((nb & 7u) == 7u) { __threadfence(); *progress = nb + 1u; }      // publish in groups: a fence per block would cost more than the hop
            } else {
                BlockRef br; br.src_off = ip; br.csize = cs; br.stored = w >> 31;
                br.dst_off = (uint64_t)nb * s.block_size; br.usize = 0; br.pad = 0;
                blocks[nb] = br;
                Piece pc; pc.src_off = ip; pc.src_len = cs;
                pc.flags = PIECE_ENDS_BLOCK | (br.stored ? PIECE_STORED : 0u);
                pc.dst_len = 0; pc.block = nb; pc.dst_rel = 0; pc.pad = 0;
                if (per_block) {
                    Chain ch; ch.dst_off = br.dst_off; ch.first_piece = nb; ch.n_pieces = 1; chains[nb] = ch;
                    if (br.stored) { pc.flags |= PIECE_BULK; stored_list[nstored++] = nb; }
                }
                pieces[nb] = pc;
            }
        }
        ++nb;
        ip += (uint64_t)cs + bc;
        if (progress) { progress[4] = (uint32_t)ip; progress[5] = (uint32_t)(ip >> 32); }   // where the prefetching warp should be ahead of
    }
    if (progress) { __threadfence(); *progress = nb < cap_blocks ? nb : cap_blocks; }
    if (err == OK && cc) {
        if (n - ip < 4) err = E_noEndMark;
        else { s.stored_content_checksum = ld_le32_dev(src + ip); ip += 4; }
    }
    s.need_blocks = nb; s.need_pieces = nb;
    if (err == OK && (nb > cap_blocks || nb > cap_pieces)) err = E_allocation_failed;
    s.err = err;
    s.n_blocks = nb; s.n_pieces = nb; s.n_stored = nstored;
    s.consumed = ip; s.consumed_with_index = ip;
    s.total_out = (uint64_t)nb * s.block_size;                   // upper bound
    if (per_block) { s.n_chains = nb; s.layout_known = 0; }
    else {
        s.n_chains = nb ? 1u : 0u; s.layout_known = 0;
        if (nb && err == OK) { Chain ch; ch.dst_off = 0; ch.first_piece = 0; ch.n_pieces = nb; chains[0] = ch; }
    }
}

__global__ void __launch_bounds__(256)
frame_scan_kernel(const uint8_t* __restrict__ src, uint64_t n, uint64_t dst_cap, uint32_t mode,
                  FrameIndex* __restrict__ fi, BlockRef* __restrict__ blocks, uint32_t cap_blocks,
                  Piece* __restrict__ pieces, uint32_t cap_pieces, Chain* __restrict__ chains,
                  uint32_t* __restrict__ stored_list, uint32_t* __restrict__ walk, DfWork W)
{
    B2_SHARED_VAR(FrameIndex, s);
    B2_SHARED_VAR(uint32_t, s_nstored);
    B2_SHARED_VAR(const uint8_t*, s_idx);
    B2_SHARED_VAR(uint32_t, s_bad);
    B2_SHARED_ARRAY(uint64_t, part_src, 256);
    B2_SHARED_ARRAY(uint64_t, part_dst, 256);
    B2_SHARED_ARRAY(uint32_t, part_cnt, 256);
    const uint32_t tid = threadIdx.x;

    if (tid == 0) {
        s = FrameIndex{};
        s_idx = nullptr; s_bad = 0; s_nstored = 0;
        uint8_t hdr[HEADER_MAX];
        uint32_t hn = n < (uint64_t)HEADER_MAX ? (uint32_t)n : (uint32_t)HEADER_MAX;
        for (uint32_t i = 0; i < hn; ++i) hdr[i] = src[i];
        s.err = header_parse(hdr, hn, s.desc);
        if (s.err == OK) {
            s.block_size = (mode & SCAN_LEGACY) ? LEGACY_BLOCK : block_size_of(s.desc.block_size_id);
            if (mode & SCAN_USE_INDEX) {
                uint64_t pb = 0;
                const uint8_t* ix = find_index(src, n, &pb);
                if (ix) {
                    IndexHead h;
                    for (uint32_t i = 0; i < sizeof(IndexHead) / 4; ++i) ((uint32_t*)&h)[i] = ld_le32_dev(ix + 4 * i);
                    uint64_t want = sizeof(IndexHead) + (uint64_t)h.n_blocks * sizeof(IndexBlock)
                                  + (uint64_t)h.n_chains * sizeof(IndexChain);
                    if (want == pb && h.block_size == s.block_size && s.desc.block_independent &&
                        h.n_chains >= h.n_blocks) {
                        s_idx = ix;
                        s.n_blocks = h.n_blocks; s.n_pieces = h.n_chains; s.n_chains = h.n_chains;
                        s.total_out = h.content_size;
                        s.consumed_with_index = n;
                        s.consumed = n - 16 - pb;
                    }
                }
            }
        }
    }
    __syncthreads();
    if (s.err != OK) { if (tid == 0) *fi = s; return; }

    const uint32_t bc = s.desc.block_checksum ? 4u : 0u;
    const uint32_t cc = s.desc.content_checksum ? 4u : 0u;

    if (s_idx) {
        // ---- indexed layout: everything is a prefix sum over the index entries (parallel).
        const uint8_t* ix = s_idx;
        const uint32_t nb = s.n_blocks, nc = s.n_chains;
        if (nb > cap_blocks || nc > cap_pieces) {
            if (tid == 0) { s.err = E_allocation_failed; s.need_blocks = nb; s.need_pieces = nc; *fi = s; }
            return;
        }
        const uint8_t* ib = ix + sizeof(IndexHead);
        const uint8_t* ic = ib + (uint64_t)nb * sizeof(IndexBlock);
        // serial-per-thread ranges + smem scan of the partial sums (blocks are few: <= U / 64 KB)
        uint32_t per = (nb + 255u) / 256u;
        uint32_t b0 = tid * per, b1 = b0 + per < nb ? b0 + per : nb;
        uint64_t a = 0, d = 0;
        uint32_t cnt = 0;
        for (uint32_t b = b0; b < b1; ++b) {
            uint32_t w = ld_le32_dev(ib + (uint64_t)b * sizeof(IndexBlock));
            a += 4ull + (w & 0x7FFFFFFFu) + bc;
            d += ld_le32_dev(ib + (uint64_t)b * sizeof(IndexBlock) + 4);
            cnt += ld_le32_dev(ib + (uint64_t)b * sizeof(IndexBlock) + 12);
        }
        part_src[tid] = a; part_dst[tid] = d; part_cnt[tid] = cnt;
        __syncthreads();
        if (tid == 0) {
            uint64_t ra = s.desc.header_size, rd = 0;
            uint32_t rc = 0;
            for (int i = 0; i < 256; ++i) {
                uint64_t ta = part_src[i], td = part_dst[i]; uint32_t tc = part_cnt[i];
                part_src[i] = ra; part_dst[i] = rd; part_cnt[i] = rc; ra += ta; rd += td; rc += tc;
            }
            // frame must end exactly where the index says: EndMark + optional checksum, then the index frame
            if (ra + 4 + cc != s.consumed || rd != s.total_out || rc != nc) s_bad = 1;
            else if (ld_le32_dev(src + ra) != 0u) s_bad = 1;
            else if (cc) s.stored_content_checksum = ld_le32_dev(src + ra + 4);
        }
        __syncthreads();
        a = part_src[tid]; d = part_dst[tid]; cnt = part_cnt[tid];
        for (uint32_t b = b0; b < b1 && !s_bad; ++b) {
            const uint8_t* e = ib + (uint64_t)b * sizeof(IndexBlock);
            uint32_t w = ld_le32_dev(e), us = ld_le32_dev(e + 4), fc = ld_le32_dev(e + 8), ncb = ld_le32_dev(e + 12);
            uint32_t cs = w & 0x7FFFFFFFu;
            // the index is only a hint: the real frame bytes decide
            if (ld_le32_dev(src + a) != w || cs == 0 || cs > s.block_size || us > s.block_size || us == 0 ||
                ncb == 0 || fc != cnt || (uint64_t)fc + ncb > nc || ((w >> 31) && (ncb != 1 || cs != us))) { s_bad = 1; break; }
            BlockRef br; br.src_off = a + 4; br.csize = cs; br.stored = w >> 31; br.dst_off = d; br.usize = us; br.pad = 0;
            blocks[b] = br;
            bool bad = false;                                       // (no exit from inside the loop: a bad entry is reported behind it)
            // the entries are read eight at a time, all loads first: one memory round trip per eight entries instead of one each
            // (this kernel is one CTA in front of the decoder: 126 K cycles for 27 K instructions before)
            constexpr uint32_t EB = 8;
            uint32_t etok[EB + 1], eout[EB + 1];
            for (uint32_t k0 = 0; k0 < ncb; k0 += EB) {
              #pragma unroll
              for (uint32_t j = 0; j <= EB; ++j) {
                  const uint32_t k = k0 + j;
                  const uint8_t* ce = ic + (uint64_t)(fc + k) * sizeof(IndexChain);
                  etok[j] = k < ncb ? ld_le32_dev(ce) : cs; eout[j] = k < ncb ? ld_le32_dev(ce + 4) : us;
              }
              #pragma unroll
              for (uint32_t j = 0; j < EB; ++j) {
                const uint32_t k = k0 + j;
                if (k >= ncb) break;
                const uint32_t tok = etok[j], out = eout[j], tok_n = etok[j + 1], out_n = eout[j + 1];
                if ((k == 0 && (tok != 0 || out != 0)) || tok >= tok_n || out > out_n || tok_n > cs || out_n > us) { bad = true; continue; }
                Piece pc; pc.src_off = a + 4 + tok; pc.src_len = tok_n - tok;
                pc.flags = PIECE_EXACT | (br.stored ? (PIECE_STORED | PIECE_BULK) : 0u) | (k + 1 == ncb ? PIECE_ENDS_BLOCK : 0u);
                if (br.stored) stored_list[atomicAdd(&s_nstored, 1u)] = fc + k;     // copied by copy_stored_kernel
                pc.dst_len = out_n - out; pc.block = b; pc.dst_rel = out; pc.pad = 0;
                pieces[fc + k] = pc;
                Chain ch; ch.dst_off = d + out; ch.first_piece = fc + k; ch.n_pieces = 1;
                chains[fc + k] = ch;
              }
            }
            if (bad) { s_bad = 1; break; }
            a += 4ull + cs + bc; d += us; cnt += ncb;
        }
        __syncthreads();
        if (!s_bad) {
            if (tid == 0) {
                s.indexed = 1; s.layout_known = 1; s.n_stored = s_nstored;
                if (s.total_out > dst_cap) s.err = E_dstMaxSize_tooSmall;
                if (s.desc.has_content_size && s.desc.content_size != s.total_out) s.err = E_frameSize_wrong;
                *fi = s;
            }
            return;
        }
        // fall through to the serial walk with the index ignored
        __syncthreads();
        if (tid == 0) { s.n_blocks = s.n_pieces = s.n_chains = 0; s.total_out = 0; s.consumed = s.consumed_with_index = 0; s_nstored = 0; }
        __syncthreads();
    }

    if (tid == 0) {
        // no (usable) index.  Dataflow decode pays where a block is long (few blocks, each one dependent chain for a lone
        // warp): linked frames, 4 MB blocks, legacy frames; many small independent blocks are faster one warp per block
        const bool few = (mode & SCAN_DF_FEW) && dst_cap / s.block_size <= (uint64_t)DF_FEW_BLOCKS;
        if ((mode & SCAN_DF) && ((mode & (SCAN_DF_ALL | SCAN_LEGACY)) || !s.desc.block_independent || s.block_size >= (4u << 20) || few)) {
            walk_blocks_df(src, n, mode, s, blocks, cap_blocks, W);
            *fi = s;
            return;
        }
        if ((mode & SCAN_OVERLAP_WALK) && s.desc.block_independent && !(mode & SCAN_SINGLE_CHAIN)) {
            // the walk is left to the decoder kernel, which runs it next to the decoding warps (walk_blocks below)
            s.walk_pending = 1;
            walk[0] = 0; walk[1] = 0; walk[2] = 0; walk[4] = s.desc.header_size; walk[5] = 0;
            *fi = s;
            return;
        }
        walk_blocks(src, n, mode, s, blocks, cap_blocks, pieces, cap_pieces, chains, stored_list, nullptr);
        *fi = s;
    }
}

}  // namespace b2
#include "decode_piece.cuh"
#include "xxh32_cta.cuh"
namespace b2 {

#ifndef B2_WALK_NAP_NS
#define B2_WALK_NAP_NS 300      // waiting for the decoder-side walk: nanoseconds of sleep per block of distance
#endif
#ifndef B2_DEC_MINCTAS
#define B2_DEC_MINCTAS 4      // 8 warps x 4 CTAs: 64 registers per thread, 32 warps per SM
#endif
__global__ void __launch_bounds__(DEC_WARPS * 32, B2_DEC_MINCTAS)
decode_chains_kernel(const uint8_t* __restrict__ src, uint64_t n_src, uint8_t* __restrict__ dst, uint64_t dst_cap, uint32_t mode,
                     FrameIndex* fi, BlockRef* blocks, Piece* pieces, Chain* chains, uint32_t cap_tables,
                     uint32_t* stored_list, PieceOut* __restrict__ outs, uint32_t* __restrict__ ticket, uint32_t* walk,
                     uint32_t hist)
{
    // hist = bytes of earlier output of the same linked-block frame that lie directly in front of dst
    // (streaming decode in batches: lz4frame.c:1286-1349 keeps the same 64 KB)
    if (fi->err != OK || fi->df) return;
    const uint32_t lane = lane_id();
    B2_SHARED_ARRAY(uint8_t, s_walk, DEC_WARPS * (1568 + TILE));           // per warp: window (272) + 5 link tables (256 each) [+ output tile]
    uint8_t* const my_win = s_walk + (threadIdx.x >> 5) * 1568u;
    uint8_t* const my_tile = s_walk + DEC_WARPS * 1568u + (threadIdx.x >> 5) * TILE;
    for (uint32_t i = lane; i < 5u * 128u; i += 32u) {                       // upper halves of the tables are identity for good
        const uint32_t t = i >> 7, x = 128u + (i & 127u);
        my_win[272u + 256u * t + x] = (uint8_t)x;
    }
    __syncwarp();
    // walk_pending: the block headers have not been walked yet (plain frame, no index).  Warp 0 of CTA 0 does it now, with
    // one thread, publishing every block as it is found; all other warps start decoding published blocks at once, so the
    // serial hops through the headers (one DRAM round trip each) hide behind the decoding instead of preceding it.
    // The walker is whichever warp draws walk[2] first — a warp that is running, whatever order the CTAs of the grid
    // are dispatched in — and it never waits for anybody: no deadlock.
    const bool pending = fi->walk_pending != 0;
    const uint32_t block_size = fi->block_size;
    const bool linked = !fi->desc.block_independent;
    uint32_t role = 0xFFFFFFFFu;                                     // 0 = walker, 1 = its prefetching companion
    if (pending) { if (lane == 0) role = atomicAdd(walk + 2, 1u); role = __shfl_sync(FULL_MASK, role, 0); }
    const uint32_t walker = role == 0u ? 1u : 0u;
    if (role == 1u) walk_prefetch(src, n_src, (volatile uint32_t*)walk, lane);
    if (walker) {
        if (lane == 0) {
            FrameIndex s = *fi;
            walk_blocks(src, n_src, mode, s, blocks, cap_tables, pieces, cap_tables, chains, stored_list, walk);
            s.walk_pending = 2;                                       // walked inside the decoder
            *fi = s;
            __threadfence();
            ((volatile uint32_t*)walk)[1] = 1u;                       // done: walk[0] is final
        }
        __syncwarp();
    }
    const uint8_t* const src_end = pending ? src + n_src : src + fi->consumed_with_index;
    const uint32_t n_chains = pending ? 0u : fi->n_chains;
    for (;;) {
        uint32_t c = 0;
        if (lane == 0) c = atomicAdd(ticket, 1u);
        c = __shfl_sync(FULL_MASK, c, 0);
        if (pending) {
            uint32_t ok = 0;
            if (lane == 0) {
                volatile uint32_t* w = (volatile uint32_t*)walk;
                for (;;) {
                    const uint32_t p = w[0];
                    if (c < p) { ok = 1; break; }
                    if (w[1]) { ok = c < w[0] ? 1u : 0u; break; }
                    // thousands of warps wait here when the blocks are small: a warp far from its turn asks rarely (the walker
                    // needs ~0.7 us per block), so that the polls do not crowd the walker's own loads out of the L2 slice
                    const uint32_t d = c - p;
                    __nanosleep(d < 8u ? 300u : d < 160u ? d * B2_WALK_NAP_NS : 160u * B2_WALK_NAP_NS);
                }
                __threadfence();
            }
            ok = __shfl_sync(FULL_MASK, ok, 0);
            if (!ok) break;
        } else if (c >= n_chains) break;
        Chain ch;
        if (pending) {
            // raw walk entry {src_off, csize, stored} -> the tables the later kernels (and the loop below) read
            const uint4 e = __ldcg((const uint4*)(chains + c));
            BlockRef br; br.src_off = (uint64_t)e.x | ((uint64_t)e.y << 32); br.csize = e.z; br.stored = e.w;
            br.dst_off = (uint64_t)c * block_size; br.usize = 0; br.pad = 0;
            Piece pc; pc.src_off = br.src_off; pc.src_len = e.z;
            pc.flags = PIECE_ENDS_BLOCK | (e.w ? (PIECE_STORED | PIECE_BULK) : 0u);
            pc.dst_len = 0; pc.block = c; pc.dst_rel = 0; pc.pad = 0;
            ch.dst_off = br.dst_off; ch.first_piece = c; ch.n_pieces = 1;
            if (lane == 0) { blocks[c] = br; pieces[c] = pc; chains[c] = ch; }
            __syncwarp();
        } else ch = ld_cg(chains + c);
        uint64_t op = ch.dst_off;
        const uint64_t chain_start = ch.dst_off;
        bool dead = false;
        for (uint32_t k = 0; k < ch.n_pieces; ++k) {
            const uint32_t pi = ch.first_piece + k;
            const Piece pc = ld_cg(pieces + pi);
            PieceOut po; po.out_len = 0; po.err = 0;
            if (dead) { po.err = E_decompressionFailed; if (lane == 0) outs[pi] = po; continue; }
            const BlockRef br = ld_cg(blocks + pc.block);
            // output window of the owning block (the reference hands maxBlockSize as capacity, lz4frame.c:1670)
            uint64_t blk_out = (pc.flags & PIECE_EXACT) ? br.dst_off : op;
            uint64_t oend = blk_out + block_size;
            if (blk_out >= dst_cap) oend = blk_out; else if (oend > dst_cap) oend = dst_cap;
            if (pc.flags & PIECE_STORED) {
                if (op + pc.src_len > dst_cap) po.err = E_dstMaxSize_tooSmall;
                else {
                    if (!(pc.flags & PIECE_BULK)) group_copy_small(dst + op, src + pc.src_off, pc.src_len, lane, 32);   // bulk ones: copy_stored_kernel
                    po.out_len = pc.src_len; op += pc.src_len;
                }
            } else {
                Cursor cu;
                cu.sp = src + br.src_off; cu.dp = dst + blk_out; cu.win = my_win; cu.tile = my_tile; cu.src_end = src_end;
                cu.ip = (uint32_t)(pc.src_off - br.src_off); cu.iend = cu.ip + pc.src_len;
                cu.iend_blk = br.csize;
                const uint32_t op_rel = (uint32_t)(op - blk_out);
                const uint32_t ocap = (uint32_t)(oend - blk_out);
                cu.op = op_rel; cu.oend = ocap;
                cu.opiece_end = (pc.flags & PIECE_EXACT) ? op_rel + pc.dst_len : ocap;
                if (cu.opiece_end > ocap && !(pc.flags & PIECE_EXACT)) cu.opiece_end = ocap;
                // window start: independent blocks/sub-blocks may only reference their own output;
                // linked blocks reach up to 64 KB into the earlier output of the chain (lz4frame.c:1286-1349)
                if (linked) { uint64_t h = blk_out - chain_start + hist; cu.reach = h > 0x40000000ull ? 0x40000000u : (uint32_t)h; }
                else cu.reach = 0u - op_rel;
                uint32_t e;
                for (;;) {
                    e = decode_piece<false>(cu, (pc.flags & PIECE_ENDS_BLOCK) != 0);
                    if (e != DEC_PAUSED) break;
                    __syncwarp();
                    warp_match_overlap(cu.dp + cu.p_dst, cu.p_off, cu.p_ml, lane);
                    __syncwarp();
                    if (cu.p_done) { e = 0; break; }
                }
                const uint32_t produced = cu.op - op_rel;
                if ((pc.flags & PIECE_EXACT) && e == 0 && produced != pc.dst_len) e = E_decompressionFailed;
                if (e && ocap < block_size) e = E_dstMaxSize_tooSmall;          // ran out of destination
                po.err = e; po.out_len = produced;
                op += produced;
            }
            if (po.err) dead = true;
            if (lane == 0) outs[pi] = po;
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------------------------------------
// stored blocks of independent-block frames (lz4frame.c:1546-1580 copies them with memcpy): plain wide copies,
// every CTA takes 64 KB slices of the blocks the scan listed.
// ------------------------------------------------------------------------------------------------
#ifndef B2_STORED_SLICE_LOG2
#define B2_STORED_SLICE_LOG2 16
#endif
__global__ void __launch_bounds__(256)
copy_stored_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint64_t dst_cap,
                   const FrameIndex* __restrict__ fi, const Piece* __restrict__ pieces,
                   const Chain* __restrict__ chains, const uint32_t* __restrict__ stored_list)
{
    if (fi->err != OK) return;
    const uint32_t ns = fi->n_stored;
    if (ns == 0) return;
    constexpr uint32_t SLICE = 1u << B2_STORED_SLICE_LOG2;
    const uint32_t per = (fi->block_size + SLICE - 1u) / SLICE;
    const uint64_t total = (uint64_t)ns * per;
    // the three dependent look-ups in front of a slice (list -> chain -> piece) cost about as much as a fifth of its copy:
    // those of the NEXT slice are made before this one is copied
    uint64_t it = blockIdx.x;
    if (it >= total) return;
    uint32_t k = (uint32_t)(it / per);
    Chain ch = chains[stored_list[k]];
    Piece pc = pieces[ch.first_piece];
    for (;;) {
        const uint32_t sl = (uint32_t)(it - (uint64_t)k * per);
        const uint64_t nit = it + gridDim.x;
        Chain nch = ch; Piece npc = pc; uint32_t nk = k;
        if (nit < total) {
            nk = (uint32_t)(nit / per);
            if (nk != k) { nch = chains[stored_list[nk]]; npc = pieces[nch.first_piece]; }
        }
        const uint32_t off = sl * SLICE;
        if (off < pc.src_len && ch.dst_off + pc.src_len <= dst_cap) {
            const uint32_t len = pc.src_len - off < SLICE ? pc.src_len - off : SLICE;
            group_copy(dst + ch.dst_off + off, src + pc.src_off + off, len, threadIdx.x, blockDim.x);
        }
        if (nit >= total) break;
        it = nit; k = nk; ch = nch; pc = npc;
    }
}

// ------------------------------------------------------------------------------------------------
// result reduction: first error in block order, total size, layout check.  One CTA.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
decode_finalize_kernel(const FrameIndex* __restrict__ fi, const Piece* __restrict__ pieces,
                       const PieceOut* __restrict__ outs, const uint32_t* __restrict__ block_crc_bad,
                       DecodeResult* __restrict__ res, uint32_t* __restrict__ ticket)
{
    B2_SHARED_VAR(uint32_t, s_first_bad);
    B2_SHARED_VAR(uint32_t, s_short);
    B2_SHARED_VAR(unsigned long long, s_total);
    const uint32_t tid = threadIdx.x;
    if (tid == 0) { s_first_bad = 0xFFFFFFFFu; s_short = 0xFFFFFFFFu; s_total = 0; *ticket = 0; }
    __syncthreads();
    if (fi->df) return;                                          // df_finalize_kernel reports
    if (fi->err != OK) {
        if (tid == 0) { DecodeResult r{}; r.err = fi->err; r.layout_ok = 1; *res = r; }
        return;
    }
    const uint32_t np = fi->n_pieces, nb = fi->n_blocks, bs = fi->block_size;
    unsigned long long tot = 0;
    for (uint32_t i0 = tid; i0 < np; i0 += 8u * blockDim.x) {       // eight loads in flight per thread (one CTA: latency is all there is)
        PieceOut po[8];
        #pragma unroll
        for (uint32_t j = 0; j < 8u; ++j) { const uint32_t i = i0 + j * blockDim.x; po[j] = i < np ? outs[i] : PieceOut{0, 0}; }
        #pragma unroll
        for (uint32_t j = 0; j < 8u; ++j) {
            const uint32_t i = i0 + j * blockDim.x;
            if (i < np && po[j].err) atomicMin(&s_first_bad, i);
            tot += po[j].out_len;
        }
    }
    // a non-final block that decodes short breaks the "block b starts at b*block_size" layout
    if (!fi->layout_known && fi->n_chains == nb) {
        for (uint32_t i = tid; i + 1 < nb; i += blockDim.x)
            if (outs[i].err == 0 && outs[i].out_len != bs) atomicMin(&s_short, i);
    }
    atomicAdd(&s_total, tot);
    __syncthreads();
    if (tid == 0) {
        DecodeResult r{};
        r.layout_ok = (s_short == 0xFFFFFFFFu);
        r.total_out = s_total;
        uint32_t crc_bad = block_crc_bad ? *block_crc_bad : 0xFFFFFFFFu;
        uint32_t bad_piece_block = s_first_bad == 0xFFFFFFFFu ? 0xFFFFFFFFu : pieces[s_first_bad].block;
        if (crc_bad != 0xFFFFFFFFu && crc_bad <= bad_piece_block) { r.err = E_blockChecksum_invalid; r.err_block = crc_bad; }
        else if (bad_piece_block != 0xFFFFFFFFu) { r.err = outs[s_first_bad].err; r.err_block = bad_piece_block; }
        if (r.err == OK && fi->desc.has_content_size && r.layout_ok && fi->desc.content_size != r.total_out)
            r.err = E_frameSize_wrong;                                      // lz4frame.c:1715-1716
        *res = r;
    }
}

// ------------------------------------------------------------------------------------------------
// block checksum verification (lz4frame.c:1604-1611): XXH32 over each block's stored bytes.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
verify_block_checksums_kernel(const uint8_t* __restrict__ src, const FrameIndex* __restrict__ fi,
                              const BlockRef* __restrict__ blocks, uint32_t* __restrict__ first_bad)
{
    if (fi->err != OK || !fi->desc.block_checksum) return;
    // A lone thread per block read 1.7 MB byte-strided at config 2 and was most of the call (8.6 GB/s for the whole decode).
    // Large blocks: one CTA per block — one hashing warp fed from shared-memory tiles the other warps prefetch (xxh32_cta.cuh,
    // 2.4 GB/s per block, all blocks at once).  Small blocks: one warp per block, coalesced loads (xxh32_warp).
    const uint32_t nb = fi->n_blocks;
    if (fi->block_size >= (256u << 10)) {
        B2_SHARED_VAR(XxhCtaShared, S);
        for (uint32_t b = blockIdx.x; b < nb; b += gridDim.x) {
            const BlockRef br = blocks[b];
            const uint32_t h = xxh32_cta(S, src + br.src_off, br.csize);
            if (threadIdx.x == 0 && h != ld32u(src + br.src_off + br.csize)) atomicMin(first_bad, b);
        }
        return;
    }
    const uint32_t lane = lane_id();
    uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t stride = (gridDim.x * blockDim.x) >> 5;
    for (; b < nb; b += stride) {
        const BlockRef br = blocks[b];
        const uint32_t h = xxh32_warp(src + br.src_off, br.csize, lane);
        if (lane == 0 && h != ld32u(src + br.src_off + br.csize)) atomicMin(first_bad, b);
    }
}

}  // namespace b2
