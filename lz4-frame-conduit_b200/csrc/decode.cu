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
__device__ __noinline__ void walk_blocks(const uint8_t* __restrict__ src, uint64_t n, uint32_t mode, FrameIndex& s,
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
                  uint32_t* __restrict__ stored_list, uint32_t* __restrict__ walk)
{
    __shared__ FrameIndex s;
    __shared__ uint32_t s_nstored;
    __shared__ const uint8_t* s_idx;
    __shared__ uint32_t s_bad;
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
        __shared__ uint64_t part_src[256], part_dst[256];
        __shared__ uint32_t part_cnt[256];
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
            for (uint32_t k = 0; k < ncb; ++k) {
                const uint8_t* ce = ic + (uint64_t)(fc + k) * sizeof(IndexChain);
                uint32_t tok = ld_le32_dev(ce), out = ld_le32_dev(ce + 4);
                uint32_t tok_n = cs, out_n = us;
                if (k + 1 < ncb) { tok_n = ld_le32_dev(ce + 8); out_n = ld_le32_dev(ce + 12); }
                if ((k == 0 && (tok != 0 || out != 0)) || tok >= tok_n || out > out_n || tok_n > cs || out_n > us) { s_bad = 1; break; }
                Piece pc; pc.src_off = a + 4 + tok; pc.src_len = tok_n - tok;
                pc.flags = PIECE_EXACT | (br.stored ? (PIECE_STORED | PIECE_BULK) : 0u) | (k + 1 == ncb ? PIECE_ENDS_BLOCK : 0u);
                if (br.stored) stored_list[atomicAdd(&s_nstored, 1u)] = fc + k;     // copied by copy_stored_kernel
                pc.dst_len = out_n - out; pc.block = b; pc.dst_rel = out; pc.pad = 0;
                pieces[fc + k] = pc;
                Chain ch; ch.dst_off = d + out; ch.first_piece = fc + k; ch.n_pieces = 1;
                chains[fc + k] = ch;
            }
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
        if ((mode & SCAN_OVERLAP_WALK) && s.desc.block_independent && !(mode & SCAN_SINGLE_CHAIN)) {
            // the walk is left to the decoder kernel, which runs it next to the decoding warps (walk_blocks below)
            s.walk_pending = 1;
            walk[0] = 0; walk[1] = 0; walk[2] = 0;
            *fi = s;
            return;
        }
        walk_blocks(src, n, mode, s, blocks, cap_blocks, pieces, cap_pieces, chains, stored_list, nullptr);
        *fi = s;
    }
}

// ------------------------------------------------------------------------------------------------
// sequence decoder: one warp per chain, pieces in order.
// All lanes keep identical parse state (uniform control flow); literal and match copies are spread
// over the 32 lanes.  Acceptance rules mirror the reference decoder's safe loop (lz4.c:1844-2055).
// ------------------------------------------------------------------------------------------------
struct Cursor {            // what the warp knows about the piece it is decoding (all offsets 32-bit, block-relative)
    const uint8_t* sp;     // first compressed byte of the owning block
    uint8_t* dp;           // first output byte of the owning block
    uint32_t ip, iend;     // current / end of this piece's compressed bytes
    uint32_t iend_blk;     // end of the owning block's compressed bytes (reference's `iend`)
    uint32_t op, oend;     // current output position / output limit of the owning block (reference's `oend`)
    uint32_t opiece_end;   // exact end of this piece's output (PIECE_EXACT) or == oend
    uint8_t* win;          // this warp's shared memory: 272-byte window of the compressed stream + 5 link tables of 256 bytes
    const uint8_t* src_end;// end of the compressed buffer (words that start at or after it are not read)
    uint32_t p_dst, p_off, p_ml;  // DEC_PAUSED: a huge overlapping match the caller fills (dp + p_dst, offset, length)
    bool p_done;           // ... and whether the piece ended with it
    uint32_t reach;        // a match at output position m may reach back `m + reach` bytes (mod 2^32):
                           // independent piece: -piece_start;  linked: bytes of earlier output in front of the block
};

// Per-lane byte copy of n bytes; loads are batched four at a time ahead of the stores so that a lane
// keeps several loads in flight.  `s` must not overlap the destination (callers guarantee it).
__device__ __forceinline__ void lane_copy(uint8_t* d, const uint8_t* s, uint32_t n) {
    uint32_t k = 0;
    for (; k + 4 <= n; k += 4) {
        uint8_t a = s[k], b = s[k + 1], c = s[k + 2], e = s[k + 3];
        d[k] = a; d[k + 1] = b; d[k + 2] = c; d[k + 3] = e;
    }
    for (; k < n; ++k) d[k] = s[k];
}

// Per-lane match copy: replicates the FINAL window [d - off, d) with period `off` (off >= n: plain copy).
// All loads come from the window, never from bytes this call writes, so they do not wait on the stores.
__device__ __forceinline__ void lane_match(uint8_t* d, uint32_t off, uint32_t n) {
    const uint8_t* s = d - off;
    uint32_t r = 0, k = 0;
    for (; k + 4 <= n; k += 4) {
        uint8_t a = s[r]; r = (r + 1 == off) ? 0u : r + 1;
        uint8_t b = s[r]; r = (r + 1 == off) ? 0u : r + 1;
        uint8_t c = s[r]; r = (r + 1 == off) ? 0u : r + 1;
        uint8_t e = s[r]; r = (r + 1 == off) ? 0u : r + 1;
        d[k] = a; d[k + 1] = b; d[k + 2] = c; d[k + 3] = e;
    }
    for (; k < n; ++k) { d[k] = s[r]; r = (r + 1 == off) ? 0u : r + 1; }
}

// Warp-cooperative match copy (long matches): the source window [d - off, d) is final.
//   off >= ml            plain copy
//   16 % off == 0        (run-length style: 1, 2, 4, 8, 16) every aligned 16-byte chunk of the run holds the same
//                        pattern, built once per lane in registers and streamed out with 16-byte stores
//   other overlaps       the first bytes are replicated directly, then the valid region doubles: each round copies a
//                        whole number of periods from what is already there (non-overlapping, so the wide copy applies)
// (rare paths are kept out of line so that their registers do not weigh on the sequence loop)
__device__ __forceinline__ void warp_match_overlap(uint8_t* d, uint32_t offset, uint64_t ml, uint32_t lane) {
    const uint8_t* m = d - offset;
    if ((16u % offset) == 0u && ml >= 64u) {
        const uint32_t head = (uint32_t)((16u - ((uintptr_t)d & 15u)) & 15u);
        if (lane < head) d[lane] = m[lane % offset];
        uint32_t pw[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t v = 0;
            #pragma unroll
            for (int k = 0; k < 4; ++k) v |= (uint32_t)m[(head + 4u * j + k) % offset] << (8 * k);
            pw[j] = v;
        }
        __syncwarp();
        const uint4 P = make_uint4(pw[0], pw[1], pw[2], pw[3]);
        uint4* d4 = (uint4*)(d + head);
        const uint64_t nq = (ml - head) >> 4;
        #pragma unroll 4
        for (uint64_t q = lane; q < nq; q += 32) d4[q] = P;
        const uint64_t done = head + (nq << 4);
        if (lane < (uint32_t)(ml - done)) d[done + lane] = m[(done + lane) % offset];
        return;
    }
    // direct replication of the first bytes (enough for the doubling rounds to use wide copies soon)
    uint64_t done = ml < 256u ? ml : 256u;
    {
        uint32_t r = lane % offset; const uint32_t step = 32u % offset;
        for (uint32_t i = lane; i < (uint32_t)done; i += 32) { d[i] = m[r]; r += step; if (r >= offset) r -= offset; }
    }
    while (done < ml) {
        __syncwarp();
        const uint64_t L = ((offset + done) / offset) * offset;         // whole periods available behind d + done
        const uint64_t cnt = (ml - done) < L ? (ml - done) : L;
        group_copy(d + done, d + done - L, cnt, lane, 32);
        done += cnt;
    }
}

constexpr uint32_t DEC_PAUSED = 0xFFFFFFFEu;   // decode_piece: stopped in front of a huge overlapping match (see Cursor::p_*)
constexpr uint32_t HUGE_MATCH = 512u;          // overlapping matches at least this long are filled outside the sequence loop
constexpr uint32_t WSTOP = 255u;       // link-table value: the token here needs the general path
#ifndef B2_DEC_PREFETCH
#define B2_DEC_PREFETCH 2      // 128-byte lines of the compressed stream requested ahead at every window refill
#endif
#ifndef B2_DEC_LANE_CAP
#define B2_DEC_LANE_CAP 24
#endif
constexpr uint32_t LANE_CAP = B2_DEC_LANE_CAP;      // copies up to this many bytes run one-lane-per-sequence; longer ones use the whole warp

// Decodes one piece.  The warp first WALKS up to 32 sequences (uniform control flow: every lane follows the
// token chain, lane i keeps the fields of sequence i), then copies them lane-parallel: literals of all 32
// sequences at once, then matches in rounds — a match may run as soon as every byte it reads is final, i.e.
// lies below the destination of the first match still pending (all literals are already in place).
// The walk has a short path for the common sequence (both nibbles < 15, far from every limit) whose only
// check is the offset; everything else takes the general path with the reference's acceptance rules.
// returns 0 or an Err; c.op is advanced
__device__ __noinline__ uint32_t decode_piece(Cursor& c, bool ends_block)
{
    const uint32_t lane = lane_id();
    const uint8_t* __restrict__ sp = c.sp;
    uint8_t* dp = c.dp;
    uint32_t ip = c.ip, op = c.op;
    const uint32_t iend = c.iend_blk, pend = c.iend, oend = c.oend, opend = c.opiece_end, reach = c.reach;
    // short path (lengths with at most one extension byte) allowed while  ip + 288 <= pend  and
    // op + 560 <= min(oend, opend): inside that region none of the general path's limit checks can fire
    uint32_t i_fast = pend >= 288u ? pend - 287u : 0u;
    const uint32_t olim = oend < opend ? oend : opend;
    uint32_t o_fast = olim >= 560u ? olim - 559u : 0u;
    asm volatile("" : "+r"(i_fast), "+r"(o_fast));                          // keep them in registers (no re-derivation per sequence)
    uint32_t err = 0;
    if (pend == ip) return E_decompressionFailed;                           // empty input (lz4.c:1702)
    // ---- lane-parallel token walk (the common case).  128 consecutive bytes of the stream are examined at once:
    // every lane computes, for 4 byte positions, where the next token would be IF a token started there (one
    // shared-memory byte, two with an extension byte).  Squaring that link table four times (T1, T2, T4, T8, T16)
    // lets lane i jump straight to the i-th token after the known entry point, so 32 sequences are located in a
    // handful of instructions instead of a 32-step dependent chain of loads.  Anything unusual (a second extension
    // byte, a literal run that leaves the window, the last 288 input / 560 output bytes of the piece) is marked
    // WSTOP and goes through the general path below, which carries every acceptance rule of the reference.
    uint8_t* const win = c.win;
    uint8_t* const T1 = win + 272, * const T2 = T1 + 256, * const T4 = T2 + 256, * const T8 = T4 + 256, * const T16 = T8 + 256;
    int32_t wb = 0; bool win_ok = false;                                    // window = stream bytes [wb, wb + 256)
    bool finished = false, paused = false;
    while (!finished) {
        uint32_t my_a = 0, my_b = 0, my_lit = 0, my_ll = 0, my_ml = 0, my_off = 1;
        bool my_packed = true, fast_batch = false;
        uint32_t my_op_fast = 0;
        const uint32_t op0 = op;
        uint32_t n = 0;
        if (ip < i_fast && op < o_fast) {
            // gather up to 32 sequences, refilling the window as often as needed: lanes n.. take the tokens that follow `tip`
            uint32_t tip = ip, my_tok = 0, my_nxt = 0, ll = 0, ml = 0, off = 1, lit = 0;
            while (tip < i_fast) {
                if (!win_ok || (uint32_t)((int32_t)tip - wb) >= 128u) {
                    // (re)fill the window; its base is 4-byte aligned in memory so that every lane moves whole words
                    wb = (int32_t)tip - (int32_t)((uintptr_t)(sp + tip) & 3u);
                    const uint8_t* g = sp + wb;
                    __syncwarp();
#if B2_DEC_PREFETCH
                    // the lines the NEXT refills (and this batch's literal copies) will want: a warp reads its stream strictly
                    // forward, one dependent DRAM round trip per window if nobody asks ahead
                    if (lane < (uint32_t)B2_DEC_PREFETCH && g + 256u + 128u * lane < c.src_end) prefetch_l1(g + 256u + 128u * lane);
#endif
                    const uint32_t w0 = (g + 4u * lane < c.src_end) ? *(const uint32_t*)(g + 4u * lane) : 0u;
                    const uint32_t w1 = (g + 128u + 4u * lane < c.src_end) ? *(const uint32_t*)(g + 128u + 4u * lane) : 0u;
                    ((uint32_t*)win)[lane] = w0; ((uint32_t*)win)[32u + lane] = w1;
                    if (lane < 4u) ((uint32_t*)win)[64u + lane] = 0u;
                    __syncwarp();
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t p = 32u * k + lane;
                        const uint32_t t = win[p];
                        uint32_t l2 = t >> 4, q = p + 1u;
                        bool stop = (wb + (int32_t)p) >= (int32_t)i_fast;
                        if (l2 == 15u) { const uint32_t e = win[q]; l2 += e; ++q; stop |= (e == 255u); }
                        q += l2;
                        if (q > 250u) { stop = true; q = 250u; }
                        q += 2u;
                        if ((t & 15u) == 15u) { const uint32_t e = win[q]; ++q; stop |= (e == 255u); }
                        T1[p] = (uint8_t)(stop ? WSTOP : q);
                    }
                    __syncwarp();
                    #define B2_SQUARE(TA, TB) { uint8_t r0 = TA[TA[lane]], r1 = TA[TA[32u + lane]], r2 = TA[TA[64u + lane]], r3 = TA[TA[96u + lane]]; \
                        TB[lane] = r0; TB[32u + lane] = r1; TB[64u + lane] = r2; TB[96u + lane] = r3; __syncwarp(); }
                    B2_SQUARE(T1, T2) B2_SQUARE(T2, T4) B2_SQUARE(T4, T8) B2_SQUARE(T8, T16)
                    #undef B2_SQUARE
                    win_ok = true;
                }
                // lane n + i: the i-th token after `tip`
                const uint32_t kx = lane - n;                                   // lanes below n already hold a sequence
                uint32_t x = tip - (uint32_t)wb;
                if (kx & 1u) x = T1[x];
                if (kx & 2u) x = T2[x];
                if (kx & 4u) x = T4[x];
                if (kx & 8u) x = T8[x];
                if (kx & 16u) x = T16[x];
                const uint32_t nxt = T1[x];                                     // entries 128..255 are identity: harmless for x >= 128
                const bool have = lane >= n && x < 128u && nxt != WSTOP;
                if (have) {
                    const uint32_t t = win[x];
                    ll = t >> 4; lit = x + 1u;
                    if (ll == 15u) { ll += win[lit]; ++lit; }
                    const uint32_t q = lit + ll;
                    off = (uint32_t)win[q] | ((uint32_t)win[q + 1u] << 8);
                    ml = (t & 15u) + 4u;
                    if (ml == 19u) ml += win[q + 2u];
                    lit = (uint32_t)(wb + (int32_t)lit);
                    my_tok = (uint32_t)(wb + (int32_t)x); my_nxt = (uint32_t)(wb + (int32_t)nxt);
                }
                const uint32_t nh = (uint32_t)__popc(__ballot_sync(FULL_MASK, have));     // `have` lanes form a run that starts at lane n
                if (nh == 0u) break;
                n += nh;
                tip = __shfl_sync(FULL_MASK, my_nxt, n - 1u);
                if (n == 32u) break;
            }
            if (n) {
                const uint32_t sz = (lane < n) ? ll + ml : 0u;
                uint32_t run = sz;
                #pragma unroll
                for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(FULL_MASK, run, d); if ((int)lane >= d) run += y; }
                const uint32_t seq_op = op0 + run - sz;                         // where this lane's sequence starts
                const bool fits = lane < n && seq_op < o_fast;                  // a prefix (lane 0 always fits)
                const uint32_t nf = (uint32_t)__popc(__ballot_sync(FULL_MASK, fits));
                const bool bad = lane < nf && (off == 0u || off > seq_op + ll + reach);              // lz4.c:1977
                if (__any_sync(FULL_MASK, bad)) { err = E_decompressionFailed; break; }
                // entry point and output position after the nf sequences taken
                const uint32_t nx_ip = __shfl_sync(FULL_MASK, my_tok, nf & 31u);
                const uint32_t nx_op = __shfl_sync(FULL_MASK, nf < 32u ? seq_op : seq_op + sz, nf < 32u ? nf : 31u);
                ip = (nf < n) ? nx_ip : tip; op = nx_op;
                n = nf;
                my_lit = lit; my_ll = ll; my_ml = ml; my_off = off; my_op_fast = seq_op;
                if (lane >= n) { my_ll = 0; my_ml = 0; my_off = 1; }
                my_packed = false; fast_batch = true;
            }
        }
        if (n == 0)
        for (; n < 32; ++n) {
            if (ip < i_fast && op < o_fast) {
                const uint32_t token = sp[ip];
                uint32_t ll = token >> 4, ml = (token & 15u) + 4u, lit = ip + 1u;
                bool more = false;
                if (ll == 15u) { const uint32_t e = sp[lit++]; ll += e; more = (e >= 240u); }   // ll stays below 256: 8 bits in the packed form
                uint32_t q = lit + ll;
                const uint32_t offset = (uint32_t)sp[q] | ((uint32_t)sp[q + 1] << 8);
                q += 2u;
                if (ml == 19u) { const uint32_t e = sp[q++]; ml += e; more |= (e == 255u); }
                if (!more) {
                    const uint32_t mop = op + ll;
                    if (offset == 0u || offset > mop + reach) { err = E_decompressionFailed; break; }   // lz4.c:1977
                    if (lane == n) { my_a = lit | (ll << 24); my_b = offset | (ml << 16); }   // lit < 2^24: legacy blocks reach 8 MB + 32 KB
                    ip = q;
                    op = mop + ml;
                    continue;
                }
            }
            // ---- general path
            if (ip >= pend) { err = E_decompressionFailed; break; }
            const uint32_t token = sp[ip++];
            uint32_t ll = token >> 4;
            if (ll == 15u) {                                                // lz4.c:1882-1889
                if (ip + 15u >= iend) { err = E_decompressionFailed; break; }
                uint32_t sbyte;
                do { sbyte = sp[ip++]; ll += sbyte; if (ip + 15u >= iend) break; } while (sbyte == 255u);
            }
            // last-sequence test (lz4.c:1906): not enough input for an offset + next token, or the
            // literals reach the last 12 bytes of the block's output window.  (ll <= 255 * 4 MB: no overflow)
            const bool last = (op + ll + 12u > oend) || (ip + ll + 8u > iend);
            if (last) {
                if (ip + ll != iend || op + ll > oend || !ends_block || op + ll > opend || iend != pend) {
                    err = E_decompressionFailed; break;
                }
                if (lane == n) { my_lit = ip; my_ll = ll; my_ml = 0; my_packed = false; }
                ip += ll; op += ll; ++n; finished = true;
                break;
            }
            if (ip + ll + 2u > pend || op + ll > opend) { err = E_decompressionFailed; break; }
            const uint32_t lit = ip;
            ip += ll;
            const uint32_t offset = (uint32_t)sp[ip] | ((uint32_t)sp[ip + 1] << 8);
            ip += 2u;
            uint32_t ml = token & 15u;
            if (ml == 15u) {                                                // lz4.c:1967-1972
                uint32_t sbyte; bool bad = false;
                do { sbyte = sp[ip++]; ml += sbyte; if (ip + 4u >= iend) { bad = true; break; } } while (sbyte == 255u);
                if (bad) { err = E_decompressionFailed; break; }
            }
            ml += 4u;
            const uint32_t mop = op + ll;
            if (offset == 0u || offset > mop + reach) { err = E_decompressionFailed; break; }           // lz4.c:1977
            if ((uint64_t)mop + ml + 5u > oend || (uint64_t)mop + ml > opend) { err = E_decompressionFailed; break; } // lz4.c:2033-2035
            const bool huge = ml >= HUGE_MATCH && offset < ml;             // run-length style: filled by the caller, out of this loop
            if (lane == n) { my_lit = lit; my_ll = ll; my_ml = huge ? 0u : ml; my_off = offset; my_packed = false; }
            op = mop + ml;
            const bool piece_ends = (ip == pend && !ends_block);           // sub-block ends after a match
            if (piece_ends && op != opend) { err = E_decompressionFailed; break; }
            if (huge) { c.p_dst = mop; c.p_off = offset; c.p_ml = ml; c.p_done = piece_ends; paused = true; ++n; finished = true; break; }
            if (piece_ends) { ++n; finished = true; break; }
        }
        if (err) break;                                                     // nothing of a bad batch is copied
        if (my_packed) { my_lit = my_a & 0xFFFFFFu; my_ll = my_a >> 24; my_off = my_b & 0xFFFFu; my_ml = my_b >> 16; }
        // output position of sequence i = op0 + sum of the sizes of the sequences before it
        uint32_t my_op = my_op_fast;
        if (!fast_batch) {
            const uint32_t sz = my_ll + my_ml;
            uint32_t x = sz;
            #pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(FULL_MASK, x, d); if ((int)lane >= d) x += y; }
            my_op = op0 + x - sz;
        }

        // ---- literals: lane i copies the literals of sequence i; long runs use the whole warp
        {
            const bool act = lane < n && my_ll != 0;
            if (act && my_ll <= LANE_CAP) lane_copy(dp + my_op, sp + my_lit, my_ll);
            uint32_t longm = __ballot_sync(FULL_MASK, act && my_ll > LANE_CAP);
            while (longm) {
                const int j = __ffs(longm) - 1; longm &= longm - 1;
                const uint32_t l = __shfl_sync(FULL_MASK, my_lit, j), o = __shfl_sync(FULL_MASK, my_op, j);
                const uint32_t len = __shfl_sync(FULL_MASK, my_ll, j);
                group_copy_small(dp + o, sp + l, len, lane, 32);
            }
        }
        __syncwarp();
        // ---- matches, in rounds
        {
            const uint32_t my_mop = my_op + my_ll;
            // end of the bytes read, relative to the match destination: -(off) + min(ml, off)  (<= 0)
            const int32_t my_need = (int32_t)(my_ml < my_off ? my_ml : my_off) - (int32_t)my_off;
            uint32_t pending = __ballot_sync(FULL_MASK, lane < n && my_ml != 0);
            while (pending) {
                const int f = __ffs(pending) - 1;
                const uint32_t W = __shfl_sync(FULL_MASK, my_mop, f);       // everything below W is final
                const uint32_t ml_f = __shfl_sync(FULL_MASK, my_ml, f);
                if (ml_f > LANE_CAP) {
                    const uint32_t off_f = __shfl_sync(FULL_MASK, my_off, f);
                    if (off_f >= ml_f) group_copy_small(dp + W, dp + W - off_f, ml_f, lane, 32);
                    else {                                                  // overlapping, < HUGE_MATCH or from the short path: period replication
                        const uint8_t* m = dp + W - off_f; uint32_t r = lane % off_f; const uint32_t step = 32u % off_f;
                        for (uint32_t i = lane; i < ml_f; i += 32) { dp[W + i] = m[r]; r += step; if (r >= off_f) r -= off_f; }
                    }
                    pending &= ~(1u << f);
                } else {
                    const bool ready = ((pending >> lane) & 1u) && my_ml <= LANE_CAP &&
                                       ((int)lane == f || (int64_t)my_mop + my_need <= (int64_t)W);
                    if (ready) { if (my_off >= my_ml) lane_copy(dp + my_mop, dp + my_mop - my_off, my_ml); else lane_match(dp + my_mop, my_off, my_ml); }
                    pending &= ~__ballot_sync(FULL_MASK, ready);
                }
                __syncwarp();
            }
        }
    }
    c.ip = ip; c.op = op;
    return err ? err : (paused ? DEC_PAUSED : 0u);
}

// table entries are read past L1: in walk_pending mode another CTA of this very kernel has just written them
template <typename T>
__device__ __forceinline__ T ld_cg(const T* p) {
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    T v;
    #pragma unroll
    for (uint32_t i = 0; i < sizeof(T) / 16; ++i) ((uint4*)&v)[i] = __ldcg((const uint4*)p + i);
    return v;
}

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
    if (fi->err != OK) return;
    const uint32_t lane = lane_id();
    __shared__ __align__(16) uint8_t s_walk[DEC_WARPS][1568];              // per warp: window (272) + 5 link tables (256 each)
    uint8_t* const my_win = s_walk[threadIdx.x >> 5];
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
    uint32_t walker = 0;
    if (pending) { if (lane == 0) walker = atomicAdd(walk + 2, 1u) == 0u ? 1u : 0u; walker = __shfl_sync(FULL_MASK, walker, 0); }
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
                    if (c < w[0]) { ok = 1; break; }
                    if (w[1]) { ok = c < w[0] ? 1u : 0u; break; }
                    __nanosleep(400);
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
                cu.sp = src + br.src_off; cu.dp = dst + blk_out; cu.win = my_win; cu.src_end = src_end;
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
                    e = decode_piece(cu, (pc.flags & PIECE_ENDS_BLOCK) != 0);
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
__global__ void __launch_bounds__(256)
copy_stored_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint64_t dst_cap,
                   const FrameIndex* __restrict__ fi, const Piece* __restrict__ pieces,
                   const Chain* __restrict__ chains, const uint32_t* __restrict__ stored_list)
{
    if (fi->err != OK) return;
    const uint32_t ns = fi->n_stored;
    if (ns == 0) return;
    constexpr uint32_t SLICE = 65536u;
    const uint32_t per = (fi->block_size + SLICE - 1u) / SLICE;
    const uint64_t total = (uint64_t)ns * per;
    for (uint64_t it = blockIdx.x; it < total; it += gridDim.x) {
        const uint32_t k = (uint32_t)(it / per), sl = (uint32_t)(it - (uint64_t)k * per);
        const Chain ch = chains[stored_list[k]];
        const Piece pc = pieces[ch.first_piece];
        const uint32_t off = sl * SLICE;
        if (off >= pc.src_len || ch.dst_off + pc.src_len > dst_cap) continue;
        const uint32_t len = pc.src_len - off < SLICE ? pc.src_len - off : SLICE;
        group_copy(dst + ch.dst_off + off, src + pc.src_off + off, len, threadIdx.x, blockDim.x);
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
    __shared__ uint32_t s_first_bad, s_short;
    __shared__ unsigned long long s_total;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) { s_first_bad = 0xFFFFFFFFu; s_short = 0xFFFFFFFFu; s_total = 0; *ticket = 0; }
    __syncthreads();
    if (fi->err != OK) {
        if (tid == 0) { DecodeResult r{}; r.err = fi->err; r.layout_ok = 1; *res = r; }
        return;
    }
    const uint32_t np = fi->n_pieces, nb = fi->n_blocks, bs = fi->block_size;
    unsigned long long tot = 0;
    for (uint32_t i = tid; i < np; i += blockDim.x) {
        PieceOut po = outs[i];
        if (po.err) atomicMin(&s_first_bad, i);
        tot += po.out_len;
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
__global__ void __launch_bounds__(128)
verify_block_checksums_kernel(const uint8_t* __restrict__ src, const FrameIndex* __restrict__ fi,
                              const BlockRef* __restrict__ blocks, uint32_t* __restrict__ first_bad)
{
    if (fi->err != OK || !fi->desc.block_checksum) return;
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t stride = gridDim.x * blockDim.x;
    for (; b < fi->n_blocks; b += stride) {
        BlockRef br = blocks[b];
        uint32_t h = xxh32_thread(src + br.src_off, br.csize);
        if (h != ld32u(src + br.src_off + br.csize)) atomicMin(first_bad, b);
    }
}

}  // namespace b2
