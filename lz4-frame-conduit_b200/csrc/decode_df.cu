// decode_df.cu — dataflow decoding of blocks that carry no side information: frames made by the reference conduit
// (linked 64 KB blocks, Conduit.hsc:248-263), by the stock lz4 CLI (4 MB independent blocks, lz4io.c:159-175), legacy
// frames, and plain frames of this codec.  Replaces LZ4_decompress_safe_usingDict / LZ4_decompress_generic
// (lz4/lib/lz4.c:2307-2320, 1653-2068) and the history hand-over of LZ4F_decompress (lz4frame.c:1286-1349, 1805-1835)
// for such frames.  A CPU decodes them as one dependent chain; here the chain is cut twice:
//
//   WHERE the sequences are  — df_links_kernel: for every byte position of every 1 KB group of a block's compressed bytes,
//       "if a sequence started here, where does the chain leave the group and how many bytes has it produced by then"
//       (per-position links in shared memory, contracted by pointer jumping).  df_stitch_kernel then needs ONE table
//       look-up per group to follow the true chain through a block; all blocks are stitched at the same time.  The
//       result is, per group, the first token in it and its output offset: a PIECE any warp can start on.
//   WHEN a match may be copied — decode_df_kernel: warps take pieces in stream order; a match whose source lies in
//       front of its piece waits until the bytes it reads are marked final in a bitmap (one bit per output byte), copies,
//       and marks its own bytes.  Waiting is per lane: a match stalls nothing but the matches that really read it.
//       A warp only ever waits for bytes of EARLIER pieces, which running warps hold (tickets are drawn in order), so
//       the lowest unfinished piece can always proceed: no deadlock, whatever the grid's residency.
//       A linked frame needs nothing more: its blocks are parsed independently, only match sources cross block borders,
//       and "earlier output of the frame" is simply a lower bit index.
// CPU models of both halves: tools/tickets_model.c, tools/lane_decode_model.c (the kernels keep their structure but use
// group-aligned pieces instead of 32-sequence tickets and a byte bitmap instead of per-batch done masks).
#include "decode_piece.cuh"

namespace b2 {

constexpr uint32_t DF_TERM = 0x80000000u, DF_KIND_SHIFT = 29u, DF_POS_MASK = 0x00FFFFFFu;
enum : uint32_t { DF_EXIT = 0u, DF_SLOW = 1u, DF_INVALID = 2u };

__device__ __forceinline__ uint32_t df_term(uint32_t kind, uint32_t pos) { return DF_TERM | (kind << DF_KIND_SHIFT) | (pos & DF_POS_MASK); }
__device__ __forceinline__ uint32_t df_kind(uint32_t x) { return (x >> DF_KIND_SHIFT) & 3u; }

// One sequence starting at block-relative position p, read from global memory with the unbounded serial rule
// (lz4_Block_format.md:27-105).  Returns the next token position (== cs: the block ends with this sequence) or
// 0xFFFFFFFF when the sequence runs off the block; *out = bytes it produces.
__device__ uint32_t df_step_serial(const uint8_t* __restrict__ sp, uint32_t cs, uint32_t p, uint32_t* out) {
    const uint32_t t = sp[p];
    uint32_t q = p + 1u, ll = t >> 4;
    if (ll == 15u) { uint32_t e; do { if (q >= cs) return 0xFFFFFFFFu; e = sp[q++]; ll += e; } while (e == 255u); }
    if (ll > cs - q) return 0xFFFFFFFFu;
    q += ll;
    if (q == cs) { *out = ll; return cs; }
    if (cs - q < 2u) return 0xFFFFFFFFu;
    q += 2u;
    uint32_t ml = t & 15u;
    if (ml == 15u) { uint32_t e; do { if (q >= cs) return 0xFFFFFFFFu; e = sp[q++]; ml += e; } while (e == 255u); }
    if (q >= cs) return 0xFFFFFFFFu;                       // a block cannot end with a match (lz4.c:1906-1941)
    *out = ll + ml + 4u;
    return q;
}

// block that owns global group / piece number x: last b with base[b] <= x  (base has nb + 1 entries, base[nb] = total)
__device__ __forceinline__ uint32_t df_owner(const uint32_t* __restrict__ base, uint32_t nb, uint32_t x) {
    uint32_t lo = 0, hi = nb;                              // invariant: base[lo] <= x < base[hi]
    while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (base[mid] <= x) lo = mid; else hi = mid; }
    return lo;
}

// ------------------------------------------------------------------------------------------------
// K1a: per-position chain tables of every group.  One warp per group, DF_WARPS warps per CTA.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t DF_WIN = DF_G + DF_LA;                  // window bytes per group (+ 8 bytes of alignment slack)
constexpr uint32_t DF_SMEM_PER_WARP = DF_G * 8u + DF_WIN + 16u;

size_t df_links_smem_bytes() { return (size_t)DF_WARPS * DF_SMEM_PER_WARP; }

__global__ void __launch_bounds__(DF_WARPS * 32)
df_links_kernel(const uint8_t* __restrict__ src, uint64_t n_src, const FrameIndex* __restrict__ fi,
                const BlockRef* __restrict__ blocks, DfWork W)
{
    if (fi->err != OK || !fi->df) return;
    B2_DYN_SHARED(uint8_t, smem);
    const uint32_t lane = lane_id(), wid = threadIdx.x >> 5;
    uint2* const ent = (uint2*)(smem + (size_t)wid * DF_SMEM_PER_WARP);
    uint8_t* const win = (uint8_t*)(ent + DF_G);
    const uint32_t nb = fi->n_blocks, ngroups = W.grp_base[nb];
    const uint8_t* const src_end = src + n_src;
    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(&W.st->ticket_links, 1u);
        g = __shfl_sync(FULL_MASK, g, 0);
        if (g >= ngroups) break;
        const uint32_t b = df_owner(W.grp_base, nb, g);
        const BlockRef br = blocks[b];
        const uint32_t cs = br.csize;
        const uint32_t g0 = (g - W.grp_base[b]) * DF_G;
        const uint32_t g1 = g0 + DF_G < cs ? g0 + DF_G : cs, gn = g1 - g0;
        const uint32_t wend = g0 + DF_WIN < cs ? g0 + DF_WIN : cs;             // bytes [g0, wend) are in the window
        const uint8_t* sp = src + br.src_off;
        // ---- window: aligned words from global memory, bytes behind the block read as zero
        const uint32_t mis = (uint32_t)((uintptr_t)(sp + g0) & 3u);
        const uint8_t* gbase = sp + g0 - mis;
        __syncwarp();
        for (uint32_t w = lane; w < (DF_WIN + 8u) / 4u; w += 32u) {
            const uint8_t* a = gbase + 4u * w;
            uint32_t v = (a < src_end && a < sp + wend) ? *(const uint32_t*)a : 0u;
            ((uint32_t*)win)[w] = v;
        }
        __syncwarp();
        const uint8_t* wb = win + mis - g0;                                  // wb[P] = byte at block-relative position P (g0 <= P < wend)
        // ---- every position: the sequence that would start there
        for (uint32_t k = 0; k < DF_G / 32u; ++k) {
            const uint32_t p = 32u * k + lane;
            if (p >= gn) break;
            const uint32_t P = g0 + p, t = wb[P];
            uint32_t q = P + 1u, ll = t >> 4, kind = DF_EXIT;
            bool term = false;
            if (ll == 15u) {
                uint32_t e, r = 0;
                do {
                    if (q >= cs) { kind = DF_INVALID; term = true; break; }
                    if (q >= wend || r == DF_RUN) { kind = DF_SLOW; term = true; break; }
                    e = wb[q++]; ll += e; ++r;
                } while (e == 255u);
            }
            uint32_t nx = 0, out = 0;
            if (!term) {
                if (ll > cs - q) { kind = DF_INVALID; term = true; }
                else {
                    q += ll;
                    if (q == cs) { nx = cs; out = ll; }                       // the block's last sequence: literals only
                    else if (cs - q < 2u) { kind = DF_INVALID; term = true; }
                    else {
                        q += 2u;
                        uint32_t ml = t & 15u;
                        if (ml == 15u) {
                            uint32_t e, r = 0;
                            do {
                                if (q >= cs) { kind = DF_INVALID; term = true; break; }
                                if (q >= wend || q < g0 || r == DF_RUN) { kind = DF_SLOW; term = true; break; }
                                e = wb[q++]; ml += e; ++r;
                            } while (e == 255u);
                        }
                        if (!term) {
                            if (q >= cs) { kind = DF_INVALID; term = true; }
                            else { nx = q; out = ll + ml + 4u; }
                        }
                    }
                }
            }
            uint2 e2;
            if (term) e2 = make_uint2(df_term(kind, P), 0u);                  // SLOW / INVALID: the chain stops AT this token
            else if (nx >= g1) e2 = make_uint2(df_term(DF_EXIT, nx), out);    // leaves the group (nx == cs: leaves the block)
            else e2 = make_uint2(nx - g0, out);
            ent[p] = e2;
        }
        __syncwarp();
        // ---- pointer jumping, in place: ent[p] = (link, bytes produced from p up to, not including, link).  An update reads
        // one consistent 8-byte entry of the successor, so it does not matter how far other lanes have got (links only
        // ever move forward along the same chain); done when every link points at a terminal entry.
        for (;;) {
            bool changed = false;
            for (uint32_t k = 0; k < DF_G / 32u; ++k) {
                const uint32_t p = 32u * k + lane;
                if (p >= gn) break;
                const uint2 e = ent[p];
                if (e.x & DF_TERM) continue;
                const uint2 s = ent[e.x];
                if (s.x & DF_TERM) continue;
                ent[p] = make_uint2(s.x, e.y + s.y);
                changed = true;
            }
            __syncwarp();
            if (!__any_sync(FULL_MASK, changed)) break;
        }
        // ---- table of the group: where the chain from p leaves it (or stops), and the bytes produced until there
        DfHead* hd = W.heads + (size_t)g * DF_G;
        for (uint32_t k = 0; k < DF_G / 32u; ++k) {
            const uint32_t p = 32u * k + lane;
            if (p >= gn) break;
            uint2 e = ent[p];
            if (!(e.x & DF_TERM)) {
                const uint2 t = ent[e.x];                                    // terminal: EXIT adds its own sequence, SLOW / INVALID stop in front of it
                e = make_uint2(t.x, e.y + t.y);
            }
            DfHead h; h.exit = e.x; h.out = e.y;
            hd[p] = h;
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------
// K1b: the true chain of every block, one thread per block: one table look-up per group.  Writes the first token of
// every group the chain enters (compact, in order) and the block's decoded size.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
df_stitch_kernel(const uint8_t* __restrict__ src, const FrameIndex* __restrict__ fi, const BlockRef* __restrict__ blocks, DfWork W)
{
    if (fi->err != OK || !fi->df) return;
    const uint32_t nb = fi->n_blocks;
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += gridDim.x * blockDim.x) {
        const BlockRef br = blocks[b];
        if (br.stored) { W.blk_nent[b] = 0; W.blk_usize[b] = br.csize; continue; }
        const uint8_t* sp = src + br.src_off;
        const uint32_t cs = br.csize, base = W.grp_base[b];
        uint32_t e = 0, o = 0, k = 0, last_g = 0xFFFFFFFFu;
        bool bad = false;
        while (e < cs) {
            const uint32_t g = e / DF_G;
            if (g != last_g) { W.ent[base + k] = make_uint2(e, o); ++k; last_g = g; }
            const DfHead h = W.heads[(size_t)(base + g) * DF_G + (e - g * DF_G)];
            const uint32_t kind = df_kind(h.exit);
            if (kind == DF_INVALID) { bad = true; break; }
            e = h.exit & DF_POS_MASK; o += h.out;
            if (kind == DF_SLOW) {                                            // a token with a long length run: the serial rule, once
                uint32_t out1 = 0;
                const uint32_t nx = df_step_serial(sp, cs, e, &out1);
                if (nx == 0xFFFFFFFFu) { bad = true; break; }
                o += out1; e = nx;
            }
            if (o > (64u << 20)) { bad = true; break; }                        // no block decodes to that much: stop adding up garbage
        }
        if (bad || e != cs) { atomicMin(&W.st->bad_key, ((unsigned long long)b << 8) | (unsigned long long)E_decompressionFailed); W.blk_nent[b] = 0; W.blk_usize[b] = 0; continue; }
        W.blk_nent[b] = k; W.blk_usize[b] = o;
    }
}

// ------------------------------------------------------------------------------------------------
// K1c: output offsets of the blocks and piece numbering (prefix sums).  One CTA.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
df_layout_kernel(FrameIndex* __restrict__ fi, BlockRef* __restrict__ blocks, uint64_t dst_cap, DfWork W)
{
    if (fi->err != OK || !fi->df) return;
    B2_SHARED_ARRAY(uint64_t, part_out, 256);
    B2_SHARED_ARRAY(uint32_t, part_pc, 256);
    B2_SHARED_VAR(uint32_t, s_big);
    const uint32_t tid = threadIdx.x, nb = fi->n_blocks, bs = fi->block_size;
    if (tid == 0) s_big = 0xFFFFFFFFu;
    __syncthreads();
    const uint32_t per = (nb + 255u) / 256u;
    const uint32_t b0 = tid * per < nb ? tid * per : nb, b1 = b0 + per < nb ? b0 + per : nb;
    uint64_t o = 0; uint32_t pc = 0;
    for (uint32_t b = b0; b < b1; ++b) {
        const uint32_t us = W.blk_usize[b];
        if (us > bs) atomicMin(&s_big, b);                                    // the reference decodes into maxBlockSize bytes (lz4frame.c:1670)
        o += us;
        pc += blocks[b].stored ? (blocks[b].csize + DF_STORED_PIECE - 1u) / DF_STORED_PIECE : W.blk_nent[b];
    }
    part_out[tid] = o; part_pc[tid] = pc;
    __syncthreads();
    if (tid == 0) {
        uint64_t ro = 0; uint32_t rp = 0;
        for (int i = 0; i < 256; ++i) { const uint64_t to = part_out[i]; const uint32_t tp = part_pc[i]; part_out[i] = ro; part_pc[i] = rp; ro += to; rp += tp; }
        if (s_big != 0xFFFFFFFFu) atomicMin(&W.st->bad_key, ((unsigned long long)s_big << 8) | (unsigned long long)E_decompressionFailed);
        W.st->total_out = ro; W.st->n_pieces = rp;
        fi->total_out = ro; fi->n_pieces = rp; fi->n_chains = rp; fi->layout_known = 1;
        fi->need_pieces = rp;
        if (W.st->bad_key == 0xFFFFFFFFFFFFFFFFull) {
            if (rp > W.cap_pieces) fi->err = E_allocation_failed;             // the host grows the piece table and comes again
            else if (ro > dst_cap) fi->err = E_dstMaxSize_tooSmall;
        }
        W.piece_base[nb] = rp;
    }
    __syncthreads();
    o = part_out[tid]; pc = part_pc[tid];
    for (uint32_t b = b0; b < b1; ++b) {
        BlockRef br = blocks[b];
        br.dst_off = o; br.usize = W.blk_usize[b];
        blocks[b] = br;
        W.piece_base[b] = pc;
        o += br.usize;
        pc += br.stored ? (br.csize + DF_STORED_PIECE - 1u) / DF_STORED_PIECE : W.blk_nent[b];
    }
}

// ------------------------------------------------------------------------------------------------
// K1d: the piece table (one warp per block) and the cleared bitmap.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
df_pieces_kernel(const FrameIndex* __restrict__ fi, const BlockRef* __restrict__ blocks, DfWork W)
{
    if (fi->err != OK || !fi->df || W.st->bad_key != 0xFFFFFFFFFFFFFFFFull) return;
    const uint32_t lane = lane_id(), nb = fi->n_blocks;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    // bitmap: one bit per output byte (+ slack), zero
    {
        const uint64_t words = (W.st->total_out >> 5) + 2u;
        uint4* b4 = (uint4*)W.bits;
        const uint64_t nq = (words + 3u) >> 2;
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nq; i += (uint64_t)gridDim.x * blockDim.x) b4[i] = z;
    }
    for (uint32_t b = warp; b < nb; b += nwarps) {
        const BlockRef br = blocks[b];
        const uint32_t p0 = W.piece_base[b];
        if (br.stored) {
            const uint32_t np = (br.csize + DF_STORED_PIECE - 1u) / DF_STORED_PIECE;
            for (uint32_t k = lane; k < np; k += 32u) {
                DfPiece pc; pc.block = b; pc.tok = k * DF_STORED_PIECE;
                pc.tok_end = (k + 1u) * DF_STORED_PIECE < br.csize ? (k + 1u) * DF_STORED_PIECE : br.csize;
                pc.out = pc.tok; pc.out_end = pc.tok_end; pc.flags = PIECE_STORED | (k + 1u == np ? PIECE_ENDS_BLOCK : 0u); pc.pad0 = pc.pad1 = 0;
                W.pieces[p0 + k] = pc;
            }
            continue;
        }
        const uint32_t ne = W.blk_nent[b], base = W.grp_base[b];
        for (uint32_t k = lane; k < ne; k += 32u) {
            const uint2 a = W.ent[base + k];
            uint2 z = make_uint2(br.csize, br.usize);
            if (k + 1u < ne) z = W.ent[base + k + 1u];
            DfPiece pc; pc.block = b; pc.tok = a.x; pc.tok_end = z.x; pc.out = a.y; pc.out_end = z.y;
            pc.flags = PIECE_EXACT | (k + 1u == ne ? PIECE_ENDS_BLOCK : 0u); pc.pad0 = pc.pad1 = 0;
            W.pieces[p0 + k] = pc;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K2: the dataflow decoder.  Warps draw pieces in stream order.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DEC_WARPS * 32, B2_DF_MINCTAS)
decode_df_kernel(const uint8_t* __restrict__ src, uint64_t n_src, uint8_t* __restrict__ dst, uint64_t dst_cap,
                 const FrameIndex* __restrict__ fi, const BlockRef* __restrict__ blocks, DfWork W, uint32_t hist)
{
    if (fi->err != OK || !fi->df || W.st->bad_key != 0xFFFFFFFFFFFFFFFFull) return;
    const uint32_t lane = lane_id();
    B2_SHARED_ARRAY(uint8_t, s_walk, DEC_WARPS * 1568);                    // per warp: window (272) + 5 link tables (256 each)
    uint8_t* const my_win = s_walk + (threadIdx.x >> 5) * 1568u;
    for (uint32_t i = lane; i < 5u * 128u; i += 32u) {                       // upper halves of the tables are identity for good
        const uint32_t t = i >> 7, x = 128u + (i & 127u);
        my_win[272u + 256u * t + x] = (uint8_t)x;
    }
    __syncwarp();
    const uint32_t np = W.st->n_pieces, block_size = fi->block_size;
    const bool linked = !fi->desc.block_independent;
    const uint8_t* const src_end = src + n_src;
    uint32_t* const bits = W.bits;
    for (;;) {
        uint32_t t = 0;
        if (lane == 0) t = atomicAdd(&W.st->ticket_decode, 1u);
        t = __shfl_sync(FULL_MASK, t, 0);
        if (t >= np) break;
        const DfPiece pc = ld_cg(W.pieces + t);
        const BlockRef br = ld_cg(blocks + pc.block);
        const uint64_t blk_out = br.dst_off;
        uint32_t e = 0, produced = 0;
        if (pc.flags & PIECE_STORED) {                                       // raw bytes (lz4frame.c:1546-1580); later blocks of a linked frame may read them
            const uint32_t len = pc.tok_end - pc.tok;
            if (blk_out + pc.out_end > dst_cap) e = E_dstMaxSize_tooSmall;
            else {
                group_copy(dst + blk_out + pc.out, src + br.src_off + pc.tok, len, lane, 32);
                __threadfence();
                __syncwarp();
                bits_mark_warp(bits, blk_out + pc.out, blk_out + pc.out_end, lane);
                produced = len;
            }
        } else {
            uint64_t oend = blk_out + block_size;
            if (blk_out >= dst_cap) oend = blk_out; else if (oend > dst_cap) oend = dst_cap;
            Cursor cu;
            cu.sp = src + br.src_off; cu.dp = dst + blk_out; cu.win = my_win; cu.src_end = src_end;
            cu.ip = pc.tok; cu.iend = pc.tok_end; cu.iend_blk = br.csize;
            cu.op = pc.out; cu.oend = (uint32_t)(oend - blk_out); cu.opiece_end = pc.out_end;
            cu.p_dst = cu.p_off = cu.p_ml = 0; cu.p_done = false;
            // window start: an independent block may only reference its own output; linked blocks reach into the earlier
            // output of the frame, and of earlier batches in front of the buffer (lz4frame.c:1286-1349)
            if (linked) { const uint64_t h = blk_out + hist; cu.reach = h > 0x40000000ull ? 0x40000000u : (uint32_t)h; }
            else cu.reach = 0u;
            cu.bits = bits; cu.abs0 = (int64_t)blk_out; cu.piece_lo = pc.out;
            const bool ends = (pc.flags & PIECE_ENDS_BLOCK) != 0;
            for (;;) {
                e = decode_piece<true>(cu, ends);
                if (e != DEC_PAUSED) break;
                // a huge overlapping match: its source window [d - off, d) may belong to other warps
                __syncwarp();
                uint8_t* d = cu.dp + cu.p_dst;
                uint32_t ml = cu.p_ml;
                const uint32_t off = cu.p_off;
                if (off > cu.p_dst - pc.out) {                               // the window starts below piece_lo: wait, then take the first period past L1
                    const int64_t a = (int64_t)blk_out + (int64_t)cu.p_dst - (int64_t)off, l = (int64_t)blk_out + pc.out;
                    while (!bits_ready_warp(bits, a, l, lane)) __nanosleep(100);
                    for (uint32_t i = lane; i < off; i += 32u) d[i] = __ldcg(d - off + i);
                    __syncwarp();
                    d += off; ml -= off;                                     // the rest repeats bytes of this warp's own
                }
                if (ml) warp_match_overlap(d, off, ml, lane);
                __threadfence();
                __syncwarp();
                bits_mark_warp(bits, blk_out + cu.p_dst, blk_out + cu.p_dst + cu.p_ml, lane);
                if (cu.p_done) { e = 0; break; }
            }
            produced = cu.op - pc.out;
            if (e == 0 && (produced != pc.out_end - pc.out || cu.ip != pc.tok_end)) e = E_decompressionFailed;
            if (e && cu.oend < block_size && e != E_dstMaxSize_tooSmall) e = E_dstMaxSize_tooSmall;   // ran out of destination
        }
        if (lane == 0) {
            if (e) atomicMin(&W.st->bad_key, ((unsigned long long)pc.block << 8) | (unsigned long long)e);
            else atomicAdd(&W.st->produced, (unsigned long long)produced);
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------
// result of a dataflow decode (decode_finalize_kernel's job for the other path).  One thread.
// ------------------------------------------------------------------------------------------------
__global__ void df_finalize_kernel(const FrameIndex* __restrict__ fi, const uint32_t* __restrict__ block_crc_bad,
                                   DecodeResult* __restrict__ res, DfWork W)
{
    if (threadIdx.x || blockIdx.x) return;
    if (fi->err != OK && !fi->df) return;                                    // not ours: decode_finalize_kernel reports
    if (!fi->df) return;
    DecodeResult r{};
    r.layout_ok = 1;
    if (fi->err != OK) { r.err = fi->err; *res = r; return; }
    const unsigned long long key = W.st->bad_key;
    const uint32_t crc_bad = block_crc_bad ? *block_crc_bad : 0xFFFFFFFFu;
    const uint32_t bad_block = key == 0xFFFFFFFFFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)(key >> 8);
    r.total_out = W.st->total_out;
    if (crc_bad != 0xFFFFFFFFu && crc_bad <= bad_block) { r.err = E_blockChecksum_invalid; r.err_block = crc_bad; }
    else if (bad_block != 0xFFFFFFFFu) { r.err = (uint32_t)(key & 0xFFu); r.err_block = bad_block; }
    else if (W.st->produced != W.st->total_out) r.err = E_decompressionFailed;
    if (r.err == OK && fi->desc.has_content_size && fi->desc.content_size != r.total_out) r.err = E_frameSize_wrong;   // lz4frame.c:1715-1716
    *res = r;
}

}  // namespace b2
