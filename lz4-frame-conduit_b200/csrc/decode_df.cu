// decode_df.cu — dataflow decoding of blocks that carry no side information: frames made by the reference conduit
// (linked 64 KB blocks, Conduit.hsc:248-263), by the stock lz4 CLI (4 MB independent blocks, lz4io.c:159-175), legacy
// frames, and plain frames of this codec.  Replaces LZ4_decompress_safe_usingDict / LZ4_decompress_generic
// (lz4/lib/lz4.c:2307-2320, 1653-2068) and the history hand-over of LZ4F_decompress (lz4frame.c:1286-1349, 1805-1835)
// for such frames.  A CPU decodes them as one dependent chain; here the chain is cut twice:
//
//   WHERE the sequences are  — df_links_kernel: for every byte position of every 1 KB group of a block's compressed bytes,
//       "if a sequence started here, where does the chain leave the group, after how many sequences and output bytes"
//       (per-position links in shared memory, contracted by pointer jumping).  df_stitch_kernel then needs ONE table
//       look-up per group to follow the true chain through a block (all blocks at the same time), and df_tickets_kernel
//       writes down the token and output offset of every 32nd sequence: a TICKET any warp can start on.
//   WHEN a match may be copied — decode_df_kernel: one CTA per unit (an independent block, or a whole linked frame).  Its
//       warps draw tickets in order; a warp parses its 32 sequences, copies the literals (they depend on nothing) and then
//       each lane copies ITS match as soon as the bytes it reads are marked final in a bitmap in shared memory (one bit per
//       output byte over a sliding window, plus a watermark below which everything is final), and marks its own bytes.
//       A match waits for nothing but its own sources, and stalls nothing but the matches that really read it.  A warp only
//       ever waits for bytes of EARLIER tickets, all held by running warps of the same CTA, so the lowest unfinished
//       ticket can always proceed: no deadlock.  Everything a match waits for is signalled through shared memory and read
//       through the SM's own L1 — a hand-over costs a few hundred cycles, not an L2 round trip (the first version of this
//       file signalled through a bitmap in global memory with pieces of ~160 sequences; it was bit-exact and 100 x slower).
//       A linked frame needs nothing more: its blocks are parsed independently, only match sources cross block borders.
// CPU models of both halves: tools/tickets_model.c, tools/lane_decode_model.c (the kernels keep their structure; progress
// is a byte bitmap instead of per-batch done masks).
#include "decode_piece.cuh"

namespace b2 {

// smem entry of df_links_kernel (uint2): non-terminal x = link (0..1023) | sequences << 11, y = bytes;
// terminal x = DF_TERM | kind << 29 | value, y = bytes of this sequence (EXIT) or 0
// table entry DfHead.exit: kind << 30 | sequences << 14 | value;  value = exit - group end (EXIT, < 16384) or the position
// in the group of the token the chain stops AT (SLOW: the stitcher steps it serially; INVALID)
constexpr uint32_t DF_TERM = 0x80000000u;
enum : uint32_t { DF_EXIT = 0u, DF_SLOW = 1u, DF_INVALID = 2u };
constexpr uint32_t DF_REL_MAX = 16384u;

__device__ __forceinline__ uint32_t df_term(uint32_t kind, uint32_t value) { return DF_TERM | (kind << 29) | value; }

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

// block that owns global group / ticket number x: last b with base[b] <= x  (base has nb + 1 entries, base[nb] = total)
__device__ __forceinline__ uint32_t df_owner(const uint32_t* __restrict__ base, uint32_t nb, uint32_t x) {
    uint32_t lo = 0, hi = nb;                              // invariant: base[lo] <= x < base[hi]
    while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (base[mid] <= x) lo = mid; else hi = mid; }
    return lo;
}

// ------------------------------------------------------------------------------------------------
// K1a: per-position chain tables of every group.  One warp per group, DF_WARPS warps per CTA.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t DF_WINB = DF_G + DF_LA;                 // window bytes per group (+ 8 bytes of alignment slack)
constexpr uint32_t DF_SMEM_PER_WARP = DF_G * 8u + DF_WINB + 16u;

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
        const uint32_t wend = g0 + DF_WINB < cs ? g0 + DF_WINB : cs;           // bytes [g0, wend) are in the window
        const uint8_t* sp = src + br.src_off;
        // ---- window: aligned words from global memory, bytes behind the block read as zero
        const uint32_t mis = (uint32_t)((uintptr_t)(sp + g0) & 3u);
        const uint8_t* gbase = sp + g0 - mis;
        __syncwarp();
        for (uint32_t w = lane; w < (DF_WINB + 8u) / 4u; w += 32u) {
            const uint8_t* a = gbase + 4u * w;
            const uint32_t v = (a < src_end && a < sp + wend) ? *(const uint32_t*)a : 0u;
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
                                if (q >= wend || r == DF_RUN) { kind = DF_SLOW; term = true; break; }
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
            if (!term && nx >= g1 && nx - g1 >= DF_REL_MAX) { kind = DF_SLOW; term = true; }   // leaves the group by a long way: the serial rule
            uint2 e2;
            if (term) e2 = make_uint2(df_term(kind, p), 0u);                  // SLOW / INVALID: the chain stops AT this token
            else if (nx >= g1) e2 = make_uint2(df_term(DF_EXIT, nx - g1), out);   // leaves the group (nx == cs: leaves the block)
            else e2 = make_uint2((nx - g0) | (1u << 11), out);
            ent[p] = e2;
        }
        __syncwarp();
        // ---- pointer jumping, in place: ent[p] = (link, sequences and bytes from p up to, not including, link).  An update
        // reads one consistent 8-byte entry of the successor, so it does not matter how far other lanes have got (links
        // only ever move forward along the same chain); done when every link points at a terminal entry.
        for (;;) {
            bool changed = false;
            for (uint32_t k = 0; k < DF_G / 32u; ++k) {
                const uint32_t p = 32u * k + lane;
                if (p >= gn) break;
                const uint2 e = ent[p];
                if (e.x & DF_TERM) continue;
                const uint2 s = ent[e.x & 0x7FFu];
                if (s.x & DF_TERM) continue;
                ent[p] = make_uint2((s.x & 0x7FFu) | ((e.x & ~0x7FFu) + (s.x & ~0x7FFu)), e.y + s.y);
                changed = true;
            }
            __syncwarp();
            if (!__any_sync(FULL_MASK, changed)) break;
        }
        // ---- table of the group: where the chain from p leaves it (or stops), the sequences and bytes until there
        DfHead* hd = W.heads + (size_t)g * DF_G;
        for (uint32_t k = 0; k < DF_G / 32u; ++k) {
            const uint32_t p = 32u * k + lane;
            if (p >= gn) break;
            const uint2 e = ent[p];
            uint32_t cnt = 0, bytes = 0; uint2 t = e;
            if (!(e.x & DF_TERM)) { cnt = e.x >> 11; bytes = e.y; t = ent[e.x & 0x7FFu]; }
            const uint32_t kind = (t.x >> 29) & 3u;
            if (kind == DF_EXIT) { ++cnt; bytes += t.y; }                      // EXIT adds its own sequence; SLOW / INVALID stop in front of the token
            DfHead h; h.exit = (kind << 30) | (cnt << 14) | (t.x & 0x3FFFu); h.out = bytes;
            hd[p] = h;
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------
// K1b: the true chain of every block, one thread per block: one table look-up per group.  Writes the first token of
// every group the chain enters (compact, in order) and the block's decoded size and sequence count.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
df_stitch_kernel(const uint8_t* __restrict__ src, const FrameIndex* __restrict__ fi, const BlockRef* __restrict__ blocks, DfWork W)
{
    if (fi->err != OK || !fi->df) return;
    const uint32_t nb = fi->n_blocks;
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += gridDim.x * blockDim.x) {
        const BlockRef br = blocks[b];
        if (br.stored) { W.blk_nent[b] = 0; W.blk_usize[b] = br.csize; W.blk_nseq[b] = 0; continue; }
        const uint8_t* sp = src + br.src_off;
        const uint32_t cs = br.csize, base = W.grp_base[b];
        uint32_t e = 0, o = 0, sq = 0, k = 0, last_g = 0xFFFFFFFFu;
        bool bad = false;
        while (e < cs) {
            const uint32_t g = e / DF_G;
            if (g != last_g) { W.ent[base + k] = make_uint2(e, o); W.ent_seq[base + k] = sq; ++k; last_g = g; }
            const DfHead h = W.heads[(size_t)(base + g) * DF_G + (e - g * DF_G)];
            const uint32_t kind = h.exit >> 30, val = h.exit & 0x3FFFu;
            if (kind == DF_INVALID) { bad = true; break; }
            o += h.out; sq += (h.exit >> 14) & 0x1FFu;
            if (kind == DF_EXIT) { const uint32_t g1 = (g + 1u) * DF_G < cs ? (g + 1u) * DF_G : cs; e = g1 + val; }
            else {                                                            // a token with a long length run: the serial rule, once
                e = g * DF_G + val;
                uint32_t out1 = 0;
                const uint32_t nx = df_step_serial(sp, cs, e, &out1);
                if (nx == 0xFFFFFFFFu) { bad = true; break; }
                o += out1; ++sq; e = nx;
            }
            if (o > (64u << 20)) { bad = true; break; }                        // no block decodes to that much: stop adding up garbage
        }
        if (bad || e != cs) {
            atomicMin(&W.st->bad_key, ((unsigned long long)b << 8) | (unsigned long long)E_decompressionFailed);
            W.blk_nent[b] = 0; W.blk_usize[b] = 0; W.blk_nseq[b] = 0; continue;
        }
        W.blk_nent[b] = k; W.blk_usize[b] = o; W.blk_nseq[b] = sq;
    }
}

// ------------------------------------------------------------------------------------------------
// K1c: output offsets of the blocks and ticket numbering (prefix sums).  One CTA.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
df_layout_kernel(FrameIndex* __restrict__ fi, BlockRef* __restrict__ blocks, uint64_t dst_cap, DfWork W)
{
    if (fi->err != OK || !fi->df) return;
    B2_SHARED_ARRAY(uint64_t, part_out, 256);
    B2_SHARED_ARRAY(uint32_t, part_tk, 256);
    B2_SHARED_VAR(uint32_t, s_big);
    const uint32_t tid = threadIdx.x, nb = fi->n_blocks, bs = fi->block_size;
    if (tid == 0) s_big = 0xFFFFFFFFu;
    __syncthreads();
    const uint32_t per = (nb + 255u) / 256u;
    const uint32_t b0 = tid * per < nb ? tid * per : nb, b1 = b0 + per < nb ? b0 + per : nb;
    uint64_t o = 0; uint32_t tk = 0;
    for (uint32_t b = b0; b < b1; ++b) {
        const uint32_t us = W.blk_usize[b];
        if (us > bs) atomicMin(&s_big, b);                                    // the reference decodes into maxBlockSize bytes (lz4frame.c:1670)
        o += us;
        tk += blocks[b].stored ? 1u : (W.blk_nseq[b] + DF_BATCH - 1u) / DF_BATCH;
    }
    part_out[tid] = o; part_tk[tid] = tk;
    __syncthreads();
    if (tid == 0) {
        uint64_t ro = 0; uint32_t rt = 0;
        for (int i = 0; i < 256; ++i) { const uint64_t to = part_out[i]; const uint32_t tt = part_tk[i]; part_out[i] = ro; part_tk[i] = rt; ro += to; rt += tt; }
        if (s_big != 0xFFFFFFFFu) atomicMin(&W.st->bad_key, ((unsigned long long)s_big << 8) | (unsigned long long)E_decompressionFailed);
        W.st->total_out = ro; W.st->n_tickets = rt;
        W.st->n_units = fi->desc.block_independent ? nb : (nb ? 1u : 0u);
        fi->total_out = ro; fi->n_pieces = rt; fi->n_chains = rt; fi->layout_known = 1;
        fi->need_tickets = rt;
        if (W.st->bad_key == 0xFFFFFFFFFFFFFFFFull) {
            if (rt > W.cap_tickets) fi->err = E_allocation_failed;            // the host grows the ticket table and comes again
            else if (ro > dst_cap) fi->err = E_dstMaxSize_tooSmall;
        }
        W.tk_base[nb] = rt;
    }
    __syncthreads();
    o = part_out[tid]; tk = part_tk[tid];
    for (uint32_t b = b0; b < b1; ++b) {
        BlockRef br = blocks[b];
        br.dst_off = o; br.usize = W.blk_usize[b];
        blocks[b] = br;
        W.tk_base[b] = tk;
        o += br.usize;
        tk += br.stored ? 1u : (W.blk_nseq[b] + DF_BATCH - 1u) / DF_BATCH;
    }
}

// ------------------------------------------------------------------------------------------------
// K1d: tickets.  One thread per entered group walks its sequences (from the group's first token to the next group's)
// and writes down (token, output offset) of every sequence whose number in the block is a multiple of DF_BATCH.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
df_tickets_kernel(const uint8_t* __restrict__ src, const FrameIndex* __restrict__ fi, const BlockRef* __restrict__ blocks, DfWork W)
{
    if (fi->err != OK || !fi->df || W.st->bad_key != 0xFFFFFFFFFFFFFFFFull) return;
    const uint32_t nb = fi->n_blocks, ngroups = W.grp_base[nb];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ngroups; i += gridDim.x * blockDim.x) {
        const uint32_t b = df_owner(W.grp_base, nb, i);
        const uint32_t k = i - W.grp_base[b], ne = W.blk_nent[b];
        if (k >= ne) continue;                                                // (slots behind the entered groups of a block are unused)
        const BlockRef br = blocks[b];
        const uint8_t* sp = src + br.src_off;
        const uint2 a = W.ent[i];
        uint32_t p = a.x, o = a.y, sq = W.ent_seq[i];
        const uint32_t pend = k + 1u < ne ? W.ent[i + 1].x : br.csize;
        uint2* tk = W.tk + W.tk_base[b];
        while (p < pend) {
            if ((sq & (DF_BATCH - 1u)) == 0u) tk[sq / DF_BATCH] = make_uint2(p, o);
            uint32_t out1 = 0;
            const uint32_t nx = df_step_serial(sp, br.csize, p, &out1);
            if (nx == 0xFFFFFFFFu) break;                                     // cannot happen: the stitch walked this chain
            o += out1; ++sq; p = nx;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K2: the dataflow decoder.  One CTA per unit; its warps draw the unit's tickets in order.
// ------------------------------------------------------------------------------------------------
struct DfCta {                                  // shared-memory control block of one CTA
    unsigned long long final_below;             // every output byte (buffer offset) below this is final
    unsigned long long ring_hi[DF_RING];        // output end of the ticket in this ring slot
    uint32_t ring_done[DF_RING];                // 1 = done, 2 = done and ran solo (no bits to clear)
    uint32_t lowest, next, lock, unit, t_end, pad;
    unsigned long long valid_from;              // B2_DF_SMEMWIN: bytes below this buffer offset are not in the data window (a solo ticket wrote them to global memory)
};

size_t decode_df_smem_bytes(uint32_t warps) { return sizeof(DfCta) + DF_WIN_BYTES / 8u + DF_DATA_BYTES + (size_t)warps * 1568u + 64u; }

// a finished ticket: publish it, then retire every finished ticket at the low end (clear their bits, raise the watermark)
__device__ __forceinline__ void df_retire(DfCta* C, const DfShared& S, uint32_t t, uint64_t hi, bool solo, uint32_t lane) {
    DF_FENCE();
    __syncwarp();
    if (lane == 0) {
        C->ring_hi[t % DF_RING] = hi;
        DF_FENCE();
        ((volatile uint32_t*)C->ring_done)[t % DF_RING] = solo ? 2u : 1u;
    }
    for (;;) {
        uint32_t got = 0;
        if (lane == 0) got = atomicCAS(&C->lock, 0u, 1u) == 0u ? 1u : 0u;
        got = __shfl_sync(FULL_MASK, got, 0);
        if (got) {
            uint32_t low = 0; unsigned long long F = 0;
            if (lane == 0) { low = *(volatile uint32_t*)&C->lowest; F = *(volatile unsigned long long*)&C->final_below; }
            low = __shfl_sync(FULL_MASK, low, 0); F = __shfl_sync(FULL_MASK, F, 0);
            for (;;) {
                uint32_t d = 0; unsigned long long h = 0;
                if (lane == 0) { d = ((volatile uint32_t*)C->ring_done)[low % DF_RING]; h = ((volatile unsigned long long*)C->ring_hi)[low % DF_RING]; }
                d = __shfl_sync(FULL_MASK, d, 0); h = __shfl_sync(FULL_MASK, h, 0);
                if (!d) break;
#if B2_DF_SMEMWIN
                DF_FENCE();
                if (d == 1u) df_flush(S, F, h, lane);                             // the ticket's bytes leave the data window for global memory, in rows
                else if (lane == 0) *(volatile unsigned long long*)&C->valid_from = h;   // a solo ticket wrote global memory itself
                __syncwarp();
#endif
                if (d == 1u) df_mark_warp(S, F, h, lane, false);                  // its bits are not needed any more: the watermark says it
                __syncwarp();
                if (lane == 0) ((volatile uint32_t*)C->ring_done)[low % DF_RING] = 0u;
                F = h; ++low;
            }
            __syncwarp();
            if (lane == 0) {
                DF_FENCE();
                *(volatile unsigned long long*)&C->final_below = F;
                DF_FENCE();
                *(volatile uint32_t*)&C->lowest = low;
                DF_FENCE();
                atomicExch(&C->lock, 0u);
            }
        }
        // a ticket may have finished while the lock holder was past it: whoever sees that takes another turn
        uint32_t again = 0;
        if (lane == 0) again = ((volatile uint32_t*)C->ring_done)[*(volatile uint32_t*)&C->lowest % DF_RING] != 0u ? 1u : 0u;
        again = __shfl_sync(FULL_MASK, again, 0);
        if (!again) break;
        if (!got) __nanosleep(32);
    }
}

__global__ void __launch_bounds__(1024, 1)
decode_df_kernel(const uint8_t* __restrict__ src, uint64_t n_src, uint8_t* __restrict__ dst, uint64_t dst_cap,
                 const FrameIndex* __restrict__ fi, const BlockRef* __restrict__ blocks, DfWork W, uint32_t hist)
{
    if (fi->err != OK || !fi->df || W.st->bad_key != 0xFFFFFFFFFFFFFFFFull) return;
    B2_DYN_SHARED(uint8_t, smem);
    DfCta* const C = (DfCta*)smem;
    uint32_t* const bits = (uint32_t*)(smem + ((sizeof(DfCta) + 15u) & ~(size_t)15u));
    const uint32_t lane = lane_id(), wid = threadIdx.x >> 5;
    uint8_t* const wdata = (uint8_t*)(bits + DF_WIN_BYTES / 32u);           // (16-byte aligned: everything in front of it is)
    uint8_t* const my_win = wdata + DF_DATA_BYTES + wid * 1568u;
    for (uint32_t i = lane; i < 5u * 128u; i += 32u) {                       // upper halves of the link tables are identity for good
        const uint32_t t = i >> 7, x = 128u + (i & 127u);
        my_win[272u + 256u * t + x] = (uint8_t)x;
    }
    DfShared S; S.bits = bits; S.final_below = &C->final_below; S.wdata = wdata; S.gdst = dst; S.wlo = (uint32_t)(uintptr_t)dst & 15u; S.gcap = dst_cap;
    const uint32_t nb = fi->n_blocks, block_size = fi->block_size, nunits = W.st->n_units;
    const bool linked = !fi->desc.block_independent;
    const uint8_t* const src_end = src + n_src;
    for (;;) {
        // ---- next unit for this CTA
        __syncthreads();
        if (threadIdx.x == 0) C->unit = atomicAdd(&W.st->ticket_units, 1u);
        __syncthreads();
        const uint32_t u = C->unit;
        if (u >= nunits) break;
        const uint32_t bf = linked ? 0u : u, bl = linked ? nb : u + 1u;       // blocks [bf, bl) form the unit
        const BlockRef br0 = ld_cg(blocks + bf);
        if (!linked && br0.stored) {                                         // raw block (lz4frame.c:1546-1580): everybody copies
            if (br0.dst_off + br0.csize > dst_cap) { if (threadIdx.x == 0) atomicMin(&W.st->bad_key, ((unsigned long long)bf << 8) | (unsigned long long)E_dstMaxSize_tooSmall); }
            else {
                group_copy(dst + br0.dst_off, src + br0.src_off, br0.csize, threadIdx.x, blockDim.x);
                if (threadIdx.x == 0) atomicAdd(&W.st->produced, (unsigned long long)br0.csize);
            }
            continue;
        }
        for (uint32_t i = threadIdx.x; i < DF_WIN_BYTES / 32u; i += blockDim.x) bits[i] = 0u;
        for (uint32_t i = threadIdx.x; i < DF_RING; i += blockDim.x) C->ring_done[i] = 0u;
        if (threadIdx.x == 0) {
            C->final_below = br0.dst_off; C->lowest = C->next = W.tk_base[bf]; C->t_end = W.tk_base[bl]; C->lock = 0u; C->valid_from = br0.dst_off;
        }
        __syncthreads();
        const uint32_t t_end = C->t_end;
        uint32_t cb = bf;                                                    // block of the warp's current ticket (tickets only grow)
        for (;;) {
            uint32_t t = 0;
            if (lane == 0) t = atomicAdd(&C->next, 1u);
            t = __shfl_sync(FULL_MASK, t, 0);
            if (t >= t_end) break;
            while (W.tk_base[cb + 1u] <= t) ++cb;
            const BlockRef br = ld_cg(blocks + cb);
            const uint64_t blk_out = br.dst_off;
            // the ticket's token / output range inside its block
            uint32_t tok, tok_end, out, out_end;
            const bool last_of_block = t + 1u == W.tk_base[cb + 1u];
            if (br.stored) { tok = 0; tok_end = br.csize; out = 0; out_end = br.csize; }
            else {
                const uint2 a = W.tk[t];
                tok = a.x; out = a.y;
                if (last_of_block) { tok_end = br.csize; out_end = br.usize; }
                else { const uint2 z = W.tk[t + 1u]; tok_end = z.x; out_end = z.y; }
            }
#ifdef B2_DF_PROF
            const long long pt0 = clock64();
#endif
            // ---- flow control: the ticket's bytes must fit the bitmap window behind the watermark, its done flag the ring;
            // a ticket too large for the window runs when it is the lowest one (everything in front of it is final then)
            uint32_t solo = 0;
            if (lane == 0) {
                uint32_t nap = DF_NAP_MIN;
                for (;;) {
                    const uint32_t low = *(volatile uint32_t*)&C->lowest;
                    const unsigned long long F = *(volatile unsigned long long*)&C->final_below;
                    const bool fits = blk_out + out_end <= F + DF_HALF && !(B2_DF_SMEMWIN && br.stored);   // (data window: a raw block goes straight to global memory, alone)
                    if (t == low) { solo = fits ? 0u : 1u; break; }
                    if (t - low < DF_RING && fits) break;
                    __nanosleep(nap); nap = nap < 4u * DF_NAP_MAX ? nap * 2u : 4u * DF_NAP_MAX;
                }
            }
            solo = __shfl_sync(FULL_MASK, solo, 0);
            DF_FENCE();
#ifdef B2_DF_PROF
            const long long pt1 = clock64(); uint32_t p_rounds = 0, p_naps = 0;
#endif
            uint32_t e = 0, produced = 0;
            if (br.stored) {                                                  // a raw block inside a linked frame: later blocks may read it
                if (blk_out + br.csize > dst_cap) e = E_dstMaxSize_tooSmall;
                else {
                    group_copy(dst + blk_out, src + br.src_off, br.csize, lane, 32);
                    if (!solo) { DF_FENCE(); __syncwarp(); df_mark_warp(S, blk_out, blk_out + br.csize, lane); }
                    produced = br.csize;
                }
            } else {
                uint64_t oend = blk_out + block_size;
                if (blk_out >= dst_cap) oend = blk_out; else if (oend > dst_cap) oend = dst_cap;
                Cursor cu;
                cu.sp = src + br.src_off; cu.dp = dst + blk_out; cu.win = my_win; cu.tile = nullptr; cu.src_end = src_end;
                cu.ip = tok; cu.iend = tok_end; cu.iend_blk = br.csize;
                cu.op = out; cu.oend = (uint32_t)(oend - blk_out); cu.opiece_end = out_end;
                cu.p_dst = cu.p_off = cu.p_ml = 0; cu.p_done = false;
                // window start: an independent block may only reference its own output; linked blocks reach into the earlier
                // output of the frame, and of earlier batches in front of the buffer (lz4frame.c:1286-1349)
                if (linked) { const uint64_t h = blk_out + hist; cu.reach = h > 0x40000000ull ? 0x40000000u : (uint32_t)h; }
                else cu.reach = 0u;
                cu.df = S; cu.abs0 = (int64_t)blk_out; cu.solo = solo != 0u;
                {   // what this ticket may read from the data window: not what a solo ticket left out of it, and not what later
                    // tickets may overwrite while this one runs (they are admitted up to DF_HALF beyond the watermark, which cannot
                    // pass this ticket's start; everything below that was flushed before this ticket was admitted)
                    unsigned long long vf = 0;
                    if (lane == 0) vf = *(volatile unsigned long long*)&C->valid_from;
                    vf = __shfl_sync(FULL_MASK, vf, 0);
                    const int64_t lo2 = (int64_t)(blk_out + out_end) - (int64_t)DF_HALF;
                    cu.df_slo = (int64_t)vf > lo2 ? (int64_t)vf : lo2;
                } cu.prof_rounds = cu.prof_naps = 0; cu.prof_cyc[0] = cu.prof_cyc[1] = cu.prof_cyc[2] = cu.prof_cyc[3] = 0;
                for (;;) {
                    e = decode_piece<true>(cu, last_of_block);
                    if (e != DEC_PAUSED) break;
                    // a huge overlapping match: wait for its source window [d - off, d), fill, publish
                    __syncwarp();
                    const int64_t a = (int64_t)blk_out + (int64_t)cu.p_dst - (int64_t)cu.p_off;
                    if (!solo) { uint32_t nap = DF_NAP_MIN; while (!df_ready_warp(S, a, (int64_t)blk_out + cu.p_dst, lane)) { __nanosleep(nap); nap = nap < DF_NAP_MAX ? nap * 2u : DF_NAP_MAX; } DF_FENCE(); }
                    if (B2_DF_SMEMWIN && !solo) {                             // into the data window: period replication from the final window
                        const uint32_t di = df_widx(S, (int64_t)blk_out + cu.p_dst);
                        uint32_t r = lane % cu.p_off; const uint32_t step = 32u % cu.p_off;
                        for (uint32_t i = lane; i < cu.p_ml; i += 32u) { wdata[(di + i) & DF_DATA_MASK] = df_getb(S, cu.df_slo, a + (int64_t)r); r += step; if (r >= cu.p_off) r -= cu.p_off; }
                    } else
                    warp_match_overlap(cu.dp + cu.p_dst, cu.p_off, cu.p_ml, lane);
                    if (!solo) { DF_FENCE(); __syncwarp(); df_mark_warp(S, blk_out + cu.p_dst, blk_out + cu.p_dst + cu.p_ml, lane); }
                    __syncwarp();
                    if (cu.p_done) { e = 0; break; }
                }
                produced = cu.op - out;
#ifdef B2_DF_PROF
                p_rounds = cu.prof_rounds; p_naps = cu.prof_naps;
                if (lane == 0) for (int q = 0; q < 4; ++q) atomicAdd(&W.st->prof[6 + q], (unsigned long long)cu.prof_cyc[q]);
#endif
                if (e == 0 && (produced != out_end - out || cu.ip != tok_end)) e = E_decompressionFailed;
                if (e && cu.oend < block_size && e != E_dstMaxSize_tooSmall) e = E_dstMaxSize_tooSmall;   // ran out of destination
            }
            if (lane == 0) {
                if (e) atomicMin(&W.st->bad_key, ((unsigned long long)cb << 8) | (unsigned long long)e);
                else atomicAdd(&W.st->produced, (unsigned long long)produced);
            }
            // (after an error the bytes may be wrong, but later tickets must not wait for them for ever: retire all the same)
#ifdef B2_DF_PROF
            const long long pt2 = clock64();
#endif
            df_retire(C, S, t, blk_out + out_end, solo != 0u, lane);
#ifdef B2_DF_PROF
            if (lane == 0) {
                const long long pt3 = clock64();
                atomicAdd(&W.st->prof[0], (unsigned long long)(pt1 - pt0)); atomicAdd(&W.st->prof[1], (unsigned long long)(pt2 - pt1));
                atomicAdd(&W.st->prof[2], (unsigned long long)(pt3 - pt2)); atomicAdd(&W.st->prof[3], 1ull);
                atomicAdd(&W.st->prof[4], (unsigned long long)p_rounds); atomicAdd(&W.st->prof[5], (unsigned long long)p_naps);
            }
#endif
        }
    }
}

// ------------------------------------------------------------------------------------------------
// result of a dataflow decode (decode_finalize_kernel's job for the other path).  One thread.
// ------------------------------------------------------------------------------------------------
__global__ void df_finalize_kernel(const FrameIndex* __restrict__ fi, const uint32_t* __restrict__ block_crc_bad,
                                   DecodeResult* __restrict__ res, DfWork W)
{
    if (threadIdx.x || blockIdx.x) return;
    if (!fi->df) return;                                                     // not ours: decode_finalize_kernel reports
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
