// decode_piece.cuh — the warp-level LZ4 sequence decoder shared by decode.cu (one warp per independent piece) and
// decode_df.cu (dataflow mode: pieces of foreign / linked blocks whose matches wait for bytes other warps produce).
// Replaces LZ4_decompress_generic (lz4/lib/lz4.c:1653-2068); acceptance rules mirror its safe loop (lz4.c:1844-2055).
#pragma once
#include "common.cuh"
#include "kernels.h"

#ifndef B2_DEC_LANE_CAP
#define B2_DEC_LANE_CAP 24
#endif
// Output tile of the one-warp-per-piece decoder (decode.cu), bytes per warp, a power of two; 0 = none.  The last B2_DEC_TILE
// bytes a warp produced live in shared memory (circular, indexed by the low bits of the global address): literals and matches
// are written THERE, a match whose source is that recent reads it from there (a shared-memory round trip instead of one
// through L2 for bytes the warp itself stored a moment ago), and every byte goes to global memory once, in 16-byte rows.
#ifndef B2_DEC_TILE
#define B2_DEC_TILE 0
#endif

namespace b2 {

constexpr uint32_t TILE = B2_DEC_TILE, TILE_MASK = TILE ? TILE - 1u : 0u, TILE_BATCH_MAX = TILE / 2u;
static_assert((TILE & TILE_MASK) == 0u && (TILE == 0u || TILE >= 1024u), "B2_DEC_TILE: 0 or a power of two >= 1024");

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
    // dataflow mode (decode_df.cu): this piece is one ticket of a unit that a whole CTA decodes; bytes in front of it are
    // written by OTHER warps of the CTA.  abs0 = offset of the owning block's first output byte in the destination buffer.
 DfShared df;
    int64_t abs0;
    uint32_t prof_rounds, prof_naps;   // -DB2_DF_PROF: match rounds / of them without progress
    uint32_t prof_cyc[4];              // ... cycles: parse, literals, chase, match rounds
    bool solo;             // the piece runs alone (every earlier byte is final, nobody reads the map): no waits, no marks
    uint8_t* tile;         // B2_DEC_TILE bytes of this warp's shared memory (16-byte aligned), unused in dataflow mode
    int64_t df_slo;        // B2_DF_SMEMWIN: sources at buffer offsets >= this are read from the CTA's data window, lower ones from global memory
};

// the CTA-scope fence of the dataflow hand-over (tuning builds: -DB2_DF_FENCE=0 none — measurement only, not correct;
// 2 = device scope)
#ifndef B2_DF_FENCE
#define B2_DF_FENCE 1
#endif
#ifndef B2_DF_PIPECOPY
#define B2_DF_PIPECOPY 0      // 1 = per-lane match copies issue all their loads before their stores (measured: +45 % time, the kernel is bound by instruction issue, not by load latency)
#endif
#if B2_DF_FENCE == 0
#define DF_FENCE() do {} while (0)
#elif B2_DF_FENCE == 2
#define DF_FENCE() __threadfence()
#else
#define DF_FENCE() __threadfence_block()
#endif

// ---- dataflow helpers.  The CTA keeps, in shared memory, one bit per output byte for a sliding window of the unit's
// output (circular, DF_WIN_BYTES long) and a watermark below which every byte is final.  A writer stores its bytes,
// fences (CTA scope: writer and readers sit on one SM and share its L1), then sets their bits; a reader that finds the
// bits set fences and then loads the bytes.  Positions are absolute offsets in the destination buffer; negative ones are
// history in front of it (earlier batches of a linked frame): final by construction.
__device__ __forceinline__ bool df_ready(const DfShared& S, int64_t a, int64_t b) {
    const int64_t F = (int64_t)*S.final_below;
    if (b <= F) return true;
    if (a < F) a = F;
    const uint32_t w0 = (uint32_t)((uint64_t)a >> 5), w1 = (uint32_t)((uint64_t)(b - 1) >> 5);
    const uint32_t m0 = 0xFFFFFFFFu << ((uint32_t)a & 31u), m1 = 0xFFFFFFFFu >> (31u - ((uint32_t)(b - 1) & 31u));
    const volatile uint32_t* bits = S.bits;
    if (w0 == w1) { const uint32_t m = m0 & m1; return (bits[w0 & DF_WIN_MASK] & m) == m; }
    if ((bits[w0 & DF_WIN_MASK] & m0) != m0) return false;
    for (uint32_t w = w0 + 1; w != w1; ++w) if (bits[w & DF_WIN_MASK] != 0xFFFFFFFFu) return false;
    return (bits[w1 & DF_WIN_MASK] & m1) == m1;
}
// the same question asked by the whole warp about a long range (uniform answer)
__device__ __forceinline__ bool df_ready_warp(const DfShared& S, int64_t a, int64_t b, uint32_t lane) {
    int64_t F = 0;
    if (lane == 0) F = (int64_t)*S.final_below;
    F = __shfl_sync(FULL_MASK, F, 0);
    if (b <= F) return true;
    if (a < F) a = F;
    const uint32_t w0 = (uint32_t)((uint64_t)a >> 5), w1 = (uint32_t)((uint64_t)(b - 1) >> 5);
    const uint32_t m0 = 0xFFFFFFFFu << ((uint32_t)a & 31u), m1 = 0xFFFFFFFFu >> (31u - ((uint32_t)(b - 1) & 31u));
    const volatile uint32_t* bits = S.bits;
    bool ok = true;
    for (uint32_t k = lane; k <= w1 - w0 && ok; k += 32u) {
        const uint32_t w = w0 + k;
        uint32_t m = 0xFFFFFFFFu;
        if (w == w0) m &= m0;
        if (w == w1) m &= m1;
        ok = (bits[w & DF_WIN_MASK] & m) == m;
    }
    return __all_sync(FULL_MASK, ok);
}
// marks [a, b) final; the caller has fenced its stores.  One lane, short ranges (a few words).
__device__ __forceinline__ void df_mark(const DfShared& S, uint64_t a, uint64_t b) {
    if (b <= a) return;
    const uint32_t w0 = (uint32_t)(a >> 5), w1 = (uint32_t)((b - 1) >> 5);
    const uint32_t m0 = 0xFFFFFFFFu << ((uint32_t)a & 31u), m1 = 0xFFFFFFFFu >> (31u - ((uint32_t)(b - 1) & 31u));
    if (w0 == w1) { atomicOr(S.bits + (w0 & DF_WIN_MASK), m0 & m1); return; }
    atomicOr(S.bits + (w0 & DF_WIN_MASK), m0);
    for (uint32_t w = w0 + 1; w != w1; ++w) atomicOr(S.bits + (w & DF_WIN_MASK), 0xFFFFFFFFu);
    atomicOr(S.bits + (w1 & DF_WIN_MASK), m1);
}
// ... by the whole warp, any length below the window (every lane has fenced its stores and the warp has synchronised);
// `set` = false clears the bits again (retiring a ticket)
__device__ __forceinline__ void df_mark_warp(const DfShared& S, uint64_t a, uint64_t b, uint32_t lane, bool set = true) {
    if (b <= a) return;
    const uint32_t w0 = (uint32_t)(a >> 5), w1 = (uint32_t)((b - 1) >> 5);
    const uint32_t m0 = 0xFFFFFFFFu << ((uint32_t)a & 31u), m1 = 0xFFFFFFFFu >> (31u - ((uint32_t)(b - 1) & 31u));
    for (uint32_t k = lane; k <= w1 - w0; k += 32u) {
        const uint32_t w = w0 + k;
        uint32_t m = 0xFFFFFFFFu;
        if (w == w0) m &= m0;
        if (w == w1) m &= m1;
        if (set) atomicOr(S.bits + (w & DF_WIN_MASK), m); else atomicAnd(S.bits + (w & DF_WIN_MASK), ~m);
    }
}

// ---- B2_DF_SMEMWIN: the data window.  Index of buffer offset p; one byte of the unit's output wherever it lives.
__device__ __forceinline__ uint32_t df_widx(const DfShared& S, int64_t p) { return ((uint32_t)p + S.wlo) & DF_DATA_MASK; }
__device__ __forceinline__ uint8_t df_getb(const DfShared& S, int64_t slo, int64_t p) {
    return p >= slo ? S.wdata[df_widx(S, p)] : S.gdst[p];
}
// writes the window's bytes [a, b) (buffer offsets) to global memory: 16-byte rows, loose bytes at the ends.  One warp.
__device__ __forceinline__ void df_flush(const DfShared& S, uint64_t a, uint64_t b, uint32_t lane) {
    if (b > S.gcap) b = S.gcap;
    if (b <= a) return;
    uint64_t head = (16u - (uint32_t)((a + S.wlo) & 15u)) & 15u;
    if (head > b - a) head = b - a;
    if (lane < head) S.gdst[a + lane] = S.wdata[df_widx(S, (int64_t)(a + lane))];
    a += head;
    const uint64_t nq = (b - a) >> 4;
    for (uint64_t q = lane; q < nq; q += 32u)
        *(uint4*)(S.gdst + a + 16u * q) = *(const uint4*)(S.wdata + df_widx(S, (int64_t)(a + 16u * q)));
    a += nq << 4;
    if (lane < b - a) S.gdst[a + lane] = S.wdata[df_widx(S, (int64_t)(a + lane))];
}

// Per-lane byte copy of n bytes; loads are batched four at a time ahead of the stores so that a lane
// keeps several loads in flight.  `s` must not overlap the destination (callers guarantee it).
#ifndef B2_LANE_COPY8
#define B2_LANE_COPY8 0       // 1 = eight loads in flight before the first store (half the dependent round trips of a long copy, same instruction count)
#endif
__device__ __forceinline__ void lane_copy(uint8_t* d, const uint8_t* s, uint32_t n) {
    uint32_t k = 0;
#if B2_LANE_COPY8
    for (; k + 8 <= n; k += 8) {
        uint8_t a = s[k], b = s[k + 1], c = s[k + 2], e = s[k + 3], f = s[k + 4], g = s[k + 5], h = s[k + 6], i = s[k + 7];
        d[k] = a; d[k + 1] = b; d[k + 2] = c; d[k + 3] = e; d[k + 4] = f; d[k + 5] = g; d[k + 6] = h; d[k + 7] = i;
    }
#endif
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

// ... the window being [s, s + off) anywhere (a chased source, decode_piece's dataflow branch)
__device__ __forceinline__ void lane_match_from(uint8_t* d, const uint8_t* s, uint32_t off, uint32_t n) {
    uint32_t r = 0;
    for (uint32_t k = 0; k < n; ++k) { d[k] = s[r]; r = (r + 1 == off) ? 0u : r + 1; }
}

// Latency-first per-lane copy of n <= LANE_CAP bytes for the dataflow decoder, where a copy sits on a chain of dependent
// copies and its source usually comes from L2 (another warp wrote it): EVERY load is issued before the first store, so the
// copy costs one memory round trip instead of one per four bytes.  period = 0: plain copy from s; else the window
// [s, s + period) is replicated.  (More instructions than lane_copy for short copies: not for the throughput-bound path.)
__device__ __forceinline__ void lane_copy_pipelined(uint8_t* d, const uint8_t* s, uint32_t n, uint32_t period) {
    uint32_t v[(B2_DEC_LANE_CAP + 3) / 4];
    #pragma unroll
    for (int j = 0; j < (B2_DEC_LANE_CAP + 3) / 4; ++j) v[j] = 0u;
    uint32_t r = 0;
    #pragma unroll
    for (int k = 0; k < B2_DEC_LANE_CAP; ++k) {
        if ((uint32_t)k < n) {
            v[k >> 2] |= (uint32_t)s[period ? r : (uint32_t)k] << (8 * (k & 3));
            r = (r + 1u == period) ? 0u : r + 1u;
        }
    }
    #pragma unroll
    for (int k = 0; k < B2_DEC_LANE_CAP; ++k) if ((uint32_t)k < n) d[k] = (uint8_t)(v[k >> 2] >> (8 * (k & 3)));
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

// table entries are read past L1: in walk_pending mode another CTA of this very kernel has just written them
template <typename T>
__device__ __forceinline__ T ld_cg(const T* p) {
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    T v;
    #pragma unroll
    for (uint32_t i = 0; i < sizeof(T) / 16; ++i) ((uint4*)&v)[i] = __ldcg((const uint4*)p + i);
    return v;
}


// ---- output tile (B2_DEC_TILE).  State: the tile holds the bytes [tv, end of the last tiled batch) of the block's output at
// index (position + lo) & TILE_MASK, lo = low bits of the block's global output address (so that 16-byte rows of the tile are
// 16-byte rows of global memory); global memory holds everything below fp; tv <= fp always.
struct TileState { uint8_t* t; uint32_t lo, tv, fp; };

// writes [T.fp, upto) to global memory — whole 16-byte rows only unless `exact`
__device__ __forceinline__ void tile_flush(TileState& T, uint8_t* dp, uint32_t upto, bool exact, uint32_t lane) {
    uint32_t lim = upto;
    if (!exact) lim -= (upto + T.lo) & 15u;
    if ((int32_t)(lim - T.fp) <= 0) return;
    uint32_t a = T.fp;
    uint32_t head = (16u - ((a + T.lo) & 15u)) & 15u;
    if (head > lim - a) head = lim - a;
    if (lane < head) dp[a + lane] = T.t[(a + T.lo + lane) & TILE_MASK];
    a += head;
    const uint32_t nq = (lim - a) >> 4;
    for (uint32_t q = lane; q < nq; q += 32u)
        *(uint4*)(dp + a + 16u * q) = *(const uint4*)(T.t + ((a + T.lo + 16u * q) & TILE_MASK));
    a += nq << 4;
    if (lane < lim - a) dp[a + lane] = T.t[(a + T.lo + lane) & TILE_MASK];
    T.fp = lim;
}
// one byte of the block's output at (signed) position pos: from the tile when it is recent enough, else from global memory
__device__ __forceinline__ uint8_t tile_getb(const TileState& T, const uint8_t* dp, int32_t pos) {
    return pos >= (int32_t)T.tv ? T.t[((uint32_t)pos + T.lo) & TILE_MASK] : dp[pos];
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
#ifndef B2_DF_LANE_CAP
#define B2_DF_LANE_CAP 16     // ... in the dataflow kernel, whose rounds run with few lanes each: a shorter per-lane loop pays (measured: 16 is
#endif                        // 4 % faster than 24 there and 6 % slower in the one-warp-per-piece kernel, profiles/r3l_lane_copy_variants.txt)

// Decodes one piece.  The warp first WALKS up to 32 sequences (uniform control flow: every lane follows the
// token chain, lane i keeps the fields of sequence i), then copies them lane-parallel: literals of all 32
// sequences at once, then matches in rounds — a match may run as soon as every byte it reads is final, i.e.
// lies below the destination of the first match still pending (all literals are already in place).
// The walk has a short path for the common sequence (both nibbles < 15, far from every limit) whose only
// check is the offset; everything else takes the general path with the reference's acceptance rules.
// returns 0 or an Err; c.op is advanced
template <bool DF>
__device__ B2_NOINLINE uint32_t decode_piece(Cursor& c, bool ends_block)
{
    constexpr uint32_t LANE_CAP = DF ? (uint32_t)B2_DF_LANE_CAP : b2::LANE_CAP;   // (shadows the namespace-scope value)
    const uint32_t lane = lane_id();
    const uint8_t* __restrict__ sp = c.sp;
    uint8_t* dp = c.dp;
    uint32_t ip = c.ip, op = c.op;
    const uint32_t iend = c.iend_blk, pend = c.iend, oend = c.oend, opend = c.opiece_end, reach = c.reach;
    // short path (lengths with at most one extension byte) allowed while  ip + 288 <= pend  and
    // op + 560 <= min(oend, opend): inside that region none of the general path's limit checks can fire
    // dataflow pieces are short (about 1 KB of the stream): their own end only bounds WHICH tokens are theirs; the limit
    // checks the short path skips are about the end of the BLOCK (input) and of its output window
    uint32_t i_fast = DF ? (iend >= 288u ? (pend < iend - 287u ? pend : iend - 287u) : 0u) : (pend >= 288u ? pend - 287u : 0u);
    const uint32_t olim = oend < opend ? oend : opend;
    uint32_t o_fast = DF ? (oend >= 560u ? (opend < oend - 559u ? opend : oend - 559u) : 0u) : (olim >= 560u ? olim - 559u : 0u);
    const int64_t abs0 = DF ? c.abs0 : 0;
    const bool solo = DF ? c.solo : true;
    asm volatile("" : "+r"(i_fast), "+r"(o_fast));                          // keep them in registers (no re-derivation per sequence)
    uint32_t err = 0;
    if (pend == ip) return E_decompressionFailed;                           // empty input (lz4.c:1702)
    TileState T{};                                                          // (dead code unless B2_DEC_TILE: the tile starts empty)
    uint32_t t_end = op;                                                    // end of the bytes the copy phases have produced
    if constexpr (!DF && TILE != 0u) { T.t = c.tile; T.lo = (uint32_t)(uintptr_t)dp & TILE_MASK; T.tv = T.fp = op; }
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
#ifdef B2_DF_PROF
        const long long q0 = clock64();
#endif
        if (DF && ip >= pend && !ends_block) {                              // the piece's last sequence took the short path
            if (ip != pend || op != opend) err = E_decompressionFailed;
            break;
        }
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
            if (DF && !ends_block && ip >= pend) break;                     // the piece's tokens are used up: copy what was gathered
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

#ifdef B2_DF_PROF
        const long long q1 = clock64();
#endif
        bool tiled = false;
        if constexpr (!DF && TILE != 0u) {
            // a batch goes through the tile when it is small enough to leave the tile's older half in place (a batch that is not
            // — long runs, a paused huge match — writes global memory directly, after everything before it has been flushed)
            tiled = !paused && op - op0 <= TILE_BATCH_MAX;
            if (!tiled) tile_flush(T, dp, t_end, true, lane);
        }
        if (tiled) {
            if ((int32_t)(op - T.tv) > (int32_t)TILE) T.tv = op - TILE;       // what this batch's bytes overwrite is gone
            uint8_t* const tl = T.t; const uint32_t lo = T.lo;
            {   // ---- literals into the tile
                const bool act = lane < n && my_ll != 0;
                if (act && my_ll <= LANE_CAP) {
                    const uint32_t di = (my_op + lo) & TILE_MASK;
                    if (di + my_ll <= TILE) lane_copy(tl + di, sp + my_lit, my_ll);
                    else for (uint32_t k = 0; k < my_ll; ++k) tl[(di + k) & TILE_MASK] = sp[my_lit + k];
                }
                uint32_t longm = __ballot_sync(FULL_MASK, act && my_ll > LANE_CAP);
                while (longm) {
                    const int j = __ffs(longm) - 1; longm &= longm - 1;
                    const uint32_t l = __shfl_sync(FULL_MASK, my_lit, j), o = __shfl_sync(FULL_MASK, my_op, j) + lo;
                    const uint32_t len = __shfl_sync(FULL_MASK, my_ll, j);
                    for (uint32_t i = lane; i < len; i += 32u) tl[(o + i) & TILE_MASK] = sp[l + i];
                }
            }
            __syncwarp();
            {   // ---- matches in rounds, same order rule as below; sources come from the tile when they are recent enough
                const uint32_t my_mop = my_op + my_ll;
                const int32_t my_need = (int32_t)(my_ml < my_off ? my_ml : my_off) - (int32_t)my_off;
                const int32_t my_a = (int32_t)(my_mop - my_off);            // first source byte (negative: history in front of the block)
                uint32_t pending = __ballot_sync(FULL_MASK, lane < n && my_ml != 0);
                while (pending) {
                    const int f = __ffs(pending) - 1;
                    const uint32_t W = __shfl_sync(FULL_MASK, my_mop, f);
                    const uint32_t ml_f = __shfl_sync(FULL_MASK, my_ml, f);
                    if (ml_f > LANE_CAP) {
                        const uint32_t off_f = __shfl_sync(FULL_MASK, my_off, f);
                        const int32_t a = (int32_t)(W - off_f);
                        if (off_f >= ml_f) {                                   // plain copy; where the source lies is the same question for every lane
                            if (a >= (int32_t)T.tv) { for (uint32_t i = lane; i < ml_f; i += 32u) tl[(W + lo + i) & TILE_MASK] = tl[((uint32_t)a + lo + i) & TILE_MASK]; }
                            else if (a + (int32_t)ml_f <= (int32_t)T.tv) { for (uint32_t i = lane; i < ml_f; i += 32u) tl[(W + lo + i) & TILE_MASK] = dp[a + (int32_t)i]; }
                            else { for (uint32_t i = lane; i < ml_f; i += 32u) tl[(W + lo + i) & TILE_MASK] = tile_getb(T, dp, a + (int32_t)i); }
                        }
                        else {
                            uint32_t r = lane % off_f; const uint32_t step = 32u % off_f;
                            for (uint32_t i = lane; i < ml_f; i += 32u) { tl[(W + lo + i) & TILE_MASK] = tile_getb(T, dp, a + (int32_t)r); r += step; if (r >= off_f) r -= off_f; }
                        }
                        pending &= ~(1u << f);
                    } else {
                        const bool ready = ((pending >> lane) & 1u) && my_ml <= LANE_CAP &&
                                           ((int)lane == f || (int64_t)my_mop + my_need <= (int64_t)W);
                        if (ready) {
                            const uint32_t di = (my_mop + lo) & TILE_MASK, si = ((uint32_t)my_a + lo) & TILE_MASK;
                            const bool in_t = my_a >= (int32_t)T.tv;
                            const bool simple = my_off >= my_ml && di + my_ml <= TILE &&
                                                (in_t ? si + my_ml <= TILE : my_a + (int32_t)my_ml <= (int32_t)T.tv);
                            if (simple) lane_copy(tl + di, in_t ? (const uint8_t*)(tl + si) : (const uint8_t*)(dp + my_a), my_ml);
                            else {
                                uint32_t r = 0;
                                for (uint32_t k = 0; k < my_ml; ++k) {
                                    tl[(di + k) & TILE_MASK] = tile_getb(T, dp, my_a + (int32_t)r);
                                    r = (r + 1u == my_off) ? 0u : r + 1u;
                                }
                            }
                        }
                        pending &= ~__ballot_sync(FULL_MASK, ready);
                    }
                    __syncwarp();
                }
            }
            t_end = op;
            tile_flush(T, dp, t_end, false, lane);
            continue;
        }
        // ---- literals: lane i copies the literals of sequence i; long runs use the whole warp
        {
            const bool act = lane < n && my_ll != 0;
            const bool wmode = DF && B2_DF_SMEMWIN && !solo;                 // the bytes go to the CTA's data window, not to global memory
            if (act && my_ll <= LANE_CAP) {
                if (wmode) {
                    const uint32_t di = df_widx(c.df, abs0 + my_op);
                    if (di + my_ll <= DF_DATA_BYTES) lane_copy(c.df.wdata + di, sp + my_lit, my_ll);
                    else for (uint32_t k = 0; k < my_ll; ++k) c.df.wdata[(di + k) & DF_DATA_MASK] = sp[my_lit + k];
                } else lane_copy(dp + my_op, sp + my_lit, my_ll);
            }
            uint32_t longm = __ballot_sync(FULL_MASK, act && my_ll > LANE_CAP);
            const uint32_t longm0 = longm;
            while (longm) {
                const int j = __ffs(longm) - 1; longm &= longm - 1;
                const uint32_t l = __shfl_sync(FULL_MASK, my_lit, j), o = __shfl_sync(FULL_MASK, my_op, j);
                const uint32_t len = __shfl_sync(FULL_MASK, my_ll, j);
                if (wmode) { const uint32_t di = df_widx(c.df, abs0 + o); for (uint32_t i = lane; i < len; i += 32u) c.df.wdata[(di + i) & DF_DATA_MASK] = sp[l + i]; }
                else group_copy_small(dp + o, sp + l, len, lane, 32);
            }
            if (DF && !solo) {                                              // literals never wait: publish them at once
                DF_FENCE();
                __syncwarp();
                if (act && my_ll <= LANE_CAP) df_mark(c.df, (uint64_t)(abs0 + my_op), (uint64_t)(abs0 + my_op) + my_ll);
                longm = longm0;
                while (longm) {
                    const int j = __ffs(longm) - 1; longm &= longm - 1;
                    const uint32_t o = __shfl_sync(FULL_MASK, my_op, j), len = __shfl_sync(FULL_MASK, my_ll, j);
                    df_mark_warp(c.df, (uint64_t)(abs0 + o), (uint64_t)(abs0 + o) + len, lane);
                }
            }
        }
        __syncwarp();
#ifdef B2_DF_PROF
        const long long q2 = clock64();
#endif
        // ---- matches, in rounds
        if (!DF || solo) {
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
        } else {
            // dataflow: every byte of the unit's output has a "final" bit (literals were marked above, matches are marked as
            // they are copied, by whichever warp of the CTA copies them).  A match goes as soon as the bits of the bytes IT reads
            // are set — it waits for nothing else, and nothing waits for it but the matches that really read it.
            const uint32_t my_mop = my_op + my_ll;
            const uint32_t my_span = my_ml < my_off ? my_ml : my_off;       // bytes of the source window that are read
            int64_t xa = abs0 + (int64_t)my_mop - (int64_t)my_off;          // the source window, in buffer coordinates: [xa, xa + span)
            uint32_t pending = __ballot_sync(FULL_MASK, lane < n && my_ml != 0);
            // ---- source chasing inside the batch: a match that reads nothing but the output of an EARLIER, still pending,
            // non-overlapping match of this batch reads that match's source instead (the bytes are the same).  Log-like data is
            // full of "this line's field = the previous line's field = the line before's ...": without this every link of
            // such a chain costs a round trip through memory; with it the lanes follow the chain back with shuffles.
            {
                const uint32_t key = lane < n ? my_op : 0xFFFFFFFFu;          // sequence starts, ascending over the lanes
                const uint32_t plain = __ballot_sync(FULL_MASK, lane < n && my_ml != 0 && my_off >= my_ml);
                for (int it = 0; it < 6; ++it) {
                    const int64_t sa = xa - abs0;                              // block-relative start of my source
                    uint32_t lo = 0;
                    #pragma unroll
                    for (uint32_t st = 16u; st; st >>= 1) {                    // last lane whose sequence starts at or below sa
                        const uint32_t cand = lo + st;
                        const uint32_t v = __shfl_sync(FULL_MASK, key, cand & 31u);
                        if (cand < 32u && (int64_t)v <= sa) lo = cand;
                    }
                    const uint32_t i_mop = __shfl_sync(FULL_MASK, my_mop, lo), i_ml = __shfl_sync(FULL_MASK, my_ml, lo);
                    const int64_t i_xa = __shfl_sync(FULL_MASK, xa, lo);
                    const bool hop = ((pending >> lane) & 1u) && sa >= (int64_t)op0 && lo < lane && ((pending & plain) >> lo & 1u) &&
                                     sa >= (int64_t)i_mop && sa + (int64_t)my_span <= (int64_t)i_mop + (int64_t)i_ml;
                    if (hop) xa = i_xa + (sa - (int64_t)i_mop);
                    if (!__any_sync(FULL_MASK, hop)) break;
                }
            }
#ifdef B2_DF_PROF
            const long long q3 = clock64();
#endif
            const int64_t xb = xa + my_span;
            const uint8_t* const my_src = dp + (xa - abs0);
            uint32_t nap = DF_NAP_MIN;                                        // a waiting warp must not eat the issue slots of the one it waits for
            while (pending) {
                const bool mine = (pending >> lane) & 1u;
                const bool small = my_ml <= LANE_CAP;
                // a lane asks about short source windows itself; long ones (rare) are looked at by the whole warp, first one first
                const bool ready = mine && my_span <= 256u && df_ready(c.df, xa, xb);
                if (ready && small) {
                    DF_FENCE();
#if B2_DF_SMEMWIN
                    {   // destination in the data window; the source there too unless it is older than the window holds (or history)
                        const DfShared& S = c.df; const int64_t slo = c.df_slo;
                        const uint32_t di = df_widx(S, abs0 + my_mop), si = df_widx(S, xa);
                        const bool in_w = xa >= slo;
                        const bool simple = my_off >= my_ml && di + my_ml <= DF_DATA_BYTES && (in_w ? si + my_ml <= DF_DATA_BYTES : xb <= slo);
                        if (simple) lane_copy(S.wdata + di, in_w ? (const uint8_t*)(S.wdata + si) : (const uint8_t*)(S.gdst + xa), my_ml);
                        else {
                            uint32_t r = 0;
                            for (uint32_t k = 0; k < my_ml; ++k) { S.wdata[(di + k) & DF_DATA_MASK] = df_getb(S, slo, xa + r); r = (r + 1u == my_off) ? 0u : r + 1u; }
                        }
                    }
#elif B2_DF_PIPECOPY
                    lane_copy_pipelined(dp + my_mop, my_src, my_ml, my_off >= my_ml ? 0u : my_off);
#else
                    if (my_off >= my_ml) lane_copy(dp + my_mop, my_src, my_ml); else lane_match_from(dp + my_mop, my_src, my_off, my_ml);
#endif
                }
                const uint32_t rb = __ballot_sync(FULL_MASK, ready && small);
                uint32_t lb = __ballot_sync(FULL_MASK, ready && !small);     // long matches whose sources are there
                const uint32_t wide = __ballot_sync(FULL_MASK, mine && my_span > 256u);
                bool progressed = rb != 0u;
                if (rb) {
                    DF_FENCE();
                    if (ready && small) df_mark(c.df, (uint64_t)(abs0 + my_mop), (uint64_t)(abs0 + my_mop) + my_ml);
                    pending &= ~rb;
                }
                if (wide) {                                                 // one long source window per round: checked by all lanes
                    const int g = __ffs(wide) - 1;
                    const int64_t a = __shfl_sync(FULL_MASK, xa, g), e = __shfl_sync(FULL_MASK, xb, g);
                    if (df_ready_warp(c.df, a, e, lane)) lb |= 1u << g;
                }
                while (lb) {                                                // long matches: the whole warp copies
                    const int g = __ffs(lb) - 1; lb &= lb - 1;
                    const uint32_t W = __shfl_sync(FULL_MASK, my_mop, g), ml_g = __shfl_sync(FULL_MASK, my_ml, g), off_g = __shfl_sync(FULL_MASK, my_off, g);
                    const int64_t xa_g = __shfl_sync(FULL_MASK, xa, g);
                    const uint8_t* m = dp + (xa_g - abs0);
                    DF_FENCE();
#if B2_DF_SMEMWIN
                    {
                        const DfShared& S = c.df; const int64_t slo = c.df_slo;
                        const uint32_t di = df_widx(S, abs0 + W);
                        if (off_g >= ml_g) { for (uint32_t i = lane; i < ml_g; i += 32u) S.wdata[(di + i) & DF_DATA_MASK] = df_getb(S, slo, xa_g + i); }
                        else {
                            uint32_t r = lane % off_g; const uint32_t step = 32u % off_g;
                            for (uint32_t i = lane; i < ml_g; i += 32u) { S.wdata[(di + i) & DF_DATA_MASK] = df_getb(S, slo, xa_g + r); r += step; if (r >= off_g) r -= off_g; }
                        }
                        (void)m;
                    }
#else
                    if (off_g >= ml_g) group_copy_small(dp + W, m, ml_g, lane, 32);
                    else {                                                  // overlapping: period replication from the final window
                        uint32_t r = lane % off_g; const uint32_t step = 32u % off_g;
                        for (uint32_t i = lane; i < ml_g; i += 32) { dp[W + i] = m[r]; r += step; if (r >= off_g) r -= off_g; }
                    }
#endif
                    DF_FENCE();
                    __syncwarp();
                    df_mark_warp(c.df, (uint64_t)(abs0 + W), (uint64_t)(abs0 + W) + ml_g, lane);
                    pending &= ~(1u << g);
                    progressed = true;
                }
                __syncwarp();
#ifdef B2_DF_PROF
                ++c.prof_rounds; if (!progressed) ++c.prof_naps;
#endif
                if (!progressed) { __nanosleep(nap); nap = nap < DF_NAP_MAX ? nap * 2u : DF_NAP_MAX; } else nap = DF_NAP_MIN;
            }
#ifdef B2_DF_PROF
            { const long long q4 = clock64(); c.prof_cyc[0] += (uint32_t)(q1 - q0); c.prof_cyc[1] += (uint32_t)(q2 - q1); c.prof_cyc[2] += (uint32_t)(q3 - q2); c.prof_cyc[3] += (uint32_t)(q4 - q3); }
#endif
        }
        if constexpr (!DF && TILE != 0u) { T.tv = T.fp = t_end = op; }      // this batch went straight to global memory: the tile is empty
    }
    if constexpr (!DF && TILE != 0u) { __syncwarp(); tile_flush(T, dp, t_end, true, lane); __syncwarp(); }
    c.ip = ip; c.op = op;
    return err ? err : (paused ? DEC_PAUSED : 0u);
}

}  // namespace b2
