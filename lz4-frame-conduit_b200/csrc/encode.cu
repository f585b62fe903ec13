// encode.cu — LZ4 frame encoding on sm_100a: warp-cooperative hash match finder + sequence emitter,
// block layout (prefix sums, stored-block decision, frame header / EndMark / index) and gather.
//
// Replaces, on this path (citations relative to /root/reference/):
//   LZ4_compress_generic (greedy hash match finder + token emitter)   lz4/lib/lz4.c:789-1177
//   LZ4F_makeBlock (block header word, stored-block fallback)         lz4/lib/lz4frame.c:739-761
//   LZ4F_compressBegin / LZ4F_compressEnd (header, EndMark)           lz4frame.c:600-702, 984-1018
// This is NOT a port of that code: the reference walks one position at a time with a 4 K-entry table;
// here one warp probes 128 positions per step against a 4 K-entry shared-memory table and weighs the
// 128 measured matches against each other (lazy evaluation) before it picks, a 4 MB frame block is cut
// into independent sub-blocks (match domains) parsed by several warps at once, 16 KB each, whose outputs
// are stitched by prefix sums,
// and the only things kept from the reference are the *format rules* every LZ4 encoder must obey
// (lz4_Block_format.md:108-129: last 5 bytes literal, last match starts >= 12 bytes before the end,
// inputs < 13 bytes are literal-only, offsets <= 65535) and the stored-block rule of the frame layer
// (a block is stored raw iff it does not compress to fewer bytes, lz4frame.c:746-752).
#include <type_traits>
#include "common.cuh"
#include "xxh32_cta.cuh"
#include "kernels.h"

namespace b2 {

constexpr uint32_t HSIZE = 1u << ENC_HLOG;

#ifndef B2_ENC_MIN_SL
#define B2_ENC_MIN_SL 0       // smallest probe stride (log2): 1 = every other position even in compressible input (tuning experiment)
#endif
#ifndef B2_ENC_WIN
#define B2_ENC_WIN 0          // 1 = the 128 positions of a stride-1 step take their own bytes from a shared-memory copy of the step window
                              // (measured on B200: 48.9 vs 58.9 GB/s — the kernel is bound by instruction issue, the window costs instructions)
#endif
constexpr uint32_t ENC_WIN_WORDS = 44;     // bytes T0 - 4 .. T0 + 127 + 12 + 19 (+ alignment) of a step: what probe, verify and the second measuring pass read at the positions
#ifndef B2_ENC_COMPACT2B
#define B2_ENC_COMPACT2B 0    // 1 = the second measuring pass (2b) works on a dense list of the matches that need it, 32 per round, instead of one
                              // per lane per round at ~30 % lane use (a 128-byte exchange area per warp in shared memory).  Same bytes out; measured
                              // on B200: 5.4 % fewer instructions, 11 % SLOWER (52.1 vs 58.8 GB/s: the exchange puts two shared-memory round trips
                              // and a global one on the step's critical path; issue slots 65 % -> 54 %)
#endif
constexpr uint32_t ENC_X_WORDS = B2_ENC_COMPACT2B ? 32u : 0u;      // exchange words per warp
size_t find_emit_smem_bytes() {
    return (size_t)ENC_SMEM_WARPS * HSIZE * sizeof(uint16_t) + (B2_ENC_WIN ? (size_t)ENC_WARPS * ENC_WIN_WORDS * 4u : 0u)
         + (size_t)ENC_WARPS * ENC_X_WORDS * 4u;
}

// the aligned word that holds *pp: in the step window (shared memory) or in global memory
template <bool WIN> __device__ __forceinline__ const uint32_t* pos_words(const uint8_t* pp, const uint32_t* wwin, uintptr_t gA) {
    if (WIN) return wwin + ((((uintptr_t)pp & ~(uintptr_t)3) - gA) >> 2);
    return (const uint32_t*)((uintptr_t)pp & ~(uintptr_t)3);
}
size_t find_emit_gtab_bytes(uint32_t grid) { return (size_t)grid * (ENC_WARPS - ENC_SMEM_WARPS) * HSIZE * sizeof(uint16_t); }

#if B2_PTX
__device__ __forceinline__ uint32_t lds_u16(uint32_t saddr) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(saddr) : "memory"); return v; }
__device__ __forceinline__ void sts_u16(uint32_t saddr, uint16_t v) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(saddr), "h"(v) : "memory"); }
#endif
__device__ __forceinline__ uint32_t hash4(uint32_t v) { return (v * 2654435761u) >> (32 - ENC_HLOG); }

__device__ __forceinline__ uint32_t len_bytes(uint32_t v) {      // extra length bytes for a field value v
    return v >= 15u ? (v - 15u) / 255u + 1u : 0u;
}

// token nibble already written by the caller; writes the 255-run + remainder bytes, returns new offset
__device__ __forceinline__ uint32_t put_len_ext(uint8_t* out, uint32_t o, uint32_t v, uint32_t lane) {
    if (v < 15u) return o;
    uint32_t r = v - 15u, n255 = r / 255u;
    for (uint32_t i = lane; i < n255; i += 32) out[o + i] = 255;
    if (lane == 0) out[o + n255] = (uint8_t)(r - n255 * 255u);
    return o + n255 + 1u;
}

#ifndef B2_ENC_PRIME_STRIDE
#define B2_ENC_PRIME_STRIDE 4
#endif
constexpr uint32_t ENC_PRIME_STRIDE = B2_ENC_PRIME_STRIDE;   // priming inserts every n-th earlier position of the domain
#ifndef B2_ENC_MINLEN
#define B2_ENC_MINLEN 4
#endif
constexpr uint32_t ENC_MINLEN = B2_ENC_MINLEN;               // matches shorter than this (forward + backward bytes seen at verify time) are ignored
#ifndef B2_ENC_LAZY
#define B2_ENC_LAZY 32
#endif
constexpr uint32_t ENC_LAZY = B2_ENC_LAZY;                   // matches shorter than this give way to a longer one 1..3 positions later (0 = off)
constexpr uint32_t ENC_PRIME_NEAR = 4096u;                   // history primed before the unit is known to be compressible
constexpr uint32_t ENC_PREMEASURE = 28u;                  // matches are measured lane-parallel up to this length
constexpr uint32_t NX_LONG = 254u, NX_NONE = 255u;     // link values: match end unknown yet / no further hit

// First hit (bit set in the 128-bit map H0..H3) at index >= j, or NX_NONE.  a1/a2/a3 = first hit in words 1..3 / 2..3 / 3.
__device__ __forceinline__ uint32_t first_hit_from(uint32_t j, uint32_t H0, uint32_t H1, uint32_t H2, uint32_t H3,
                                                   uint32_t a1, uint32_t a2, uint32_t a3) {
    if (j >= 128u) return NX_NONE;
    const uint32_t w = j >> 5;
    const uint32_t m = (w == 0 ? H0 : w == 1 ? H1 : w == 2 ? H2 : H3) & (0xFFFFFFFFu << (j & 31u));
    if (m) return (w << 5) + (uint32_t)__ffs(m) - 1u;
    return w == 0 ? a1 : w == 1 ? a2 : w == 2 ? a3 : NX_NONE;
}

// bits [j0, j1) of a 128-bit map, restricted to the 32-bit word that starts at bit `w0`
__device__ __forceinline__ uint32_t range_mask(uint32_t j0, uint32_t j1, uint32_t w0) {
    // clamp both ends into 0..32 relative to the word, then (2^hi - 1) & ~(2^lo - 1) in 64-bit so that 32 works
    const uint32_t lo = min(j0 > w0 ? j0 - w0 : 0u, 32u), hi = min(j1 > w0 ? j1 - w0 : 0u, 32u);
    return (uint32_t)(((1ull << hi) - 1ull) & ~((1ull << lo) - 1ull));
}

constexpr uint32_t ENC_LANE_CAP = 32;     // literal runs up to this many bytes are copied one-lane-per-sequence

// Up to 32 parsed sequences wait here, one per lane, until the batch is written out.
struct Pending { uint32_t anchor, ll, off, mlc; };

// Writes `npend` sequences (lane i holds sequence i) at out + o: token, literal-length bytes, literals, offset,
// match-length bytes (lz4_Block_format.md:27-105).  Sizes are prefix-summed so every lane writes its own
// sequence; only literal runs longer than ENC_LANE_CAP are copied by the whole warp.  Returns the new o.
__device__ __forceinline__ uint32_t emit_batch(uint8_t* __restrict__ out, uint32_t o, const uint8_t* __restrict__ base,
                                               uint32_t npend, const Pending& pd, uint32_t lane)
{
    const bool act = lane < npend;
    const uint32_t hdr = 1u + len_bytes(pd.ll);
    const uint32_t sz = act ? hdr + pd.ll + 2u + len_bytes(pd.mlc) : 0u;
    uint32_t x = sz;
    #pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(FULL_MASK, x, d); if ((int)lane >= d) x += y; }
    const uint32_t my_o = o + x - sz;
    const uint32_t total = __shfl_sync(FULL_MASK, x, 31);
    if (act) {
        uint8_t* w = out + my_o;
        *w++ = (uint8_t)(((pd.ll < 15u ? pd.ll : 15u) << 4) | (pd.mlc < 15u ? pd.mlc : 15u));
        if (pd.ll >= 15u) { uint32_t r = pd.ll - 15u; while (r >= 255u) { *w++ = 255; r -= 255u; } *w++ = (uint8_t)r; }
        if (pd.ll <= ENC_LANE_CAP) {
            const uint8_t* l = base + pd.anchor;
            uint32_t k = 0;
            for (; k + 4 <= pd.ll; k += 4) { uint8_t a = l[k], b = l[k + 1], c = l[k + 2], e = l[k + 3]; w[k] = a; w[k + 1] = b; w[k + 2] = c; w[k + 3] = e; }
            for (; k < pd.ll; ++k) w[k] = l[k];
        }
        w += pd.ll;
        w[0] = (uint8_t)pd.off; w[1] = (uint8_t)(pd.off >> 8); w += 2;
        if (pd.mlc >= 15u) { uint32_t r = pd.mlc - 15u; while (r >= 255u) { *w++ = 255; r -= 255u; } *w++ = (uint8_t)r; }
    }
    uint32_t longm = __ballot_sync(FULL_MASK, act && pd.ll > ENC_LANE_CAP);
    while (longm) {
        const int j = __ffs(longm) - 1; longm &= longm - 1;
        const uint32_t a = __shfl_sync(FULL_MASK, pd.anchor, j), l = __shfl_sync(FULL_MASK, pd.ll, j);
        const uint32_t at = __shfl_sync(FULL_MASK, my_o + hdr, j);
        group_copy(out + at, base + a, l, lane, 32);
    }
    return o + total;
}

// ------------------------------------------------------------------------------------------------
// match finder + emitter: one warp per independent sub-block, 128 positions per step.
//   1. probe   every lane hashes 4 positions and reads their table entries (all 128 look-ups happen before
//              any insert of the step, so 128 candidate loads are in flight together)
//   2. verify  4-byte compare at the candidates -> a 128-bit hit map (4 ballots)
//   3. select  greedy, in position order: first hit at or after the parse position, extended backwards
//              to the pending literals and forwards as far as it goes (warp-wide compares); the sequence
//              is parked in a lane (Pending) and written out 32 at a time (emit_batch)
//   4. insert  positions that did not end up inside a match go into the table (inserting match
//              interiors would evict the field starts the following lines need)
// After a step without a hit the probe stride doubles (1, 2, 4, 8 bytes; 16 and 32 after 8 and 16 empty steps),
// as incompressible data should cost little; a hit resets it.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ENC_WARPS * 32, ENC_CTAS_PER_SM)
find_emit_kernel(const uint8_t* __restrict__ src, EncParams P, uint8_t* __restrict__ scratch,
                 SubOut* __restrict__ subs, uint32_t* __restrict__ ticket, uint16_t* __restrict__ gtab)
{
    B2_DYN_SHARED(uint16_t, s_tab);
    const uint32_t lane = lane_id();
    // ENC_SMEM_WARPS (6 of 8) warps keep their hash table in shared memory, the others in a global table that
    // lives in L1/L2: 32 warps per SM fit that way (192 KB of tables), and measured against 24 warps with every
    // table in shared memory the occupancy wins (59 vs 55.5 GB/s)
    const uint32_t wid = threadIdx.x >> 5;
#if B2_ENC_WIN
    uint32_t* const wwin = (uint32_t*)(s_tab + (size_t)ENC_SMEM_WARPS * HSIZE) + wid * ENC_WIN_WORDS;   // this warp's step window
#else
    uint32_t* const wwin = nullptr;
#endif
    [[maybe_unused]] uint32_t* const xch = (uint32_t*)(s_tab + (size_t)ENC_SMEM_WARPS * HSIZE) + (B2_ENC_WIN ? ENC_WARPS * ENC_WIN_WORDS : 0u) + wid * ENC_X_WORDS;
    uint16_t* tab = wid < (uint32_t)ENC_SMEM_WARPS ? s_tab + wid * HSIZE
                  : gtab + ((size_t)blockIdx.x * (ENC_WARPS - ENC_SMEM_WARPS) + (wid - ENC_SMEM_WARPS)) * HSIZE;

    // tuning switch: read and write the global tables past L1 (.cg) so that they do not evict the input window there
#ifndef B2_ENC_TAB_CG
#define B2_ENC_TAB_CG 0      // measured: .cg table accesses 45.7 GB/s, cached ones 47.3 GB/s
#endif
    const bool tab_global = wid >= (uint32_t)ENC_SMEM_WARPS;
#ifndef B2_ENC_TABSPLIT
#define B2_ENC_TABSPLIT 0     // (measured: 55.3 vs 58.9 GB/s)  1 = every table access is written twice, through a shared and through a global pointer (warp-uniform
                              // choice): LDS / STS and LDG / STG instead of generic LD / ST through one pointer
#endif
#if B2_ENC_TABSPLIT && B2_PTX
    const uint32_t stab = (uint32_t)__cvta_generic_to_shared(s_tab + (wid < (uint32_t)ENC_SMEM_WARPS ? wid : 0u) * HSIZE);   // (only used when !tab_global)
    #define TAB_LD(p) (tab_global ? (uint32_t)*(p) : lds_u16(stab + 2u * (uint32_t)((p) - tab)))
    #define TAB_ST(p, v) do { if (tab_global) *(p) = (v); else sts_u16(stab + 2u * (uint32_t)((p) - tab), (uint16_t)(v)); } while (0)
#else
    #define TAB_LD(p) ((B2_ENC_TAB_CG && tab_global) ? (uint32_t)__ldcg((const unsigned short*)(p)) : (uint32_t)*(p))
    #define TAB_ST(p, v) do { if (B2_ENC_TAB_CG && tab_global) __stcg((unsigned short*)(p), (unsigned short)(v)); else *(p) = (v); } while (0)
#endif
    for (;;) {
        uint32_t s = 0;
        if (lane == 0) s = atomicAdd(ticket, 1u);
        s = __shfl_sync(FULL_MASK, s, 0);
        if (s >= P.n_subs) break;

        const uint32_t blk = s / P.subs_per_block, j = s - blk * P.subs_per_block;
        const uint64_t blk_start = (uint64_t)blk * P.block_size;
        // linked blocks (and dictionaries, LZ4F_compressBegin_usingCDict, lz4frame.c:594-702): the bytes in front of the block
        // — earlier blocks of the stream, the dictionary in front of the first one — are part of the match domain.  Every
        // position below is counted from `base`, which then lies hb bytes in front of the block.
        uint32_t hb = 0;
        if (P.linked) { const uint64_t h = blk_start + P.hist; hb = h > 65536ull ? 65536u : (uint32_t)h; }
        const uint32_t usize = hb + (uint32_t)((P.n - blk_start) < P.block_size ? (P.n - blk_start) : P.block_size);
        // the warp parses the unit [b0, b1); its matches may reach back to d0: the start of the match domain the unit
        // belongs to (dom_units units = the piece ONE decoder warp decodes, so nothing may reach before it), or — linked —
        // 64 KB in front of the unit wherever that lies
        const uint32_t b0 = hb + j * P.sub_size;
        const uint32_t b1 = (b0 + P.sub_size < usize) ? b0 + P.sub_size : usize;
        const uint32_t d0 = P.linked ? (b0 > 65536u ? b0 - 65536u : 0u) : (j / P.dom_units) * P.dom_units * P.sub_size;
        const bool last_sub = (b1 == usize);
        const uint8_t* __restrict__ base = src + blk_start - hb;
        uint8_t* out = scratch + (uint64_t)s * P.scratch_stride;

        // zero the table so the parse is deterministic, then prime it with the part of the domain in front of the unit:
        // every ENC_PRIME_STRIDE-th earlier position (at most 65535 back) is inserted, the later one winning — what a serial parse of the
        // domain would have left there, only denser.  That is what lets the units of one domain be parsed by different
        // warps at the same time without paying for it in ratio (tools/enc_model.c: SEG / PRIME).
        { uint4 z = make_uint4(0, 0, 0, 0); uint4* t4 = (uint4*)tab;
          for (uint32_t i = lane; i < HSIZE * 2 / 16; i += 32) t4[i] = z; }
        __syncwarp();
        if (b0 > d0) {
            // priming 64 KB costs more than parsing a unit of incompressible data, so it is done in two steps: the newest
            // 4 KB first; if none of the unit's first 256 positions finds a match there, the unit is taken to be
            // incompressible and parsed with that much history only (its own matches are still found)
            const uint32_t from = (b0 - d0 > 65535u) ? b0 - 65535u : d0;
            const uint32_t near = (b0 - from > ENC_PRIME_NEAR) ? b0 - ENC_PRIME_NEAR : from;
            auto prime = [&](uint32_t lo) {
                for (uint32_t q = lo; q < b0; q += 128u * ENC_PRIME_STRIDE) {
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t pp = q + (32u * k + lane) * ENC_PRIME_STRIDE;
                        const uint32_t vv = pp < b0 ? ld32u_ro(base + pp) : 0u;
                        // inside a run the neighbouring position has the same hash: only the later one needs the slot
                        const uint32_t vn = __shfl_down_sync(FULL_MASK, vv, 1);
                        if (pp < b0 && (lane == 31u || vn != vv || pp + ENC_PRIME_STRIDE >= b0)) TAB_ST(tab + hash4(vv), (uint16_t)(pp - d0));
                    }
                    __syncwarp();
                }
            };
            prime(near);
            if (near > from) {
                bool hit = false;
                #pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t pp = b0 + 32u * k + lane;
                    if (pp + 4u <= b1) {
                        const uint32_t vv = ld32u_ro(base + pp);
                        const uint32_t cl = TAB_LD(tab + hash4(vv));
                        int32_t cc = (int32_t)(((pp - d0) & ~0xFFFFu) | cl);
                        if ((uint32_t)cc >= pp - d0) cc -= 65536;
                        if (cc >= 0 && ld32u_ro(base + d0 + cc) == vv) hit = true;
                    }
                }
                if (__any_sync(FULL_MASK, hit)) {                           // compressible: the whole history, oldest first
                    { uint4 z = make_uint4(0, 0, 0, 0); uint4* t4 = (uint4*)tab;
                      for (uint32_t i = lane; i < HSIZE * 2 / 16; i += 32) t4[i] = z; }
                    __syncwarp();
                    prime(from);
                }
            }
        }

        // positions that may start a match: p + 4 <= b1 and p <= usize - 12 (block end rule)
        uint32_t limit = 0;
        if (usize - hb >= 13u) {                                    // (the block's own size decides: lz4_Block_format.md:108-129)
            uint32_t la = b1 >= 3u ? b1 - 3u : 0u, lb = usize - 11u;
            limit = la < lb ? la : lb;
        }
        const uint32_t matchlimit = (usize - hb >= 5u) ? ((b1 < usize - 5u) ? b1 : usize - 5u) : 0u;

        uint32_t ip = b0, anchor = b0, o = 0, T0 = b0;
        uint32_t miss = 0, nseq = 0, first_ll = 0, first_hdr = 0, npend = 0;
        Pending pd = {0, 0, 0, 0};

        while (T0 < limit) {
            // probe stride 1, 2, 4, 8 after 0..3 steps without a hit, 16 after 8, 32 after 16 (the reference's skip
            // acceleration grows the same way: one more byte per 64 failed attempts, lz4.c:895-905)
            uint32_t sl = miss < 3u ? miss : miss < 8u ? 3u : miss < 16u ? 4u : 5u;
#if B2_ENC_MIN_SL > 0
            if (sl < (uint32_t)B2_ENC_MIN_SL) sl = (uint32_t)B2_ENC_MIN_SL;
#endif
            const uint32_t span = 128u << sl;
            if (lane < 2u && T0 + span + 128u * (lane + 1u) <= usize) prefetch_l1(base + T0 + span + 128u * lane);   // next step's input
            // ---- 0. stride-1 steps (compressible input: almost all of them): the bytes at and around the 128 positions are loaded
            // ONCE, coalesced, into the warp's window in shared memory; probe, verify and the second measuring pass read their
            // position-side words there (44 shared-memory words instead of 28 global loads per lane).  Words are fetched whole, so
            // only words that hold a byte of [base, base + usize) are touched (same-word slack, as everywhere in this kernel).
            const bool use_win = B2_ENC_WIN && sl == 0u;
            uintptr_t gA = 0;
            if (use_win) {
                gA = ((uintptr_t)(base + T0) - 4u) & ~(uintptr_t)3;
                const uintptr_t lo = (uintptr_t)base & ~(uintptr_t)3, hi = (uintptr_t)(base + usize);
                __syncwarp();
                for (uint32_t w = lane; w < ENC_WIN_WORDS; w += 32u) {
                    const uintptr_t a = gA + 4u * w;
                    wwin[w] = (a >= lo && a < hi) ? ldg32((const uint32_t*)a) : 0u;
                }
                __syncwarp();
            }
            // ---- 1. probe
            uint32_t v[4], h[4], rel[4]; int32_t c[4]; bool ok[4];
            #pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t p = T0 + ((32u * k + lane) << sl);
                ok[k] = p < limit;
                rel[k] = p - d0;
                if (use_win) {
                    const uint8_t* pp = base + p;
                    const uint32_t* w = pos_words<true>(pp, wwin, gA);
                    const uint32_t sh = (uint32_t)((uintptr_t)pp & 3u) * 8u;
                    v[k] = ok[k] ? __funnelshift_r(w[0], w[1], sh) : 0u;
                } else
                v[k] = ok[k] ? ld32u_ro(base + p) : 0u;
                h[k] = hash4(v[k]);
            }
            #pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t cl = TAB_LD(tab + h[k]);
                int32_t cc = (int32_t)((rel[k] & ~0xFFFFu) | cl);
                if ((uint32_t)cc >= rel[k]) cc -= 65536;
                c[k] = (ok[k] && cc >= 0 && rel[k] - (uint32_t)cc < 65536u) ? cc : -1;
            }
            // ---- 2. verify: one round trip fetches 4 bytes before and 12 bytes after every candidate, so that
            // most matches are fully measured here (forward length up to 12, backward up to 4) and the serial
            // selection below rarely has to touch memory
            uint32_t H[4], pk[4];
            auto verify = [&](auto WINC) {
            constexpr bool WIN = decltype(WINC)::value;
            #define PW(i) (WIN ? pw[(i)] : ldg32(pw + (i)))      /* window: shared memory; else global */
            #pragma unroll
            for (int k = 0; k < 4; ++k) {
                bool hit = false;
                pk[k] = 0;
                if (c[k] >= 0) {
                    const uint8_t* cp = base + d0 + c[k];
                    const uint8_t* pp = base + d0 + rel[k];
                    const uint32_t* cw = (const uint32_t*)((uintptr_t)cp & ~(uintptr_t)3);
                    const uint32_t* pw = pos_words<WIN>(pp, wwin, gA);
                    const uint32_t csh = (uint32_t)((uintptr_t)cp & 3u) * 8u, psh = (uint32_t)((uintptr_t)pp & 3u) * 8u;
                    const bool bk_ok = c[k] >= 4;                             // 4 bytes in front of the candidate are inside the sub-block
                    const uint32_t c0 = ldg32(cw), c1 = ldg32(cw + 1), c2 = ldg32(cw + 2), c3 = csh ? ldg32(cw + 3) : 0u, cm = bk_ok ? ldg32(cw - 1) : 0u;
                    const uint32_t p0 = PW(0), p1 = PW(1), p2 = PW(2), p3 = psh ? PW(3) : 0u, pm = bk_ok ? PW(-1) : 0u;
                    const uint32_t x1 = __funnelshift_r(c0, c1, csh) ^ __funnelshift_r(p0, p1, psh);
                    if (x1 == 0u) {
                        hit = true;
                        const uint32_t x2 = __funnelshift_r(c1, c2, csh) ^ __funnelshift_r(p1, p2, psh);
                        const uint32_t x3 = __funnelshift_r(c2, c3, csh) ^ __funnelshift_r(p2, p3, psh);
                        uint32_t fl = 4u + (x2 ? ((uint32_t)(__ffs(x2) - 1) >> 3) : 4u + (x3 ? ((uint32_t)(__ffs(x3) - 1) >> 3) : 4u));
                        const uint32_t room = matchlimit - (d0 + rel[k]);     // >= 7 for every probed position
                        if (fl > room) fl = room;
                        uint32_t bk = 0;
                        if (bk_ok) {
                            const uint32_t x0 = __funnelshift_r(cm, c0, csh) ^ __funnelshift_r(pm, p0, psh);
                            bk = x0 ? ((uint32_t)__clz(x0) >> 3) : 4u;
                        }
                        pk[k] = (uint32_t)c[k] | (fl << 22) | (bk << 27);
                        // short matches are not worth a sequence: 3 bytes of token + offset, a broken literal run, and the
                        // greedy parse would miss the longer match that often starts a byte or two later (enc_model MINLEN)
                        if (fl + bk < ENC_MINLEN) { hit = false; pk[k] = 0; }
                    }
                }
                H[k] = __ballot_sync(FULL_MASK, hit);
            }
            // ---- 2b. matches that filled the first 12 bytes are measured for 16 more
            // (41 % of the matches of log-like data are 12+ bytes, 5 % are 28+; only those reach the warp-wide loop)
            // `measure16`: bytes 12.. of the match (packed pv) at relative position rl -> pv with its length field updated
            auto measure16 = [&](uint32_t pv, uint32_t rl) -> uint32_t {
                const uint32_t room = matchlimit - (d0 + rl);
                if (room <= 12u) return pv;
                const uint8_t* cp = base + d0 + (pv & 0x3FFFFFu) + 12u;
                const uint8_t* pp = base + d0 + rl + 12u;
                const uint32_t* cw = (const uint32_t*)((uintptr_t)cp & ~(uintptr_t)3);
                const uint32_t* pw = pos_words<WIN>(pp, wwin, gA);
                const uint32_t csh = (uint32_t)((uintptr_t)cp & 3u) * 8u, psh = (uint32_t)((uintptr_t)pp & 3u) * 8u;
                // only words that hold bytes below matchlimit + 3 are touched (same-word slack only)
                const uint32_t nw = (room - 12u + 3u) >> 2;                      // words of interest: 1..4+
                const uint32_t c0 = ldg32(cw), p0 = PW(0);
                const uint32_t c1 = ldg32(cw + 1), p1 = PW(1);
                const uint32_t c2 = nw > 1u ? ldg32(cw + 2) : 0u, p2 = nw > 1u ? PW(2) : 0u;
                const uint32_t c3 = nw > 2u ? ldg32(cw + 3) : 0u, p3 = nw > 2u ? PW(3) : 0u;
                const uint32_t c4 = (nw > 3u && csh) ? ldg32(cw + 4) : 0u, p4 = (nw > 3u && psh) ? PW(4) : 0u;
                const uint32_t x0 = __funnelshift_r(c0, c1, csh) ^ __funnelshift_r(p0, p1, psh);
                const uint32_t x1 = __funnelshift_r(c1, c2, csh) ^ __funnelshift_r(p1, p2, psh);
                const uint32_t x2 = __funnelshift_r(c2, c3, csh) ^ __funnelshift_r(p2, p3, psh);
                const uint32_t x3 = __funnelshift_r(c3, c4, csh) ^ __funnelshift_r(p3, p4, psh);
                uint32_t more = x0 ? ((uint32_t)(__ffs(x0) - 1) >> 3)
                              : x1 ? 4u + ((uint32_t)(__ffs(x1) - 1) >> 3)
                              : x2 ? 8u + ((uint32_t)(__ffs(x2) - 1) >> 3)
                              : x3 ? 12u + ((uint32_t)(__ffs(x3) - 1) >> 3) : 16u;
                if (12u + more > room) more = room - 12u;
                return (pv & ~(31u << 22)) | ((12u + more) << 22);
            };
#if B2_ENC_COMPACT2B
            {
                // the matches that need it are numbered across the warp (position order: k major, lane minor) and handled 32 at a
                // time, one per lane: owner -> exchange word "who owns item i" -> the lane that works on item i fetches the packed
                // match with shuffles, measures, and hands the result back through the same words
                bool need[4]; uint32_t idx[4]; uint32_t tot = 0;
                #pragma unroll
                for (int k = 0; k < 4; ++k) {
                    need[k] = ((pk[k] >> 22) & 31u) == 12u;
                    const uint32_t Bk = __ballot_sync(FULL_MASK, need[k]);
                    idx[k] = tot + (uint32_t)__popc(Bk & ((1u << lane) - 1u));
                    tot += (uint32_t)__popc(Bk);
                }
                const uint32_t rel0 = T0 - d0;
                for (uint32_t r0 = 0; r0 < tot; r0 += 32u) {
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) if (need[k] && idx[k] - r0 < 32u) xch[idx[k] - r0] = ((uint32_t)k << 5) | lane;
                    __syncwarp();
                    const bool valid = r0 + lane < tot;
                    const uint32_t code = valid ? xch[lane] : 0u;
                    const uint32_t own = code & 31u, kk = code >> 5;
                    const uint32_t g0 = __shfl_sync(FULL_MASK, pk[0], own), g1 = __shfl_sync(FULL_MASK, pk[1], own);
                    const uint32_t g2 = __shfl_sync(FULL_MASK, pk[2], own), g3 = __shfl_sync(FULL_MASK, pk[3], own);
                    uint32_t pv = kk == 0 ? g0 : kk == 1 ? g1 : kk == 2 ? g2 : g3;
                    if (valid) pv = measure16(pv, rel0 + ((32u * kk + own) << sl));
                    __syncwarp();
                    xch[lane] = pv;
                    __syncwarp();
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) if (need[k] && idx[k] - r0 < 32u) pk[k] = xch[idx[k] - r0];
                    __syncwarp();
                }
            }
#else
            {
                uint32_t todo = 0;
                #pragma unroll
                for (int k = 0; k < 4; ++k) if (((pk[k] >> 22) & 31u) == 12u) todo |= 1u << k;
                while (__any_sync(FULL_MASK, todo != 0u)) {
                    if (todo) {
                        const uint32_t k = (uint32_t)__ffs(todo) - 1u;
                        todo &= todo - 1u;
                        const uint32_t pv = k == 0 ? pk[0] : k == 1 ? pk[1] : k == 2 ? pk[2] : pk[3];
                        const uint32_t rl = k == 0 ? rel[0] : k == 1 ? rel[1] : k == 2 ? rel[2] : rel[3];
                        const uint32_t nv = measure16(pv, rl);
                        if (k == 0) pk[0] = nv; else if (k == 1) pk[1] = nv; else if (k == 2) pk[2] = nv; else pk[3] = nv;
                    }
                }
            }
#endif
            #undef PW
            };
            if (use_win) verify(std::true_type{}); else verify(std::false_type{});
            // ---- 2c. lazy evaluation, lane-parallel: a short match is dropped when one of the next three positions holds a
            // match at least two bytes longer — the greedy parse would otherwise take the short one and step over the
            // start of the long one (tools/enc_model.c LAZY: ratio +4 % on log-like, +1 % on text, 15 % fewer sequences)
            if (ENC_LAZY && sl == 0u) {
                #pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t f = pk[k] ? ((pk[k] >> 22) & 31u) : 0u;
                    const uint32_t t = f + (pk[k] >> 27);
                    uint32_t best = 0;
                    #pragma unroll
                    for (int d = 1; d <= 3; ++d) {
                        const uint32_t nf = __shfl_down_sync(FULL_MASK, f, d);
                        if (lane + d < 32u && nf > best) best = nf;
                    }
                    if (f && t < ENC_LAZY && best >= t + 2u) pk[k] = 0;
                    H[k] = __ballot_sync(FULL_MASK, pk[k] != 0u);
                }
            }
            uint32_t C0 = 0, C1 = 0, C2 = 0, C3 = 0;                          // samples swallowed by a match (uniform)
            if ((H[0] | H[1] | H[2] | H[3]) != 0u) {
                miss = 0;
                // ---- 3a. link: every hit computes where the greedy parse goes next if it is taken — the first hit
                // at or after its end (lane-parallel; ends of 12+-byte matches are not known yet -> NX_LONG)
                const uint32_t a3 = H[3] ? 95u + (uint32_t)__ffs(H[3]) : NX_NONE;          // first hit in word 3
                const uint32_t a2 = H[2] ? 63u + (uint32_t)__ffs(H[2]) : a3;               // ... in words 2..3
                const uint32_t a1 = H[1] ? 31u + (uint32_t)__ffs(H[1]) : a2;               // ... in words 1..3
                uint32_t NX = 0;
                #pragma unroll
                for (int k = 0; k < 4; ++k) {
                    uint32_t nx = NX_NONE;
                    if (pk[k]) {
                        const uint32_t fl = (pk[k] >> 22) & 31u;
                        if (fl >= ENC_PREMEASURE) nx = NX_LONG;
                        else nx = first_hit_from(32u * k + lane + ((fl + (1u << sl) - 1u) >> sl), H[0], H[1], H[2], H[3], a1, a2, a3);
                    }
                    NX |= nx << (8 * k);
                }
                // ---- 3b. select: follow the links from the first hit; lane `npend` takes the next match
                uint32_t cur = H[0] ? (uint32_t)__ffs(H[0]) - 1u : a1;
                while (cur != NX_NONE) {
                    const uint32_t s0 = npend;
                    uint32_t my_idx = 0, nx;
                    for (;;) {
                        nx = (__shfl_sync(FULL_MASK, NX, cur & 31u) >> ((cur >> 5) * 8u)) & 255u;
                        if (lane == npend) my_idx = cur;
                        ++npend;
                        if (nx >= NX_LONG || npend == 32u) break;
                        cur = nx;
                    }
                    // ---- 3c. the lanes s0 .. npend-1 turn their match into a sequence
                    const bool mine = lane >= s0 && lane < npend;
                    const uint32_t own = my_idx & 31u, kk = my_idx >> 5;
                    const uint32_t g0 = __shfl_sync(FULL_MASK, pk[0], own), g1 = __shfl_sync(FULL_MASK, pk[1], own);
                    const uint32_t g2 = __shfl_sync(FULL_MASK, pk[2], own), g3 = __shfl_sync(FULL_MASK, pk[3], own);
                    const uint32_t pkv = kk == 0 ? g0 : kk == 1 ? g1 : kk == 2 ? g2 : g3;
                    const uint32_t pos = T0 + (my_idx << sl);
                    uint32_t fl = (pkv >> 22) & 31u;
                    uint32_t mend = pos + fl;
                    if (nx == NX_LONG) {
                        // the last match taken is 12+ bytes: measure it with the whole warp (4 bytes per lane per round;
                        // bytes at and after matchlimit never count)
                        const uint32_t lp = __shfl_sync(FULL_MASK, pos, npend - 1u);
                        const uint32_t lc = d0 + (__shfl_sync(FULL_MASK, pkv, npend - 1u) & 0x3FFFFFu);
                        uint32_t len = ENC_PREMEASURE;
                        for (uint32_t round = 0;; ++round) {
                            uint32_t nbytes = 0, per;
                            if (round == 0) {                                 // most long matches end within 128 more bytes
                                per = 4u;
                                const uint32_t qq = lp + len + lane * 4u;
                                const uint32_t avail = matchlimit > qq ? matchlimit - qq : 0u;
                                if (avail) {
                                    const uint32_t xx = ld32u_ro(base + qq) ^ ld32u_ro(base + lc + len + lane * 4u);
                                    nbytes = xx ? (uint32_t)(__ffs(xx) - 1) >> 3 : 4u;
                                    if (nbytes > avail) nbytes = avail;
                                }
                            } else {                                          // runs: 16 bytes per lane, 512 per round
                                per = 16u;
                                const uint32_t qq = lp + len + lane * 16u;
                                const uint32_t avail = matchlimit > qq ? matchlimit - qq : 0u;
                                if (avail) {
                                    const uint8_t* pa = base + qq;
                                    const uint8_t* pb = base + lc + len + lane * 16u;
                                    nbytes = 16u;
                                    #pragma unroll
                                    for (int j = 3; j >= 0; --j) {
                                        if (4u * j < avail) {
                                            const uint32_t xx = ld32u_ro(pa + 4 * j) ^ ld32u_ro(pb + 4 * j);
                                            if (xx) nbytes = 4u * j + ((uint32_t)(__ffs(xx) - 1) >> 3);
                                        }
                                    }
                                    if (nbytes > avail) nbytes = avail;
                                }
                            }
                            const uint32_t part = __ballot_sync(FULL_MASK, nbytes != per);
                            if (part == 0) { len += 32u * per; continue; }
                            const int g = __ffs(part) - 1;
                            len += (uint32_t)g * per + __shfl_sync(FULL_MASK, nbytes, g);
                            break;
                        }
                        if (lane == npend - 1u) { fl = len; mend = pos + len; }
                    }
                    uint32_t prev_end = __shfl_up_sync(FULL_MASK, mend, 1);
                    if (lane == s0) prev_end = anchor;
                    uint32_t j0 = 0, j1 = 0;
                    if (mine) {
                        uint32_t back = pkv >> 27;
                        if (back > pos - prev_end) back = pos - prev_end;
                        const uint32_t mpos = pos - back;
                        pd.anchor = prev_end; pd.ll = mpos - prev_end; pd.off = pos - (d0 + (pkv & 0x3FFFFFu)); pd.mlc = fl + back - 4u;
                        // sample indices strictly inside the match: ((mpos - T0) >> sl) + 1 .. first index at or after mend
                        j0 = mpos >= T0 ? ((mpos - T0) >> sl) + 1u : 0u;
                        j1 = (mend - T0 + (1u << sl) - 1u) >> sl;
                    }
                    C0 |= __reduce_or_sync(FULL_MASK, range_mask(j0, j1, 0u));
                    C1 |= __reduce_or_sync(FULL_MASK, range_mask(j0, j1, 32u));
                    C2 |= __reduce_or_sync(FULL_MASK, range_mask(j0, j1, 64u));
                    C3 |= __reduce_or_sync(FULL_MASK, range_mask(j0, j1, 96u));
                    anchor = ip = __shfl_sync(FULL_MASK, mend, npend - 1u);
                    if (nx == NX_LONG) {
                        const uint32_t jl = __shfl_sync(FULL_MASK, j1, npend - 1u);
                        cur = first_hit_from(jl, H[0], H[1], H[2], H[3], a1, a2, a3);
                    } else cur = nx;                                          // NX_NONE, or the next hit when the batch filled up
                    if (npend == 32u) {
                        if (nseq == 0) { first_ll = __shfl_sync(FULL_MASK, pd.ll, 0); first_hdr = 1u + len_bytes(first_ll); }
                        o = emit_batch(out, o, base, npend, pd, lane);
                        nseq += npend; npend = 0;
                    }
                }
            } else ++miss;
            // ---- 4. insert the positions that stayed outside every match
            {
                const uint32_t Cm[4] = {C0, C1, C2, C3};
                #pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (ok[k] && !((Cm[k] >> lane) & 1u)) TAB_ST(tab + h[k], (uint16_t)rel[k]);
#if B2_ENC_MIN_SL == 1
                // stride-2 steps of compressible input: the positions in between are not probed, but they are inserted (a repeat
                // at an odd distance is then still found, from its even positions), unless both neighbours lie inside a match
                if (sl == 1u) {
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t cj = (Cm[k] >> lane) & 1u;
                        const uint32_t cn = lane < 31u ? (Cm[k] >> (lane + 1u)) & 1u : (k < 3 ? Cm[k < 3 ? k + 1 : 3] & 1u : 0u);
                        const uint32_t p1 = d0 + rel[k] + 1u;
                        if (p1 < limit && !(cj & cn)) TAB_ST(tab + hash4(ld32u_ro(base + p1)), (uint16_t)(rel[k] + 1u));
                    }
                }
#endif
            }
            __syncwarp();
            T0 += span;
            if (T0 < ip) T0 = ip;
        }
        if (npend) {
            if (nseq == 0) { first_ll = __shfl_sync(FULL_MASK, pd.ll, 0); first_hdr = 1u + len_bytes(first_ll); }
            o = emit_batch(out, o, base, npend, pd, lane);
            nseq += npend; npend = 0;
        }

        uint32_t tail = b1 - anchor;
        if (last_sub) {                            // final literal-only sequence of the block
            if (nseq == 0) { first_ll = tail; first_hdr = 1u + len_bytes(tail); }
            if (lane == 0) out[o] = (uint8_t)((tail < 15u ? tail : 15u) << 4);
            o = put_len_ext(out, o + 1, tail, lane);
            if (tail <= 32) { if (lane < tail) out[o + lane] = base[anchor + lane]; }
            else group_copy(out + o, base + anchor, tail, lane, 32);
            o += tail; tail = 0; ++nseq;
        }
        if (lane == 0) {
            SubOut so; so.out_len = o; so.first_ll = first_ll; so.first_hdr = first_hdr; so.tail = tail;
            so.emitted = nseq ? 1u : 0u; so.n_seq = nseq;
            subs[s] = so;
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------
// layout, part 1: the sub-blocks of every block are stitched — one WARP per block, a lane per sub-block (32 at a time), the
// serial rules "literals left over by a sub-block that emitted nothing (or by a tail) are carried into the next one that
// emits; the first emitting unit of every match domain is a decoder entry point" written as warp scans:
//   carry   segmented sum of the tails, restarting behind every sub-block that emitted
//   dst_rel prefix sum of the final sizes (which depend on the carry through the re-encoded first token)
//   entry   segmented OR over the units of a domain -> first emitting unit, ranked by a ballot
// (the previous version walked the 256 sub-blocks of a 4 MB block with one thread: 485 us per GiB, 2.7 % of the encoder)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
layout_blocks_kernel(EncParams P, const SubOut* __restrict__ subs, SubPlace* __restrict__ places, BlockPlace* __restrict__ bplaces)
{
    const uint32_t lane = lane_id();
    const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t b = gw; b < P.n_blocks; b += nwarps) {
        const uint64_t blk_start = (uint64_t)b * P.block_size;
        const uint32_t usize = (uint32_t)((P.n - blk_start) < P.block_size ? (P.n - blk_start) : P.block_size);
        const uint32_t ns = (usize + P.sub_size - 1u) / P.sub_size;
        uint32_t carry_run = 0, off_run = 0, nc_run = 0, dom_any_run = 0;
        for (uint32_t j0 = 0; j0 < ns; j0 += 32u) {
            const uint32_t j = j0 + lane;
            const bool valid = j < ns;
            const uint32_t s = b * P.subs_per_block + j;
            SubOut so{};
            if (valid) so = subs[s];
            const bool e = valid && so.emitted != 0u;
            // carry behind sub-block j: its tail if it emitted, else the carry behind j - 1 plus its tail
            uint32_t v = valid ? so.tail : 0u, f = e ? 1u : 0u;
            #pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t pv = __shfl_up_sync(FULL_MASK, v, d), pf = __shfl_up_sync(FULL_MASK, f, d);
                if ((int)lane >= d) { if (!f) v += pv; f |= pf; }
            }
            const uint32_t c_incl = f ? v : v + carry_run;
            uint32_t c_in = __shfl_up_sync(FULL_MASK, c_incl, 1);
            if (lane == 0) c_in = carry_run;
            uint32_t size = 0, new_hdr = 0;
            if (e) { new_hdr = 1u + len_bytes(so.first_ll + c_in); size = so.out_len - so.first_hdr + new_hdr + c_in; }
            uint32_t x = size;
            #pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(FULL_MASK, x, d); if ((int)lane >= d) x += y; }
            // has a unit of j's domain emitted before j?
            const bool dstart = (j % P.dom_units) == 0u;
            uint32_t a = e ? 1u : 0u, g = dstart ? 1u : 0u;
            #pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t pa = __shfl_up_sync(FULL_MASK, a, d), pg = __shfl_up_sync(FULL_MASK, g, d);
                if ((int)lane >= d) { if (!g) a |= pa; g |= pg; }
            }
            const uint32_t any_incl = g ? a : (a | dom_any_run);
            uint32_t any_excl = __shfl_up_sync(FULL_MASK, any_incl, 1);
            if (lane == 0) any_excl = dom_any_run;
            if (dstart) any_excl = 0u;
            const bool is_first = e && !any_excl;
            const uint32_t fb = __ballot_sync(FULL_MASK, is_first);
            if (valid) {
                SubPlace sp; sp.dst_rel = off_run + x - size; sp.carry_in = c_in; sp.new_hdr = new_hdr;
                sp.size = is_first ? nc_run + (uint32_t)__popc(fb & ((1u << lane) - 1u)) + 1u : 0u;   // rank of the entry point + 1
                places[s] = sp;
            }
            carry_run = __shfl_sync(FULL_MASK, c_incl, 31); off_run += __shfl_sync(FULL_MASK, x, 31);
            dom_any_run = __shfl_sync(FULL_MASK, any_incl, 31); nc_run += (uint32_t)__popc(fb);
        }
        if (lane == 0) {
            BlockPlace bp; bp.frame_off = 0; bp.usize = usize; bp.first_chain = 0; bp.pad = 0;
            bp.stored = (off_run >= usize) ? 1u : 0u;                    // lz4frame.c:746-752
            bp.csize = bp.stored ? usize : off_run;
            bp.n_chain = bp.stored ? 1u : nc_run;
            bplaces[b] = bp;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// layout, part 2: frame offsets of the blocks (scan), header / EndMark / index head and block entries.
// One CTA of LAYOUT_THREADS threads — no more than one match-finder CTA, so that it can start in any slot a retiring
// match-finder CTA of ANOTHER chunk frees (the host-buffer pipeline keeps several chunks in flight on different streams).
// Blocks are taken LAYOUT_THREADS at a time, thread t the t-th of them (coalesced), with a CTA-wide scan per round.
// The index entries of the decoder entry points are written by gather_kernel, which visits every sub-block anyway.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(LAYOUT_THREADS)
layout_kernel(EncParams P, BlockPlace* __restrict__ bplaces, uint8_t* __restrict__ dst, uint64_t dst_cap,
              EncodeResult* __restrict__ res, uint32_t* __restrict__ ticket, uint8_t* __restrict__ idx_frag)
{
    constexpr uint32_t NT = LAYOUT_THREADS, NW = NT / 32u;
    B2_SHARED_ARRAY(uint64_t, w_bytes, NW);
    B2_SHARED_ARRAY(uint32_t, w_chain, NW);
    B2_SHARED_VAR(EncodeResult, s_res);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const uint32_t nb = P.n_blocks;
    const uint32_t bc = P.block_checksum ? 4u : 0u;
    if (tid == 0) *ticket = 0;
    const uint32_t hs = P.write_header ? (7u + (P.desc.has_content_size ? 8u : 0u) + (P.desc.has_dict_id ? 4u : 0u)) : 0u;
    uint64_t run = hs; uint32_t rc = 0;                              // totals of the rounds so far (uniform)
    for (uint32_t r0 = 0; r0 < nb; r0 += NT) {
        const uint32_t b = r0 + tid;
        BlockPlace bp{};
        if (b < nb) bp = bplaces[b];
        const uint64_t my_bytes = b < nb ? 4ull + bp.csize + bc : 0ull;
        const uint32_t my_chain = b < nb ? bp.n_chain : 0u;
        uint64_t xb = my_bytes; uint32_t xc = my_chain;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t yb = __shfl_up_sync(FULL_MASK, xb, d); const uint32_t yc = __shfl_up_sync(FULL_MASK, xc, d);
            if ((int)lane >= d) { xb += yb; xc += yc; }
        }
        if (lane == 31u) { w_bytes[wid] = xb; w_chain[wid] = xc; }
        __syncthreads();
        uint64_t pre_b = 0, tot_b = 0; uint32_t pre_c = 0, tot_c = 0;
        #pragma unroll
        for (uint32_t w = 0; w < NW; ++w) {
            const uint64_t wb = w_bytes[w]; const uint32_t wc = w_chain[w];
            if (w < wid) { pre_b += wb; pre_c += wc; }
            tot_b += wb; tot_c += wc;
        }
        __syncthreads();
        if (b < nb) {
            bp.frame_off = run + pre_b + xb - my_bytes; bp.first_chain = rc + pre_c + xc - my_chain;
            bplaces[b] = bp;
        }
        run += tot_b; rc += tot_c;
    }
    if (tid == 0) {
        EncodeResult r{};
        r.n_blocks = nb; r.n_chains = rc;
        r.blocks_len = run - hs;
        uint64_t end = run;
        if (P.write_header) {
            end += 4;                                                // EndMark
            if (P.content_checksum) { r.content_checksum_off = end; end += 4; }
            if (P.write_index) {
                r.index_off = end;
                end += 8 + sizeof(IndexHead) + (uint64_t)nb * sizeof(IndexBlock) + (uint64_t)rc * sizeof(IndexChain) + 8;
            }
        }
        r.frame_len = end;
        r.err = end > dst_cap ? (uint32_t)E_dstMaxSize_tooSmall : (uint32_t)OK;
        s_res = r; *res = r;
        if (r.err == OK && P.write_header) {
            uint8_t hdr[HEADER_MAX];
            uint32_t n = header_write(hdr, P.desc);
            for (uint32_t i = 0; i < n; ++i) dst[i] = hdr[i];
            st_le32(dst + run, 0u);                                  // EndMark (lz4frame.c:998-1000)
            if (r.index_off) {
                uint8_t* x = dst + r.index_off;
                uint32_t pb = (uint32_t)(sizeof(IndexHead) + (uint64_t)nb * sizeof(IndexBlock) + (uint64_t)rc * sizeof(IndexChain));
                st_le32(x, IDX_SKIP_MAGIC); st_le32(x + 4, pb + 8u);
                x += 8;
                st_le32(x, IDX_MAGIC); st_le32(x + 4, 1u); st_le32(x + 8, P.block_size); st_le32(x + 12, nb);
                st_le32(x + 16, rc); st_le32(x + 20, 0u);
                st_le32(x + 24, (uint32_t)P.n); st_le32(x + 28, (uint32_t)(P.n >> 32));
                st_le32(x + pb, pb); st_le32(x + pb + 4, IDX_MAGIC);
            }
        }
    }
    __syncthreads();
    if (s_res.err != OK) return;
    // index entries go into the frame's own index (one-shot call) or, when the frame is produced in chunks, into a
    // fragment buffer [nb x IndexBlock][n_chains x IndexChain] that the host rebases and concatenates
    uint8_t* ib = s_res.index_off ? dst + s_res.index_off + 8 + sizeof(IndexHead) : idx_frag;
    if (!ib) return;
    uint8_t* ic = ib + (uint64_t)nb * sizeof(IndexBlock);
    for (uint32_t b = tid; b < nb; b += NT) {
        const BlockPlace bp = bplaces[b];
        uint8_t* e = ib + (uint64_t)b * sizeof(IndexBlock);
        st_le32(e, bp.csize | (bp.stored << 31)); st_le32(e + 4, bp.usize);
        st_le32(e + 8, bp.first_chain); st_le32(e + 12, bp.n_chain);
        if (bp.stored) { uint8_t* ce = ic + (uint64_t)bp.first_chain * sizeof(IndexChain); st_le32(ce, 0u); st_le32(ce + 4, 0u); }
    }
}

// ------------------------------------------------------------------------------------------------
// gather: the sub-blocks' bytes go to their final place in the frame.  A CTA takes 8 consecutive sub-blocks at a time: one
// WARP per compressed sub-block (a few KB each: eight of them in flight per CTA hide the look-ups in front of every copy),
// the whole CTA for sub-blocks of stored blocks (16 KB of raw input each).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_kernel(const uint8_t* __restrict__ src, EncParams P, const uint8_t* __restrict__ scratch,
              const SubOut* __restrict__ subs, const SubPlace* __restrict__ places,
              const BlockPlace* __restrict__ bplaces, uint8_t* __restrict__ dst,
              const EncodeResult* __restrict__ res, uint8_t* __restrict__ idx_frag)
{
    if (res->err != OK) return;
    const uint32_t tid = threadIdx.x, nt = blockDim.x, lane = tid & 31u, wid = tid >> 5, nw = nt >> 5;
    // index entries of the decoder entry points (layout_kernel wrote the head and the block entries)
    const uint64_t index_off = res->index_off;
    uint8_t* const ib = index_off ? dst + index_off + 8 + sizeof(IndexHead) : idx_frag;
    uint8_t* const ic = ib ? ib + (uint64_t)P.n_blocks * sizeof(IndexBlock) : nullptr;
    for (uint32_t g = blockIdx.x * nw; g < P.n_subs; g += gridDim.x * nw) {
        const uint32_t s = g + wid;
        if (s < P.n_subs) {
            const uint32_t blk = s / P.subs_per_block, j = s - blk * P.subs_per_block;
            const BlockPlace bp = bplaces[blk];
            if (j == 0 && lane == 0) st_le32(dst + bp.frame_off, bp.csize | (bp.stored << 31));
            const SubOut so = subs[s];
            if (!bp.stored && so.emitted) {
                const uint64_t blk_start = (uint64_t)blk * P.block_size;
                const uint32_t b0 = j * P.sub_size;
                const SubPlace sp = places[s];
                const uint8_t* sc = scratch + (uint64_t)s * P.scratch_stride;
                uint8_t* out = dst + bp.frame_off + 4 + sp.dst_rel;
                if (ic && sp.size && lane == 0) {                            // first emitting unit of its domain: (token offset, output offset)
                    uint8_t* ce = ic + (uint64_t)(bp.first_chain + sp.size - 1u) * sizeof(IndexChain);
                    st_le32(ce, sp.dst_rel); st_le32(ce + 4, b0 - sp.carry_in);
                }
                if (sp.carry_in == 0) group_copy(out, sc, so.out_len, lane, 32);
                else {
                    // re-encode the first token with the literals carried in from the previous sub-block(s)
                    const uint32_t ll2 = so.first_ll + sp.carry_in;
                    if (lane == 0) out[0] = (uint8_t)(((ll2 < 15u ? ll2 : 15u) << 4) | (sc[0] & 15u));
                    if (ll2 >= 15u) {
                        uint32_t r = ll2 - 15u, n255 = r / 255u;
                        for (uint32_t i = lane; i < n255; i += 32u) out[1 + i] = 255;
                        if (lane == 0) out[1 + n255] = (uint8_t)(r - n255 * 255u);
                    }
                    group_copy(out + sp.new_hdr, src + blk_start + b0 - sp.carry_in, sp.carry_in, lane, 32);
                    group_copy(out + sp.new_hdr + sp.carry_in, sc + so.first_hdr, so.out_len - so.first_hdr, lane, 32);
                }
            }
        }
        for (uint32_t q = 0; q < nw && g + q < P.n_subs; ++q) {            // stored blocks: raw input, copied by the whole CTA
            const uint32_t s2 = g + q, blk = s2 / P.subs_per_block, j = s2 - blk * P.subs_per_block;
            const BlockPlace bp = bplaces[blk];
            if (!bp.stored) continue;
            const uint32_t b0 = j * P.sub_size;
            if (b0 >= bp.usize) continue;
            const uint32_t b1 = (b0 + P.sub_size < bp.usize) ? b0 + P.sub_size : bp.usize;
            group_copy(dst + bp.frame_off + 4 + b0, src + (uint64_t)blk * P.block_size + b0, b1 - b0, tid, nt);
        }
    }
}

// block checksums (lz4frame.c:756-759): XXH32 over the block's stored bytes, appended after them
__global__ void __launch_bounds__(256)
block_checksums_kernel(EncParams P, const BlockPlace* __restrict__ bplaces, uint8_t* __restrict__ dst,
                       const EncodeResult* __restrict__ res)
{
    if (res->err != OK || !P.block_checksum) return;
    if (P.block_size >= (256u << 10)) {                              // one CTA per block (xxh32_cta.cuh), see verify_block_checksums_kernel
        B2_SHARED_VAR(XxhCtaShared, S);
        for (uint32_t b = blockIdx.x; b < P.n_blocks; b += gridDim.x) {
            const BlockPlace bp = bplaces[b];
            const uint32_t h = xxh32_cta(S, dst + bp.frame_off + 4, bp.csize);
            if (threadIdx.x == 0) st_le32(dst + bp.frame_off + 4 + bp.csize, h);
        }
        return;
    }
    const uint32_t lane = lane_id();                                 // small blocks: one warp per block (xxh32_warp)
    uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (; b < P.n_blocks; b += (gridDim.x * blockDim.x) >> 5) {
        const BlockPlace bp = bplaces[b];
        const uint32_t h = xxh32_warp(dst + bp.frame_off + 4, bp.csize, lane);
        if (lane == 0) st_le32(dst + bp.frame_off + 4 + bp.csize, h);
    }
}

}  // namespace b2
