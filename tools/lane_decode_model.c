/* FUNCTIONAL CPU model of the decode scheme DESIGN.md §9 item 1 proposes for foreign (unindexed) independent blocks — the
 * companion of tools/dep_model.c (which only models its timing).  Not product code.  It decodes every block of a stock-CLI
 * frame the way the planned kernel would and compares the bytes with a plain serial decode:
 *
 *   - the block's sequences are cut into batches of 32 (the pre-pass, tools/prepass_model.c, supplies the entry points);
 *   - K "warps" take batches in ticket order; batch b is handed out only when batch b - RING is final, so a reader never needs
 *     state older than the ring (anything in front of it is final by construction);
 *   - round 1 of a batch: parse, copy all literals (they come from the compressed stream, no dependency), publish the 33
 *     output offsets of the batch;
 *   - following rounds: lane i owns match i; it copies when every byte of ITS source range is final — a literal byte is
 *     final once its batch is parsed, a match byte once that match's bit is set in its batch's done mask (lanes finish out
 *     of order, so progress is a mask, not a position); an overlapping match (offset < length) waits for the bytes in
 *     front of it and then copies forward, byte by byte, inside the lane;
 *   - rounds are concurrent: readiness is evaluated against the state at the START of the round (all warps), then the
 *     copies of the round are done.  Unwritten output is poisoned, so a missed dependency shows up as a byte mismatch.
 *
 *   gcc -O2 -o /tmp/lane_decode_model tools/lane_decode_model.c && /tmp/lane_decode_model frame.lz4 [K=32] [RING=128]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

#define BATCH 32u
static uint32_t rd32(const uint8_t* p) { return p[0] | p[1] << 8 | p[2] << 16 | (uint32_t)p[3] << 24; }

typedef struct { uint32_t lit_src, ll, o, off, ml; } Seq;      /* o = output offset of the sequence (its literals come first) */

/* which sequence writes output byte s?  (the last one whose o <= s) */
static uint32_t seq_of(const Seq* q, uint32_t nseq, uint32_t s) {
    uint32_t lo = 0, hi = nseq - 1;
    while (lo < hi) { uint32_t mid = (lo + hi + 1) / 2; if (q[mid].o <= s) lo = mid; else hi = mid - 1; }
    return lo;
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s frame.lz4 [K] [RING]\n", argv[0]); return 2; }
    const uint32_t K = argc > 2 ? (uint32_t)atoi(argv[2]) : 32u, RING = argc > 3 ? (uint32_t)atoi(argv[3]) : 128u;
    if (K == 0 || K > 1024 || RING == 0) return 2;
    FILE* f = fopen(argv[1], "rb"); if (!f) { perror("open"); return 2; }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    uint8_t* d = malloc(sz ? sz : 1); if (fread(d, 1, sz, f) != (size_t)sz) return 2; fclose(f);
    if (sz < 7 || rd32(d) != 0x184D2204u) { fprintf(stderr, "not an LZ4 frame\n"); return 2; }
    const uint8_t flg = d[4];
    if (!(flg & 0x20)) { fprintf(stderr, "linked-block frame\n"); return 3; }
    size_t at = 7 + ((flg & 8) ? 8 : 0) + ((flg & 1) ? 4 : 0);
    const int bchk = (flg >> 4) & 1;
    const uint32_t cap = 4u << 20;
    uint8_t* ref = malloc(cap); uint8_t* out = malloc(cap);
    Seq* q = malloc((size_t)(cap / 3 + 16) * sizeof(Seq));
    uint32_t* done = malloc((size_t)(cap / 3 / BATCH + 16) * 4);       /* per batch: bit i = match i copied */
    uint8_t* parsed = malloc(cap / 3 / BATCH + 16); uint8_t* final_ = malloc(cap / 3 / BATCH + 16);
    uint32_t* rounds_of = malloc((size_t)(cap / 3 / BATCH + 16) * 4);
    uint32_t* wb = malloc((size_t)K * 4); uint32_t* rdy = malloc((size_t)K * 4);
    uint64_t blocks = 0, ubytes = 0, tot_rounds = 0, tot_batches = 0, batch_rounds = 0, stall_ring = 0, long_matches = 0;
    uint32_t batch_rounds_max = 0;
    while (at + 4 <= (size_t)sz) {
        uint32_t hw = rd32(d + at); at += 4;
        if (hw == 0) break;
        const uint32_t n = hw & 0x7FFFFFFFu;
        const uint8_t* c = d + at; at += n + (bchk ? 4 : 0);
        if (hw & 0x80000000u) continue;
        /* ---- serial truth + the sequence list (what the pre-pass + per-batch parse produce) */
        uint32_t p = 0, o = 0, nseq = 0;
        for (;;) {
            uint32_t t = c[p++], ll = t >> 4;
            if (ll == 15) { uint32_t b; do { b = c[p++]; ll += b; } while (b == 255); }
            if (o + ll > cap) return 1;
            Seq* s = &q[nseq++]; s->lit_src = p; s->ll = ll; s->o = o; s->off = 0; s->ml = 0;
            memcpy(ref + o, c + p, ll); p += ll; o += ll;
            if (p >= n) break;
            s->off = c[p] | c[p + 1] << 8; p += 2;
            uint32_t ml = t & 15;
            if (ml == 15) { uint32_t b; do { b = c[p++]; ml += b; } while (b == 255); }
            ml += 4; s->ml = ml;
            if (s->off == 0 || s->off > o || o + ml > cap) { fprintf(stderr, "malformed block\n"); return 1; }
            for (uint32_t i = 0; i < ml; ++i) ref[o + i] = ref[o - s->off + i];
            if (ml > 4096) ++long_matches;
            o += ml;
        }
        const uint32_t usize = o, nb = (nseq + BATCH - 1) / BATCH;
        /* ---- the modelled execution */
        memset(out, 0xCC, usize); memset(done, 0, (size_t)nb * 4); memset(parsed, 0, nb); memset(final_, 0, nb); memset(rounds_of, 0, (size_t)nb * 4);
        for (uint32_t k = 0; k < K; ++k) wb[k] = UINT32_MAX;
        uint32_t ticket = 0, nfinal = 0; uint64_t rounds = 0;
        while (nfinal < nb) {
            ++rounds;
            /* phase A, on the state at the start of the round: idle warps take a ticket; busy warps find their ready lanes */
            for (uint32_t k = 0; k < K; ++k) {
                rdy[k] = 0;
                if (wb[k] == UINT32_MAX) {
                    if (ticket < nb) { if (ticket >= RING && !final_[ticket - RING]) { ++stall_ring; continue; } wb[k] = ticket++; }
                    continue;
                }
                const uint32_t b = wb[k];
                if (!parsed[b]) continue;
                const uint32_t s0 = b * BATCH, s1 = s0 + BATCH < nseq ? s0 + BATCH : nseq;
                for (uint32_t j = s0; j < s1; ++j) {
                    if (q[j].ml == 0 || (done[b] >> (j - s0) & 1u)) continue;
                    const uint32_t mo = q[j].o + q[j].ll, src = mo - q[j].off, span = q[j].ml < q[j].off ? q[j].ml : q[j].off;
                    int ok = 1;
                    for (uint32_t x = seq_of(q, nseq, src); ok && x <= j && q[x].o < src + span; ++x) {      /* every sequence the range touches */
                        const uint32_t xb = x / BATCH;
                        if (xb + RING <= b) continue;                                   /* in front of the ring: final by construction */
                        if (!parsed[xb]) { ok = 0; break; }                             /* literals (and offsets) not there yet */
                        const uint32_t xm = q[x].o + q[x].ll;                           /* where x's match part starts */
                        const int touches_match = q[x].ml && src + span > xm && src < xm + q[x].ml;
                        if (touches_match && x != j && !(done[xb] >> (x - xb * BATCH) & 1u)) ok = 0;
                    }
                    if (ok) rdy[k] |= 1u << (j - s0);
                }
            }
            /* phase B: the work of the round */
            for (uint32_t k = 0; k < K; ++k) {
                const uint32_t b = wb[k];
                if (b == UINT32_MAX) continue;
                ++rounds_of[b];
                const uint32_t s0 = b * BATCH, s1 = s0 + BATCH < nseq ? s0 + BATCH : nseq;
                if (!parsed[b]) {
                    for (uint32_t j = s0; j < s1; ++j) memcpy(out + q[j].o, c + q[j].lit_src, q[j].ll);
                    parsed[b] = 1;
                    for (uint32_t j = s0; j < s1; ++j) if (q[j].ml == 0) done[b] |= 1u << (j - s0);           /* the block's last sequence has no match */
                } else {
                    for (uint32_t j = s0; j < s1; ++j) if (rdy[k] >> (j - s0) & 1u) {
                        const uint32_t mo = q[j].o + q[j].ll;
                        for (uint32_t i = 0; i < q[j].ml; ++i) out[mo + i] = out[mo - q[j].off + i];
                        done[b] |= 1u << (j - s0);
                    }
                }
                const uint32_t full = (s1 - s0 == 32) ? 0xFFFFFFFFu : ((1u << (s1 - s0)) - 1u);
                if (parsed[b] && done[b] == full) {
                    final_[b] = 1; ++nfinal; wb[k] = UINT32_MAX;
                    batch_rounds += rounds_of[b]; if (rounds_of[b] > batch_rounds_max) batch_rounds_max = rounds_of[b];
                }
            }
            if (rounds > (uint64_t)nb * 40 + 1000) { fprintf(stderr, "no progress (deadlock) in block %llu\n", (unsigned long long)blocks); return 1; }
        }
        if (memcmp(out, ref, usize) != 0) {
            uint32_t x = 0; while (out[x] == ref[x]) ++x;
            fprintf(stderr, "block %llu: byte %u differs from the serial decode\n", (unsigned long long)blocks, x); return 1;
        }
        ++blocks; ubytes += usize; tot_rounds += rounds; tot_batches += nb;
    }
    if (!blocks) { printf("no compressed blocks\n"); return 0; }
    printf("blocks %llu, %llu B out: lane-dataflow decode equals the serial decode: OK\n", (unsigned long long)blocks, (unsigned long long)ubytes);
    printf("K = %u warps, ring %u batches: %.0f batches per block done in %.0f rounds per block (%.1f batches per round; the "
           "longest copy chain, tools/dep_model.c, is the floor); rounds a batch stays resident: avg %.1f, max %u; ticket stalls on the ring %llu; "
           "matches longer than 4 KB (want a warp-wide copy) %llu\n", K, RING, (double)tot_batches / blocks, (double)tot_rounds / blocks,
           (double)tot_batches / tot_rounds, (double)batch_rounds / tot_batches, batch_rounds_max, (unsigned long long)stall_ring,
           (unsigned long long)long_matches);
    return 0;
}
