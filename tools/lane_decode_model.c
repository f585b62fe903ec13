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
 *   - a LINKED-block frame (what the reference conduit writes, Conduit.hsc:248-263) is one unit: sequences of all its blocks in
 *     one list, matches reach into the previous blocks' output, batches may straddle block borders (a block's last sequence
 *     has no match; a stored block is one all-literal sequence) — DESIGN.md §9 item 2.
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

typedef struct { uint64_t units, ubytes, rounds, batches, batch_rounds, stall_ring, long_matches, checks, touched, matches; uint32_t batch_rounds_max; } Stats;

/* decode one unit (an independent block, or a whole linked frame) the modelled way into `out` and compare with `ref` */
static int run_model(const uint8_t* d, const Seq* q, uint32_t nseq, uint32_t usize, const uint8_t* ref, uint8_t* out, uint32_t K, uint32_t RING, Stats* st) {
    const uint32_t nb = (nseq + BATCH - 1) / BATCH;
    uint32_t* done = calloc((size_t)nb + 1, 4);                       /* per batch: bit i = match i copied */
    uint8_t* parsed = calloc((size_t)nb + 1, 1); uint8_t* final_ = calloc((size_t)nb + 1, 1);
    uint32_t* rounds_of = calloc((size_t)nb + 1, 4);
    uint32_t* wb = malloc((size_t)K * 4); uint32_t* rdy = malloc((size_t)K * 4);
    memset(out, 0xCC, usize);
    for (uint32_t k = 0; k < K; ++k) wb[k] = UINT32_MAX;
    uint32_t ticket = 0, nfinal = 0; uint64_t rounds = 0;
    while (nfinal < nb) {
        ++rounds;
        /* phase A, on the state at the start of the round: idle warps take a ticket; busy warps find their ready lanes */
        for (uint32_t k = 0; k < K; ++k) {
            rdy[k] = 0;
            if (wb[k] == UINT32_MAX) {
                if (ticket < nb) { if (ticket >= RING && !final_[ticket - RING]) { ++st->stall_ring; continue; } wb[k] = ticket++; }
                continue;
            }
            const uint32_t b = wb[k];
            if (!parsed[b]) continue;
            const uint32_t s0 = b * BATCH, s1 = s0 + BATCH < nseq ? s0 + BATCH : nseq;
            for (uint32_t j = s0; j < s1; ++j) {
                if (q[j].ml == 0 || (done[b] >> (j - s0) & 1u)) continue;
                const uint32_t mo = q[j].o + q[j].ll, src = mo - q[j].off, span = q[j].ml < q[j].off ? q[j].ml : q[j].off;
                int ok = 1; ++st->checks;
                for (uint32_t x = seq_of(q, nseq, src); ok && x <= j && q[x].o < src + span; ++x) {      /* every sequence the range touches */
                    const uint32_t xb = x / BATCH; ++st->touched;
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
                for (uint32_t j = s0; j < s1; ++j) memcpy(out + q[j].o, d + q[j].lit_src, q[j].ll);
                parsed[b] = 1;
                for (uint32_t j = s0; j < s1; ++j) if (q[j].ml == 0) done[b] |= 1u << (j - s0);           /* a block's last sequence has no match */
            } else {
                for (uint32_t j = s0; j < s1; ++j) if (rdy[k] >> (j - s0) & 1u) {
                    const uint32_t mo = q[j].o + q[j].ll;
                    for (uint32_t i = 0; i < q[j].ml; ++i) out[mo + i] = out[mo - q[j].off + i];
                    done[b] |= 1u << (j - s0); ++st->matches;
                }
            }
            const uint32_t full = (s1 - s0 == 32) ? 0xFFFFFFFFu : ((1u << (s1 - s0)) - 1u);
            if (parsed[b] && done[b] == full) {
                final_[b] = 1; ++nfinal; wb[k] = UINT32_MAX;
                st->batch_rounds += rounds_of[b]; if (rounds_of[b] > st->batch_rounds_max) st->batch_rounds_max = rounds_of[b];
            }
        }
        if (rounds > (uint64_t)nb * 40 + 1000) { fprintf(stderr, "no progress (deadlock) in unit %llu\n", (unsigned long long)st->units); return 1; }
    }
    if (memcmp(out, ref, usize) != 0) {
        uint32_t x = 0; while (out[x] == ref[x]) ++x;
        fprintf(stderr, "unit %llu: byte %u differs from the serial decode\n", (unsigned long long)st->units, x); return 1;
    }
    ++st->units; st->ubytes += usize; st->rounds += rounds; st->batches += nb;
    free(done); free(parsed); free(final_); free(rounds_of); free(wb); free(rdy);
    return 0;
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
    const int linked = !(flg & 0x20), bchk = (flg >> 4) & 1;
    const size_t at0 = 7 + ((flg & 8) ? 8 : 0) + ((flg & 1) ? 4 : 0);
    /* the serial truth needs the output size first: one pass over the sequences without copying */
    uint64_t cap = 0, nseq_cap = 0, cur = 0, cur_seq = 0;
    for (size_t at = at0; at + 4 <= (size_t)sz;) {
        uint32_t hw = rd32(d + at); at += 4;
        if (hw == 0) break;
        const uint32_t n = hw & 0x7FFFFFFFu;
        if (at + n > (size_t)sz) { fprintf(stderr, "truncated frame\n"); return 1; }
        if (hw & 0x80000000u) { if (linked) { cur += n; ++cur_seq; } at += n + (bchk ? 4 : 0); continue; }
        if (!linked) { cur = 0; cur_seq = 0; }
        const uint8_t* c = d + at;
        for (uint32_t p = 0;;) {
            uint32_t t = c[p++], ll = t >> 4;
            if (ll == 15) { uint32_t b; do { b = c[p++]; ll += b; } while (b == 255 && p < n); }
            p += ll; cur += ll; ++cur_seq;
            if (p >= n) break;
            p += 2;
            uint32_t ml = t & 15;
            if (ml == 15) { uint32_t b; do { b = c[p++]; ml += b; } while (b == 255 && p < n); }
            cur += ml + 4;
        }
        at += n + (bchk ? 4 : 0);
        if (cur > cap) cap = cur;
        if (cur_seq > nseq_cap) nseq_cap = cur_seq;
    }
    if (cur > cap) cap = cur;
    if (cur_seq > nseq_cap) nseq_cap = cur_seq;
    if (cap >= UINT32_MAX) { fprintf(stderr, "unit larger than 4 GiB\n"); return 2; }
    uint8_t* ref = malloc(cap + 1); uint8_t* out = malloc(cap + 1);
    Seq* q = malloc((size_t)(nseq_cap + 1) * sizeof(Seq));
    Stats st; memset(&st, 0, sizeof st);
    uint32_t o = 0, nseq = 0;
    for (size_t at = at0; at + 4 <= (size_t)sz;) {
        uint32_t hw = rd32(d + at); at += 4;
        if (hw == 0) break;
        const uint32_t n = hw & 0x7FFFFFFFu;
        const uint8_t* c = d + at;
        if (hw & 0x80000000u) {                                    /* stored block: one all-literal sequence of a linked frame's list */
            if (linked) { Seq* s = &q[nseq++]; s->lit_src = (uint32_t)at; s->ll = n; s->o = o; s->off = 0; s->ml = 0; memcpy(ref + o, c, n); o += n; }
            at += n + (bchk ? 4 : 0); continue;
        }
        if (!linked) { o = 0; nseq = 0; }
        for (uint32_t p = 0;;) {                                   /* serial truth + the sequence list (what pre-pass + per-batch parse produce) */
            uint32_t t = c[p++], ll = t >> 4;
            if (ll == 15) { uint32_t b; do { b = c[p++]; ll += b; } while (b == 255); }
            Seq* s = &q[nseq++]; s->lit_src = (uint32_t)(at + p); s->ll = ll; s->o = o; s->off = 0; s->ml = 0;
            memcpy(ref + o, c + p, ll); p += ll; o += ll;
            if (p >= n) break;
            s->off = c[p] | c[p + 1] << 8; p += 2;
            uint32_t ml = t & 15;
            if (ml == 15) { uint32_t b; do { b = c[p++]; ml += b; } while (b == 255); }
            ml += 4; s->ml = ml;
            if (s->off == 0 || s->off > o) { fprintf(stderr, "malformed block\n"); return 1; }
            for (uint32_t i = 0; i < ml; ++i) ref[o + i] = ref[o - s->off + i];
            if (ml > 4096) ++st.long_matches;
            o += ml;
        }
        at += n + (bchk ? 4 : 0);
        if (!linked && run_model(d, q, nseq, o, ref, out, K, RING, &st)) return 1;
    }
    if (linked && nseq && run_model(d, q, nseq, o, ref, out, K, RING, &st)) return 1;
    if (!st.units) { printf("no compressed blocks\n"); return 0; }
    printf("%s %llu, %llu B out: lane-dataflow decode equals the serial decode: OK\n", linked ? "linked frames" : "blocks",
           (unsigned long long)st.units, (unsigned long long)st.ubytes);
    printf("K = %u warps, ring %u batches: %.0f batches per unit done in %.0f rounds per unit (%.1f batches per round; the "
           "longest copy chain, tools/dep_model.c, is the floor); rounds a batch stays resident: avg %.1f, max %u; ticket stalls on the ring %llu; "
           "matches longer than 4 KB (want a warp-wide copy) %llu\n", K, RING, (double)st.batches / st.units, (double)st.rounds / st.units,
           (double)st.batches / st.rounds, (double)st.batch_rounds / st.batches, st.batch_rounds_max, (unsigned long long)st.stall_ring,
           (unsigned long long)st.long_matches);
    printf("readiness checks per match %.2f (1 = ready at first look), sequences examined per check %.2f\n",
           st.matches ? (double)st.checks / st.matches : 0.0, st.checks ? (double)st.touched / st.checks : 0.0);
    return 0;
}
