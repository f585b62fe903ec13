/* CPU model of the pre-pass kernels K1a / K1b / K1c of DESIGN.md §9 item 1 (foreign, unindexed blocks): from the compressed
 * bytes of a block to the TICKET LIST the lane-dataflow decoder (tools/lane_decode_model.c) hands out — the compressed and
 * output offset of every 32nd sequence.  Not product code; it pins the details before the kernels are written:
 *
 *   K1a  per group of G compressed bytes: for every byte position p, IF a sequence started at p — next[p], out[p]; length
 *        runs are followed for at most RUN bytes (a kernel loop bound), longer ones mark the position SLOW; a position whose
 *        sequence runs off the block is INVALID.  Then (exit, out, count, slow-at) per position = follow next[] to the first
 *        position outside the group or to a SLOW token (the kernel squares the link table; the result is the same).
 *   K1b  one walker per block: entry of group g = where the true chain enters it, with the output offset and the sequence
 *        number there; one table look-up per group, a SLOW token is stepped with the unbounded serial rule and the walk
 *        re-enters the same group when the token's successor still lies in it.
 *   K1c  per group, from its entry: emit (position, output offset) of every sequence whose number is a multiple of 32.
 *
 * Checked against a serial parse: the tickets are exactly the boundaries of sequences 0, 32, 64, ...; the last ticket range
 * ends at the block's end.  Prints what sizes the kernels: groups entered, SLOW tokens, INVALID positions.
 *   gcc -O2 -o /tmp/tickets_model tools/tickets_model.c && /tmp/tickets_model frame.lz4 [G=1024] [RUN=64]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

#define INVALID 0xFFFFFFFFu
#define SLOW    0xFFFFFFFEu
static uint32_t rd32(const uint8_t* p) { return p[0] | p[1] << 8 | p[2] << 16 | (uint32_t)p[3] << 24; }

/* one sequence starting at p; run = longest length run followed (0 = unbounded).  Returns the next token position (n = the block
 * ends with this sequence), INVALID when it runs off the block, SLOW when a run is longer than `run`. */
static uint32_t step(const uint8_t* c, uint32_t n, uint32_t p, uint32_t run, uint32_t* produced) {
    uint32_t t = c[p], ll = t >> 4, q = p + 1, k;
    if (ll == 15) { uint32_t b; k = 0; do { if (q >= n) return INVALID; if (run && k++ == run) return SLOW; b = c[q++]; ll += b; } while (b == 255); }
    if (ll > n - q) return INVALID;
    q += ll;
    if (q == n) { *produced = ll; return n; }
    if (q + 2 > n) return INVALID;
    q += 2;
    uint32_t ml = t & 15;
    if (ml == 15) { uint32_t b; k = 0; do { if (q >= n) return INVALID; if (run && k++ == run) return SLOW; b = c[q++]; ml += b; } while (b == 255); }
    *produced = ll + ml + 4;
    return q;
}

typedef struct { uint32_t exit, out, count, slow_at; } Row;      /* slow_at = position of a SLOW token the chain stops at, or INVALID */
typedef struct { uint32_t pos, out; } Ticket;

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s frame.lz4 [G] [RUN]\n", argv[0]); return 2; }
    const uint32_t G = argc > 2 ? (uint32_t)atoi(argv[2]) : 1024u, RUN = argc > 3 ? (uint32_t)atoi(argv[3]) : 64u;
    if (G < 8 || RUN == 0) return 2;
    FILE* f = fopen(argv[1], "rb"); if (!f) { perror("open"); return 2; }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    uint8_t* d = malloc(sz ? sz : 1); if (fread(d, 1, sz, f) != (size_t)sz) return 2; fclose(f);
    if (sz < 7 || rd32(d) != 0x184D2204u) { fprintf(stderr, "not an LZ4 frame\n"); return 2; }
    const uint8_t flg = d[4];
    size_t at = 7 + ((flg & 8) ? 8 : 0) + ((flg & 1) ? 4 : 0);
    const int bchk = (flg >> 4) & 1;
    uint64_t blocks = 0, seqs = 0, groups = 0, entered = 0, slow_hit = 0, slow_pos = 0, invalid_pos = 0, positions = 0, tickets = 0, reenter = 0;
    while (at + 4 <= (size_t)sz) {
        uint32_t hw = rd32(d + at); at += 4;
        if (hw == 0) break;
        const uint32_t n = hw & 0x7FFFFFFFu;
        const uint8_t* c = d + at; at += n + (bchk ? 4 : 0);
        if ((hw & 0x80000000u) || n == 0) continue;
        ++blocks;
        /* ---- serial truth: boundaries of sequences 0, 32, 64, ... */
        Ticket* truth = malloc((size_t)(n / 3 / 32 + 2) * sizeof(Ticket)); uint32_t ntruth = 0, nseq = 0, total = 0;
        for (uint32_t p = 0; p < n;) {
            uint32_t o = 0; const uint32_t nx = step(c, n, p, 0, &o);
            if (nx == INVALID) { fprintf(stderr, "malformed block\n"); return 1; }
            if ((nseq & 31u) == 0) { truth[ntruth].pos = p; truth[ntruth].out = total; ++ntruth; }
            ++nseq; total += o; p = nx;
        }
        seqs += nseq;
        /* ---- K1a: tables of every group (the model keeps them all; the kernel writes them to scratch) */
        Row* row = malloc((size_t)n * sizeof(Row));
        const uint32_t ng = (n + G - 1) / G;
        for (uint32_t g = 0; g < ng; ++g) {
            const uint32_t ws = g * G, we = ws + G < n ? ws + G : n;
            for (uint32_t p = we; p-- > ws;) {                       /* back to front = the fixed point of squaring next[] inside the group */
                uint32_t o = 0; const uint32_t nx = step(c, n, p, RUN, &o);
                ++positions;
                if (nx == INVALID) { row[p] = (Row){ INVALID, 0, 0, INVALID }; ++invalid_pos; }
                else if (nx == SLOW) { row[p] = (Row){ p, 0, 0, p }; ++slow_pos; }
                else if (nx >= we) row[p] = (Row){ nx, o, 1, INVALID };
                else if (row[nx].exit == INVALID) row[p] = (Row){ INVALID, 0, 0, INVALID };
                else row[p] = (Row){ row[nx].exit, o + row[nx].out, 1 + row[nx].count, row[nx].slow_at };
            }
        }
        groups += ng;
        /* ---- K1b: stitch; entry per group */
        uint32_t* ent_pos = malloc((size_t)ng * 4); uint32_t* ent_out = malloc((size_t)ng * 4); uint32_t* ent_seq = malloc((size_t)ng * 4);
        for (uint32_t g = 0; g < ng; ++g) ent_pos[g] = INVALID;
        uint32_t e = 0, produced = 0, sq = 0;
        while (e < n) {
            const uint32_t g = e / G, we = (g + 1) * G < n ? (g + 1) * G : n;
            if (ent_pos[g] == INVALID) { ent_pos[g] = e; ent_out[g] = produced; ent_seq[g] = sq; ++entered; } else ++reenter;
            const Row r = row[e];
            if (r.exit == INVALID) { fprintf(stderr, "true chain reached an INVALID row\n"); return 1; }
            produced += r.out; sq += r.count; e = r.exit;
            if (r.slow_at != INVALID) {                              /* the unbounded serial rule for this one token */
                uint32_t o = 0; const uint32_t nx = step(c, n, e, 0, &o);
                if (nx == INVALID) { fprintf(stderr, "SLOW token is malformed\n"); return 1; }
                produced += o; ++sq; e = nx; ++slow_hit;
            }
            (void)we;
        }
        if (e != n || produced != total || sq != nseq) { fprintf(stderr, "stitch: end %u/%u, bytes %u/%u, sequences %u/%u\n", e, n, produced, total, sq, nseq); return 1; }
        /* ---- K1c: tickets from the entries (a group the chain jumped over has none) */
        uint32_t nt = 0;
        for (uint32_t g = 0; g < ng; ++g) {
            if (ent_pos[g] == INVALID) continue;
            const uint32_t we = (g + 1) * G < n ? (g + 1) * G : n;
            uint32_t p = ent_pos[g], o = ent_out[g], s = ent_seq[g];
            while (p < we) {                                         /* the kernel ranks this chain by pointer jumping; SLOW tokens use the serial rule */
                if ((s & 31u) == 0) {
                    if (nt >= ntruth || truth[nt].pos != p || truth[nt].out != o) { fprintf(stderr, "ticket %u: (%u, %u) is not the boundary of sequence %u\n", nt, p, o, nt * 32); return 1; }
                    ++nt;
                }
                uint32_t po = 0; const uint32_t nx = step(c, n, p, 0, &po);
                o += po; ++s; p = nx;
            }
        }
        if (nt != ntruth) { fprintf(stderr, "tickets %u, expected %u\n", nt, ntruth); return 1; }
        tickets += nt;
        free(truth); free(row); free(ent_pos); free(ent_out); free(ent_seq);
    }
    if (!blocks) { printf("no compressed blocks\n"); return 0; }
    printf("blocks %llu, sequences %llu: tickets %llu = every 32nd sequence boundary: OK\n", (unsigned long long)blocks, (unsigned long long)seqs, (unsigned long long)tickets);
    printf("G = %u, RUN = %u: groups %llu, entered by the chain %llu (one dependent look-up each), re-entered after a SLOW token %llu; SLOW tokens on the "
           "chain %llu; of all %llu positions %.2f %% parse as SLOW and %.1f %% as INVALID\n", G, RUN, (unsigned long long)groups, (unsigned long long)entered,
           (unsigned long long)reenter, (unsigned long long)slow_hit, (unsigned long long)positions, 100.0 * slow_pos / positions, 100.0 * invalid_pos / positions);
    return 0;
}
