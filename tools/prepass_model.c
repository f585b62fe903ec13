/* CPU model of the token-link pre-pass planned for foreign (unindexed) blocks — DESIGN.md §9 item 1.  Not product code:
 * it checks the idea on real frames before any kernel is written.
 *
 *   next[p], out[p]   where the next token starts and how many bytes come out IF a sequence starts at byte p of the block
 *                     (what decode.cu's window code computes per 128-byte window as table T1)
 *   per window of W bytes: exit(e), sum(e) for every entry e — follow `next` until it leaves the window
 *   stitch            e = 0; for each window the chain enters: (e, produced) += table[e]   -> one dependent look-up per window
 *
 * The model verifies that the stitched entry points are exactly the sequence boundaries of a serial parse, and prints
 * what sizes the kernel: hops per window (squaring depth), windows skipped by long literal runs, entry points per block.
 *   gcc -O2 -o /tmp/prepass_model tools/prepass_model.c && /tmp/prepass_model frame.lz4 [W=128]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

#define INVALID 0xFFFFFFFFu
static uint32_t rd32(const uint8_t* p) { return p[0] | p[1] << 8 | p[2] << 16 | (uint32_t)p[3] << 24; }

/* one sequence starting at p: returns next token position (n = block ends here), INVALID when it runs off the block */
static uint32_t step(const uint8_t* c, uint32_t n, uint32_t p, uint32_t* produced) {
    uint32_t t = c[p], ll = t >> 4, q = p + 1;
    if (ll == 15) { uint32_t b; do { if (q >= n) return INVALID; b = c[q++]; ll += b; } while (b == 255); }
    if (q + ll > n) return INVALID;
    q += ll;
    if (q == n) { *produced = ll; return n; }                       /* last sequence: literals only */
    if (q + 2 > n) return INVALID;
    q += 2;
    uint32_t ml = t & 15;
    if (ml == 15) { uint32_t b; do { if (q >= n) return INVALID; b = c[q++]; ml += b; } while (b == 255); }
    *produced = ll + ml + 4;
    return q;
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s frame.lz4 [W]\n", argv[0]); return 2; }
    const uint32_t W = argc > 2 ? (uint32_t)atoi(argv[2]) : 128;
    FILE* f = fopen(argv[1], "rb"); if (!f) { perror("open"); return 2; }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    uint8_t* d = malloc(sz); if (fread(d, 1, sz, f) != (size_t)sz) return 2; fclose(f);
    if (rd32(d) != 0x184D2204u) { fprintf(stderr, "not an LZ4 frame\n"); return 2; }
    const uint8_t flg = d[4];
    size_t at = 7 + ((flg & 8) ? 8 : 0) + ((flg & 1) ? 4 : 0);
    const int bchk = (flg >> 4) & 1;
    uint64_t blocks = 0, seqs = 0, windows = 0, entered = 0, hops_max = 0, hops_sum = 0, skipped = 0, bad = 0, cbytes = 0, ubytes = 0;
    uint64_t false_valid = 0, positions = 0;
    while (at + 4 <= (size_t)sz) {
        uint32_t hw = rd32(d + at); at += 4;
        if (hw == 0) break;
        uint32_t n = hw & 0x7FFFFFFFu;
        const uint8_t* c = d + at; at += n + (bchk ? 4 : 0);
        if (hw & 0x80000000u) continue;                              /* stored block: nothing to parse */
        ++blocks; cbytes += n;
        uint32_t* nx = malloc((size_t)n * 4); uint32_t* ou = malloc((size_t)n * 4);
        for (uint32_t p = 0; p < n; ++p) { uint32_t o = 0; nx[p] = step(c, n, p, &o); ou[p] = o; positions++; if (nx[p] != INVALID) ++false_valid; }
        /* serial truth */
        uint8_t* boundary = calloc(n + 1, 1);
        uint64_t total = 0;
        for (uint32_t p = 0; p < n;) { boundary[p] = 1; ++seqs; if (nx[p] == INVALID) { ++bad; break; } total += ou[p]; p = nx[p]; }
        ubytes += total;
        /* per-window tables + stitch */
        const uint32_t nw = (n + W - 1) / W;
        uint32_t* ex = malloc((size_t)W * 4); uint64_t* sm = malloc((size_t)W * 8);
        uint32_t e = 0; uint64_t produced = 0;
        for (uint32_t w = 0; w < nw; ++w) {
            ++windows;
            const uint32_t ws = w * W, we = ws + W < n ? ws + W : n;
            if (e >= we) { ++skipped; continue; }                    /* a long literal run jumped over this window */
            /* exit / sum for every entry of the window, back to front (the kernel does this by squaring the link table) */
            for (uint32_t p = we; p-- > ws;) {
                uint32_t k = p - ws;
                if (nx[p] == INVALID) { ex[k] = INVALID; sm[k] = 0; }
                else if (nx[p] >= we) { ex[k] = nx[p]; sm[k] = ou[p]; }
                else { ex[k] = ex[nx[p] - ws]; sm[k] = ou[p] + sm[nx[p] - ws]; }
            }
            /* hops the true chain makes inside this window = squaring depth needed */
            uint64_t hops = 0; for (uint32_t p = e; p < we && p != INVALID; p = nx[p]) { ++hops; if (nx[p] == INVALID) break; }
            hops_sum += hops; if (hops > hops_max) hops_max = hops;
            ++entered;
            if (!boundary[e]) { fprintf(stderr, "entry %u of window %u is not a sequence boundary\n", e, w); return 1; }
            if (ex[e - ws] == INVALID) { ++bad; break; }
            produced += sm[e - ws]; e = ex[e - ws];
        }
        if (e != n || produced != total) { fprintf(stderr, "block %llu: stitched end %u/%u, bytes %llu/%llu\n", (unsigned long long)blocks, e, n, (unsigned long long)produced, (unsigned long long)total); return 1; }
        free(nx); free(ou); free(boundary); free(ex); free(sm);
    }
    printf("blocks %llu, compressed %llu B -> %llu B, sequences %llu (%.1f B out each), malformed %llu\n", (unsigned long long)blocks,
           (unsigned long long)cbytes, (unsigned long long)ubytes, (unsigned long long)seqs, seqs ? (double)ubytes / seqs : 0.0, (unsigned long long)bad);
    printf("W = %u: windows %llu, entered by the chain %llu, jumped over by long literal runs %llu; sequences per entered window avg %.1f, max %llu "
           "(squaring depth %d)\n", W, (unsigned long long)windows, (unsigned long long)entered, (unsigned long long)skipped,
           entered ? (double)hops_sum / entered : 0.0, (unsigned long long)hops_max, hops_max > 32 ? 6 : hops_max > 16 ? 5 : hops_max > 8 ? 4 : 3);
    printf("positions that parse as a sequence although most are not boundaries: %.1f %% (speculative work is done for every byte)\n",
           positions ? 100.0 * false_valid / positions : 0.0);
    printf("stitch: one dependent look-up per entered window = %.0f per 4 MB of output; every stitched entry is a true sequence boundary: OK\n",
           ubytes ? (double)entered * (4u << 20) / ubytes : 0.0);
    return 0;
}
