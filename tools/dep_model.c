/* CPU model of the MATCH dependencies inside foreign (unindexed) independent blocks — DESIGN.md §9 items 1-2.  Not product
 * code.  The token-link pre-pass (tools/prepass_model.c) finds entry points, but decode.cu gives an independent piece
 * `reach = -piece_start` (decode.cu: Cursor::reach): a match may not read output in front of its own piece.  That holds for
 * frames our encoder wrote (matches never leave their sub-block) and NOT for blocks another encoder wrote.  This model says
 * how much of a foreign block depends on bytes in front of a piece of P output bytes, and how deep the copy chains are:
 *
 *   direct     matches whose source starts in front of their piece            (would fail the reach check today)
 *   tainted    output bytes that depend, through any number of copies, on bytes in front of their piece
 *              (what a "decode pieces in parallel, defer what reaches back" scheme has to redo serially)
 *   depth      longest chain of match copies feeding one another (literals = 0): sequences / depth is the parallelism a
 *              dataflow execution of the match copies could reach at best; a piece-ordered execution needs `waves` rounds
 *              (a piece can run once every piece it reads from is final)
 *
 *   pipeline   timing model of the scheme DESIGN.md §9 proposes: one CTA per block, K warps take batches of 32 sequences in
 *              order (entry points from the pre-pass); a warp parses its batch and copies the literals at once (Tp), then does
 *              its match copies in order (Tc each); a match whose source lies in front of the batch waits until the
 *              "final up to" watermark (end of the longest prefix of finished batches, seen Th after the fact) has passed it.
 *              Printed as speed-up over one warp doing everything (batches x (Tp + matches x Tc)).
 *
 *   gcc -O2 -o /tmp/dep_model tools/dep_model.c && /tmp/dep_model frame.lz4 [P=131072]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

static uint32_t rd32(const uint8_t* p) { return p[0] | p[1] << 8 | p[2] << 16 | (uint32_t)p[3] << 24; }

#define BATCH 32u
typedef struct { uint32_t src, src_end, own, o, o_end; } Mt;            /* a match: end of its source range; own = 1 when the source starts inside its batch */

/* K warps, batches in ticket order; returns the time the last batch is final.  mt[mfirst[b] .. mfirst[b+1]) are batch b's matches,
 * bend[b] = output position at the end of batch b. */
/* when is output byte s final?  fine = 1: as soon as the sequence that writes it is done (every batch publishes its own progress);
 * fine = 0: when the watermark (longest prefix of finished batches) has passed it */
static double ready_at(uint32_t s, uint32_t b, uint32_t nm_before, const Mt* mt, const uint32_t* mbatch, const uint32_t* bend, const double* pm,
                       const double* parsed, const double* ft, int fine) {
    if (!fine) {
        uint32_t lo = 0, hi = b - 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (bend[mid] > s) hi = mid; else lo = mid + 1; }
        return pm[lo];
    }
    uint32_t lo = 0, hi = nm_before;                       /* first match whose output ends behind s (nm_before = the asking match) */
    while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (mt[mid].o_end > s) hi = mid; else lo = mid + 1; }
    if (lo < nm_before && s >= mt[lo].o) return ft[lo];     /* inside match lo */
    return parsed[lo < nm_before ? mbatch[lo] : b];         /* a literal of the sequence that match lo closes */
}

static double pipeline(uint32_t nb, const uint32_t* mfirst, const Mt* mt, const uint32_t* bend, uint32_t K, double Tp, double Tc, double Th,
                       double* done, double* pm, double* wfree, uint32_t* mbatch, double* parsed, double* ft, int fine) {
    const double Tl = 4.0 * Tc;                            /* a lane copying its match byte by byte: a few times the warp-wide copy */
    for (uint32_t k = 0; k < K; ++k) wfree[k] = 0.0;
    for (uint32_t b = 0; b < nb; ++b) {
        uint32_t kmin = 0; for (uint32_t k = 1; k < K; ++k) if (wfree[k] < wfree[kmin]) kmin = k;
        double t = wfree[kmin] + Tp;
        parsed[b] = t;
        for (uint32_t i = mfirst[b]; i < mfirst[b + 1]; ++i) {
            mbatch[i] = b;
            if (!mt[i].own && b > 0) {
                double r0 = ready_at(mt[i].src, b, i, mt, mbatch, bend, pm, parsed, ft, fine);
                double r1 = ready_at(mt[i].src_end - 1, b, i, mt, mbatch, bend, pm, parsed, ft, fine);
                if (mt[i].src_end - 1 >= (b ? bend[b - 1] : 0)) r1 = 0.0;   /* the end lies in the own batch: in order, no wait */
                const double ready = (r0 > r1 ? r0 : r1) + Th;
                if (ready > t) t = ready;
            }
            t += Tc; ft[i] = t;
        }
        if (fine == 2) {                                   /* one match per lane: each waits for its own sources only (rounds inside the warp) */
            t = parsed[b];
            for (uint32_t i = mfirst[b]; i < mfirst[b + 1]; ++i) {
                double r = parsed[b];
                const uint32_t ends[2] = { mt[i].src, mt[i].src_end - 1 };
                for (int e = 0; e < 2; ++e) {
                    const uint32_t own_start = b ? bend[b - 1] : 0;
                    double q = ready_at(ends[e], b, i, mt, mbatch, bend, pm, parsed, ft, 1);
                    if (ends[e] < own_start) q += Th;
                    if (q > r) r = q;
                }
                ft[i] = r + Tl; if (ft[i] > t) t = ft[i];
            }
        }
        done[b] = t; pm[b] = b && pm[b - 1] > t ? pm[b - 1] : t; wfree[kmin] = t;
    }
    return nb ? pm[nb - 1] : 0.0;
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s frame.lz4 [P]\n", argv[0]); return 2; }
    const uint32_t P = argc > 2 ? (uint32_t)atoi(argv[2]) : 131072u;
    FILE* f = fopen(argv[1], "rb"); if (!f) { perror("open"); return 2; }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    uint8_t* d = malloc(sz ? sz : 1); if (fread(d, 1, sz, f) != (size_t)sz) return 2; fclose(f);
    if (sz < 7 || rd32(d) != 0x184D2204u) { fprintf(stderr, "not an LZ4 frame\n"); return 2; }
    const uint8_t flg = d[4];
    if (!(flg & 0x20)) { fprintf(stderr, "linked-block frame: every block depends on the one before it (one chain per frame)\n"); return 3; }
    size_t at = 7 + ((flg & 8) ? 8 : 0) + ((flg & 1) ? 4 : 0);
    const int bchk = (flg >> 4) & 1;
    const uint32_t cap = 4u << 20;
    uint8_t* taint = malloc(cap); uint32_t* depth = malloc((size_t)cap * 4);
    uint32_t* piece_wave = malloc((size_t)(cap / (P ? P : 1) + 2) * 4);
    uint64_t blocks = 0, seqs = 0, matches = 0, direct = 0, ubytes = 0, mbytes = 0, tbytes = 0, pieces = 0, waves_sum = 0;
    uint64_t depth_sum = 0, depth_max = 0, far = 0;
    Mt* mt = malloc((size_t)(cap / 4 + 16) * sizeof(Mt)); uint32_t* mfirst = malloc((size_t)(cap / 4 / BATCH + 16) * 4);
    uint32_t* bend = malloc((size_t)(cap / 4 / BATCH + 16) * 4);
    double* done = malloc((size_t)(cap / 4 / BATCH + 16) * 8); double* pm = malloc((size_t)(cap / 4 / BATCH + 16) * 8); double wfree[64];
    static const uint32_t Ks[4] = { 4, 8, 16, 32 };
    double t_serial = 0.0, t_pipe[4] = { 0, 0, 0, 0 }, t_fine[4] = { 0, 0, 0, 0 }, t_lane[4] = { 0, 0, 0, 0 };
    uint32_t* mbatch = malloc((size_t)(cap / 4 + 16) * 4); double* ft = malloc((size_t)(cap / 4 + 16) * 8); double* parsed = malloc((size_t)(cap / 4 / BATCH + 16) * 8);
    const double Tp = 4.0, Tc = 0.12, Th = 0.3;            /* microseconds: a lone warp spends ~8 us per 32 sequences today */
    while (at + 4 <= (size_t)sz) {
        uint32_t hw = rd32(d + at); at += 4;
        if (hw == 0) break;
        uint32_t n = hw & 0x7FFFFFFFu;
        const uint8_t* c = d + at; at += n + (bchk ? 4 : 0);
        if (hw & 0x80000000u) continue;
        ++blocks;
        uint32_t p = 0, o = 0, bdepth = 0, cur_piece = 0, pstart = 0;
        uint32_t npieces = 1; piece_wave[0] = 1;
        uint32_t nm = 0, nbat = 0, bseq = 0, bstart_o = 0; mfirst[0] = 0;
        for (;;) {
            /* an entry point is a sequence boundary: a new piece starts at the first sequence at or behind a multiple of P */
            if (o >= (uint64_t)(cur_piece + 1) * P) { cur_piece = o / P; pstart = o; piece_wave[npieces++] = 1; }
            if (bseq == BATCH) { bend[nbat] = o; mfirst[++nbat] = nm; bseq = 0; bstart_o = o; }
            ++bseq;
            uint32_t t = c[p++], ll = t >> 4;
            if (ll == 15) { uint32_t b; do { b = c[p++]; ll += b; } while (b == 255); }
            if (o + ll > cap) { fprintf(stderr, "block larger than 4 MB\n"); return 1; }
            memset(taint + o, 0, ll); memset(depth + o, 0, (size_t)ll * 4);
            p += ll; o += ll; ++seqs;
            if (p >= n) break;
            uint32_t off = c[p] | c[p + 1] << 8; p += 2;
            uint32_t ml = t & 15;
            if (ml == 15) { uint32_t b; do { b = c[p++]; ml += b; } while (b == 255); }
            ml += 4;
            if (off == 0 || off > o || o + ml > cap) { fprintf(stderr, "malformed block\n"); return 1; }
            ++matches; mbytes += ml;
            const uint32_t src = o - off;
            mt[nm].src = src; mt[nm].src_end = src + (ml < off ? ml : off); mt[nm].own = src >= bstart_o; mt[nm].o = o; mt[nm].o_end = o + ml; ++nm;
            if (src < pstart) {
                ++direct; if (off > 4096) ++far;
                /* piece-ordered execution: this piece runs one wave after the latest piece it reads from.  Pieces start at
                 * sequence boundaries, so find the source's piece by position among the piece starts seen so far: the one
                 * just in front is by far the common case; older ones only make the wave smaller, never larger. */
                uint32_t w = piece_wave[npieces - 2] + 1;
                if (w > piece_wave[npieces - 1]) piece_wave[npieces - 1] = w;
            }
            uint32_t dmax = 0;
            const uint32_t span = ml < off ? ml : off;
            for (uint32_t i = 0; i < span; ++i) if (depth[src + i] > dmax) dmax = depth[src + i];
            ++dmax; if (dmax > bdepth) bdepth = dmax;
            for (uint32_t i = 0; i < ml; ++i) {
                const uint32_t s = src + i;
                const uint8_t tn = s < pstart ? 1 : taint[s];
                taint[o + i] = tn; tbytes += tn; depth[o + i] = dmax;
            }
            o += ml;
        }
        bend[nbat] = o; mfirst[++nbat] = nm;
        t_serial += nbat * Tp + nm * Tc;
        for (int k = 0; k < 4; ++k) {
            t_pipe[k] += pipeline(nbat, mfirst, mt, bend, Ks[k], Tp, Tc, Th, done, pm, wfree, mbatch, parsed, ft, 0);
            t_fine[k] += pipeline(nbat, mfirst, mt, bend, Ks[k], Tp, Tc, Th, done, pm, wfree, mbatch, parsed, ft, 1);
            t_lane[k] += pipeline(nbat, mfirst, mt, bend, Ks[k], Tp, Tc, Th, done, pm, wfree, mbatch, parsed, ft, 2);
        }
        ubytes += o; depth_sum += bdepth; if (bdepth > depth_max) depth_max = bdepth;
        pieces += npieces;
        uint32_t wv = 0; for (uint32_t i = 0; i < npieces; ++i) if (piece_wave[i] > wv) wv = piece_wave[i];
        waves_sum += wv;
    }
    if (!blocks) { printf("no compressed blocks\n"); return 0; }
    printf("blocks %llu, %llu B out, sequences %llu, matches %llu (%.1f %% of the output bytes)\n", (unsigned long long)blocks,
           (unsigned long long)ubytes, (unsigned long long)seqs, (unsigned long long)matches, 100.0 * mbytes / (ubytes ? ubytes : 1));
    printf("P = %u: pieces %llu; matches that start in front of their piece: %llu = %.2f %% of all matches (%.0f per piece; %.0f %% of them "
           "more than 4 KB back)\n", P, (unsigned long long)pieces, (unsigned long long)direct, 100.0 * direct / (matches ? matches : 1),
           (double)direct / (pieces ? pieces : 1), 100.0 * far / (direct ? direct : 1));
    printf("output bytes that depend on bytes in front of their piece (transitively): %.1f %% of the output, %.1f %% of the match bytes\n",
           100.0 * tbytes / (ubytes ? ubytes : 1), 100.0 * tbytes / (mbytes ? mbytes : 1));
    printf("piece-ordered execution: %.1f waves per block on average for %.1f pieces per block\n", (double)waves_sum / blocks, (double)pieces / blocks);
    printf("copy-chain depth per block: avg %.0f, max %llu; sequences per block %.0f -> dataflow parallelism at best %.0f x\n",
           (double)depth_sum / blocks, (unsigned long long)depth_max, (double)seqs / blocks, depth_sum ? (double)seqs / depth_sum : 0.0);
    printf("pipeline model (Tp %.1f us per batch of %u, Tc %.2f us per match, Th %.1f us): one warp %.1f ms per block = %.0f MB/s", Tp, BATCH, Tc, Th,
           t_serial / blocks / 1e3, (double)ubytes / t_serial);
    printf("\n  watermark = longest prefix of finished batches:");
    for (int k = 0; k < 4; ++k) printf(" K=%u: %.1f x", Ks[k], t_serial / t_pipe[k]);
    printf("\n  every batch publishes its own progress after each match:");
    for (int k = 0; k < 4; ++k) printf(" K=%u: %.1f x (%.0f MB/s per block)", Ks[k], t_serial / t_fine[k], (double)ubytes / t_fine[k]);
    printf("\n  one match per lane, each waiting for its own sources only (%.2f us per lane copy):", 4.0 * Tc);
    for (int k = 0; k < 4; ++k) printf(" K=%u: %.1f x (%.0f MB/s per block)", Ks[k], t_serial / t_lane[k], (double)ubytes / t_lane[k]);
    printf("\n");
    return 0;
}
