/* tools/enc_model.c — CPU model of find_emit_kernel's PARSE DECISIONS (design exploration only: it
 * computes compressed sizes for parameter studies, it is not a codec and nothing links it).
 * Build: gcc -O2 -o /tmp/enc_model tools/enc_model.c tools/b2gen.c   Usage: enc_model MiB hlog sublog2 [variant] */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
void b2gen_loglike(uint8_t*, size_t, uint64_t);
static inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static uint32_t lenb(uint32_t v) { return v >= 15 ? (v - 15) / 255 + 1 : 0; }
int HLOG = 13, VARIANT = 0;
unsigned SEG = 0, PRIME = 1, MINLEN = 4, LAZY = 0;   /* SEG: encoder parse unit inside a sub-block (0 = whole sub-block); PRIME: 1 = insert every earlier position, 2/4 = every 2nd/4th, 9 = literal-ish heuristic */
static uint32_t hash4(uint32_t v) { return (v * 2654435761u) >> (32 - HLOG); }

/* returns compressed size of sub-block [b0,b1) of block `base` (usize bytes); *tail = carried literals */
static size_t sub(const uint8_t* base, uint32_t usize, uint32_t b0, uint32_t b1x, int last, uint32_t carry, uint32_t* tail, size_t* nseq, uint32_t s0, uint32_t b1)
{ (void)b1x;
    static uint16_t tab[1 << 16];
    memset(tab, 0, sizeof(uint16_t) << HLOG);
    if (s0 > b0) {   /* prime the table with the history of this sub-block */
        uint32_t from = s0 - b0 > 65535 ? s0 - 65535 : b0;
        for (uint32_t p = from; p < s0; p += PRIME) tab[hash4(rd32(base + p))] = (uint16_t)(p - b0);
    }
    uint32_t limit = 0;
    if (usize >= 13) { uint32_t la = b1 >= 3 ? b1 - 3 : 0, lb = usize - 11; limit = la < lb ? la : lb; }
    uint32_t matchlimit = usize >= 5 ? (b1 < usize - 5 ? b1 : usize - 5) : 0;
    uint32_t ip = s0, anchor = s0, miss = 0; size_t o = 0; int first = 1;
    while (ip < limit) {
        uint32_t sl = (miss >> 2) < 3 ? (miss >> 2) : 3;
        int f = -1; uint32_t cf = 0;
        uint32_t hs[32], ps[32]; int nv = 0;
        for (int lane = 0; lane < 32; ++lane) {
            uint32_t p = ip + ((uint32_t)lane << sl);
            if (p >= limit) break;
            uint32_t v = rd32(base + p), h = hash4(v), rel = p - b0;
            int32_t c = (int32_t)((rel & ~0xFFFFu) | tab[h]);
            if ((uint32_t)c >= rel) c -= 65536;
            if (f < 0 && c >= 0 && rel - (uint32_t)c < 65536 && rd32(base + b0 + c) == v) { f = lane; cf = (uint32_t)c; }
            hs[nv] = h; ps[nv] = rel; nv++;
            if (VARIANT == 1) tab[h] = (uint16_t)rel;        /* variant 1: serial visibility inside the batch */
            if (VARIANT == 2 && f >= 0) break;               /* variant 2: insert only up to the match start */
        }
        if (VARIANT != 1) for (int i = 0; i < nv; ++i) tab[hs[i]] = (uint16_t)ps[i];
        if (f < 0) { ip += 32u << sl; ++miss; continue; }
        miss = 0;
        uint32_t mpos = ip + ((uint32_t)f << sl), cpos = b0 + cf;
        while (mpos > anchor && cpos > b0 && base[mpos - 1] == base[cpos - 1]) { --mpos; --cpos; }
        uint32_t len = 0;
        while (mpos + len < matchlimit && base[mpos + len] == base[cpos + len]) ++len;
        if (len < MINLEN) { ip = mpos + 1; continue; }
        if (LAZY && len < LAZY) {
            int better = 0;
            uint32_t p0 = ip + ((uint32_t)f << sl);
            for (uint32_t d = 1; d <= 3 && !better; ++d) {
                uint32_t p = p0 + d; if (p >= limit) break;
                uint32_t v = rd32(base + p), rel = p - b0;
                int32_t c = (int32_t)((rel & ~0xFFFFu) | tab[hash4(v)]); if ((uint32_t)c >= rel) c -= 65536;
                if (c >= 0 && rel - (uint32_t)c < 65536 && rd32(base + b0 + c) == v) {
                    uint32_t l2 = 0; while (p + l2 < matchlimit && base[p + l2] == base[b0 + c + l2]) ++l2;
                    if (l2 >= len + 2) better = 1;
                }
            }
            if (better) { ip = p0 + 1; continue; }
        }
        uint32_t ll = mpos - anchor + (first ? carry : 0); first = 0;
        o += 1 + lenb(ll) + ll + 2 + lenb(len - 4); ++*nseq;
        ip = mpos + len; anchor = ip;
        if (ip - 2 < limit) tab[hash4(rd32(base + ip - 2))] = (uint16_t)(ip - 2 - b0);
    }
    uint32_t t = b1 - anchor;
    if (last) { uint32_t ll = t + (first ? carry : 0); o += 1 + lenb(ll) + ll; t = 0; ++*nseq; }
    else if (first) t += carry;
    *tail = t;
    return o;
}

int main(int argc, char** argv) {
    size_t n = (size_t)(argc > 1 ? atoi(argv[1]) : 32) << 20;
    HLOG = argc > 2 ? atoi(argv[2]) : 13;
    uint32_t S = 1u << (argc > 3 ? atoi(argv[3]) : 18);
    VARIANT = argc > 4 ? atoi(argv[4]) : 0;
    if (getenv("SEG")) SEG = (unsigned)atoi(getenv("SEG"));
    if (getenv("PRIME")) PRIME = (unsigned)atoi(getenv("PRIME"));
    if (getenv("LAZY")) LAZY = (unsigned)atoi(getenv("LAZY"));
    if (getenv("MINLEN")) MINLEN = (unsigned)atoi(getenv("MINLEN"));
    uint8_t* d = malloc(n);
    if (argc > 5) { FILE* f = fopen(argv[5], "rb"); n = fread(d, 1, n, f); fclose(f); }
    else b2gen_loglike(d, n, 0x10C5);
    size_t total = 7 + 4, nseq = 0;
    const uint32_t BS = 4u << 20;
    for (size_t b = 0; b < n; b += BS) {
        uint32_t usize = (uint32_t)(n - b < BS ? n - b : BS), carry = 0; size_t c = 0;
        for (uint32_t b0 = 0; b0 < usize; b0 += S) {
            uint32_t b1 = b0 + S < usize ? b0 + S : usize;
            uint32_t seg = SEG ? SEG : S;
            for (uint32_t s0 = b0; s0 < b1; s0 += seg) {
                uint32_t s1 = s0 + seg < b1 ? s0 + seg : b1;
                c += sub(d + b, usize, b0, b1, s1 == usize, carry, &carry, &nseq, s0, s1);
            }
        }
        total += 4 + (c >= usize ? usize : c);
    }
    printf("n=%zu hlog=%d sub=%u variant=%d -> %zu bytes ratio %.4f  seqs %zu (%.1f B/seq)\n", n, HLOG, S, VARIANT, total, (double)n / total, nseq, (double)n / nseq);
    return 0;
}
