/* oracle/oracle_lz4.c — CPU restatement of the reference hot path (TEST INFRASTRUCTURE ONLY).
 * See oracle_lz4.h for the parity status.  Citations are relative to /root/reference/.
 * Little-endian hosts only (like the GPU).
 */
#include "oracle_lz4.h"
#include <string.h>
#include <stdlib.h>

static inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static inline uint64_t rd64(const uint8_t* p) { uint64_t v; memcpy(&v, p, 8); return v; }
static inline uint16_t rd16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
static inline void wr32(uint8_t* p, uint32_t v) { memcpy(p, &v, 4); }
static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

/* ------------------------------------------------------------------ XXH32 (xxhash.c:263-554) */
#define P1 0x9E3779B1U   /* xxhash.c:263-267 */
#define P2 0x85EBCA77U
#define P3 0xC2B2AE3DU
#define P4 0x27D4EB2FU
#define P5 0x165667B1U

static inline uint32_t xxh_round(uint32_t acc, uint32_t in) {      /* xxhash.c:269-275 */
    return rotl32(acc + in * P2, 13) * P1;
}
static uint32_t xxh_tail(uint32_t h, const uint8_t* p, size_t len) { /* finalize + avalanche, :277-348 */
    len &= 15;
    while (len >= 4) { h = rotl32(h + rd32(p) * P3, 17) * P4; p += 4; len -= 4; }
    while (len > 0)  { h = rotl32(h + (*p++) * P5, 11) * P1; len--; }
    h ^= h >> 15; h *= P2; h ^= h >> 13; h *= P3; h ^= h >> 16;
    return h;
}
uint32_t orc_xxh32(const void* data, size_t len, uint32_t seed) {   /* xxhash.c:350-416 */
    const uint8_t* p = (const uint8_t*)data;
    const uint8_t* end = p + len;
    uint32_t h;
    if (len >= 16) {
        uint32_t v1 = seed + P1 + P2, v2 = seed + P2, v3 = seed, v4 = seed - P1;
        const uint8_t* lim = end - 15;
        do {
            v1 = xxh_round(v1, rd32(p)); v2 = xxh_round(v2, rd32(p + 4));
            v3 = xxh_round(v3, rd32(p + 8)); v4 = xxh_round(v4, rd32(p + 12));
            p += 16;
        } while (p < lim);
        h = rotl32(v1, 1) + rotl32(v2, 7) + rotl32(v3, 12) + rotl32(v4, 18);
    } else {
        h = seed + P5;
    }
    h += (uint32_t)len;
    return xxh_tail(h, p, len);
}
void orc_xxh32_reset(orc_xxh32_state* s, uint32_t seed) {           /* xxhash.c:437-449 */
    memset(s, 0, sizeof *s);
    s->v[0] = seed + P1 + P2; s->v[1] = seed + P2; s->v[2] = seed; s->v[3] = seed - P1;
}
void orc_xxh32_update(orc_xxh32_state* s, const void* data, size_t len) { /* xxhash.c:452-526 */
    const uint8_t* p = (const uint8_t*)data;
    const uint8_t* end = p + len;
    if (!p) return;
    s->total_len += (uint32_t)len;                                   /* wraps mod 2^32 by design */
    s->large |= (uint32_t)((len >= 16) | (s->total_len >= 16));
    if (s->memsize + len < 16) {
        memcpy((uint8_t*)s->mem + s->memsize, p, len);
        s->memsize += (uint32_t)len;
        return;
    }
    if (s->memsize) {
        memcpy((uint8_t*)s->mem + s->memsize, p, 16 - s->memsize);
        for (int i = 0; i < 4; i++) s->v[i] = xxh_round(s->v[i], s->mem[i]);
        p += 16 - s->memsize;
        s->memsize = 0;
    }
    while (p + 16 <= end) {
        for (int i = 0; i < 4; i++) s->v[i] = xxh_round(s->v[i], rd32(p + 4 * i));
        p += 16;
    }
    if (p < end) {
        memcpy(s->mem, p, (size_t)(end - p));
        s->memsize = (uint32_t)(end - p);
    }
}
uint32_t orc_xxh32_digest(const orc_xxh32_state* s) {               /* xxhash.c:529-554 */
    uint32_t h;
    if (s->large) h = rotl32(s->v[0], 1) + rotl32(s->v[1], 7) + rotl32(s->v[2], 12) + rotl32(s->v[3], 18);
    else          h = s->v[2] /* == seed */ + P5;
    h += s->total_len;
    return xxh_tail(h, (const uint8_t*)s->mem, s->memsize);
}

/* ------------------------------------------------------------------ block compressor
 * Restates LZ4_compress_generic (lz4.c:789-1177) as instantiated by
 * LZ4_compress_fast_extState_fastReset (lz4.c:1211-1245) for the frame layer's independent blocks:
 * limitedOutput, noDict, acceleration 1, table freshly zeroed, currentOffset 0.
 *   srcSize <  65547 : u16 table, 8192 entries, 4-byte hash      (lz4.c:631-637, :566)
 *   otherwise        : u32 table, 4096 entries, 5-byte hash      (lz4.c:639-650)
 */
enum { MINMATCH = 4, LASTLITERALS = 5, MFLIMIT = 12, MIN_LENGTH = MFLIMIT + 1,
       LIMIT_64K = 65536 + (MFLIMIT - 1), MAX_DISTANCE = 65535, SKIP_TRIGGER = 6,
       ML_BITS = 4, ML_MASK = 15, RUN_MASK = 15 };

typedef struct { int small; uint32_t t32[4096]; uint16_t* t16; } ctab;   /* 16 KB either way */

static inline uint32_t hash_at(const ctab* t, const uint8_t* p) {
    if (t->small) return (rd32(p) * 2654435761U) >> (32 - 13);
    return (uint32_t)(((rd64(p) << 24) * 889523592379ULL) >> (64 - 12));
}
static inline uint32_t tab_get(const ctab* t, uint32_t h) { return t->small ? t->t16[h] : t->t32[h]; }
static inline void tab_put(ctab* t, uint32_t h, uint32_t idx) {
    if (t->small) t->t16[h] = (uint16_t)idx; else t->t32[h] = idx;
}
static unsigned count_eq(const uint8_t* a, const uint8_t* b, const uint8_t* alimit) { /* lz4.c:536-559 */
    const uint8_t* s = a;
    while (a < alimit && *a == *b) { a++; b++; }
    return (unsigned)(a - s);
}

int orc_block_compress(const uint8_t* src, int n, uint8_t* dst, int cap)
{
    ctab tab;
    if (n < 0 || (uint32_t)n > 0x7E000000U) return 0;                 /* lz4.c:840 */
    memset(tab.t32, 0, sizeof tab.t32);
    tab.small = n < LIMIT_64K;
    tab.t16 = (uint16_t*)tab.t32;

    const uint8_t* ip = src;
    const uint8_t* anchor = src;
    const uint8_t* const iend = src + n;
    const uint8_t* const mflimit1 = iend - MFLIMIT + 1;
    const uint8_t* const matchlimit = iend - LASTLITERALS;
    uint8_t* op = dst;
    uint8_t* const olimit = dst + cap;
    uint8_t* token;
    const uint8_t* match;
    uint32_t forwardH;

    if (n < MIN_LENGTH) goto last_literals;                           /* lz4.c:859 */

    tab_put(&tab, hash_at(&tab, ip), 0);                              /* lz4.c:862 */
    ip++; forwardH = hash_at(&tab, ip);

    for (;;) {
        {   /* search, lz4.c:893-951 */
            const uint8_t* fwd = ip;
            int step = 1;
            int probes = 1 << SKIP_TRIGGER;
            for (;;) {
                uint32_t h = forwardH;
                uint32_t cur = (uint32_t)(fwd - src);
                uint32_t cand = tab_get(&tab, h);
                ip = fwd;
                fwd += step;
                step = (probes++ >> SKIP_TRIGGER);
                if (fwd > mflimit1) goto last_literals;
                match = src + cand;
                forwardH = hash_at(&tab, fwd);
                tab_put(&tab, h, cur);
                if (!tab.small && cand + MAX_DISTANCE < cur) continue; /* too far, lz4.c:939-942 */
                if (rd32(match) == rd32(ip)) break;
            }
        }
        /* catch up, lz4.c:956-957 */
        while (ip > anchor && match > src && ip[-1] == match[-1]) { ip--; match--; }

        {   /* literals, lz4.c:960-984 */
            unsigned lit = (unsigned)(ip - anchor);
            token = op++;
            if (op + lit + (2 + 1 + LASTLITERALS) + (lit / 255) > olimit) return 0;
            if (lit >= RUN_MASK) {
                int len = (int)(lit - RUN_MASK);
                *token = RUN_MASK << ML_BITS;
                for (; len >= 255; len -= 255) *op++ = 255;
                *op++ = (uint8_t)len;
            } else *token = (uint8_t)(lit << ML_BITS);
            memcpy(op, anchor, lit);
            op += lit;
        }
    next_match:
        {   /* offset + match length, lz4.c:1004-1074 */
            unsigned off = (unsigned)(ip - match);
            op[0] = (uint8_t)off; op[1] = (uint8_t)(off >> 8); op += 2;
            unsigned code = count_eq(ip + MINMATCH, match + MINMATCH, matchlimit);
            ip += code + MINMATCH;
            if (op + (1 + LASTLITERALS) + (code + 240) / 255 > olimit) return 0;
            if (code >= ML_MASK) {
                *token += ML_MASK;
                code -= ML_MASK;
                for (; code >= 255; code -= 255) *op++ = 255;
                *op++ = (uint8_t)code;
            } else *token += (uint8_t)code;
        }
        anchor = ip;
        if (ip >= mflimit1) break;                                    /* lz4.c:1081 */
        tab_put(&tab, hash_at(&tab, ip - 2), (uint32_t)(ip - 2 - src)); /* lz4.c:1084 */
        {   /* immediate re-probe, lz4.c:1096-1132 */
            uint32_t h = hash_at(&tab, ip);
            uint32_t cur = (uint32_t)(ip - src);
            uint32_t cand = tab_get(&tab, h);
            match = src + cand;
            tab_put(&tab, h, cur);
            if ((tab.small || cand + MAX_DISTANCE >= cur) && rd32(match) == rd32(ip)) {
                token = op++; *token = 0;
                goto next_match;
            }
        }
        forwardH = hash_at(&tab, ++ip);                               /* lz4.c:1136 */
    }

last_literals:
    {   size_t last = (size_t)(iend - anchor);                        /* lz4.c:1140-1166 */
        if (op + last + 1 + ((last + 255 - RUN_MASK) / 255) > olimit) return 0;
        if (last >= RUN_MASK) {
            size_t acc = last - RUN_MASK;
            *op++ = RUN_MASK << ML_BITS;
            for (; acc >= 255; acc -= 255) *op++ = 255;
            *op++ = (uint8_t)acc;
        } else *op++ = (uint8_t)(last << ML_BITS);
        memcpy(op, anchor, last);
        op += last;
    }
    return (int)(op - dst);
}

/* ------------------------------------------------------------------ block decoder
 * Restates the safe loop of LZ4_decompress_generic (lz4.c:1844-2055) for endOnInputSize / full
 * decoding with an optional prefix (history immediately before dst), including its acceptance
 * rules: a non-final sequence must leave >= 8 input bytes after its literals and >= 12 output bytes
 * of capacity (lz4.c:1906); the final sequence must consume the input exactly (:1941); a match may
 * not reach into the last 5 bytes of the capacity (:2035); offsets may not reach before the
 * history (:1977).  Deviation (malformed input only): offset 0 is rejected here, the reference
 * v1.9.2 lets it through and reproduces zero bytes (lz4.c:2027).
 */
int orc_block_decompress(const uint8_t* src, int csize, uint8_t* dst, int cap, size_t prefix)
{
    if (!src || csize < 0 || cap < 0) return -1;
    const uint8_t* ip = src;
    const uint8_t* const iend = src + csize;
    uint8_t* op = dst;
    uint8_t* const oend = dst + cap;
    const uint8_t* const low = dst - prefix;
#define FAIL() return (int)(-(ip - src)) - 1
    if (cap == 0) return (csize == 1 && *ip == 0) ? 0 : -1;           /* lz4.c:1696-1700 */
    if (csize == 0) return -1;                                        /* lz4.c:1702 */
    for (;;) {
        unsigned token = *ip++;
        size_t len = token >> ML_BITS;
        if (len == RUN_MASK) {                                        /* lz4.c:1882-1889, :1625-1645 */
            unsigned s;
            if (ip >= iend - RUN_MASK) FAIL();
            /* a hit of the input guard inside the loop only stops the accumulation (loop_error is
             * ignored for literal lengths, lz4.c:1893-1894); the exact-consumption check below decides */
            do { s = *ip++; len += s; if (ip >= iend - RUN_MASK) break; } while (s == 255);
        }
        uint8_t* cpy = op + len;
        if (cpy > oend - MFLIMIT || ip + len > iend - (2 + 1 + LASTLITERALS)) {  /* lz4.c:1906 */
            if (ip + len != iend || cpy > oend) FAIL();               /* lz4.c:1941 */
            memmove(op, ip, len);
            op += len; ip += len;
            break;
        }
        memcpy(op, ip, len);
        ip += len; op = cpy;

        size_t offset = rd16(ip); ip += 2;                            /* lz4.c:1960 */
        const uint8_t* match = op - offset;
        len = token & ML_MASK;
        if (len == ML_MASK) {                                         /* lz4.c:1967-1972 */
            unsigned s;
            do { s = *ip++; len += s; if (ip >= iend - LASTLITERALS + 1) FAIL(); } while (s == 255);
        }
        len += MINMATCH;
        if (offset == 0) FAIL();
        if (match < low) FAIL();                                      /* lz4.c:1977 */
        cpy = op + len;
        if (cpy > oend - LASTLITERALS) FAIL();                        /* lz4.c:2033-2035 */
        while (op < cpy) *op++ = *match++;                            /* overlap-safe byte copy */
    }
#undef FAIL
    return (int)(op - dst);
}

/* ------------------------------------------------------------------ frame layer */
#define MAGIC            0x184D2204U   /* lz4frame.c:210 */
#define MAGIC_SKIP_START 0x184D2A50U   /* lz4frame.c:209 */
#define UNCOMPRESSED_FLAG 0x80000000U  /* lz4frame.c:211 */

size_t orc_block_size(int id) {                                       /* lz4frame.c:278-287 */
    if (id == 0) id = 4;
    if (id < 4 || id > 7) return ORC_ERROR(ORC_ERR_maxBlockSize_invalid);
    return (size_t)65536 << (2 * (id - 4));
}

size_t orc_frame_header_write(uint8_t* dst, size_t cap, const orc_frame_desc* d) { /* lz4frame.c:663-701 */
    if (cap < 19) return ORC_ERROR(ORC_ERR_dstMaxSize_tooSmall);
    int id = d->block_size_id ? d->block_size_id : 4;
    if (id < 4 || id > 7) return ORC_ERROR(ORC_ERR_maxBlockSize_invalid);
    uint8_t* p = dst;
    /* the reference only emits the field when contentSize != 0 (lz4frame.c:679, :686-690) */
    int has_cs = d->has_content_size && d->content_size != 0;
    wr32(p, MAGIC); p += 4;
    uint8_t* desc = p;
    *p++ = (uint8_t)((1 << 6) | ((d->block_independent & 1) << 5) | ((d->block_checksum & 1) << 4)
                     | (has_cs << 3) | ((d->content_checksum & 1) << 2)
                     | (d->has_dict_id & 1));
    *p++ = (uint8_t)((id & 7) << 4);
    if (has_cs) { uint64_t v = d->content_size; memcpy(p, &v, 8); p += 8; }
    if (d->has_dict_id) { wr32(p, d->dict_id); p += 4; }
    *p = (uint8_t)(orc_xxh32(desc, (size_t)(p - desc), 0) >> 8);      /* lz4frame.c:294-298 */
    p++;
    return (size_t)(p - dst);
}

size_t orc_frame_header_parse(const uint8_t* src, size_t n, orc_frame_desc* d) { /* lz4frame.c:1109-1198 */
    memset(d, 0, sizeof *d);
    if (n < 7) return ORC_ERROR(ORC_ERR_frameHeader_incomplete);
    if (rd32(src) != MAGIC) return ORC_ERROR(ORC_ERR_frameType_unknown);
    unsigned flg = src[4];
    if ((flg >> 1) & 1) return ORC_ERROR(ORC_ERR_reservedFlag_set);
    if (((flg >> 6) & 3) != 1) return ORC_ERROR(ORC_ERR_headerVersion_wrong);
    d->block_independent = (flg >> 5) & 1;
    d->block_checksum = (flg >> 4) & 1;
    d->has_content_size = (flg >> 3) & 1;
    d->content_checksum = (flg >> 2) & 1;
    d->has_dict_id = flg & 1;
    size_t hs = 7 + (d->has_content_size ? 8 : 0) + (d->has_dict_id ? 4 : 0);
    if (n < hs) return ORC_ERROR(ORC_ERR_frameHeader_incomplete);
    unsigned bd = src[5];
    if (bd & 0x80) return ORC_ERROR(ORC_ERR_reservedFlag_set);
    d->block_size_id = (bd >> 4) & 7;
    if (d->block_size_id < 4) return ORC_ERROR(ORC_ERR_maxBlockSize_invalid);
    if (bd & 0x0F) return ORC_ERROR(ORC_ERR_reservedFlag_set);
    if ((uint8_t)(orc_xxh32(src + 4, hs - 5, 0) >> 8) != src[hs - 1])
        return ORC_ERROR(ORC_ERR_headerChecksum_invalid);
    if (d->has_content_size) memcpy(&d->content_size, src + 6, 8);
    if (d->has_dict_id) d->dict_id = rd32(src + hs - 5);
    return hs;
}

size_t orc_frame_compress_bound(size_t n, const orc_frame_desc* d) {
    size_t bs = orc_block_size(d->block_size_id);
    size_t nb = (n + bs - 1) / bs;
    return 19 + nb * (4 + (d->block_checksum ? 4 : 0)) + n + 4 + 4;
}

size_t orc_frame_compress(const uint8_t* src, size_t n, uint8_t* dst, size_t cap, const orc_frame_desc* d)
{
    if (!d->block_independent) return ORC_ERROR(ORC_ERR_blockMode_invalid); /* linked: use oracle/_ref */
    if (cap < orc_frame_compress_bound(n, d)) return ORC_ERROR(ORC_ERR_dstMaxSize_tooSmall);
    size_t bs = orc_block_size(d->block_size_id);
    if (orc_is_error(bs)) return bs;
    size_t o = orc_frame_header_write(dst, cap, d);
    if (orc_is_error(o)) return o;
    for (size_t pos = 0; pos < n; pos += bs) {                        /* LZ4F_makeBlock, lz4frame.c:739-761 */
        size_t sz = n - pos < bs ? n - pos : bs;
        uint32_t c = (uint32_t)orc_block_compress(src + pos, (int)sz, dst + o + 4, (int)sz - 1);
        if (c == 0) {
            c = (uint32_t)sz;
            wr32(dst + o, c | UNCOMPRESSED_FLAG);
            memcpy(dst + o + 4, src + pos, sz);
        } else wr32(dst + o, c);
        o += 4 + c;
        if (d->block_checksum) { wr32(dst + o, orc_xxh32(dst + o - c, c, 0)); o += 4; }
    }
    wr32(dst + o, 0); o += 4;                                         /* EndMark, lz4frame.c:998-1000 */
    if (d->content_checksum) { wr32(dst + o, orc_xxh32(src, n, 0)); o += 4; } /* lz4frame.c:1002-1007 */
    return o;
}

size_t orc_frame_decompress(const uint8_t* src, size_t n, uint8_t* dst, size_t cap, size_t* consumed)
{
    orc_frame_desc d;
    size_t ip = orc_frame_header_parse(src, n, &d);
    if (consumed) *consumed = 0;
    if (orc_is_error(ip)) return ip;
    size_t bs = orc_block_size(d.block_size_id);
    size_t op = 0;
    for (;;) {                                                        /* lz4frame.c:1470-1705 */
        if (n - ip < 4) return ORC_ERROR(ORC_ERR_frameHeader_incomplete);
        uint32_t hdr = rd32(src + ip); ip += 4;
        uint32_t c = hdr & 0x7FFFFFFFU;
        if (c == 0) break;                                            /* EndMark, :1485-1488 */
        if (c > bs) return ORC_ERROR(ORC_ERR_maxBlockSize_invalid);   /* :1489-1490 */
        if (n - ip < (size_t)c + (d.block_checksum ? 4 : 0)) return ORC_ERROR(ORC_ERR_frameHeader_incomplete);
        if (d.block_checksum && rd32(src + ip + c) != orc_xxh32(src + ip, c, 0))
            return ORC_ERROR(ORC_ERR_blockChecksum_invalid);          /* :1604-1611, :1541-1548 */
        if (hdr & UNCOMPRESSED_FLAG) {                                /* :1491-1530 */
            if (cap - op < c) return ORC_ERROR(ORC_ERR_dstMaxSize_tooSmall);
            memcpy(dst + op, src + ip, c);
            op += c;
        } else {
            size_t room = cap - op < bs ? cap - op : bs;
            size_t prefix = d.block_independent ? 0 : (op < 65536 ? op : 65536);
            int r = orc_block_decompress(src + ip, (int)c, dst + op, (int)room, prefix);
            if (r < 0) return ORC_ERROR(room < bs ? ORC_ERR_dstMaxSize_tooSmall : ORC_ERR_decompressionFailed);
            op += (size_t)r;
        }
        ip += c + (d.block_checksum ? 4 : 0);
    }
    if (d.has_content_size && d.content_size != op) return ORC_ERROR(ORC_ERR_frameSize_wrong); /* :1715-1716 */
    if (d.content_checksum) {                                         /* :1734-1745 */
        if (n - ip < 4) return ORC_ERROR(ORC_ERR_frameHeader_incomplete);
        if (rd32(src + ip) != orc_xxh32(dst, op, 0)) return ORC_ERROR(ORC_ERR_contentChecksum_invalid);
        ip += 4;
    }
    if (consumed) *consumed = ip;
    return op;
}
