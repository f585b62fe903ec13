/* tools/b2gen.c — deterministic synthetic inputs for the benchmark configs (bench/test utility, not
 * part of the codec).  Shapes follow SURVEY.md §8(d):
 *   b2gen_loglike : config 2/3 "log-like" text lines, splitmix64, seed 0x10C5 by default
 *   b2gen_random  : config 5 incompressible bytes (every block ends up stored)
 *   b2gen_textlike: a generator with the same statistical recipe the reference's datagen documents
 *                   (lz4/programs/datagen.c:54-153: skewed literal alphabet, 50 % match probability,
 *                   offsets <= 32 K, short/long lengths); NOT byte-identical to datagen — tests that
 *                   need the reference's exact generator run oracle/_ref/datagen instead.
 * Build: gcc -O2 -fPIC -shared -o tools/libb2gen.so tools/b2gen.c
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdio.h>

static inline uint64_t sm64(uint64_t* s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

void b2gen_random(uint8_t* dst, size_t n, uint64_t seed) {
    uint64_t s = seed;
    size_t i = 0;
    for (; i + 8 <= n; i += 8) { uint64_t v = sm64(&s); memcpy(dst + i, &v, 8); }
    if (i < n) { uint64_t v = sm64(&s); memcpy(dst + i, &v, n - i); }
}

static const char* LEVELS[5] = { "INFO", "DEBUG", "WARN", "ERROR", "TRACE" };
static const char* SVCS[24] = {
    "auth", "billing", "cart", "catalog", "checkout", "edge-proxy", "email", "fraud", "gateway", "geo",
    "identity", "inventory", "ledger", "metrics", "notify", "orders", "payments", "pricing", "search",
    "session", "shipping", "storage", "users", "webhooks" };
/* 64 message templates; '#' = decimal field, '$' = hex field */
static const char* TPL[64] = {
    "request completed status=200 bytes=# duration_ms=#",
    "request completed status=404 bytes=# duration_ms=#",
    "request failed status=500 upstream=$ duration_ms=#",
    "cache hit key=$ ttl=#",
    "cache miss key=$ backend_ms=#",
    "db query ok table=orders rows=# elapsed_us=#",
    "db query ok table=users rows=# elapsed_us=#",
    "db query slow table=ledger rows=# elapsed_us=# plan=$",
    "connection accepted peer=10.#.#.# port=#",
    "connection closed peer=10.#.#.# reason=idle_timeout",
    "retrying operation attempt=# backoff_ms=#",
    "circuit breaker opened target=payments failures=#",
    "circuit breaker closed target=payments after_ms=#",
    "queue depth=# consumers=# lag_ms=#",
    "published event topic=order.created partition=# offset=#",
    "published event topic=user.updated partition=# offset=#",
    "consumed event topic=order.created partition=# offset=# age_ms=#",
    "token validated subject=$ scope=read",
    "token validated subject=$ scope=write",
    "token rejected subject=$ reason=expired",
    "rate limit exceeded client=$ limit=# window_s=60",
    "gc pause young_ms=# old_ms=# heap_mb=#",
    "heartbeat ok uptime_s=# load=#",
    "config reloaded version=# keys=#",
    "scheduled job started name=reconcile run=#",
    "scheduled job finished name=reconcile run=# items=# elapsed_ms=#",
    "payment authorized amount_cents=# currency=USD txn=$",
    "payment captured amount_cents=# currency=USD txn=$",
    "payment declined amount_cents=# currency=EUR code=#",
    "shipment created order=# carrier=ups tracking=$",
    "shipment updated order=# status=in_transit",
    "inventory reserved sku=# qty=# warehouse=#",
    "inventory released sku=# qty=# warehouse=#",
    "search executed terms=# hits=# took_ms=#",
    "index refresh shard=# docs=# took_ms=#",
    "email queued template=welcome recipient_hash=$",
    "email delivered provider=ses message=$ latency_ms=#",
    "webhook delivered endpoint=$ status=200 attempt=#",
    "webhook failed endpoint=$ status=# attempt=#",
    "session created user=# device=$",
    "session expired user=# idle_s=#",
    "feature flag evaluated flag=new_checkout user=# result=true",
    "feature flag evaluated flag=new_checkout user=# result=false",
    "tls handshake ok version=1.3 cipher=aes256gcm ms=#",
    "dns lookup host=internal.svc.cluster.local ms=#",
    "upstream timeout target=$ after_ms=#",
    "thread pool saturated active=# queued=#",
    "file uploaded bucket=media size=# etag=$",
    "file deleted bucket=media key=$",
    "price recalculated sku=# old_cents=# new_cents=#",
    "fraud score computed order=# score=# model=$",
    "geo lookup ip=10.#.#.# country=US ms=#",
    "metrics flushed series=# points=# ms=#",
    "ledger entry posted account=# debit_cents=# credit_cents=#",
    "ledger balanced batch=# entries=#",
    "user profile updated user=# fields=#",
    "password reset requested user=#",
    "audit record written actor=# action=update resource=$",
    "snapshot started volume=$ size_gb=#",
    "snapshot finished volume=$ took_s=#",
    "leader elected node=# term=#",
    "replication lag follower=# bytes=#",
    "compaction finished level=# files=# bytes=# ms=#",
    "shutdown requested signal=15 draining_s=#" };

static char* put_str(char* p, const char* s) { while (*s) *p++ = *s++; return p; }
static char* put_dec(char* p, uint64_t v, int width) {   /* zero padded if width > 0 */
    char tmp[24]; int n = 0;
    do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (n < width) tmp[n++] = '0';
    while (n) *p++ = tmp[--n];
    return p;
}
static char* put_hex(char* p, uint64_t v, int digits) {
    for (int i = digits - 1; i >= 0; --i) *p++ = "0123456789abcdef"[(v >> (4 * i)) & 15];
    return p;
}

/* "2026-03-01T00:00:00.000Z LEVEL host-07 svc req=0123456789abcdef message...\n" */
void b2gen_loglike(uint8_t* dst, size_t n, uint64_t seed) {
    uint64_t s = seed;
    uint64_t ms = 0;                       /* milliseconds since 2026-03-01T00:00:00Z */
    size_t o = 0;
    char line[512];
    while (o < n) {
        uint64_t r = sm64(&s);
        ms += r % 37; r /= 37;
        uint64_t t = ms;
        unsigned milli = (unsigned)(t % 1000); t /= 1000;
        unsigned sec = (unsigned)(t % 60); t /= 60;
        unsigned min = (unsigned)(t % 60); t /= 60;
        unsigned hour = (unsigned)(t % 24); t /= 24;
        unsigned day = 1 + (unsigned)(t % 28);
        char* p = line;
        p = put_str(p, "2026-03-"); p = put_dec(p, day, 2); *p++ = 'T';
        p = put_dec(p, hour, 2); *p++ = ':'; p = put_dec(p, min, 2); *p++ = ':'; p = put_dec(p, sec, 2);
        *p++ = '.'; p = put_dec(p, milli, 3); p = put_str(p, "Z ");
        unsigned lv = (unsigned)(r % 16); r /= 16;       /* skewed: INFO 8/16, DEBUG 4, WARN 2, ERROR 1, TRACE 1 */
        p = put_str(p, LEVELS[lv < 8 ? 0 : lv < 12 ? 1 : lv < 14 ? 2 : lv < 15 ? 3 : 4]); *p++ = ' ';
        p = put_str(p, "host-"); p = put_dec(p, r % 32, 2); r /= 32; *p++ = ' ';
        p = put_str(p, SVCS[r % 24]); r /= 24;
        p = put_str(p, " req="); p = put_hex(p, sm64(&s), 16); *p++ = ' ';
        const char* tp = TPL[r % 64];
        for (; *tp; ++tp) {
            if (*tp == '#') { uint64_t q = sm64(&s); unsigned k = (unsigned)(q & 3); q >>= 2;
                              p = put_dec(p, k == 0 ? q % 10 : k == 1 ? q % 256 : k == 2 ? q % 5000 : q % 1000000, 0); }
            else if (*tp == '$') p = put_hex(p, sm64(&s), 8);
            else *p++ = *tp;
        }
        *p++ = '\n';
        size_t len = (size_t)(p - line);
        if (len > n - o) len = n - o;
        memcpy(dst + o, line, len);
        o += len;
    }
}

/* text-like: literals from a skewed 64-symbol alphabet, matches copied from <= 32 KB back */
void b2gen_textlike(uint8_t* dst, size_t n, uint64_t seed) {
    uint64_t s = seed;
    uint8_t lut[4096];
    {   /* skewed literal table over '('..'}' */
        size_t k = 0; unsigned c = '(';
        double w = 0.22;
        while (k < sizeof lut) {
            size_t cnt = (size_t)(w * (double)(sizeof lut - k)) + 1;
            for (size_t i = 0; i < cnt && k < sizeof lut; ++i) lut[k++] = (uint8_t)c;
            if (++c > '}') c = '(';
        }
    }
    size_t o = 0;
    while (o < n) {
        uint64_t r = sm64(&s);
        int is_match = (r & 1) && o > 16; r >>= 1;
        size_t len = (r & 1) ? (size_t)((r >> 1) & 15) + 1 : (size_t)((r >> 1) & 511) + 15;
        r >>= 11;
        if (is_match) {
            size_t maxoff = o < 32768 ? o : 32768;
            size_t off = (size_t)(r % maxoff) + 1;
            len += 4;
            if (len > n - o) len = n - o;
            for (size_t i = 0; i < len; ++i) dst[o + i] = dst[o + i - off];
            o += len;
        } else {
            len = (len & 63) + 1;
            if (len > n - o) len = n - o;
            for (size_t i = 0; i < len; ++i) { dst[o + i] = lut[sm64(&s) & 4095]; }
            o += len;
        }
    }
}
