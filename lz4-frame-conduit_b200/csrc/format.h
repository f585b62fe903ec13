// format.h — LZ4 frame wire format: constants, header encode/parse, tiny XXH32.
// Shared by host code and device code (B2_HD functions compile for both).
//
// Replaces, on this path, the reference's frame-header logic:
//   header write + header checksum   /root/reference/lz4/lib/lz4frame.c:663-701, :294-298
//   header parse / validation        lz4frame.c:1109-1198, :1205-1229
//   block size table                 lz4frame.c:278-287
// Wire format: lz4/doc/lz4_Frame_format.md.  Written from the format description, not from the
// reference sources.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD inline
#endif

namespace b2 {

// ---- error codes: same numbering as the reference's list (lz4frame.h:503-524) so that the
// (size_t)-code convention and LZ4F_getErrorName stay drop-in.  Codes >= 20 are ours.
enum Err : uint32_t {
    OK = 0, E_GENERIC, E_maxBlockSize_invalid, E_blockMode_invalid, E_contentChecksumFlag_invalid,
    E_compressionLevel_invalid, E_headerVersion_wrong, E_blockChecksum_invalid, E_reservedFlag_set,
    E_allocation_failed, E_srcSize_tooLarge, E_dstMaxSize_tooSmall, E_frameHeader_incomplete,
    E_frameType_unknown, E_frameSize_wrong, E_srcPtr_wrong, E_decompressionFailed,
    E_headerChecksum_invalid, E_contentChecksum_invalid, E_frameDecoding_alreadyStarted,
    E_cudaFailure,          // a CUDA runtime call or kernel launch failed
    E_noEndMark,            // input ended before the EndMark (Conduit.hsc:689 throws for this)
    E_maxCode
};

constexpr uint32_t MAGIC_FRAME     = 0x184D2204u;
constexpr uint32_t MAGIC_SKIP_LO   = 0x184D2A50u;   // .. 0x184D2A5F
constexpr uint32_t MAGIC_LEGACY    = 0x184C2102u;
constexpr uint32_t BLOCK_STORED    = 0x80000000u;
constexpr int      HEADER_MIN      = 7;
constexpr int      HEADER_MAX      = 19;

struct FrameDesc {
    uint32_t block_size_id;      // 4..7
    uint32_t block_independent;  // FLG bit 5
    uint32_t block_checksum;     // FLG bit 4
    uint32_t has_content_size;   // FLG bit 3
    uint32_t content_checksum;   // FLG bit 2
    uint32_t has_dict_id;        // FLG bit 0
    uint32_t dict_id;
    uint32_t header_size;        // filled by parse
    uint64_t content_size;
};

B2_HD uint32_t block_size_of(uint32_t id) {          // 64 KB << 2*(id-4); id 0 = default = 64 KB
    if (id == 0) id = 4;
    return (id < 4 || id > 7) ? 0u : (65536u << (2u * (id - 4u)));
}

B2_HD uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

constexpr uint32_t XP1 = 0x9E3779B1u, XP2 = 0x85EBCA77u, XP3 = 0xC2B2AE3Du, XP4 = 0x27D4EB2Fu, XP5 = 0x165667B1u;

B2_HD uint32_t xxh32_round(uint32_t acc, uint32_t in) { return rotl32(acc + in * XP2, 13) * XP1; }
B2_HD uint32_t xxh32_merge(uint32_t v1, uint32_t v2, uint32_t v3, uint32_t v4) {
    return rotl32(v1, 1) + rotl32(v2, 7) + rotl32(v3, 12) + rotl32(v4, 18);
}
B2_HD uint32_t xxh32_avalanche(uint32_t h) {
    h ^= h >> 15; h *= XP2; h ^= h >> 13; h *= XP3; h ^= h >> 16; return h;
}
// tail: `h` already includes the total length; p[0..n) with n < 16 are the bytes after the last stripe
B2_HD uint32_t xxh32_tail(uint32_t h, const uint8_t* p, uint32_t n) {
    while (n >= 4) {
        uint32_t w = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
        h = rotl32(h + w * XP3, 17) * XP4; p += 4; n -= 4;
    }
    while (n) { h = rotl32(h + (uint32_t)(*p++) * XP5, 11) * XP1; --n; }
    return xxh32_avalanche(h);
}
// XXH32 of fewer than 16 bytes, seed 0 (enough for the <= 14 descriptor bytes of a frame header)
B2_HD uint32_t xxh32_short(const uint8_t* p, uint32_t n) { return xxh32_tail(XP5 + n, p, n); }

B2_HD uint32_t ld_le32(const uint8_t* p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
B2_HD void st_le32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

// Writes magic + descriptor + HC.  Returns the header size (7..19).  `dst` needs HEADER_MAX bytes.
B2_HD uint32_t header_write(uint8_t* dst, const FrameDesc& d) {
    st_le32(dst, MAGIC_FRAME);
    uint8_t* q = dst + 4;
    uint32_t bsid = d.block_size_id ? d.block_size_id : 4;
    q[0] = (uint8_t)((1u << 6) | ((d.block_independent & 1u) << 5) | ((d.block_checksum & 1u) << 4) |
                     ((d.has_content_size & 1u) << 3) | ((d.content_checksum & 1u) << 2) | (d.has_dict_id & 1u));
    q[1] = (uint8_t)((bsid & 7u) << 4);
    uint32_t n = 2;
    if (d.has_content_size) {
        for (int i = 0; i < 8; ++i) q[n + i] = (uint8_t)(d.content_size >> (8 * i));
        n += 8;
    }
    if (d.has_dict_id) { st_le32(q + n, d.dict_id); n += 4; }
    q[n] = (uint8_t)(xxh32_short(q, n) >> 8);
    return 4 + n + 1;
}

// Size of the header that starts at src (needs >= 5 bytes).  0 = not a normal LZ4 frame.
B2_HD uint32_t header_size_from_flg(uint8_t flg) {
    return 7u + ((flg >> 3) & 1u) * 8u + (flg & 1u) * 4u;
}

// Parses and validates a normal frame header.  Returns OK or an Err; fills d (incl. header_size).
B2_HD Err header_parse(const uint8_t* src, size_t n, FrameDesc& d) {
    if (n < (size_t)HEADER_MIN) return E_frameHeader_incomplete;
    uint32_t magic = ld_le32(src);
    if (magic != MAGIC_FRAME) return E_frameType_unknown;
    uint8_t flg = src[4];
    uint32_t hs = header_size_from_flg(flg);
    if (n < hs) return E_frameHeader_incomplete;
    uint8_t bd = src[5];
    if (((flg >> 6) & 3u) != 1u) return E_headerVersion_wrong;
    if ((flg >> 1) & 1u) return E_reservedFlag_set;
    if ((bd >> 7) & 1u) return E_reservedFlag_set;
    uint32_t bsid = (bd >> 4) & 7u;
    if (bsid < 4) return E_maxBlockSize_invalid;
    if (bd & 0x0Fu) return E_reservedFlag_set;
    uint8_t hc = (uint8_t)(xxh32_short(src + 4, hs - 5) >> 8);
    if (hc != src[hs - 1]) return E_headerChecksum_invalid;
    d.block_size_id = bsid;
    d.block_independent = (flg >> 5) & 1u;
    d.block_checksum = (flg >> 4) & 1u;
    d.has_content_size = (flg >> 3) & 1u;
    d.content_checksum = (flg >> 2) & 1u;
    d.has_dict_id = flg & 1u;
    d.content_size = 0; d.dict_id = 0;
    uint32_t o = 6;
    if (d.has_content_size) {
        uint64_t v = 0;
        for (int i = 0; i < 8; ++i) v |= (uint64_t)src[o + i] << (8 * i);
        d.content_size = v; o += 8;
    }
    if (d.has_dict_id) { d.dict_id = ld_le32(src + o); o += 4; }
    d.header_size = hs;
    return OK;
}

}  // namespace b2
