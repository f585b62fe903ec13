// b200lz4c — command-line twin of the reference's `haskell-lz4c` (app/Main.hs:20-64) over libb200lz4.so.
//
//   b200lz4c [INPUT_FILE] [OUTPUT_FILE] [-d|--decompress]        ("-" or nothing = stdin / stdout, as in app/Main.hs:48-58)
//
// Compression follows the `compress` conduit's call sequence (src/Codec/Compression/LZ4/Conduit.hsc:457-533):
// LZ4F_compressBegin -> header, LZ4F_compressUpdate per input chunk, LZ4F_compressEnd; decompression follows
// `decompress` (Conduit.hsc:598-701): LZ4F_decompress until it returns 0, "stream ended before EndMark" otherwise.
// Beyond the reference program:
//   -B4..-B7, --block-checksum, --content-checksum   the Preferences the conduit API cannot reach (Conduit.hsc:347)
//   -BD              linked blocks (blockMode = linked taken literally: matches cross block borders, like `lz4 -BD`)
//   -D FILE          dictionary (LZ4F_compressBegin_usingCDict / LZ4F_decompress_usingDict; frames become linked-block frames)
//   --all-frames     decode every concatenated frame, skip skippable frames, decode legacy frames — what `lz4 -d`
//                    does (lz4/programs/lz4io.c:1094-1135, 895-940); default = first frame only, like the conduit
// All codec work happens on the GPU inside the library; this file only moves bytes between files and the C ABI.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "b200lz4.h"
#include "lz4frame.h"

namespace {

[[noreturn]] void die(const std::string& msg) { fprintf(stderr, "b200lz4c: %s\n", msg.c_str()); exit(1); }
size_t check(size_t r) {                      // handleLz4Error (Conduit.hsc:145-160)
    if (LZ4F_isError(r)) die(std::string("lz4frame error: ") + LZ4F_getErrorName(r));
    return r;
}
void put(FILE* f, const void* p, size_t n) { if (n && fwrite(p, 1, n, f) != n) die("write error"); }

struct Reader {                               // buffered input with look-ahead
    FILE* f; std::vector<uint8_t> buf; size_t pos = 0, end = 0; bool eof = false;
    explicit Reader(FILE* f_) : f(f_), buf(8u << 20) {}
    size_t avail() const { return end - pos; }
    const uint8_t* ptr() const { return buf.data() + pos; }
    void consume(size_t k) { pos += k; }
    bool fill(size_t want) {                  // make at least `want` bytes available if the file has them
        if (want > buf.size()) { buf.resize(want + (want >> 2)); }
        if (avail() >= want) return true;
        if (pos) { memmove(buf.data(), buf.data() + pos, end - pos); end -= pos; pos = 0; }
        while (!eof && avail() < want) {
            size_t k = fread(buf.data() + end, 1, buf.size() - end, f);
            if (k == 0) { if (ferror(f)) die("read error"); eof = true; }
            end += k;
        }
        return avail() >= want;
    }
    void refill() { if (avail() == 0) { pos = end = 0; } fill(1); }
};

uint32_t le32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

std::vector<uint8_t> g_dict; bool g_have_dict = false;

void do_compress(FILE* in, FILE* out, const LZ4F_preferences_t& prefs, bool literal_mode) {
    LZ4F_cctx* c = nullptr;
    check(LZ4F_createCompressionContext(&c, LZ4F_VERSION));
    const size_t CHUNK = 4u << 20;
    std::vector<uint8_t> ibuf(CHUNK), obuf(check(LZ4F_compressBound(CHUNK, &prefs)) + LZ4F_HEADER_SIZE_MAX);
    LZ4F_CDict* cd = g_have_dict ? LZ4F_createCDict(g_dict.data(), g_dict.size()) : nullptr;
    if (literal_mode) put(out, obuf.data(), check(LZ4F_compressBegin_usingCDict(c, obuf.data(), obuf.size(), cd, &prefs)));
    else
    put(out, obuf.data(), check(LZ4F_compressBegin(c, obuf.data(), obuf.size(), &prefs)));
    for (;;) {
        size_t n = fread(ibuf.data(), 1, CHUNK, in);
        if (n == 0) { if (ferror(in)) die("read error"); break; }
        put(out, obuf.data(), check(LZ4F_compressUpdate(c, obuf.data(), obuf.size(), ibuf.data(), n, nullptr)));
    }
    put(out, obuf.data(), check(LZ4F_compressEnd(c, obuf.data(), obuf.size(), nullptr)));
    check(LZ4F_freeCompressionContext(c));
    LZ4F_freeCDict(cd);
}

// one legacy frame (after its magic): [u32 size][LZ4 block]... until EOF or a word that cannot be a block size
void decode_legacy(Reader& rd, FILE* out, b2lz4_ctx*& gpu) {
    const size_t LB = 8u << 20, BOUND = LB + LB / 255 + 16;
    if (!gpu) { const char* e = getenv("B2LZ4_DEVICE"); check(b2lz4_create(&gpu, e ? atoi(e) : 0, nullptr)); }
    b2lz4_frame_info info; memset(&info, 0, sizeof info);
    info.block_size_id = B2LZ4_LEGACY_BLOCK_ID; info.block_independent = 1; info.block_size = LB;
    std::vector<uint8_t> batch, obuf;
    size_t nblocks = 0; bool first = true;
    auto flush = [&]() {
        if (!nblocks) return;
        obuf.resize(nblocks * LB);
        size_t r = check(b2lz4_decompress_blocks_host(gpu, &info, batch.data(), batch.size(), obuf.data(), obuf.size(),
                                                      first ? B2LZ4_STREAM_FIRST : 0u));
        put(out, obuf.data(), r);
        batch.clear(); nblocks = 0; first = false;
    };
    for (;;) {
        if (!rd.fill(4)) { if (rd.avail()) die("legacy frame: truncated block header"); break; }
        const uint32_t cs = le32(rd.ptr());
        if (cs > BOUND) break;                                   // the next frame's magic number (lz4io.c:913-918)
        if (!rd.fill(4 + (size_t)cs)) die("legacy frame: truncated block");
        batch.insert(batch.end(), rd.ptr(), rd.ptr() + 4 + cs);
        rd.consume(4 + (size_t)cs);
        if (++nblocks == 8) flush();
    }
    flush();
}

void do_decompress(FILE* in, FILE* out, bool all_frames) {
    LZ4F_dctx* d = nullptr;
    check(LZ4F_createDecompressionContext(&d, LZ4F_VERSION));
    b2lz4_ctx* gpu = nullptr;
    Reader rd(in);
    std::vector<uint8_t> obuf(64u << 20);
    size_t frames = 0;
    for (;;) {
        // ---- frame boundary: which kind of frame comes next?
        if (!rd.fill(4)) {
            if (frames == 0 || rd.avail()) die("decompress: not enough bytes for header");      // Conduit.hsc:615-616
            break;
        }
        const uint32_t magic = le32(rd.ptr());
        if (magic == 0x184C2102u) {
            if (!all_frames && frames == 0) die("lz4frame error: ERROR_frameType_unknown");     // the conduit does not read legacy frames
            rd.consume(4);
            decode_legacy(rd, out, gpu);
            ++frames;
            continue;
        }
        if (magic != 0x184D2204u && (magic & 0xFFFFFFF0u) != 0x184D2A50u) {
            if (frames == 0) die("lz4frame error: ERROR_frameType_unknown");
            fprintf(stderr, "b200lz4c: stream followed by undecodable data\n");                 // lz4io.c:1117-1122
            break;
        }
        const bool skippable = magic != 0x184D2204u;
        // ---- one frame through LZ4F_decompress until it returns 0 (a skippable frame is consumed inside it and ends with 0 too)
        size_t hint = 1;
        while (hint != 0) {
            if (rd.avail() == 0) { rd.refill(); if (rd.avail() == 0) {
                if (skippable) die("skippable frame: truncated");
                die("decompress: stream ended before EndMark"); } }                            // Conduit.hsc:689
            size_t ssz = rd.avail(), dsz = obuf.size();
            hint = check(g_have_dict ? LZ4F_decompress_usingDict(d, obuf.data(), &dsz, rd.ptr(), &ssz, g_dict.data(), g_dict.size(), nullptr)
                                     : LZ4F_decompress(d, obuf.data(), &dsz, rd.ptr(), &ssz, nullptr));
            put(out, obuf.data(), dsz);
            rd.consume(ssz);
            if (hint > obuf.size() && hint < (1ull << 31)) obuf.resize(hint);
        }
        if (!skippable) ++frames;
        if (!all_frames && frames) break;                                                       // the conduit stops after the first frame
    }
    LZ4F_freeDecompressionContext(d);
    b2lz4_destroy(gpu);
}

}  // namespace

int main(int argc, char** argv) {
    const char* in_path = nullptr; const char* out_path = nullptr;
    bool decompress = false, all_frames = false, literal_mode = false;
    LZ4F_preferences_t prefs; memset(&prefs, 0, sizeof prefs);
    int npos = 0;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-d" || a == "--decompress") decompress = true;
        else if (a == "--all-frames") all_frames = true;
        else if (a == "-BD") { prefs.frameInfo.blockMode = LZ4F_blockLinked; literal_mode = true; }
        else if (a == "-D") {
            if (++i >= argc) die("-D needs a file");
            FILE* f = fopen(argv[i], "rb"); if (!f) die(std::string("cannot open dictionary ") + argv[i]);
            uint8_t tmp[65536]; size_t k;
            while ((k = fread(tmp, 1, sizeof tmp, f)) > 0) g_dict.insert(g_dict.end(), tmp, tmp + k);
            fclose(f); g_have_dict = true; literal_mode = true;
        }
        else if (a.size() == 3 && a[0] == '-' && a[1] == 'B' && a[2] >= '4' && a[2] <= '7') prefs.frameInfo.blockSizeID = (LZ4F_blockSizeID_t)(a[2] - '0');
        else if (a == "--block-checksum" || a == "-BX") prefs.frameInfo.blockChecksumFlag = LZ4F_blockChecksumEnabled;
        else if (a == "--content-checksum") prefs.frameInfo.contentChecksumFlag = LZ4F_contentChecksumEnabled;
        else if (a == "-h" || a == "--help") {
            printf("Usage: b200lz4c [INPUT_FILE] [OUTPUT-FILE] [-d|--decompress] [-B4..-B7] [--block-checksum] [--content-checksum] [-BD] [-D DICT_FILE] [--all-frames]\n"
                   "  Compress or decompress .lz4 files\n");
            return 0;
        }
        else if (a == "-" || a[0] != '-') { if (npos == 0) in_path = argv[i]; else if (npos == 1) out_path = argv[i]; else die("too many arguments"); ++npos; }
        else die("unknown option " + a);
    }
    FILE* in = (!in_path || !strcmp(in_path, "-")) ? stdin : fopen(in_path, "rb");
    if (!in) die(std::string("cannot open ") + in_path);
    FILE* out = (!out_path || !strcmp(out_path, "-")) ? stdout : fopen(out_path, "wb");
    if (!out) die(std::string("cannot open ") + out_path);
    setvbuf(out, nullptr, _IOFBF, 4u << 20);
    if (decompress) do_decompress(in, out, all_frames); else do_compress(in, out, prefs, literal_mode);
    if (fflush(out) != 0) die("write error");
    if (out != stdout) fclose(out);
    if (in != stdin) fclose(in);
    return 0;
}
