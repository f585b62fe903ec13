"""Streaming mirror of Codec.Compression.LZ4.Conduit over the Tier-1 C ABI (include/lz4frame.h).

The reference's conduits are written against twelve LZ4F_* symbols (src/Codec/Compression/LZ4/Conduit.hsc:150-591).
This module is those conduits in Python, call for call, bound to the SAME symbols of libb200lz4.so — so it runs in
bounded memory, accepts any chunking of its input and raises the reference's exceptions:

    compress                    Conduit.hsc:336-337  (= compressWithOutBufferSize 0)
    compressYieldImmediately    Conduit.hsc:364-425
    compressWithOutBufferSize   Conduit.hsc:457-533
    decompress                  Conduit.hsc:598-701
    bsChunksOf                  Conduit.hsc:428-433

A Haskell `ConduitT ByteString ByteString m ()` becomes a generator over an iterable of `bytes`.  Nothing here compresses:
every byte of codec work happens behind LZ4F_compressUpdate / LZ4F_decompress on the GPU.
"""
import ctypes as C

from . import _lib
from .codec import Lz4FrameError, bs_chunks_of

_sz = C.c_size_t
LZ4F_VERSION = 100                      # lz4frame.h:236
LZ4F_HEADER_SIZE_MAX = 19               # Conduit.hsc:367, 466
_BS_CHUNK = 16 * 1024                   # bsChunksOf 16 KiB, Conduit.hsc:464, 501


class Preferences(C.Structure):
    """LZ4F_preferences_t, 56 bytes (CTypes.hsc:155-232; lz4frame.h:169-192)."""
    _fields_ = [("blockSizeID", C.c_int), ("blockMode", C.c_int), ("contentChecksumFlag", C.c_int), ("frameType", C.c_int),
                ("contentSize", C.c_ulonglong), ("dictID", C.c_uint), ("blockChecksumFlag", C.c_int),
                ("compressionLevel", C.c_int), ("autoFlush", C.c_uint), ("favorDecSpeed", C.c_uint), ("reserved", C.c_uint * 3)]


class FrameInfo(C.Structure):
    """LZ4F_frameInfo_t, 32 bytes."""
    _fields_ = [("blockSizeID", C.c_int), ("blockMode", C.c_int), ("contentChecksumFlag", C.c_int), ("frameType", C.c_int),
                ("contentSize", C.c_ulonglong), ("dictID", C.c_uint), ("blockChecksumFlag", C.c_int)]


_bound = None


def _f():
    """The twelve symbols with their C signatures (lz4frame.h of this repo)."""
    global _bound
    if _bound is None:
        lib = _lib.load()
        vp = C.c_void_p
        sig = {
            "LZ4F_isError": (C.c_uint, [_sz]), "LZ4F_getErrorName": (C.c_char_p, [_sz]),
            "LZ4F_createCompressionContext": (_sz, [C.POINTER(vp), C.c_uint]), "LZ4F_freeCompressionContext": (_sz, [vp]),
            "LZ4F_compressBegin": (_sz, [vp, vp, _sz, C.POINTER(Preferences)]),
            "LZ4F_compressBound": (_sz, [_sz, C.POINTER(Preferences)]),
            "LZ4F_compressUpdate": (_sz, [vp, vp, _sz, C.c_char_p, _sz, vp]),
            "LZ4F_compressEnd": (_sz, [vp, vp, _sz, vp]),
            "LZ4F_createDecompressionContext": (_sz, [C.POINTER(vp), C.c_uint]), "LZ4F_freeDecompressionContext": (_sz, [vp]),
            "LZ4F_getFrameInfo": (_sz, [vp, C.POINTER(FrameInfo), C.c_char_p, C.POINTER(_sz)]),
            "LZ4F_decompress": (_sz, [vp, vp, C.POINTER(_sz), C.c_char_p, C.POINTER(_sz), vp]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _bound = lib
    return _bound


def _handle(lib, r):
    """handleLz4Error (Conduit.hsc:145-160): error code -> exception "lz4frame error: NAME"."""
    if lib.LZ4F_isError(r):
        raise Lz4FrameError(lib.LZ4F_getErrorName(r).decode(), r)
    return r


def lz4_default_preferences():
    """lz4DefaultPreferences (Conduit.hsc:248-263): everything zero = library defaults."""
    return Preferences()


class _CCtx:
    """withScopedLz4fCompressionContext (Conduit.hsc:194-212)."""
    def __enter__(self):
        self.lib = _f()
        self.p = C.c_void_p()
        _handle(self.lib, self.lib.LZ4F_createCompressionContext(C.byref(self.p), LZ4F_VERSION))
        return self

    def __exit__(self, *a):
        self.lib.LZ4F_freeCompressionContext(self.p)


def compress_yield_immediately(chunks, prefs=None):
    """compressYieldImmediately (Conduit.hsc:364-425): header, then the result of every LZ4F_compressUpdate as it comes
    (possibly empty: the library buffers), then the footer."""
    prefs = prefs or lz4_default_preferences()
    with _CCtx() as ctx:
        lib = ctx.lib
        hdr = C.create_string_buffer(LZ4F_HEADER_SIZE_MAX)
        n = _handle(lib, lib.LZ4F_compressBegin(ctx.p, hdr, LZ4F_HEADER_SIZE_MAX, C.byref(prefs)))
        yield hdr.raw[:n]
        # the reference allocates LZ4F_compressBound(len) bytes per piece (allocaBytes, :385, :400); this library's bound covers
        # two whole batches, so ONE buffer of the largest bound is kept for the stream instead of a fresh one per 16 KB
        cap = _handle(lib, lib.LZ4F_compressBound(_BS_CHUNK, C.byref(prefs)))
        out = (C.c_char * cap)()
        base = C.addressof(out)
        for bs in chunks:
            for piece in bs_chunks_of(_BS_CHUNK, bytes(bs)):
                w = _handle(lib, lib.LZ4F_compressUpdate(ctx.p, base, cap, piece, len(piece), None))
                if w:
                    yield C.string_at(base, w)
        w = _handle(lib, lib.LZ4F_compressEnd(ctx.p, base, cap, None))
        yield C.string_at(base, w)


def compress_with_out_buffer_size(buffer_size, chunks, prefs=None):
    """compressWithOutBufferSize (Conduit.hsc:457-533): one reusable output buffer of
    max(bufferSize, LZ4F_compressBound(19 + 16 KiB)) bytes, flushed downstream whenever the next compressUpdate might
    not fit (`remainingCapacity < compressBound`, :506-512)."""
    prefs = prefs or lz4_default_preferences()
    with _CCtx() as ctx:
        lib = ctx.lib
        bound = _handle(lib, lib.LZ4F_compressBound(LZ4F_HEADER_SIZE_MAX + _BS_CHUNK, C.byref(prefs)))     # :466
        size = max(buffer_size, bound)
        buf = (C.c_char * size)()
        base = C.addressof(buf)
        used = _handle(lib, lib.LZ4F_compressBegin(ctx.p, base, size, C.byref(prefs)))
        for bs in chunks:
            for piece in bs_chunks_of(_BS_CHUNK, bytes(bs)):
                if size - used < bound:                                                                     # :506-512
                    yield C.string_at(base, used)
                    used = 0
                used += _handle(lib, lib.LZ4F_compressUpdate(ctx.p, base + used, size - used, piece, len(piece), None))
        if size - used < bound:
            yield C.string_at(base, used)
            used = 0
        used += _handle(lib, lib.LZ4F_compressEnd(ctx.p, base + used, size - used, None))                   # :518-525
        yield C.string_at(base, used)


def compress(chunks, prefs=None):
    """compress = compressWithOutBufferSize 0 (Conduit.hsc:336-337)."""
    return compress_with_out_buffer_size(0, chunks, prefs)


def decompress(chunks):
    """decompress (Conduit.hsc:598-701): reads the header (7, or 15 bytes when the content-size flag is set, :614-627),
    LZ4F_getFrameInfo, then LZ4F_decompress over every upstream chunk with a destination of the size the library hinted
    (:634-635, 665), until it returns 0; raises where the reference throws (:616, 689).  Bytes after the frame are left
    unread, like leftovers in the conduit."""
    lib = _f()
    d = C.c_void_p()
    _handle(lib, lib.LZ4F_createDecompressionContext(C.byref(d), LZ4F_VERSION))
    try:
        it = iter(chunks)
        buf = b""

        def need(k, what):
            nonlocal buf
            while len(buf) < k:
                nxt = next(it, None)
                if nxt is None:
                    raise Lz4FrameError(what)
                buf += bytes(nxt)

        need(5, "decompress: not enough bytes for header")                                                 # :615-616
        flg = buf[4]
        hs = 7 + (8 if flg & 0x08 else 0) + (4 if flg & 0x01 else 0)      # the reference forgets the dictID bytes (:621-623)
        need(hs, "decompress: not enough bytes for header")
        info = FrameInfo()
        used = _sz(hs)
        _handle(lib, lib.LZ4F_getFrameInfo(d, C.byref(info), buf[:hs], C.byref(used)))                     # :629-632
        buf = buf[used.value:]
        dst_cap = 16 * 1024                                                                                # :634-635
        hint = 1
        while hint != 0:
            if not buf:
                nxt = next(it, None)
                if nxt is None:
                    raise Lz4FrameError("decompress: stream ended before EndMark")                         # :689
                buf = bytes(nxt)
                if not buf:
                    continue
            out = C.create_string_buffer(dst_cap)
            dsz, ssz = _sz(dst_cap), _sz(len(buf))
            hint = _handle(lib, lib.LZ4F_decompress(d, out, C.byref(dsz), buf, C.byref(ssz), None))        # :591, 664-676
            if dsz.value:
                yield out.raw[:dsz.value]
            buf = buf[ssz.value:]
            if 0 < hint <= (1 << 31):
                dst_cap = max(16 * 1024, hint)                                                             # :665 uses the hint as the next size
    finally:
        lib.LZ4F_freeDecompressionContext(d)                                                               # :545, 701
