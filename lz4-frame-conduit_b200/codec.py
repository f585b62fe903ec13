"""Host-side mirror of the reference's conduit API over the C ABI.

    compress   :: ConduitT ByteString ByteString m ()      (Conduit.hsc:336-337)
    decompress :: ConduitT ByteString ByteString m ()      (Conduit.hsc:598-701)
    bsChunksOf :: Int -> ByteString -> [ByteString]        (Conduit.hsc:428-433)

A Haskell conduit is a stream transformer; the Python mirror is a generator over an iterable of
`bytes`.  `compress(chunks)` yields the bytes of ONE LZ4 frame; `decompress(chunks)` yields the decoded
bytes of the first frame of the stream and raises Lz4FrameError("stream ended before EndMark") on a
truncated stream, exactly where the reference throws (Conduit.hsc:616, 689).  All codec work happens in
libb200lz4.so on the GPU; nothing here compresses or decompresses on the CPU.
"""
import ctypes as C

from . import _lib
from ._lib import Prefs, FrameInfo, Stats


class Lz4FrameError(Exception):
    """Mirrors `throwString ("lz4frame error: " ++ name)` (Conduit.hsc:145-160)."""
    def __init__(self, name, code=None):
        super().__init__(f"lz4frame error: {name}")
        self.name, self.code = name, code


def _check(lib, r):
    if lib.b2lz4_is_error(r):
        raise Lz4FrameError(lib.b2lz4_error_name(r).decode(), r)
    return r


def bs_chunks_of(n, bs):
    """bsChunksOf (Conduit.hsc:428-433): split into chunks of n bytes, last one shorter."""
    if n < 1:
        raise ValueError(f"bsChunksOf: chunkSize < 1: {n}")
    if len(bs) <= n:                     # first equation of the reference: [bs], also for the empty ByteString
        return [bs]
    return [bs[i:i + n] for i in range(0, len(bs), n)]


def default_prefs(**kw):
    lib = _lib.load()
    p = Prefs()
    lib.b2lz4_default_prefs(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise TypeError(f"unknown preference {k!r}")
        setattr(p, k, int(v))
    return p


def frame_header_parse(buf):
    """LZ4F_getFrameInfo equivalent (host only).  Returns (FrameInfo, header_size)."""
    lib = _lib.load()
    info = FrameInfo()
    b = bytes(buf[:19])
    r = _check(lib, lib.b2lz4_frame_header_parse(b, len(b), C.byref(info)))
    return info, r


def frame_header_write(prefs=None, content_size=0):
    lib = _lib.load()
    out = C.create_string_buffer(19)
    r = _check(lib, lib.b2lz4_frame_header_write(out, 19, C.byref(prefs) if prefs is not None else None, content_size))
    return out.raw[:r]


def compress_bound(n, prefs=None):
    lib = _lib.load()
    return _check(lib, lib.b2lz4_compress_bound(n, C.byref(prefs) if prefs is not None else None))


class Codec:
    """One b2lz4 context: a CUDA device, its streams and workspaces."""

    def __init__(self, device=0, **prefs):
        self.lib = _lib.load()
        self.prefs = default_prefs(**prefs)
        h = C.c_void_p()
        _check(self.lib, self.lib.b2lz4_create(C.byref(h), device, C.byref(self.prefs)))
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.b2lz4_destroy(self.h)
            self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_prefs(self, **kw):
        for k, v in kw.items():
            setattr(self.prefs, k, int(v))
        _check(self.lib, self.lib.b2lz4_set_prefs(self.h, C.byref(self.prefs)))

    def set_devices(self, devices):
        """Host-buffer calls shard their blocks over these CUDA devices (b2lz4_set_devices); [] = own device only."""
        arr = (C.c_int * max(1, len(devices)))(*devices)
        _check(self.lib, self.lib.b2lz4_set_devices(self.h, arr, len(devices)))

    def _ck(self, r):
        if self.lib.b2lz4_is_error(r):
            name = self.lib.b2lz4_error_name(r).decode()
            if name == "ERROR_cudaFailure":
                name += ": " + self.lib.b2lz4_last_cuda_error(self.h).decode()
            raise Lz4FrameError(name, r)
        return r

    def stats(self):
        s = Stats()
        self.lib.b2lz4_last_stats(self.h, C.byref(s))
        return s

    # ---- device-resident (raw device pointers, e.g. torch tensor .data_ptr())
    def compress_device(self, d_src, n, d_dst, cap):
        return self._ck(self.lib.b2lz4_compress_device(self.h, d_src, n, d_dst, cap))

    def decompress_device(self, d_src, n, d_dst, cap):
        consumed = C.c_size_t(0)
        r = self._ck(self.lib.b2lz4_decompress_device(self.h, d_src, n, d_dst, cap, C.byref(consumed)))
        return r, consumed.value

    # ---- host buffers
    def compress_host_ptr(self, src, n, dst, cap):
        return self._ck(self.lib.b2lz4_compress_host(self.h, src, n, dst, cap))

    def decompress_host_ptr(self, src, n, dst, cap):
        consumed = C.c_size_t(0)
        r = self._ck(self.lib.b2lz4_decompress_host(self.h, src, n, dst, cap, C.byref(consumed)))
        return r, consumed.value

    def compress_bytes(self, data):
        """One frame from one bytes object."""
        data = bytes(data)
        cap = compress_bound(len(data), self.prefs)
        out = C.create_string_buffer(cap)
        r = self.compress_host_ptr(data, len(data), out, cap)
        return out.raw[:r]

    def decompress_bytes(self, frame, max_out=None):
        """Decode the first frame in `frame`.  Returns (decoded bytes, compressed bytes consumed)."""
        frame = bytes(frame)
        if len(frame) < 5:
            raise Lz4FrameError("decompress: not enough bytes for header")      # Conduit.hsc:615-616
        info, _ = frame_header_parse(frame)
        if max_out is None:                 # blocks in the frame (host walk over the header words) x block size
            max_out = _check(self.lib, self.lib.b2lz4_decompress_bound(frame, len(frame)))
            if info.has_content_size:
                max_out = min(max_out, info.content_size) if info.content_size else max_out
        out = C.create_string_buffer(max(1, max_out))
        r, consumed = self.decompress_host_ptr(frame, len(frame), out, max_out)
        return out.raw[:r], consumed

    def xxh32(self, data):
        data = bytes(data)
        h = C.c_uint32(0)
        self._ck(self.lib.b2lz4_xxh32_host(self.h, data, len(data), C.byref(h)))
        return h.value


# ------------------------------------------------------------------ conduit-shaped generators
_default = {}


def _codec(codec):
    if codec is not None:
        return codec
    if "c" not in _default:
        _default["c"] = Codec()
    return _default["c"]


def compress(chunks, codec=None):
    """compress (Conduit.hsc:336-337): yields the bytes of one LZ4 frame for the whole input stream.

    The reference re-cuts its input into 16 KB FFI calls; here the incoming ByteStrings are batched
    and handed to the GPU as whole frame blocks (north_star).  Output chunking differs from the
    reference (it is not part of the contract, Conduit.hsc:327-333); the concatenated bytes form one
    standard LZ4 frame that `lz4 -d` and the reference `decompress` accept.
    """
    c = _codec(codec)
    data = b"".join(chunks)
    yield c.compress_bytes(data)


def decompress(chunks, codec=None):
    """decompress (Conduit.hsc:598-701): yields the decoded bytes of the first frame in the stream."""
    c = _codec(codec)
    data = b"".join(chunks)
    try:
        out, _ = c.decompress_bytes(data)
    except Lz4FrameError as e:
        if e.name in ("ERROR_noEndMark", "ERROR_frameHeader_incomplete") and len(data) >= 5:
            raise Lz4FrameError("decompress: stream ended before EndMark") from e   # Conduit.hsc:689
        raise
    yield out
