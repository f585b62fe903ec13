"""b2lz4_xxh32_ranges_static (csrc/noalloc.cu, include/b200lz4.h): the entry point that allocates nothing — block checksums
of many ranges of a device buffer through a file-scope __device__ scratch array.  It has no counterpart in the reference,
so the checker is the EXISTING entry point b2lz4_xxh32_device on the same
ranges (and the oracle's XXH32): integer work, bit-exact.  Two calls in a row, because every call shares the array and
finds the previous call's contents in it.  The file sorts last on purpose."""
import ctypes as C
import random

import pytest

pytestmark = pytest.mark.gpu

MAX_RANGES = 65536          # B2LZ4_STATIC_MAX_RANGES


def _ranges(rnd, n_src, n):
    offs, lens = [], []
    for _ in range(n):
        kind = rnd.randrange(5)
        ln = (0, rnd.randrange(1, 16), rnd.randrange(16, 300), rnd.randrange(300, 70000), 65536)[kind]
        ln = min(ln, n_src)
        offs.append(rnd.randrange(0, n_src - ln + 1))
        lens.append(ln)
    return offs, lens


def _static(lib, dev_ptr, n_src, offs, lens, device=0):
    n = len(offs)
    a = (C.c_uint64 * max(n, 1))(*offs)
    b = (C.c_uint32 * max(n, 1))(*lens)
    out = (C.c_uint32 * max(n, 1))(*([0xDEADBEEF] * max(n, 1)))
    r = lib.b2lz4_xxh32_ranges_static(device, None, C.c_void_p(dev_ptr), n_src, a, b, n, out)
    return r, list(out)[:n]


def test_static_scratch_checksums_match_existing_entry_point(pkg, oracle, gen):
    import torch
    from conftest import gen_bytes
    lib = pkg.load()
    data = gen_bytes(gen, "loglike", 3 << 20) + gen_bytes(gen, "random", 1 << 20, 7) + bytes(100000)
    n_src = len(data)
    dev = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    torch.cuda.synchronize()
    base = dev.data_ptr()
    rnd = random.Random(0xB200)
    with pkg.Codec(0) as c:
        for call, n in enumerate((1500, 700)):                      # the second call is smaller: stale entries stay behind it
            offs, lens = _ranges(rnd, n_src, n)
            if call == 0:
                offs[:3], lens[:3] = [0, n_src, 1], [n_src, 0, 3]    # the whole buffer, an empty range at the end, misaligned
            r, got = _static(lib, base, n_src, offs, lens)
            assert not lib.b2lz4_is_error(r), (lib.b2lz4_error_name(r), lib.b2lz4_static_last_error())
            assert lib.b2lz4_static_last_error() == b""
            h = C.c_uint32(0)
            for i in range(n):
                want = oracle.orc_xxh32(data[offs[i]:offs[i] + lens[i]], lens[i], 0)
                assert got[i] == want, (call, i, offs[i], lens[i])
                if i < 200:                                          # the existing device entry point, range by range
                    rr = lib.b2lz4_xxh32_device(c.h, C.c_void_p(base + offs[i]), lens[i], C.byref(h))
                    assert not lib.b2lz4_is_error(rr)
                    assert got[i] == h.value, (call, i)
        # the limit: one range too many is refused with a clear error and nothing is launched
        big = MAX_RANGES + 1
        r, _ = _static(lib, base, n_src, [0] * big, [1] * big)
        assert lib.b2lz4_is_error(r) and lib.b2lz4_error_name(r) == b"ERROR_srcSize_tooLarge"
        assert b"B2LZ4_STATIC_MAX_RANGES" in lib.b2lz4_static_last_error()
        # exactly the limit works, and a range that leaves the buffer is refused
        offs, lens = _ranges(rnd, n_src, MAX_RANGES)
        lens = [min(x, 64) for x in lens]
        r, got = _static(lib, base, n_src, offs, lens)
        assert not lib.b2lz4_is_error(r), lib.b2lz4_static_last_error()
        for i in range(0, MAX_RANGES, 257):
            assert got[i] == oracle.orc_xxh32(data[offs[i]:offs[i] + lens[i]], lens[i], 0), i
        r, _ = _static(lib, base, n_src, [n_src - 3], [4])
        assert lib.b2lz4_is_error(r) and b"leaves the buffer" in lib.b2lz4_static_last_error()
        # and the codec's own paths are untouched by the shared array
        fr = c.compress_bytes(data[:300000])
        back, used = c.decompress_bytes(fr)
        assert back == data[:300000] and used == len(fr)
    del dev
