"""Host-buffer calls sharded over several lanes / devices: the frame bytes must not depend on the device
list, every chunk boundary must reassemble in order, and the streamed content checksum must survive hopping between
devices.  On a one-GPU box the device list names the same GPU several times (same code path: one lane per list entry,
checksum state in pinned host memory); with >= 2 GPUs the real devices are used."""
import os

import pytest

from conftest import gen_bytes

pytestmark = pytest.mark.gpu


def _devices():
    import torch
    n = torch.cuda.device_count()
    return list(range(n)) if n >= 2 else [0, 0, 0]


def test_sharded_host_calls_match_single_device(pkg, gen, oracle):
    os.environ["B2LZ4_PIPE_MB"] = "8"                      # many chunks even for a modest input
    try:
        data = gen_bytes(gen, "loglike", (150 << 20) + 12345) + gen_bytes(gen, "random", 9 << 20, 7) + bytes(13 << 20)
        for kw in (dict(), dict(content_checksum=1, block_checksum=1), dict(write_index=1, content_checksum=1, block_size_id=6)):
            with pkg.Codec(0, **kw) as one, pkg.Codec(0, **kw) as many:
                many.set_devices(_devices())
                f1 = one.compress_bytes(data)
                f2 = many.compress_bytes(data)
                assert f1 == f2, kw                         # same bytes whatever the device list
                for c in (one, many):
                    back, used = c.decompress_bytes(f2, max_out=len(data))
                    assert back == data and used == len(f2), kw
                if kw.get("content_checksum"):               # a flipped bit must be caught by the checksum that hopped between the lanes
                    bad = bytearray(f2); bad[len(bad) // 2] ^= 1
                    with pytest.raises(pkg.Lz4FrameError):
                        many.decompress_bytes(bytes(bad), max_out=len(data))
                many.set_devices([])                        # back to the context's own device
                assert many.compress_bytes(data[:5 << 20]) == one.compress_bytes(data[:5 << 20])
    finally:
        del os.environ["B2LZ4_PIPE_MB"]


def test_lz4f_layer_over_several_devices(replay, gen, ref, monkeypatch):
    """The unmodified conduit's call sequence (oracle/conduit_replay.c on our LZ4F_* symbols) with B2LZ4_DEVICES set:
    batches of the stream shard over the lanes; output decodes with the reference library and with ourselves."""
    import ctypes as C
    from conftest import ref_conduit_decode
    monkeypatch.setenv("B2LZ4_DEVICES", ",".join(map(str, _devices())))
    monkeypatch.setenv("B2LZ4_PIPE_MB", "4")
    data = gen_bytes(gen, "textlike", (40 << 20) + 77)
    cap = len(data) + (len(data) >> 6) + (80 << 20)
    out = C.create_string_buffer(cap)
    r = replay.ref_conduit_compress(data, len(data), out, cap, 16384, 0)
    assert r < (1 << 62)
    frame = out.raw[:r]
    got, used = ref_conduit_decode(ref, frame, len(data) + 64)
    assert got == data and used == len(frame)
    got, used = ref_conduit_decode(replay, frame, len(data) + 64)
    assert got == data and used == len(frame)
