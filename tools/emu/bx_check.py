import os, sys, importlib.util, ctypes as C
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT,'tests'))
os.environ["B2LZ4_LIB"]=os.path.join(ROOT,"tests/_build/libb200lz4_emu.so")
pkgdir=os.path.join(ROOT,"lz4-frame-conduit_b200")
spec=importlib.util.spec_from_file_location("b2emu", os.path.join(pkgdir,"__init__.py"), submodule_search_locations=[pkgdir])
mod=importlib.util.module_from_spec(spec); sys.modules["b2emu"]=mod; spec.loader.exec_module(mod); mod.load()
import conftest, subprocess
gen=C.CDLL(os.path.join(ROOT,"tools/libb2gen.so")); gen.b2gen_loglike.argtypes=[C.c_void_p,C.c_size_t,C.c_uint64]
b=C.create_string_buffer(5<<20); gen.b2gen_loglike(b,5<<20,3); d=b.raw[:(5<<20)-777]+os.urandom(300001)
cli=os.path.join(ROOT,"oracle/_ref/lz4")
with mod.Codec(0) as c:
    for bid in (4,5,6,7):
        c.set_prefs(block_size_id=bid, sub_block_log2=16, block_checksum=1, content_checksum=1, write_index=0)
        fr=c.compress_bytes(d)
        r=subprocess.run([cli,"-d","-c"],input=fr,capture_output=True); assert r.returncode==0 and r.stdout==d,(bid,r.stderr[:200])
        back,_=c.decompress_bytes(fr); assert back==d
        # a frame the stock CLI made with block checksums, decoded (and verified) here; and a damaged checksum must be caught
        f2=subprocess.run([cli,"-c","-BX",f"-B{bid}"],input=d,capture_output=True,check=True).stdout
        back,_=c.decompress_bytes(f2); assert back==d
        bad=bytearray(f2); bad[len(bad)//2]^=0x40
        try:
            c.decompress_bytes(bytes(bad)); raise SystemExit("damage not caught")
        except mod.Lz4FrameError as e: pass
        print("bid",bid,"ok",len(fr))
