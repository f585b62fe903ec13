import os, sys, random, importlib.util, ctypes as C, subprocess
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT,'tests'))
os.environ["B2LZ4_LIB"]=os.path.join(ROOT,"tests/_build/libb200lz4_emu.so")
pkgdir=os.path.join(ROOT,"lz4-frame-conduit_b200")
spec=importlib.util.spec_from_file_location("b2emu", os.path.join(pkgdir,"__init__.py"), submodule_search_locations=[pkgdir])
mod=importlib.util.module_from_spec(spec); sys.modules["b2emu"]=mod; spec.loader.exec_module(mod); mod.load()
import conftest
sz=C.c_size_t
lib=C.CDLL(os.path.join(ROOT,"oracle/liboracle_lz4.so"))
lib.orc_frame_decompress.restype=sz; lib.orc_frame_decompress.argtypes=[C.c_char_p,sz,C.c_char_p,sz,C.POINTER(sz)]
gen=C.CDLL(os.path.join(ROOT,"tools/libb2gen.so"))
for f in (gen.b2gen_loglike, gen.b2gen_textlike): f.argtypes=[C.c_void_p,C.c_size_t,C.c_uint64]
def mk(fn,n,seed):
    b=C.create_string_buffer(n); fn(b,n,seed); return b.raw
rnd=random.Random(int(sys.argv[1]) if len(sys.argv)>1 else 1)
cli=os.path.join(ROOT,"oracle/_ref/lz4")
d=mk(gen.b2gen_loglike,700000,1)+bytes(5000)+mk(gen.b2gen_textlike,400000,2)+os.urandom(70000)
nbad=nok=0
with mod.Codec(0,use_index=0) as c:
    for flags in (["-B4"],["-B5","-BX"],["-B4","-BD"],["-B6","--no-frame-crc"]):
        good=subprocess.run([cli,"-c",*flags],input=d,capture_output=True,check=True).stdout
        for it in range(int(sys.argv[2]) if len(sys.argv)>2 else 25):
            b=bytearray(good)
            for _ in range(rnd.randint(1,3)): b[rnd.randrange(7,len(b))]=rnd.getrandbits(8)
            b=bytes(b)
            want,used=conftest.orc_decode(lib,b,len(d)+300000)
            try: got,used2=c.decompress_bytes(b,max_out=len(d)+300000)
            except mod.Lz4FrameError: got=None
            if isinstance(want,int): assert got is None,(flags,it,want); nbad+=1
            else: assert got==want and used2==used,(flags,it); nok+=1
        st=c.stats(); print(flags, "chains", st.n_chains, "blocks", st.n_blocks, flush=True)
print("ok bad",nbad,"accepted",nok)
