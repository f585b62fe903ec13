import os, sys, importlib.util, ctypes as C, subprocess, time
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT,'tests'))
os.environ["B2LZ4_LIB"]=os.path.join(ROOT,"tests/_build/libb200lz4_emu.so")
pkgdir=os.path.join(ROOT,"lz4-frame-conduit_b200")
spec=importlib.util.spec_from_file_location("b2emu", os.path.join(pkgdir,"__init__.py"), submodule_search_locations=[pkgdir])
mod=importlib.util.module_from_spec(spec); sys.modules["b2emu"]=mod; spec.loader.exec_module(mod); mod.load()
import conftest
gen=C.CDLL(os.path.join(ROOT,"tools/libb2gen.so")); 
n=int(sys.argv[1]) if len(sys.argv)>1 else 4
def loglike(n):
    b=C.create_string_buffer(n); gen.b2gen_loglike.argtypes=[C.c_void_p,C.c_size_t,C.c_uint64]; gen.b2gen_loglike(b,n,0x10C5); return b.raw
ref=conftest._bind_replay(C.CDLL(os.path.join(ROOT,"oracle/_ref/libref.so")))
data={"loglike":loglike(n<<20), "datagen":subprocess.run([os.path.join(ROOT,"oracle/_ref/datagen"),f"-g{n<<20}","-s1"],capture_output=True,check=True).stdout}
with mod.Codec(0) as c:
    c.set_prefs(block_size_id=7, sub_block_log2=int(os.environ.get("SUB_LOG2","17")), write_index=0)
    for name,d in data.items():
        cap=len(d)+len(d)//100+100000; o=C.create_string_buffer(cap)
        r=ref.ref_conduit_compress(d,len(d),o,cap,0,0)
        t0=time.time(); fr=c.compress_bytes(d); t=time.time()-t0
        back,_=c.decompress_bytes(fr); assert back==d
        print(f"{name}: ours {len(fr)} ref {r} ({100.0*(len(fr)/r-1):+.2f} %)  [{t:.0f}s]", flush=True)
