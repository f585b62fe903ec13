import os, sys, importlib.util, ctypes as C, subprocess, hashlib
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT,'tests'))
os.environ["B2LZ4_LIB"]=os.path.join(ROOT,"tests/_build/libb200lz4_emu.so")
pkgdir=os.path.join(ROOT,"lz4-frame-conduit_b200")
spec=importlib.util.spec_from_file_location("b2emu", os.path.join(pkgdir,"__init__.py"), submodule_search_locations=[pkgdir])
mod=importlib.util.module_from_spec(spec); sys.modules["b2emu"]=mod; spec.loader.exec_module(mod); mod.load()
gen=C.CDLL(os.path.join(ROOT,"tools/libb2gen.so"))
for f in (gen.b2gen_loglike, gen.b2gen_random, gen.b2gen_textlike): f.argtypes=[C.c_void_p,C.c_size_t,C.c_uint64]
def mk(fn,n,seed):
    b=C.create_string_buffer(n); fn(b,n,seed); return b.raw
data=[mk(gen.b2gen_loglike,3<<20,1), mk(gen.b2gen_textlike,2<<20,2), bytes(1<<20)+mk(gen.b2gen_random,300000,3)+mk(gen.b2gen_loglike,700001,4),
      subprocess.run([os.path.join(ROOT,"oracle/_ref/datagen"),"-g3000000","-s7"],capture_output=True,check=True).stdout, (b"abcdefghijklmnopqrstuvwxyz0123456789"*40000)[:1234567]]
h=hashlib.sha256()
with mod.Codec(0) as c:
    for prefs in (dict(block_size_id=7, sub_block_log2=17, write_index=1), dict(block_size_id=4, sub_block_log2=16, write_index=0), dict(block_size_id=5, sub_block_log2=18, write_index=0, block_linked=1)):
        c.set_prefs(**prefs)
        for d in data:
            fr=c.compress_bytes(d); h.update(fr)
            back,_=c.decompress_bytes(fr); assert back==d
            print(len(fr), end=" ")
print(); print(h.hexdigest())
