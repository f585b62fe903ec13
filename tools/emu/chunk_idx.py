import os, sys, importlib.util, ctypes as C
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B2LZ4_LIB"]=os.path.join(ROOT,"tests/_build/libb200lz4_emu.so"); os.environ["B2_EMU_DEVICES"]="2"
pkgdir=os.path.join(ROOT,"lz4-frame-conduit_b200")
spec=importlib.util.spec_from_file_location("b2emu", os.path.join(pkgdir,"__init__.py"), submodule_search_locations=[pkgdir])
mod=importlib.util.module_from_spec(spec); sys.modules["b2emu"]=mod; spec.loader.exec_module(mod); mod.load()
gen=C.CDLL(os.path.join(ROOT,"tools/libb2gen.so")); gen.b2gen_loglike.argtypes=[C.c_void_p,C.c_size_t,C.c_uint64]; gen.b2gen_random.argtypes=[C.c_void_p,C.c_size_t,C.c_uint64]
def mk(fn,n,seed):
    b=C.create_string_buffer(n); fn(b,n,seed); return b.raw
d=mk(gen.b2gen_loglike,9<<20,5)+mk(gen.b2gen_random,(1<<20)+77,6)+bytes(3<<20)+mk(gen.b2gen_loglike,(2<<20)+12345,7)
frames=[]
for devs in ([0],[0,1]):
    with mod.Codec(0) as c:
        c.set_devices(devs)
        for prefs in (dict(block_size_id=6, sub_block_log2=17, write_index=1, block_checksum=1), dict(block_size_id=4, sub_block_log2=16, write_index=1)):
            c.set_prefs(**prefs)
            fr=c.compress_bytes(d); frames.append(fr)
            back,used=c.decompress_bytes(fr); assert back==d and used==len(fr)
            st=c.stats(); print(devs, prefs['block_size_id'], len(fr), "indexed", st.indexed, "chains", st.n_chains, "blocks", st.n_blocks)
assert frames[0]==frames[2] and frames[1]==frames[3]; print("identical across device counts")
