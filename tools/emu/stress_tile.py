import os, sys, random, importlib.util, ctypes as C
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT,'tests'))
os.environ["B2LZ4_LIB"]=os.path.join(ROOT,"tests/_build/libb200lz4_emu.so")
pkgdir=os.path.join(ROOT,"lz4-frame-conduit_b200")
spec=importlib.util.spec_from_file_location("b2emu", os.path.join(pkgdir,"__init__.py"), submodule_search_locations=[pkgdir])
mod=importlib.util.module_from_spec(spec); sys.modules["b2emu"]=mod; spec.loader.exec_module(mod); mod.load()
import conftest, json
rnd=random.Random(int(sys.argv[1]) if len(sys.argv)>1 else 1)
def periodic(n):
    out=bytearray()
    while len(out)<n:
        k=rnd.randrange(6)
        if k==0: out+=bytes(rnd.getrandbits(8) for _ in range(rnd.randrange(1,40)))
        elif k==1:
            p=rnd.randrange(1,40); pat=bytes(rnd.getrandbits(8) for _ in range(p)); L=rnd.choice([5,9,17,30,70,300,700,3000])
            out+=(pat*(L//p+1))[:L]
        elif k==2 and len(out)>10:
            off=rnd.randrange(1,min(len(out),70000)); L=rnd.randrange(4,60); st=len(out)-off
            for i in range(L): out.append(out[st+i])
        elif k==3 and len(out)>5000:
            off=rnd.randrange(2000,min(len(out),9000)); L=rnd.randrange(4,200); st=len(out)-off
            for i in range(L): out.append(out[st+i])
        elif k==4: out+=bytes(rnd.randrange(2000))
        else:
            w=[b"GET ",b"POST ",b"/index.html ",b"200 ",b"404 ",b"user=",b"alice ",b"bob ",b"\n",b"2026-11-05 "]
            out+=b"".join(rnd.choice(w) for _ in range(rnd.randrange(3,30)))
    return bytes(out[:n])
n_ok=0
with mod.Codec(0) as c, mod.Codec(0,use_index=0) as c0:
    for it in range(int(sys.argv[2]) if len(sys.argv)>2 else 12):
        d=periodic(rnd.choice([1000,70000,200000,400000]))
        for prefs in (dict(block_size_id=4,sub_block_log2=16,write_index=1), dict(block_size_id=5,sub_block_log2=17,write_index=1),
                      dict(block_size_id=4,sub_block_log2=16,write_index=0), dict(block_size_id=6,sub_block_log2=18,write_index=0,block_linked=1)):
            c.set_prefs(**prefs)
            fr=c.compress_bytes(d)
            for cc in (c,c0):
                back,used=cc.decompress_bytes(fr)
                assert back==d and (used==len(fr) or cc is c0),(it,prefs,len(d))
                n_ok+=1
    man=json.load(open(os.path.join(ROOT,"tests/golden/manifest.json")))
    for name,e in man.items():
        if name.startswith("_"): continue
        raw=conftest.golden_bytes(e["raw"])[:e["size"]]
        for kind,fr in e["frames"].items():
            f=conftest.golden_bytes(fr["file"])
            out,used=c0.decompress_bytes(f)
            assert out==raw and used==len(f),(name,kind)
            n_ok+=1
print("ok",n_ok,"DF",os.environ.get("B2LZ4_DF"))
