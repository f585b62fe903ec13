import os, sys, random, ctypes as C
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B2LZ4_BATCH_MB"]="4"
lib=C.CDLL(os.path.join(ROOT,"tests/_build/libb200lz4_emu.so"))
sz=C.c_size_t
lib.LZ4F_createDecompressionContext.argtypes=[C.POINTER(C.c_void_p),C.c_uint]; lib.LZ4F_createDecompressionContext.restype=sz
lib.LZ4F_freeDecompressionContext.argtypes=[C.c_void_p]; lib.LZ4F_freeDecompressionContext.restype=sz
lib.LZ4F_decompress.argtypes=[C.c_void_p,C.c_void_p,C.POINTER(sz),C.c_void_p,C.POINTER(sz),C.c_void_p]; lib.LZ4F_decompress.restype=sz
lib.LZ4F_isError.argtypes=[sz]; lib.LZ4F_getErrorName.argtypes=[sz]; lib.LZ4F_getErrorName.restype=C.c_char_p
import subprocess
gen=C.CDLL(os.path.join(ROOT,"tools/libb2gen.so")); gen.b2gen_loglike.argtypes=[C.c_void_p,C.c_size_t,C.c_uint64]
def mk(n,seed):
    b=C.create_string_buffer(n); gen.b2gen_loglike(b,n,seed); return b.raw
cli=os.path.join(ROOT,"oracle/_ref/lz4")
rnd=random.Random(int(sys.argv[1]) if len(sys.argv)>1 else 1)
d=mk(21<<20,9)+os.urandom(1<<20)+bytes(3<<20)+mk(777777,10)
frames={}
for flags in (["-B4"],["-B5","-BX"],["-B6"],["-B7","--no-frame-crc"],["-B4","-BD"]):
    frames[" ".join(flags)]=subprocess.run([cli,"-c",*flags],input=d,capture_output=True,check=True).stdout
def decode(fr, srcmax, dstmax):
    ctx=C.c_void_p(); assert lib.LZ4F_createDecompressionContext(C.byref(ctx),100)==0
    out=bytearray(); pos=0; dstbuf=C.create_string_buffer(dstmax); srcb=C.create_string_buffer(fr,len(fr))
    hint=1; guard=0
    while True:
        take=min(len(fr)-pos, rnd.randint(1,srcmax)); cap=rnd.randint(0 if take else 1,dstmax)
        s=sz(take); o=sz(cap)
        hint=lib.LZ4F_decompress(ctx, dstbuf, C.byref(o), C.addressof(srcb)+pos, C.byref(s), None)
        assert not lib.LZ4F_isError(hint), lib.LZ4F_getErrorName(hint)
        out+=dstbuf.raw[:o.value]; pos+=s.value
        guard+=1; assert guard<10_000_000
        if hint==0: break
        assert not (pos==len(fr) and o.value==0 and s.value==0 and take==0 and cap>0) or guard<5_000_000
    lib.LZ4F_freeDecompressionContext(ctx)
    return bytes(out), pos
for name,fr in frames.items():
    for srcmax,dstmax in ((65536,65536),(1<<20,16384),(4096,1<<20),(3<<20,5<<20),(100000,100)):
        if dstmax==100 and len(d)>(4<<20): 
            fr2=subprocess.run([cli,"-c",*name.split()],input=d[:3_000_000],capture_output=True,check=True).stdout; want=d[:3_000_000]
        else: fr2=fr; want=d
        got,pos=decode(fr2,srcmax,dstmax)
        assert got==want and pos==len(fr2),(name,srcmax,dstmax,len(got),pos,len(fr2))
    print(name,"ok",flush=True)
