"""Debug helper (GPU box): per-golden-frame decode status."""
import importlib, json, os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
pkg = importlib.import_module("lz4-frame-conduit_b200")
man = json.load(open(os.path.join(ROOT, "tests/golden/manifest.json")))
c = pkg.Codec(0)
for name, e in man.items():
    if name.startswith("_"): continue
    d = open(os.path.join(ROOT, "tests/golden", e["raw"]), "rb").read()[:e["size"]]
    for kind, fr in e["frames"].items():
        b = open(os.path.join(ROOT, "tests/golden", fr["file"]), "rb").read()
        for extra in (64, 1 << 20):
            try:
                out, consumed = c.decompress_bytes(b, max_out=len(d) + extra)
                st = c.stats()
                ok = "OK " if out == d and consumed == len(b) else f"MISMATCH len={len(out)} consumed={consumed}/{len(b)}"
            except Exception as ex:
                st = c.stats(); ok = f"ERR {ex}"
            print(f"{name:12s} {kind:14s} +{extra:<8d} blocks={st.n_blocks} chains={st.n_chains} retries={st.retries} {ok}")
