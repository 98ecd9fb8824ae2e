"""Scratch diagnostic: which block coordinates differ between ours / reference / oracle."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H

def dec(k):
    def s16(v):
        v &= 0xFFFF
        return v - 0x10000 if v >= 0x8000 else v
    return (s16(int(k)), s16(int(k) >> 16), s16(int(k) >> 32))

def h(b, K):
    return (((b[0] * 73856093) & 0xFFFFFFFF) ^ ((b[1] * 19349669) & 0xFFFFFFFF) ^ ((b[2] * 83492791) & 0xFFFFFFFF)) % K

name = sys.argv[1] if len(sys.argv) > 1 else "small"
scn = H.make_scenario(name, (0,), "depth", 1.2)
ours = H.run_ours(scn)
orc = H.run_oracle(scn)
print("ours", len(ours["keys"]), "oracle", len(orc["keys"]))
for trial in range(4):
    for nofmad in (True, False):
        ref = H.run_ref(scn, nofmad=nofmad)
        a, b = set(ours["keys"].tolist()), set(ref["keys"].tolist())
        extra_ref = sorted(b - a); extra_ours = sorted(a - b)
        print("trial", trial, "nofmad", nofmad, "ref", len(ref["keys"]), "only_ref", [dec(k) for k in extra_ref][:10],
              "only_ours", [dec(k) for k in extra_ours][:10])
        allb = {dec(k) for k in b}
        for k in extra_ref[:10]:
            c = dec(k); hb = h(c, scn["M"])
            same = [d for d in allb if h(d, scn["M"]) == hb]
            print("   bucket", hb, "blocks in bucket:", same)
