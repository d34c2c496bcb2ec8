"""Per-kernel counts of the Blackwell mnemonics in the built library, with one example line each:
python scripts/sass_listing.py > profiles/rNN_sass.txt   (cuobjdump -sass dino_b200/libdino_b200.so | c++filt)"""
import collections, re, subprocess
MN = ("UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UTMAPF", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "SYNCS", "REDG", "UTCCP")
raw = subprocess.run("cuobjdump -sass dino_b200/libdino_b200.so | c++filt", shell=True, capture_output=True, text=True).stdout
print("# SASS evidence (cuobjdump -sass dino_b200/libdino_b200.so, sm_100a): tcgen05 / TMEM / TMA mnemonics per kernel\n")
name, cnt, ex = None, None, None
def flush():
    if name and any(k in cnt for k in ("UTCHMMA", "UTMALDG", "LDTM")):
        print(name)
        print("    " + ", ".join(f"{k} x{cnt[k]}" for k in sorted(cnt)))
        for k in ("UTCHMMA", "LDTM", "STTM", "UTMALDG", "REDG"):
            if k in ex:
                print("      e.g. " + ex[k])
        print()
for ln in raw.split("\n"):
    m = re.match(r"\s*Function : (.*)", ln)
    if m:
        flush()
        name, cnt, ex = m.group(1), collections.Counter(), {}
        continue
    if name is None:
        continue
    for k in MN:
        if re.search(r"\b" + k + r"\b|\b" + k + r"\.", ln):
            cnt[k] += 1
            ex.setdefault(k, ln.strip())
flush()
