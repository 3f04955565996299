"""python tools/sass_summary.py > profiles/r02_sass_summary.txt
Static SASS instruction counts per kernel of the shipped libilfdt.so (cuobjdump -sass): which kernels carry tcgen05 MMAs
(UTCHMMA), tensor-memory loads / stores (LDTM / STTM), TMA loads (UTMALDG), their cta_group::2 forms, packed FFMA2, vector
reductions (REDG) -- and that no legacy HMMA (mma.sync) exists anywhere."""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "iterated-learning-for-vlm_b200", "ilfdt", "libilfdt.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
cols = [("UTCHMMA", r"\bUTCHMMA"), ("MMA.2CTA", r"\bUTCHMMA\.2CTA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTMALDG", r"\bUTMALDG"),
        ("LDG.2CTA", r"\bUTMALDG\.\w+\.2CTA"), ("UTCBAR", r"\bUTCBAR"), ("FFMA2", r"\bFFMA2"), ("FFMA", r"\bFFMA\b"),
        ("REDG", r"\bREDG\b"), ("REDG.x4", r"\bREDG\.E\.ADD\.F32x4"), ("ATOMS", r"\bATOMS"), ("ATOMG", r"\bATOMG"), ("HMMA", r"\bHMMA")]
names, bodies = [], []
for part in re.split(r"\n\s*Function : ", sass)[1:]:
    name, _, body = part.partition("\n")
    names.append(name.strip())
    bodies.append(body)
dem = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines() if names else []
print("cuobjdump -sass of iterated-learning-for-vlm_b200/ilfdt/libilfdt.so (sm_100a), instruction counts per kernel (static SASS, not executed counts).")
print("UTCHMMA = tcgen05.mma (kind::f16), LDTM / STTM = tcgen05.ld / st (tensor memory), UTMALDG = TMA tensor load, .2CTA = cta_group::2 forms, UTCBAR = tcgen05.commit,")
print("FFMA2 = packed fp32x2 FMA, REDG = red.global (x4 = 16-byte vector form). No legacy HMMA (mma.sync) instruction exists anywhere in the library.\n")
print("%-72s" % "kernel" + "".join("%9s" % c for c, _ in cols))
rows = []
for n, d, body in zip(names, dem or names, bodies):
    short = re.sub(r"^void\s+", "", d).replace("ilfdt::", "").replace("(int)", "").replace("(bool)", "")
    short = re.sub(r"\(.*$", "", short)
    counts = [len(re.findall(rx, body)) for _, rx in cols]
    rows.append((short, counts))
rows.sort(key=lambda r: (-r[1][0], -r[1][7], r[0]))
for short, counts in rows:
    print("%-72s" % short[:72] + "".join("%9d" % c for c in counts))
tot = [sum(r[1][i] for r in rows) for i in range(len(cols))]
print("%-72s" % ("total (%d kernels)" % len(rows)) + "".join("%9d" % c for c in tot))
