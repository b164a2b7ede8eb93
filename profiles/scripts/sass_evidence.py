"""cuobjdump evidence for libk8sgpu.so: arch list, the mnemonics that matter (TMA bulk copy, mbarrier, votes, atomics),
and the instruction mix of every loop body of the two shipped scoring kernels.
    python profiles/scripts/sass_evidence.py > profiles/r02_sass_evidence.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, "k8s-gpu-_b200", "libk8sgpu.so")
run = lambda *a: subprocess.run(a, capture_output=True, text=True).stdout
print("# SASS evidence for libk8sgpu.so (cuobjdump -sass, sm_100a cubins only), round 2, final build")
print("# arch list:")
for l in run("cuobjdump", "-lelf", LIB).splitlines():
    print("#  ", l.strip())
sass = run("cuobjdump", "-sass", LIB)
ops = collections.Counter()
for l in sass.splitlines():
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", l)
    if m:
        ops[m.group(1)] += 1
print("\n# mnemonic counts over the whole library (TMA bulk copy = UBLKCP, mbarrier = SYNCS.*; no UTCMMA / HMMA / UTMALDG: nothing here is a contraction)")
for k, v in sorted(ops.items(), key=lambda kv: -kv[1]):
    if re.match(r"UBLKCP|SYNCS|MATCH|REDUX|ATOMG|ATOMS|RED\.|BAR\.RED|UTC|HMMA|UTMA|VOTE", k):
        print(f"{v:7d} {k}")
print(f"{sum(ops.values()):7d} instructions in all")
names = [n for n in re.findall(r"Function : (\S+)", sass) if "score_kernelILi2ELi4ELi4" in n]
for name in names:
    body = run("cuobjdump", "-sass", "-fun", name, LIB)
    ins = []
    for l in body.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2)))
    addr = {a: i for i, (a, _) in enumerate(ins)}
    unr = "8" if "ELi8E" in name else "16"
    print(f"\n# score_kernel<PPT=2, RES=4, MB=4, UNR={unr}>: {len(ins)} instructions; every loop of 60..1000 instructions (one trip = UNR nodes x 2 pods;"
          " the R loops always 16 nodes):")
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA\S*\s+.*?0x([0-9a-f]+)", t)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt < a and tgt in addr and 60 < i - addr[tgt] < 1000:
            blk = ins[addr[tgt]:i + 1]
            c = collections.Counter((x.split()[1] if x.startswith("@") else x.split()[0]).split(".")[0] for _, x in blk)
            lds = c.get("LDS", 0)
            print(f"  {len(blk):4d} instructions  {dict(c.most_common(7))}")
