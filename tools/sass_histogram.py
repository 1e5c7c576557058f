"""Opcode histogram of selected kernels of libhalo_b200.so from `cuobjdump -sass` (no GPU needed).
usage: tools/sass_histogram.py [regex ...]     default: the multiplier / squaring probes and the MSM hot loop
Prints, per kernel: instruction count, the wide multiplies (IMAD.WIDE*), other IMAD forms, the predicate-spill
family (P2R / R2P / SEL / LOP3 on predicates) and the ten most frequent opcodes."""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "halo_b200", "libhalo_b200.so")
pats = sys.argv[1:] or [r"sass_probe_mul", r"sass_probe_sqr", r"msm_accumulate", r"bench_modmul_kernel", r"ntt_last_pass|ntt_pass"]
txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)[1:]
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    if not any(re.search(p, name) for p in pats):
        continue
    dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
    ops = collections.Counter()
    for line in f.splitlines():
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            ops[m.group(1)] += 1
    tot = sum(ops.values())
    wide = sum(v for k, v in ops.items() if k.startswith("IMAD.WIDE"))
    imad_other = sum(v for k, v in ops.items() if k.startswith("IMAD") and not k.startswith("IMAD.WIDE"))
    spill = sum(v for k, v in ops.items() if k.split(".")[0] in ("P2R", "R2P", "SEL", "PLOP3"))
    print(f"== {dem[:150]}")
    print(f"   instructions {tot}   IMAD.WIDE* {wide}   other IMAD {imad_other}   P2R/R2P/SEL/PLOP3 {spill}   "
          f"IADD3* {sum(v for k, v in ops.items() if k.startswith('IADD3'))}   LOP3 {ops.get('LOP3.LUT', 0)}   "
          f"ISETP* {sum(v for k, v in ops.items() if k.startswith('ISETP'))}   SHF* {sum(v for k, v in ops.items() if k.startswith('SHF'))}")
    print("   top: " + ", ".join(f"{k} {v}" for k, v in ops.most_common(10)))
