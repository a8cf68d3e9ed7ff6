"""Static SASS of the Goldilocks field operations the NTT kernels are built from (commit the excerpts the
instruction counts are quoted from). Takes the loop body of each k_field<OP> kernel of profiles/int_peak.cu (8 independent
chains per iteration, so per-operation counts = body / 8 after removing the 3 loop-control instructions and the 16 LOP3 of
the harness's operand mixing) and prints the opcode histogram per pipe plus the body itself.
usage: python profiles/sass_excerpts.py /tmp/int_peak > profiles/r2_sass_field_ops.txt"""
import collections
import re
import subprocess
import sys

ALU = {"IADD3", "LOP3", "SHF", "PRMT", "SEL", "ISETP", "LEA", "MOV", "VIADD", "PLOP3", "P2R", "R2P", "CS2R"}
FMA = {"IMAD", "HFMA2"}
exe = sys.argv[1]
out = subprocess.run(["cuobjdump", "-sass", exe], capture_output=True, text=True).stdout
for f in re.split(r"\n\s+Function : ", out)[1:]:
    name = f.split("\n", 1)[0]
    m = re.search(r"k_fieldI\d+(\w+?)EvPmm", name)
    if not m:
        continue
    ins = [(int(a, 16), t.strip()) for a, t in re.findall(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", f)]
    best = None
    for addr, txt in ins:
        b = re.search(r"BRA\S*\s+.*?(0x[0-9a-f]+)", txt)
        if b and int(b.group(1), 16) < addr and (best is None or addr - int(b.group(1), 16) > best[1] - best[0]):
            best = (int(b.group(1), 16), addr)
    body = [t for a, t in ins if best[0] <= a <= best[1]]
    c = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for t in body)
    alu = sum(v for k, v in c.items() if k in ALU)
    fma = sum(v for k, v in c.items() if k in FMA)
    wide = sum(1 for t in body if "IMAD.WIDE" in t or "IMAD.HI" in t)
    print(f"==== {m.group(1)}: loop body {len(body)} instructions for 8 operations: ALU pipe {alu} (incl. 16 harness LOP3), FMA pipe {fma} "
          f"(of which IMAD.WIDE/HI {wide}), other {len(body) - alu - fma}  -> per operation ~{(alu - 16) / 8:.1f} ALU + {fma / 8:.1f} FMA")
    for t in body:
        print("    " + t)
