"""Instruction mix of the hot kernels from `cuobjdump -sass` of the built library (no GPU needed): opcode counts per
kernel, grouped, so that a change in the inner loops shows up as numbers.  Usage: python tools/sass_mix.py [lib.so]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "photon_ml_b200", "lib", "libphoton_ml_b200.so")
WANT = [("seg_spmv_tiled_kernel<float, false>", r"seg_spmv_tiled_kernelIfLb0E"),
        ("seg_spmv_warp_kernel<float,8,0,0,1,1024> (CSR stream)", r"seg_spmv_warp_kernelIfLi8ELb0ELb0ELb1ELi1024E"),
        ("seg_spmv_warp_kernel<float,8,0,0,0,256> (cold CSC)", r"seg_spmv_warp_kernelIfLi8ELb0ELb0ELb0ELi256E"),
        ("row_epilogue_kernel", r"row_epilogue_kernel"),
        ("peer_merge_kernel<8>", r"peer_merge_kernelILi8E")]
GROUPS = [("LDG/LD global", r"^LDG|^LD\."), ("STG/ST global", r"^STG|^ST\."), ("LDS", r"^LDS"), ("STS", r"^STS"),
          ("SHFL", r"^SHFL"), ("VOTE/MATCH", r"^VOTE|^MATCH"), ("DFMA/DADD/DMUL", r"^DFMA|^DADD|^DMUL"),
          ("F2F/I2F/F2I", r"^F2F|^I2F|^F2I|^I2FP"), ("integer/logic", r"^IADD|^IMAD|^LOP|^SHF|^LEA|^ISETP|^SEL|^PRMT|^POPC|^FLO|^BREV|^IABS|^VIADD"),
          ("FSEL/FSETP/DSETP", r"^FSEL|^FSETP|^DSETP"), ("BRA/BSSY/BSYNC/EXIT", r"^BRA|^BSSY|^BSYNC|^EXIT|^CALL|^RET|^WARPSYNC"),
          ("BAR/MEMBAR/ATOM", r"^BAR|^MEMBAR|^ATOM|^RED|^FENCE|^ERRBAR|^CCTL"), ("MOV/S2R/CS2R/R2UR", r"^MOV|^S2R|^CS2R|^R2UR|^S2UR|^UMOV|^LDC|^ULDC")]
text = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", text)
for label, pat in WANT:
    body = next((f for f in funcs if re.match(r"\S*" + pat, f)), None)
    if body is None:
        print("%s: not found" % label)
        continue
    ops = collections.Counter()
    for m in re.finditer(r"^\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]*)", body, re.M):
        ops[m.group(1)] += 1
    total = sum(ops.values())
    print("%s: %d SASS instructions" % (label, total))
    seen = 0
    for name, gp in GROUPS:
        n = sum(c for op, c in ops.items() if re.match(gp, op))
        seen += n
        print("    %-24s %5d" % (name, n))
    print("    %-24s %5d" % ("other", total - seen))
    tc = [op for op in ops if re.match(r"^UTC|^HMMA|^LDTM|^STTM|^UTMA|^UBLKCP", op)]
    print("    tensor / TMA opcodes: %s" % (", ".join(tc) if tc else "none (by design: sparse fp64 contraction, HBM-bound)"))
