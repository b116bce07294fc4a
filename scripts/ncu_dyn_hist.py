#!/usr/bin/env python
"""Dynamic SASS histogram of one kernel from an `ncu --page source --csv` export (per-instruction "Instructions Executed"):
warp-instructions per sample by opcode and by issue pipe, plus shared-memory wavefronts and stall samples per opcode class.
  python scripts/ncu_dyn_hist.py file_source.csv <samples per launch> [top]"""
import collections
import csv
import re
import sys

ALU = ("PRMT", "SHF", "LOP3", "ISETP", "VIMNMX", "VIADD", "VIADDMNMX", "IADD3", "LEA", "SEL", "IABS", "PLOP3", "SGXT", "IMNMX", "FLO", "POPC", "BREV", "VABSDIFF", "I2I", "MOV", "VOTE", "FSETP", "FMNMX", "FSEL")
FMA = ("IMAD", "FFMA", "FADD", "FMUL", "FFMA2", "FADD2", "FMUL2", "HFMA2")
LSU = ("LDS", "STS", "LDG", "STG", "LD", "ST", "LDL", "STL", "ATOMS", "ATOMG", "RED", "LDGSTS", "SHFL", "LDSM", "ATOM")
OTHER_PIPES = {"BAR": "adu/cbu", "BRA": "cbu", "BSYNC": "cbu", "BSSY": "cbu", "EXIT": "cbu", "WARPSYNC": "cbu", "S2R": "adu", "LDC": "adu/lsu", "LDCU": "uniform", "R2UR": "uniform",
               "MUFU": "xu", "I2F": "xu", "F2I": "xu", "F2F": "xu", "I2FP": "alu", "F2FP": "alu", "CS2R": "adu", "NOP": "none", "DEPBAR": "none", "ERRBAR": "none", "MEMBAR": "lsu"}


def pipe_of(op):
    base = op.split(".")[0]
    if base.startswith("U") and base not in ("UNKNOWN",):
        return "uniform"
    if base in FMA:
        return "fma"
    if base in ALU:
        return "alu"
    if base in LSU:
        return "lsu"
    return OTHER_PIPES.get(base, "other")


def main():
    path, samples = sys.argv[1], float(sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    rows = list(csv.reader(open(path)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {h: i for i, h in enumerate(hdr)}
    ops, pipes, wav, wav_ideal, stall = collections.Counter(), collections.Counter(), collections.Counter(), collections.Counter(), collections.Counter()
    total = 0
    for r in rows[hdr_i + 1:]:
        if len(r) < len(hdr):
            continue
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[col["Source"]])
        if not m:
            continue
        op = ".".join(m.group(2).split(".")[:2])
        n = int(r[col["Instructions Executed"]] or 0)
        ops[op] += n; pipes[pipe_of(op)] += n; total += n
        wav[op] += int(r[col["L1 Wavefronts Shared"]] or 0)
        wav_ideal[op] += int(r[col["L1 Wavefronts Shared Ideal"]] or 0)
        stall[op] += int(r[col["# Samples"]] or 0)
    print(f"{rows[0][1]}\nwarp-instructions executed: {total}  = {total / samples:.3f} per sample = {32 * total / samples:.1f} thread-instructions per sample")
    print("by pipe (warp-instr per sample, share):")
    for p, n in pipes.most_common():
        print(f"  {p:10s}{n / samples:8.3f}  {100.0 * n / total:5.1f} %")
    print("by opcode (warp-instr per sample, share, stall samples share, shared wavefronts / ideal):")
    st = sum(stall.values()) or 1
    for op, n in ops.most_common(top):
        w = f"  smem wavefronts {wav[op] / samples:.3f}/sample (ideal {wav_ideal[op] / samples:.3f})" if wav[op] else ""
        print(f"  {op:22s}{n / samples:8.3f}  {100.0 * n / total:5.1f} %   stalls {100.0 * stall[op] / st:5.1f} %{w}")


if __name__ == "__main__":
    main()
