"""Static pipe-load estimate of a compiled kernel's loops (no GPU needed): which of the two half-rate integer pipes of an
sm_100a SM sub-partition bounds a kernel, from its SASS.

    python scripts/sass_pipes.py <lib.so | .cubin> <kernel-name substring> [more substrings ...]

For every backward branch (a loop) of the matching function it prints the instruction count of the loop body and the
cycles each pipe is busy per iteration, with the issue costs measured by scripts/int_peak.cu on a B200:
    FMA-heavy pipe : IMAD / IMAD.X / IMAD.MOV / IMAD.IADD / IMAD.SHL      2 cycles per warp instruction (18.4 T/s)
                     IMAD.WIDE / IMAD.HI                                  4 cycles                      (9.2 / 8.7 T/s)
    ALU pipe       : IADD3 / LOP3 / SHF / LEA / ISETP / SEL / MOV / ...   2 cycles
                     IMAD.WIDE / IMAD.HI                                  2 cycles MORE: scripts/exp/pipes.cu measures that
                     every wide multiply also holds the ALU pipe (or its dispatch / write-back port) for 2 cycles -- one
                     IMAD.WIDE + n independent IADD3 take 2 + 2n cycles, not max(4, 2n)
    issue          : 1 cycle per warp instruction of any kind
The larger of FMA and ALU' (ALU + the wide-multiply share) is the floor of an iteration; a kernel with the two close
runs at about 0.8 of it (two narrow instructions, one per pipe, co-issue at 2.4 cycles instead of 2.0).
"""
import re
import subprocess
import sys
from collections import Counter

FMA4 = ("IMAD.WIDE", "IMAD.HI")
ALU = ("IADD3", "LOP3", "SHF", "LEA", "ISETP", "SEL", "MOV", "VIADD", "PRMT", "IABS", "VIADDMNMX", "IMNMX", "PLOP3", "FSEL", "VIMNMX", "P2R", "R2P", "BMSK", "SGXT", "FLO", "POPC")
OTHER = Counter()


def classify(op):
    if op.startswith(FMA4):
        return "fma", 4
    if op.startswith("IMAD") or op.startswith("FFMA") or op.startswith("FMUL") or op.startswith("FADD"):
        return "fma", 2
    base = op.split(".")[0]
    if base in ALU:
        return "alu", 2
    return "other", 0


def functions(path):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    cur, body = None, []
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if cur:
                yield cur, body
            cur, body = m.group(1), []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur:
            body.append((int(m.group(1), 16), m.group(2).strip()))
    if cur:
        yield cur, body


def opcode(text):
    t = re.sub(r"^@!?U?P\w+\s+", "", text)
    return t.split()[0]


def tally(instrs):
    c = Counter()
    fma = alu = 0
    for _, t in instrs:
        op = opcode(t)
        pipe, cyc = classify(op)
        if pipe == "fma":
            fma += cyc
            if cyc == 4:
                alu += 2
        elif pipe == "alu":
            alu += cyc
        key = op
        for k in ("IMAD.WIDE", "IMAD.HI", "IMAD.X", "IMAD.MOV", "IMAD.IADD", "IMAD.SHL", "IADD3.X", "IADD3", "LEA.HI", "LEA", "LOP3", "SHF", "ISETP", "LDS", "STS", "LDG", "STG"):
            if op.startswith(k):
                key = k
                break
        else:
            if op.startswith("IMAD"):
                key = "IMAD"
        c[key] += 1
    return len(instrs), fma, alu, c


def loops_of(body):
    addr = {a: i for i, (a, _) in enumerate(body)}
    loops = []
    for i, (a, t) in enumerate(body):
        m = re.search(r"BRA(?:\.\w+)*\s+(?:!?U?P\w+,\s*)?0x([0-9a-f]+)", t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < a and tgt in addr and i - addr[tgt] > 40:
                loops.append((addr[tgt], i))
    return loops


def hot_loop(path, pats, min_wide=100):
    """The smallest loop of the (first) matching kernel that holds at least `min_wide` wide multiplies:
    {"kernel", "instr", "fma_cycles", "alu_cycles", "wide"} per iteration, or None."""
    for name, body in functions(path):
        if not all(p in name for p in pats):
            continue
        best = None
        for lo, hi in loops_of(body):
            n, fma, alu, c = tally(body[lo:hi + 1])
            wide = c["IMAD.WIDE"] + c["IMAD.HI"]
            if wide >= min_wide and (best is None or n < best["instr"]):
                best = {"kernel": name, "instr": n, "fma_cycles": fma, "alu_cycles": alu, "wide": wide}
        return best
    return None


def main():
    path, pats = sys.argv[1], sys.argv[2:]
    for name, body in functions(path):
        if not all(p in name for p in pats):
            continue
        print(name)
        n, fma, alu, c = tally(body)
        print(f"  whole function: {n} instr, FMA {fma} cyc, ALU' {alu} cyc")
        for lo, hi in loops_of(body):
            n, fma, alu, c = tally(body[lo:hi + 1])
            top = ", ".join(f"{k} {v}" for k, v in c.most_common(14))
            print(f"  loop {body[lo][0]:#x}..{body[hi][0]:#x}: {n} instr (issue {n} cyc), FMA {fma} cyc, ALU' {alu} cyc\n     {top}")


if __name__ == "__main__":
    main()
