"""Developer tool: static issue-time estimate of the softmax warps' full-tile pass in the pair kernel.

Walks the kernel's SASS (cuobjdump -sass) from the first score load of the optimistic pass to the "P ready" arrive,
taking every forward branch that skips a cold block (mask arithmetic of partial tiles, refresh sweeps, spin-wait
retries), and sums the instructions and the stall counts of their control codes: the time ONE warp needs for the
pass when nothing else competes for its scheduler (variable-latency waits excluded).  Validated against the ncu
source page of round 2 (539 instructions / 1131 clk for the build of commit d41d91c).

usage: python tools/sass_hot_path.py [libtmx.so] [kernel-name substring]
"""
import re
import subprocess
import sys


def load(lib, name):
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.split("\n")
    out, on = [], False
    for i, line in enumerate(txt):
        if "Function :" in line:
            on = name in line
            continue
        if not on:
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);\s*/\* (0x[0-9a-f]+) \*/", line)
        if m and i + 1 < len(txt):
            m2 = re.match(r"\s*/\* (0x[0-9a-f]+) \*/", txt[i + 1])
            if m2:
                hi = int(m2.group(1), 16)
                out.append((int(m.group(1), 16), m.group(2).strip(), (hi >> 41) & 0xF))
    return out


COLD = ("VIADDMNMX", "LDG", "VIMNMX", "LDTM", "BAR", "WARPSYNC.ALL", "NANOSLEEP")


def walk(ins, start, stop):
    idx = {a: k for k, (a, _, _) in enumerate(ins)}
    k = idx[start]
    n = cyc = mufu = 0
    trace = []
    while ins[k][0] != stop:
        a, s, st = ins[k]
        n += 1
        cyc += max(1, st)
        mufu += "MUFU" in s
        trace.append((a, s, st))
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", s)
        if m and "BRA" in s.split()[0 if not s.startswith("@") else 1]:
            tgt = int(m.group(1), 16)
            if tgt > a and tgt in idx:
                skipped = [x[1] for x in ins[k + 1:idx[tgt]]]
                uncond = not s.startswith("@") and "UP" not in s.split(",")[0]
                to_trap = any("BPT.TRAP" in x[1] for x in ins[idx[tgt]:idx[tgt] + 6])      # spin-wait time-out
                if not to_trap and (uncond or any(any(c in x for c in COLD) for x in skipped)):
                    k = idx[tgt]
                    continue
        k += 1
    return n, cyc, mufu, trace


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("-")]
    lib = args[0] if len(args) > 0 else "vllm_trianglemix_b200/libtmx.so"
    name = args[1] if len(args) > 1 else "pair_kernelILi4ELb1ELb0"
    ins = load(lib, name)
    ldtm = [a for a, s, _ in ins if s.startswith("LDTM")]
    sttm = [a for a, s, _ in ins if s.startswith("STTM")]
    # the pass (heuristics keyed to this kernel's layout): from the first score load after the refresh sweep's four
    # O-rescale stores to the last TMEM store of the kernel (the second half of P; the epilogue only loads)
    start = [a for a in ldtm if a > sttm[3]][0]
    stop = sttm[-1]
    n, cyc, mufu, trace = walk(ins, start, stop)
    print("pass %#x .. %#x: %d instructions, %d MUFU, static issue time %d clk" % (start, stop, n, mufu, cyc))
    if "-v" in sys.argv:
        for a, s, st in trace:
            print("%6x %2d %s" % (a, st, s[:80]))
    if "-c" in sys.argv:                                   # compact: opcode + stall count, 20 per line
        ops = []
        for a, s, st in trace:
            w = s.split()
            ops.append("%s%d" % ((w[1] if w[0].startswith("@") else w[0]).split(".")[0][:5], st))
        for k in range(0, len(ops), 20):
            print(" ".join(ops[k:k + 20]))


if __name__ == "__main__":
    main()
