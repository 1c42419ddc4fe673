"""Instruction mix of the frame kernel from an `ncu --set full --import-source on` capture (needs -lineinfo).

    python tools/sass_hist.py gpurun_out/x.ncu-rep <warp_steps> [top_lines]

Reads the source page (CUDA lines with their SASS), attributes every executed warp instruction to a stage of the
frame (by source file / line range of the function it was inlined from) and prints, per stage, warp instructions per
32-particle step (= thread instructions per particle), the opcode histogram of the stage and the pipe mix -- the
evidence VERDICT r01 asked for ("where do 41.5 warp-instructions per particle go").
warp_steps = particles * trackers / 32 of the captured launch (x score passes, 1 in the bench scenes)."""
import collections
import csv
import re
import subprocess
import sys


def load(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    recs, cur, hdr, line, src = [], None, None, None, None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if r[0] in ("Function Name", "Kernel Name"):
            continue
        if r[0] != "":
            line, src = int(r[0]), r[1]
            continue
        if len(r) > 7 and r[2].startswith("0x"):
            d = dict(zip(hdr[2:], r[2:]))
            try:
                n = float(d.get("Instructions Executed", 0))
                s = float(r[4])
            except ValueError:
                continue
            recs.append(dict(file=cur, line=line, src=src, addr=int(r[2], 16), sass=r[3].strip(), n=n, smp=s,
                             thr=float(d.get("Thread Instructions Executed", 0) or 0)))
    return recs


def stage_of(f, ln, ranges):
    for name, ff, a, b in ranges:
        if f == ff and a <= ln <= b:
            return name
    return f


def function_ranges(root):
    """(stage, file, first line, last line) from the function definitions of the csrc headers."""
    import os
    out = []
    pat = re.compile(r"^(?:template\s*<[^>]*>\s*)?(?:PF_HD|__device__|__global__|static|inline)[^;{]*?\b(pf_\w+)\s*\(")
    for fn in ("pf_kernels.cuh", "pf_math.cuh", "pf_tail.cuh", "pf_tail_dev.cuh", "pf_meancov.cuh", "pf_p3p.cu"):
        p = os.path.join(root, "pf_monocular_pose_estimator_b200", "csrc", fn)
        if not os.path.exists(p):
            continue
        lines = open(p).read().split("\n")
        starts = []
        for i, l in enumerate(lines):
            m = pat.match(l.strip()) if not l.startswith(" ") else None
            if m and not l.rstrip().endswith(";"):
                starts.append((i + 1, m.group(1)))
        for k, (ln, name) in enumerate(starts):
            end = starts[k + 1][0] - 1 if k + 1 < len(starts) else len(lines)
            out.append((name, fn, ln, end))
    return out


PIPE = [("fma", r"^(FFMA|FMUL|FADD|FSEL|FMNMX|FSET|FSETP|FCHK)"), ("fp64", r"^(DADD|DMUL|DFMA|DSETP|DMNMX)"),
        ("xu", r"^(MUFU|F2I|I2F|F2F|I2I|FRND|F2FP)"), ("int/alu", r"^(IADD|IMAD|IMNMX|ISETP|LOP|SHF|SHL|SHR|LEA|SEL|PRMT|FLO|POPC|BREV|IABS|ICMP|PLOP|VIADD|VIMNMX|MOV|CS2R|S2R|R2P|P2R|UMOV|UIADD|ULOP|USHF|ULEA|UISETP|USEL|UIMAD|UPLOP|R2UR|S2UR|UFLO|UPOPC|VOTE|VOTEU|NOP|IDP|LDC|ULDC)"),
        ("shfl/redux", r"^(SHFL|REDUX|MATCH)"), ("lds/sts", r"^(LDS|STS|LDSM|ATOMS)"), ("ldg/stg", r"^(LDG|STG|LD|ST|ATOM|RED|ATOMG|LDL|STL|CCTL|MEMBAR|ERRBAR)"),
        ("branch", r"^(BRA|BSSY|BSYNC|CALL|RET|EXIT|WARPSYNC|BAR|BRX|JMP|YIELD|BREAK|BMOV|DEPBAR|NANOSLEEP)")]


def pipe_of(op):
    for name, pat in PIPE:
        if re.match(pat, op):
            return name
    return "other"


def main():
    rep, steps = sys.argv[1], float(sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ranges = function_ranges(root)
    recs = load(rep)
    tot = sum(r["n"] for r in recs)
    smp = sum(r["smp"] for r in recs)
    print("warp instructions executed %.4g (%.1f per 32-particle step = thread instructions per particle), stall samples %d" % (tot, tot / steps, smp))
    st = collections.OrderedDict()
    for r in recs:
        op = re.sub(r"^@!?U?P\d+\s+", "", r["sass"]).split()[0].split(".")[0]
        r["op"] = op
        s = st.setdefault(stage_of(r["file"], r["line"], ranges), dict(n=0.0, smp=0.0, ops=collections.Counter(), pipes=collections.Counter()))
        s["n"] += r["n"]; s["smp"] += r["smp"]; s["ops"][op] += r["n"]; s["pipes"][pipe_of(op)] += r["n"]
    print("\n%-26s %9s %6s %6s  %s" % ("stage (inlined function)", "ins/step", "ins%", "smp%", "top opcodes (per step)"))
    for name, s in sorted(st.items(), key=lambda kv: -kv[1]["n"]):
        if s["n"] / tot < 0.002:
            continue
        print("%-26s %9.1f %5.1f%% %5.1f%%  %s" % (name, s["n"] / steps, 100 * s["n"] / tot, 100 * s["smp"] / max(smp, 1),
                                                  " ".join("%s=%.0f" % (o, c / steps) for o, c in s["ops"].most_common(7))))
    pipes = collections.Counter()
    for s in st.values():
        pipes.update(s["pipes"])
    print("\npipe mix (share of executed warp instructions): " + "  ".join("%s %.1f%%" % (k, 100 * v / tot) for k, v in pipes.most_common()))
    ops = collections.Counter()
    for r in recs:
        ops[r["op"]] += r["n"]
    print("opcodes per step: " + " ".join("%s=%.1f" % (o, c / steps) for o, c in ops.most_common(30)))
    agg = collections.OrderedDict()
    for r in recs:
        a = agg.setdefault((r["file"], r["line"], r["src"][:90]), [0.0, 0.0])
        a[0] += r["n"]; a[1] += r["smp"]
    print("\ntop source lines by instructions:")
    for (f, ln, src), (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%7.1f/step %5.1f%% smp %5.1f%%  %s:%d  %s" % (n / steps, 100 * n / tot, 100 * s / max(smp, 1), f, ln, src))


if __name__ == "__main__":
    main()
