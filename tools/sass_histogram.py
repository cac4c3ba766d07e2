#!/usr/bin/env python
"""SASS evidence from the built library (no GPU needed): an opcode histogram per kernel and, for selected kernels,
the listing of the instructions that prove what they are (tcgen05 / TMA / TMEM, peer and multimem traffic, packed fp32).

    python tools/sass_histogram.py            # writes profiles/sass/opcode_histogram.md and profiles/sass/r2_*.sass
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "pytorch_distributed_rnn_b200" / "_lib" / "libb200rnn.so"
OUT = ROOT / "profiles" / "sass"
KEY = ("UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTCBAR", "LDTM", "STTM", "SYNCS", "UCGABAR",
       "LDGMC", "FFMA2", "FFMA", "MUFU", "LDS", "STS", "SHFL", "BAR", "MEMBAR", "LDG", "STG", "RED", "ATOMG", "CCTL")
LISTINGS = {   # file name -> (substring of the mangled kernel name, regex of the instructions to keep)
    "r2_comm_oneshot_ll.sass": ("oneshot_ll_kernelILb1ELb1", r"ST\.E.*SYS|LD\.E.*SYS|LDG.*SYS|STG.*SYS|STG\.E\.128|LDG\.E\.128|BAR\.SYNC|FADD|NANOSLEEP"),
    "r2_reducer_wait.sass": ("reducer_wait_kernel", r"LDG|LD\.E|NANOSLEEP|CS2R|S2UR|STG"),
    "r2_head_ce.sass": ("head_ce_sample_kernel", r"SHFL|MUFU|RED|STG|LDG|BAR"),
    "r2_lstm_bwd_ffma2.sass": ("lstm_bwd_kernelILi32ELi4ELi12ELi1ELi3ELb0ELb1E", r"FFMA2|LDS\.128|BAR\.SYNC|SHFL|MUFU|ST\.E|STG|LDG.*STRONG"),
}


def demangle_short(name: str) -> str:
    try:
        full = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    except OSError:
        full = name
    full = full.replace("b200rnn::(anonymous namespace)::", "").replace("void ", "")
    return full.split("(")[0][:90]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
            continue
        if cur is not None and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
            kernels[cur].append(line)
    OUT.mkdir(parents=True, exist_ok=True)
    lines = ["# SASS opcode histogram per kernel (`python tools/sass_histogram.py`: cuobjdump -sass libb200rnn.so, sm_100a)", "",
             "`UTCHMMA` = tcgen05.mma (`.2CTA` = cta_group::2), `LDTM`/`STTM` = tcgen05.ld/st, `UTMALDG` / `UTMASTG` / `UTMAREDG` = TMA load / "
             "store / reduce-store, `UBLKCP` = 1-D bulk copy, `UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier ops, `UCGABAR` = cluster "
             "barrier, `LDGMC` = multimem.ld_reduce, `FFMA2` = packed fp32 FMA.", "",
             "| kernel | instr | " + " | ".join(KEY) + " |", "|---|---|" + "---|" * len(KEY)]
    for name, ins in kernels.items():
        if not ins:
            continue
        ops = collections.Counter()
        for l in ins:
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
            if m:
                ops[m.group(1)] += 1
        row = [str(sum(v for k, v in ops.items() if k == key or (key in ("RED",) and k.startswith("RED")))) for key in KEY]
        # FFMA column must not double count FFMA2 (exact opcode match above takes care of it)
        lines.append(f"| `{demangle_short(name)}` | {len(ins)} | " + " | ".join(row) + " |")
    (OUT / "opcode_histogram.md").write_text("\n".join(lines) + "\n")
    for fname, (needle, pat) in LISTINGS.items():
        for name, ins in kernels.items():
            if needle in name:
                keep = [l for l in ins if re.search(pat, l)]
                (OUT / fname).write_text(f"// {name}\n// instructions matching /{pat}/ of {len(ins)}\n" + "\n".join(keep) + "\n")
                break
        else:
            print("kernel not found:", needle, file=sys.stderr)
    print("kernels:", sum(1 for v in kernels.values() if v))


if __name__ == "__main__":
    main()
