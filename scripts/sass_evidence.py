"""Per-kernel SASS evidence: which Blackwell-native instructions each kernel of libainic_cuda.so
contains (B200_PROFILING.md: tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG,
multimem -> LDGMC / STG.MC ...).  Writes profiles/sass_evidence.md; the full listing can be
regenerated with `python -m fpga_ai_nic_b200.utils.build --sass`."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fpga_ai_nic_b200", "_lib", "libainic_cuda.so")
KEYS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTCATOMSWS", "UCGABAR", "SYNCS", "LDGMC",
        "REDG", "HMMA", "LDG.E", "STG.E", "ERRBAR", "MEMBAR", "FFMA", "SHFL", "STL", "LDL", "ACQBULK", "NANOSLEEP"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            for k in KEYS:
                if op.startswith(k):
                    kernels[cur][k] += 1
            kernels[cur]["_total"] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    lines = ["# SASS evidence per kernel (`cuobjdump -sass fpga_ai_nic_b200/_lib/libainic_cuda.so`)\n",
             "Counts of the instructions that identify the Blackwell-native paths; `STL`/`LDL` are local-memory",
             "(spill / stack) accesses.\n", "| kernel | instr | " + " | ".join(KEYS) + " |", "|---|---|" + "---|" * len(KEYS)]
    for (name, c), dm in zip(kernels.items(), demangle):
        short = re.sub(r"\(.*", "", dm).replace("ainic::", "").replace("(anonymous namespace)::", "")
        short = short.replace("void ", "")[:70]
        lines.append(f"| `{short}` | {c['_total']} | " + " | ".join(str(c[k]) if c[k] else "" for k in KEYS) + " |")
    path = os.path.join(ROOT, "profiles", "sass_evidence.md")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    open(path, "w").write("\n".join(lines) + "\n")
    print(path, len(kernels), "kernels")


if __name__ == "__main__":
    sys.exit(main())
