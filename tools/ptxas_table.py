"""Registers / stack / spills per kernel from `nvcc -Xptxas -v` (no GPU needed).

    python tools/ptxas_table.py tc_gemm tc_wgrad edge_bf16 segment geom > profiles/rNN_ptxas_registers.md
"""
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PAT = re.compile(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?Function properties for \1\n\s*(\d+) bytes stack frame, "
                 r"(\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers", re.S)


def main(files):
    print("| file | kernel | registers | stack B | spill st B | spill ld B |\n|---|---|---|---|---|---|")
    with tempfile.TemporaryDirectory() as tmp:
        for f in files:
            extra = ["-fmad=false"] if f.startswith("geom") else []
            r = subprocess.run(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17",
                                "-lineinfo", "-I", os.path.join(ROOT, "include"), "-Xptxas", "-v", "-c",
                                os.path.join(ROOT, "aether_b200", "csrc", f + ".cu"), "-o", os.path.join(tmp, f + ".o")] + extra,
                               capture_output=True, text=True, check=True)
            for m in PAT.finditer(r.stderr):
                name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
                name = name.replace("aether::(anonymous namespace)::", "").replace("void ", "")
                i = name.find(">(")
                name = name[:i + 1] if i >= 0 else name[:name.find("(")]
                name = name.replace("(int)", "").replace("(bool)1", "true").replace("(bool)0", "false")
                print("| %s | `%s` | %s | %s | %s | %s |" % (f, name, m.group(5), m.group(2), m.group(3), m.group(4)))


if __name__ == "__main__":
    main(sys.argv[1:] or ["tc_gemm", "tc_wgrad", "edge_bf16", "segment", "geom"])
