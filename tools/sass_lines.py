#!/usr/bin/env python
"""Instructions per source line of one kernel in the built library (needs -lineinfo; no GPU needed).

    python tools/sass_lines.py <lib.so> <substring of the mangled kernel name> [top]

Extracts the cubin (cuobjdump -xelf), disassembles it with line info (nvdisasm --print-line-info) and counts the SASS
instructions attributed to each file:line -- the greedy kernel is sensitive to code size (instruction-cache tiers at
32 KB), so this is the map of where its bytes are."""
import collections
import os
import re
import subprocess
import sys
import tempfile


def main():
    lib, pat = os.path.abspath(sys.argv[1]), sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=d, check=True, stdout=subprocess.DEVNULL)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, cubin)], capture_output=True, text=True).stdout
    inside, cur = False, "?"
    per_line, per_file, total = collections.Counter(), collections.Counter(), 0
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            inside = pat in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]*)", line (\d+)', ln)
        if m:
            cur = "%s:%s" % (os.path.basename(m.group(1)), m.group(2))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            per_line[cur] += 1
            per_file[cur.split(":")[0]] += 1
            total += 1
    print("total %d instructions = %.1f KB" % (total, total * 16 / 1024))
    for f, c in per_file.most_common():
        print("  %5d  %s" % (c, f))
    print("top lines:")
    for k, c in per_line.most_common(top):
        print("  %5d  %s" % (c, k))


if __name__ == "__main__":
    main()
