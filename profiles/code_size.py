"""Instructions per source line of one kernel (nvdisasm -g of the cubin inside libpmfs.so): where the code size goes.
usage: python profiles/code_size.py <mangled kernel substring> [top]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

here = os.path.dirname(os.path.abspath(__file__))
lib = os.path.join(here, "..", "panman_b200", "libpmfs.so")
kern = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, check=True, capture_output=True)
    cub = [f for f in os.listdir(d) if f.startswith("pmfs.") and f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", os.path.join(d, cub)], capture_output=True, text=True).stdout
inside, cur, counts, total = False, None, collections.Counter(), 0
for line in txt.splitlines():
    if line.startswith("//---") and ".text." in line:
        inside = kern in line
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        counts[cur] += 1
        total += 1
print("instructions", total, "=", total * 16 // 1024, "KB")
src = {}
for (f, ln), n in counts.most_common(top):
    if f not in src:
        for root in (os.path.join(here, "..", "panman_b200", "csrc"),):
            pth = os.path.join(root, f)
            src[f] = open(pth).read().splitlines() if os.path.exists(pth) else []
    text = src[f][ln - 1].strip()[:100] if 0 < ln <= len(src[f]) else ""
    print(f"{n:5d}  {f}:{ln}  {text}")
