"""Static SASS instruction count per source line of one kernel (no GPU needed).
usage: python scripts/static_sass_by_line.py <lib.so> <mangled kernel name> [top_n]"""
import collections, os, re, subprocess, sys, tempfile
lib, fun = sys.argv[1], sys.argv[2]; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
d = tempfile.mkdtemp(); subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, capture_output=True)
cub = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-g", "-fun", fun, cub], capture_output=True, text=True).stdout
cur = None; cnt = collections.Counter(); tot = 0
for line in txt.splitlines():
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', line): cnt[cur] += 1; tot += 1
print("static SASS instructions:", tot)
bf = collections.Counter()
for k, v in cnt.items(): bf[k[0] if k else None] += v
print(bf.most_common())
src = {}
for k, v in cnt.most_common(topn):
    text = ""
    if k:
        for base in ("intent_radio_sched_multi_slice_b200/csrc", "/usr/local/cuda/include"):
            p = os.path.join(base, k[0])
            if os.path.exists(p):
                L = open(p).read().splitlines(); text = L[k[1]-1].strip()[:80] if k[1] <= len(L) else ""; break
    print(k, v, "|", text)
