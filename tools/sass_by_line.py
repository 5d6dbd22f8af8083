"""SASS instruction count of one kernel attributed to source lines (needs -lineinfo):
   python tools/sass_by_line.py build/wavefront.o k_extendILb0ELb1 [bucket]"""
import collections, os, re, subprocess, sys, tempfile
obj, pat = os.path.abspath(sys.argv[1]), sys.argv[2]
bucket = int(sys.argv[3]) if len(sys.argv) > 3 else 10
with tempfile.TemporaryDirectory() as d:
    subprocess.run(['cuobjdump', '-xelf', 'all', obj], cwd=d, capture_output=True)
    cubin = [f for f in os.listdir(d) if f.endswith('.cubin')][0]
    lines = subprocess.run(['nvdisasm', '-g', '-c', cubin], cwd=d, capture_output=True, text=True).stdout.splitlines()
start = end = None
for i, l in enumerate(lines):
    if l.startswith('.text.'):
        if start is None and pat in l: start = i
        elif start is not None: end = i; break
seg = lines[start:end]
cur = None; cnt = collections.Counter()
for l in seg:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l): cnt[cur] += 1
print("instructions", sum(cnt.values()))
byfile = collections.Counter()
for (f, ln), c in cnt.items(): byfile[f] += c
print(byfile.most_common())
b = collections.Counter()
for (f, ln), c in cnt.items(): b[(f, ln // bucket * bucket)] += c
for k, c in sorted(b.items()):
    if c >= 8: print(k, c)
