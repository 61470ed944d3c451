"""SASS instruction count per CUDA source line for one kernel of libyolohead.so.
usage: python tools/sass_lines.py <cubin-name-fragment> <kernel-name-fragment> [top]"""
import collections, glob, os, re, subprocess, sys, tempfile
frag, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(here, "yolo_v1_b200", "libyolohead.so")], cwd=tmp, capture_output=True)
cub = [f for f in glob.glob(os.path.join(tmp, "*.cubin")) if os.path.basename(f).startswith(frag + ".")][0]
dis = subprocess.run(["nvdisasm", "-g", cub], capture_output=True, text=True).stdout
cnt = collections.Counter(); cur = None; infn = False; sec = collections.Counter(); name = None
for line in dis.splitlines():
    m = re.match(r'\s*\.section\s+(\.\S+?),', line)
    if m:
        name = m.group(1); infn = name.startswith('.text.') and kern in name; continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', line):
        sec[name] += 1
        if infn: cnt[cur] += 1
for n, v in sec.most_common(8): print(f"{v*16/1024:7.1f} KB  {n[:90]}")
tot = sum(cnt.values()); print("kernel instr", tot, f"{tot*16/1024:.1f} KB")
for k, v in cnt.most_common(top): print(v, k)
