"""Static SASS instruction count per source line of one kernel (development tool; needs -lineinfo).
usage: python tools/sass_lines.py file.cubin kernel-substring [topN]"""
import collections, re, subprocess, sys
out = subprocess.run(["nvdisasm", "--print-line-info", sys.argv[1]], capture_output=True, text=True).stdout
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
on, cur = False, None
cnt = collections.Counter()
for ln in out.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+)", ln)
    if m:
        on = sys.argv[2] in m.group(1); continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        cnt[cur] += 1
print("total", sum(cnt.values()))
for k, v in cnt.most_common(top):
    print(v, k)
