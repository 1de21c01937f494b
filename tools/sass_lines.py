"""Instructions per source line of one kernel: python tools/sass_lines.py <nvdisasm --print-line-info output> <kernel substring> [top]"""
import collections, re, sys
lines = open(sys.argv[1]).read().split("\n")
key = sys.argv[2]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
start = next(i for i, l in enumerate(lines) if re.match(r"\s*\.section\s+\.text\.", l) and key in l)
cnt = collections.Counter(); sub = collections.Counter(); cur = None; fn = "(kernel body)"; n = 0
for l in lines[start + 1:]:
    if re.match(r"\s*\.section", l): break
    m = re.search(r'//## File "(.*?)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m2 = re.match(r"\s*(\$[^:]+):", l)
    if m2: fn = m2.group(1).split("$")[-1]
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+\S", l): cnt[cur] += 1; sub[fn] += 1; n += 1
print(n, "instructions")
for k, v in sub.most_common(): print("%6d  %s" % (v, k))
for k, v in cnt.most_common(top):
    src = ""
    try: src = open("diceoptim_b200/csrc/" + k[0]).read().split("\n")[k[1] - 1].strip()[:90]
    except Exception: pass
    print("%6d  %s:%d  %s" % (v, k[0], k[1], src))
