"""Per-source-line executed-instruction profile: joins `ncu --page source --print-source sass` of a kernel with
`nvdisasm --print-line-info` of a cubin built from the same sources.
usage: sass_lines.py report.ncu-rep kernel_regex cubin mangled_substring [positions]"""
import collections, csv, re, subprocess, sys
rep, kre, cubin, mangled = sys.argv[1:5]
npos = float(sys.argv[5]) if len(sys.argv) > 5 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre, "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; ie = hdr.index("Instructions Executed"); ia = hdr.index("Address")
prof = []
for r in rows[2:]:
    if len(r) <= ie or r[ia] == "Address":
        break
    prof.append(r)
dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout.split("\n")
st = [i for i, l in enumerate(dis) if l.startswith("//---") and ".text." in l and mangled in l][0]
en = [i for i, l in enumerate(dis) if l.startswith("//---") and i > st][0]
cur = None; seq = []
for l in dis[st:en]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        seq.append((m.group(2), cur))
assert len(seq) == len(prof), (len(seq), len(prof))
agg = collections.Counter(); tot = 0
for (ins, loc), r in zip(seq, prof):
    n = int(r[ie]); tot += n; agg[loc] += n
print("total warp instructions", tot, "per position", tot / npos if npos else "")
for (f, ln), n in agg.most_common(40):
    print(f"{f}:{ln}  {n}  {100 * n / tot:.1f}%" + (f"  {n / npos:.1f}/pos" if npos else ""))
