"""Per-source-line warp-stall samples: joins `ncu --page source --print-source sass` of a kernel with
`nvdisasm --print-line-info` of a cubin built from the same sources.
usage: sass_stalls.py report.ncu-rep kernel_regex cubin mangled_substring [top]"""
import collections, csv, re, subprocess, sys
rep, kre, cubin, mangled = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre, "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ia, ins_i, smp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
prof = []
for r in rows[2:]:
    if len(r) <= smp or r[ia] == "Address":
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
agg = collections.Counter(); why = collections.defaultdict(collections.Counter); tot = 0; inst = collections.Counter()
for (ins, loc), r in zip(seq, prof):
    n = int(r[smp] or 0); tot += n; agg[loc] += n; inst[loc] += int(r[ins_i] or 0)
    for c in stall_cols:
        v = int(r[c] or 0)
        if v:
            why[loc][hdr[c]] += v
print("total samples", tot)
for (f, ln), n in agg.most_common(top):
    reasons = ", ".join(f"{k[6:]} {v}" for k, v in why[(f, ln)].most_common(3))
    print(f"{f}:{ln}  {100 * n / tot:5.1f}%  inst {inst[(f, ln)]}  [{reasons}]")
