"""Per-source-line executed instructions and stall samples from a saved `ncu --page source --csv --print-source sass`
(gzipped; may hold several kernels), joined with `nvdisasm --print-line-info` of the cubin of the same build.
usage: sass_lines_csv.py source.csv.gz kernel_substring cubin mangled_substring [positions] [top]"""
import collections, csv, gzip, re, subprocess, sys
path, ksub, cubin, mangled = sys.argv[1:5]
npos = float(sys.argv[5]) if len(sys.argv) > 5 else None
top = int(sys.argv[6]) if len(sys.argv) > 6 else 30
rows = list(csv.reader(gzip.open(path, "rt")))
prof = []; take = False; done = False
for r in rows:
    if r and r[0] == "Kernel Name":
        if take and prof: break
        take = ksub in r[1]; continue
    if not take: continue
    if r and r[0] == "Address": hdr = r; continue
    prof.append(r)
ie, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout.split("\n")
st = [i for i, l in enumerate(dis) if l.startswith("//---") and ".text." in l and mangled in l][0]
en = [i for i, l in enumerate(dis) if l.startswith("//---") and i > st][0]
cur = None; seq = []
for l in dis[st:en]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: seq.append((m.group(2), cur))
assert len(seq) == len(prof), (len(seq), len(prof))
agg = collections.Counter(); smp = collections.Counter(); why = collections.defaultdict(collections.Counter); tot = 0; tots = 0
for (ins, loc), r in zip(seq, prof):
    n = int(r[ie]); tot += n; agg[loc] += n
    s = int(r[ismp] or 0); smp[loc] += s; tots += s
    for i in stall: why[loc][hdr[i][6:]] += int(r[i] or 0)
print("total warp instructions", tot, ("per position %.2f" % (tot / npos)) if npos else "", "samples", tots)
for (f, ln), n in agg.most_common(top):
    w = ", ".join(f"{k} {v}" for k, v in why[(f, ln)].most_common(2))
    print(f"{f}:{ln}  inst {100 * n / tot:.1f}%  samples {100 * smp[(f, ln)] / max(tots, 1):.1f}%  [{w}]")
