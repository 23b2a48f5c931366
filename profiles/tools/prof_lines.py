"""map an ncu report's SASS-level samples of a kernel to CUDA source lines (needs cuobjdump/nvdisasm)"""
import csv, re, collections, subprocess, sys, os, glob, tempfile
rep = sys.argv[1]; kern = sys.argv[2] if len(sys.argv) > 2 else "k_turbo_decode"; cu = sys.argv[3] if len(sys.argv) > 3 else "fec_rx"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", "/root/repo/dect-nr-plus-sdr_b200/libdnrp_cuda.so"], cwd=tmp, capture_output=True)
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cu + ".sm_100a.cubin")], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(sass) if l.startswith(".text.") and kern in l][0]
cur = None; seq = []
for l in sass[start + 1:]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m: seq.append((cur, m.group(2)))
    if (l.startswith(".text.") or l.startswith("//-----")) and seq: break
extra = os.environ.get("NCU_ARGS", "").split()
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + extra, capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n"))); h = rows[1]; data = [r for r in rows[2:] if len(r) > 10]; idx = {n: i for i, n in enumerate(h)}
print("sass", len(seq), "profile", len(data))
ss = idx["# Samples"]; ie = idx["Instructions Executed"]
st = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
for i in range(min(len(seq), len(data))):
    k = seq[i][0]; r = data[i]
    agg[k][0] += int(r[ss] or 0); agg[k][1] += int(r[ie] or 0)
    for s in st:
        v = int(r[idx[s]] or 0)
        if v: agg[k][2][s] += v
tot = sum(v[0] for v in agg.values()); toti = sum(v[1] for v in agg.values())
src = {}
for f in glob.glob("/root/repo/dect-nr-plus-sdr_b200/csrc/*"):
    src[os.path.basename(f)] = open(f).read().split("\n")
for k, v in sorted(agg.items(), key=lambda x: -x[1][0])[:int(os.environ.get("TOP", "30"))]:
    txt = src[k[0]][k[1] - 1].strip()[:64] if k and k[0] in src else ""
    print("%5.1f%% smp %5.1f%% ins %-12s:%4s %-64s %s" % (100 * v[0] / tot, 100 * v[1] / toti, (k[0] if k else "?")[:12], k[1] if k else "", txt, dict(v[2].most_common(2))))
