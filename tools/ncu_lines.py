"""Aggregates an .ncu-rep source page by CUDA source line: stall samples and executed warp instructions."""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; skip = sys.argv[2] if len(sys.argv) > 2 else "0"; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
agg = collections.defaultdict(lambda: [0, 0, "", collections.Counter()])
cur_file = ""; hdr = None; cur_line = None; cur_src = ""
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; iS = hdr.index("Warp Stall Sampling (All Samples)"); iI = hdr.index("Instructions Executed"); 
    elif hdr:
        if r[0] != "": cur_line = r[0]; cur_src = r[1]
        try: s = int(r[iS]); n = int(r[iI])
        except Exception: continue
        key = (cur_file, cur_line)
        a = agg[key]; a[0] += s; a[1] += n; a[2] = cur_src
        for name in ("stall_barrier", "stall_long_sb", "stall_short_sb", "stall_wait", "stall_math", "stall_branch_resolving", "stall_no_inst","stall_not_selected","stall_selected","stall_dispatch","stall_mio"):
            try: a[3][name] += int(r[hdr.index(name)])
            except Exception: pass
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print("total samples", ts, "total warp instr", ti)
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    st = ",".join(f"{k[6:]}:{v}" for k, v in a[3].most_common(3))
    print(f"{key[0][:22]:22s}:{key[1]:>4s} smp {100*a[0]/ts:5.1f}% ins {100*a[1]/ti:5.1f}% [{st}] {a[2].strip()[:80]}")
