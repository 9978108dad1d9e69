"""Reads an .ncu-rep (one kernel) and prints headline metrics + warp-stall samples per SASS region.
usage: python tools/ncu_regions.py report.ncu-rep [step]"""
import collections, csv, subprocess, sys, io
rep = sys.argv[1]; step = int(sys.argv[2]) if len(sys.argv) > 2 else 150
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units, r = rows[0], rows[1], rows[2]
for k in ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
          "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
          "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
          "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.max", "launch__registers_per_thread",
          "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct"]:
    if k in hdr:
        print(f"{k:85s} {r[hdr.index(k)]:>16s} {units[hdr.index(k)]}")
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(sass))); hdr = rows[1]; data = [x for x in rows[2:] if len(x) > 5]
iS, iSrc = hdr.index("# Samples"), hdr.index("Source")
stall = [h for h in hdr if h.startswith("stall_") and "Not" not in h]
tot = sum(int(x[iS] or 0) for x in data)
print("total samples", tot, "instructions", len(data))
for s in range(0, len(data), step):
    ch = data[s:s + step]; n = sum(int(x[iS] or 0) for x in ch)
    if n < 0.01 * tot: continue
    st = collections.Counter(); ops = collections.Counter()
    for x in ch:
        for h in stall: st[h[6:]] += int(x[hdr.index(h)] or 0)
        t = x[iSrc].split(); o = t[1] if t[0].startswith('@') else t[0]; ops[o.split('.')[0]] += 1
    print(f"{s:5d} {n:6d} {100*n/tot:5.1f}%", st.most_common(4), ops.most_common(4))
top = sorted(data, key=lambda x: -int(x[iS] or 0))[:12]
for x in top: print(x[iS].rjust(6), x[iSrc][:80])
