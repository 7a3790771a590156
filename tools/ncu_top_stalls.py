import csv, subprocess, sys
rep=sys.argv[1]
raw=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr,vals=rows[0],rows[2]
d=dict(zip(hdr,vals))
print(d.get("Kernel Name"))
for k in ["gpu__time_duration.sum","smsp__inst_executed.sum","smsp__issue_active.avg.pct_of_peak_sustained_active","l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed","gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed","lts__throughput.avg.pct_of_peak_sustained_elapsed","dram__bytes_read.sum","dram__bytes_write.sum","launch__registers_per_thread","smsp__warps_eligible.avg.per_cycle_active"]:
    print(k, d.get(k))
for k,v in d.items():
    if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio"):
        try:
            if float(v)>0.3: print("  ",k.split("stalled_")[1].split("_per_issue")[0], v)
        except: pass
src=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","sass"],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hdr=rows[1]; ix={h:i for i,h in enumerate(hdr)}; data=rows[2:]
def f(r,k):
    try: return float(r[ix[k]])
    except: return 0.0
tot=sum(f(r,'# Samples') for r in data)
print("total samples",tot)
for r in sorted(data,key=lambda r:-f(r,'# Samples'))[:int(sys.argv[2]) if len(sys.argv)>2 else 22]:
    print("%6s %5.1f%% long=%4d short=%4d wait=%4d mio=%4d math=%3d lg=%3d exec=%9s  %s" % (r[ix['Address']][-5:], 100*f(r,'# Samples')/tot, f(r,'stall_long_sb'), f(r,'stall_short_sb'), f(r,'stall_wait'), f(r,'stall_mio'), f(r,'stall_math'), f(r,'stall_lg'), r[ix['Instructions Executed']], r[ix['Source']][:60]))
