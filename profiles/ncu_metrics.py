import csv,sys,subprocess
rep=sys.argv[1]
out=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr,units,vals=rows[0],rows[1],rows[2]
pats=sys.argv[2:]
for i,h in enumerate(hdr):
    if any(p in h for p in pats):
        print(f"{h:90s} {units[i]:12s} {vals[i]}")
