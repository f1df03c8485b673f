import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]; data=rows[2:]
ia=hdr.index("Address"); isrc=hdr.index("Source"); ismp=hdr.index("# Samples"); iex=hdr.index("Instructions Executed")
tot_s=sum(int(r[ismp]) for r in data); tot_i=sum(int(r[iex]) for r in data)
print("instrs",len(data),"samples",tot_s,"inst_exec",tot_i)
# segment into regions of 100 instructions
seg=int(sys.argv[2]) if len(sys.argv)>2 else 100
for a in range(0,len(data),seg):
    blk=data[a:a+seg]
    s=sum(int(r[ismp]) for r in blk); i=sum(int(r[iex]) for r in blk)
    ops={}
    for r in blk:
        op=r[isrc].split()[0] if not r[isrc].strip().startswith('@') else r[isrc].split()[1]
        ops[op.split('.')[0]]=ops.get(op.split('.')[0],0)+1
    top=sorted(ops.items(),key=lambda x:-x[1])[:5]
    print(f"{a:5d}-{a+len(blk):5d} samples {100*s/tot_s:5.1f}% inst {100*i/tot_i:5.1f}%  {top}")
