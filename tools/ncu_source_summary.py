import csv,sys,collections
rep=sys.argv[1]
import subprocess
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','sass,cuda'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
cur_file=None; agg=collections.OrderedDict(); hdr=None
tot=0
for r in rows:
    if len(r)==2 and r[0]=='File Path': cur_file=r[1].split('/')[-1]; continue
    if len(r)>5 and r[0]=='Line No': hdr=r; iI=hdr.index('Instructions Executed'); iS=hdr.index('# Samples'); continue
    if hdr and len(r)>iI and r[0]!='' :
        key=(cur_file,int(r[0]),r[1].strip()[:90])
        try: v=int(r[iI]); s=int(r[iS])
        except: continue
        a=agg.setdefault(key,[0,0]); a[0]+=v; a[1]+=s
tot=sum(a[0] for a in agg.values()); tots=sum(a[1] for a in agg.values())
print('total inst',tot,'samples',tots)
for k,a in sorted(agg.items(), key=lambda kv:-kv[1][0])[:int(sys.argv[2]) if len(sys.argv)>2 else 40]:
    print(f"{k[0]:>18}:{k[1]:<4} inst {100*a[0]/tot:5.1f}%  samples {100*a[1]/tots:5.1f}%  {k[2]}")
