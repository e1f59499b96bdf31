import csv,sys,collections,subprocess
rep,pat=sys.argv[1],sys.argv[2]
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--kernel-name','regex:'+pat],capture_output=True,text=True).stdout
r=list(csv.reader(out.splitlines()))
rows=[];seen=0
for row in r:
    if row and row[0]=='Kernel Name':
        seen+=1
        if seen>1: break
        continue
    if row and row[0]=='Address': hdr=row; continue
    rows.append(row)
ie=hdr.index('Instructions Executed'); ss=hdr.index('# Samples')
tot=sum(int(x[ie]) for x in rows); print('total inst',tot,'sass rows',len(rows))
ops=collections.Counter(); samp=collections.Counter()
for x in rows:
    t=x[1].split()
    op=t[1] if t[0].startswith('@') else t[0]
    ops[op]+=int(x[ie]); samp[op]+=int(x[ss])
for k,v in ops.most_common(int(sys.argv[3]) if len(sys.argv)>3 else 25): print('%-22s %10d %5.1f%%  samples %d'%(k,v,100*v/tot,samp[k]))
