import re,csv,collections,sys,subprocess
rep=sys.argv[1]; func=sys.argv[2]; srcfile=sys.argv[3] if len(sys.argv)>3 else 'star_kernels.cuh'
subprocess.run("cd /tmp && rm -f *.cubin && cuobjdump -xelf all /root/repo/fishman_b200/libfishman_b200.so >/dev/null 2>&1 && nvdisasm -g /tmp/fishman_b200.sm_100a.cubin > /tmp/all.sass 2>/dev/null", shell=True)
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout
lines=open('/tmp/all.sass').read().split('\n')
start=next(i for i,l in enumerate(lines) if '.text.'+func in l)
cur=None; amap={}
for l in lines[start+1:]:
    if '.section' in l and amap: break
    m=re.search(r'//## File "([^"]+)", line (\d+)',l)
    if m: cur=(m.group(1).split('/')[-1],int(m.group(2))); continue
    m=re.search(r'/\*([0-9a-f]{4,})\*/\s+(\S.*?);',l)
    if m: amap[int(m.group(1),16)]=(cur,m.group(2))
rows=list(csv.reader(out.split('\n')))
hi=next(i for i,r in enumerate(rows) if r and r[0]=='Address')
h=rows[hi]; R=[r for r in rows[hi+1:] if r and r[0].startswith('0x')]
def I(x):
    try: return int(x)
    except: return 0
ia=h.index('Address'); iall=h.index('Warp Stall Sampling (All Samples)'); iex=h.index('Instructions Executed'); ith=h.index('Avg. Threads Executed')
ils=h.index('stall_long_sb') if 'stall_long_sb' in h else None
seen=set(); agg=collections.defaultdict(lambda:[0,0,0.0,0])
base=int(R[0][ia],16)
for r in R:
    if r[ia] in seen: continue
    seen.add(r[ia])
    a=int(r[ia],16)-base
    key=amap.get(a,(None,''))[0]
    d=agg[key]; d[0]+=I(r[iall]); d[1]+=I(r[iex]); d[2]+=float(r[ith] or 0)*I(r[iex]); d[3]+=I(r[ils]) if ils else 0
tot=sum(d[0] for d in agg.values()); totex=sum(d[1] for d in agg.values())
print('samples',tot,'warp instr',totex, 'avg thr %.1f'%(sum(d[2] for d in agg.values())/totex))
src=open('/root/repo/fishman_b200/csrc/'+srcfile).read().split('\n')
for k,d in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[4]) if len(sys.argv)>4 else 30]:
    txt = src[k[1]-1].strip()[:64] if k and k[0]==srcfile else ''
    print(k, '%5.1f%% samp'%(100*d[0]/tot), '(long_sb %4.1f%%)'%(100*d[3]/tot), '%5.1f%% instr'%(100*d[1]/totex), 'thr %.1f'%(d[2]/max(d[1],1)), txt)
