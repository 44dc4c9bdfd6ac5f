"""Stall reasons of knn_lane_kernel<0> per source region, from one `ncu --set full --import-source on` report.
    ncu -i report.ncu-rep --page source --csv > src.csv
    cuobjdump -xelf all open3d-py-extension_b200/libo3dpp_b200.so && nvdisasm -g o3dpp.sm_100a.cubin > all.sass
    python tools/ncu_regions.py src.csv all.sass
The SASS addresses of the report are mapped to CUDA source lines through nvdisasm's line info (the library must be the
profiled build), the lines to named regions of knn_lane.cuh; columns are percentages of all warp samples."""
import csv,re,collections,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]
names=['# Samples','stall_no_inst','Instructions Executed','stall_wait','stall_long_sb','stall_short_sb','stall_branch_resolving','stall_not_selected','stall_math']
ix={n:hdr.index(n) for n in ['Address']+names}
lines=open(sys.argv[2]).read().split('\n')
start=[i for i,l in enumerate(lines) if l.startswith('.text._Z15knn_lane_kernelILi0')][0]
addr2line={}; cur=None
for l in lines[start+1:]:
    if l.startswith('.text.') or l.startswith('.section'): break
    m=re.search(r'//## File "([^"]+)", line (\d+)',l)
    if m: cur=(m.group(1).split('/')[-1],int(m.group(2))); continue
    m=re.match(r'\s+/\*([0-9a-f]{4,6})\*/',l)
    if m: addr2line[int(m.group(1),16)]=cur
import os
src=open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),'open3d-py-extension_b200','csrc','knn_lane.cuh')).read().split('\n')
def find(pat): return [i+1 for i,l in enumerate(src) if pat in l][0]
marks=[('helpers',0),('gap',find('__device__ __forceinline__ float gap')),('collect',find('template <int LVL>')),('collect_levels',find('bool collect_levels')),('append+stage',find('void append(')),('kl_bin',find('int kl_bin(')),('sel-setup',find('KlAnswer kl_select')),('passA',find('// pass A')),('binscan',find('int c0 = 0, bstar')),('passC',find('// pass C:')),('passF',find('// pass F')),('certify',find('const float ubound')),('boundary-pick',find('const int need = K - c0')),('kernel-prolog',find('knn_lane_kernel(const GridDesc g')),('parts',find('int start = 0, len = 32')),('window',find('bool window = false')),('selcall',find('const int ncand =')),('output',find('// OUTPUT: sorted positions')),('stats',find('unsigned s1 = __reduce_add_sync')-2)]
marks.sort(key=lambda x:x[1])
def bucket(k):
    if not k: return 'none'
    f,l=k
    if f!='knn_lane.cuh': return f
    name=marks[0][0]
    for n,s0 in marks:
        if l>=s0: name=n
    return 'lane:'+name
agg=collections.defaultdict(lambda:[0.0]*9)
base=None
def fl(x):
    try: return float(x)
    except: return 0.0
for r in rows[2:]:
    if len(r)<len(hdr): continue
    try: a=int(r[ix['Address']],16)
    except: continue
    if base is None: base=a
    k=bucket(addr2line.get(a-base))
    g=agg[k]
    for j,n in enumerate(names): g[j]+=fl(r[ix[n]])
ts=sum(g[0] for g in agg.values()); ti=sum(g[2] for g in agg.values())
print('samples',ts,'inst',ti)
print(f"{'region':24s}  smp%  inst% | noinst wait long short branch notsel math (% of all samples)")
for k,g in sorted(agg.items(), key=lambda x:-x[1][0]):
    print(f"{k:24s} {100*g[0]/ts:6.2f} {100*g[2]/ti:6.2f} | {100*g[1]/ts:5.2f} {100*g[3]/ts:5.2f} {100*g[4]/ts:5.2f} {100*g[5]/ts:5.2f} {100*g[6]/ts:5.2f} {100*g[7]/ts:5.2f} {100*g[8]/ts:5.2f}")
