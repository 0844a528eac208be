import itertools
NS=20; NR=11
def word(x,ty): return (x>>1)*20 + (x>>4) + (ty>>1)
def wavefronts(lanes, c0):
    tot=0
    for dx,dy in ((-1,-1),(0,-1),(-1,0),(0,0)):
        banks={}
        for (k,sr) in lanes:
            x=3*k+2+dx; ty=3*sr+c0+dy
            w=word(x,ty); banks.setdefault(w%32,set()).add(w)
        tot+=max(len(v) for v in banks.values())
    return tot
def evaluate(name, mapping):   # mapping: list of (k,sr) in thread order
    warps=[mapping[i:i+32] for i in range(0,len(mapping),32)]
    res=[]
    for c0 in (2,3,4):
        res.append(sum(wavefronts(w,c0) for w in warps))
    print(f"{name:28s} warps {len(warps)} wavefronts per 4 loads (c0=2,3,4): {res}  ideal {4*len(warps)}")
rowmajor=[(k,sr) for sr in range(NR) for k in range(NS)]
evaluate("row-major (current)", rowmajor)
colmajor=[(k,sr) for k in range(NS) for sr in range(NR)]
evaluate("column-major", colmajor)
for RB,KB in ((4,8),(2,16),(8,4),(3,10),(6,5),(11,2),(11,3),(1,20),(2,10),(4,5),(5,4)):
    m=[]
    for r0 in range(0,NR,RB):
        for k0 in range(0,NS,KB):
            for sr in range(r0,min(r0+RB,NR)):
                for k in range(k0,min(k0+KB,NS)):
                    m.append((k,sr))
    evaluate(f"blocks {RB} rows x {KB} k", m)
# interleaved: even ks then odd ks
for RB in (2,3,4):
    m=[]
    for r0 in range(0,NR,RB):
        for par in (0,1):
            for sr in range(r0,min(r0+RB,NR)):
                for k in range(par,NS,2): m.append((k,sr))
    evaluate(f"{RB} rows x 10 k of one parity", m)
print("---- skew search (row-major mapping)")
best=[]
def word2(x,ty,sk,TRP=20): return (x>>1)*TRP + sk[x>>4] + (ty>>1)
def wf2(lanes,c0,sk,TRP=20):
    tot=0
    for dx,dy in ((-1,-1),(0,-1),(-1,0),(0,0)):
        banks={}
        for (k,sr) in lanes:
            w=word2(3*k+2+dx,3*sr+c0+dy,sk,TRP); banks.setdefault(w%32,set()).add(w)
        tot+=max(len(v) for v in banks.values())
    return tot
for name,mp in (("row-major",rowmajor),):
    warps=[mp[i:i+32] for i in range(0,len(mp),32)]
    for sk in itertools.product(range(8),repeat=3):
        sk=(0,)+sk
        tot=max(sum(wf2(w,c0,sk) for w in warps) for c0 in (2,3,4))
        best.append((tot,sk))
best.sort(); print(best[:8])
print("---- 16 k x 2 rows + leftovers")
def ev(name, warps):
    res=[sum(wavefronts(w,c0) for w in warps) for c0 in (2,3,4)]
    print(f"{name:40s} warps {len(warps)} {res} ideal {4*len(warps)} max lanes {max(len(w) for w in warps)}")
for KB in (16,12,10,14):
    warps=[]
    for r0 in range(0,NR,2):
        warps.append([(k,sr) for sr in range(r0,min(r0+2,NR)) for k in range(KB)])
    left=[(k,sr) for sr in range(NR) for k in range(KB,NS)]
    for i in range(0,len(left),32): warps.append(left[i:i+32])
    ev(f"{KB} k x 2 rows + leftovers row-major", warps)
    warps2=warps[:6 if NR>10 else 5]
    left=[(k,sr) for k in range(KB,NS) for sr in range(NR)]
    w3=list(warps[: (NR+1)//2])
    for i in range(0,len(left),32): w3.append(left[i:i+32])
    ev(f"{KB} k x 2 rows + leftovers col-major", w3)
# rows paired with distance d instead of adjacent
for d in (2,3,4,5,6):
    used=set(); warps=[]
    for r in range(NR):
        if r in used: continue
        r2=r+d
        if r2<NR and r2 not in used:
            used|={r,r2}; warps.append([(k,sr) for sr in (r,r2) for k in range(16)])
        else:
            used.add(r); warps.append([(k,r) for k in range(16)])
    left=[(k,sr) for sr in range(NR) for k in range(16,NS)]
    for i in range(0,len(left),32): warps.append(left[i:i+32])
    ev(f"16 k x rows (r, r+{d}) + leftovers", warps)
