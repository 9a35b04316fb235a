"""Count shared-memory wavefronts (64-bit accesses, half a warp per wavefront, 16 bank pairs) of the FFT passes, the
untangle and the contiguous walks of csrc/fftg.cuh for candidate row paddings.  usage: python scripts/bank_sim.py"""
import math, itertools
def wavefronts(addrs):  # addrs: list of 32 physical float2 indices (or None for inactive)
    tot=0
    for hw in (addrs[:16], addrs[16:]):
        cnt={}
        seen=set()
        for a in hw:
            if a is None or a in seen: continue
            seen.add(a); cnt[a%16]=cnt.get(a%16,0)+1
        tot+=max(cnt.values()) if cnt else 0
    return tot
def passes(L):
    n=(L+3)//4; base=L//n; extra=L-base*n
    return [base+(1 if i<extra else 0) for i in range(n)]
def sim(L, frames, pad, zs_of):
    n=1<<L; zs=zs_of(n); total=0; ideal=0
    lh=L-1
    for R in passes(L):
        P=1<<R; lhs=lh-(R-1); hs=1<<lhs; lq=L-R
        items=frames<<lq
        for w0 in range(0, items, 32):
            for m in range(P):
                ad=[]
                for idx in range(w0, w0+32):
                    if idx>=items: ad.append(None); continue
                    f=idx>>lq; q=idx&((1<<lq)-1); g=q>>lhs; j=q&(hs-1)
                    base=(g<<(lh+1))+j
                    ad.append(f*zs+pad(base+(m<<lhs)))
                wf=wavefronts(ad); total+=2*wf; ideal+=2*2*(sum(a is not None for a in ad)/32)
        lh-=R
    # untangle reads: lane i -> positions brev(k), brev(n-k)
    def brev(x,bits):
        r=0
        for b in range(bits): r|=((x>>b)&1)<<(bits-1-b)
        return r
    half=n>>1
    un=0
    for f in range(frames):
        for w0 in range(0, half, 32):
            a1=[];a2=[]
            for i in range(w0,w0+32):
                k=brev(i,L-1)
                a1.append(f*zs+pad(brev(k,L))); a2.append(f*zs+pad(brev((n-k)&(n-1),L)))
            un+=wavefronts(a1)+wavefronts(a2)
    # contiguous (loader store / ola read)
    co=0
    for f in range(frames):
        for w0 in range(0,n,32):
            co+=wavefronts([f*zs+pad(i) for i in range(w0,w0+32)])
    return total, ideal, un, co
pads={
 'none': (lambda p:p, lambda n:n),
 '>>3': (lambda p:p+(p>>3), lambda n:n+(n>>3)),
 '>>4': (lambda p:p+(p>>4), lambda n:n+(n>>4)),
 '>>5': (lambda p:p+(p>>5), lambda n:n+(n>>5)),
 '>>4+>>8': (lambda p:p+(p>>4)+(p>>8), lambda n:n+(n>>4)+(n>>8)),
 '>>5+>>9': (lambda p:p+(p>>5)+(p>>9), lambda n:n+(n>>5)+(n>>9)),
 '>>4 x2?': (lambda p:p+2*(p>>5), lambda n:n+2*(n>>5)),
}
for L in (10,11,9,12):
    print("n =",1<<L, "passes",passes(L))
    for name,(pf,zf) in pads.items():
        t,i,u,c=sim(L,2,pf,zf)
        print("  %-9s fft wavefronts %7d (ideal %7d)  untangle %6d  contiguous %6d   sum(fft + untangle + 2*contig) %d"%(name,t,i,u,c,t+u+2*c))
