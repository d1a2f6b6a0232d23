"""Lane-level numpy model of the single-sweep FILIP epilogue (csrc/filip_tc1.cu): the tcgen05.ld.16x256b register layout,
the shared 5-bit key code, the per-tile row update, the column exchange and the finaliser -- checked against argmax / max
of a random score matrix.  Debug aid written before the kernel first ran (it caught nothing, which was the point):
python tools/filip_sweep_model.py"""
import numpy as np
rng=np.random.default_rng(0)
def f2u(x): return np.float32(x).view(np.uint32)
def u2f(x): return np.uint32(x).view(np.float32)
NEGKEY=np.uint32(0xff7fffe0); NEGMAX=np.uint32(0xff7fffff)
def reg_of(j,k,e): return 16*(j>>1)+4*k+2*(j&1)+e
def load_chunk(S, lane):  # S: 32x32 chunk (rows of warp quarter, cols of chunk)
    g,c=lane>>2,lane&3
    r=np.zeros(32,np.uint32)
    for ld in range(2):
        for k in range(4):
            for jj in range(2):
                for e in range(2):
                    r[16*ld+4*k+2*jj+e]=f2u(S[16*ld+g+8*jj, 8*k+2*c+e])
    return r
def chunk_both(R, cc, nv, rowmasks, st, rowwin):
    # R: [32 lanes][32 regs]; st: dict mk[lane][4], base[lane][4]
    keys=np.zeros((32,32),np.uint32); w=np.zeros((32,8),np.uint32)
    for lane in range(32):
        g,c=lane>>2,lane&3
        r=R[lane].copy()
        if nv<32:
            for k in range(4):
                for e in range(2):
                    if 8*k+2*c+e>=nv:
                        for j in range(4): r[reg_of(j,k,e)]=NEGMAX
        for I in range(32):
            j=2*(I>>4)+((I>>1)&1); k=(I>>2)&3; e=I&1
            keys[lane,I]=(r[I]&np.uint32(0xffffffe0))|np.uint32(31-(8*j+2*k+e))
        for j in range(4):
            m=max(u2f(keys[lane,reg_of(j,k,e)]) for k in range(4) for e in range(2))
            m=f2u(m)
            v=(m&np.uint32(0xffffffe7))|np.uint32((3-cc)<<3)
            rowwin[lane][j]=v if cc==0 else f2u(max(u2f(rowwin[lane][j]),u2f(v)))
        kk=keys[lane].copy()
        for j in range(4):
            if rowmasks[lane]&(1<<j):
                for k in range(4):
                    for e in range(2): kk[reg_of(j,k,e)]=NEGKEY
        for k in range(4):
            for e in range(2):
                m=f2u(max(u2f(kk[reg_of(j,k,e)]) for j in range(4)))
                w[lane,2*k+e]=(m&np.uint32(0xfffffff8))|np.uint32(7-g)
    # exchange through the [column slot][g] patch: lane (g, c) stores word 8c + g + 32e + 64k; lane L reads slot L
    mem=np.zeros(256,np.uint32)
    for lane in range(32):
        g,c=lane>>2,lane&3
        for k in range(4):
            for e in range(2): mem[8*c+g+32*e+64*k]=w[lane,2*k+e]
    return np.array([f2u(max(u2f(mem[8*L+i]) for i in range(8))) for L in range(32)],np.uint32)

def run(lx, ly, Tx=512):
    nh=(lx+127)//128
    S=rng.standard_normal((512, ((ly+127)//128)*128)).astype(np.float32)
    S[lx:]=0  # pad rows give zero scores
    argX=np.full(Tx,-1); valX=np.zeros(Tx,np.float32); argY=np.full(ly,-1); valY=np.zeros(ly,np.float32)
    ntile=(ly+127)//128
    colbuf=np.zeros((ntile,16,128),np.uint32)
    for h in range(nh):
        for q in range(4):
            st={'mk':[[NEGKEY]*4 for _ in range(32)],'base':[[0]*4 for _ in range(32)]}
            rowmasks=[sum((1<<j) for j in range(4) if h*128+q*32+(l>>2)+8*j>=lx) for l in range(32)]
            for n in range(ntile):
                s0=n*128; nvalid=min(ly-s0,128)
                cc=0
                rowwin=[[np.uint32(0)]*4 for _ in range(32)]
                while cc*32<nvalid:
                    Sc=S[h*128+q*32:h*128+q*32+32, s0+cc*32:s0+cc*32+32]
                    R=[load_chunk(Sc,l) for l in range(32)]
                    wv=chunk_both(R,cc,min(nvalid-cc*32,32),rowmasks,st,rowwin)
                    for l in range(32):
                        g,c=l>>2,l&3
                        colbuf[n,h*4+q,cc*32+8*(l>>3)+2*c+((l>>2)&1)]=wv[l]
                    cc+=1
                for l in range(32):
                    for j in range(4):
                        m=rowwin[l][j]
                        if u2f(m|np.uint32(31))>u2f(st['mk'][l][j]|np.uint32(31)):
                            st['mk'][l][j]=m; st['base'][l][j]=s0
            # finish pair
            sb=np.zeros((32,4),np.uint32); col=np.zeros((32,4),np.int64)
            for l in range(32):
                c=l&3
                for j in range(4):
                    mk=st['mk'][l][j]; v=7-int(mk&np.uint32(7)); cc2=3-int((mk>>np.uint32(3))&np.uint32(3)); sb[l,j]=mk|np.uint32(31); col[l,j]=st['base'][l][j]+32*cc2+8*(v>>1)+2*c+(v&1)
            for j in range(4):
                for o in (1,2):
                    nsb=sb[:,j].copy(); ncol=col[:,j].copy()
                    for l in range(32):
                        osb,ocol=sb[l^o,j],col[l^o,j]
                        take=u2f(osb)>u2f(sb[l,j]) or (osb==sb[l,j] and ocol<col[l,j])
                        if take: nsb[l]=osb; ncol[l]=ocol
                    sb[:,j]=nsb; col[:,j]=ncol
            for l in range(32):
                g,c=l>>2,l&3
                t=h*128+q*32+g+8*c
                if t<Tx:
                    argX[t]=col[l,c]; valX[t]=u2f((sb[l,c]&np.uint32(0xffffffe0))|np.uint32(16))
    for n in range(ntile):
        s0=n*128; nvalid=min(ly-s0,128)
        for colc in range(nvalid):
            best=colbuf[n,0,colc]; slot=0
            for wq in range(1,nh*4):
                k=colbuf[n,wq,colc]
                if u2f(k|np.uint32(31))>u2f(best|np.uint32(31)): best=k; slot=wq
            argY[s0+colc]=slot*32+31-int(best&np.uint32(31)); valY[s0+colc]=u2f((best&np.uint32(0xffffffe0))|np.uint32(16))
    Sv=S[:lx,:ly]
    rX=Sv.argmax(1); rY=Sv.argmax(0)
    okX=(argX[:lx]==rX).mean(); okY=(argY==rY).mean()
    eX=np.abs(valX[:lx]-Sv.max(1)).max(); eY=np.abs(valY-Sv.max(0)).max()
    print(lx,ly,"argX agree",okX,"argY agree",okY,"val err",eX,eY)
    assert okX>0.99 and okY>0.99 and eX<1e-4 and eY<1e-4
if __name__ == '__main__':
    run(300,200); run(77,131); run(129,33); run(1,1); run(512,512)
