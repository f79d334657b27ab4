"""Numpy prototype of the chunk-local recursions and their scan (carry) algebra -- development aid.
Validates the math the CUDA kernels implement against the oracle."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'tests'))
import numpy as np, oracle
from helpers import *
from celerite2_b200 import workloads as wl
P = oracle.port()

def pvec(t, c, n, lower=True):
    N = len(t)
    if lower:
        return np.exp(c*(t[n-1]-t[n])) if n > 0 else np.ones_like(c)
    return np.exp(c*(t[n]-t[n+1])) if n < N-1 else np.ones_like(c)

# ---------------- factor: chunk-local apply from carry Sigma (pre-P) ---------------------------
def factor_chunk(t,c,a,U,V,n0,n1,Sig):
    J=len(c); d=np.zeros(n1-n0); W=np.zeros((n1-n0,J)); S=np.zeros((n1-n0,J*J))
    Sig=Sig.copy()
    for n in range(n0,n1):
        p=pvec(t,c,n)
        St=p[:,None]*Sig
        S[n-n0]=St.T.reshape(-1)   # column-major flatten
        Sn=St*p[None,:]
        tmp=U[n]@Sn
        dn=a[n]-tmp@U[n]
        w=(V[n]-tmp)/dn
        Sig=Sn+dn*np.outer(w,w)
        d[n-n0]=dn; W[n-n0]=w
    return d,W,S,Sig

def factor_summary(t,c,a,U,V,n0,n1):
    """LFT composite (A,C,Jm) acting on X=-Sigma_in, returning X_out=-Sigma_out:  X_out = A X (I+Jm X)^-1 A^T + C"""
    J=len(c); A=np.eye(J); C=np.zeros((J,J)); Jm=np.zeros((J,J))
    for n in range(n0,n1):
        p=pvec(t,c,n); u=U[n]; v=V[n]
        # incoming composite maps Sigma_in -> Sigma_n (pre-P). This point: S = P Sigma P ; Sigma' = S + (v-Su)(v-Su)^T/(a-u S u)
        # in X=-Sigma terms: X' = Xp - (v+Xp u)(v+Xp u)^T/(a+u Xp u), Xp = P X P
        # first fold the scaling P: (A,C,Jm) -> (P A, P C P, Jm)
        A=p[:,None]*A; C=p[:,None]*C*p[None,:]
        q=C@u; gam=a[n]+u@q; r=A.T@u; g=(q+v)/gam
        A=A-np.outer(g,r)
        C=C-gam*np.outer(g,g)
        Jm=Jm+np.outer(r,r)/gam
    return A,C,Jm

def lft_apply(comp,X):
    A,C,Jm=comp; J=len(A)
    return A@X@np.linalg.solve(np.eye(J)+Jm@X, A.T)+C   # X (I+Jm X)^-1 = (I + X Jm)^-1 X; use sym

def lft_apply2(comp,X):
    A,C,Jm=comp; J=len(A)
    Z=np.linalg.solve(np.eye(J)+X@Jm, X)  # (I+XJ)^-1 X = X (I+JX)^-1
    return A@Z@A.T+C

def lft_combine(c1,c2):
    """apply c1 then c2"""
    A1,C1,J1=c1; A2,C2,J2=c2; J=len(A1)
    Minv=np.linalg.inv(np.eye(J)+C1@J2)
    A=A2@Minv@A1
    C=A2@Minv@C1@A2.T+C2
    Jn=A1.T@np.linalg.solve(np.eye(J)+J2@C1, J2@A1)+J1
    return A,C,Jn

def test_factor(N=3000,L=97,co=None):
    cfg=wl.config_c2(N)
    co=co or cfg['coeffs']
    t,y,diag=wl.light_curve(N,1)
    c,a,U,V=P.get_celerite_matrices(*co,t,diag)
    J=len(c)
    fl,d,W,S=P.factor(t,c,a,U,V)
    bounds=list(range(0,N,L))+[N]
    comps=[factor_summary(t,c,a,U,V,bounds[i],bounds[i+1]) for i in range(len(bounds)-1)]
    # sequential carry
    X=np.zeros((J,J)); errs=[]; carr=[]
    for i,cp in enumerate(comps):
        carr.append(-X)
        X=lft_apply2(cp,X)
    # tree: inclusive prefix via combine
    pref=[comps[0]]
    for cp in comps[1:]:
        pref.append(lft_combine(pref[-1],cp))
    d2=np.zeros(N);W2=np.zeros((N,J));S2=np.zeros((N,J*J))
    for i in range(len(bounds)-1):
        Sig = carr[i]
        if i>0:
            Sig_tree = -pref[i-1][1]  # apply to X=0 -> C
            errs.append(np.max(np.abs(Sig_tree-Sig))/max(1e-300,np.max(np.abs(Sig))))
        dd,WW,SS,_=factor_chunk(t,c,a,U,V,bounds[i],bounds[i+1],Sig)
        d2[bounds[i]:bounds[i+1]]=dd;W2[bounds[i]:bounds[i+1]]=WW;S2[bounds[i]:bounds[i+1]]=SS
    print('factor chunked: d',relerr(d2,d),'W',relerr(W2,W),'S',relerr(S2,S),'tree-vs-seq carry',max(errs))
    return t,c,a,U,V,d,W,S,y

# ---------------- forward sweep (lower) chunk-local, K columns -------------------------------
def sweep_chunk(is_solve,lower,t,c,A_,B_,Y,Zin,n0,n1,G):
    """A_=out-vector rows, B_=in-vector rows. G carry (J,K) pre-P. returns Z rows, F rows, G_out"""
    J=len(c);K=Y.shape[1]; G=G.copy()
    Z=np.zeros((n1-n0,K));F=np.zeros((n1-n0,J*K))
    rng = range(n0,n1) if lower else range(n1-1,n0-1,-1)
    for n in rng:
        p=pvec(t,c,n,lower)
        F[n-n0]=G.T.reshape(-1)  # col-major (J,K): index j+J*k
        Fp=p[:,None]*G
        if is_solve:
            z=Y[n]-A_[n]@Fp; G=Fp+np.outer(B_[n],z)
        else:
            z=Zin[n]+A_[n]@Fp; G=Fp+np.outer(B_[n],Y[n])
        Z[n-n0]=z
    return Z,F,G

def test_sweeps(t,c,U,W,V,Y,L=97):
    N,J=U.shape;K=Y.shape[1]
    bounds=list(range(0,N,L))+[N]
    for op in oracle.OPS:
        is_solve=op.startswith('solve'); lower=op.endswith('lower')
        Bm = W if is_solve else V
        Z0=np.cos(Y)
        Zr,Fr=P.sweep(op,t,c,U,Bm,Y,Z0=None if is_solve else Z0)
        A_,B_=(U,Bm) if lower else (Bm,U)
        # sequential chunk carry
        Z=np.zeros_like(Y);F=np.zeros((N,J*K))
        G=np.zeros((J,K))
        idx = range(len(bounds)-1) if lower else range(len(bounds)-2,-1,-1)
        carries={}
        for i in idx:
            carries[i]=G
            z,f,G=sweep_chunk(is_solve,lower,t,c,A_,B_,Y,Z0,bounds[i],bounds[i+1],G)
            Z[bounds[i]:bounds[i+1]]=z;F[bounds[i]:bounds[i+1]]=f
        e1=(relerr(Z,Zr),relerr(F,Fr))
        # summary-based carries: M via J basis columns, b via zero carry
        errs=[]
        G=np.zeros((J,K))
        for i in idx:
            errs.append(np.max(np.abs(G-carries[i])))
            if is_solve:
                _,_,M=sweep_chunk(True,lower,t,c,A_,B_,np.zeros((N,J)),None,bounds[i],bounds[i+1],np.eye(J))
            else:
                n0,n1=bounds[i],bounds[i+1]
                if lower:
                    M=np.diag(np.exp(c*(t[max(n0-1,0)]-t[n1-1])))
                else:
                    M=np.diag(np.exp(c*(t[n0]-t[min(n1,N-1)])))
            _,_,b=sweep_chunk(is_solve,lower,t,c,A_,B_,Y,Z0,bounds[i],bounds[i+1],np.zeros((J,K)))
            G=M@G+b
        print(op,'chunked',e1,'summary carry err',max(errs))

# ---------------- rev sweep chunk-local ---------------------------------------------------------
def sweep_rev_chunk(is_solve,lower,t,c,A_,B_,Y,Z,F,bZ,n0,n1,bF):
    """returns dict of row outputs for rows n0..n1-1, bc partial, bt contributions (own rows + edge), bF_out"""
    N=len(t);J=len(c);K=Y.shape[1];bF=bF.copy()
    bA=np.zeros((n1-n0,J));bB=np.zeros((n1-n0,J));bY=np.zeros((n1-n0,K));bt=np.zeros(N);bc=np.zeros(J)
    rng = range(n1-1,n0-1,-1) if lower else range(n0,n1)
    s = 1.0 if lower else -1.0
    for n in rng:
        if is_solve:
            bz=bZ[n]+B_[n]@bF; bB[n-n0]=bF@Z[n]; bY[n-n0]=bz
        else:
            bz=bZ[n]; bB[n-n0]=bF@Y[n]; bY[n-n0]=B_[n]@bF
        o = n-1 if lower else n+1
        if 0<=o<N:
            dt=s*(t[o]-t[n]); p=np.exp(c*dt); Fn=F[n].reshape(K,J).T
            sg = -1.0 if is_solve else 1.0
            bA[n-n0]=sg*(p[:,None]*Fn)@bz
            bF=bF+sg*np.outer(A_[n],bz)
            bp=np.sum(Fn*bF,axis=1)*p
            bc+=dt*bp; fac=c@bp
            bt[n]-=s*fac; bt[o]+=s*fac
            bF=p[:,None]*bF
    return bA,bB,bY,bt,bc,bF

def test_sweeps_rev(t,c,U,W,V,Y,L=97):
    N,J=U.shape;K=Y.shape[1]
    bounds=list(range(0,N,L))+[N]
    rng=np.random.default_rng(5); bZ=rng.standard_normal(Y.shape)
    for op in oracle.OPS:
        is_solve=op.startswith('solve'); lower=op.endswith('lower')
        Bm = W if is_solve else V
        Zr,Fr=P.sweep(op,t,c,U,Bm,Y)
        ref=P.sweep_rev(op,t,c,U,Bm,Y,Zr,Fr,bZ)  # bt,bc,bU,bW,bY
        A_,B_=(U,Bm) if lower else (Bm,U)
        bA=np.zeros((N,J));bB=np.zeros((N,J));bY=np.zeros((N,K));bt=np.zeros(N);bc=np.zeros(J)
        bF=np.zeros((J,K))
        idx = range(len(bounds)-2,-1,-1) if lower else range(len(bounds)-1)
        errs=[]; G=np.zeros((J,K))
        for i in idx:
            n0,n1=bounds[i],bounds[i+1]
            # summary route
            errs.append(np.max(np.abs(G-bF)))
            if is_solve:
                M=np.zeros((J,J))
                for j in range(J):
                    e=np.zeros((J,1));e[j]=1
                    *_,Mj=sweep_rev_chunk(True,lower,t,c,A_,B_,np.zeros((N,1)),np.zeros((N,1)),np.zeros((N,J)),np.zeros((N,1)),n0,n1,e)
                    M[:,j]=Mj[:,0]
            else:
                if lower: M=np.diag(np.exp(c*(t[max(n0-1,0)]-t[n1-1]))) if n0>0 else np.diag(np.exp(c*(t[0]-t[n1-1])))
                else: M=np.diag(np.exp(c*(t[n0]-t[min(n1,N-1)])))
            *_,b=sweep_rev_chunk(is_solve,lower,t,c,A_,B_,Y,Zr,Fr,bZ,n0,n1,np.zeros((J,K)))
            G=M@G+b
            a_,b_,y_,bt_,bc_,bF=sweep_rev_chunk(is_solve,lower,t,c,A_,B_,Y,Zr,Fr,bZ,n0,n1,bF)
            bA[n0:n1]=a_;bB[n0:n1]=b_;bY[n0:n1]=y_;bt+=bt_;bc+=bc_
        bU,bW=(bA,bB) if lower else (bB,bA)
        print(op,'_rev chunked',[relerr(u,v) for u,v in zip((bt,bc,bU,bW,bY),ref)],'summary carry err',max(errs))

# ---------------- factor_rev chunk-local with symmetric G ---------------------------------------
def factor_rev_chunk(t,c,U,d,W,S,bd,bW,n0,n1,G):
    N=len(t);J=len(c);G=G.copy()
    ba=np.zeros(n1-n0);bU=np.zeros((n1-n0,J));bV=np.zeros((n1-n0,J));bt=np.zeros(N);bc=np.zeros(J)
    for n in range(n1-1,n0-1,-1):
        w=W[n];u=U[n]
        b0=bW[n]/d[n]
        wG=w@G
        bv=b0+2*wG
        ban=bd[n]+wG@w-w@bv
        ba[n-n0]=ban;bV[n-n0]=bv
        if n>=1:
            dt=t[n-1]-t[n];p=np.exp(c*dt);Sn=S[n].reshape(J,J).T
            x=bv+2*ban*u
            bU[n-n0]=-(x@Sn)*p
            yv=bv+ban*u
            G=G-0.5*(np.outer(u,yv)+np.outer(yv,u))
            bp=2*p*np.sum(Sn*G,axis=0)
            bc+=dt*bp;fac=c@bp;bt[n]-=fac;bt[n-1]+=fac
            G=p[:,None]*G*p[None,:]
    return ba,bU,bV,bt,bc,G

def test_factor_rev(t,c,a,U,V,d,W,S,L=97):
    N,J=U.shape
    rng=np.random.default_rng(3);bd=rng.standard_normal(N);bW=rng.standard_normal((N,J))
    ref=P.factor_rev(t,c,a,U,V,d,W,S,bd,bW) # bt,bc,ba,bU,bV
    bounds=list(range(0,N,L))+[N]
    ba=np.zeros(N);bU=np.zeros((N,J));bV=np.zeros((N,J));bt=np.zeros(N);bc=np.zeros(J)
    G=np.zeros((J,J));G2=np.zeros((J,J));errs=[]
    for i in range(len(bounds)-2,-1,-1):
        n0,n1=bounds[i],bounds[i+1]
        errs.append(np.max(np.abs(G2-G)))
        # summary: Kmat = T_{n1-1}...T_{n0} via solve_lower chunk on identity (forward) ; B = run with zero
        _,_,Kmat=sweep_chunk(True,True,t,c,U,W,np.zeros((N,J)),None,n0,n1,np.eye(J))
        *_,Bc=factor_rev_chunk(t,c,U,d,W,S,bd,bW,n0,n1,np.zeros((J,J)))
        if n0==0:
            pass
        G2=Kmat.T@G2@Kmat+Bc
        a_,u_,v_,bt_,bc_,G=factor_rev_chunk(t,c,U,d,W,S,bd,bW,n0,n1,G)
        ba[n0:n1]=a_;bU[n0:n1]=u_;bV[n0:n1]=v_;bt+=bt_;bc+=bc_
    print('factor_rev chunked',[relerr(u,v) for u,v in zip((bt,bc,ba,bU,bV),ref)],'summary carry err',max(errs[:-1]) if len(errs)>1 else 0)

if __name__=='__main__':
    N=3000
    t,c,a,U,V,d,W,S,y=test_factor(N)
    Y=np.stack([y,np.cos(t)],axis=1)
    test_sweeps(t,c,U,W,V,Y)
    test_sweeps_rev(t,c,U,W,V,Y)
    test_factor_rev(t,c,a,U,V,d,W,S)
