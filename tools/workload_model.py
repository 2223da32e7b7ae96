import numpy as np
N=2048; H=1024
rng=np.random.default_rng(0)
def negacyclic_ref(a,b):
    # a int digits, b int (python ints) mod 2^64
    res=[0]*N
    for i in range(N):
        if a[i]==0: continue
        for j in range(N):
            k=i+j
            if k<N: res[k]+=a[i]*b[j]
            else: res[k-N]-=a[i]*b[j]
    return res
c=np.exp(2j*np.pi*np.arange(32)/128)
T=np.exp(2j*np.pi*np.outer(np.arange(32)+0.25,np.arange(32))/1024)  # T[k2][n1]
def dft32(x,sign):
    W=np.exp(sign*2j*np.pi*np.outer(np.arange(32),np.arange(32))/32)
    return W@x
def fwd(p):
    z=p[:H]+1j*p[H:]
    zz=z.reshape(32,32)  # zz[n2][n1]
    Y=np.zeros((32,32),complex) # Y[n1][k2]
    for n1 in range(32):
        Y[n1,:]=dft32(zz[:,n1]*c,+1)
    Y=Y*T.T
    X=np.zeros((32,32),complex) # X[k1][k2]
    for k2 in range(32):
        X[:,k2]=dft32(Y[:,k2],+1)
    return X
def inv(X):
    W=np.zeros((32,32),complex) # W[n1][k2]
    for k2 in range(32):
        W[:,k2]=dft32(X[:,k2],-1)
    W=W*np.conj(T.T)
    u=np.zeros((32,32),complex) # u[n2][n1]
    for n1 in range(32):
        u[:,n1]=dft32(W[n1,:],-1)*np.conj(c)
    u=u.reshape(H)/1024
    return np.concatenate([u.real,u.imag])
a=rng.integers(-5,6,N).astype(float); b=rng.integers(-100,100,N).astype(float)
r=inv(fwd(a)*fwd(b))
ref=np.array(negacyclic_ref([int(x) for x in a],[int(x) for x in b]),float)
print(np.abs(r-ref).max())
# direct formula check
z=a[:H]+1j*a[H:]
j=np.arange(H)
X=fwd(a)
k=5*32+7
print(abs(X[5,7]-np.sum(z*np.exp(2j*np.pi*j*(k+0.25)/1024))))
