"""TEST INFRASTRUCTURE: elb200_gemm_c32 / _c64 (csrc/capi_complex.cpp) executed on the hostsim build — the CPU twin of
tests/test_gpu_complex.py::test_local_complex_gemm_matches_definition. Run by tests/test_hostsim.py."""
import sys, ctypes as C, itertools
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from tests.hostsim import sim_el as el
from oracle import gemm as OG
lib=el.lib
lib.cudaMalloc.argtypes=[C.POINTER(C.c_void_p), C.c_size_t]; lib.cudaMalloc.restype=C.c_int
lib.cudaFree.argtypes=[C.c_void_p]; lib.cudaDeviceSynchronize.restype=C.c_int
def dev(a):
    p=C.c_void_p(); assert lib.cudaMalloc(C.byref(p), max(a.nbytes,16))==0
    C.memmove(p, a.ctypes.data, a.nbytes); return p
def colmajor(rows, cols, ld, seed, dt, fill=0.0):
    rng = np.random.default_rng(seed)
    buf = np.full((cols, ld), fill, dtype=dt)
    buf[:, :rows] = (rng.uniform(-1, 1, (cols, rows)) + 1j * rng.uniform(-1, 1, (cols, rows))).astype(dt)
    return buf
def op(X,t): return X if t==0 else (X.T if t==1 else X.conj().T)
cnt=0
for dt in (np.complex64, np.complex128):
  R=C.c_float if dt==np.complex64 else C.c_double
  fn=lib.elb200_gemm_c32 if dt==np.complex64 else lib.elb200_gemm_c64
  for ta,tb in itertools.product([0,1,2],[0,1,2]):
    for (m,n,k) in [(37,29,53),(1,7,5),(130,1,16),(64,64,32)]:
      for pad,alpha,beta in ((0,0.5-1.5j,0.0),(3,2.0,-0.5),(1,1.0+0.25j,0.75-2.0j)):
        ar,ac=(m,k) if ta==0 else (k,m); br,bc=(k,n) if tb==0 else (n,k)
        lda,ldb,ldc=ar+pad,br+pad,m+pad
        A,B,C0=colmajor(ar,ac,lda,31,dt),colmajor(br,bc,ldb,32,dt),colmajor(m,n,ldc,33,dt,fill=5.0)
        if beta==0: C0[:,:m]=np.nan+1j*np.nan
        dA,dB,dC=dev(A),dev(B),dev(C0)
        al,be=complex(alpha),complex(beta)
        st=fn(ta,tb,m,n,k,(R*2)(al.real,al.imag),dA,lda,dB,ldb,(R*2)(be.real,be.imag),dC,ldc,None)
        assert st==0,(st,lib.elb200_last_error())
        lib.cudaDeviceSynchronize()
        out=np.empty_like(C0); C.memmove(out.ctypes.data, dC, out.nbytes)
        a,b=A[:,:ar].T.astype(np.complex128),B[:,:br].T.astype(np.complex128)
        ref=alpha*(op(a,ta)@op(b,tb))+(beta*C0[:,:m].T.astype(np.complex128) if beta!=0 else 0.0)
        got=out[:,:m].T.astype(np.complex128)
        assert np.isfinite(got).all()
        err=np.linalg.norm(got-ref)/max(np.linalg.norm(ref),1e-300)
        assert err<=OG.gemm_tolerance(dt,k),(m,n,k,ta,tb,alpha,beta,err)
        if pad: assert np.all(out[:,m:]==dt(5.0))
        for p in (dA,dB,dC): lib.cudaFree(p)
        cnt+=1
print('COMPLEX_LOCAL_PASS', cnt)
