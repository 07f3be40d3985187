import time, ctypes as C, os, sys
t0=time.time()
L=C.CDLL("dekupl-run_b200/libdekupl_b200.so")
t1=time.time()
ctx=C.c_void_p()
L.dk_init.argtypes=[C.POINTER(C.c_void_p), C.c_int]
rc=L.dk_init(C.byref(ctx),0)
t2=time.time()
print("dlopen %.3f s, dk_init %.3f s rc=%d"%(t1-t0,t2-t1,rc))
