import numpy as np
from numpy.polynomial import chebyshev as C
# minimax-ish via weighted least squares at many Chebyshev nodes, then Remez-lite iteration (Lawson)
a=np.pi/4*1.0001
def lawson(fun, basis, n=4000, iters=200):
    x=np.cos(np.pi*(np.arange(n)+0.5)/n)*a
    x=x[x>1e-9]
    A=np.stack([b(x) for b in basis],1); y=fun(x)
    w=np.ones_like(x)
    for _ in range(iters):
        W=np.sqrt(w)[:,None]
        c,*_=np.linalg.lstsq(A*W,y*W[:,0],rcond=None)
        e=np.abs(A@c-y); w=w*e; w/=w.sum()
    return c, e.max()
# sin(x) = x + x^3*(S1+S2 s+S3 s^2) ; target g(s) = (sin(x)-x)/x^3 ; weight relative to sin: error in sin = x^3*err_g ; relative = x^2*err_g/(sin/x)
cs,es=lawson(lambda x:(np.sin(x)-x)/x**3*x**2, [lambda x:x**2, lambda x:x**4, lambda x:x**6])
print('sin',[repr(np.float32(v)) for v in cs],es)
# cos(x) = 1 -0.5 s + s^2*(C1 + C2 s + C3 s^2)
cc,ec=lawson(lambda x:(np.cos(x)-1+0.5*x*x), [lambda x:x**4, lambda x:x**6, lambda x:x**8])
print('cos',[repr(np.float32(v)) for v in cc],ec)
for v in list(cs)+list(cc): print(float(np.float32(v)).hex(), repr(float(np.float32(v))))
p=np.pi/2
hi=np.float32(p); hi=np.frombuffer(np.array([hi.view(np.uint32)&0xFFFFF000],dtype=np.uint32).tobytes(),dtype=np.float32)[0]
mid=np.float32(p-float(hi)); mid=np.frombuffer(np.array([mid.view(np.uint32)&0xFFFFF000],dtype=np.uint32).tobytes(),dtype=np.float32)[0]
lo=np.float32(p-float(hi)-float(mid))
print('pio2',repr(float(hi)),repr(float(mid)),repr(float(lo)), float(hi).hex(), float(mid).hex(), float(lo).hex())
print('2/pi', repr(float(np.float32(2/np.pi))), float(np.float32(2/np.pi)).hex())
