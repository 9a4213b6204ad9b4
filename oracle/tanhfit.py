import numpy as np
from numpy.polynomial import chebyshev as C
# fit g(t) = (tanh(x)/x - 1)/x^2  as polynomial in t=x^2 on [0, 0.55^2], relative error weighting
T=0.55
n=4000
k=np.arange(n)
xs = T*np.sqrt((np.cos(np.pi*(k+0.5)/n)+1)/2)  # chebyshev nodes in t
xs = xs[xs>1e-4]
t=xs**2
g=(np.tanh(xs)/xs-1)/t
for deg in (3,4,5):
    A=np.vander(t,deg+1,increasing=True)
    # minimize relative error of tanh: error in tanh/x = t*err(g) ; weight by t
    w=t
    c,*_=np.linalg.lstsq(A*w[:,None], g*w, rcond=None)
    # evaluate in float32 horner
    c32=c.astype(np.float32)
    xx=np.linspace(1e-6,T,2000001).astype(np.float32)
    x2=(xx*xx).astype(np.float32)
    p=np.full_like(xx,c32[-1])
    for ci in c32[-2::-1]:
        p=(p*x2+ci).astype(np.float32)   # not fma but close
    r=(xx + (xx*x2).astype(np.float32)*p).astype(np.float32)
    ref=np.tanh(xx.astype(np.float64))
    ulp=np.abs(r-ref)/np.spacing(ref.astype(np.float32))
    print(deg,[float(v) for v in c32], 'max ulp',ulp.max())
    print('   ', [v.hex() for v in c32.astype(float)])
