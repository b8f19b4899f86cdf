import numpy as np
from scipy.special import erfc, erf
amax = 6.0
a = np.linspace(0, amax, 20001)
M = np.log2(erfc(a/np.sqrt(2))) - 1.0     # includes the 0.5 factor
w = a*0.5*erfc(a/np.sqrt(2)) + 1e-4
xs = np.linspace(-9, 9, 900001)
ref = xs*0.5*(1+erf(xs/np.sqrt(2)))
for deg in (5,6,7,8,9):
    # iterative reweighted LS toward minimax of weighted error
    W = w.copy()
    for it in range(30):
        co = np.polyfit(a, M, deg, w=W)
        err = np.abs((np.polyval(co, a)-M)*w)
        W = W*(1+ 2*err/err.max())
        W /= W.max()
    co32 = co.astype(np.float32)
    x = xs.astype(np.float32); aa = np.minimum(np.abs(x), np.float32(amax))
    r = np.full_like(aa, co32[0])
    for c in co32[1:]:
        r = (r*aa + c).astype(np.float32)
    E = np.exp2(r.astype(np.float32)).astype(np.float32)
    out = (np.maximum(x,0) - aa*E).astype(np.float32)
    # note for |x|>amax: uses aa (clamped) - negligible
    e = np.abs(out-ref)
    print(deg, "max abs %.2e at %.3f" % (e.max(), xs[e.argmax()]), "max rel(|x|<3) %.2e" % (e[np.abs(xs)<3]/np.maximum(np.abs(ref[np.abs(xs)<3]),1e-3)).max())
    print("   ", ", ".join("%.9ef" % c for c in co32))
