import sys; sys.path.insert(0,'.')
import numpy as np, oracle, low_light_image_enhancement_b200 as L
x = oracle.synth_lowlight(400, 600, 0)
y = L.enhance(x, "msrcr")
xc = np.ascontiguousarray(x.transpose(2, 0, 1))[None]
yc = L.enhance(xc, "msrcr", dict(layout="NCHW"))[0].transpose(1,2,0)
d = np.abs(yc - y)
print('max', d.max(), 'count>1e-4', (d>1e-4).sum(), 'of', d.size)
ys, xs, cs = np.nonzero(d > 1e-4)
if len(ys):
    print('rows', ys.min(), ys.max(), 'cols', xs.min(), xs.max(), 'ch', np.bincount(cs, minlength=3))
    print('row hist', np.bincount(ys // 32)); print('col hist', np.bincount(xs // 64))
    print(d[:4,:4,:])
