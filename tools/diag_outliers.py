"""Where are the largest |delta_gpu - delta_oracle| of a big grid?  usage: diag_outliers.py n [streamed]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, scipy.fft
import oracle
from randomfield_b200 import _lib, powertools
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
shape = (n, n, n)
power = oracle.load_default_power()
eng = _lib.Engine(*shape)
k, sig = powertools.sigma_table(power, shape, 2.0)
eng.set_power(k, sig, 2.0)
print("kernels", eng.kernel_names, flush=True)
eng.generate_delta(123)
d = eng.copy_field()
kf = oracle.philox_kfield_big(shape, 2.0, power, 123)
t = scipy.fft.irfftn(kf, s=shape, axes=(0, 1, 2), workers=-1, overwrite_x=True)
del kf
err = np.empty(shape, dtype=np.float32)
for x0 in range(0, n, 64):
    err[x0:x0 + 64] = d[x0:x0 + 64] - t[x0:x0 + 64]
std = t.std()
print("rel rms %.3e  max %.3e (x std)" % (np.sqrt(np.mean(err.astype(np.float64) ** 2)) / std, np.abs(err).max() / std))
flat = np.argsort(np.abs(err).ravel())[-20:][::-1]
for f in flat:
    ix, iy, iz = np.unravel_index(f, shape)
    print(ix, iy, iz, "err %.3e gpu %.5f oracle %.5f" % (err[ix, iy, iz], d[ix, iy, iz], t[ix, iy, iz]))
big = np.abs(err) > 20 * np.sqrt(np.mean(err.astype(np.float64) ** 2))
print("cells above 20 sigma_err:", int(big.sum()))
if big.sum():
    w = np.argwhere(big)
    for ax, name in enumerate("xyz"):
        u, c = np.unique(w[:, ax], return_counts=True)
        print(name, "distinct", len(u), "top", sorted(zip(c, u), reverse=True)[:8])
# error spectrum along z of the worst line
ix, iy, iz = np.unravel_index(flat[0], shape)
line = err[ix, iy, :].astype(np.float64)
print("worst line rms %.3e, field-wide rms %.3e" % (line.std(), err.std()))
fz = np.abs(np.fft.rfft(line))
print("worst line |FFT| top modes", np.argsort(fz)[-5:][::-1], np.sort(fz)[-5:][::-1])
plane = err[ix].astype(np.float64)
print("worst x-plane rms %.3e" % plane.std())
