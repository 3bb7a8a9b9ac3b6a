"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

Every array is produced by oracle/ref_replay.py, i.e. by /root/reference/randomfield's own
transform.py / powertools.py / random.py.  The fixtures are small (sub-sampled where the
grid is large) and pin oracle/oracle.py on machines that do not have the reference tree.
Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_replay  # noqa: E402

CASES = {
    # name: (shape, spacing, seed, smoothing, dtype, sub-sampling strides for delta / kfield)
    "ref_64_f32": ((64, 64, 64), 2.0, 123, 0.0, np.complex64, (4, 4, 4), (8, 8, 4)),
    "ref_64_f32_smooth": ((64, 64, 64), 2.0, 124, 3.0, np.complex64, (4, 4, 4), (8, 8, 4)),
    "ref_128x64x256_f32": ((128, 64, 256), 2.0, 5, 0.0, np.complex64, (8, 8, 16), (16, 8, 16)),
    "ref_16x8x12_f64": ((16, 8, 12), 2.5, 7, 1.0, np.complex128, (1, 1, 1), (1, 1, 1)),
    "ref_12x10x8_f32": ((12, 10, 8), 2.5, 9, 0.0, np.complex64, (1, 1, 1), (1, 1, 1)),
}


def main():
    for name, (shape, spacing, seed, smoothing, dtype, sd, sk) in CASES.items():
        r = ref_replay.replay(*shape, spacing, seed=seed, smoothing=smoothing, dtype=dtype,
                              save_potential=True)
        np.savez(os.path.join(HERE, name + ".npz"),
                 shape=np.array(shape), spacing=spacing, seed=seed, smoothing=smoothing,
                 dtype=np.dtype(dtype).name, delta_stride=np.array(sd), kfield_stride=np.array(sk),
                 delta=r["delta"][::sd[0], ::sd[1], ::sd[2]],
                 kfield=r["kfield"][::sk[0], ::sk[1], ::sk[2]],
                 potential=r["potential"][::sk[0], ::sk[1], ::sk[2]],
                 rms=r["rms"], mean=np.mean(r["delta"], dtype=np.float64))
        print(name, r["rms"])
    # symmetrize, packed and unpacked (transform.py:114-158)
    transform = ref_replay.modules()[0]
    rng = np.random.RandomState(11)
    out = {}
    for packed, shp in ((True, (8, 6, 5)), (False, (8, 6, 8))):
        for dt in (np.complex64, np.complex128):
            a = (rng.normal(size=shp) + 1j * rng.normal(size=shp)).astype(dt)
            b = a.copy()
            transform.symmetrize(b, packed=packed)
            key = "%s_%s" % ("packed" if packed else "full", np.dtype(dt).name)
            out[key + "_in"], out[key + "_out"] = a, b
    np.savez(os.path.join(HERE, "ref_symmetrize.npz"), **out)
    # calculate_lensing_potential (generate.py:353-418): the reference's own loop, unmodified, on stub
    # attributes (oracle/ref_replay.py: lensing_replay); scipy.integrate.simps is the restated rule
    out = {}
    rng = np.random.RandomState(17)
    for name, shp, imin, ok0, dt in (("flat_f32", (5, 4, 48), None, 0.0, np.float32), ("open_f32", (4, 3, 33), 3, 0.02, np.float32),
                                     ("closed_f64", (3, 4, 40), 0, -0.015, np.float64), ("odd_f64", (2, 3, 17), 1, 0.0, np.float64)):
        nz = shp[2]
        dphi = (1e-33 * rng.normal(size=shp)).astype(dt)
        DC = np.arange(nz) * 2.5
        DA = DC * (1 + 0.002 * np.arange(nz))
        psi = ref_replay.lensing_replay(dphi, DC, DA, Ok0=ok0, i_min=imin)
        out[name + "_dphi"], out[name + "_DC"], out[name + "_DA"], out[name + "_psi"] = dphi, DC, DA, psi
        out[name + "_par"] = np.array([-1 if imin is None else imin, ok0])
    np.savez(os.path.join(HERE, "ref_lensing.npz"), **out)
    # first deviates of RandomState(123)
    np.savez(os.path.join(HERE, "ref_misc.npz"),
             normal123=np.random.RandomState(123).normal(size=8))
    print("done")


if __name__ == "__main__":
    main()
