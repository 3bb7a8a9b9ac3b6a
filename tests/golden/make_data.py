"""Regenerate randomfield_b200/data/default_power.npy from the reference's bundled table
(randomfield/data/default_power.dat, powertools.py:167-197).  Build-container only."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_replay
p = ref_replay.default_power()
out = np.stack([p['k'], p['Pk']], axis=1).astype(np.float64)
np.save(os.path.join(ROOT, "randomfield_b200", "data", "default_power.npy"), out)
print(out.shape, out[0], out[-1])
