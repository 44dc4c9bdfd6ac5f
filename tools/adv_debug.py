"""Diagnostic: worst normal-angle offenders of the tier chain against the oracle on the adversarial cloud of
tests/test_gpu_parity.py (O3DPP_KNN_LIST=1 runs the reference-order list kernel instead)."""
import sys, os
sys.path.insert(0, 'open3d-py-extension_b200'); sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from open3dpypro import _native as nat
from oracle import orc
from test_gpu_parity import _adversarial, _angle
pts = _adversarial()
c = nat.Context(0)
c.upload(nat.ATTR_POINTS, pts)
for mode, radius in ((0, 0.0), (2, 0.3)):
    c.drop(nat.ATTR_NORMALS)
    c.estimate_normals(mode, 30, radius)
    got = c.download(nat.ATTR_NORMALS)
    ref, cov = orc.estimate_normals(pts, mode={0: "knn", 2: "hybrid"}[mode], k=30, radius=radius, want_cov=True)
    Cm = np.empty((len(pts), 3, 3))
    Cm[:, 0, 0], Cm[:, 0, 1], Cm[:, 0, 2] = cov[:, 0], cov[:, 1], cov[:, 2]
    Cm[:, 1, 0], Cm[:, 1, 1], Cm[:, 1, 2] = cov[:, 1], cov[:, 3], cov[:, 4]
    Cm[:, 2, 0], Cm[:, 2, 1], Cm[:, 2, 2] = cov[:, 2], cov[:, 4], cov[:, 5]
    lam = np.linalg.eigvalsh(Cm)
    scale = np.maximum(np.abs(lam).max(1), 1e-300)
    gap = (lam[:, 1] - lam[:, 0]) / scale
    ang = _angle(got, ref)
    well = gap > 1e-6
    bad = np.nonzero(well & (ang > 1e-4))[0]
    print("mode", mode, "bad", len(bad), "env list", os.environ.get("O3DPP_KNN_LIST"))
    for i in bad[np.argsort(-ang[bad])][:8]:
        print("  i", i, "p", pts[i], "ang %.3e gap %.3e scale %.3e" % (ang[i], gap[i], scale[i]), "cancel %.1e" % ((pts[i] ** 2).max() / scale[i]))
