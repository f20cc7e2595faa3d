"""The reference's own known-answer test for the quadrature rules (tests/test_metrics.py:37-45,
`TestMetrics.test_weight_normalization`: the integral of 1 over the sphere is 1 at rtol 1e-5 for naive /
clenshaw-curtiss / legendre-gauss weights on 20 x 101 and 721 x 1440), run UNMODIFIED: its `GridQuadrature`
(makani/utils/grids.py:26-97, unmodified) takes `legendre_gauss_weights` / `clenshaw_curtiss_weights` from
`torch_harmonics.quadrature`, which the harness resolves to makani_b200.quadrature.
The file's other tests need xarray / xskillscore (not installed, unrelated to the path): inert stand-ins satisfy the
two imports and only the quadrature test is selected."""
import os
import sys
import types
import unittest
import warnings

warnings.filterwarnings("ignore")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import refharness as H                                                     # noqa: E402

H.install("product")
for name in ("xarray", "xskillscore"):
    if name not in sys.modules:
        sys.modules[name] = types.ModuleType(name)
H._path_package("makani.utils.metrics", os.path.join(H.REF_ROOT, "makani", "utils", "metrics"))
mod = H.load_file("tests/test_metrics.py", "reference_test_metrics")
grids = sys.modules["makani.utils.grids"]
import makani_b200.quadrature as q                                          # noqa: E402
assert grids.__file__.startswith("/root/reference/")
assert grids.legendre_gauss_weights is q.legendre_gauss_weights and grids.clenshaw_curtiss_weights is q.clenshaw_curtiss_weights
suite = unittest.TestSuite([mod.TestMetrics("test_weight_normalization")])
res = unittest.TextTestRunner(stream=sys.stderr, verbosity=2).run(suite)
ok = res.wasSuccessful() and res.testsRun == 1
print("reference test_weight_normalization on product quadrature:", "OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
