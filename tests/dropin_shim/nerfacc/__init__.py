"""`nerfacc` as a thin alias of nerfacc_b200 — what a user does to switch (`import nerfacc_b200
as nerfacc`), spelled as a package so that `from nerfacc.volrend import ...` also resolves.
Only used by tests/test_gpu_dropin.py (on PYTHONPATH of the subprocess that runs the reference's
own, unmodified test files)."""
import sys

import nerfacc_b200
from nerfacc_b200 import *  # noqa: F401,F403  (the 22 public names, nerfacc/__init__.py:23-46)
from nerfacc_b200 import __all__, __version__  # noqa: F401
from nerfacc_b200 import data_specs, estimators, grid, pack, pdf, scan, volrend
from nerfacc_b200.estimators import occ_grid, prop_net

for _name, _mod in (("data_specs", data_specs), ("estimators", estimators), ("grid", grid), ("pack", pack),
                    ("pdf", pdf), ("scan", scan), ("volrend", volrend), ("estimators.occ_grid", occ_grid),
                    ("estimators.prop_net", prop_net)):
    sys.modules[f"{__name__}.{_name}"] = _mod
