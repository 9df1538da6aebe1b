import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ani1xnr_path():
    p = os.path.join(ROOT, "lammps_ani_b200", "data", "ani1x_nr.aniw")
    if not os.path.exists(p):
        if os.path.exists("/root/reference/external/ani-1xnr/model/ani-1xnr.info"):
            from lammps_ani_b200 import extract
            os.makedirs(os.path.dirname(p), exist_ok=True)
            extract.extract("/root/reference/external/ani-1xnr/model/ani-1xnr.info", p)
        else:
            pytest.skip("ANI-1xnr parameters not extracted (run tools/make_golden.py where /root/reference exists)")
    return p


@pytest.fixture(scope="session")
def ani2x_synth_path(tmp_path_factory):
    """ANI-2x geometry with seeded random weights (real ANI-2x parameters are not available offline)."""
    from lammps_ani_b200 import aniw, extract
    p = str(tmp_path_factory.mktemp("models") / "ani2x_synth.aniw")
    aniw.save(extract.synthetic_ani2x(seed=2024), p)
    return p


@pytest.fixture(scope="session")
def water30():
    import json
    import numpy as np
    with open(os.path.join(ROOT, "tests", "golden", "water30_reference.json")) as f:
        ref = json.load(f)
    x = np.array(ref["coords"]).reshape(-1, 3)
    sp = np.array(ref["species"], dtype=np.int32)
    idx12 = np.array(ref["atom_index12"], dtype=np.int64).reshape(2, -1)
    return dict(x=x, species=sp, atom_index12=idx12, ref=ref)
