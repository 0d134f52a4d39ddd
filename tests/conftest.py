import os
import sys

import pytest

# build every device code layout so the tests can force each scan variant (A/B parity)
os.environ.setdefault("LB200_LAYOUTS", "all")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle as o
    o.build()
    return o


_LANCE_DIR = None


def lance_files_dir():
    """Directory holding the reference-written `index.idx` fixtures as files (`<tag>_<uuid8>_index.idx`) next to the
    datasets' raw vectors.  The index files' bytes are kept in one archive (tests/golden/lance_files/
    reference_index_files.npz, made by tests/golden/make_lance_fixture.py) and written out here unchanged."""
    global _LANCE_DIR
    if _LANCE_DIR is None:
        import shutil
        import tempfile

        import numpy as np
        src = os.path.join(ROOT, "tests", "golden", "lance_files")
        out = tempfile.mkdtemp(prefix="lance_files_")
        z = np.load(os.path.join(src, "reference_index_files.npz"))
        for key in z.files:
            with open(os.path.join(out, f"{key}_index.idx"), "wb") as f:
                f.write(z[key].tobytes())
        for name in os.listdir(src):
            if name.endswith("_raw_vectors.npz"):
                shutil.copyfile(os.path.join(src, name), os.path.join(out, name))
        _LANCE_DIR = out
    return _LANCE_DIR
