import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
for p in (str(REPO), str(REPO / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Checkers (oracle, emulator) are test infrastructure: build them on demand.  The product
    libraries are built by __graft_entry__.build() / `make product`."""
    import subprocess
    lib = REPO / "owl_path_tracer_b200" / "lib"
    if not (lib / "libb200pt.so").exists() or not (lib / "libb200pt_host.so").exists():
        subprocess.run(["make", "-s", "product"], cwd=REPO, check=True)
    from checkers import ensure_built
    ensure_built()
