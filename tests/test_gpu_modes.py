"""GPU: k_search exists in three instantiations — MODE 1 (throughput kernel, production configuration fixed at compile time),
MODE 2 (latency kernel for batches that fit the GPU at once, 16 KB row stage) and MODE 0 (run-time knobs, A/B paths). The host
picks one per call from the batch size and the environment (read once per process), so the parity tests of test_gpu_search.py /
test_gpu_fuzz.py — whose batches are small — are repeated here in child processes that force each instantiation."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SELECT = "golden or readme or reference_built or tie_heavy or smaller_than_the_beam or removed_members or duplicate_vectors"


@pytest.mark.parametrize("name,env", [
    ("throughput kernel (MODE 1)", {"GVSS_SEARCH_LATENCY": "0"}),
    ("generic kernel (MODE 0)", {"GVSS_SEARCH_LATENCY": "0", "GVSS_SEARCH_GENERIC": "1"}),
    ("generic kernel, usearch's heap as the queue", {"GVSS_SEARCH_LATENCY": "0", "GVSS_LIST_QUEUE": "0"}),
])
def test_search_parity_under_every_kernel_instantiation(name, env):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_search.py"), "-x", "-q", "-m", "gpu",
                        "-k", SELECT, "-p", "no:cacheprovider"], capture_output=True, text=True, env=e, cwd=ROOT, timeout=900)
    assert r.returncode == 0, f"{name}:\n{r.stdout[-3000:]}\n{r.stderr[-1500:]}"
    assert " passed" in r.stdout
