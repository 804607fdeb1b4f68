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
    # (debugging knob: the heap prunes dead entries when its shared-memory part fills up, which may reorder EXACT ties — so the
    # all-ties cases stay out of this one)
    ("generic kernel, usearch's heap as the queue", {"GVSS_SEARCH_LATENCY": "0", "GVSS_LIST_QUEUE": "0", "SELECT": "reference_built or smaller_than_the_beam"}),
])
def test_search_parity_under_every_kernel_instantiation(name, env):
    e = dict(os.environ)
    e.update(env)
    select = e.pop("SELECT", SELECT)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_search.py"), "-x", "-q", "-m", "gpu",
                        "-k", select, "-p", "no:cacheprovider"], capture_output=True, text=True, env=e, cwd=ROOT, timeout=900)
    assert r.returncode == 0, f"{name}:\n{r.stdout[-3000:]}\n{r.stderr[-1500:]}"
    assert " passed" in r.stdout


_CHILD = r"""
import sys, hashlib, numpy as np
sys.path.insert(0, %r)
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
base, q = synth.clustered(30000, 32, 300, 8, 5, 400)
ix = gv.Index(32, "l2sq", M=16)
ix.reserve(30000, 1); ix.set_build_concurrency(1); ix.add_batch(base); ix.flush()
keys, dist, counts, st = ix.ef_search(q, 10, 512, stats=True)
print("RESULT", hashlib.sha1(keys.tobytes() + dist.tobytes() + st["computed"].tobytes()).hexdigest(), float(np.mean((st["flags"] & 2) != 0)), int((st["flags"] & 1).sum()))
"""


def test_overfilled_visited_table_falls_back_to_the_exact_bitmap_with_identical_results():
    """A query that would push its hashed visited table past 75 % load is redone inside the kernel with the exact per-warp bitmap
    (stats flag bit1). GVSS_VIS_FACTOR=1 makes the table far too small for ef = 512, so most queries take that path: ids,
    distance bits and traversal counters must equal the run with the normally sized table."""
    outs = []
    for env in ({}, {"GVSS_VIS_FACTOR": "1"}):
        e = dict(os.environ)
        e.update(env)
        r = subprocess.run([sys.executable, "-c", _CHILD % ROOT], capture_output=True, text=True, env=e, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")][0].split()
        outs.append((line[1], float(line[2]), int(line[3])))
    assert outs[0][1] == 0.0 and outs[1][1] > 0.2, outs   # the small table really overflowed
    assert outs[0][2] == 0 and outs[1][2] == 0, outs      # and no query was cut short
    assert outs[0][0] == outs[1][0], outs
