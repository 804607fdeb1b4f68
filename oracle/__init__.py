"""TEST INFRASTRUCTURE ONLY — the CPU checkers for the HNSW hot path.

Two checkers live here and neither is ever imported by the product package ``duckdb_vss_b200``:

* ``oracle.ref``  — ctypes binding of ``oracle/_ref/libref_usearch.so``: the UNMODIFIED reference
  (vendored usearch v2.12 + DuckDB ``array_kernels.hpp``) compiled by ``oracle/Makefile`` from the
  sources where they lie under /root/reference. Present whenever the .so was built (it travels to
  the GPU box as a prebuilt file).
* ``oracle.port`` — ctypes binding of ``oracle/libhnsw_oracle.so``: a plain-C restatement of the same
  algorithms (``oracle/hnsw_oracle.c``), pinned against ``oracle.ref`` and the committed golden
  vectors by ``tests/test_oracle.py``.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu-baseline / ``--impl reference``
legs may import this package.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libref_usearch.so")
PORT_SO = os.path.join(HERE, "libhnsw_oracle.so")


def build(verbose: bool = False) -> None:
    """Compile the C restatement and (when /root/reference is mounted) the reference .so."""
    r = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout, r.stderr)
    if r.returncode:
        raise RuntimeError("oracle build failed")


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def have_port() -> bool:
    return os.path.exists(PORT_SO)
